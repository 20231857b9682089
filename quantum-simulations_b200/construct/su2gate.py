"""Single-qubit rotation gate ``R_n(angle) = cos(angle/2) I - i sin(angle/2) (n . sigma)``.

Mirror of the reference's ``RGate``/``RAxis`` (``quompiler/construct/su2gate.py:11-132``)
and of ``rot`` (``quompiler/utils/su2fun.py:145-157``): same constructor
arguments (principal axis enum, 3-vector, or spherical ``(theta, phi)`` pair),
same snapping of (anti)parallel-to-principal axes, same ``.matrix``.
"""
import numpy as np

from .types import PrincipalAxis

_PAULI = np.array([[[0, 1], [1, 0]], [[0, -1j], [1j, 0]], [[1, 0], [0, -1]]], dtype=np.complex128)


def rot(n, angle: float) -> np.ndarray:
    n = np.asarray(n, dtype=float)
    assert n.shape == (3,)
    n = n / np.linalg.norm(n)
    sigma = np.tensordot(n, _PAULI, axes=1)
    return np.cos(angle / 2) * np.eye(2) - 1j * np.sin(angle / 2) * sigma


class RAxis:
    def __init__(self, axis):
        if isinstance(axis, PrincipalAxis):
            v = np.array(axis.value, dtype=float)
        else:
            v = np.asarray(axis, dtype=float)
            if v.shape == (2,):
                th, ph = v
                v = np.array([np.sin(th) * np.cos(ph), np.sin(th) * np.sin(ph), np.cos(th)])
            elif v.shape == (3,):
                norm = np.linalg.norm(v)
                if np.isclose(norm, 0):
                    raise ValueError("axis must not be zero.")
                v = v / norm
            else:
                raise ValueError('axis must be either a 2-vector or 3-vector')
        self.nvec = v
        self.theta = float(np.arccos(np.clip(v[2], -1, 1)))
        self.phi = float(np.arctan2(v[1], v[0]))
        self.principal = PrincipalAxis.get(v)

    def isprincipal(self) -> bool:
        return self.principal is not None

    def spherical(self):
        return self.theta, self.phi

    def __eq__(self, other):
        if not isinstance(other, RAxis):
            return False
        if (self.isprincipal() or other.isprincipal()) and self.principal == other.principal:
            return True
        return bool(np.allclose(self.nvec, other.nvec))

    def __repr__(self):
        return repr(self.principal) if self.isprincipal() else f"({self.theta:.6g}, {self.phi:.6g})"


class RGate:
    def __init__(self, angle: float, axis):
        if isinstance(axis, RAxis):
            self.axis = axis
        elif isinstance(axis, PrincipalAxis):
            self.axis = RAxis(axis)
        elif isinstance(axis, (np.ndarray, tuple, list)):
            hit = PrincipalAxis.get_prop(axis)
            if hit is not None:
                principal, ratio = hit
                if np.real(ratio) < 0:
                    angle = -angle
                self.axis = RAxis(principal)
            else:
                self.axis = RAxis(axis)
        else:
            raise ValueError('axis must be either an np.array, sequence, or enum PrincipalAxis')
        assert np.isclose(np.imag(angle), 0)
        self.angle = float(np.real(angle))
        self.matrix = rot(self.axis.nvec, self.angle)

    def __array__(self, dtype=None, copy=None):
        return self.matrix if dtype is None else self.matrix.astype(dtype)

    def __matmul__(self, other):
        if isinstance(other, np.ndarray):
            return self.matrix @ other
        if not isinstance(other, RGate):
            raise NotImplementedError(f'matmul only implemented for {other}')
        if self.axis == other.axis:
            return RGate(self.angle + other.angle, self.axis)
        raise NotImplementedError('product of rotations about different axes is a compile-time helper, not part of the simulation path')

    def __repr__(self):
        return f"R{self.axis!r}({self.angle:.6g})"
