"""Enumerations shared by the gate types.

Host-side mirror of the reference's ``quompiler/construct/types.py``:
``QType`` (102-123), ``UnivGate`` and its exact matrices (10-20, 23-99),
``GateGrain`` (126-143), ``QompilePlatform`` (146-152, extended here with the
``B200`` target) and ``PrincipalAxis`` (175-219).  Only what the simulation path
consumes is kept; semantics (names, values, ordering, matrices) are identical.
"""
from enum import Enum, IntEnum, IntFlag

import numpy as np

_SQ = 1.0 / np.sqrt(2.0)

_GATE_MATRICES = {
    'I': np.array([[1, 0], [0, 1]], dtype=np.complex128),
    'X': np.array([[0, 1], [1, 0]], dtype=np.complex128),
    'Y': np.array([[0, -1j], [1j, 0]], dtype=np.complex128),
    'Z': np.array([[1, 0], [0, -1]], dtype=np.complex128),
    'H': np.array([[1, 1], [1, -1]], dtype=np.complex128) / np.sqrt(2),
    'S': np.array([[1, 0], [0, 1j]], dtype=np.complex128),
    'T': np.array([[1, 0], [0, np.exp(1j * np.pi / 4)]], dtype=np.complex128),
    'SD': np.array([[1, 0], [0, -1j]], dtype=np.complex128),
    'TD': np.array([[1, 0], [0, np.exp(-1j * np.pi / 4)]], dtype=np.complex128),
}

_HERM = {'S': 'SD', 'SD': 'S', 'T': 'TD', 'TD': 'T'}


def proportional(a: np.ndarray, b: np.ndarray, rtol=1e-5, atol=1e-8):
    """``a == lam * b`` for a scalar ``lam``?  Returns ``(ok, lam)``.

    Same test as the reference's ``allprop`` (``quompiler/utils/mfun.py:34-52``):
    the ratio is taken at the first entry where ``b`` is not ~0.
    """
    a = np.asarray(a)
    b = np.asarray(b)
    if a.shape != b.shape:
        return False, 1.0
    nz = ~np.isclose(b, 0)
    if not nz.any():
        return bool(np.allclose(a, 0, rtol=rtol, atol=atol)), 0
    lam = (a[nz] / b[nz])[0]
    return bool(np.allclose(a, b * lam, rtol=rtol, atol=atol)), lam


class UnivGate(Enum):
    I = 'I'
    X = 'X'
    Y = 'Y'
    Z = 'Z'
    H = 'H'
    S = 'S'
    T = 'T'
    SD = 'SD'
    TD = 'TD'

    def __array__(self, dtype=None, copy=None):
        m = _GATE_MATRICES[self.value]
        return m.astype(dtype) if dtype is not None else m.copy()

    def __repr__(self):
        return self.name

    def __lt__(self, other):
        return self.name < other.name

    def __reduce__(self):
        return self.__class__, (self.name,)

    def __matmul__(self, other):
        return np.array(self) @ np.asarray(other)

    def __rmatmul__(self, other):
        return np.asarray(other) @ np.array(self)

    def herm(self) -> 'UnivGate':
        return UnivGate[_HERM.get(self.name, self.name)]

    @staticmethod
    def _unitary2(m) -> bool:
        return m.shape == (2, 2) and np.allclose(m.conj() @ m.T, np.eye(2))

    @staticmethod
    def get(m):
        """Exact (to tolerance) match of a 2x2 matrix against the nine gates."""
        m = np.asarray(m)
        if not UnivGate._unitary2(m):
            return None
        for g in UnivGate:
            if np.allclose(m, _GATE_MATRICES[g.value]):
                return g
        return None

    @staticmethod
    def get_prop(m):
        """Match up to a unit scalar, first hit in declaration order (types.py:76-87)."""
        m = np.asarray(m)
        if not UnivGate._unitary2(m):
            return None
        for g in UnivGate:
            ref = _GATE_MATRICES[g.value]
            if np.allclose(m, ref) or proportional(m, ref)[0]:
                return g
        return None

    def rotation(self, theta):
        """exp(-i theta/2 G) for the involutory gates (G^2 = I) in closed form."""
        g = np.array(self)
        if np.allclose(g @ g, np.eye(2)):
            return np.cos(theta / 2) * np.eye(2) - 1j * np.sin(theta / 2) * g
        from scipy.linalg import expm
        return expm(-.5j * theta * g)

    @staticmethod
    def cliffordt():
        return [UnivGate.I, UnivGate.X, UnivGate.H, UnivGate.Z, UnivGate.S, UnivGate.T, UnivGate.SD, UnivGate.TD]


class QType(IntFlag):
    """Role of a qubit in a gate; ``base`` = the bit values the role spans."""
    TARGET = (1, (0, 1))
    CONTROL0 = (2, (0,))
    CONTROL1 = (4, (1,))
    IDLER = (8, (0, 1))

    def __new__(cls, value, base=None):
        obj = int.__new__(cls, value)
        obj._value_ = value
        obj.base = base
        return obj

    def __repr__(self):
        return self.name if self.name else repr(int(self))


def qtype_code(c) -> int:
    """Integer role code of a QType from this package or from the reference."""
    return int(c)


TARGET, CONTROL0, CONTROL1, IDLER = 1, 2, 4, 8


class GateGrain(IntEnum):
    INVALID = 0
    UNITARY = 0x20
    TWO_LEVEL = 0x40
    MULTI_TARGET = 0x60
    SINGLET = 0x80
    CTRL_PRUNED = 0xa0
    PRINCIPAL = 0xb0
    UNIV_GATE = 0xc0
    CLIFFORD_T = 0xe0

    def __repr__(self):
        return self.name


class QompilePlatform(Enum):
    CIRQ = 'CirqBuilder'
    QISKIT = 'QiskitBuilder'
    QUIMB = 'QuimbBuilder'
    B200 = 'B200Builder'  # new target: the CUDA statevector engine of this repo
    QASM = 'QasmBuilder'  # new target: OpenQASM 3 text (what the demos get from qiskit's qasm.dumps)

    def __repr__(self):
        return self.name


class PrincipalAxis(Enum):
    X = (1, 0, 0)
    Y = (0, 1, 0)
    Z = (0, 0, 1)

    def __repr__(self):
        return self.name.lower()

    @staticmethod
    def _vec(axis):
        axis = np.asarray(axis, dtype=float)
        assert axis.shape in ((2,), (3,))
        if axis.shape == (2,):
            th, ph = axis
            axis = np.array([np.sin(th) * np.cos(ph), np.sin(th) * np.sin(ph), np.cos(th)])
        return axis

    @staticmethod
    def get(axis):
        v = PrincipalAxis._vec(axis)
        for k in PrincipalAxis:
            if np.allclose(v, np.array(k.value)):
                return k
        return None

    @staticmethod
    def get_prop(axis):
        """``(axis, factor)`` if ``axis`` is (anti)parallel to x, y or z, else ``None``."""
        v = PrincipalAxis._vec(axis)
        for k in PrincipalAxis:
            ok, lam = proportional(v, np.array(k.value, dtype=float))
            if ok:
                return k, lam
        return None
