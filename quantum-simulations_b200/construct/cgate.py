"""``CtrlGate``: a k-qubit unitary with per-qubit roles (target / control-on-0 / control-on-1).

Host-side mirror of ``quompiler/construct/cgate.py:18-331``.  Constructor
contract, role accessors, ``sorted`` (160-184), ``expand`` (195-218),
``promote`` (307-331), ``herm`` and ``convert`` (133-158) behave as in the
reference but never build anything larger than the gate's own ``2^t x 2^t``
core.  Everything the reference computes with dense ``2^K x 2^K`` products --
``__matmul__`` across different qubit sets (104-121), ``inflate`` (61-62),
``dela``/``trace`` with |0> (228-305) -- is deferred to a lazy
:class:`~.chain.GateChain` that the CUDA statevector engine evaluates in one run.
"""
import numpy as np

from .qontroller import core2control, ctrl2core
from .qspace import Qubit, qubit_key
from .su2gate import RGate
from .types import CONTROL0, CONTROL1, TARGET, IDLER, QType, UnivGate, proportional


def project_factor(core: np.ndarray, idx: int, state) -> np.ndarray:
    """``<state| core |state>`` on tensor factor ``idx`` (0 = most significant) of a ``2^t x 2^t`` matrix: the
    reference's ``qproject`` (``utils/inter_product.py:318-363``) as one tensor contraction."""
    t = core.shape[0].bit_length() - 1
    assert core.shape == (1 << t, 1 << t) and t >= 2 and 0 <= idx < t
    s = np.asarray(state, dtype=np.complex128)
    ten = np.asarray(core, dtype=np.complex128).reshape((2,) * (2 * t))
    ten = np.tensordot(np.conj(s), ten, axes=([0], [idx]))            # row factor gone: axes shift down by one
    ten = np.tensordot(ten, s, axes=([t - 1 + idx], [0]))
    return ten.reshape(1 << (t - 1), 1 << (t - 1))


def _is_rgate(g) -> bool:
    return isinstance(g, RGate) or (type(g).__name__ == 'RGate' and hasattr(g, 'angle') and hasattr(g, 'axis'))


class CtrlGate:
    def __init__(self, gate, control, qspace=None, phase: complex = 1.0):
        control = list(control)
        assert all(isinstance(c, QType) or type(c).__name__ == 'QType' for c in control), 'control contains non-QType items'
        mat = np.array(gate)
        self.gate = UnivGate.get_prop(mat)
        if self.gate is not None:
            # snap to the exact named gate and move the scalar into the phase (cgate.py:33-37)
            exact = np.array(self.gate)
            phase = phase * proportional(mat, exact)[1]
            mat = exact
        elif _is_rgate(gate):
            self.gate = gate
        n_target = sum(1 for c in control if int(c) in (TARGET, IDLER))
        assert mat.shape[0] == (1 << n_target), 'matrix shape does not match the control sequence'
        assert mat.ndim == 2 and mat.shape[0] == mat.shape[1]
        assert np.isclose(abs(phase), 1), 'phase factor must be normalized.'
        assert np.allclose(mat @ mat.conj().T, np.eye(mat.shape[0])), f'Matrix is not unitary {mat}'
        if qspace is None:
            qspace = [Qubit(i) for i in range(len(control))]
        else:
            qspace = list(qspace)
            assert all(hasattr(q, 'qid') and hasattr(q, 'ancilla') for q in qspace), 'qspace contains non-Qubit items'
            assert len(qspace) == len(set(qspace))
            assert len(qspace) == len(control)
        self.qspace = qspace
        self._qontrol = dict(zip(qspace, control))
        self._matrix = mat
        self._phase = phase

    def __setstate__(self, state):
        """Accept both this class's own pickles and the reference's (which keep the core matrix and the phase
        inside a ``_unitary`` UnitaryM, ``cgate.py:41``)."""
        if "_unitary" in state:
            u = state["_unitary"]
            self.gate = state.get("gate")
            self.qspace = list(state["qspace"])
            self._qontrol = dict(state["_qontrol"])
            self._matrix = np.asarray(u.matrix)
            self._phase = u.phase
        else:
            self.__dict__.update(state)

    # ---- structure -------------------------------------------------------------------
    def order(self) -> int:
        return 1 << len(self.qspace)

    def matrix(self) -> np.ndarray:
        return self._matrix

    def phase(self):
        return self._phase

    def core(self) -> tuple:
        return tuple(ctrl2core(self.controls()))

    def controls(self) -> list:
        return [self._qontrol[q] for q in self.qspace]

    def qontrol(self) -> dict:
        return self._qontrol

    def qids(self) -> list:
        return self.qspace

    def control_qids(self) -> list:
        return [q for q in self.qspace if int(self._qontrol[q]) in (CONTROL0, CONTROL1)]

    def target_qids(self) -> list:
        return [q for q in self.qspace if int(self._qontrol[q]) == TARGET]

    def is_std(self) -> bool:
        return isinstance(self.gate, UnivGate)

    def is_principal(self) -> bool:
        return _is_rgate(self.gate) and bool(self.gate.axis.principal)

    def is2l(self) -> bool:
        return self._matrix.shape[0] <= 2

    def issinglet(self) -> bool:
        return len(self.target_qids()) == 1

    def isid(self) -> bool:
        return bool(np.allclose(self._matrix, np.eye(self._matrix.shape[0])))

    def herm(self) -> 'CtrlGate':
        return CtrlGate(np.conj(self._matrix.T), self.controls(), self.qspace, self._phase)

    def __copy__(self):
        return CtrlGate(self._matrix, self.controls(), self.qspace, self._phase)

    def __deepcopy__(self, memo=None):
        return CtrlGate(self._matrix.copy(), list(self.controls()), list(self.qspace), self._phase)

    def __repr__(self):
        g = self.gate if self.gate else self._matrix
        return f'CtrlGate{{{g!r},controls={self.controls()!r},qspace={self.qspace!r}}}'

    # ---- small host-side transforms (index shuffles on the 2^t core only) ------------
    def sorted(self, sorting=None) -> 'CtrlGate':
        """Same operator with the qubits listed in ascending (or the given) order."""
        k = len(self.qspace)
        if sorting is None:
            sorting = sorted(range(k), key=lambda i: qubit_key(self.qspace[i]))
        else:
            sorting = [int(i) for i in sorting]
            assert sorted(sorting) == list(range(k))
        qspace = [self.qspace[i] for i in sorting]
        controls = [self._qontrol[q] for q in qspace]
        old_t = self.target_qids()
        new_t = [q for q, c in zip(qspace, controls) if int(c) == TARGET]
        t = len(new_t)
        src_pos = [old_t.index(q) for q in new_t]  # new target j was old target src_pos[j]
        idx = np.zeros(1 << t, dtype=np.int64)
        for v in range(1 << t):
            o = 0
            for j, p in enumerate(src_pos):
                if (v >> (t - 1 - j)) & 1:
                    o |= 1 << (t - 1 - p)
            idx[v] = o
        return CtrlGate(self._matrix[np.ix_(idx, idx)], controls, qspace, self._phase)

    def expand(self, qspace, ctrls=None) -> 'CtrlGate':
        """Append qubits (as extra targets the gate ignores, or as further controls)."""
        qspace = list(qspace)
        if not qspace:
            return self
        assert not set(qspace) & set(self.qspace), 'Extended qspace must not overlap with existing qspace.'
        if ctrls is None:
            ctrls = [QType.TARGET] * len(qspace)
        else:
            ctrls = list(ctrls)
            assert all(int(c) != IDLER for c in ctrls)
            assert len(qspace) == len(ctrls)
        extra = sum(1 for c in ctrls if int(c) == TARGET)
        mat = np.kron(self._matrix, np.eye(1 << extra)) if extra else self._matrix
        return CtrlGate(mat, self.controls() + ctrls, self.qspace + qspace, self._phase)

    def promote(self, qubits) -> 'CtrlGate':
        """Turn the listed control qubits into targets (block-diagonal embedding).

        The phase is absorbed into the matrix first because, as in
        ``unitary.py:76``, it only multiplies the controlled block.
        """
        qubits = set(qubits)
        if not qubits or all(int(self._qontrol[q]) == TARGET for q in qubits):
            return self
        assert qubits <= set(self.qspace), 'Promoting qubits must be a subset of existing qspace.'
        old = self.controls()
        new = [QType.TARGET if (q in qubits and int(c) in (CONTROL0, CONTROL1)) else c for q, c in zip(self.qspace, old)]
        # positions (among the new targets, MSB first) and activation of each promoted control
        new_t = [i for i, c in enumerate(new) if int(c) == TARGET]
        t = len(new_t)
        big = np.eye(1 << t, dtype=np.complex128)
        need_mask = need_val = 0
        keep_bits = []
        for j, i in enumerate(new_t):
            bit = t - 1 - j
            if int(old[i]) == TARGET:
                keep_bits.append(bit)
            else:
                need_mask |= 1 << bit
                if int(old[i]) == CONTROL1:
                    need_val |= 1 << bit
        rows = []
        for v in range(1 << len(keep_bits)):
            r = need_val
            for j, bit in enumerate(keep_bits):
                if (v >> (len(keep_bits) - 1 - j)) & 1:
                    r |= 1 << bit
            rows.append(r)
        big[np.ix_(rows, rows)] = self._matrix * self._phase
        return CtrlGate(big, new, self.qspace)

    @classmethod
    def convert(cls, u, qspace=None) -> 'CtrlGate':
        """``UnitaryM`` of power-of-two order -> ``CtrlGate`` (bits shared by all core
        indexes become controls, the rest targets; cgate.py:133-158)."""
        if qspace is None:
            assert u.order() & (u.order() - 1) == 0
            n = u.order().bit_length() - 1
            qspace = [Qubit(i) for i in range(n)]
        else:
            qspace = list(qspace)
            assert len(qspace) == len(set(qspace))
            n = len(qspace)
            assert u.order() == 1 << n
        controls = core2control(n, u.core)
        span = ctrl2core(controls)
        where = {idx: i for i, idx in enumerate(span)}
        sel = [where[c] for c in u.core]
        m = np.eye(len(span), dtype=np.complex128)
        m[np.ix_(sel, sel)] = u.matrix
        return cls(m, controls, qspace=qspace)

    # ---- dense semantics: evaluated on the GPU -----------------------------------------
    def __matmul__(self, other):
        from .chain import GateChain
        if isinstance(other, GateChain):
            return GateChain([self]) @ other
        if not hasattr(other, 'qspace') or not hasattr(other, 'controls'):
            return NotImplemented
        if self.qspace == list(other.qspace) and [int(c) for c in self.controls()] == [int(c) for c in other.controls()]:
            return CtrlGate(self._matrix @ other.matrix(), self.controls(), self.qspace, self._phase * other.phase())
        return GateChain([self, other])

    def inflate(self) -> np.ndarray:
        """Dense ``2^k x 2^k`` operator over ``self.qspace`` *in the order given*."""
        from .chain import GateChain
        return GateChain([self], register=self.qspace).inflate()

    def __array__(self, dtype=None, copy=None):
        m = self.inflate()
        return m if dtype is None else m.astype(dtype)

    def dela(self, state=None):
        from .chain import GateChain
        return GateChain([self]).dela(state)

    def trace(self, qubit, state) -> 'CtrlGate':
        """Fix one qubit of the gate in the pure state ``state`` and return the gate left on the others
        (``cgate.py:228-287``).  Work on the gate's own ``2^t x 2^t`` core only:

        * control qubit: the controlled block acts with weight ``|<active|state>|^2``, the identity with the rest;
        * target qubit: ``<state| core |state>`` on that tensor factor (``qproject``, ``inter_product.py:318-363``);
          a gate whose only target is traced leaves the identity, as in the reference.

        Like the reference, the result goes through the constructor, which rejects a non-unitary outcome.
        """
        assert qubit in self._qontrol, f'Qubit {qubit} not in qspace.'
        state = np.asarray(state)
        assert state.shape == (2,), f'state vector must be a 1D array of length 2, but got {state.shape}.'
        assert np.isclose(np.linalg.norm(state), 1), f'state vector must normalized but got {state}.'
        role = int(self._qontrol[qubit])
        if role == IDLER:
            raise NotImplementedError(f"CtrlGate {self._qontrol[qubit]} not implemented.")
        rest = [q for q in self.qspace if q != qubit]
        roles = [self._qontrol[q] for q in rest]
        if role in (CONTROL0, CONTROL1):
            w_idle, w_act = abs(state[0]) ** 2, abs(state[1]) ** 2
            if role == CONTROL0:
                w_idle, w_act = w_act, w_idle
            core = w_idle * np.eye(self._matrix.shape[0]) + w_act * self._matrix
            return CtrlGate(core, roles, rest, self._phase)
        targets = self.target_qids()
        if len(targets) == 1:
            return CtrlGate(UnivGate.I, [QType.TARGET], rest, self._phase)
        return CtrlGate(project_factor(self._matrix, targets.index(qubit), state), roles, rest, self._phase)

    def is_idler(self, qubit) -> bool:
        return qubit not in self._qontrol
