"""Flattener: quompiler gate objects -> POD ``qsv_gate_t`` records over the sorted register.

Conventions restated from the reference (verified against it by the golden
fixtures in ``tests/golden``):
 * register = sorted qubits, ordinary before ancilla, then by id
   (``quompiler/construct/qspace.py:31-36``, ``optimize/code_analyze.py:39-46``);
 * the qubit at sorted position ``p`` of ``N`` owns index bit ``N-1-p``
   (big-endian, ``construct/qontroller.py:7-15``);
 * matrix row/column bits run over the TARGET qubits in the gate's own ``qspace``
   order, first = most significant (``qontroller.py:112-120``);
 * the phase multiplies only the controlled block (``construct/unitary.py:76``).
Works on this package's gate classes and, by duck typing, on the reference's own
``CtrlGate`` / ``UnitaryM`` instances.
"""
import numpy as np

from ._lib import DENSE1, DENSEK, DIAG1, GATE_DTYPE, MAX_TARGETS, PERMX, ROWS2
from .construct.qspace import qubit_key
from .construct.types import CONTROL0, CONTROL1

_X = np.array([[0, 1], [1, 0]], dtype=np.complex128)


def sorted_register(qubits):
    return sorted(set(qubits), key=qubit_key)


def classify_1q(m: np.ndarray) -> int:
    """Kind of a 2x2 block: exact X -> swap (bit-exact), exact zero off-diagonal -> diagonal."""
    if np.array_equal(m, _X):
        return PERMX
    if m[0, 1] == 0 and m[1, 0] == 0:
        return DIAG1
    return DENSE1


class Program:
    """A gate stream in circuit TIME order (first record acts first) plus its matrix pool."""

    def __init__(self, register):
        self.register = list(register)
        self.n_qubits = len(self.register)
        self._pos = {q: i for i, q in enumerate(self.register)}
        self._recs = []
        self._mats = []
        self._pool_len = 0
        self.dimension = None  # set for UnitaryM (row) programs

    def bit_of(self, q) -> int:
        return self.n_qubits - 1 - self._pos[q]

    @property
    def ancilla_mask(self) -> int:
        m = 0
        for q in self.register:
            if q.ancilla:
                m |= 1 << self.bit_of(q)
        return m

    def __len__(self):
        return len(self._recs)

    def _push_matrix(self, m) -> int:
        m = np.ascontiguousarray(m, dtype=np.complex128)
        off = self._pool_len
        self._mats.append(m.reshape(-1))
        self._pool_len += m.size
        return off

    def add_ctrl_gate(self, g):
        """Append one CtrlGate (reference semantics: cgate.py:24-52 + unitary.py:70-77)."""
        mask = value = 0
        tbits = []
        for q, c in zip(g.qspace, g.controls()):
            bit = self.bit_of(q)
            code = int(c)
            if code == CONTROL1:
                mask |= 1 << bit
                value |= 1 << bit
            elif code == CONTROL0:
                mask |= 1 << bit
            else:  # TARGET or IDLER: both span (0, 1) in the matrix index
                tbits.append(bit)
        m = np.asarray(g.matrix(), dtype=np.complex128) * complex(g.phase())
        t = len(tbits)
        if t > MAX_TARGETS:
            raise NotImplementedError(f"gate with {t} targets; the device path supports {MAX_TARGETS}")
        if t == 0:
            raise ValueError("gate without target")
        rec = np.zeros((), dtype=GATE_DTYPE)
        rec["n_targets"] = t
        rec["target_bits"][:t] = tbits
        rec["ctrl_mask"] = mask
        rec["ctrl_value"] = value
        if t == 1:
            kind = classify_1q(m)
            rec["kind"] = kind
            rec["matrix_offset"] = 0 if kind == PERMX else self._push_matrix(m)
        else:
            rec["kind"] = DENSEK
            rec["matrix_offset"] = self._push_matrix(m)
        self._recs.append(rec)

    def add_raw(self, target_bits, ctrl_mask, ctrl_value, matrix):
        """Append a gate given directly by register bit positions (no CtrlGate object, hence none of
        its tolerance-based snapping to named gates, cgate.py:33-37)."""
        m = np.asarray(matrix, dtype=np.complex128)
        t = len(target_bits)
        assert m.shape == (1 << t, 1 << t) and 1 <= t <= MAX_TARGETS
        rec = np.zeros((), dtype=GATE_DTYPE)
        rec["n_targets"] = t
        rec["target_bits"][:t] = list(target_bits)
        rec["ctrl_mask"] = ctrl_mask
        rec["ctrl_value"] = ctrl_value
        kind = classify_1q(m) if t == 1 else DENSEK
        rec["kind"] = kind
        rec["matrix_offset"] = 0 if kind == PERMX else self._push_matrix(m)
        self._recs.append(rec)

    def add_two_level(self, u):
        """Append a UnitaryM with a 2-index core (unitary.py:26-46) as a row gate."""
        assert len(u.core) == 2
        rec = np.zeros((), dtype=GATE_DTYPE)
        rec["kind"] = ROWS2
        rec["n_targets"] = 0
        rec["row0"], rec["row1"] = u.core
        rec["matrix_offset"] = self._push_matrix(np.asarray(u.matrix, dtype=np.complex128) * complex(u.phase))
        self._recs.append(rec)

    def records(self, batch_bits: int = 0) -> np.ndarray:
        """Records with bit positions shifted up by the batch bits (batch-innermost layout)."""
        if not self._recs:
            return np.zeros(0, dtype=GATE_DTYPE)
        recs = np.array(self._recs, dtype=GATE_DTYPE)
        if batch_bits:
            bitk = recs["kind"] != ROWS2
            recs["ctrl_mask"][bitk] <<= np.uint64(batch_bits)
            recs["ctrl_value"][bitk] <<= np.uint64(batch_bits)
            for i in np.nonzero(bitk)[0]:
                t = int(recs["n_targets"][i])
                recs["target_bits"][i][:t] += batch_bits
        return recs

    def pool(self) -> np.ndarray:
        if not self._mats:
            return np.zeros(0, dtype=np.complex128)
        return np.concatenate(self._mats)


def flatten(gates_time_order, register=None) -> Program:
    """Flatten CtrlGates given in circuit time order (first acts first)."""
    gates = list(gates_time_order)
    if register is None:
        qs = set()
        for g in gates:
            qs.update(g.qspace)
        register = sorted_register(qs)
    prog = Program(register)
    for g in gates:
        prog.add_ctrl_gate(g)
    return prog


def flatten_two_level(unitaries_time_order, dimension: int) -> Program:
    """Flatten two-level UnitaryMs (time order) into QSV_ROWS2 records of a [dimension][batch] array."""
    prog = Program([])
    prog.dimension = int(dimension)
    for u in unitaries_time_order:
        if u.dimension != dimension:
            raise ValueError('matmul: Input operands have dimension mismatch.')
        prog.add_two_level(u)
    return prog
