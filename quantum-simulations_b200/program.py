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


class RecordWriter:
    """Columnar writer of gate records over a fixed register; ``take()`` hands out the chunk written so far."""

    def __init__(self, register, batch_bits: int = 0, capacity: int = 1 << 16):
        self.register = list(register)
        self.n_qubits = len(self.register)
        self.batch_bits = int(batch_bits)
        n = self.n_qubits
        # qubits are keyed by the integer 2*qid + ancilla: hashing / comparing ints is C speed, Qubit objects are not
        self._bit = {2 * int(q.qid) + bool(q.ancilla): n - 1 - i + self.batch_bits for i, q in enumerate(self.register)}
        self._cache = {}
        self._cap = max(16, int(capacity))
        self._alloc()
        self.total = 0

    def _alloc(self):
        c = self._cap
        self._kind = np.zeros(c, dtype=np.uint32)
        self._ntar = np.zeros(c, dtype=np.uint32)
        self._tbit = np.zeros((c, 8), dtype=np.uint8)
        self._mask = np.zeros(c, dtype=np.uint64)
        self._val = np.zeros(c, dtype=np.uint64)
        self._row0 = np.zeros(c, dtype=np.uint64)
        self._row1 = np.zeros(c, dtype=np.uint64)
        self._off = np.zeros(c, dtype=np.uint64)
        self._pool = np.empty(4 * c, dtype=np.complex128)
        self._n = 0
        self._pn = 0

    def __len__(self):
        return self._n

    def _pattern(self, qspace, controls):
        key = (tuple([2 * q.qid + q.ancilla for q in qspace]), tuple([int(c) for c in controls]))
        hit = self._cache.get(key)
        if hit is None:
            mask = value = 0
            tbits = []
            for q, code in zip(key[0], key[1]):
                bit = self._bit[q]
                if code == CONTROL1:
                    mask |= 1 << bit
                    value |= 1 << bit
                elif code == CONTROL0:
                    mask |= 1 << bit
                else:  # TARGET or IDLER: both span (0, 1) in the matrix index
                    tbits.append(bit)
            if not tbits:
                raise ValueError("gate without target")
            if len(tbits) > MAX_TARGETS:
                raise NotImplementedError(f"gate with {len(tbits)} targets; the device path supports {MAX_TARGETS}")
            hit = (mask, value, tuple(tbits))
            if len(self._cache) < (1 << 20):
                self._cache[key] = hit
        return hit

    def _grow(self):
        old_n = self._cap
        self._cap *= 2
        for name in ("_kind", "_ntar", "_tbit", "_mask", "_val", "_row0", "_row1", "_off"):
            a = getattr(self, name)
            b = np.zeros((self._cap,) + a.shape[1:], dtype=a.dtype)
            b[:old_n] = a
            setattr(self, name, b)

    def _pool_room(self, need):
        if need > self._pool.size:
            grown = np.empty(max(need, 2 * self._pool.size), dtype=np.complex128)
            grown[:self._pn] = self._pool[:self._pn]
            self._pool = grown

    def add_ctrl_gate(self, g):
        mask, value, tbits = self._pattern(g.qspace, g.controls())
        m = g.matrix()
        ph = complex(g.phase())
        t = len(tbits)
        i = self._n
        if i == self._cap:
            self._grow()
        self._ntar[i] = t
        self._mask[i] = mask
        self._val[i] = value
        if t == 1:
            self._tbit[i, 0] = tbits[0]
            if ph == 1:
                a, b, c, d = m.flat
            else:  # (the array product, not four scalar ones: numpy's vector loop rounds differently in the last bit)
                a, b, c, d = (np.asarray(m, dtype=np.complex128) * ph).flat
            if a == 0 and d == 0 and b == 1 and c == 1:
                self._kind[i] = PERMX
                self._off[i] = 0
            else:
                self._kind[i] = DIAG1 if (b == 0 and c == 0) else DENSE1
                p = self._pn
                self._pool_room(p + 4)
                self._off[i] = p
                self._pool[p:p + 4] = (a, b, c, d)
                self._pn = p + 4
        else:
            self._tbit[i, :t] = tbits
            self._kind[i] = DENSEK
            size = 1 << (2 * t)
            self._pool_room(self._pn + size)
            self._off[i] = self._pn
            self._pool[self._pn:self._pn + size] = (np.asarray(m, dtype=np.complex128) * ph).reshape(-1)
            self._pn += size
        self._n = i + 1
        self.total += 1

    def add_two_level(self, u):
        if len(u.core) != 2:
            raise ValueError("only two-level UnitaryM gates stream as row records")
        i = self._n
        if i == self._cap:
            self._grow()
        self._pool_room(self._pn + 4)
        self._kind[i] = ROWS2
        self._ntar[i] = 0
        self._mask[i] = 0
        self._val[i] = 0
        self._row0[i], self._row1[i] = u.core
        self._off[i] = self._pn
        self._pool[self._pn:self._pn + 4] = (np.asarray(u.matrix, dtype=np.complex128) * complex(u.phase)).reshape(-1)
        self._pn += 4
        self._n = i + 1
        self.total += 1

    def add(self, g):
        if hasattr(g, "controls"):
            self.add_ctrl_gate(g)
        else:
            self.add_two_level(g)

    def add_raw(self, target_bits, ctrl_mask, ctrl_value, matrix):
        """Append a gate given directly by index bit positions (already including any batch shift)."""
        m = np.asarray(matrix, dtype=np.complex128)
        t = len(target_bits)
        assert m.shape == (1 << t, 1 << t) and 1 <= t <= MAX_TARGETS
        i = self._n
        if i == self._cap:
            self._grow()
        self._ntar[i] = t
        self._mask[i] = ctrl_mask
        self._val[i] = ctrl_value
        self._tbit[i, :t] = list(target_bits)
        kind = classify_1q(m) if t == 1 else DENSEK
        self._kind[i] = kind
        if kind == PERMX:
            self._off[i] = 0
        else:
            self._pool_room(self._pn + m.size)
            self._off[i] = self._pn
            self._pool[self._pn:self._pn + m.size] = m.reshape(-1)
            self._pn += m.size
        self._n = i + 1
        self.total += 1

    def snapshot(self, shift: int = 0):
        """(records, pool) written so far, WITHOUT resetting; bit positions of the bit-addressed records moved up by
        ``shift`` (the batch bits of a batch-innermost layout)."""
        n = self._n
        recs = np.zeros(n, dtype=GATE_DTYPE)
        recs["kind"] = self._kind[:n]
        recs["n_targets"] = self._ntar[:n]
        recs["target_bits"] = self._tbit[:n]
        recs["ctrl_mask"] = self._mask[:n]
        recs["ctrl_value"] = self._val[:n]
        recs["row0"] = self._row0[:n]
        recs["row1"] = self._row1[:n]
        recs["matrix_offset"] = self._off[:n]
        if shift and n:
            bitk = recs["kind"] != ROWS2
            recs["ctrl_mask"][bitk] <<= np.uint64(shift)
            recs["ctrl_value"][bitk] <<= np.uint64(shift)
            tb = recs["target_bits"]
            used = (np.arange(8)[None, :] < recs["n_targets"][:, None]) & bitk[:, None]
            tb[used] += np.uint8(shift)
        return recs, self._pool[:self._pn].copy()

    def take(self):
        """(records, pool) of the chunk written since the last ``take`` (copies; the writer is reset)."""
        recs, pool = self.snapshot()
        n = self._n
        self._tbit[:n] = 0
        self._row0[:n] = 0
        self._row1[:n] = 0
        self._n = 0
        self._pn = 0
        return recs, pool



class Program:
    """A gate stream in circuit TIME order (first record acts first) plus its matrix pool -- a register, and a
    ``RecordWriter`` over it."""

    def __init__(self, register):
        self.register = list(register)
        self.n_qubits = len(self.register)
        self._pos = {q: i for i, q in enumerate(self.register)}
        self._w = RecordWriter(self.register, 0, capacity=256)
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
        return len(self._w)

    def add_ctrl_gate(self, g):
        """Append one CtrlGate (reference semantics: cgate.py:24-52 + unitary.py:70-77)."""
        self._w.add_ctrl_gate(g)

    def add_raw(self, target_bits, ctrl_mask, ctrl_value, matrix):
        """Append a gate given directly by register bit positions (no CtrlGate object, hence none of
        its tolerance-based snapping to named gates, cgate.py:33-37)."""
        self._w.add_raw(target_bits, ctrl_mask, ctrl_value, matrix)

    def add_two_level(self, u):
        """Append a UnitaryM with a 2-index core (unitary.py:26-46) as a row gate."""
        assert len(u.core) == 2
        self._w.add_two_level(u)

    def records(self, batch_bits: int = 0) -> np.ndarray:
        """Records with bit positions shifted up by the batch bits (batch-innermost layout)."""
        return self._w.snapshot(batch_bits)[0]

    def pool(self) -> np.ndarray:
        return self._w.snapshot()[1]


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
