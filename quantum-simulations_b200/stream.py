"""Streaming flattener: Bytecode trees / ``.qco`` files -> ``qsv_gate_t`` record chunks.

The reference replays a compiled program leaf by leaf (``quompiler/circuits/render.py:21-32``,
``optimize/code_analyze.py:49-89``); ``program.flatten`` mirrors that with one numpy record object per
gate (~12 us/gate -- a minute for the 4.7 M-gate stream of BASELINE config 4).  This module does the
same conversion in COLUMNS: records are written into preallocated structured arrays, the
control/target decoding of a gate is cached per distinct ``(qspace, controls)`` pattern and the 2x2
classification works on scalars, and the stream is cut into bounded chunks that an executor consumes
while the next chunk is being flattened -- the GPU path uploads, fuses and launches a chunk
asynchronously, so host flattening overlaps device execution and memory stays bounded for programs
of any length.

Conventions are exactly those of ``program.Program`` (same records, bit for bit -- tests compare them).
Pickle compatibility: trees written by the reference load through ``compat.read_qco``.
"""
import numpy as np

from ._lib import DENSE1, DENSEK, DIAG1, GATE_DTYPE, MAX_TARGETS, PERMX, ROWS2
from .construct.qspace import qubit_key
from .construct.types import CONTROL0, CONTROL1


def iter_leaves(code):
    """Leaves of a Bytecode tree in circuit TIME order (``BytecodeRevIter``), skipping flagged subtrees' nodes
    the way ``QRenderer.render`` does (a node with ``skip`` is ignored, its children are still visited)."""
    stack = [code]
    while stack:
        node = stack.pop()
        children = node.children
        if children:
            stack.extend(children)
        elif not node.skip:
            yield node.data


def collect_register(code):
    """Sorted qubit set of a tree (``render.collect_qspace``), found without building any per-leaf object."""
    seen = set()
    stack = [code]
    while stack:
        node = stack.pop()
        qs = getattr(node.data, "qspace", None)
        if qs is not None:
            seen.update(qs)
        stack.extend(node.children)
    return sorted(seen, key=qubit_key)


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

    def take(self):
        """(records, pool) of the chunk written since the last ``take`` (copies; the writer is reset)."""
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
        pool = self._pool[:self._pn].copy()
        self._tbit[:n] = 0
        self._row0[:n] = 0
        self._row1[:n] = 0
        self._n = 0
        self._pn = 0
        return recs, pool


def flatten_stream(source, register=None, batch_bits: int = 0, chunk_gates: int = 1 << 18):
    """Yield ``(records, pool)`` chunks, in circuit time order, for a Bytecode tree, a ``.qco`` path or any iterable
    of gates given in time order.  ``register`` defaults to the sorted qubit set of the tree (a plain iterable of
    gates needs it given, or is materialised once to find it)."""
    if isinstance(source, (str, bytes)) or hasattr(source, "__fspath__"):
        from .compat import read_qco
        source = read_qco(source)
    if hasattr(source, "children"):
        if register is None:
            register = collect_register(source)
        leaves = iter_leaves(source)
    else:
        if register is None:
            source = list(source)
            qs = set()
            for g in source:
                qs.update(getattr(g, "qspace", ()))
            register = sorted(qs, key=qubit_key)
        leaves = iter(source)
    w = RecordWriter(register, batch_bits, capacity=min(chunk_gates, 1 << 16))
    for g in leaves:
        w.add(g)
        if len(w) >= chunk_gates:
            yield w.take()
    if len(w):
        yield w.take()


class GpuChunkExecutor:
    """Applies record chunks to a device-resident state: upload -> fuse -> run per chunk, all asynchronous on the
    engine's stream, so the host flattens chunk i+1 while the device executes chunk i."""

    def __init__(self, engine, state, n_bits, batch_bits=0, fuse=True):
        self.eng, self.state, self.n_bits, self.batch_bits, self.fuse = engine, state, n_bits, batch_bits, fuse
        self.spare = None

    def apply(self, recs, pool):
        from ._lib import QsvError
        eng = self.eng
        eng.upload_records(recs, pool, self.n_bits, self.batch_bits)
        spare = None
        if self.fuse:
            try:
                eng.fuse()
                if self.spare is None:
                    self.spare = eng.empty(self.state.numel())
                spare = self.spare
            except QsvError as e:  # the fusion pass refuses this chunk -> one kernel per gate (still CUDA)
                if "unsupported" not in str(e).lower():
                    raise
        res = eng.run(self.state, spare)
        if res is not self.state:
            self.state, self.spare = res, self.state


def apply_stream(source, states, register=None, dtype=np.complex128, chunk_gates: int = 1 << 18, fuse=True, engine=None):
    """``engine.apply_circuit`` for programs too long to flatten in one piece: same result, bounded host memory."""
    from .engine import _pow2_ceil_bits, get_engine
    eng = engine or get_engine(dtype)
    states = np.asarray(states)
    single = states.ndim == 1
    x = states.reshape(states.shape[0], -1)
    dim, batch = x.shape
    b = _pow2_ceil_bits(batch)
    chunks = flatten_stream(source, register, b, chunk_gates)
    first = next(chunks, None)
    n = (dim - 1).bit_length()
    if dim != 1 << n:
        raise ValueError(f"state has {dim} rows, not a power of two")
    padded = np.zeros((dim, 1 << b), dtype=eng.np_dtype)
    padded[:, :batch] = x
    ex = GpuChunkExecutor(eng, eng.from_host(padded), n + b, b, fuse)
    while first is not None:
        ex.apply(*first)
        first = next(chunks, None)
    eng.synchronize()
    out = ex.state.view(dim, 1 << b).cpu().numpy()[:, :batch]
    return out[:, 0].copy() if single else np.ascontiguousarray(out)
