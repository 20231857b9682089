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

from .construct.qspace import qubit_key
from .program import RecordWriter  # noqa: F401  (the columnar writer lives with Program; re-exported here)


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
        from .engine import _try_fuse
        if self.fuse and _try_fuse(eng):                 # (a refused chunk runs one kernel per gate, with a warning)
            if self.spare is None:
                self.spare = eng.empty(self.state.numel())
            spare = self.spare
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
