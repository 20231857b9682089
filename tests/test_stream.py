"""Streaming flattener (quantum_simulations_b200/stream.py): same records as ``program.flatten`` bit for
bit, in bounded chunks; chunked application == one-shot application; reference-written ``.qco`` files stream without
the reference."""
import os
import random

import numpy as np
import pytest

import quantum_simulations_b200 as qs
from oracle import sv_oracle as orc
from quantum_simulations_b200.compat import read_qco
from quantum_simulations_b200.construct.bytecode import Bytecode
from quantum_simulations_b200.program import flatten, flatten_two_level
from quantum_simulations_b200.stream import RecordWriter, collect_register, flatten_stream, iter_leaves
from quantum_simulations_b200.utils.mgen import random_ctrlgate, random_state

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def _same_records(chunks, prog, batch_bits):
    want, wpool = prog.records(batch_bits), prog.pool()
    i = 0
    for recs, pool in chunks:
        assert recs.dtype == want.dtype
        for r in recs:
            w = want[i]
            for f in ("kind", "n_targets", "ctrl_mask", "ctrl_value", "row0", "row1"):
                assert r[f] == w[f], (i, f)
            assert np.array_equal(r["target_bits"], w["target_bits"]), i
            if int(r["kind"]) != 1:
                size = 4 if int(r["kind"]) == 4 else 1 << (2 * int(r["n_targets"]))
                a, b = int(r["matrix_offset"]), int(w["matrix_offset"])
                assert np.array_equal(pool[a:a + size], wpool[b:b + size]), i
            i += 1
    assert i == len(want)


def _random_gates(n, count, seed):
    random.seed(seed)
    np.random.seed(seed)
    gates = []
    for _ in range(count):
        k = random.randint(1, min(5, n))
        t = random.randint(1, min(k, 3))
        g = random_ctrlgate(k, t, qnum=n)
        r = random.random()
        if t == 1 and r < 0.3:
            g = qs.CtrlGate(random.choice(list(qs.UnivGate)), g.controls(), g.qspace)
        elif t == 1 and r < 0.5:
            g = qs.CtrlGate(np.diag(np.exp(1j * np.random.uniform(0, 6.28, 2))), g.controls(), g.qspace, phase=np.exp(0.3j))
        gates.append(g)
    return gates


@pytest.mark.parametrize("batch_bits,chunk", [(0, 7), (3, 64), (2, 1 << 18)])
def test_stream_records_equal_flatten(batch_bits, chunk):
    n = 7
    gates = _random_gates(n, 300, 11)
    reg = [qs.Qubit(i) for i in range(n)]
    chunks = list(flatten_stream(gates, reg, batch_bits, chunk))
    assert all(len(r) <= chunk for r, _ in chunks) and sum(len(r) for r, _ in chunks) == len(gates)
    _same_records(chunks, flatten(gates, reg), batch_bits)


def test_stream_of_a_tree_walks_time_order_and_honours_skip():
    gates = _random_gates(5, 12, 3)
    # matrix-product order tree: root -> [A, B, C], B -> [b0, b1]; time order is right to left
    leaf = [Bytecode(g) for g in gates]
    skipped = Bytecode(gates[5], skip=True)
    inner = Bytecode(gates[0], [leaf[1], leaf[2], skipped])
    inner_skip = Bytecode(gates[0], [leaf[6], leaf[7]], skip=True)   # a skipped inner node: its leaves still count
    root = Bytecode(gates[0], [leaf[3], inner, leaf[4], inner_skip])
    order = list(iter_leaves(root))
    want = [n.data for n in qs.construct.bytecode.BytecodeRevIter(root) if n.is_leaf() and not n.skip]
    assert [id(g) for g in order] == [id(g) for g in want]
    assert all(g is not gates[5] for g in order)
    reg = collect_register(root)
    assert set(reg) >= {q for g in order for q in g.qspace} and reg == sorted(reg, key=lambda q: (q.ancilla, q.qid))
    chunks = list(flatten_stream(root, chunk_gates=3))
    _same_records(chunks, flatten(order, reg), 0)


def test_chunked_application_equals_one_shot():
    n = 8
    gates = _random_gates(n, 200, 5)
    reg = [qs.Qubit(i) for i in range(n)]
    x = random_state(1 << n)
    want = orc.simulate(gates, reg, x.reshape(-1, 1))[:, 0]
    got = x.copy()
    n_chunks = 0
    for recs, pool in flatten_stream(gates, reg, 0, 37):
        orc.run_records(recs, pool, got, n)
        n_chunks += 1
    assert n_chunks == 6
    assert np.abs(got - want).max() < 1e-12


def test_two_level_stream():
    random.seed(2)
    np.random.seed(2)
    d = 6
    us = []
    for _ in range(20):
        i, j = sorted(random.sample(range(d), 2))
        q, _ = np.linalg.qr(np.random.randn(2, 2) + 1j * np.random.randn(2, 2))
        us.append(qs.UnitaryM(d, (i, j), q))
    w = RecordWriter([], 0, capacity=4)   # capacity smaller than the stream: exercises growth
    for u in us:
        w.add(u)
    _same_records([w.take()], flatten_two_level(us, d), 0)
    assert len(w) == 0 and w.total == 20


@pytest.mark.parametrize("name", ["qco_r4", "qco_cyc8", "qco_r8p"])
def test_reference_qco_files_stream(name):
    """A .qco written by the reference streams straight from the file: same records as rendering it gate by gate, and
    <0|U|0> on the ancillas reproduces the compiled input."""
    path = os.path.join(GOLDEN, name + ".qco")
    code = read_qco(path)
    circ = qs.QRenderer(qs.B200Builder()).render(code)
    prog = circ.program()
    n, am = prog.n_qubits, prog.ancilla_mask
    keep = [i for i in range(1 << n) if not (i & am)]
    k = len(keep).bit_length() - 1
    chunks = list(flatten_stream(path, batch_bits=k, chunk_gates=100))
    _same_records(chunks, prog, k)
    st = np.zeros((1 << n, 1 << k), dtype=np.complex128)
    st[keep, np.arange(len(keep))] = 1
    flat = st.reshape(-1)
    for recs, pool in chunks:
        orc.run_records(recs, pool, flat, n + k)
    got = flat.reshape(1 << n, 1 << k)[keep, :]
    d = np.load(os.path.join(GOLDEN, "qco_inputs.npz"))
    assert np.abs(got - d[name + "_input"]).max() < 1e-10


def test_writer_rejects_what_the_device_cannot_run():
    reg = [qs.Qubit(i) for i in range(7)]
    w = RecordWriter(reg)
    with pytest.raises(NotImplementedError):
        w.add(random_ctrlgate(6, 6, qnum=7))
    with pytest.raises(KeyError):
        w.add(qs.CtrlGate(qs.UnivGate.H, [qs.QType.TARGET], [qs.Qubit(9)]))


@pytest.mark.gpu
def test_apply_stream_on_the_gpu_equals_apply_circuit():
    from quantum_simulations_b200.engine import apply_circuit
    from quantum_simulations_b200.stream import apply_stream
    n = 10
    gates = [g for g in _random_gates(n, 400, 9)]
    reg = [qs.Qubit(i) for i in range(n)]
    x = np.stack([random_state(1 << n) for _ in range(3)], axis=1)
    want = apply_circuit(flatten(gates, reg), x)
    for chunk in (64, 1 << 18):
        got = apply_stream(gates, x, reg, chunk_gates=chunk)
        assert np.abs(got - want).max() < 1e-10
    ref = orc.simulate(gates, reg, x)
    assert np.abs(want - ref).max() < 1e-10


class _HostEngine:
    """Stand-in for ``engine.Engine`` with the oracle behind it (TEST ONLY): lets the chunk executor's bookkeeping --
    buffer ping-pong, spare allocation, the fall-back when the fusion pass refuses a chunk -- run on a CPU box."""

    def __init__(self, refuse_fuse_every=0):
        import torch
        self.torch = torch
        self.np_dtype = np.dtype(np.complex128)
        self.device = 0
        self.calls = []
        self.refuse = refuse_fuse_every
        self.n_fuse = 0

    def from_host(self, array):
        return self.torch.from_numpy(np.ascontiguousarray(array, dtype=self.np_dtype).reshape(-1).copy())

    def empty(self, n):
        return self.torch.empty(int(n), dtype=self.torch.complex128)

    def synchronize(self):
        pass

    def upload_records(self, recs, pool, n_bits, batch_bits=0):
        self.prog = (recs.copy(), pool.copy(), n_bits)
        self.fused = False

    def fuse(self, max_tile_bits=0):
        from quantum_simulations_b200._lib import QsvError
        self.n_fuse += 1
        if self.refuse and self.n_fuse % self.refuse == 0:
            raise QsvError("unsupported: refused by the test")
        self.fused = True
        return 1

    def run(self, state, spare=None):
        recs, pool, n_bits = self.prog
        # like the device: a fused plan may leave its result in the spare buffer
        out = spare if (self.fused and spare is not None and len(self.calls) % 2 == 0) else state
        if out is not state:
            out.copy_(state)
        orc.run_records(recs, pool, out.numpy(), n_bits)
        self.calls.append(self.fused)
        return out


@pytest.mark.parametrize("refuse", [0, 2])
def test_apply_stream_bookkeeping_with_a_host_engine(refuse):
    from quantum_simulations_b200.stream import apply_stream
    n = 7
    gates = [g for g in _random_gates(n, 150, 21) if len([c for c in g.controls() if int(c) in (1, 8)]) == 1]
    reg = [qs.Qubit(i) for i in range(n)]
    x = np.stack([random_state(1 << n) for _ in range(3)], axis=1)
    eng = _HostEngine(refuse)
    got = apply_stream(gates, x, reg, chunk_gates=16, engine=eng)
    want = orc.simulate(gates, reg, x)
    assert got.shape == x.shape and np.abs(got - want).max() < 1e-12
    assert len(eng.calls) == -(-len(gates) // 16) and (all(eng.calls) if not refuse else not all(eng.calls))
    one = apply_stream(gates, x[:, 0], reg, chunk_gates=1 << 18, engine=_HostEngine())
    assert one.shape == (1 << n,) and np.abs(one - want[:, 0]).max() < 1e-12


def _naive_records(gates, reg, batch_bits):
    """Per-gate restatement of the record conventions (program.py header), independent of RecordWriter."""
    from quantum_simulations_b200._lib import GATE_DTYPE
    n = len(reg)
    pos = {q: i for i, q in enumerate(reg)}
    recs, pool = [], []
    for g in gates:
        r = np.zeros((), dtype=GATE_DTYPE)
        mask = value = 0
        tb = []
        for q, c in zip(g.qspace, g.controls()):
            bit = n - 1 - pos[q] + batch_bits
            if int(c) == 4:
                mask |= 1 << bit
                value |= 1 << bit
            elif int(c) == 2:
                mask |= 1 << bit
            else:
                tb.append(bit)
        m = np.asarray(g.matrix(), dtype=np.complex128) * complex(g.phase())
        r["n_targets"], r["ctrl_mask"], r["ctrl_value"] = len(tb), mask, value
        r["target_bits"][:len(tb)] = tb
        if len(tb) == 1 and np.array_equal(m, [[0, 1], [1, 0]]):
            r["kind"] = 1
        else:
            r["kind"] = 3 if len(tb) > 1 else (2 if (m[0, 1] == 0 and m[1, 0] == 0) else 0)
            r["matrix_offset"] = sum(p.size for p in pool)
            pool.append(m.reshape(-1))
        recs.append(r)
    return np.array(recs, dtype=GATE_DTYPE), (np.concatenate(pool) if pool else np.zeros(0, complex))


@pytest.mark.parametrize("batch_bits", [0, 4])
def test_program_records_against_a_naive_flattener(batch_bits):
    n = 7
    gates = _random_gates(n, 250, 33)
    reg = [qs.Qubit(i) for i in range(n)]
    want, wpool = _naive_records(gates, reg, batch_bits)
    prog = flatten(gates, reg)
    assert prog.records(batch_bits).tobytes() == want.tobytes()
    assert np.array_equal(prog.pool(), wpool)
    assert prog.records(batch_bits).tobytes() == want.tobytes()      # records() does not consume the program
