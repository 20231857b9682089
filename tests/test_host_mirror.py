"""Host logic (no GPU): the mirror gate types and the flattener, against reference outputs."""
import ctypes
import os
import re

import numpy as np
import pytest

import quantum_simulations_b200 as qs
from oracle import sv_oracle as orc
from quantum_simulations_b200 import _lib
from quantum_simulations_b200.construct.qontroller import core2control, ctrl2core
from quantum_simulations_b200.program import flatten
from tests.golden_io import STREAMS, load, unpack_gates

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def same_gate(a, b, tol=1e-12):
    return (a.qspace == b.qspace and [int(c) for c in a.controls()] == [int(c) for c in b.controls()]
            and a.matrix().shape == b.matrix().shape
            and np.abs(a.matrix() * a.phase() - b.matrix() * b.phase()).max() < tol)


def test_qubit_order_and_repr():
    qs_ = [qs.Qubit(3, True), qs.Qubit(5), qs.Qubit(1), qs.Qubit(0, True)]
    assert [repr(q) for q in sorted(qs_)] == ["q1", "q5", "a0", "a3"]
    assert qs.Qubit(2) == qs.Qubit(2) and qs.Qubit(2) != qs.Qubit(2, True)
    assert len({qs.Qubit(2), qs.Qubit(2), qs.Qubit(2, True)}) == 2
    with pytest.raises(AssertionError):
        qs.Qubit(-1)


def test_qtype_flags_and_univgate():
    assert qs.QType.CONTROL1 in (qs.QType.CONTROL0 | qs.QType.CONTROL1)
    assert qs.QType.TARGET.base == (0, 1) and qs.QType.CONTROL0.base == (0,)
    assert sorted([qs.QType.CONTROL1, qs.QType.TARGET, qs.QType.CONTROL0])[0] == qs.QType.TARGET
    for g in qs.UnivGate:
        m = np.array(g)
        assert np.allclose(m @ m.conj().T, np.eye(2))
        assert qs.UnivGate.get(m) == g
        assert np.allclose(np.array(g.herm()), m.conj().T)
    assert qs.UnivGate.get_prop(1j * np.array(qs.UnivGate.H)) == qs.UnivGate.H


def test_ctrlgate_snaps_named_gates_and_keeps_phase():
    g = qs.CtrlGate(np.exp(0.7j) * np.array(qs.UnivGate.S), [qs.QType.CONTROL1, qs.QType.TARGET])
    assert g.is_std() and g.gate == qs.UnivGate.S
    assert np.isclose(g.phase(), np.exp(0.7j))
    assert np.array_equal(g.matrix(), np.array(qs.UnivGate.S))


def test_ctrlgate_validation_errors():
    with pytest.raises(AssertionError):
        qs.CtrlGate(np.eye(2), [qs.QType.TARGET, qs.QType.TARGET])          # shape mismatch
    with pytest.raises(AssertionError):
        qs.CtrlGate(np.array([[1, 1], [0, 1]]), [qs.QType.TARGET])          # not unitary
    with pytest.raises(AssertionError):
        qs.CtrlGate(np.eye(2), [qs.QType.TARGET], [qs.Qubit(0), qs.Qubit(1)])  # qspace length
    with pytest.raises(AssertionError):
        qs.UnitaryM(3, (0, 0), np.eye(2))                                   # duplicate core
    with pytest.raises(AssertionError):
        qs.UnitaryM(1, (0, 1), np.eye(2))                                   # dimension too small


def test_index_math_matches_reference():
    d = load("index_math")
    role = {1: qs.QType.TARGET, 2: qs.QType.CONTROL0, 4: qs.QType.CONTROL1}
    for i in range(len(d["n"])):
        n = int(d["n"][i])
        ctrl = [role[int(c)] for c in d["ctrl"][i][:n]]
        core = d["core"][d["core_off"][i]:d["core_off"][i + 1]].tolist()
        assert ctrl2core(ctrl) == core
        if len(core) > 1:
            assert [int(c) for c in core2control(n, core)] == [int(c) for c in d["back"][i][:n]]


def test_sorted_expand_promote_match_reference():
    d = load("transforms")
    src = unpack_gates(d, "src")
    for g, s, p, e in zip(src, unpack_gates(d, "sorted"), unpack_gates(d, "promote"), unpack_gates(d, "expand")):
        assert same_gate(g.sorted(), s)
        ctrls = g.control_qids()
        assert same_gate(g.promote(ctrls[:1]), p)
        assert same_gate(g.expand([qs.Qubit(10)], [qs.QType.CONTROL1]), e)
        # all three leave the operator unchanged (checked with the oracle's dense inflate)
        reg = orc.register_of([g])
        assert np.abs(orc.lift(g.sorted(), reg) - orc.lift(g, reg)).max() < 1e-12
        assert np.abs(orc.lift(g.promote(ctrls[:1]), reg) - orc.lift(g, reg)).max() < 1e-12


def test_convert_two_level_to_ctrlgate():
    d = load("two_level")
    for c, m, p in zip(d["d8_core"], d["d8_mat"], d["d8_phase"]):
        u = qs.UnitaryM(8, tuple(c), m, p)
        g = qs.CtrlGate.convert(u)
        assert np.abs(orc.inflate_gate(g) - orc.inflate_unitarym(8, u.core, u.matrix, u.phase)).max() < 1e-12


def test_matmul_is_lazy_and_same_space_product_is_small():
    a = qs.CtrlGate(qs.UnivGate.H, [qs.QType.TARGET])
    b = qs.CtrlGate(qs.UnivGate.T, [qs.QType.TARGET])
    ab = a @ b
    assert isinstance(ab, qs.CtrlGate) and np.allclose(ab.matrix() * ab.phase(), np.array(qs.UnivGate.H) @ np.array(qs.UnivGate.T))
    c = qs.CtrlGate(qs.UnivGate.X, [qs.QType.CONTROL1, qs.QType.TARGET])
    chain = a @ c @ b
    assert isinstance(chain, qs.GateChain) and len(chain.gates) == 3
    assert chain.qspace == [qs.Qubit(0), qs.Qubit(1)] and chain.order() == 4
    assert chain.dela().qspace == [qs.Qubit(0), qs.Qubit(1)]


@pytest.mark.parametrize("name", STREAMS)
def test_flattened_records_reproduce_reference(name):
    """flatten() -> qsv_gate_t records, executed by the CPU oracle's record runner, equals the
    reference value: pins bit positions, target order, control masks and phase folding."""
    d = load("streams")
    leaves = unpack_gates(d, name)
    prog = flatten(leaves[::-1])                     # time order
    n = prog.n_qubits
    am = prog.ancilla_mask
    keep = [i for i in range(1 << n) if not (i & am)]
    b = len(keep).bit_length() - 1
    st = np.zeros((1 << n, 1 << b), dtype=np.complex128)
    st[keep, np.arange(len(keep))] = 1
    flat = st.reshape(-1)
    orc.run_records(prog.records(b), prog.pool(), flat, n + b)
    got = flat.reshape(1 << n, 1 << b)[keep, :]
    assert np.abs(got - d[f"{name}_value"]).max() < 1e-11
    kinds = set(prog.records()["kind"].tolist())
    if name in ("cfg2s", "cyc16"):
        assert kinds == {_lib.PERMX}                 # pure swaps -> bit-exact path
        assert np.array_equal(got, d[f"{name}_value"])


def test_builder_protocol():
    b = qs.get_builder(qs.QompilePlatform.B200)
    assert isinstance(b, qs.CircuitBuilder)
    g = qs.CtrlGate(qs.UnivGate.S, (qs.QType.TARGET, qs.QType.CONTROL0, qs.QType.CONTROL1))
    b.register(g.qspace)
    b.register(g.qspace[:2])                          # repeated / overlapping registration is fine
    b.build_gate(g)
    b.build_group(g)
    c1 = b.finish()
    b.build_gate(g)
    c2 = b.finish()                                   # finish may be called repeatedly
    assert len(c1) == 1 and len(c2) == 2
    assert b.all_qubits() == sorted(g.qspace)
    with pytest.raises(KeyError):
        b.build_gate(qs.CtrlGate(qs.UnivGate.X, [qs.QType.TARGET], [qs.Qubit(9)]))
    with pytest.raises(ValueError):
        qs.get_builder(qs.QompilePlatform.CIRQ)


def test_renderer_walks_time_order():
    calls = []

    class Rec(qs.CircuitBuilder):
        def register(self, qspace): calls.append(("reg", list(qspace)))
        def build_gate(self, m): calls.append(("gate", m))
        def build_group(self, m): calls.append(("group", m))
        def finish(self, optimized=False): return "done"
        def all_qubits(self): return []

    g = [qs.CtrlGate(u, [qs.QType.TARGET], [qs.Qubit(i)]) for i, u in enumerate([qs.UnivGate.H, qs.UnivGate.T, qs.UnivGate.X])]
    tree = qs.Bytecode(g[0], [qs.Bytecode(g[0]), qs.Bytecode(g[1], skip=True), qs.Bytecode(g[2])])
    assert qs.QRenderer(Rec()).render(tree) == "done"
    assert calls[0] == ("reg", [qs.Qubit(0), qs.Qubit(1), qs.Qubit(2)])
    assert [c[0] for c in calls[1:]] == ["group", "gate", "gate"]
    assert calls[2][1] is g[2] and calls[3][1] is g[0]   # right-most leaf first = circuit time order
    assert [n.data for n in qs.BytecodeIter(tree) if n.is_leaf()] == [g[0], g[1], g[2]]


def test_library_exports_every_declared_symbol():
    """The C-ABI library loads on a CPU-only box and exports exactly what include/qsv.h declares."""
    header = open(os.path.join(ROOT, "include", "qsv.h")).read()
    declared = set(re.findall(r"\b(qsv_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations parsed"
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in qsv.h but not exported"
    assert declared == set(_lib.exported_symbols()), "ctypes binding and header out of sync"
    # the test-only host check of the fusion pass has its own header and is bound under tests/, not in the package
    hc = open(os.path.join(ROOT, "include", "qsv_hostcheck.h")).read()
    hc_declared = set(re.findall(r"\b(qsv_[a-z0-9_]+)\s*\(", hc))
    assert hc_declared == {"qsv_hostcheck_fused", "qsv_hostcheck_fused_layout"}
    assert not hasattr(lib, "qsv_hostcheck_fused"), "the host interpreter must not be part of the product library"
    from tests.hostcheck import HOSTCHECK_PATH
    assert hasattr(ctypes.CDLL(HOSTCHECK_PATH), "qsv_hostcheck_fused")
    assert "qsv_hostcheck_fused" not in _lib.exported_symbols()
    assert b"sm_100a" in ctypes.cast(lib.qsv_version, ctypes.CFUNCTYPE(ctypes.c_char_p))()


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    g = qs.CtrlGate(qs.UnivGate.H, [qs.QType.TARGET])
    with pytest.raises(_lib.QsvError):
        g.inflate()
    with pytest.raises(_lib.QsvError):
        (g @ qs.CtrlGate(qs.UnivGate.X, [qs.QType.CONTROL1, qs.QType.TARGET])).inflate()


@pytest.mark.skipif(not os.path.isdir("/root/reference/quompiler"), reason="the reference tree exists only in the build container")
def test_builder_accepts_the_reference_own_objects():
    """Drop-in check: the REAL quompiler CtrlGate/Qubit/QType objects go through B200Builder and the
    flattener unchanged (duck typing), and give the reference's own value."""
    import subprocess
    import sys
    code = r'''
import sys, random
import scipy, sklearn
import numpy as np
np.complexfloating = np.complex128
sys.path.insert(0, "/root/reference"); sys.path.insert(0, %r)
from functools import reduce
from quompiler.circuits.qdevice import QDevice
from quompiler.circuits.qompiler import Qompiler
from quompiler.config.config_manager import create_config
from quompiler.construct.bytecode import BytecodeRevIter
from quompiler.utils.mgen import random_unitary
import quantum_simulations_b200 as qs
from oracle import sv_oracle as orc
random.seed(42); np.random.seed(42)
u = random_unitary(4)
cfg = create_config(emit="CTRL_PRUNED", ancilla_offset=100)
bc = Qompiler(cfg, QDevice(cfg.device)).decompose(u)
b = qs.B200Builder()
qs.QRenderer(b).render(bc)                      # our renderer walking the reference's Bytecode tree
circ = b.finish()
prog = circ.program()
n = prog.n_qubits; am = prog.ancilla_mask
keep = [i for i in range(1 << n) if not (i & am)]
k = len(keep).bit_length() - 1
st = np.zeros((1 << n, 1 << k), dtype=np.complex128); st[keep, np.arange(len(keep))] = 1
flat = st.reshape(-1)
orc.run_records(prog.records(k), prog.pool(), flat, n + k)
got = flat.reshape(1 << n, 1 << k)[keep, :]
assert np.abs(got - u).max() < 1e-10, np.abs(got - u).max()
print("ok", len(circ), [repr(q) for q in circ.all_qubits()])
''' % ROOT
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "ok" in out.stdout, out.stderr[-2000:]


def test_fast_two_level_producer_matches_reference():
    """O(N^3) Givens producer == the reference's mat2l_decompose output (tests/golden/two_level.npz)."""
    from quantum_simulations_b200.utils.producers import two_level_stream
    d = load("two_level")
    for name in ("d8", "d3", "d6"):
        stream = two_level_stream(d[f"{name}_input"])
        assert [(i, j) for i, j, _ in stream] == [tuple(c) for c in d[f"{name}_core"].tolist()]
        for (_, _, c), ref in zip(stream, d[f"{name}_mat"]):
            assert np.abs(c - ref).max() < 1e-12


def test_fast_singlet_producer_reproduces_the_unitary():
    """SINGLET records written directly (Gray chains) rebuild the input unitary through the oracle."""
    from quantum_simulations_b200.utils.producers import singlet_records, two_level_records, two_level_stream
    d = load("streams")
    u = d["r32_input"]
    stream = two_level_stream(u)
    n = 5
    recs, pool = singlet_records(stream, n, batch_bits=n)
    assert set(recs["kind"].tolist()) <= {_lib.PERMX, _lib.DENSE1, _lib.DIAG1}
    full = ((1 << n) - 1) << n
    assert all((int(r["ctrl_mask"]) | (1 << int(r["target_bits"][0]))) == full for r in recs)   # every gate fully controlled
    st = np.eye(32, dtype=np.complex128).reshape(-1).copy()
    orc.run_records(recs, pool, st, 2 * n)
    assert np.abs(st.reshape(32, 32) - u).max() < 1e-12
    # the two-level (ROWS2) form of the same stream
    r2, p2 = two_level_records(stream)
    st2 = np.eye(32, dtype=np.complex128)
    for r in r2:
        off = int(r["matrix_offset"])
        orc.apply_two_level(st2, (int(r["row0"]), int(r["row1"])), p2[off:off + 4].reshape(2, 2))
    assert np.abs(st2 - u).max() < 1e-12


def test_trace_matches_reference_outcomes():
    """CtrlGate.trace(qubit, state) on the gate's own core (cgate.py:228-287): same gate out -- qspace, roles, core matrix,
    phase -- as the reference on 130 (gate, qubit, state) cases shaped like test_cgate.py:673-806, and an exception exactly
    where the reference raises (its constructor rejects a non-unitary result)."""
    from tests.golden_io import load, unpack_gates
    d = load("trace")
    src, res = unpack_gates(d, "src"), unpack_gates(d, "res")
    accepted = 0
    for i, (g, r) in enumerate(zip(src, res)):
        q = g.qspace[int(d["which"][i])]
        try:
            got = g.trace(q, d["state"][i])
            ok = True
        except (AssertionError, NotImplementedError):
            ok = False
        assert ok == bool(d["ok"][i]), i
        if ok:
            accepted += 1
            assert got.qspace == r.qspace
            assert [int(c) for c in got.controls()] == [int(c) for c in r.controls()]
            assert np.abs(got.matrix() - r.matrix()).max() < 1e-12
            assert abs(got.phase() - r.phase()) < 1e-12
    assert accepted == int(d["ok"].sum()) and accepted > 40


def test_trace_argument_errors():
    """Messages of test_cgate.py:716-741."""
    import quantum_simulations_b200 as qs
    g = qs.CtrlGate(np.array(qs.UnivGate.H), [qs.QType.CONTROL1, qs.QType.TARGET, qs.QType.CONTROL0],
                    [qs.Qubit(0), qs.Qubit(1), qs.Qubit(2)])
    with pytest.raises(AssertionError) as e:
        g.trace(g.qspace[0], np.array([[1], [0]]))
    assert str(e.value) == 'state vector must be a 1D array of length 2, but got (2, 1).'
    with pytest.raises(AssertionError) as e:
        g.trace(g.qspace[0], np.array([1, 1]))
    assert str(e.value) == 'state vector must normalized but got [1 1].'
    with pytest.raises(AssertionError) as e:
        g.trace(qs.Qubit(10), np.array([1, 0]))
    assert str(e.value) == 'Qubit q10 not in qspace.'


def test_column_at_a_time_givens_producer_matches_the_sequential_one():
    """two_level_stream_scan (closed form per QR column: prefix norms + one cumulative sum over the rows, torch; the GPU
    producer of config 4) against two_level_stream (one rotation at a time, itself pinned to the reference's
    mat2l_decompose by tests/golden/two_level.npz): same rotations in the same order to 1e-12, skip rules included."""
    from quantum_simulations_b200.utils.mgen import cyclic_matrix, random_unitary
    from quantum_simulations_b200.utils.producers import two_level_stream, two_level_stream_scan
    np.random.seed(3)
    cases = [random_unitary(n) for n in (3, 6, 8, 32, 64)]
    cases += [np.array(cyclic_matrix(8, 1), dtype=complex), np.array(cyclic_matrix(16, 2), dtype=complex), np.eye(8, dtype=complex),
              np.kron(np.eye(2), random_unitary(4)), np.kron(random_unitary(4), np.eye(2))]
    d = __import__("tests.golden_io", fromlist=["load"]).load("two_level")
    cases += [d[k] for k in d.files if k.endswith("_input")]
    for u in cases:
        a, b = two_level_stream(u), two_level_stream_scan(u, "cpu")
        assert [(x[0], x[1]) for x in a] == [(y[0], y[1]) for y in b]
        for x, y in zip(a, b):
            assert np.abs(x[2] - y[2]).max() < 1e-12


def test_vectorised_singlet_records_match_the_loop():
    from quantum_simulations_b200.utils.mgen import cyclic_matrix, random_unitary
    from quantum_simulations_b200.utils.producers import singlet_records, singlet_records_fast, two_level_stream
    np.random.seed(4)
    for n, bb, u in [(3, 0, random_unitary(8)), (5, 5, random_unitary(32)), (6, 2, random_unitary(64)),
                     (4, 0, np.array(cyclic_matrix(16, 2), dtype=complex))]:
        st = two_level_stream(u)
        r1, p1 = singlet_records(st, n, bb)
        r2, p2 = singlet_records_fast(st, n, bb)
        assert len(r1) == len(r2)
        for f in ("kind", "n_targets", "ctrl_mask", "ctrl_value", "target_bits"):
            assert np.array_equal(r1[f], r2[f]), f
        m1 = [p1[int(o):int(o) + 4] for o, k in zip(r1["matrix_offset"], r1["kind"]) if k != _lib.PERMX]
        m2 = [p2[int(o):int(o) + 4] for o, k in zip(r2["matrix_offset"], r2["kind"]) if k != _lib.PERMX]
        assert np.array_equal(np.array(m1), np.array(m2))
