"""Peephole pass over gate records (quantum_simulations_b200/optimize/peephole.py) against the reference's
own results: golden fixtures written by ``oracle/gen_golden.py optimizer`` hold, for compiled streams and hand-made
sequences, the leaves before and after the REAL ``SlidingWindowCombiner``; the record-level pass must leave the same
gates.  The operator must not change (oracle on CPU; ``verify.equal_on_device`` on the GPU)."""
import numpy as np
import pytest

import quantum_simulations_b200 as qs
from oracle import sv_oracle as orc
from quantum_simulations_b200.construct.bytecode import Bytecode, BytecodeIter
from quantum_simulations_b200.optimize import RecordPeephole, grain_of, optimize_leaves, peephole_records
from tests.golden_io import load, unpack_gates

D = load("optimizer")
CASES = [str(n) for n in D["names"]]
T, C1 = qs.QType.TARGET, qs.QType.CONTROL1


def _run(leaves, window, emit):
    return optimize_leaves(leaves, window, qs.GateGrain[emit])


@pytest.mark.parametrize("name", CASES)
def test_same_leaves_as_the_reference(name):
    before = unpack_gates(D, name + "_before")
    want = unpack_gates(D, name + "_after")
    got = _run(before, int(D[name + "_window"]), str(D[name + "_emit"]))
    assert len(got) == len(want), (len(before), len(got), len(want))
    for i, (g, w) in enumerate(zip(got, want)):
        assert g.qspace == w.qspace and [int(c) for c in g.controls()] == [int(c) for c in w.controls()], i
        assert np.abs(np.asarray(g.matrix()) * g.phase() - np.asarray(w.matrix()) * w.phase()).max() < 1e-12, i


@pytest.mark.parametrize("name", [c for c in CASES if c.startswith(("r4", "hand"))])
def test_operator_is_preserved(name):
    before = unpack_gates(D, name + "_before")
    after = _run(before, int(D[name + "_window"]), str(D[name + "_emit"]))
    reg = sorted({q for g in before for q in g.qspace}, key=lambda q: (q.ancilla, q.qid))
    plain = {q: qs.Qubit(i) for i, q in enumerate(reg)}
    ren = lambda gs: [qs.CtrlGate(g.matrix(), g.controls(), [plain[q] for q in g.qspace], g.phase()) for g in gs]  # noqa: E731
    order = [plain[q] for q in reg]
    eye = np.eye(1 << len(reg), dtype=np.complex128)
    # leaves are in matrix-product order: the last one acts first
    a = orc.simulate(ren(before)[::-1], order, eye)
    b = orc.simulate(ren(after)[::-1], order, eye)
    assert np.abs(a - b).max() < 1e-10


def test_textbook_rewrites():
    q = [qs.Qubit(i) for i in range(2)]
    g1 = lambda g, t: qs.CtrlGate(g, [T], [q[t]])  # noqa: E731
    U = qs.UnivGate
    # HXH -> Z needs an emit level that tolerates the intermediate product XH (not a named gate): at CLIFFORD_T the
    # reference's operator stops at the first product that is coarser than ``emit`` (window.py:97-99)
    assert len(_run([g1(U.H, 0), g1(U.X, 0), g1(U.H, 0)], 0, "CLIFFORD_T")) == 3
    out = _run([g1(U.H, 0), g1(U.X, 0), g1(U.H, 0)], 0, "CTRL_PRUNED")
    assert len(out) == 1 and out[0].gate == U.Z
    out = _run([g1(U.T, 0), g1(U.T, 0)], 0, "CLIFFORD_T")
    assert len(out) == 1 and out[0].gate == U.S
    out = _run([g1(U.S, 1), g1(U.S, 1)], 0, "CLIFFORD_T")
    assert len(out) == 1 and out[0].gate == U.Z
    assert _run([g1(U.X, 0), g1(U.X, 0)], 0, "CLIFFORD_T") == []
    cx = qs.CtrlGate(U.X, [C1, T], q)
    assert _run([cx, qs.CtrlGate(U.X, [C1, T], q)], 0, "CLIFFORD_T") == []
    # a gate on another qubit in between does not block; a gate sharing a qubit with different roles does
    assert len(_run([g1(U.T, 0), g1(U.H, 1), g1(U.TD, 0)], 0, "CLIFFORD_T")) == 1
    assert len(_run([g1(U.T, 1), cx, g1(U.TD, 1)], 0, "CLIFFORD_T")) == 3
    # window = 1 lets a leaf see one predecessor only: T T T T -> S S (pairs), not Z
    assert [g.gate for g in _run([g1(U.T, 0)] * 4, 1, "CLIFFORD_T")] == [U.S, U.S]
    assert [g.gate for g in _run([g1(U.T, 0)] * 4, 0, "CLIFFORD_T")] == [U.Z]
    assert grain_of(np.array(U.X), 1) == qs.GateGrain.CLIFFORD_T and grain_of(np.array(U.X), 2) == qs.GateGrain.SINGLET
    assert grain_of(np.array(U.Y), 0) == qs.GateGrain.CTRL_PRUNED and grain_of(np.eye(4), 0) == qs.GateGrain.MULTI_TARGET


def test_record_level_interface():
    """The pass proper sees nothing but qsv_gate_t records and the matrix pool."""
    from quantum_simulations_b200.program import flatten
    q = [qs.Qubit(i) for i in range(3)]
    U = qs.UnivGate
    g1 = lambda g, t: qs.CtrlGate(g, [T], [q[t]])  # noqa: E731
    cx = qs.CtrlGate(U.X, [C1, T], [q[0], q[1]])
    prog = flatten([g1(U.T, 2), cx, g1(U.T, 2), cx, g1(U.H, 0), g1(U.H, 0), g1(U.S, 1)])
    keep, mats, absorbed = peephole_records(prog.records(0), prog.pool())
    assert keep == [2, 6] and absorbed[2] == [0] and absorbed[6] == []
    assert np.allclose(mats[2], np.array(U.S)) and np.allclose(mats[6], np.array(U.S))
    # the tree-level wrapper keeps non-gate leaves as fences
    root = Bytecode(np.eye(2), [Bytecode(g1(U.X, 0)), Bytecode(qs.UnitaryM(8, (0, 3), np.array(U.X))), Bytecode(g1(U.X, 0))])
    out = RecordPeephole().optimize(root)
    assert len([a for a in BytecodeIter(out) if a.is_leaf()]) == 3


@pytest.mark.gpu
def test_optimiser_checked_on_the_gpu():
    from quantum_simulations_b200.optimize import equal_on_device, optimize_checked
    for name in ("r4ct", "r4pr", "hand_w0", "r8sg"):
        before = unpack_gates(D, name + "_before")
        root = Bytecode(np.eye(2), [Bytecode(g) for g in before])
        opt = RecordPeephole(window=int(D[name + "_window"]), emit=qs.GateGrain[str(D[name + "_emit"])])
        out = optimize_checked(root, opt)
        after = [a.data for a in BytecodeIter(out) if a.is_leaf() and not a.skip]
        assert len(after) == len(unpack_gates(D, name + "_after"))
    # and the check does reject a changed operator, exactly and up to a phase
    a = unpack_gates(D, "hand_w0_before")
    q0 = a[0].qspace[0]
    broken = a + [qs.CtrlGate(qs.UnivGate.T, [T], [q0])]
    assert not equal_on_device(a[::-1], broken[::-1])
    phased = a + [qs.CtrlGate(np.exp(0.3j) * np.eye(2), [T], [q0])]
    assert not equal_on_device(a[::-1], phased[::-1])
    assert equal_on_device(a[::-1], phased[::-1], up_to_phase=True)


def test_verify_plumbing_with_a_host_engine(monkeypatch):
    """``optimize.verify`` above oracle-backed stand-ins (tests/host_engine.py): same calls as the GPU test."""
    from quantum_simulations_b200.optimize import equal_on_device, optimize_checked
    from tests import host_engine
    host_engine.install(monkeypatch)
    for name in ("r4ct", "r4pr", "hand_w0", "r8sg"):
        before = unpack_gates(D, name + "_before")
        root = Bytecode(np.eye(2), [Bytecode(g) for g in before])
        opt = RecordPeephole(window=int(D[name + "_window"]), emit=qs.GateGrain[str(D[name + "_emit"])])
        out = optimize_checked(root, opt)
        after = [a.data for a in BytecodeIter(out) if a.is_leaf() and not a.skip]
        assert len(after) == len(unpack_gates(D, name + "_after"))
    a = unpack_gates(D, "hand_w0_before")
    q0 = a[0].qspace[0]
    broken = a + [qs.CtrlGate(qs.UnivGate.T, [T], [q0])]
    info = {}
    assert not equal_on_device(a[::-1], broken[::-1], info=info) and info["violations"] > 0
    phased = a + [qs.CtrlGate(np.exp(0.3j) * np.eye(2), [T], [q0])]
    assert not equal_on_device(a[::-1], phased[::-1])
    assert equal_on_device(a[::-1], phased[::-1], up_to_phase=True)
    with pytest.raises(AssertionError):
        class Bad:
            def optimize(self, root):
                root.children.pop()
                return root
        optimize_checked(Bytecode(np.eye(2), [Bytecode(g) for g in a]), Bad())
