"""Generate tests/golden/*.npz by running the REAL reference.

Run in the build container only (needs /root/reference; it does not exist on the
GPU box):   python oracle/gen_golden.py
The reference pins numpy 1.26.4; under numpy 2.x the abstract dtype
``np.complexfloating`` it passes to ``np.eye`` must be aliased first (shim below).
Everything is seeded with 42 like the reference's ``tests/conftest.py:7-11``.
"""
import os
import random
import sys
from functools import reduce

import scipy  # noqa: F401  (import before the shim)
import sklearn  # noqa: F401
import numpy as np

np.complexfloating = np.complex128
sys.path.insert(0, "/root/reference")

from quompiler.circuits.qdevice import QDevice  # noqa: E402
from quompiler.circuits.qompiler import Qompiler  # noqa: E402
from quompiler.config.config_manager import create_config  # noqa: E402
from quompiler.construct.bytecode import BytecodeIter  # noqa: E402
from quompiler.construct.cgate import CtrlGate  # noqa: E402
from quompiler.construct.qontroller import core2control, ctrl2core  # noqa: E402
from quompiler.construct.qspace import Qubit  # noqa: E402
from quompiler.construct.types import QType, UnivGate  # noqa: E402
from quompiler.construct.unitary import UnitaryM  # noqa: E402
from quompiler.optimize.basic_optimizer import SlidingWindowCombiner  # noqa: E402
from quompiler.utils.mat2l_decompose import mat2l_decompose  # noqa: E402
from quompiler.utils.cnot_decompose import cnot_decompose  # noqa: E402
from quompiler.utils.mgen import (cyclic_matrix, fft_matrix, random_control, random_ctrlgate, random_state,  # noqa: E402
                                  random_unitary, random_UnitaryM, random_UnitaryM_2l)
from quompiler.utils.permute import Permuter  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")
MAXK = 8


def seed(s=42):
    random.seed(s)
    np.random.seed(s)


def pack_gates(gates):
    """CtrlGates -> flat arrays: qid/ancilla/role per slot (padded with -1), matrix pool, phase."""
    G = len(gates)
    qid = -np.ones((G, MAXK), dtype=np.int64)
    anc = np.zeros((G, MAXK), dtype=np.int8)
    role = np.zeros((G, MAXK), dtype=np.int8)
    off = np.zeros(G + 1, dtype=np.int64)
    phase = np.zeros(G, dtype=np.complex128)
    mats = []
    for i, g in enumerate(gates):
        k = len(g.qspace)
        assert k <= MAXK
        for j, (q, c) in enumerate(zip(g.qspace, g.controls())):
            qid[i, j], anc[i, j], role[i, j] = q.qid, q.ancilla, int(c)
        m = np.asarray(g.matrix(), dtype=np.complex128).reshape(-1)
        mats.append(m)
        off[i + 1] = off[i] + m.size
        phase[i] = g.phase()
    return dict(qid=qid, anc=anc, role=role, off=off, phase=phase, pool=np.concatenate(mats) if mats else np.zeros(0, complex))


def prefixed(prefix, d):
    return {f"{prefix}_{k}": v for k, v in d.items()}


def compile_leaves(u, emit, optimize=True, ancilla_offset=100):
    cfg = create_config(emit=emit, ancilla_offset=ancilla_offset)
    qc = Qompiler(cfg, QDevice(cfg.device))
    qc.optimizers.append(SlidingWindowCombiner(window=0, emit=cfg.emit))
    bc = qc.decompose(u)
    if optimize:
        bc = qc.optimize(bc)
    return [a.data for a in BytecodeIter(bc) if a.is_leaf() and not a.skip]


def reference_value(leaves):
    v = reduce(lambda a, b: a @ b, leaves)
    return v.sorted().dela().inflate()


def gen_single_gates():
    seed()
    gates, inflated = [], []
    for _ in range(40):
        n = random.randint(1, 5)
        k = random.randint(1, min(n, 3))
        g = random_ctrlgate(n, k, qnum=8)
        gates.append(g)
        inflated.append(g.inflate())
    for ug in UnivGate:
        n = random.randint(1, 4)
        g = CtrlGate(np.array(ug), random_control(n, 1))
        gates.append(g)
        inflated.append(g.inflate())
    # phase lives only inside the controlled block (test_cgate.py:817-933)
    for _ in range(6):
        ph = np.exp(1j * np.random.uniform(0, 2 * np.pi))
        g = CtrlGate(random_unitary(2), random_control(3, 1), [Qubit(5), Qubit(1), Qubit(3)], phase=ph)
        gates.append(g)
        inflated.append(g.inflate())
    d = pack_gates(gates)
    d["inflate_off"] = np.cumsum([0] + [m.size for m in inflated])
    d["inflate_pool"] = np.concatenate([m.reshape(-1) for m in inflated])
    np.savez_compressed(os.path.join(OUT, "single_gates.npz"), **d)
    print("single_gates", len(gates))


def gen_transforms():
    """sorted / expand / promote / dela / matmul identities on random gates (test_cgate.py:186-806)."""
    seed()
    out = {}
    src, srt, prm, exp = [], [], [], []
    for _ in range(20):
        g = random_ctrlgate(4, 2, qnum=6)
        src.append(g)
        srt.append(g.sorted())
        ctrls = g.control_qids()
        prm.append(g.promote(ctrls[:1]))
        exp.append(g.expand([Qubit(10)], [QType.CONTROL1]))
    out.update(prefixed("src", pack_gates(src)))
    out.update(prefixed("sorted", pack_gates(srt)))
    out.update(prefixed("promote", pack_gates(prm)))
    out.update(prefixed("expand", pack_gates(exp)))
    # pairwise products over different qubit sets (cgate.py:104-121)
    a_list, b_list, prods = [], [], []
    for _ in range(12):
        a = random_ctrlgate(3, 1, qnum=5)
        b = random_ctrlgate(3, 2, qnum=5)
        a_list.append(a)
        b_list.append(b)
        prods.append((a @ b).sorted().inflate())
    out.update(prefixed("mulA", pack_gates(a_list)))
    out.update(prefixed("mulB", pack_gates(b_list)))
    out["mul_off"] = np.cumsum([0] + [m.size for m in prods])
    out["mul_pool"] = np.concatenate([m.reshape(-1) for m in prods])
    # ancilla projection (test_cgate.py:673-713)
    dl_src, dl_out = [], []
    for _ in range(8):
        qs = [Qubit(0), Qubit(1), Qubit(7, ancilla=True)]
        random.shuffle(qs)
        ctrl = [QType.TARGET, QType.TARGET, QType.TARGET]
        ai = [i for i, q in enumerate(qs) if q.ancilla][0]
        ctrl[ai] = random.choice([QType.CONTROL0, QType.CONTROL1])
        g = CtrlGate(random_unitary(4), ctrl, qs)
        dl_src.append(g)
        dl_out.append(g.sorted().dela().inflate())
    out.update(prefixed("dela", pack_gates(dl_src)))
    out["delaval_off"] = np.cumsum([0] + [m.size for m in dl_out])
    out["delaval_pool"] = np.concatenate([m.reshape(-1) for m in dl_out])
    np.savez_compressed(os.path.join(OUT, "transforms.npz"), **out)
    print("transforms ok")


def gen_index_math():
    seed()
    rows = []
    for _ in range(30):
        n = random.randint(1, 6)
        ctrl = random_control(n)
        core = ctrl2core(ctrl)
        back = core2control(n, core) if len(core) > 1 else ctrl
        rows.append((n, [int(c) for c in ctrl], core, [int(c) for c in back]))
    perms = []
    for _ in range(20):
        n = random.randint(2, 6)
        seqp = random.sample(range(20), n)
        p = Permuter(seqp)
        perms.append((seqp, p.bitsortall(range(1 << n)), p.bitpermuteall(range(1 << n))))
    np.savez_compressed(os.path.join(OUT, "index_math.npz"),
                        ctrl=np.array([r[1] + [0] * (6 - r[0]) for r in rows], dtype=np.int8),
                        n=np.array([r[0] for r in rows]),
                        core_off=np.cumsum([0] + [len(r[2]) for r in rows]),
                        core=np.concatenate([np.array(r[2], dtype=np.int64) for r in rows]),
                        back=np.array([r[3] + [0] * (6 - r[0]) for r in rows], dtype=np.int8),
                        perm_n=np.array([len(p[0]) for p in perms]),
                        perm_seq=np.array([p[0] + [-1] * (6 - len(p[0])) for p in perms]),
                        perm_off=np.cumsum([0] + [len(p[1]) for p in perms]),
                        bitsort=np.concatenate([np.array(p[1], dtype=np.int64) for p in perms]),
                        bitpermute=np.concatenate([np.array(p[2], dtype=np.int64) for p in perms]))
    print("index_math ok")


def gen_streams():
    out = {}
    # cfg1: compile_unitary_demo -i 3 (demo/compile_unitary_demo.py:19-30), seed 42
    seed()
    u = random_unitary(8)
    leaves = compile_leaves(u, "PRINCIPAL")
    out.update(prefixed("cfg1", pack_gates(leaves)))
    out["cfg1_input"] = u
    out["cfg1_value"] = reference_value(leaves)
    print("cfg1", len(leaves), np.abs(out["cfg1_value"] - u).max())
    # cfg2: cyclic 8x8 (demo/compile_cyclic_matrix_demo.py), SINGLET (bit-exact) and CLIFFORD_T
    seed()
    u = cyclic_matrix(8, 1)
    leaves = compile_leaves(u, "SINGLET")
    out.update(prefixed("cfg2s", pack_gates(leaves)))
    out["cfg2s_input"] = u.astype(np.complex128)
    out["cfg2s_value"] = reference_value(leaves)
    print("cfg2 SINGLET", len(leaves), np.abs(out["cfg2s_value"] - u).max())
    seed()
    leaves = compile_leaves(u, "CLIFFORD_T")
    out.update(prefixed("cfg2c", pack_gates(leaves)))
    out["cfg2c_input"] = u.astype(np.complex128)
    out["cfg2c_value"] = reference_value(leaves)
    print("cfg2 CLIFFORD_T", len(leaves), np.abs(out["cfg2c_value"] - u).max())
    # cyclic 16 SINGLET (bit-exact)
    seed()
    u = cyclic_matrix(16, 1)
    leaves = compile_leaves(u, "SINGLET")
    out.update(prefixed("cyc16", pack_gates(leaves)))
    out["cyc16_input"] = u.astype(np.complex128)
    out["cyc16_value"] = reference_value(leaves)
    print("cyclic16", len(leaves), np.abs(out["cyc16_value"] - u).max())
    # random 4x4 at every emit level of test_qompiler.py:138-167, random 32x32 at SINGLET
    for emit in ["SINGLET", "CTRL_PRUNED", "PRINCIPAL", "UNIV_GATE", "MULTI_TARGET"]:
        seed()
        u = random_unitary(4)
        try:
            leaves = compile_leaves(u, emit)
        except Exception as e:  # noqa
            print("skip", emit, type(e).__name__, e)
            continue
        leaves = [l for l in leaves if isinstance(l, CtrlGate)]
        key = f"r4{emit.lower()}"
        out.update(prefixed(key, pack_gates(leaves)))
        out[f"{key}_input"] = u
        out[f"{key}_value"] = reference_value(leaves)
        print(key, len(leaves), np.abs(out[f"{key}_value"] - u).max())
    seed()
    u = random_unitary(32)
    leaves = compile_leaves(u, "SINGLET")
    out.update(prefixed("r32", pack_gates(leaves)))
    out["r32_input"] = u
    out["r32_value"] = reference_value(leaves)
    print("r32", len(leaves), np.abs(out["r32_value"] - u).max())
    np.savez_compressed(os.path.join(OUT, "streams.npz"), **out)


def gen_qft():
    """QFT gate circuit (tests/qompiler/utils/test_mgen.py:84-99 generalised) through the reference's @."""
    out = {}
    for n in (3, 4, 5, 6):
        q = [Qubit(i) for i in range(n)]
        T, C1 = QType.TARGET, QType.CONTROL1
        time_order = []
        for i in range(n):
            time_order.append(CtrlGate(UnivGate.H, [T], [q[i]]))
            for j in range(i + 1, n):
                ph = np.exp(2j * np.pi / (1 << (j - i + 1)))
                time_order.append(CtrlGate(np.array([[1, 0], [0, ph]]), [T, C1], [q[i], q[j]]))
        for i in range(n // 2):
            a, b = q[i], q[n - 1 - i]
            time_order += [CtrlGate(UnivGate.X, [C1, T], [a, b]), CtrlGate(UnivGate.X, [T, C1], [a, b]),
                           CtrlGate(UnivGate.X, [C1, T], [a, b])]
        val = reduce(lambda a, b: a @ b, time_order[::-1]).sorted().inflate()
        err = np.abs(val - fft_matrix(n)).max()
        print("qft", n, len(time_order), err)
        assert err < 1e-12
        out[f"qft{n}_value"] = val
        out[f"qft{n}_fft"] = fft_matrix(n)
    seed()
    out["state64"] = random_state(64)
    np.savez_compressed(os.path.join(OUT, "qft.npz"), **out)


def gen_two_level():
    out = {}
    for name, dim in (("d8", 8), ("d3", 3), ("d6", 6)):
        seed()
        u = random_unitary(dim)
        comps = mat2l_decompose(UnitaryM(dim, tuple(range(dim)), u))
        out[f"{name}_input"] = u
        out[f"{name}_core"] = np.array([c.core for c in comps], dtype=np.int64)
        out[f"{name}_mat"] = np.array([c.matrix for c in comps], dtype=np.complex128)
        out[f"{name}_phase"] = np.array([c.phase for c in comps], dtype=np.complex128)
        val = reduce(lambda a, b: a @ b, comps).inflate()
        out[f"{name}_value"] = val
        print("two-level", name, len(comps), np.abs(val - u).max())
    # general cores with a phase (test_unitary.py:30-71)
    seed()
    us = [random_UnitaryM(5, (0, 2, 4), phase=np.exp(0.3j)), random_UnitaryM_2l(5, 1, 3, phase=np.exp(1.1j)), random_UnitaryM(5, (4, 1, 0, 3))]
    out["gen_dim"] = np.array(5)
    out["gen_ncore"] = np.array([len(u.core) for u in us])
    out["gen_core"] = np.array([list(u.core) + [-1] * (4 - len(u.core)) for u in us])
    out["gen_mat"] = np.concatenate([u.matrix.reshape(-1) for u in us])
    out["gen_phase"] = np.array([u.phase for u in us])
    out["gen_inflate"] = np.concatenate([u.inflate().reshape(-1) for u in us])
    out["gen_value"] = reduce(lambda a, b: a @ b, us).inflate()
    # Gray-code chain of one two-level matrix (cnot_decompose.py:84-107), 3 qubits
    seed()
    m2 = random_UnitaryM_2l(8, 1, 6)
    chain = list(cnot_decompose(m2))
    out.update(prefixed("gray", pack_gates(chain)))
    out["gray_value"] = m2.inflate()
    print("gray chain", len(chain))
    np.savez_compressed(os.path.join(OUT, "two_level.npz"), **out)


def gen_qco():
    """Compiled-code files exactly as the reference writes them (Qompiler.compile -> pickle, file_io.py:9-12)."""
    import pickle
    out = {}
    for name, u, emit in (("qco_r4", None, "CTRL_PRUNED"), ("qco_cyc8", cyclic_matrix(8, 1), "SINGLET"), ("qco_r8p", None, "PRINCIPAL")):
        seed()
        if u is None:
            u = random_unitary(4 if name == "qco_r4" else 8)
        cfg = create_config(emit=emit, ancilla_offset=100, output=os.path.join(OUT, name + ".qco"))
        qc = Qompiler(cfg, QDevice(cfg.device))
        qc.optimizers.append(SlidingWindowCombiner(window=0, emit=cfg.emit))
        qc.compile(u)
        out[name + "_input"] = np.asarray(u, dtype=np.complex128)
        print(name, os.path.getsize(os.path.join(OUT, name + ".qco")), "bytes")
    np.savez_compressed(os.path.join(OUT, "qco_inputs.npz"), **out)


def gen_optimizer():
    """SlidingWindowCombiner (optimize/basic_optimizer.py:11-49, window.py:25-104) run by the real reference on flat
    leaf lists: the un-optimised compiled leaves of several unitaries / emit levels / window sizes, plus hand-made
    sequences (HXH, TT, SS, XX, identities, interleaved qubits).  Stored: leaves before, leaves after."""
    from quompiler.construct.bytecode import Bytecode
    from quompiler.construct.types import GateGrain
    out = {}
    names = []

    def run_case(name, leaves, window, emit):
        leaves = [l for l in leaves if isinstance(l, CtrlGate)]
        root = Bytecode(np.eye(2), [Bytecode(CtrlGate(np.array(g.matrix()), g.controls(), g.qspace, g.phase())) for g in leaves])
        opt = SlidingWindowCombiner(window=window, emit=GateGrain[emit])
        res = opt.optimize(root)
        after = [a.data for a in BytecodeIter(res) if a.is_leaf() and not a.skip and isinstance(a.data, CtrlGate)]
        out.update(prefixed(name + "_before", pack_gates(leaves)))
        out.update(prefixed(name + "_after", pack_gates(after)))
        out[name + "_window"] = np.int64(window)
        out[name + "_emit"] = np.str_(emit)
        names.append(name)
        print("opt", name, len(leaves), "->", len(after))

    for tag, dim, emit, window in (("r4ct", 4, "CLIFFORD_T", 0), ("r4ug", 4, "UNIV_GATE", 0), ("r4pr", 4, "PRINCIPAL", 0),
                                   ("r4cp", 4, "CTRL_PRUNED", 0), ("r8pr", 8, "PRINCIPAL", 0), ("r8pr_w2", 8, "PRINCIPAL", 2),
                                   ("r8cp_w1", 8, "CTRL_PRUNED", 1), ("r8sg", 8, "SINGLET", 0)):
        seed()
        u = random_unitary(dim)
        run_case(tag, compile_leaves(u, emit, optimize=False), window, emit)
    seed()
    run_case("cyc8ct", compile_leaves(cyclic_matrix(8, 1), "CLIFFORD_T", optimize=False), 0, "CLIFFORD_T")
    T_, C1_, C0_ = QType.TARGET, QType.CONTROL1, QType.CONTROL0
    q = [Qubit(i) for i in range(4)]
    g1 = lambda g, t: CtrlGate(g, [T_], [q[t]])  # noqa: E731
    cx = lambda c, t: CtrlGate(UnivGate.X, [C1_, T_], [q[c], q[t]])  # noqa: E731
    hand = [g1(UnivGate.H, 0), g1(UnivGate.X, 0), g1(UnivGate.H, 0), g1(UnivGate.T, 1), g1(UnivGate.T, 1), g1(UnivGate.S, 2),
            g1(UnivGate.S, 2), g1(UnivGate.X, 3), g1(UnivGate.X, 3), g1(UnivGate.I, 0), cx(0, 1), cx(0, 1), cx(1, 2), g1(UnivGate.H, 1),
            cx(1, 2), g1(UnivGate.T, 0), g1(UnivGate.H, 3), g1(UnivGate.TD, 0), CtrlGate(UnivGate.Z, [C0_, T_], [q[2], q[3]]),
            CtrlGate(UnivGate.Z, [T_, C0_], [q[3], q[2]]), g1(UnivGate.Y, 1), g1(UnivGate.H, 1), g1(UnivGate.Y, 1), g1(UnivGate.H, 1)]
    for w in (0, 1, 2):
        run_case(f"hand_w{w}", hand, w, "CLIFFORD_T")
    run_case("hand_ug", hand, 0, "UNIV_GATE")
    out["names"] = np.array(names)
    np.savez_compressed(os.path.join(OUT, "optimizer.npz"), **out)


def gen_trace():
    """CtrlGate.trace(qubit, state) / dela(state) with states other than |0> (cgate.py:228-305, qproject
    inter_product.py:318-363; cases shaped like test_cgate.py:673-806).

    (a) single gates: every (gate, qubit, state) is recorded with the reference's outcome -- the traced gate, or the
        fact that the reference raised (its constructor rejects a non-unitary result);
    (b) products with ancillas: the sandwich <psi|V|psi> on every ancilla of V = reduce(@), taken with the reference's
        own qproject on V.sorted().inflate() (valid for any state), and V.dela(psi).inflate() where the reference
        accepts it."""
    from quompiler.utils.inter_product import qproject, mykron
    seed()
    out = {}
    states = [np.array([1, 0]), np.array([0, 1]), np.array([1, 1]) / np.sqrt(2), random_state(2), random_state(2)]
    src, res, which, st_used, ok = [], [], [], [], []
    cases = []
    for ctrl in (QType.CONTROL0, QType.CONTROL1):
        for qidx in (0, 1):
            u = random_unitary(2)
            mat = np.kron(np.eye(2), u) if qidx == 0 else np.kron(u, np.eye(2))
            cases.append(CtrlGate(mat, [ctrl, QType.TARGET, QType.TARGET]))
        cases.append(CtrlGate(random_unitary(2), [ctrl, QType.TARGET], phase=np.exp(0.7j)))
        cases.append(CtrlGate(random_unitary(4), [QType.TARGET, ctrl, QType.TARGET], [Qubit(4), Qubit(2), Qubit(9, ancilla=True)]))
    cases.append(CtrlGate(mykron(np.eye(2), random_unitary(2), np.eye(2)), [QType.CONTROL1] + [QType.TARGET] * 3,
                          [Qubit(10), Qubit(12, ancilla=True), Qubit(7), Qubit(1, ancilla=True)]))
    for g in cases:
        for qi, q in enumerate(g.qspace):
            for si, st in enumerate(states):
                try:
                    r = g.trace(q, st)
                    good = True
                except (AssertionError, NotImplementedError):
                    r = CtrlGate(np.eye(2), [QType.TARGET])
                    good = False
                src.append(g)
                res.append(r)
                which.append(qi)
                st_used.append(st.astype(np.complex128))
                ok.append(good)
    out.update(prefixed("src", pack_gates(src)))
    out.update(prefixed("res", pack_gates(res)))
    out["which"] = np.array(which)
    out["state"] = np.array(st_used)
    out["ok"] = np.array(ok)
    # (b) products
    chains, vals, cstate, ref_ok, ref_vals = [], [], [], [], []
    a1, a2 = Qubit(101, ancilla=True), Qubit(102, ancilla=True)
    T, C0, C1 = QType.TARGET, QType.CONTROL0, QType.CONTROL1
    for trial in range(10):
        qs_ = [Qubit(0), Qubit(1), Qubit(2)]
        anc = [a1] if trial % 2 == 0 else [a1, a2]
        gates = []
        for _ in range(4):
            pick = random.sample(qs_ + anc, 3)
            gates.append(CtrlGate(random_unitary(2), random_control(3, 1), pick))
        if trial >= 6:   # ancillas only ever used as controls: with a basis state the reference accepts the result
            gates = [CtrlGate(random_unitary(2), [random.choice([C0, C1]), T], [anc[0], random.choice(qs_)]) for _ in range(3)]
            gates.append(CtrlGate(random_unitary(2), [T, C1], [qs_[0], qs_[1]]))
        st = states[trial % len(states)] if trial < 6 else states[trial % 2]
        V = reduce(lambda a, b: a @ b, gates).sorted()
        m = V.inflate()
        space = list(V.qspace)
        for q in [q for q in space if q.ancilla][::-1]:
            m = qproject(m, space.index(q), st)
            space.remove(q)
        try:
            rv = V.dela(st).inflate()
            r_ok = True
        except AssertionError:
            rv = np.zeros_like(m)
            r_ok = False
        chains.append(gates)
        vals.append(m)
        cstate.append(st.astype(np.complex128))
        ref_ok.append(r_ok)
        ref_vals.append(rv)
    for i, gates in enumerate(chains):
        out.update(prefixed(f"chain{i}", pack_gates(gates)))
        out[f"chain{i}_value"] = vals[i]
        out[f"chain{i}_refdela"] = ref_vals[i]
    out["chain_state"] = np.array(cstate)
    out["chain_ref_ok"] = np.array(ref_ok)
    out["n_chains"] = np.array(len(chains))
    np.savez_compressed(os.path.join(OUT, "trace.npz"), **out)
    print("trace", len(src), "single cases,", int(np.sum(ok)), "accepted;", len(chains), "chains,", int(np.sum(ref_ok)), "accepted by the reference")


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    if len(sys.argv) > 1 and sys.argv[1] == "optimizer":
        gen_optimizer()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "trace":
        gen_trace()
        sys.exit(0)
    gen_single_gates()
    gen_transforms()
    gen_index_math()
    gen_streams()
    gen_qft()
    gen_two_level()
    gen_qco()
    gen_optimizer()
    gen_trace()
    print("done ->", os.path.abspath(OUT))
