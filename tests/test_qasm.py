"""OpenQASM 3 exporter (circuits/qasm_builder.py): the text parses back into the gate program it was
built from, and that program reproduces the circuit's unitary on the oracle (CPU) and on the CUDA engine (GPU).
Parity with qiskit's ``qasm.dumps`` is UNPINNED (qiskit is not installed)."""
import cmath
import random
import warnings

import numpy as np
import pytest

import quantum_simulations_b200 as qs
from oracle import sv_oracle as orc
from quantum_simulations_b200.circuits.qasm_builder import QasmBuilder, parse_qasm, u_matrix, zyz_angles
from quantum_simulations_b200.utils.mgen import random_ctrlgate, random_unitary
from tests.golden_io import load, unpack_gates

T, C0, C1 = qs.QType.TARGET, qs.QType.CONTROL0, qs.QType.CONTROL1


def _unitary_of(prog):
    """Dense operator of a Program over all its qubits, by the oracle (columns = basis states)."""
    n = prog.n_qubits
    st = np.eye(1 << n, dtype=np.complex128).reshape(-1).copy()
    orc.run_records(prog.records(n), prog.pool(), st, 2 * n)
    return st.reshape(1 << n, 1 << n)


def test_zyz_angles_cover_the_degenerate_cases():
    np.random.seed(4)
    mats = [random_unitary(2) for _ in range(50)]
    mats += [np.diag(np.exp(1j * np.random.uniform(0, 6, 2))) for _ in range(10)]
    mats += [np.array([[0, 1], [1, 0]]) @ np.diag(np.exp(1j * np.random.uniform(0, 6, 2))) for _ in range(10)]
    mats += [np.array(g) for g in qs.UnivGate]
    for m in mats:
        t, p, l, g = zyz_angles(m)
        assert np.abs(cmath.exp(1j * g) * u_matrix(t, p, l) - m).max() < 1e-13


def _random_one_target_gates(n, count, seed):
    random.seed(seed)
    np.random.seed(seed)
    gates = []
    for _ in range(count):
        k = random.randint(1, min(4, n))
        g = random_ctrlgate(k, 1, qnum=n)
        r = random.random()
        ph = np.exp(1j * np.random.uniform(0, 6.28)) if random.random() < 0.5 else 1.0
        if r < 0.3:
            g = qs.CtrlGate(random.choice(list(qs.UnivGate)), g.controls(), g.qspace, phase=ph)
        elif r < 0.5:
            g = qs.CtrlGate(qs.RGate(np.random.uniform(-3, 3), random.choice(list(qs.PrincipalAxis))), g.controls(), g.qspace, phase=ph)
        elif r < 0.6:
            g = qs.CtrlGate(np.diag(np.exp(1j * np.random.uniform(0, 6.28, 2))), g.controls(), g.qspace)
        gates.append(g)
    return gates


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_text_round_trips_to_the_same_operator(seed):
    n = 5
    gates = _random_one_target_gates(n, 60, seed)
    reg = [qs.Qubit(i) for i in range(n)]
    b = QasmBuilder()
    b.register(reg)
    for g in gates:
        b.build_gate(g)
    circ = b.finish()
    text = circ.qasm()
    assert text.startswith("OPENQASM 3.0;") and f"qubit[{n}] q;" in text and len(circ) == len(gates)
    got = _unitary_of(parse_qasm(text))
    want = orc.simulate(gates, reg, np.eye(1 << n, dtype=np.complex128))
    assert np.abs(got - want).max() < 1e-12
    # names: standard gates and principal rotations are written by name, the rest as U(...)
    assert any(re_ in text for re_ in ("rx(", "ry(", "rz(")) and "U(" in text and "negctrl @" in text and "gphase(" in text


@pytest.mark.parametrize("name", ["cfg1", "cfg2s", "cfg2c", "r4singlet", "r4principal", "r4univ_gate"])
def test_compiled_streams_export_and_reproduce_their_unitary(name):
    """The reference's compiled gate streams (golden fixtures): exported, parsed back, simulated with ancillas in |0>
    and projected on <0| -- equals the unitary that was compiled."""
    d = load("streams")
    leaves = unpack_gates(d, name)                  # matrix-product order
    b = QasmBuilder()
    for g in leaves[::-1]:                           # circuit time order, as QRenderer feeds a builder
        b.build_gate(g)
    circ = b.finish()
    prog = circ.program()
    n, am = prog.n_qubits, prog.ancilla_mask
    keep = [i for i in range(1 << n) if not (i & am)]
    k = len(keep).bit_length() - 1
    st = np.zeros((1 << n, 1 << k), dtype=np.complex128)
    st[keep, np.arange(len(keep))] = 1
    flat = st.reshape(-1)
    orc.run_records(prog.records(k), prog.pool(), flat, n + k)
    got = flat.reshape(1 << n, 1 << k)[keep, :]
    assert np.abs(got - d[f"{name}_value"]).max() < 1e-10
    if am:
        assert "] a;" in circ.qasm()


def test_builder_protocol():
    b = qs.get_builder(qs.QompilePlatform.QASM)
    assert isinstance(b, QasmBuilder)
    b.register([qs.Qubit(1), qs.Qubit(0)])
    b.register([qs.Qubit(1), qs.Qubit(7, True)])            # overlapping registrations are fine
    assert [repr(q) for q in b.all_qubits()] == ["q0", "q1", "a7"]
    b.build_gate(qs.CtrlGate(qs.UnivGate.H, [T], [qs.Qubit(1)]))
    first = b.finish().qasm()
    b.build_gate(qs.CtrlGate(qs.UnivGate.X, [C1, T], [qs.Qubit(7, True), qs.Qubit(0)]))
    second = b.finish()                                      # finish may be called more than once
    assert first.count("\n") + 1 == second.qasm().count("\n") and "ctrl @ x a[0], q[0];" in second.qasm()
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        b.build_gate(qs.UnitaryM(4, (1, 2), random_unitary(2)))
    assert w and "ignored" in str(w[0].message)
    with pytest.raises(NotImplementedError):
        b.build_gate(random_ctrlgate(3, 2, qnum=3))
        b.finish()
    with pytest.raises(ValueError):
        parse_qasm("OPENQASM 3.0;\nqubit[1] q;\nfoo q[0];\n")
    with pytest.raises(ValueError):
        parse_qasm(second.qasm(), qubits=[qs.Qubit(0)])


def test_qubit_order_permutes_the_operator():
    gates = _random_one_target_gates(3, 10, 8)
    reg = [qs.Qubit(i) for i in range(3)]
    b = QasmBuilder()
    b.register(reg)
    for g in gates:
        b.build_gate(g)
    circ = b.finish()
    order = [reg[2], reg[0], reg[1]]
    got = _unitary_of(circ.program(order))
    want = orc.simulate(gates, order, np.eye(8, dtype=np.complex128))
    assert np.abs(got - want).max() < 1e-12


@pytest.mark.gpu
def test_text_evaluated_by_the_cuda_engine():
    d = load("streams")
    for name in ("cfg1", "cfg2s"):
        b = QasmBuilder()
        for g in unpack_gates(d, name)[::-1]:
            b.build_gate(g)
        got = b.finish().unitary(drop_ancilla=True)
        assert np.abs(got - d[f"{name}_value"]).max() < 1e-10
    gates = _random_one_target_gates(4, 40, 5)
    reg = [qs.Qubit(i) for i in range(4)]
    b = QasmBuilder()
    b.register(reg)
    for g in gates:
        b.build_gate(g)
    got = b.finish().unitary()
    want = orc.simulate(gates, reg, np.eye(16, dtype=np.complex128))
    assert np.abs(got - want).max() < 1e-10


def test_unitary_plumbing_with_a_host_engine(monkeypatch):
    """``QasmCircuit.unitary`` above an oracle-backed ``simulate_columns`` (tests/host_engine.py): the same calls the GPU
    test makes, so its argument plumbing (ancilla handling, qubit order) is exercised on a CPU box too."""
    from tests import host_engine
    host_engine.install(monkeypatch)
    d = load("streams")
    for name in ("cfg1", "cfg2s"):
        b = QasmBuilder()
        for g in unpack_gates(d, name)[::-1]:
            b.build_gate(g)
        assert np.abs(b.finish().unitary(drop_ancilla=True) - d[f"{name}_value"]).max() < 1e-10
    gates = _random_one_target_gates(4, 40, 5)
    reg = [qs.Qubit(i) for i in range(4)]
    b = QasmBuilder()
    b.register(reg)
    for g in gates:
        b.build_gate(g)
    want = orc.simulate(gates, reg, np.eye(16, dtype=np.complex128))
    assert np.abs(b.finish().unitary() - want).max() < 1e-10
    order = [reg[3], reg[1], reg[0], reg[2]]
    assert np.abs(b.finish().unitary(order) - orc.simulate(gates, order, np.eye(16, dtype=np.complex128))).max() < 1e-10
    # ancillas kept as ordinary qubits: the full 16x16 operator of a 3+1 register
    gs = [qs.CtrlGate(qs.UnivGate.H, [T], [qs.Qubit(0)]), qs.CtrlGate(qs.UnivGate.X, [C1, T], [qs.Qubit(0), qs.Qubit(9, True)])]
    b = QasmBuilder()
    for g in gs:
        b.build_gate(g)
    full = b.finish().unitary()
    assert full.shape == (4, 4) and abs(full[0b11, 0] - 1 / np.sqrt(2)) < 1e-12
    assert b.finish().unitary(drop_ancilla=True).shape == (2, 2)


# ---- hand-written known-answer files (tests/golden/qasm/README.md) -------------------------------------------------
def _kron(*ms):
    out = np.eye(1)
    for m in ms:
        out = np.kron(out, m)
    return out


def _ctrl(n, controls, target, m):
    """Operator over n qubits (qubit 0 most significant): `m` on `target` where every (qubit, value) in `controls` holds."""
    dim = 1 << n
    out = np.zeros((dim, dim), dtype=complex)
    for col in range(dim):
        bit = lambda q, x=col: (x >> (n - 1 - q)) & 1  # noqa: E731
        if all(bit(q) == v for q, v in controls):
            for r in (0, 1):
                row = (col & ~(1 << (n - 1 - target))) | (r << (n - 1 - target))
                out[row, col] += m[r, bit(target)]
        else:
            out[col, col] = 1
    return out


def _known_answers():
    I2 = np.eye(2)
    X = np.array([[0, 1], [1, 0]], dtype=complex)
    Y = np.array([[0, -1j], [1j, 0]])
    Z = np.diag([1, -1]).astype(complex)
    H = np.array([[1, 1], [1, -1]]) / np.sqrt(2)
    rx = lambda t: np.array([[np.cos(t / 2), -1j * np.sin(t / 2)], [-1j * np.sin(t / 2), np.cos(t / 2)]])  # noqa: E731
    ry = lambda t: np.array([[np.cos(t / 2), -np.sin(t / 2)], [np.sin(t / 2), np.cos(t / 2)]], dtype=complex)  # noqa: E731
    rz = lambda t: np.diag([np.exp(-0.5j * t), np.exp(0.5j * t)])  # noqa: E731
    ph = lambda t: np.diag([1, np.exp(1j * t)])  # noqa: E731
    t_, p_, l_ = 1.2, -0.4, 2.2
    u = np.array([[np.cos(t_ / 2), -np.exp(1j * l_) * np.sin(t_ / 2)],
                  [np.exp(1j * p_) * np.sin(t_ / 2), np.exp(1j * (p_ + l_)) * np.cos(t_ / 2)]])   # OpenQASM 3.0 spec, built-in U
    return {
        "bell": _ctrl(2, [(0, 1)], 1, X) @ _kron(H, I2),
        "mixed_controls": _ctrl(3, [(0, 1), (1, 0)], 2, X),
        "controlled_phase": _kron(np.diag([np.exp(0.5j * np.pi), 1]), I2) @ _kron(I2, ph(np.pi / 4)),
        "u_and_gphase": np.exp(0.25j) * u,
        "rotations": _kron(np.diag([1, np.exp(-0.25j * np.pi)]), I2) @ _kron(I2, np.diag([1, -1j])) @ _ctrl(2, [(1, 0)], 0, rz(0.5))
                     @ _kron(rz(2.0), I2) @ _kron(I2, ry(-1.1)) @ _kron(rx(0.3), I2),
        "ancilla_register": _kron(Y, I2, I2) @ _ctrl(3, [(0, 1), (1, 1)], 2, Z) @ _ctrl(3, [(2, 1)], 1, H),
    }


def _golden_text(name):
    import os
    return open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "qasm", name + ".qasm")).read()


@pytest.mark.parametrize("name", ["bell", "mixed_controls", "controlled_phase", "u_and_gphase", "rotations", "ancilla_register"])
def test_reader_on_hand_written_known_answers(name):
    """The reader is pinned to files it did not write: each hand-written OpenQASM 3 program must parse into the
    operator worked out independently from the specification (ancillas counted as ordinary qubits)."""
    prog = parse_qasm(_golden_text(name), ancilla_as_plain=True)
    assert np.abs(_unitary_of(prog) - _known_answers()[name]).max() < 1e-13


def test_writer_emits_the_hand_written_statements():
    """... and the writer to the same files: gate objects -> exactly the statements of the hand-written text."""
    q = [qs.Qubit(i) for i in range(3)]
    a0 = qs.Qubit(7, ancilla=True)
    U = qs.UnivGate

    def body(text):
        return [ln.strip() for ln in text.splitlines() if ln.strip() and not ln.startswith(("OPENQASM", "include", "qubit", "//"))]

    def emit(gates, extra=()):
        b = QasmBuilder()
        b.register(list(extra))
        for g in gates:
            b.register(g.qspace)
            b.build_gate(g)
        return body(b.finish().qasm())

    assert emit([qs.CtrlGate(U.H, [T], [q[0]]), qs.CtrlGate(U.X, [C1, T], [q[0], q[1]])]) == body(_golden_text("bell"))
    assert emit([qs.CtrlGate(U.X, [C1, C0, T], q)]) == body(_golden_text("mixed_controls"))
    rot = [qs.CtrlGate(qs.RGate(0.3, qs.PrincipalAxis.X), [T], [q[0]]), qs.CtrlGate(qs.RGate(-1.1, qs.PrincipalAxis.Y), [T], [q[1]]),
           qs.CtrlGate(qs.RGate(2.0, qs.PrincipalAxis.Z), [T], [q[0]]), qs.CtrlGate(qs.RGate(0.5, qs.PrincipalAxis.Z), [C0, T], [q[1], q[0]]),
           qs.CtrlGate(U.SD, [T], [q[1]]), qs.CtrlGate(U.TD, [T], [q[0]])]
    assert emit(rot) == body(_golden_text("rotations"))
    anc = [qs.CtrlGate(U.H, [C1, T], [a0, q[1]]), qs.CtrlGate(U.Z, [C1, C1, T], [q[0], q[1], a0]), qs.CtrlGate(U.Y, [T], [q[0]])]
    assert emit(anc) == body(_golden_text("ancilla_register"))
    # a general one-qubit unitary goes out as gphase + U with the angles of the specification's U
    t_, p_, l_ = 1.2, -0.4, 2.2
    lines = emit([qs.CtrlGate(np.exp(0.25j) * u_matrix(t_, p_, l_), [T], [q[0]])])
    assert len(lines) == 2 and lines[0].startswith("gphase(") and lines[1].startswith("U(") and lines[1].endswith(" q[0];")
    got = [float(x) for x in lines[1][2:lines[1].index(")")].split(",")] + [float(lines[0][7:lines[0].index(")")])]
    assert np.allclose(got, [t_, p_, l_, 0.25], atol=1e-12)
