"""Pins the oracle (oracle/sv_oracle.py) to outputs of the real reference (tests/golden)."""
import numpy as np
import pytest

from oracle import sv_oracle as orc
from tests.golden_io import STREAMS, load, unpack_gates, unpack_mats

TOL = 1e-12


def test_single_gate_inflate_matches_reference():
    d = load("single_gates")
    gates = unpack_gates(d)
    expected = unpack_mats(d, "inflate_off", "inflate_pool")
    assert len(gates) == len(expected) == 55
    for g, e in zip(gates, expected):
        assert np.abs(orc.inflate_gate(g) - e).max() < TOL
        # statevector restatement over the gate's own qubit order
        assert np.abs(orc.unitary_of([g], register=g.qspace, drop_ancilla=False) - e).max() < TOL


def test_index_math_matches_reference():
    d = load("index_math")
    for i in range(len(d["n"])):
        n = int(d["n"][i])
        ctrl = [int(c) for c in d["ctrl"][i][:n]]
        core = d["core"][d["core_off"][i]:d["core_off"][i + 1]].tolist()
        assert orc.ctrl2core(ctrl) == core
        if len(core) > 1:
            assert [int(c) for c in orc.core2control(n, core)] == [int(c) for c in d["back"][i][:n]]


@pytest.mark.parametrize("name", STREAMS)
def test_stream_value_matches_reference(name):
    d = load("streams")
    leaves = unpack_gates(d, name)
    expected = d[f"{name}_value"]
    got = orc.unitary_of(leaves)                      # statevector restatement
    assert got.shape == expected.shape
    assert np.abs(got - expected).max() < 1e-11
    if len(leaves) <= 500:                            # dense restatement, independent route
        reg = orc.register_of(leaves)
        dense = orc.project_ancilla(orc.dense_product(leaves, reg), reg)
        assert np.abs(dense - expected).max() < 1e-11


@pytest.mark.parametrize("name", ["cfg2s", "cyc16"])
def test_permutation_streams_are_bit_exact(name):
    d = load("streams")
    leaves = unpack_gates(d, name)
    assert np.array_equal(orc.unitary_of(leaves), d[f"{name}_value"])
    assert np.array_equal(d[f"{name}_value"], d[f"{name}_input"])


def test_pairwise_products_and_dela():
    d = load("transforms")
    a, b = unpack_gates(d, "mulA"), unpack_gates(d, "mulB")
    for x, y, e in zip(a, b, unpack_mats(d, "mul_off", "mul_pool")):
        assert np.abs(orc.unitary_of([x, y], drop_ancilla=False) - e).max() < TOL
    for g, e in zip(unpack_gates(d, "dela"), unpack_mats(d, "delaval_off", "delaval_pool")):
        assert np.abs(orc.unitary_of([g]) - e).max() < TOL
        reg = orc.register_of([g])
        assert np.abs(orc.project_ancilla(orc.dense_product([g], reg), reg) - e).max() < TOL


def test_qft_circuit_matches_reference_and_fft():
    from quantum_simulations_b200.utils.mgen import fft_matrix, qft_circuit
    d = load("qft")
    for n in (3, 4, 5, 6):
        circ = qft_circuit(n)
        got = orc.unitary_of(circ[::-1], drop_ancilla=False)
        assert np.abs(got - d[f"qft{n}_value"]).max() < TOL
        assert np.abs(got - d[f"qft{n}_fft"]).max() < TOL
        assert np.abs(fft_matrix(n) - d[f"qft{n}_fft"]).max() < 1e-15
    x = d["state64"]
    reg = orc.register_of(qft_circuit(6))
    assert np.abs(orc.simulate(qft_circuit(6), reg, x) - orc.qft_reference(x)).max() < TOL


def test_two_level_streams():
    import quantum_simulations_b200 as qs
    d = load("two_level")
    for name, dim in (("d8", 8), ("d3", 3), ("d6", 6)):
        us = [qs.UnitaryM(dim, tuple(c), m, p) for c, m, p in zip(d[f"{name}_core"], d[f"{name}_mat"], d[f"{name}_phase"])]
        got = orc.unitary_of_rows(us, dim)
        assert np.abs(got - d[f"{name}_value"]).max() < TOL
        assert np.abs(got - d[f"{name}_input"]).max() < 1e-11
    # general cores with phases
    dim = int(d["gen_dim"])
    off = 0
    ioff = 0
    us = []
    for k, core, ph in zip(d["gen_ncore"], d["gen_core"], d["gen_phase"]):
        k = int(k)
        m = d["gen_mat"][off:off + k * k].reshape(k, k)
        off += k * k
        u = qs.UnitaryM(dim, tuple(int(c) for c in core[:k]), m, complex(ph))
        e = d["gen_inflate"][ioff:ioff + dim * dim].reshape(dim, dim)
        ioff += dim * dim
        assert np.abs(orc.inflate_unitarym(dim, u.core, u.matrix, u.phase) - e).max() < TOL
        us.append(u)
    assert np.abs(orc.unitary_of_rows(us, dim) - d["gen_value"]).max() < TOL
    chain = unpack_gates(d, "gray")
    assert np.abs(orc.unitary_of(chain, drop_ancilla=False) - d["gray_value"]).max() < TOL


@pytest.mark.parametrize("name", ["cfg1", "cfg2s", "cfg2c", "r32", "r4multi_target"])
def test_c_oracle_matches_reference(name):
    """The C/OpenMP restatement (timed CPU baseline) run on flattened records equals the reference value."""
    from oracle import c_oracle
    from quantum_simulations_b200.program import flatten
    d = load("streams")
    leaves = unpack_gates(d, name)
    prog = flatten(leaves[::-1])
    n = prog.n_qubits
    am = prog.ancilla_mask
    keep = [i for i in range(1 << n) if not (i & am)]
    b = len(keep).bit_length() - 1
    st = np.zeros((1 << n, 1 << b), dtype=np.complex128)
    st[keep, np.arange(len(keep))] = 1
    flat = st.reshape(-1)
    c_oracle.run_records(prog.records(b), prog.pool(), flat, n + b, b)
    got = flat.reshape(1 << n, 1 << b)[keep, :]
    assert np.abs(got - d[f"{name}_value"]).max() < 1e-11
    if name == "cfg2s":
        assert np.array_equal(got, d[f"{name}_value"])
