"""The fusion pass (host code of libqsv) checked WITHOUT a GPU: ``qsv_hostcheck_fused`` interprets the fused plan on
the host with the same per-op source the CUDA kernel is compiled from (include/qsv_hostcheck.h), and the result must
match the CPU oracle.  Covers what the GPU fuzz covers (tools/fuzz_fused.py) plus circuits built to hit the FFT-shaped
launches (OP_FSTAGE) in all their shapes."""
import importlib.util
import os

import numpy as np
import pytest

from tests.hostcheck import ROOT, run_fused_on_host
from oracle import sv_oracle
from quantum_simulations_b200.construct.qspace import Qubit
from quantum_simulations_b200.program import Program
from quantum_simulations_b200.utils.mgen import qft_program

H = np.array([[1, 1], [1, -1]]) / np.sqrt(2)
X = np.array([[0, 1], [1, 0]], dtype=complex)


def _fuzz_module():
    spec = importlib.util.spec_from_file_location("fuzz_fused", os.path.join(ROOT, "tools", "fuzz_fused.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def _state(rng, n):
    x = rng.standard_normal(1 << n) + 1j * rng.standard_normal(1 << n)
    return x / np.linalg.norm(x)


def _check(prog, n, x, tile_bits, tol=1e-12, dtype=np.complex128):
    recs, pool = prog.records(0), prog.pool()
    got, sweeps, fft = run_fused_on_host(recs, pool, x, n, tile_bits, dtype=dtype)
    ref = x.astype(np.complex128).copy()
    sv_oracle.run_records(recs, pool, ref, n)
    err = float(np.abs(got - ref).max())
    assert err < tol, (n, tile_bits, len(prog), sweeps, fft, err)
    return sweeps, fft


@pytest.mark.parametrize("n,tile_bits", [(5, 0), (9, 0), (12, 0), (12, 9), (13, 11), (14, 9), (16, 11), (17, 10), (18, 12)])
def test_qft_fused_plan_matches_oracle_and_fft(n, tile_bits):
    rng = np.random.default_rng(100 + n)
    x = _state(rng, n)
    prog = qft_program(n)
    sweeps, fft = _check(prog, n, x, tile_bits)
    got, _, _ = run_fused_on_host(prog.records(0), prog.pool(), x, n, tile_bits)
    assert np.abs(got - np.fft.ifft(x, norm="ortho")).max() < 1e-12
    if n >= 13:
        assert fft >= 1, "QFT sweeps are expected to run in the FFT-shaped image"


def _fft_like_program(rng, n, ordered):
    """Hadamards in random qubit order, each followed by a random subset of controlled phases (control on 1 or 0, with or
    without a phase on the control alone) and plain phase gates.  ``ordered``: phases couple a qubit only to qubits whose
    Hadamard comes later (the QFT discipline, in any qubit order) -- the shape OP_FSTAGE is made for, in all its variants;
    otherwise any partner, which the fusion pass must recognise as NOT FFT-shaped where it matters."""
    prog = Program([Qubit(i) for i in range(n)])
    order = rng.permutation(n)
    done = []
    for t in order:
        t = int(t)
        if ordered or rng.random() < 0.85:
            prog.add_raw([t], 0, 0, H)
        others = [int(b) for b in range(n) if b != t and not (ordered and b in done)]
        for c in others:
            if rng.random() < 0.55:
                d = np.array([1.0, np.exp(1j * rng.uniform(0, 2 * np.pi))])
                val = 1 << c
                if not ordered:
                    if rng.random() < 0.1:
                        d[0] = np.exp(1j * rng.uniform(0, 2 * np.pi))
                    if rng.random() < 0.15:
                        val = 0
                prog.add_raw([t], 1 << c, val, np.diag(d))
        if rng.random() < 0.3:
            prog.add_raw([t], 0, 0, np.diag(np.exp(1j * rng.uniform(0, 2 * np.pi, 2))))
        done.append(t)
    return prog


@pytest.mark.parametrize("ordered", [True, False])
@pytest.mark.parametrize("seed", range(12))
def test_fft_shaped_circuits(seed, ordered):
    rng = np.random.default_rng(seed)
    n = int(rng.integers(9, 17))
    prog = _fft_like_program(rng, n, ordered)
    x = _state(rng, n)
    total_fft = 0
    for tb in (0, 9, 11):
        if tb > n:
            continue
        _, fft = _check(prog, n, x, tb)
        total_fft += fft
    if ordered:
        assert total_fft >= 1, "circuits with the QFT discipline are expected to run in the FFT-shaped image"


def test_hadamard_layers_and_scale():
    """Pure Hadamard layers: butterflies without tables; the owed 1/sqrt2^k scale is settled once per sweep."""
    rng = np.random.default_rng(5)
    for n in (10, 13, 15):
        prog = Program([Qubit(i) for i in range(n)])
        for t in rng.permutation(n):
            prog.add_raw([int(t)], 0, 0, H)
        x = _state(rng, n)
        _, fft = _check(prog, n, x, 11 if n > 11 else 0)
        assert fft >= 1


@pytest.mark.parametrize("seed,tile_bits", [(1, 0), (2, 9), (3, 10), (4, 12), (5, 0), (6, 9)])
def test_random_circuits_every_op_kind(seed, tile_bits):
    fz = _fuzz_module()
    rng = np.random.default_rng(seed)
    for _ in range(10):
        n = int(rng.integers(4, 15))
        n_gates = int(rng.choice([5, 60, 200, 800])) if n <= 11 else int(rng.choice([20, 200]))
        prog = fz.random_program(rng, n, n_gates)
        _check(prog, n, _state(rng, n), tile_bits, tol=1e-11)


def test_complex64_fused_plan():
    rng = np.random.default_rng(9)
    for n in (12, 14):
        x = _state(rng, n)
        _check(qft_program(n), n, x, 11, tol=2e-6, dtype=np.complex64)


def test_hostcheck_rejects_bad_input():
    prog = qft_program(5)
    recs = prog.records(0).copy()
    recs["target_bits"][0][0] = 9
    with pytest.raises(RuntimeError):
        run_fused_on_host(recs, prog.pool(), np.ones(32, dtype=complex), 5)


@pytest.mark.parametrize("n,sweeps", [(24, 4), (28, 4), (30, 5), (34, 5)])
def test_full_size_qft_plans(n, sweeps):
    """Plan shape at BASELINE's sizes (nothing is executed): the whole QFT, final SWAP layer included, takes
    ceil(n / 7) sweeps and every one of them is an FFT-shaped launch."""
    from tests.hostcheck import plan_shape
    prog = qft_program(n)
    got, fft = plan_shape(prog.records(0), prog.pool(), n)
    assert (got, fft) == (sweeps, sweeps)


def test_sharded_qft34_local_segments_are_fft_shaped():
    """The per-rank segments of QFT-34 over 8 GPUs (31 local bits): planner + localisation + fusion, plan only."""
    from quantum_simulations_b200.sharded import localize_records, plan_sharded
    from tests.hostcheck import plan_shape
    n, g = 34, 3
    prog = qft_program(n)
    recs, pool = prog.records(0), prog.pool()
    plan = plan_sharded(recs, n, g)
    locals_ = [s for s in plan.steps if s[0] == "local"]
    assert plan.n_exchanges == 3 and len(locals_) >= 2
    total = fft = 0
    for rank in (0, 5):
        for li, s in enumerate(locals_):
            lrecs, lpool = localize_records(recs, pool, s[1], plan.where_at[li], n - g, rank)
            if len(lrecs):
                a, b = plan_shape(lrecs, lpool, n - g)
                total += a
                fft += b
    assert total > 0 and fft == total


def test_batched_qft_config3_shape_and_semantics():
    """BASELINE config 3 (QFT-24 on 64 states, batch innermost = 6 low index bits no gate touches): the plan at full size
    is 4 FFT-shaped sweeps; at small size every column of the batch comes out as the inverse FFT of its input."""
    from tests.hostcheck import plan_shape
    p = qft_program(24)
    assert plan_shape(p.records(6), p.pool(), 30) == (4, 4)
    n, b = 10, 3
    p = qft_program(n)
    rng = np.random.default_rng(0)
    x = rng.standard_normal((1 << n, 1 << b)) + 1j * rng.standard_normal((1 << n, 1 << b))
    x /= np.linalg.norm(x, axis=0)
    got, _, fft = run_fused_on_host(p.records(b), p.pool(), x.reshape(-1), n + b, 11)
    assert fft >= 1 and np.abs(got.reshape(1 << n, 1 << b) - np.fft.ifft(x, axis=0, norm="ortho")).max() < 1e-12


@pytest.mark.parametrize("carry", [False, True])
@pytest.mark.parametrize("n,tile", [(13, 9), (14, 11), (16, 11)])
def test_requested_end_layout(n, tile, carry):
    """qsv_fuse_to_layout: the plan leaves logical bit b at physical position final_pos[b] -- the layout a sharded run's next
    exchange needs -- whether the permutation rides on a compute sweep or needs a sweep of its own; with carry_exchange the
    last sweep is an out-of-place one in an image whose store can fan out over peers."""
    from oracle import sv_oracle as orc
    from tests.hostcheck import run_fused_to_layout_on_host
    rng = np.random.default_rng(n * 7 + carry)
    for case in range(4):
        prog = qft_program(n) if case == 0 else _fft_like_program(rng, n, ordered=bool(case & 1))
        final = list(range(n))
        if case != 3:                                  # park three random bits in the top positions, keep the rest in order
            victims = sorted(rng.choice(n, size=3, replace=False).tolist())
            rest = [b for b in range(n) if b not in victims]
            for i, b in enumerate(rest):
                final[b] = i
            for i, b in enumerate(victims):
                final[b] = n - 3 + i
        x = rng.standard_normal(1 << n) + 1j * rng.standard_normal(1 << n)
        x /= np.linalg.norm(x)
        got, sweeps, _ = run_fused_to_layout_on_host(prog.records(0), prog.pool(), x, n, final, carry, tile)
        ref = x.copy()
        orc.run_records(prog.records(0), prog.pool(), ref, n)
        idx = np.arange(1 << n)
        phys = np.zeros_like(idx)
        for b in range(n):
            phys |= ((idx >> b) & 1) << final[b]
        want = np.empty_like(ref)
        want[phys] = ref
        assert np.abs(got - want).max() < 1e-12, (case, final)


def test_fuzz_random_programs_with_requested_layouts():
    """Random programs over every op kind (tools/fuzz_fused.py) through the commuting-gate scheduler, the radix-8 generic
    loop and -- every other case -- a requested end layout with or without an exchange-carrying last sweep."""
    import importlib.util
    import os
    from oracle import sv_oracle
    from tests.hostcheck import run_fused_to_layout_on_host
    spec = importlib.util.spec_from_file_location("fuzz_fused", os.path.join(ROOT, "tools", "fuzz_fused.py"))
    fz = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(fz)
    rng = np.random.default_rng(11)
    for case in range(24):
        n = int(rng.integers(10, 15))
        prog = fz.random_program(rng, n, int(rng.choice([5, 30, 120, 300])))
        x = rng.standard_normal(1 << n) + 1j * rng.standard_normal(1 << n)
        x /= np.linalg.norm(x)
        ref = x.copy()
        sv_oracle.run_records(prog.records(0), prog.pool(), ref, n)
        tile = int(rng.choice([9, 10, 11]))
        final = list(range(n))
        if case % 2:
            v = sorted(rng.choice(n, size=2, replace=False).tolist())
            rest = [b for b in range(n) if b not in v]
            for i, b in enumerate(rest):
                final[b] = i
            for i, b in enumerate(v):
                final[b] = n - 2 + i
        got, _, _ = run_fused_to_layout_on_host(prog.records(0), prog.pool(), x, n, final, bool(case % 3 == 0), tile)
        idx = np.arange(1 << n)
        phys = np.zeros_like(idx)
        for b in range(n):
            phys |= ((idx >> b) & 1) << final[b]
        want = np.empty_like(ref)
        want[phys] = ref
        assert np.abs(got - want).max() < 1e-12, (case, n, tile, final)


def test_mixed_target_caps_plan(tmp_path):
    """Mixed target caps: the first sweeps of a plan are held to six targets (radix-8 / TMA image) when that adds no launch
    -- QFT-19 at 2^11-amplitude tiles becomes 6+6+7 instead of 7+7+5 (``QSV_NO_MIXED_CAPS`` restores the balanced plan).
    The plan's shape is read from the debug output of a subprocess; the emulated result is compared with the oracle there."""
    import os
    import subprocess
    import sys
    code = (
        "import sys, numpy as np\n"
        "from quantum_simulations_b200.utils.mgen import qft_program\n"
        "from tests.hostcheck import run_fused_on_host\n"
        "from oracle import sv_oracle as orc\n"
        "n = 19\n"
        "prog = qft_program(n)\n"
        "rng = np.random.default_rng(3)\n"
        "x = rng.standard_normal(1 << n) + 1j * rng.standard_normal(1 << n); x /= np.linalg.norm(x)\n"
        "got, sweeps, fft = run_fused_on_host(prog.records(0), prog.pool(), x, n, tile_bits=11)\n"
        "want = x.copy(); orc.run_records(prog.records(0), prog.pool(), want, n)\n"
        "print('RESULT', sweeps, fft, float(np.abs(got - want).max()))\n")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for switch, n_radix8 in (({}, 2), ({"QSV_NO_MIXED_CAPS": "1"}, 1)):
        env = dict(os.environ, QSV_PLAN_DEBUG="1", PYTHONPATH=root, **switch)
        env.pop("QSV_NO_MIXED_CAPS", None) if not switch else None
        run = subprocess.run([sys.executable, "-c", code], env=env, cwd=root, capture_output=True, text=True, check=True)
        sweeps, fft, err = [ln for ln in run.stdout.splitlines() if ln.startswith("RESULT")][0].split()[1:]
        assert int(sweeps) == 3 and int(fft) == 3 and float(err) < 1e-12
        images = [ln for ln in run.stderr.splitlines() if "sweep: image" in ln][-3:]
        assert sum("fft8-tma" in ln for ln in images) == n_radix8, images
