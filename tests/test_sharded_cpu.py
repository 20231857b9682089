"""Sharded-state logic on CPU: the planner + rank localisation driven by the oracle as local executor.

 * single-process emulation of 2/4/8 ranks (fake exchanger moving halves between in-memory shards);
 * a real two-process ``gloo`` run through ``TorchExchanger`` (the N>1 code path of bench.py / the product).
"""
import os
import random

import numpy as np
import pytest

import quantum_simulations_b200 as qs
from oracle import sv_oracle as orc
from quantum_simulations_b200.program import flatten
from quantum_simulations_b200.sharded import ShardedCircuit, plan_sharded
from quantum_simulations_b200.utils.mgen import qft_program, random_ctrlgate, random_state


class OracleExecutor:
    """Local executor backed by the CPU oracle (test infrastructure)."""

    def __init__(self):
        self.segs = {}

    def prepare(self, key, recs, pool, n_loc):
        self.segs[key] = (recs, pool, n_loc)

    def run_segment(self, key, state, spare):
        recs, pool, n_loc = self.segs[key]
        arr = state.numpy() if hasattr(state, "numpy") else state
        orc.run_records(recs, pool, arr, n_loc)
        return state, spare

    def permute(self, state, spare, n_loc, perm):
        a = state.numpy() if hasattr(state, "numpy") else state
        b = spare.numpy() if hasattr(spare, "numpy") else spare
        idx = np.arange(1 << n_loc)
        y = np.zeros_like(idx)
        for i, p in enumerate(perm):
            y |= ((idx >> i) & 1) << p
        b[y] = a
        return spare, state


class FakeWorld:
    """All ranks in one process: exchange(gbit) swaps halves between partner shards."""

    def __init__(self, shards):
        self.shards = shards

    def exchange_pair(self, gbit):
        done = set()
        for r in range(len(self.shards)):
            p = r ^ (1 << gbit)
            if r in done or p in done:
                continue
            lo, hi = (r, p) if not (r >> gbit) & 1 else (p, r)
            half = self.shards[lo].size // 2
            tmp = self.shards[lo][half:].copy()
            self.shards[lo][half:] = self.shards[hi][:half]
            self.shards[hi][:half] = tmp
            done.update((r, p))

    def exchange_all(self):
        w = len(self.shards)
        chunk = self.shards[0].size // w
        new = [np.empty_like(s) for s in self.shards]
        for r in range(w):
            for c in range(w):
                new[r][c * chunk:(c + 1) * chunk] = self.shards[c][r * chunk:(r + 1) * chunk]
        self.shards[:] = new


def run_emulated(recs, pool, n, world, x, restore=True):
    g = world.bit_length() - 1
    n_loc = n - g
    shards = [x[r << n_loc:(r + 1) << n_loc].copy() for r in range(world)]
    spares = [np.zeros_like(s) for s in shards]
    fw = FakeWorld(shards)

    class NullEx:
        def exchange(self, state, spare, gbit):
            return state

        def exchange_all(self, state, spare):
            return state, spare

    circs = [ShardedCircuit(recs, pool, n, r, world, OracleExecutor(), NullEx(), restore=restore) for r in range(world)]
    plan = circs[0].plan
    assert all(c.plan.steps == plan.steps for c in circs), "plan must be rank independent"
    li = 0
    for s in plan.steps:
        if s[0] == "local":
            for r in range(world):
                fw.shards[r], spares[r] = circs[r].executor.run_segment(li, fw.shards[r], spares[r])
            li += 1
        elif s[0] == "exchange":
            fw.exchange_pair(s[1])
        elif s[0] == "xall":
            fw.exchange_all()
        else:
            for r in range(world):
                fw.shards[r], spares[r] = circs[r].executor.permute(fw.shards[r], spares[r], n_loc, s[1])
    return np.concatenate(fw.shards), plan


@pytest.mark.parametrize("world", [2, 4, 8])
def test_qft_sharded_emulation(world):
    n = 9
    prog = qft_program(n)
    x = random_state(1 << n)
    got, plan = run_emulated(prog.records(), prog.pool(), n, world, x)
    assert np.abs(got - orc.qft_reference(x)).max() < 1e-12
    assert plan.final_where == list(range(n))          # canonical layout restored
    assert plan.n_exchanges >= world.bit_length() - 1   # every global H needs its qubit local once


@pytest.mark.parametrize("world", [2, 4, 8])
def test_qft_sharded_without_layout_restore(world):
    """restore=False: two exchange steps instead of three; the result is the FFT of the input read through the plan's
    bit-position table (logical bit b at physical position final_where[b]) -- QFT's SWAP layer became output ordering."""
    n = 10
    prog = qft_program(n)
    x = random_state(1 << n)
    got, plan = run_emulated(prog.records(), prog.pool(), n, world, x, restore=False)
    assert plan.n_exchanges == 2
    assert sorted(plan.final_where) == list(range(n)) and plan.final_where != list(range(n))
    p = np.arange(1 << n)
    k = np.zeros_like(p)
    for b in range(n):
        k |= ((p >> plan.final_where[b]) & 1) << b
    assert np.abs(got - orc.qft_reference(x)[k]).max() < 1e-12
    full, plan3 = run_emulated(prog.records(), prog.pool(), n, world, x, restore=True)
    assert plan3.n_exchanges == 3 and np.abs(full - orc.qft_reference(x)).max() < 1e-12


@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("seed", [1, 2, 3])
def test_random_circuit_sharded_emulation(world, seed):
    random.seed(seed)
    np.random.seed(seed)
    n = 7
    gates = []
    for _ in range(40):
        k = random.randint(1, 4)
        t = random.randint(1, min(k, 2))
        g = random_ctrlgate(k, t, qnum=n)
        if t == 1 and random.random() < 0.4:
            g = qs.CtrlGate(np.diag(np.exp(1j * np.random.uniform(0, 6.28, 2))), g.controls(), g.qspace)
        elif t == 1 and random.random() < 0.3:
            g = qs.CtrlGate(qs.UnivGate.X, g.controls(), g.qspace)
        gates.append(g)
    reg = [qs.Qubit(i) for i in range(n)]
    prog = flatten(gates, reg)
    x = random_state(1 << n)
    got, plan = run_emulated(prog.records(), prog.pool(), n, world, x)
    want = orc.simulate(gates, reg, x)
    assert np.abs(got - want).max() < 1e-12


def test_controls_and_diagonals_on_global_bits_need_no_exchange():
    n, world = 6, 4
    T, C1, C0 = qs.QType.TARGET, qs.QType.CONTROL1, qs.QType.CONTROL0
    q = [qs.Qubit(i) for i in range(n)]
    gates = [qs.CtrlGate(qs.UnivGate.H, [C1, C0, T], [q[0], q[1], q[4]]),            # controls on both global qubits
             qs.CtrlGate(qs.UnivGate.T, [T, C1], [q[0], q[3]]),                       # diagonal target on a global qubit
             qs.CtrlGate(np.diag([np.exp(0.3j), np.exp(-1.1j)]), [T], [q[1]])]        # uncontrolled diagonal on a global qubit
    prog = flatten(gates, q)
    plan = plan_sharded(prog.records(), n, 2)
    assert plan.n_exchanges == 0
    x = random_state(1 << n)
    got, _ = run_emulated(prog.records(), prog.pool(), n, world, x)
    assert np.abs(got - orc.simulate(gates, q, x)).max() < 1e-13
    # an X on a global qubit is NOT free
    plan2 = plan_sharded(flatten([qs.CtrlGate(qs.UnivGate.X, [T], [q[0]])], q).records(), n, 2, restore=False)
    assert plan2.n_exchanges == 1


def _gloo_worker(rank, world, port, n, out_dir):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from quantum_simulations_b200.sharded import TorchExchanger
    np.random.seed(7)
    x = random_state(1 << n)
    prog = qft_program(n)
    n_loc = n - (world.bit_length() - 1)
    state = torch.from_numpy(x[rank << n_loc:(rank + 1) << n_loc].copy())
    spare = torch.zeros_like(state)
    circ = ShardedCircuit(prog.records(), prog.pool(), n, rank, world, OracleExecutor(), TorchExchanger(rank, world))
    state, spare = circ.apply(state, spare)
    np.save(os.path.join(out_dir, f"shard{rank}.npy"), state.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_two_process_gloo_half_state_swap(tmp_path):
    import torch.multiprocessing as mp
    n, world = 8, 2
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_gloo_worker, args=(world, port, n, str(tmp_path)), nprocs=world, join=True)
    np.random.seed(7)
    x = random_state(1 << n)
    got = np.concatenate([np.load(tmp_path / f"shard{r}.npy") for r in range(world)])
    assert np.abs(got - orc.qft_reference(x)).max() < 1e-12


class FusedHostExecutor(OracleExecutor):
    """Local executor that runs every segment through the REAL fusion pass, interpreted on the host
    (tests/hostcheck.py): the per-rank records of ``localize_records`` go through the same planner -- tile choice,
    phase polynomials, SWAP relabels, FFT-shaped launches -- that the GPU executor feeds to k_sweep."""

    def __init__(self):
        super().__init__()
        self.fft_sweeps = 0

    def run_segment(self, key, state, spare):
        from tests.hostcheck import run_fused_on_host
        recs, pool, n_loc = self.segs[key]
        if len(recs) == 0:
            return state, spare
        out, _, fft = run_fused_on_host(recs, pool, state, n_loc)
        self.fft_sweeps += fft
        state[:] = out
        return state, spare


class LayoutHostExecutor(FusedHostExecutor):
    """... and accepts a requested end layout like GpuExecutor.prepare: the park permutation of the exchange that follows
    rides on the segment's last sweep (qsv_fuse_to_layout, interpreted on the host)."""
    fuse = True

    def __init__(self):
        super().__init__()
        self.final = {}

    def prepare(self, key, recs, pool, n_loc, final_pos=None, carry_exchange=False):
        super().prepare(key, recs, pool, n_loc)
        if final_pos is not None:
            self.final[key] = list(final_pos)
        return 1 if final_pos is not None else 0

    def agree_on_buffers(self, keys, flips, all_reduce_min):
        pass

    def run_segment(self, key, state, spare):
        if key not in self.final:
            return super().run_segment(key, state, spare)
        from tests.hostcheck import run_fused_to_layout_on_host
        recs, pool, n_loc = self.segs[key]
        out, _, _ = run_fused_to_layout_on_host(recs, pool, state, n_loc, self.final[key])
        state[:] = out
        return state, spare


@pytest.mark.parametrize("world,n", [(2, 13), (4, 14), (8, 15)])
def test_park_permutation_rides_on_the_segments_last_sweep(world, n):
    """With an executor that can end a segment in a requested layout, ShardedCircuit drops the gate-free permuting sweep
    in front of an exchange: same result, one plan step fewer per merged exchange."""
    prog = qft_program(n)
    recs, pool = prog.records(), prog.pool()
    x = random_state(1 << n)
    g = world.bit_length() - 1
    n_loc = n - g
    fw = FakeWorld([x[r << n_loc:(r + 1) << n_loc].copy() for r in range(world)])
    spares = [np.zeros_like(s) for s in fw.shards]

    class NullEx:
        def exchange_all(self, state, spare):
            return state, spare

    circs = [ShardedCircuit(recs, pool, n, r, world, LayoutHostExecutor(), NullEx(), restore=False) for r in range(world)]
    assert circs[0]._merged, "no exchange was merged"
    steps = circs[0].plan.steps
    k = li = 0
    permutes_run = 0
    while k < len(steps):
        s = steps[k]
        merged = circs[0]._merged.get(k) if s[0] == "local" else None
        k += 1
        if s[0] == "local":
            for r in range(world):
                fw.shards[r], spares[r] = circs[r].executor.run_segment(li, fw.shards[r], spares[r])
            li += 1
            if merged:
                k += merged[0]
                if merged[0] == 1 and steps[k - 1][0] == "permute":
                    pass
        elif s[0] == "xall":
            fw.exchange_all()
        else:
            permutes_run += 1
            for r in range(world):
                fw.shards[r], spares[r] = circs[r].executor.permute(fw.shards[r], spares[r], n_loc, s[1])
    got = np.concatenate(fw.shards)
    where = circs[0].final_where
    p = np.arange(1 << n)
    kk = np.zeros_like(p)
    for b in range(n):
        kk |= ((p >> where[b]) & 1) << b
    assert np.abs(got - orc.qft_reference(x)[kk]).max() < 1e-12
    assert permutes_run < sum(1 for s in steps if s[0] == "permute")


@pytest.mark.parametrize("world,n", [(2, 13), (4, 14), (8, 15), (8, 16)])
def test_qft_sharded_through_the_fusion_pass(world, n):
    """Sharding planner + rank localisation + fusion pass end to end on CPU: every rank's segments are planned and
    interpreted by libqsv's host check, exchanges are emulated in memory, the result is the FFT of the input."""
    prog = qft_program(n)
    recs, pool = prog.records(), prog.pool()
    x = random_state(1 << n)
    g = world.bit_length() - 1
    n_loc = n - g
    shards = [x[r << n_loc:(r + 1) << n_loc].copy() for r in range(world)]
    spares = [np.zeros_like(s) for s in shards]
    fw = FakeWorld(shards)

    class NullEx:
        def exchange(self, state, spare, gbit):
            return state

        def exchange_all(self, state, spare):
            return state, spare

    circs = [ShardedCircuit(recs, pool, n, r, world, FusedHostExecutor(), NullEx()) for r in range(world)]
    plan = circs[0].plan
    li = 0
    for s in plan.steps:
        if s[0] == "local":
            for r in range(world):
                fw.shards[r], spares[r] = circs[r].executor.run_segment(li, fw.shards[r], spares[r])
            li += 1
        elif s[0] == "exchange":
            fw.exchange_pair(s[1])
        elif s[0] == "xall":
            fw.exchange_all()
        else:
            for r in range(world):
                fw.shards[r], spares[r] = circs[r].executor.permute(fw.shards[r], spares[r], n_loc, s[1])
    got = np.concatenate(fw.shards)
    assert np.abs(got - np.fft.ifft(x, norm="ortho")).max() < 1e-12
    assert sum(c.executor.fft_sweeps for c in circs) >= world, "local QFT segments are expected to be FFT-shaped launches"


@pytest.mark.parametrize("world,seed", [(2, 4), (4, 5), (8, 6)])
def test_random_circuit_sharded_through_the_fusion_pass(world, seed):
    """Random one-target circuits (controls anywhere, diagonal gates, X/CNOT/Toffoli, SWAP triples) sharded over 2/4/8
    emulated ranks, every rank's segments planned and interpreted by the host check of the fusion pass."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("fuzz_fused", os.path.join(os.path.dirname(os.path.dirname(
        os.path.abspath(__file__))), "tools", "fuzz_fused.py"))
    fz = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(fz)
    rng = np.random.default_rng(seed)
    n = 10
    prog = fz.random_program(rng, n, 120)
    recs, pool = prog.records(), prog.pool()
    x = random_state(1 << n)
    g = world.bit_length() - 1
    n_loc = n - g
    shards = [x[r << n_loc:(r + 1) << n_loc].copy() for r in range(world)]
    spares = [np.zeros_like(s) for s in shards]
    fw = FakeWorld(shards)

    class NullEx:
        def exchange(self, state, spare, gbit):
            return state

        def exchange_all(self, state, spare):
            return state, spare

    circs = [ShardedCircuit(recs, pool, n, r, world, FusedHostExecutor(), NullEx()) for r in range(world)]
    li = 0
    for s in circs[0].plan.steps:
        if s[0] == "local":
            for r in range(world):
                fw.shards[r], spares[r] = circs[r].executor.run_segment(li, fw.shards[r], spares[r])
            li += 1
        elif s[0] == "exchange":
            fw.exchange_pair(s[1])
        elif s[0] == "xall":
            fw.exchange_all()
        else:
            for r in range(world):
                fw.shards[r], spares[r] = circs[r].executor.permute(fw.shards[r], spares[r], n_loc, s[1])
    want = x.copy()
    orc.run_records(recs, pool, want, n)
    assert np.abs(np.concatenate(fw.shards) - want).max() < 1e-11
