"""The sharded path's exchange kernels on ONE GPU: every rank is a thread of this process with its shard on cuda:0, and the
"peer" buffers a rank stores into are the other threads' buffers on the same device.  The kernels cannot tell (a peer is an
address), so the fused permute+exchange kernel (``qsv_permute_exchange``) and the compute sweep whose store is the exchange
(``qsv_run_exchange``, the cp.async image and -- ``QSV_CARRY8=1`` -- the radix-8 image with thread stores) run on a single-GPU box, where
``tests/test_gpu_sharded.py`` (real NVLink peers, NCCL) is skipped.  Checked against ``torch.fft.ifft`` and the oracle."""
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


class ThreadPeers:
    """What the ranks share: every rank's two buffers, a barrier, a scratch list for the min-reduction."""

    def __init__(self, world, n_loc, device):
        import torch
        self.world = world
        self.bufs = [[torch.empty(1 << n_loc, dtype=torch.complex128, device=device) for _ in range(2)] for _ in range(world)]
        self.barrier = threading.Barrier(world)
        self.scratch = [None] * world

    def all_reduce_min(self, rank, values):
        self.scratch[rank] = list(values)
        self.barrier.wait()
        out = [min(self.scratch[r][i] for r in range(self.world)) for i in range(len(values))]
        self.barrier.wait()
        return out


class ThreadExchanger:
    """SymmExchanger's interface (``sharded.py``) over same-device buffers: the signal-pad barriers become a thread barrier
    around a device synchronisation."""

    def __init__(self, rank, shared):
        self.rank, self.world, self.shared = rank, shared.world, shared
        self.g = self.world.bit_length() - 1
        self.bytes_sent = 0

    def _peers_of(self, spare):
        mine = self.shared.bufs[self.rank]
        i = 0 if spare.data_ptr() == mine[0].data_ptr() else 1
        assert spare.data_ptr() == mine[i].data_ptr()
        return [int(self.shared.bufs[r][i].data_ptr()) for r in range(self.world)]

    def _sync(self):
        import torch
        torch.cuda.synchronize()
        self.shared.barrier.wait()

    def permute_exchange(self, executor, state, spare, n_loc, perm):
        import ctypes as C
        from quantum_simulations_b200._lib import check
        eng = executor.any_engine()
        arr = (C.c_void_p * self.world)(*self._peers_of(spare))
        p = np.ascontiguousarray(perm, dtype=np.int32)
        self._sync()
        check(eng.lib.qsv_permute_exchange(eng._ctx, C.c_void_p(state.data_ptr()), arr, self.g, self.rank, n_loc,
                                           p.ctypes.data_as(C.c_void_p), eng._stream()), eng._ctx)
        self._sync()
        return spare, state

    def segment_exchange(self, executor, key, state, spare):
        eng, _ = executor.engines[key]
        peers = self._peers_of(spare)
        self._sync()
        eng.run_exchange(state, spare, peers, self.g, self.rank)
        self._sync()
        return spare, state


def _where_index(n, where):
    import torch
    p = torch.arange(1 << n, dtype=torch.int64)
    k = torch.zeros_like(p)
    for b in range(n):
        k |= ((p >> int(where[b])) & 1) << b
    return k


def _run_ranks(world, n, make_program, restore, env=None):
    """Runs the sharded circuit with one thread per rank; returns (result gathered in rank order, circuits)."""
    import torch
    from quantum_simulations_b200.sharded import GpuExecutor, ShardedCircuit
    g = world.bit_length() - 1
    n_loc = n - g
    gen = torch.Generator().manual_seed(11)
    x = torch.randn(1 << n, dtype=torch.complex128, generator=gen)
    x /= torch.linalg.vector_norm(x)
    shared = ThreadPeers(world, n_loc, "cuda:0")
    prog = make_program(n)
    recs, pool = prog.records(), prog.pool()
    circs, results, errors = [None] * world, [None] * world, []

    def rank_main(rank):
        try:
            torch.cuda.set_device(0)
            state, spare = shared.bufs[rank]
            state.copy_(x[rank << n_loc:(rank + 1) << n_loc])
            circ = ShardedCircuit(recs, pool, n, rank, world, GpuExecutor(0), ThreadExchanger(rank, shared), restore=restore,
                                  all_reduce_min=lambda v: shared.all_reduce_min(rank, v))
            circs[rank] = circ
            state, spare = circ.apply(state, spare)
            torch.cuda.synchronize()
            results[rank] = state.cpu()
        except BaseException as e:  # a dead rank must not leave the others at the barrier
            errors.append((rank, e))
            shared.barrier.abort()

    threads = [threading.Thread(target=rank_main, args=(r,)) for r in range(world)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    real = [e for e in errors if not isinstance(e[1], threading.BrokenBarrierError)]
    if real:
        raise real[0][1]
    assert not errors, errors
    return x, torch.cat(results), circs


@pytest.mark.parametrize("restore", [False, True])
@pytest.mark.parametrize("world,n", [(2, 20), (4, 21), (8, 22), (2, 23)])
def test_qft_ranks_as_threads(world, n, restore):
    import torch
    from quantum_simulations_b200.utils.mgen import qft_program
    x, got, circs = _run_ranks(world, n, qft_program, restore)
    full = torch.fft.ifft(x.cuda(), norm="ortho").cpu()
    want = full if restore else full[_where_index(n, circs[0].final_where)]
    assert float((got - want).abs().max()) < 1e-10
    c = circs[0]
    assert c.plan.n_exchanges == (3 if restore else 2)
    assert c.carrying_sweeps > 0, "an exchange should ride on a compute sweep"
    assert all(ci.final_where == c.final_where for ci in circs)


def test_carrying_sweep_images_agree(monkeypatch):
    """The exchange-carrying sweep in the radix-8 image (thread stores of the output-ordered tile, ``QSV_CARRY8=1`` -- read
    when the library first plans such a sweep, hence a subprocess for that run) against the default cp.async image."""
    import subprocess
    import sys
    import torch
    from quantum_simulations_b200.utils.mgen import qft_program
    x, got, _ = _run_ranks(4, 21, qft_program, False)
    code = ("import sys, torch; sys.path.insert(0, '.'); from tests.test_gpu_sharded_one_device import _run_ranks\n"
            "from quantum_simulations_b200.utils.mgen import qft_program\n"
            "x, got, _ = _run_ranks(4, 21, qft_program, False); torch.save(got, sys.argv[1])")
    import os
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "got.pt")
        # (QSV_NO_MIXED_CAPS: the balanced plan, whose carrying sweep has five targets -- eligible for the radix-8 image)
        env = dict(os.environ, QSV_CARRY8="1", QSV_NO_MIXED_CAPS="1", QSV_PLAN_DEBUG="1")
        run = subprocess.run([sys.executable, "-c", code, path], check=True, env=env, capture_output=True, text=True,
                             cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        other = torch.load(path)
    # (the radix-8 image did carry it: a permuting store of that image appears in the plans)
    assert any("fft8-tma" in ln and "permuting store" in ln for ln in (run.stderr + run.stdout).splitlines())
    assert float((got - other).abs().max()) < 1e-14


def test_uneven_rank_plans_as_threads():
    """A SWAP of two local qubits controlled by the GLOBAL qubit: the ranks whose control is not satisfied drop the gates,
    the others relabel -- per-rank plans of different buffer parity before a fused exchange (``agree_on_buffers``)."""
    import torch
    import quantum_simulations_b200 as qs
    from oracle import sv_oracle
    from quantum_simulations_b200.program import Program

    def make(n):
        prog = Program([qs.Qubit(i) for i in range(n)])
        hm, xm = np.array(qs.UnivGate.H), np.array(qs.UnivGate.X)
        top, a, b = n - 1, 3, n - 5
        for q in (a, b, 5):
            prog.add_raw([q], 0, 0, hm)
        prog.add_raw([b], (1 << top) | (1 << a), (1 << top) | (1 << a), xm)
        prog.add_raw([a], (1 << top) | (1 << b), (1 << top) | (1 << b), xm)
        prog.add_raw([b], (1 << top) | (1 << a), (1 << top) | (1 << a), xm)
        prog.add_raw([top], 0, 0, hm)
        prog.add_raw([a], 1 << top, 1 << top, hm)
        return prog

    n = 20
    x, got, _ = _run_ranks(2, n, make, True)
    ref = x.numpy().copy()
    prog = make(n)
    sv_oracle.run_records(prog.records(), prog.pool(), ref, n)
    assert np.abs(got.numpy() - ref).max() < 1e-12
