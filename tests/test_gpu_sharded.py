"""Sharded state vector over >= 2 GPUs (NCCL half-state swaps); skipped on single-GPU boxes."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _where_index(n, where):
    """physical index -> logical index for a bit-position table (logical bit b sits at physical position where[b])."""
    import torch
    p = torch.arange(1 << n, dtype=torch.int64)
    k = torch.zeros_like(p)
    for b in range(n):
        k |= ((p >> int(where[b])) & 1) << b
    return k


def _worker(rank, world, port, n, out_dir, mode="nccl"):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device(f"cuda:{rank}"))
    from quantum_simulations_b200.sharded import GpuExecutor, ShardedCircuit, TorchExchanger
    from quantum_simulations_b200.utils.mgen import qft_program
    g = torch.Generator().manual_seed(5)
    x = torch.randn(1 << n, dtype=torch.complex128, generator=g)
    x /= torch.linalg.vector_norm(x)
    n_loc = n - (world.bit_length() - 1)
    if mode.startswith("fused"):
        from quantum_simulations_b200.sharded import SymmExchanger
        try:
            exch = SymmExchanger(rank, world, 1 << n_loc, torch.device(f"cuda:{rank}"))
        except Exception as e:  # torch symmetric memory not usable on this box
            with open(os.path.join(out_dir, f"r{rank}.txt"), "w") as f:
                f.write(f"skip {type(e).__name__}: {e}")
            dist.destroy_process_group()
            return
        state, spare = exch.buffers()
        state.copy_(x[rank << n_loc:(rank + 1) << n_loc])
    else:
        exch = TorchExchanger(rank, world)
        state = x[rank << n_loc:(rank + 1) << n_loc].clone().cuda()
        spare = torch.empty_like(state)
    def all_reduce_min(values):
        t = torch.tensor(values if values else [0], device=f"cuda:{rank}", dtype=torch.int32)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        return [int(v) for v in t.tolist()][:len(values)]

    if mode.endswith("cswap"):
        # three Toffolis = a SWAP of two local qubits controlled by the GLOBAL qubit: the ranks whose control is not
        # satisfied drop the gates, the others relabel the SWAP (an out-of-place sweep) -- plans of different parity
        # per rank, which the fused exchange that follows must survive (agree_on_buffers)
        import quantum_simulations_b200 as qs
        from quantum_simulations_b200.program import Program
        T, C1 = qs.QType.TARGET, qs.QType.CONTROL1
        prog = Program([qs.Qubit(i) for i in range(n)])
        hm, xm = np.array(qs.UnivGate.H), np.array(qs.UnivGate.X)
        top, a, b = n - 1, 3, n - 5
        for q in (a, b, 5):
            prog.add_raw([q], 0, 0, hm)
        prog.add_raw([b], (1 << top) | (1 << a), (1 << top) | (1 << a), xm)
        prog.add_raw([a], (1 << top) | (1 << b), (1 << top) | (1 << b), xm)
        prog.add_raw([b], (1 << top) | (1 << a), (1 << top) | (1 << a), xm)
        prog.add_raw([top], 0, 0, hm)          # global target: forces the exchange right after the uneven segment
        prog.add_raw([a], 1 << top, 1 << top, hm)
        from oracle import sv_oracle
        ref = x.numpy().copy()
        sv_oracle.run_records(prog.records(), prog.pool(), ref, n)
        circ = ShardedCircuit(prog.records(), prog.pool(), n, rank, world, GpuExecutor(rank), exch, all_reduce_min=all_reduce_min)
        state, spare = circ.apply(state, spare)
        torch.cuda.synchronize()
        want = torch.from_numpy(ref)[rank << n_loc:(rank + 1) << n_loc].cuda()
    else:
        prog = qft_program(n)
        restore = not mode.endswith("norestore")
        circ = ShardedCircuit(prog.records(), prog.pool(), n, rank, world, GpuExecutor(rank), exch, restore=restore,
                              all_reduce_min=all_reduce_min)
        if restore:
            for _ in range(2):          # twice: buffers trade roles between applications
                state, spare = circ.apply(state, spare)
            x = torch.fft.ifft(x.cuda(), norm="ortho").cpu()
            torch.cuda.synchronize()
            want = torch.fft.ifft(x.cuda(), norm="ortho")[rank << n_loc:(rank + 1) << n_loc]
        else:
            # the result stays in the layout the plan ends in: the checker reads it through the bit-position table
            state, spare = circ.apply(state, spare)
            torch.cuda.synchronize()
            full = torch.fft.ifft(x.cuda(), norm="ortho").cpu()
            want = full[_where_index(n, circ.final_where)][rank << n_loc:(rank + 1) << n_loc].cuda()
            assert circ.plan.n_exchanges == 2, circ.plan.n_exchanges
    err = float((state - want).abs().max())
    fused = circ.executor.total_stats().get("fused_sweeps", 0)
    with open(os.path.join(out_dir, f"r{rank}.txt"), "w") as f:
        f.write(f"{err} {fused} {circ.plan.n_exchanges}")
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("mode", ["nccl", "fused", "nccl_norestore", "fused_norestore", "fused_cswap", "nccl_cswap"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_qft_sharded(world, mode, tmp_path):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp
    n = 22
    port = 29600 + (os.getpid() % 2000) + world
    port += 10 * ["nccl", "fused", "nccl_norestore", "fused_norestore", "fused_cswap", "nccl_cswap"].index(mode)
    mp.spawn(_worker, args=(world, port, n, str(tmp_path), mode), nprocs=world, join=True)
    for r in range(world):
        txt = open(tmp_path / f"r{r}.txt").read()
        if txt.startswith("skip"):
            pytest.skip(txt)
        err, fused, nex = txt.split()
        assert float(err) < 1e-10, f"rank {r}: max abs err {err}"
        assert int(fused) > 0 and int(nex) >= 1
