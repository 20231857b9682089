"""Sharded state vector over >= 2 GPUs (NCCL half-state swaps); skipped on single-GPU boxes."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


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
    if mode == "fused":
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
    prog = qft_program(n)
    circ = ShardedCircuit(prog.records(), prog.pool(), n, rank, world, GpuExecutor(rank), exch)
    for _ in range(2):          # twice: buffers trade roles between applications
        state, spare = circ.apply(state, spare)
    x = torch.fft.ifft(x.cuda(), norm="ortho").cpu()
    torch.cuda.synchronize()
    want = torch.fft.ifft(x.cuda(), norm="ortho")[rank << n_loc:(rank + 1) << n_loc]
    err = float((state - want).abs().max())
    fused = circ.executor.total_stats().get("fused_sweeps", 0)
    with open(os.path.join(out_dir, f"r{rank}.txt"), "w") as f:
        f.write(f"{err} {fused} {circ.plan.n_exchanges}")
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("mode", ["nccl", "fused"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_qft_sharded(world, mode, tmp_path):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp
    n = 22
    port = 29600 + (os.getpid() % 2000) + world
    port += 10 * (mode == "fused")
    mp.spawn(_worker, args=(world, port, n, str(tmp_path), mode), nprocs=world, join=True)
    for r in range(world):
        txt = open(tmp_path / f"r{r}.txt").read()
        if txt.startswith("skip"):
            pytest.skip(txt)
        err, fused, nex = txt.split()
        assert float(err) < 1e-10, f"rank {r}: max abs err {err}"
        assert int(fused) > 0 and int(nex) >= world.bit_length() - 1
