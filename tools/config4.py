#!/usr/bin/env python
"""BASELINE config 4 at full size: random 2^10 x 2^10 unitary -> SINGLET stream (4.7 M fully controlled gates)
-> full unitary reconstruction on the 1024 basis columns, checked against the input on the device.

  python tools/config4.py [--n 10]                                   one GPU, all 1024 columns
  python -m torch.distributed.run --nproc-per-node 8 ... tools/config4.py    8 GPUs, 128 columns each

Under torchrun every rank rebuilds ITS OWN block of columns -- columns are independent, so there is no collective on
the data path; the ranks only reduce their timings and error figures for the one JSON line rank 0 prints.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=10)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--as-rank", default="", help="R/W: on ONE GPU, do what rank R of W would do (its block of columns only)")
    args = ap.parse_args()
    import torch
    from quantum_simulations_b200.engine import Engine
    from quantum_simulations_b200.utils.mgen import random_unitary
    from quantum_simulations_b200.utils.producers import singlet_records_fast, two_level_stream_scan
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    solo = None
    if args.as_rank:
        assert world == 1
        r_, w_ = (int(x) for x in args.as_rank.split("/"))
        solo = (r_, w_)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local_rank}"))
    n = args.n
    N = 1 << n
    if solo:
        rank, world = solo
    assert N % world == 0
    cols = N // world
    cb = cols.bit_length() - 1
    np.random.seed(42)
    u = random_unitary(N)                      # (every rank draws the same matrix)
    t0 = time.time()
    stream = two_level_stream_scan(u, f"cuda:{local_rank}")     # Givens QR one column at a time, on the GPU
    torch.cuda.synchronize()
    t1 = time.time()
    recs, pool = singlet_records_fast(stream, n, batch_bits=cb)
    t2 = time.time()
    eng = Engine(local_rank, np.complex128)
    st = eng.empty(N * cols)
    expected = torch.from_numpy(np.ascontiguousarray(u[:, rank * cols:(rank + 1) * cols]).reshape(-1)).cuda()
    eng.upload_records(recs, pool, n + cb, cb)
    t3 = time.time()
    sweeps = eng.fuse()
    t4 = time.time()
    times = []
    for i in range(args.steps + 1):
        eng.init_basis_columns(st, n + cb, cb, rank * cols, 0)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        res = eng.run(st, None)
        b.record()
        torch.cuda.synchronize()
        if i:
            times.append(a.elapsed_time(b))
    mx, bad = eng.compare(res, expected, rtol=1e-5, atol=1e-8)        # K8: np.allclose(v, u) on the device
    ms = float(np.median(times))
    if world > 1 and not solo:
        t = torch.tensor([ms, mx, float(bad)], device=f"cuda:{local_rank}", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, mx, bad = float(t[0]), float(t[1]), int(t[2])
    if rank == 0 or solo:
        flops = 28.0 * len(stream) * N          # 28 real flops per 2x2 complex butterfly, per column, whole job
        print(json.dumps({"config": f"random 2^{n} unitary, SINGLET stream, {N}-column reconstruction, complex128, {world} B200" + (f" [rank {rank} alone on one GPU]" if solo else "")
                                    + (f" ({cols} columns each, no collective)" if world > 1 else ""),
                          "gates": int(len(recs)), "two_level_ops": len(stream), "fused_passes": sweeps, "n_gpus": world, "ms": ms,
                          "gates_per_sec": len(recs) / (ms / 1e3), "fp64_GFLOPs": flops / (ms / 1e3) / 1e9,
                          "max_abs_err_vs_input": mx, "allclose_violations": bad,
                          "producer_s": {"two_level_stream_scan (GPU)": round(t1 - t0, 2), "singlet_records_fast (host)": round(t2 - t1, 2),
                                     "plan (fold relabels, level schedule, upload)": round(t4 - t3, 2)}}))
    if world > 1 and not solo:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
