#!/usr/bin/env python
"""BASELINE config 4 at full size: random 2^10 x 2^10 unitary -> SINGLET stream (4.7 M fully controlled gates)
-> full unitary reconstruction on the 1024 basis columns, checked against the input on the device.

  python tools/config4.py [--n 10] [--shards 8]

--shards S additionally rebuilds the unitary as S independent column blocks (what S GPUs would each do, with no
communication) and checks the concatenation.
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
    ap.add_argument("--shards", type=int, default=8)
    ap.add_argument("--steps", type=int, default=5)
    args = ap.parse_args()
    import torch
    from quantum_simulations_b200.engine import Engine
    from quantum_simulations_b200.utils.mgen import random_unitary
    from quantum_simulations_b200.utils.producers import singlet_records, two_level_stream
    n = args.n
    N = 1 << n
    np.random.seed(42)
    u = random_unitary(N)
    t0 = time.time()
    stream = two_level_stream(u)
    t1 = time.time()
    recs, pool = singlet_records(stream, n, batch_bits=n)
    t2 = time.time()
    eng = Engine(0, np.complex128)
    st = eng.empty(N * N)
    expected = torch.from_numpy(u.reshape(-1)).cuda()
    eng.upload_records(recs, pool, 2 * n, n)
    sweeps = eng.fuse()
    times = []
    for i in range(args.steps + 1):
        eng.init_basis_columns(st, 2 * n, n, 0, 0)
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
    out = {"config": f"random 2^{n} unitary, SINGLET stream, {N}-column reconstruction, complex128, 1 B200",
           "gates": int(len(recs)), "two_level_ops": len(stream), "fused_passes": sweeps, "ms": ms,
           "gates_per_sec": len(recs) / (ms / 1e3), "max_abs_err_vs_input": mx, "allclose_violations": bad,
           "producer_s": {"two_level_stream": round(t1 - t0, 1), "singlet_records": round(t2 - t1, 1)}}
    # column sharding: S independent blocks of N/S columns (no communication between them)
    S = args.shards
    if S > 1 and N % S == 0:
        cb = (N // S).bit_length() - 1
        recs_s, pool_s = singlet_records(stream, n, batch_bits=cb)
        eng.upload_records(recs_s, pool_s, n + cb, cb)
        eng.fuse()
        blocks = []
        worst = 0.0
        for s in range(S):
            blk = eng.empty(N * (N // S))
            eng.init_basis_columns(blk, n + cb, cb, s * (N // S), 0)
            r = eng.run(blk, None).view(N, N // S)
            worst = max(worst, float((r - expected.view(N, N)[:, s * (N // S):(s + 1) * (N // S)]).abs().max()))
        out["column_shards"] = {"shards": S, "columns_each": N // S, "max_abs_err": worst}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
