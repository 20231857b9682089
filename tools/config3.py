#!/usr/bin/env python
"""BASELINE config 3 at full size: QFT-24 gate circuit on a batch of 64 random input states, complex128, 1 B200
(`qft_benchmark`: 336 gates, state array [2^24][64] batch-innermost = 16 GiB), checked on the device against the
inverse FFT of the inputs (fft_matrix(n) @ x == ifft(x, norm="ortho"), quompiler/utils/mgen.py:128-137).

  python tools/config3.py [--qubits 24] [--batch 64] [--steps 5]

cuFFT is the CHECKER here (torch.fft on the same device), never the thing measured.
"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--qubits", type=int, default=24)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--steps", type=int, default=5)
    args = ap.parse_args()
    import torch
    from quantum_simulations_b200.engine import Engine
    from quantum_simulations_b200.utils.mgen import qft_program
    n, batch = args.qubits, args.batch
    b = (batch - 1).bit_length()
    assert 1 << b == batch, "batch must be a power of two (it becomes the low index bits)"
    prog = qft_program(n)
    recs, pool = prog.records(b), prog.pool()
    eng = Engine(0, np.complex128)
    amps = 1 << (n + b)
    x = eng.empty(amps)
    g = torch.Generator(device="cuda:0")
    g.manual_seed(42)
    view = torch.view_as_real(x)
    for s in range(0, amps, 1 << 26):
        view[s:s + (1 << 26)].normal_(generator=g)
    cols = x.view(1 << n, batch)
    cols /= torch.linalg.vector_norm(cols, dim=0, keepdim=True)          # every input state L2-normalised (mgen.random_state)
    # reference result of 4 of the columns (cuFFT on a copy), kept small so that the 16 GiB state has room for its spare
    pick = sorted({0, 1, batch // 2, batch - 1})
    want = torch.fft.ifft(cols[:, pick].clone(), dim=0, norm="ortho")
    eng.upload_records(recs, pool, n + b, b)
    sweeps = eng.fuse()
    spare = eng.empty(amps)
    cur, other = x, spare
    times = []
    for i in range(args.steps + 1):
        a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        res = eng.run(cur, other)
        e.record()
        torch.cuda.synchronize()
        if i == 0:
            err = float((res.view(1 << n, batch)[:, pick] - want).abs().max())
            norms = torch.linalg.vector_norm(res.view(1 << n, batch), dim=0)
            norm_dev = float((norms - 1).abs().max())
        else:
            times.append(a.elapsed_time(e))
        if res is not cur:
            cur, other = res, cur
    ms = float(np.median(times))
    st_bytes = 2 * amps * 16
    print(json.dumps({"config": f"QFT-{n} gate circuit, batch of {batch} random states, complex128, 1 B200",
                      "gates": int(len(recs)), "state_GiB": amps * 16 / 2**30, "fused_sweeps": sweeps, "ms_per_step": ms,
                      "gates_per_sec": len(recs) / (ms / 1e3), "state_gates_per_sec": batch * len(recs) / (ms / 1e3),
                      "amp_GBps": sweeps * st_bytes / (ms / 1e3) / 1e9, "max_abs_err_vs_ifft": err, "max_norm_deviation": norm_dev}))


if __name__ == "__main__":
    main()
