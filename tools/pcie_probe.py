#!/usr/bin/env python
"""What the host link gives on this box: pinned H2D alone, D2H alone and both at once (GB/s per direction), for the
buffer sizes bench.py's e2e leg moves.  Says how far the e2e figure is from the link.

  python tools/pcie_probe.py [--gib 8] > gpurun_out/pcie_probe.json
"""
import argparse
import json
import os

import torch


def timed(fn, streams, reps):
    torch.cuda.synchronize()
    t0 = torch.cuda.Event(enable_timing=True)
    t1 = torch.cuda.Event(enable_timing=True)
    t0.record()
    for s in streams:
        s.wait_event(t0)
    for _ in range(reps):
        fn()
    for s in streams:
        torch.cuda.current_stream().wait_stream(s)
    t1.record()
    torch.cuda.synchronize()
    return t0.elapsed_time(t1) / reps / 1e3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gib", type=float, default=8.0)
    ap.add_argument("--reps", type=int, default=3)
    args = ap.parse_args()
    n = int(args.gib * (1 << 30)) // 16
    nbytes = n * 16
    dev_a = torch.empty(n, dtype=torch.complex128, device="cuda:0")
    dev_b = torch.empty(n, dtype=torch.complex128, device="cuda:0")
    host_in = torch.empty(n, dtype=torch.complex128, pin_memory=True)
    host_out = torch.empty(n, dtype=torch.complex128, pin_memory=True)
    host_in.view(torch.float64).fill_(1.0)
    host_out.view(torch.float64).fill_(0.0)
    s_up, s_down = torch.cuda.Stream(), torch.cuda.Stream()
    out = {"bytes": nbytes, "cpus": len(os.sched_getaffinity(0))}

    def up():
        with torch.cuda.stream(s_up):
            dev_a.copy_(host_in, non_blocking=True)

    def down():
        with torch.cuda.stream(s_down):
            host_out.copy_(dev_b, non_blocking=True)

    def both():
        up()
        down()

    up(); down(); torch.cuda.synchronize()
    out["h2d_alone_GBps"] = nbytes / timed(up, [s_up], args.reps) / 1e9
    out["d2h_alone_GBps"] = nbytes / timed(down, [s_down], args.reps) / 1e9
    out["both_GBps_per_direction"] = nbytes / timed(both, [s_up, s_down], args.reps) / 1e9

    # the same in pieces of 256 MiB, alternating directions on two streams
    piece = (256 << 20) // 16
    def both_pieces():
        for o in range(0, n, piece):
            with torch.cuda.stream(s_up):
                dev_a[o:o + piece].copy_(host_in[o:o + piece], non_blocking=True)
            with torch.cuda.stream(s_down):
                host_out[o:o + piece].copy_(dev_b[o:o + piece], non_blocking=True)
    out["both_256MiB_pieces_GBps_per_direction"] = nbytes / timed(both_pieces, [s_up, s_down], args.reps) / 1e9

    # two upload streams + two download streams (halves)
    s_up2, s_down2 = torch.cuda.Stream(), torch.cuda.Stream()
    h = n // 2
    def both_split():
        with torch.cuda.stream(s_up):
            dev_a[:h].copy_(host_in[:h], non_blocking=True)
        with torch.cuda.stream(s_up2):
            dev_a[h:].copy_(host_in[h:], non_blocking=True)
        with torch.cuda.stream(s_down):
            host_out[:h].copy_(dev_b[:h], non_blocking=True)
        with torch.cuda.stream(s_down2):
            host_out[h:].copy_(dev_b[h:], non_blocking=True)
    out["both_two_streams_each_GBps_per_direction"] = nbytes / timed(both_split, [s_up, s_up2, s_down, s_down2], args.reps) / 1e9

    # write-combined / portable host allocations through the runtime (torch's pin_memory is cudaHostAlloc default flags)
    try:
        from cuda.bindings import runtime as rt
        for name, flags in (("default", rt.cudaHostAllocDefault), ("write_combined", rt.cudaHostAllocWriteCombined)):
            err, p = rt.cudaHostAlloc(nbytes, flags)
            if int(err) != 0:
                out["hostalloc_" + name] = f"cudaHostAlloc failed: {err}"
                continue
            st = s_up.cuda_stream
            def up_raw():
                rt.cudaMemcpyAsync(dev_a.data_ptr(), p, nbytes, rt.cudaMemcpyKind.cudaMemcpyHostToDevice, st)
            # touch the pages once
            rt.cudaMemcpyAsync(p, dev_b.data_ptr(), nbytes, rt.cudaMemcpyKind.cudaMemcpyDeviceToHost, st)
            torch.cuda.synchronize()
            out["h2d_alone_hostalloc_" + name + "_GBps"] = nbytes / timed(up_raw, [s_up], args.reps) / 1e9
            def both_raw():
                up_raw()
                down()
            out["both_hostalloc_" + name + "_GBps_per_direction"] = nbytes / timed(both_raw, [s_up, s_down], args.reps) / 1e9
            rt.cudaFreeHost(p)
    except Exception as e:  # the probe is a diagnostic: say what was not measured
        out["hostalloc_variants"] = f"not measured: {type(e).__name__}: {e}"
    print(json.dumps(out))


if __name__ == "__main__":
    main()
