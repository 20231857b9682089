#!/usr/bin/env python
"""Benchmark of the hot path: applying a compiled gate stream to a statevector on B200.

Contract (see the task brief):  python bench.py --gpus N --steps K --warmup W  prints ONE JSON line.
 * workload (N=1): QFT-30 gate circuit (30 H, 435 controlled-phase, 45 CNOT) on one complex128
   statevector of 2^30 amplitudes (16 GiB, >> L2) -- the configuration BASELINE.json's metric names.
   N>1: the same circuit and state sharded by the top log2(N) qubits (strong scaling), half-state
   swaps over NCCL.  `--qubits n` selects another size (e.g. 34 on 8 GPUs).
 * a step = one application of the whole circuit to the resident state.
 * value = source gates / second (whole job); amp_GBps = algorithmic sweep bytes / second.
 * e2e   = the same through host buffers: pinned H2D of the state, run, D2H of the result, every step.
 * roofline: the dominant kernel's algorithmic bytes per launch / its CUDA-event time per launch.
 * cpu_baseline / --impl reference: the oracle's C+OpenMP port of the reference's gate action, timed on
   the host cores on a bounded sample (first G gates of the same circuit on the same-size state).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


CIRCUIT = "qft"
CHECK_X = 0x2B992DDF          # basis state used by the QFT self-check (odd, mixed bits; masked to n bits)


def workload_kind():
    return CIRCUIT.split(":")[0]


def workload_name(n):
    kind = workload_kind()
    if kind == "qft":
        return f"qft{n}_statevector_c128"
    return f"{CIRCUIT.replace(':', '_')}_{n}q_statevector_c128"


def qft_records(n, batch_bits=0):
    """Flattened benchmark circuit (time order) without building Python gate objects for every gate.

    --circuit qft            QFT-n gate circuit (the headline workload)
    --circuit random:D[:S]   D dense 2x2 gates (Haar random) with 0..3 controls (either polarity) on random distinct qubits,
                             seed S -- the shape of the compiler's SINGLET / CTRL_PRUNED output for a random unitary
                             (quompiler/utils/mgen.py:52-76,140-145: random_control + random_CtrlGate)
    --circuit cyclic         cyclic shift by one of the 2^n basis states (quompiler/utils/mgen.py:112-125, cyclic_matrix(2^n))
                             as n multi-controlled X gates: a pure permutation stream, checked BIT-EXACTLY
    --circuit hlayer:LO:HI   Hadamards on bits LO..HI-1 (diagnostic)"""
    from quantum_simulations_b200.construct.qspace import Qubit
    from quantum_simulations_b200.construct.types import UnivGate
    from quantum_simulations_b200.program import Program
    from quantum_simulations_b200.utils.mgen import qft_program
    kind = workload_kind()
    if kind == "qft":
        prog = qft_program(n)
        return prog, prog.records(batch_bits), prog.pool()
    prog = Program([Qubit(i) for i in range(n)])
    if kind == "hlayer":
        _, lo, hi = CIRCUIT.split(":")
        for b in range(int(lo), int(hi)):
            prog.add_raw([b], 0, 0, np.array(UnivGate.H))
    elif kind == "random":
        parts = CIRCUIT.split(":")
        depth = int(parts[1]) if len(parts) > 1 else 64
        rng = np.random.default_rng(int(parts[2]) if len(parts) > 2 else 42)
        for _ in range(depth):
            c = int(rng.integers(0, 4))
            bits = rng.choice(n, size=c + 1, replace=False)
            a = rng.standard_normal((2, 2)) + 1j * rng.standard_normal((2, 2))
            q, r = np.linalg.qr(a)
            u = q @ np.diag(np.diag(r) / np.abs(np.diag(r)))
            mask = value = 0
            for b in bits[1:]:
                mask |= 1 << int(b)
                value |= int(rng.integers(0, 2)) << int(b)
            prog.add_raw([int(bits[0])], mask, value, u)
    elif kind == "cyclic":
        x = np.array(UnivGate.X)
        for k in range(n - 1, -1, -1):      # +1 mod 2^n: bit k flips iff all lower bits are 1 (highest bit first)
            low = (1 << k) - 1
            prog.add_raw([k], low, low, x)
    else:
        raise SystemExit(f"unknown --circuit {CIRCUIT}")
    return prog, prog.records(batch_bits), prog.pool()


def qft_basis_error(shard, n, x0, base_index=0, where=None, chunk=1 << 24):
    """max |shard[p] - exp(2 pi i k x0 / 2^n) / sqrt(2^n)| over a (shard of a) state that should hold QFT|x0>
    (fft_matrix(n) column x0, quompiler/utils/mgen.py:128-137: omega = e^{+2 pi i / N}).  ``p + base_index`` is the
    physical index; ``where`` (logical bit -> physical position) maps it to the logical index k when the layout is not
    canonical.  Checker only: runs after the timed region."""
    import math
    import torch
    dev = shard.device
    mask = (1 << n) - 1
    amp = 2.0 ** (-n / 2)
    worst = 0.0
    for s in range(0, shard.numel(), chunk):
        e = min(s + chunk, shard.numel())
        p = torch.arange(s, e, device=dev, dtype=torch.int64) + base_index
        if where is not None and list(where) != list(range(n)):
            k = torch.zeros_like(p)
            for b in range(n):
                k |= ((p >> int(where[b])) & 1) << b
        else:
            k = p
        m = (k * int(x0)) & mask                      # int64 wrap-around is harmless: 2^n divides 2^64
        ang = m.to(torch.float64) * (2.0 * math.pi / (1 << n))
        want = torch.polar(torch.full_like(ang, amp), ang)
        worst = max(worst, float((shard[s:e] - want).abs().max()))
        del p, k, m, ang, want
    return worst


def algorithmic_bytes(recs, n_bits, amp_bytes=16):
    """Unfused lower bound: every gate reads+writes the amplitudes it touches."""
    total = 0
    for r in recs:
        c = bin(int(r["ctrl_mask"])).count("1")
        touched = (1 << n_bits) >> c
        if int(r["kind"]) == 2:  # diagonal: diag(1, d) touches only target=1
            touched >>= 1
        total += 2 * touched * amp_bytes
    return total


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                for nm, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def ncu_traffic(n):
    """(DRAM bytes per k_sweep launch, source) from the committed `ncu --set full` capture under profiles/ -- a constant
    read from a file, NOT a measurement of this run (ncu cannot run inside a timed bench)."""
    p = os.path.join(ROOT, "profiles", "ncu_k_sweep_traffic.json")
    try:
        rec = json.load(open(p)).get(f"qft{n}", {})
        return rec.get("dram_bytes_per_launch"), rec.get("source", "profiles/ncu_k_sweep_traffic.json")
    except Exception:
        return None, None


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def workload_config(n, n_gates):
    """The workload's own description: identical on the GPU arm and on the reference arm."""
    return {"workload": workload_name(n), "circuit": CIRCUIT, "qubits": n, "amplitudes": 1 << n, "gates": n_gates,
            "l2": "state (16 B x 2^n) is far larger than the 126 MB L2; nothing is flushed between steps"}


def cpu_sample(n, n_gates_sample):
    """Oracle C/OpenMP port on the first `n_gates_sample` gates of the workload on a full-size state."""
    from oracle import c_oracle
    cores = c_oracle.set_threads(0)                      # every core, whatever OMP_NUM_THREADS says (torchrun sets it to 1)
    prog, recs, pool = qft_records(n)
    g = min(n_gates_sample, len(recs))
    st = np.full(1 << n, 2.0 ** (-n / 2), dtype=np.complex128)
    c_oracle.run_records(recs[:1], pool, st, n)          # touch pages / warm threads
    t0 = time.perf_counter()
    c_oracle.run_records(recs[:g], pool, st, n)
    dt = time.perf_counter() - t0
    return g / dt, g, dt, cores


def run_reference(args):
    """--impl reference: the CPU port (the reference itself is Python and cannot reach these sizes:
    its dense-operator path needs 2^n x 2^n matrices)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n = args.qubits or 30
    from oracle import c_oracle
    cores = c_oracle.set_threads(0)
    prog, recs, pool = qft_records(n)
    g = min(args.cpu_gates, len(recs))
    st = np.full(1 << n, 2.0 ** (-n / 2), dtype=np.complex128)
    for _ in range(max(1, min(args.warmup, 1))):
        c_oracle.run_records(recs[:2], pool, st, n)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        c_oracle.run_records(recs[:g], pool, st, n)
    dt = time.perf_counter() - t0
    value = args.steps * g / dt
    sample = f"first {g} of {len(recs)} gates of {workload_name(n)} on the full 2^{n} complex128 state, per step"
    line = {"impl": "reference", "metric": "gates_per_sec", "value": value, "unit": "gates/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "c128", "data": "synthetic",
            "config": workload_config(n, len(recs)),
            "cpu_baseline": {"value": value, "unit": "gates/s", "cores": cores, "kind": "port", "sample": sample,
                             "omp_num_threads_env": os.environ.get("OMP_NUM_THREADS")},
            "e2e": {"value": value, "unit": "gates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def fill_random_state(state, dev, seed=42):
    """complex standard normal, L2-normalised (mgen.random_state), generated on the device."""
    import torch
    g = torch.Generator(device=f"cuda:{dev}")
    g.manual_seed(seed)
    view = torch.view_as_real(state)
    chunk = 1 << 26
    for s in range(0, state.numel(), chunk):
        view[s:s + chunk].normal_(generator=g)
    state.div_(torch.linalg.vector_norm(state))


def chunked_max_abs_diff(a, b, chunk=1 << 26):
    worst = 0.0
    for s in range(0, a.numel(), chunk):
        worst = max(worst, float((a[s:s + chunk] - b[s:s + chunk]).abs().max()))
    return worst


def self_check(eng, n, recs, pool, bufs, fused, tile_bits):
    """Correctness of the run at benchmark size, OUTSIDE the timed region.  Returns a dict whose "ok" says whether
    every figure is inside the parity bar (1e-10 complex128; bit-exact for the permutation stream)."""
    import torch
    amps = 1 << n
    a, b, c = bufs
    kind = workload_kind()
    out = {"tolerance": 1e-10}

    def run(x, spare):
        res = eng.run(x, spare if fused else None)
        torch.cuda.synchronize()
        return res

    if kind == "qft":
        x0 = CHECK_X & (amps - 1)
        a.zero_()
        a[x0] = 1.0
        res = run(a, b)
        err = qft_basis_error(res, n, x0)
        out.update({"check": f"QFT|x0={x0:#x}> against the closed form exp(2 pi i k x0/2^n)/sqrt(2^n) on all 2^{n} amplitudes",
                    "max_abs_err": err, "max_rel_err": err * 2.0 ** (n / 2)})
        ok = err <= 1e-10 and err * 2.0 ** (n / 2) <= 1e-9
        if n <= 30:
            # second check on generic data: cuFFT (torch.fft) as the checker, fft_matrix(n) @ x == ifft(x, norm="ortho")
            fill_random_state(a, eng.device, seed=7)
            c.copy_(a)
            res = run(a, b)
            want = torch.fft.ifft(c, norm="ortho")
            e2 = chunked_max_abs_diff(res, want)
            del want
            out["max_abs_err_vs_ifft_random_state"] = e2
            ok = ok and e2 <= 1e-10
        out["ok"] = bool(ok)
    elif kind == "cyclic":
        fill_random_state(a, eng.device, seed=7)
        c.copy_(a)
        res = run(a, b)
        want = torch.roll(c, 1)
        same = True
        for s in range(0, amps, 1 << 26):
            same = same and bool(torch.equal(torch.view_as_real(res[s:s + (1 << 26)]), torch.view_as_real(want[s:s + (1 << 26)])))
        del want
        out.update({"check": "out[(x+1) mod 2^n] == in[x] for a random state, compared bit for bit", "bit_exact": same,
                    "max_abs_err": 0.0 if same else chunked_max_abs_diff(res, torch.roll(c, 1)), "ok": same})
    else:
        # generic stream: the fused sweeps against the one-kernel-per-gate path (itself pinned to the oracle at small
        # sizes by tests/test_gpu_parity.py) on the same random state, plus the norm
        from quantum_simulations_b200.engine import Engine
        fill_random_state(a, eng.device, seed=7)
        c.copy_(a)
        res = run(a, b)
        other = b if res is a else a
        eng2 = Engine(eng.device, np.complex128)
        eng2.upload_records(recs, pool, n, 0)
        ref = eng2.run(c, None)
        torch.cuda.synchronize()
        err = chunked_max_abs_diff(res, ref)
        nrm = float(torch.linalg.vector_norm(res))
        eng2.close()
        del other
        out.update({"check": "fused sweeps against the per-gate kernels on the same random state, all amplitudes; norm",
                    "max_abs_err": err, "norm_after": nrm, "ok": bool(err <= 1e-10 and abs(nrm - 1) <= 1e-10)})
    return out


def run_single(args):
    import torch
    from quantum_simulations_b200.engine import Engine
    n = args.qubits or 30
    dev = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(dev)
    eng = Engine(dev, np.complex128)
    prog, recs, pool = qft_records(n)
    n_gates = len(recs)
    amps = 1 << n
    state = eng.empty(amps)
    fill_random_state(state, dev)
    eng.upload_records(recs, pool, n, 0)
    fused = False
    spare = None
    n_sweeps = None
    if not args.no_fuse:
        n_sweeps = eng.fuse(args.tile_bits)              # raises when the fusion pass refuses: no silent per-gate downgrade
        spare = eng.empty(amps)
        fused = True
    cur = state
    other = spare

    def step():
        nonlocal cur, other
        res = eng.run(cur, other)
        if res is not cur:
            cur, other = res, cur

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    eng.reset_stats()
    if fused:
        eng.set_profiling(True)
    sampler = ClockSampler(dev)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop()
    st = eng.stats()
    sec = ms / 1e3
    value = args.steps * n_gates / sec
    # ---- roofline of the dominant kernel ----------------------------------------------------
    peak, peak_src = measured_peak()
    launches = st["kernel_launches"]
    traffic, traffic_src = (None, None)
    if fused:
        kern = "k_sweep (fused tile sweep)"
        per_launch_bytes = 2 * amps * 16
        prof = eng.sweep_profile()
        eng.set_profiling(False)
        per_step = max(1, len(prof) // max(1, args.steps))
        sweep_ms = [round(x, 3) for x in prof[:per_step]] if prof else None
        avg_ms = float(np.mean(prof)) if prof else ms / max(1, launches)
        achieved = per_launch_bytes / (avg_ms * 1e-3) / 1e9
        if workload_kind() == "qft":
            traffic, traffic_src = ncu_traffic(n)
    else:
        sweep_ms = None
        kern = "k_gate1/k_phase1 (one kernel per gate; aggregate over all launches)"
        per_launch_bytes = st["amp_bytes"] / max(1, launches)
        avg_ms = ms / max(1, launches)
        achieved = st["amp_bytes"] / sec / 1e9
    amp_gbps = st["amp_bytes"] / sec / 1e9
    # arithmetic of the circuit itself: 28 real flops per 2x2 complex butterfly on the amplitude pairs a gate touches
    # (SURVEY.md 8d); diagonal gates 6 per touched amplitude.  Dense-gate streams are bound by this, not by HBM.
    flop = 0.0
    for r in recs:
        c = bin(int(r["ctrl_mask"])).count("1")
        kind = int(r["kind"])
        if kind == 0:
            flop += 28.0 * (amps >> (c + 1))
        elif kind == 2:
            flop += 6.0 * (amps >> c)
    compute = None
    if workload_kind() == "random":   # (diagonal runs of QFT-like streams are merged by the fusion pass: their gate-by-gate count says nothing)
        compute = {"fp64_flop_per_step": flop, "achieved_TFLOPs": flop * args.steps / sec / 1e12,
                   "note": "algorithmic FP64 work of the gate stream / step time; B200's FP64 vector rate is nominally ~37-40 TFLOP/s (not measured on this pool)"}
    # ---- e2e through host buffers -------------------------------------------------------------
    # Every step uploads its input state from pinned host memory, applies the circuit and downloads the result.
    # Steps are independent jobs, so they are pipelined over four device buffers (two per step in flight: state + the
    # fused plan's second buffer) and three streams: the download of step i overlaps the upload of step i+1 (PCIe is full
    # duplex) and the compute in between never waits for a download.  Copies go in 256 MiB pieces: with both directions
    # busy that measured 49.8 GB/s per direction against 46.0 for one 8 GiB copy each way (profiles/r2x_pcie_probe.json).
    e2e = None
    third = eng.empty(amps)
    if args.e2e_steps > 0:
        fourth = None
        if other is not None and torch.cuda.mem_get_info()[0] > amps * 16 + (8 << 30):   # (a 32 GiB state: three buffers must do)
            fourth = eng.empty(amps)
        piece = (256 << 20) // 16

        def copy_in_pieces(dst_t, src_t):
            for o in range(0, amps, piece):
                dst_t[o:o + piece].copy_(src_t[o:o + piece], non_blocking=True)

        host_in = torch.empty(amps, dtype=torch.complex128, pin_memory=True)
        host_out = torch.empty(amps, dtype=torch.complex128, pin_memory=True)
        host_in.copy_(cur)
        torch.cuda.synchronize()
        free = [b for b in (cur, other, third, fourth) if b is not None]
        s_h2d, s_run, s_d2h = torch.cuda.Stream(), torch.cuda.Stream(), torch.cuda.Stream()

        def pipeline(count):
            bufs = list(free)
            busy_until = {id(b): None for b in bufs}    # event after which the buffer may be overwritten
            last = None
            for _ in range(count):
                a = bufs.pop(0)
                sp = bufs.pop(0) if (fused and len(bufs)) else None
                with torch.cuda.stream(s_h2d):
                    for b in (a, sp):
                        if b is not None and busy_until[id(b)] is not None:
                            s_h2d.wait_event(busy_until[id(b)])
                    copy_in_pieces(a, host_in)
                    ev_in = torch.cuda.Event()
                    ev_in.record(s_h2d)
                with torch.cuda.stream(s_run):
                    s_run.wait_event(ev_in)
                    if sp is not None and busy_until[id(sp)] is not None:
                        s_run.wait_event(busy_until[id(sp)])
                    res = eng.run(a, sp)
                    ev_run = torch.cuda.Event()
                    ev_run.record(s_run)
                with torch.cuda.stream(s_d2h):
                    s_d2h.wait_event(ev_run)
                    copy_in_pieces(host_out, res)
                    ev_out = torch.cuda.Event()
                    ev_out.record(s_d2h)
                # the buffer holding the result is busy until its download ends; the other one is free after compute
                # and goes back first: it takes the next upload
                for b in sorted((x for x in (a, sp) if x is not None), key=lambda x: x is res):
                    busy_until[id(b)] = ev_out if b is res else ev_run
                    bufs.append(b)
                last = ev_out
            return last

        pipeline(1).synchronize()
        torch.cuda.synchronize()
        t0 = torch.cuda.Event(enable_timing=True)
        t1 = torch.cuda.Event(enable_timing=True)
        t0.record()
        for st_ in (s_h2d, s_run, s_d2h):
            st_.wait_event(t0)
        pipeline(args.e2e_steps)
        for st_ in (s_h2d, s_run, s_d2h):
            torch.cuda.current_stream().wait_stream(st_)
        t1.record()
        torch.cuda.synchronize()
        e2e_ms = t0.elapsed_time(t1) / args.e2e_steps
        e2e = {"value": n_gates / (e2e_ms / 1e3), "unit": "gates/s", "h2d_bytes_per_step": amps * 16, "d2h_bytes_per_step": amps * 16,
               "ms_per_step": e2e_ms, "steps": args.e2e_steps,
               "note": f"steps pipelined over {len(free)} device buffers / 3 streams, copies in 256 MiB pieces: D2H of step i overlaps H2D + compute of step i+1"}
        del host_in, host_out, fourth
    # ---- correctness at this size (outside the timed region) ------------------------------------
    check = None
    if not args.no_check:
        check = self_check(eng, n, recs, pool, (cur, other if other is not None else eng.empty(amps), third), fused, args.tile_bits)
    # ---- CPU baseline on the host cores ---------------------------------------------------------
    cpu = None
    if args.cpu_gates > 0:
        v, gcount, dt, cores = cpu_sample(n, args.cpu_gates)
        cpu = {"value": v, "unit": "gates/s", "cores": cores, "kind": "port",
               "sample": f"first {gcount} of {n_gates} gates of {workload_name(n)} on the full 2^{n} complex128 state ({dt:.1f} s, C+OpenMP oracle)"}
    line = {"metric": "gates_per_sec", "value": value, "unit": "gates/s", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "c128",
            "data": "synthetic",
            "config": workload_config(n, n_gates),
            "impl_config": {"fused": fused, "sweeps_per_step": n_sweeps, "tile_bits": args.tile_bits or "default"},
            "amp_GBps": amp_gbps, "hbm_frac_of_measured": amp_gbps / peak, "hbm_frac_of_nominal_8TBps": amp_gbps / 8000.0,
            "roofline": {"bound": "hbm", "kernel": kern, "achieved": achieved, "peak": peak, "peak_source": peak_src, "unit": "GB/s",
                         "frac": achieved / peak, "frac_of_nominal_8TBps": achieved / 8000.0, "bytes_per_launch": per_launch_bytes,
                         "avg_launch_ms": avg_ms, "traffic": traffic, "traffic_source": traffic_src},
            "compute": compute,
            "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks, "sweep_ms_first_step": sweep_ms,
            "check": check, "max_abs_err": check.get("max_abs_err") if check else None}
    print(json.dumps(line))
    if check is not None and not check["ok"]:
        print(f"[bench] SELF-CHECK FAILED: {check}", file=sys.stderr)
        sys.exit(3)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--qubits", type=int, default=0)
    ap.add_argument("--tile-bits", type=int, default=0)
    ap.add_argument("--no-fuse", action="store_true")
    ap.add_argument("--no-check", action="store_true", help="skip the correctness check that follows the timed region")
    ap.add_argument("--no-qft34", action="store_true", help="on 8 GPUs: do not append the QFT-34 (BASELINE config 5) record")
    ap.add_argument("--no-merge", action="store_true", help="sharded runs: keep park permutation / exchange as sweeps of their own (A/B)")
    ap.add_argument("--restore-layout", action="store_true",
                    help="sharded runs: bring the result back to the canonical layout (one more exchange per QFT) instead of "
                         "reporting the final bit-position table")
    ap.add_argument("--e2e-steps", type=int, default=32)
    ap.add_argument("--cpu-gates", type=int, default=12, help="gates in the CPU baseline sample (0 = skip)")
    ap.add_argument("--exchange", default="auto", choices=["auto", "fused", "nccl"], help="multi-GPU exchange engine")
    ap.add_argument("--circuit", default="qft", help="qft | random:DEPTH[:SEED] | cyclic | hlayer:LO:HI (diagnostic)")
    args = ap.parse_args()
    global CIRCUIT
    CIRCUIT = args.circuit
    if args.impl == "reference":
        return run_reference(args)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1 or args.gpus > 1:
        from quantum_simulations_b200 import sharded
        return sharded.bench_main(args)
    return run_single(args)


if __name__ == "__main__":
    main()
