#!/usr/bin/env python
"""Long differential fuzz of the fusion planner on the CPU: random programs over every op kind (tools/fuzz_fused.py's
generator), random tile sizes, every other case a requested end layout (1-3 victim bits parked at the top) with or without
an exchange-carrying last sweep, emulated by libqsv_hostcheck.so and compared with the oracle.  No GPU needed.

  python tools/fuzz_hostcheck.py SEED CASES        (QSV_CARRY8=1 / QSV_NO_MIXED_CAPS=1 / QSV_TILE_BITS=k select plan variants)

Round 2 ran 2 800 cases over six seeds and the three switches: no failure.
"""
import importlib.util, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import sv_oracle
from tests.hostcheck import run_fused_to_layout_on_host
spec = importlib.util.spec_from_file_location("fuzz_fused", os.path.join(ROOT, "tools", "fuzz_fused.py"))
fz = importlib.util.module_from_spec(spec); spec.loader.exec_module(fz)
seed = int(sys.argv[1]); cases = int(sys.argv[2])
rng = np.random.default_rng(seed)
t0 = time.time(); bad = 0
for case in range(cases):
    n = int(rng.integers(10, 19))
    prog = fz.random_program(rng, n, int(rng.choice([5, 30, 120, 300, 600])))
    x = rng.standard_normal(1 << n) + 1j * rng.standard_normal(1 << n); x /= np.linalg.norm(x)
    ref = x.copy(); sv_oracle.run_records(prog.records(0), prog.pool(), ref, n)
    tile = int(rng.choice([9, 10, 11]))
    final = list(range(n)); g = 0
    if case % 2:
        g = int(rng.integers(1, 4))
        v = sorted(rng.choice(n, size=g, replace=False).tolist())
        rest = [b for b in range(n) if b not in v]
        for i, b in enumerate(rest): final[b] = i
        for i, b in enumerate(v): final[b] = n - g + i
    carry = bool(case % 3 == 0)
    try:
        got, sw, fft = run_fused_to_layout_on_host(prog.records(0), prog.pool(), x, n, final, carry, tile)
    except RuntimeError as e:
        if "unsupported" in str(e):
            continue
        print("ERR", seed, case, n, tile, final, carry, e); bad += 1; continue
    idx = np.arange(1 << n); phys = np.zeros_like(idx)
    for b in range(n): phys |= ((idx >> b) & 1) << final[b]
    want = np.empty_like(ref); want[phys] = ref
    err = np.abs(got - want).max()
    if err > 1e-12:
        print("BAD", seed, case, n, tile, final, carry, err); bad += 1
print("seed", seed, "cases", cases, "bad", bad, "time", round(time.time() - t0, 1))
