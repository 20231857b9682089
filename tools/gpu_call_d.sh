#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
B="python bench.py --steps 5 --warmup 3 --e2e-steps 0 --cpu-gates 0"
for pf in 0 1 2 3; do
QSV_L2_PREFETCH=$pf timeout 300 $B > $O/r2d_bench_pf$pf.json 2> $O/r2d_bench_pf$pf.err; echo "rc=$?"
done
QSV_L2_PREFETCH=0 timeout 300 $B > $O/r2d_bench_pf0b.json 2> $O/r2d_bench_pf0b.err
for f in pf0 pf1 pf2 pf3 pf0b; do python - <<PY
import json
try:
    d=json.load(open("$O/r2d_bench_$f.json"))
    print("$f", round(d["ms_per_step"],3), round(d["roofline"]["frac"],4), d["sweep_ms_first_step"], d["check"]["ok"], d["clocks"]["sm_mhz"])
except Exception as e:
    print("$f", "FAILED", e); print(open("$O/r2d_bench_$f.err").read()[-1500:])
PY
done
