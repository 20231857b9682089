#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
B="python bench.py --steps 5 --warmup 3 --e2e-steps 0 --cpu-gates 0"
for h in 0 1 2 3 0; do
QSV_L2_HINT=$h timeout 300 $B > $O/r2r_hint$h.json 2> $O/r2r_hint$h.err; echo "rc=$?"
python - <<PY
import json
try:
    d=json.load(open("$O/r2r_hint$h.json"))
    print("hint$h", round(d["ms_per_step"],3), round(d["roofline"]["frac"],4), d["sweep_ms_first_step"], d["check"]["ok"], d["clocks"]["sm_mhz"])
except Exception as e:
    print("hint$h", "FAILED", e); print(open("$O/r2r_hint$h.err").read()[-1000:])
PY
done
