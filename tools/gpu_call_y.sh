#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
timeout 600 python bench.py > $O/r2y_bench_default.json 2> $O/r2y_bench_default.err; echo "bench rc=$?"
tail -3 $O/r2y_bench_default.err
python - <<PY
import json
d=json.load(open("$O/r2y_bench_default.json"))
print(round(d["ms_per_step"],3), d["value"], d["roofline"]["frac"], d["e2e"], d["check"]["ok"])
PY
timeout 300 python bench.py --e2e-steps 6 --cpu-gates 0 --no-check > $O/r2y_bench_e2e6.json 2>/dev/null
python - <<PY
import json
d=json.load(open("$O/r2y_bench_e2e6.json"))
print("e2e 6 steps", d["e2e"])
PY
