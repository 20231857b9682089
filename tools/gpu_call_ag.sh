#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
timeout 600 python bench.py > $O/r2ag_bench_default.json 2> $O/r2ag_bench_default.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open("$O/r2ag_bench_default.json"))
print(round(d["ms_per_step"],3), d["roofline"]["frac"], d["e2e"], d["check"]["ok"])
PY
