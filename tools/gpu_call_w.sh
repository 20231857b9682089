#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
python -c "import __graft_entry__ as g; g.smoke()" > $O/r2w_smoke.txt 2>&1; echo "smoke rc=$?"; tail -2 $O/r2w_smoke.txt
timeout 1200 python -m pytest tests -m gpu -x -q > $O/r2w_pytest_gpu.txt 2>&1; echo "pytest rc=$?" >> $O/r2w_pytest_gpu.txt; tail -3 $O/r2w_pytest_gpu.txt
timeout 600 python bench.py > $O/r2w_bench_default.json 2> $O/r2w_bench_default.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference > $O/r2w_bench_reference.json 2> $O/r2w_bench_reference.err; echo "ref rc=$?"
python - <<PY
import json
d=json.load(open("$O/r2w_bench_default.json")); r=json.load(open("$O/r2w_bench_reference.json"))
print(round(d["ms_per_step"],3), d["value"], d["roofline"]["frac"], d["e2e"]["value"], d["check"]["ok"], d["config"]==r["config"], r["value"], r["cpu_baseline"]["cores"])
print("e2e ratio", d["e2e"]["value"]/r["value"], "device ratio", d["value"]/r["value"])
PY
