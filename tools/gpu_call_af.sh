#!/bin/bash
# final tree (mixed-cap plans on), one GPU: smoke, the whole GPU suite, both bench arms, the cyclic stream
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
python -c "import __graft_entry__ as g; g.smoke()" > $O/r2af_smoke.txt 2>&1; echo "smoke rc=$?"; tail -2 $O/r2af_smoke.txt
timeout 1200 python -m pytest tests -m gpu -x -q > $O/r2af_pytest_gpu.txt 2>&1; echo "pytest rc=$?" >> $O/r2af_pytest_gpu.txt; tail -3 $O/r2af_pytest_gpu.txt
timeout 600 python bench.py > $O/r2af_bench_default.json 2> $O/r2af_bench_default.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference > $O/r2af_bench_reference.json 2> $O/r2af_bench_reference.err; echo "ref rc=$?"
timeout 300 python bench.py --circuit cyclic --steps 3 --warmup 3 --e2e-steps 0 --cpu-gates 0 > $O/r2af_bench_cyclic.json 2> $O/r2af_bench_cyclic.err; echo "cyclic rc=$?"
python - <<PY
import json
d=json.load(open("$O/r2af_bench_default.json")); r=json.load(open("$O/r2af_bench_reference.json")); c=json.load(open("$O/r2af_bench_cyclic.json"))
print(round(d["ms_per_step"],3), d["value"], d["roofline"]["frac"], d["e2e"]["value"], d["e2e"]["ms_per_step"], d["check"]["ok"], d["config"]==r["config"], r["value"], r["cpu_baseline"]["cores"])
print("e2e ratio", d["e2e"]["value"]/r["value"], "device ratio", d["value"]/r["value"])
print("cyclic", round(c["ms_per_step"],3), c["roofline"]["frac"], c["check"])
PY
