#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
timeout 400 python -m pytest tests/test_gpu_sharded_one_device.py -m gpu -x -q > $O/r2z_pytest_one_device.txt 2>&1; echo "pytest rc=$?" >> $O/r2z_pytest_one_device.txt; tail -30 $O/r2z_pytest_one_device.txt
timeout 300 python bench.py --e2e-steps 0 --cpu-gates 0 > $O/r2z_bench_qft30.json 2> $O/r2z_bench.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open("$O/r2z_bench_qft30.json"))
print(round(d["ms_per_step"],3), d["roofline"]["frac"], d["check"]["ok"], d["sweep_ms_first_step"])
PY
