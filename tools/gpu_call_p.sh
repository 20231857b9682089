#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q > $O/r2p_pytest_gpu.txt 2>&1; echo "pytest rc=$?" >> $O/r2p_pytest_gpu.txt; tail -3 $O/r2p_pytest_gpu.txt
B="python bench.py --steps 3 --warmup 3 --e2e-steps 0 --cpu-gates 0"
for c in random:96 random:400:7 cyclic qft; do
  t=$(echo $c | tr ':' '_')
  QSV_NO_SCHEDULE=1 timeout 300 $B --circuit $c > $O/r2p_${t}_nosched.json 2> $O/r2p_${t}_nosched.err
  timeout 300 $B --circuit $c > $O/r2p_${t}_sched.json 2> $O/r2p_${t}_sched.err
done
for f in $O/r2p_*.json; do python - <<PY
import json
try:
    d=json.load(open("$f"))
    print("$f".split("/")[-1], round(d["ms_per_step"],3), "gates/s", round(d["value"]), "frac", round(d["roofline"]["frac"],4), d["impl_config"]["sweeps_per_step"], d["check"]["ok"], d["check"]["max_abs_err"])
except Exception as e:
    print("$f", "FAILED", e)
PY
done
