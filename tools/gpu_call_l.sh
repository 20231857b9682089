#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > $O/r2l_pytest.txt 2>&1; echo "pytest rc=$?" >> $O/r2l_pytest.txt; tail -3 $O/r2l_pytest.txt
B="python bench.py --steps 3 --warmup 3 --e2e-steps 0 --cpu-gates 0"
for c in random:96 cyclic random:400:7 qft; do
  t=$(echo $c | tr ':' '_')
  QSV_NO_COST_MODEL=1 timeout 300 $B --circuit $c > $O/r2l_${t}_nocost.json 2> $O/r2l_${t}_nocost.err
  timeout 300 $B --circuit $c > $O/r2l_${t}_cost.json 2> $O/r2l_${t}_cost.err
done
for f in $O/r2l_*.json; do python - <<PY
import json
try:
    d=json.load(open("$f"))
    print("$f".split("/")[-1], round(d["ms_per_step"],3), "gates/s", round(d["value"]), "frac", round(d["roofline"]["frac"],4), d["impl_config"]["sweeps_per_step"], d["sweep_ms_first_step"][:14], d["check"]["ok"], d["check"]["max_abs_err"])
except Exception as e:
    print("$f", "FAILED", e)
PY
done
