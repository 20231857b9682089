#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_stream.py tests/test_qasm.py tests/test_optimizer.py tests/test_qco_compat.py -m gpu -x -q > $O/r2k_pytest.txt 2>&1; echo "pytest rc=$?" >> $O/r2k_pytest.txt; tail -3 $O/r2k_pytest.txt
QSV_GEN8_MAX_TARGETS=9 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "fuzz or random_circuit or stream_golden" > $O/r2k_pytest9.txt 2>&1; echo "pytest9 rc=$?" >> $O/r2k_pytest9.txt; tail -2 $O/r2k_pytest9.txt
python -c "import __graft_entry__ as g; g.smoke()" > $O/r2k_smoke.txt 2>&1; echo "smoke rc=$?"; tail -2 $O/r2k_smoke.txt
B="python bench.py --steps 3 --warmup 3 --e2e-steps 0 --cpu-gates 0"
for c in random:96 cyclic random:400:7; do
  t=$(echo $c | tr ':' '_')
  QSV_NO_GEN8=1 timeout 300 $B --circuit $c > $O/r2k_${t}_gen16.json 2> $O/r2k_${t}_gen16.err
  timeout 300 $B --circuit $c > $O/r2k_${t}_gen8_6.json 2> $O/r2k_${t}_gen8_6.err
  QSV_GEN8_MAX_TARGETS=9 timeout 300 $B --circuit $c > $O/r2k_${t}_gen8_9.json 2> $O/r2k_${t}_gen8_9.err
done
for f in $O/r2k_*.json; do python - <<PY
import json
try:
    d=json.load(open("$f"))
    print("$f".split("/")[-1], round(d["ms_per_step"],3), "frac", round(d["roofline"]["frac"],4), d["impl_config"]["sweeps_per_step"], d["sweep_ms_first_step"][:12], d["check"]["ok"], d["check"]["max_abs_err"])
except Exception as e:
    print("$f", "FAILED", e)
PY
done
