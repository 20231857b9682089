#!/bin/bash
# 1 GPU: whole GPU test suite, warp-shuffle A/B, bench lines of the other workload families
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q > $O/r2f_pytest_gpu.txt 2>&1; echo "pytest rc=$?" >> $O/r2f_pytest_gpu.txt; tail -4 $O/r2f_pytest_gpu.txt
B="python bench.py --steps 3 --warmup 3 --e2e-steps 0 --cpu-gates 0"
for c in hlayer:0:1 hlayer:3:4 hlayer:0:4; do
  t=$(echo $c | tr ':' '_')
  timeout 300 $B --circuit $c --no-fuse > $O/r2f_${t}_pair.json 2> $O/r2f_${t}_pair.err
  QSV_SHFL=1 timeout 300 $B --circuit $c --no-fuse > $O/r2f_${t}_shfl.json 2> $O/r2f_${t}_shfl.err
  timeout 300 $B --circuit $c > $O/r2f_${t}_fused.json 2> $O/r2f_${t}_fused.err
done
timeout 300 $B --circuit random:96 > $O/r2f_bench_random96.json 2> $O/r2f_bench_random96.err
timeout 300 $B --circuit cyclic > $O/r2f_bench_cyclic.json 2> $O/r2f_bench_cyclic.err
for f in $O/r2f_*.json; do python - <<PY
import json
try:
    d=json.load(open("$f"))
    print("$f".split("/")[-1], round(d["ms_per_step"],3), "frac", round(d["roofline"]["frac"],4), "amp_GBps", round(d["amp_GBps"]), d["sweep_ms_first_step"], d["check"]["ok"], d["gpu_launches"])
except Exception as e:
    print("$f", "FAILED", e)
PY
done
