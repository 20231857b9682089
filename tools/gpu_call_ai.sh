#!/bin/bash
# 2 GPUs: per-rank plan 7+7+7+8 (largest sweep carries the exchange) against 8+7+7+7 (QSV_NO_MIXED_CAPS=1), plus the fused parity tests
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29759"
timeout 200 $T bench.py --gpus 2 --steps 5 --warmup 3 --e2e-steps 0 > $O/r2ai_bench_2gpu.json 2> $O/r2ai_bench_2gpu.err; echo "rc=$?"
QSV_NO_MIXED_CAPS=1 timeout 200 $T bench.py --gpus 2 --steps 5 --warmup 3 --e2e-steps 0 > $O/r2ai_bench_2gpu_balanced.json 2> $O/r2ai_bench_2gpu_balanced.err; echo "rc=$?"
python - <<PY
import json
for f in ("r2ai_bench_2gpu.json","r2ai_bench_2gpu_balanced.json"):
    d=json.loads(open("$O/"+f).read().strip().splitlines()[-1])
    print(f, round(d["ms_per_step"],3), d["check"]["ok"], d["check"]["max_abs_err"], d["nvlink"]["frac"], d["ms_split_one_step"])
PY
timeout 150 python -m pytest tests/test_gpu_sharded.py -x -q -k "2-fused" > $O/r2ai_pytest_sharded_2gpu.txt 2>&1; tail -2 $O/r2ai_pytest_sharded_2gpu.txt
