#!/bin/bash
# 4 GPUs, final tree: the world-4 peer-memory parity tests and the default bench line
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
timeout 200 python -m pytest tests/test_gpu_sharded.py -x -q -k "4-fused" > $O/r2ah_pytest_sharded_4gpu.txt 2>&1; echo "pytest rc=$?" >> $O/r2ah_pytest_sharded_4gpu.txt; tail -3 $O/r2ah_pytest_sharded_4gpu.txt
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29757"
timeout 300 $T bench.py --gpus 4 --steps 5 --warmup 3 --e2e-steps 8 > $O/r2ah_bench_4gpu.json 2> $O/r2ah_bench_4gpu.err; echo "rc=$?"
python - <<PY
import json
d=json.loads(open("$O/r2ah_bench_4gpu.json").read().strip().splitlines()[-1])
print(round(d["ms_per_step"],3), d["check"]["ok"], d["check"]["max_abs_err"], d["e2e"] and d["e2e"]["ms_per_step"], d["nvlink"]["frac"], d["ms_split_one_step"])
PY
