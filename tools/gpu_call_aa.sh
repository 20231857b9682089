#!/bin/bash
# 2-GPU re-check after the scheduler / generic-loop changes: sharded parity tests + the default bench line
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_sharded.py -x -q > $O/r2aa_pytest_sharded_2gpu.txt 2>&1; echo "pytest rc=$?" >> $O/r2aa_pytest_sharded_2gpu.txt; tail -3 $O/r2aa_pytest_sharded_2gpu.txt
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29743"
timeout 600 $T bench.py --gpus 2 --steps 5 --warmup 3 > $O/r2aa_bench_2gpu.json 2> $O/r2aa_bench_2gpu.err; echo "rc=$?"
python - <<PY
import json
d=json.loads(open("$O/r2aa_bench_2gpu.json").read().strip().splitlines()[-1])
print(round(d["ms_per_step"],2), d["check"]["ok"], d["check"]["max_abs_err"], d["e2e"]["ms_per_step"], d["nvlink"]["frac"], d["ms_split_one_step"])
PY
