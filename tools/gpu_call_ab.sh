#!/bin/bash
# 8 GPUs, final tree: the world-8 peer-memory parity tests, the default bench line (QFT-30 + the QFT-34 block) and an A/B of
# the exchange-carrying sweep's image
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_sharded.py -x -q -k "8-fused" > $O/r2ab_pytest_sharded_8gpu.txt 2>&1; echo "pytest rc=$?" >> $O/r2ab_pytest_sharded_8gpu.txt; tail -3 $O/r2ab_pytest_sharded_8gpu.txt
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29751"
timeout 420 $T bench.py --gpus 8 --steps 5 --warmup 3 > $O/r2ab_bench_8gpu.json 2> $O/r2ab_bench_8gpu.err; echo "rc=$?"
tail -3 $O/r2ab_bench_8gpu.err
QSV_NO_CARRY8=1 timeout 240 $T bench.py --gpus 8 --steps 5 --warmup 3 --e2e-steps 0 > $O/r2ab_bench_8gpu_nocarry8.json 2> $O/r2ab_bench_8gpu_nocarry8.err; echo "rc=$?"
python - <<PY
import json
for f in ("r2ab_bench_8gpu.json", "r2ab_bench_8gpu_nocarry8.json"):
    try:
        d=json.loads(open("$O/"+f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "unreadable", e); continue
    print(f, round(d["ms_per_step"],3), d["check"]["ok"], d["check"]["max_abs_err"], d["e2e"] and d["e2e"]["ms_per_step"], d["nvlink"]["frac"], d["ms_split_one_step"])
    q=d.get("qft34")
    if q: print("   qft34", round(q["ms_per_step"],2), q["check"]["ok"], q["check"]["max_abs_err"], q["ms_split_one_step"])
PY
