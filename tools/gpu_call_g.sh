#!/bin/bash
# 8-GPU call: sharded parity tests on 4 and 8 GPUs, bench lines on 8 and 4 GPUs (QFT-30 + the QFT-34 block), config 4 column-sharded
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_sharded.py -x -q -k "4- or 8-" > $O/r2g_pytest_sharded_8gpu.txt 2>&1; echo "pytest rc=$?" >> $O/r2g_pytest_sharded_8gpu.txt; tail -4 $O/r2g_pytest_sharded_8gpu.txt
T8="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29721"
T4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29722"
timeout 900 $T8 bench.py --gpus 8 --steps 5 --warmup 3 > $O/r2g_bench_8gpu.json 2> $O/r2g_bench_8gpu.err; echo "rc=$?"
timeout 600 $T8 bench.py --gpus 8 --steps 5 --warmup 3 --exchange fused --no-qft34 --e2e-steps 0 > $O/r2g_bench_8gpu_fused.json 2> $O/r2g_bench_8gpu_fused.err; echo "rc=$?"
timeout 600 $T4 bench.py --gpus 4 --steps 5 --warmup 3 > $O/r2g_bench_4gpu.json 2> $O/r2g_bench_4gpu.err; echo "rc=$?"
timeout 600 $T4 bench.py --gpus 4 --steps 5 --warmup 3 --exchange nccl --e2e-steps 0 > $O/r2g_bench_4gpu_nccl.json 2> $O/r2g_bench_4gpu_nccl.err; echo "rc=$?"
timeout 600 $T8 tools/config4.py > $O/r2g_config4_8gpu.json 2> $O/r2g_config4_8gpu.err; echo "rc=$?"; cat $O/r2g_config4_8gpu.json
for f in 8gpu 8gpu_fused 4gpu 4gpu_nccl; do python - <<PY
import json
try:
    d=json.loads(open("$O/r2g_bench_$f.json").read().strip().splitlines()[-1])
    print("$f", round(d["ms_per_step"],2), "check", d["check"]["ok"], d["check"]["max_abs_err"], "e2e", d["e2e"] and round(d["e2e"]["ms_per_step"],1), "nvlink", d["nvlink"]["frac"], d["impl_config"]["exchange"][:20], d["ms_split_one_step"])
    if "qft34" in d:
        q=d["qft34"]; print("  qft34", round(q["ms_per_step"],1), q["check"], q["ms_split_one_step"], q["local_sweep_roofline"]["frac"], q["nvlink"]["frac"])
except Exception as e:
    print("$f", "FAILED", e); print(open("$O/r2g_bench_$f.err").read()[-2500:])
PY
done
