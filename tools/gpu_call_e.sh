#!/bin/bash
# 2-GPU call: sharded parity tests + bench lines (fused and NCCL exchange engines)
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
nvidia-smi topo -m > $O/r2e_topo.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_sharded.py -x -q > $O/r2e_pytest_sharded_2gpu.txt 2>&1; echo "pytest rc=$?" >> $O/r2e_pytest_sharded_2gpu.txt; tail -5 $O/r2e_pytest_sharded_2gpu.txt
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29711"
timeout 600 $T bench.py --gpus 2 --steps 5 --warmup 3 > $O/r2e_bench_2gpu.json 2> $O/r2e_bench_2gpu.err; echo "rc=$?"
timeout 600 $T bench.py --gpus 2 --steps 5 --warmup 3 --exchange nccl --e2e-steps 0 > $O/r2e_bench_2gpu_nccl.json 2> $O/r2e_bench_2gpu_nccl.err; echo "rc=$?"
timeout 600 $T bench.py --gpus 2 --steps 5 --warmup 3 --restore-layout --e2e-steps 0 > $O/r2e_bench_2gpu_restore.json 2> $O/r2e_bench_2gpu_restore.err; echo "rc=$?"
timeout 600 $T bench.py --gpus 2 --steps 5 --warmup 3 --no-merge --e2e-steps 0 > $O/r2e_bench_2gpu_nomerge.json 2> $O/r2e_bench_2gpu_nomerge.err; echo "rc=$?"
timeout 300 $T bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > $O/r2e_bench_ref_2gpu.json 2> $O/r2e_bench_ref_2gpu.err; echo "rc=$?"
for f in 2gpu 2gpu_nccl 2gpu_restore 2gpu_nomerge; do python - <<PY
import json
try:
    d=json.loads(open("$O/r2e_bench_$f.json").read().strip().splitlines()[-1])
    print("$f", round(d["ms_per_step"],3), d["ms_split_one_step"], d["check"], d["e2e"] and d["e2e"]["ms_per_step"], d["nvlink"]["frac"], d["local_sweep_roofline"]["frac"])
except Exception as e:
    print("$f", "FAILED", e); print(open("$O/r2e_bench_$f.err").read()[-2500:])
PY
done
tail -c 600 $O/r2e_bench_ref_2gpu.json
