#!/bin/bash
# 8 GPUs, final tree with the mixed-cap plans: the default bench line (QFT-30 + the QFT-34 block)
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29753"
timeout 420 $T bench.py --gpus 8 --steps 5 --warmup 3 --e2e-steps 4 > $O/r2ae_bench_8gpu.json 2> $O/r2ae_bench_8gpu.err; echo "rc=$?"
python - <<PY
import json
d=json.loads(open("$O/r2ae_bench_8gpu.json").read().strip().splitlines()[-1])
print(round(d["ms_per_step"],3), d["check"]["ok"], d["check"]["max_abs_err"], d["e2e"] and d["e2e"]["ms_per_step"], d["nvlink"]["frac"], d["ms_split_one_step"])
q=d.get("qft34")
if q: print("   qft34", round(q["ms_per_step"],2), q["check"]["ok"], q["check"]["max_abs_err"], q["local_sweep_roofline"]["frac"], q["ms_split_one_step"])
PY
