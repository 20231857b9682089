#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
T8="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29731"
timeout 900 $T8 bench.py --gpus 8 --steps 5 --warmup 3 --e2e-steps 2 > $O/r2h_bench_8gpu.json 2> $O/r2h_bench_8gpu.err; echo "rc=$?"
python - <<PY
import json
try:
    d=json.loads(open("$O/r2h_bench_8gpu.json").read().strip().splitlines()[-1])
    print("8gpu", round(d["ms_per_step"],2), "check", d["check"]["ok"], "e2e", d["e2e"] and round(d["e2e"]["ms_per_step"],1), "nvlink", d["nvlink"]["frac"], d["impl_config"]["exchange"][:20], d["ms_split_one_step"])
    q=d["qft34"]; print("  qft34", round(q["ms_per_step"],1), q["check"]["ok"], q["check"]["max_abs_err"], q["ms_split_one_step"], q["local_sweep_roofline"]["frac"], q["nvlink"]["frac"], q["impl_config"]["exchange"][:20])
except Exception as e:
    print("FAILED", e); print(open("$O/r2h_bench_8gpu.err").read()[-3000:])
PY
