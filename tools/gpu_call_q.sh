#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
CMD="python tools/config4.py --steps 1"
timeout 300 $CMD > $O/r2q_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_rowlevels -s 1 -c 1 -o $O/r2q_ncu_full_rowlevels_config4 $CMD > $O/r2q_ncu.log 2>&1
CMD2="python bench.py --steps 2 --warmup 3 --cpu-gates 0 --e2e-steps 0 --no-check"
timeout 300 $CMD2 > $O/r2q_plain_launch.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2q_ncu_launches_qft30.csv $CMD2 > $O/r2q_ncu_launch.log 2>&1
timeout 600 python bench.py --steps 5 --warmup 3 > $O/r2q_bench_qft30.json 2> $O/r2q_bench_qft30.err; echo "bench rc=$?"
tail -c 600 $O/r2q_bench_qft30.json
