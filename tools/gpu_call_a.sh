#!/bin/bash
# Round-2 GPU call A (1 GPU): tests, bench lines for the three workload families, configs 3 and 4, ncu evidence.
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/r2a_smi.txt
timeout 900 python -m pytest tests -m gpu -x -q > $O/r2a_pytest_gpu.txt 2>&1; echo "pytest rc=$?" >> $O/r2a_pytest_gpu.txt
tail -3 $O/r2a_pytest_gpu.txt
timeout 600 python bench.py --steps 5 --warmup 3 > $O/r2a_bench_qft30.json 2> $O/r2a_bench_qft30.err; echo "bench rc=$?"
tail -c 1500 $O/r2a_bench_qft30.json
timeout 300 python bench.py --circuit random:96 --steps 3 --warmup 3 --e2e-steps 0 --cpu-gates 0 > $O/r2a_bench_random96.json 2> $O/r2a_bench_random96.err; echo "random rc=$?"
timeout 300 python bench.py --circuit cyclic --steps 3 --warmup 3 --e2e-steps 0 --cpu-gates 0 > $O/r2a_bench_cyclic.json 2> $O/r2a_bench_cyclic.err; echo "cyclic rc=$?"
timeout 300 python tools/config3.py > $O/r2a_config3_qft24x64.json 2> $O/r2a_config3.err; echo "cfg3 rc=$?"
cat $O/r2a_config3_qft24x64.json
timeout 400 python tools/config4.py --shards 1 > $O/r2a_config4_1gpu.json 2> $O/r2a_config4.err; echo "cfg4 rc=$?"
cat $O/r2a_config4_1gpu.json
# ncu: launch list of the shipped FFT image (the same command exited 0 above without ncu, minus the check / CPU legs)
CMD="python bench.py --steps 2 --warmup 3 --cpu-gates 0 --e2e-steps 0 --no-check"
timeout 300 $CMD > $O/r2a_plain_launch.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2a_ncu_launches_fft_qft30.csv $CMD > $O/r2a_ncu_launch.log 2>&1
CMD1="python bench.py --steps 1 --warmup 1 --cpu-gates 0 --e2e-steps 0 --no-check"
timeout 300 $CMD1 > $O/r2a_plain_full.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_sweep -s 5 -c 5 -o $O/r2a_ncu_full_fft_qft30 $CMD1 > $O/r2a_ncu_full.log 2>&1
CMD2="python bench.py --circuit random:96 --steps 1 --warmup 1 --cpu-gates 0 --e2e-steps 0 --no-check"
timeout 300 $CMD2 > $O/r2a_plain_full_random.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_sweep -s 14 -c 3 -o $O/r2a_ncu_full_generic_random96 $CMD2 > $O/r2a_ncu_full_random.log 2>&1
ls -la $O
