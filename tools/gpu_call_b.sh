#!/bin/bash
# k_sweep8 (radix-8 / TMA image): parity first, then A/B bench lines against the radix-16 cp.async image.
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q > $O/r2b_pytest_parity.txt 2>&1; echo "pytest rc=$?" >> $O/r2b_pytest_parity.txt
tail -5 $O/r2b_pytest_parity.txt
B="python bench.py --steps 5 --warmup 3 --e2e-steps 0 --cpu-gates 0"
timeout 300 $B > $O/r2b_bench_fft8.json 2> $O/r2b_bench_fft8.err; echo "rc=$?"
QSV_NO_FFT8=1 timeout 300 $B > $O/r2b_bench_fft16.json 2> $O/r2b_bench_fft16.err; echo "rc=$?"
timeout 300 python tools/config3.py > $O/r2b_config3_fft8.json 2> $O/r2b_config3.err; echo "rc=$?"; cat $O/r2b_config3_fft8.json
for f in fft8 fft16; do python - <<PY
import json
try:
    d=json.load(open("$O/r2b_bench_$f.json"))
    print("$f", round(d["ms_per_step"],3), round(d["roofline"]["frac"],4), d["sweep_ms_first_step"], d["check"]["ok"], d["check"]["max_abs_err"], d["check"].get("max_abs_err_vs_ifft_random_state"))
except Exception as e:
    print("$f", "FAILED", e); print(open("$O/r2b_bench_$f.err").read()[-1500:])
PY
done
