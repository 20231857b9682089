#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "fft_shaped or qft_matches or large_state or complex64_fused or fuzz" > $O/r2c_pytest.txt 2>&1; echo "pytest rc=$?" >> $O/r2c_pytest.txt; tail -3 $O/r2c_pytest.txt
B="python bench.py --steps 5 --warmup 3 --e2e-steps 0 --cpu-gates 0"
timeout 300 $B > $O/r2c_bench_fft8_mid.json 2> $O/r2c_bench_fft8_mid.err; echo "rc=$?"
QSV_NO_MIDLAYOUT=1 timeout 300 $B > $O/r2c_bench_fft8_nomid.json 2> $O/r2c_bench_fft8_nomid.err; echo "rc=$?"
QSV_NO_FFT8=1 timeout 300 $B > $O/r2c_bench_fft16.json 2> $O/r2c_bench_fft16.err; echo "rc=$?"
for f in fft8_mid fft8_nomid fft16; do python - <<PY
import json
try:
    d=json.load(open("$O/r2c_bench_$f.json"))
    print("$f", round(d["ms_per_step"],3), round(d["roofline"]["frac"],4), d["sweep_ms_first_step"], d["check"]["ok"], d["check"]["max_abs_err"], d["clocks"])
except Exception as e:
    print("$f", "FAILED", e); print(open("$O/r2c_bench_$f.err").read()[-1500:])
PY
done
CMD1="python bench.py --steps 1 --warmup 1 --cpu-gates 0 --e2e-steps 0 --no-check"
timeout 300 $CMD1 > $O/r2c_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_sweep -s 5 -c 5 -o $O/r2c_ncu_full_fft8 $CMD1 > $O/r2c_ncu.log 2>&1
ls -la $O | tail -3
