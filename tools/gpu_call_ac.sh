#!/bin/bash
# final tree, one GPU: smoke, the whole GPU suite, both bench arms, then (after those exited 0) the ncu launch list and one
# --set full capture of the five sweeps of a QFT-30 step
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
python -c "import __graft_entry__ as g; g.smoke()" > $O/r2ac_smoke.txt 2>&1; echo "smoke rc=$?"; tail -2 $O/r2ac_smoke.txt
timeout 1200 python -m pytest tests -m gpu -x -q > $O/r2ac_pytest_gpu.txt 2>&1; echo "pytest rc=$?" >> $O/r2ac_pytest_gpu.txt; tail -3 $O/r2ac_pytest_gpu.txt
timeout 600 python bench.py > $O/r2ac_bench_default.json 2> $O/r2ac_bench_default.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference > $O/r2ac_bench_reference.json 2> $O/r2ac_bench_reference.err; echo "ref rc=$?"
python - <<PY
import json
d=json.load(open("$O/r2ac_bench_default.json")); r=json.load(open("$O/r2ac_bench_reference.json"))
print(round(d["ms_per_step"],3), d["value"], d["roofline"]["frac"], d["e2e"]["value"], d["e2e"]["ms_per_step"], d["check"]["ok"], d["config"]==r["config"], r["value"], r["cpu_baseline"]["cores"])
print("e2e ratio", d["e2e"]["value"]/r["value"], "device ratio", d["value"]/r["value"])
PY
CMD="python bench.py --steps 2 --warmup 3 --cpu-gates 0 --e2e-steps 0 --no-check"
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2ac_ncu_launches_qft30.csv $CMD > $O/r2ac_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_sweep8 -s 15 -c 5 -o $O/r2ac_ncu_full_sweep8_qft30 $CMD > $O/r2ac_ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la $O/r2ac_ncu_full_sweep8_qft30.ncu-rep
