#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q > $O/r2i_pytest_gpu.txt 2>&1; echo "pytest rc=$?" >> $O/r2i_pytest_gpu.txt; tail -4 $O/r2i_pytest_gpu.txt
timeout 400 python tools/config4.py > $O/r2i_config4_1gpu.json 2> $O/r2i_config4.err; echo "cfg4 rc=$?"; cat $O/r2i_config4_1gpu.json
QSV_NO_ROWSLAB=1 timeout 100 python -c "print('skip')"
timeout 600 python bench.py --steps 5 --warmup 3 > $O/r2i_bench_qft30.json 2> $O/r2i_bench_qft30.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open("$O/r2i_bench_qft30.json"))
print(round(d["ms_per_step"],3), d["roofline"]["frac"], d["sweep_ms_first_step"], d["e2e"]["ms_per_step"], d["check"], d["cpu_baseline"], d["clocks"])
PY
CMD="python bench.py --steps 2 --warmup 3 --cpu-gates 0 --e2e-steps 0 --no-check"
timeout 300 $CMD > $O/r2i_plain_launch.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2i_ncu_launches_qft30.csv $CMD > $O/r2i_ncu_launch.log 2>&1
