#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
CMD="python bench.py --circuit random:96 --steps 1 --warmup 1 --cpu-gates 0 --e2e-steps 0 --no-check"
timeout 300 $CMD > $O/r2m_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_sweep -s 12 -c 3 -o $O/r2m_ncu_generic16_random96 $CMD > $O/r2m_ncu.log 2>&1
export QSV_COST_MODEL=1
timeout 300 $CMD > $O/r2m_plain2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_sweep8 -s 14 -c 3 -o $O/r2m_ncu_generic8_random96 $CMD > $O/r2m_ncu2.log 2>&1
ls -la $O | grep r2m
