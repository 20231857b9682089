#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "two_level or rows or producer or singlet or slab or stream_golden" > $O/r2j_pytest.txt 2>&1; echo "pytest rc=$?" >> $O/r2j_pytest.txt; tail -3 $O/r2j_pytest.txt
timeout 400 python tools/config4.py > $O/r2j_config4_1gpu.json 2> $O/r2j_config4.err; echo "cfg4 rc=$?"; cat $O/r2j_config4_1gpu.json; tail -3 $O/r2j_config4.err
