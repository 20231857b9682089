#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "two_level or rows or rowslab or config4" > $O/r2o_pytest.txt 2>&1; echo "pytest rc=$?" >> $O/r2o_pytest.txt; tail -3 $O/r2o_pytest.txt
timeout 400 python tools/config4.py --as-rank 3/8 > $O/r2o_config4_rank3of8.json 2> $O/r2o_config4.err; echo "rc=$?"; cat $O/r2o_config4_rank3of8.json
timeout 400 python tools/config4.py > $O/r2o_config4_1gpu.json 2>> $O/r2o_config4.err; echo "rc=$?"; cat $O/r2o_config4_1gpu.json
