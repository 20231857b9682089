#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
/usr/bin/time -v timeout 400 python tools/config4.py > $O/r2u_config4_1gpu.json 2> $O/r2u_config4.err; echo "rc=$?"; cat $O/r2u_config4_1gpu.json; grep -E "Elapsed|Maximum resident" $O/r2u_config4.err
timeout 400 python tools/config4.py --as-rank 5/8 > $O/r2u_config4_rank5of8.json 2>> $O/r2u_config4.err; echo "rc=$?"; cat $O/r2u_config4_rank5of8.json
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "config4 or rowslab or repeat_bit" > $O/r2u_pytest.txt 2>&1; tail -2 $O/r2u_pytest.txt
