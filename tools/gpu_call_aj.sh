#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
python -c "import __graft_entry__ as g; g.smoke()" > $O/r2aj_smoke.txt 2>&1; echo "smoke rc=$?"; tail -1 $O/r2aj_smoke.txt
timeout 300 python -m pytest tests -m gpu -x -q > $O/r2aj_pytest_gpu.txt 2>&1; echo "pytest rc=$?" >> $O/r2aj_pytest_gpu.txt; tail -2 $O/r2aj_pytest_gpu.txt
