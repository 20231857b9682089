#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
timeout 120 python tools/sanitize_small.py > $O/r2s_plain.txt 2>&1; echo "plain rc=$?"; cat $O/r2s_plain.txt
timeout 400 compute-sanitizer --tool racecheck --print-limit 20 python tools/sanitize_small.py > $O/r2s_racecheck.txt 2>&1; echo "racecheck rc=$?"; tail -15 $O/r2s_racecheck.txt
nvidia-smi --query-gpu=name,memory.used --format=csv
