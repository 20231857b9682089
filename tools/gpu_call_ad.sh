#!/bin/bash
# A/B of the mixed-cap plans (first m sweeps held to six targets = radix-8 image) on one GPU: QFT-31 and QFT-32/33 (32-128 GiB states
# do not all fit: 31 only), random:96, plus the GPU parity tests with the new planner
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
for v in mixed nomixed; do
  if [ $v = mixed ]; then export QSV_MIXED_CAPS=1; else unset QSV_MIXED_CAPS; fi
  timeout 300 python bench.py --qubits 31 --steps 3 --warmup 3 --e2e-steps 0 --cpu-gates 0 > $O/r2ad_qft31_$v.json 2> $O/r2ad_qft31_$v.err; echo "rc=$?"
  timeout 300 python bench.py --circuit random:96 --steps 3 --warmup 3 --e2e-steps 0 --cpu-gates 0 > $O/r2ad_random96_$v.json 2> $O/r2ad_random96_$v.err; echo "rc=$?"
  timeout 300 python bench.py --circuit random:400:7 --steps 2 --warmup 3 --e2e-steps 0 --cpu-gates 0 > $O/r2ad_random400_$v.json 2> $O/r2ad_random400_$v.err; echo "rc=$?"
done
export QSV_MIXED_CAPS=1
python - <<PY
import json
for w in ("qft31","random96","random400"):
    for v in ("mixed","nomixed"):
        try:
            d=json.load(open("$O/r2ad_%s_%s.json"%(w,v)))
            print(w, v, round(d["ms_per_step"],3), d["roofline"]["frac"], d["check"]["ok"], d.get("sweep_ms_first_step"))
        except Exception as e:
            print(w, v, "unreadable", e)
PY
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sharded_one_device.py -m gpu -x -q > $O/r2ad_pytest.txt 2>&1; tail -3 $O/r2ad_pytest.txt
