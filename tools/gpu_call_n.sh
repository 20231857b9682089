#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
mkdir -p $O
B="python bench.py --steps 3 --warmup 2 --e2e-steps 0 --cpu-gates 0"
timeout 600 $B --qubits 31 > $O/r2n_qft31.json 2> $O/r2n_qft31.err; echo rc=$?
timeout 600 $B --qubits 29 > $O/r2n_qft29.json 2> $O/r2n_qft29.err; echo rc=$?
QSV_PLAN_DEBUG=1 timeout 600 $B --qubits 31 --circuit hlayer:25:31 > $O/r2n_h31_hi.json 2> $O/r2n_h31_hi.err; echo rc=$?
timeout 600 $B --qubits 31 --circuit hlayer:5:11 > $O/r2n_h31_lo.json 2> $O/r2n_h31_lo.err; echo rc=$?
timeout 600 $B --qubits 30 --circuit hlayer:24:30 > $O/r2n_h30_hi.json 2> $O/r2n_h30_hi.err; echo rc=$?
for f in qft31 qft29 h31_hi h31_lo h30_hi; do python - <<PY
import json
try:
    d=json.load(open("$O/r2n_$f.json"))
    print("$f", round(d["ms_per_step"],3), "frac", round(d["roofline"]["frac"],4), d["sweep_ms_first_step"], d["check"]["ok"])
except Exception as e:
    print("$f", "FAILED", e); print(open("$O/r2n_$f.err").read()[-800:])
PY
done
grep "qsv" $O/r2n_h31_hi.err | head
