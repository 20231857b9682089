#!/usr/bin/env python
"""Join an `ncu --page source --csv` SASS listing (per-instruction executed counts and stall samples) with
`nvdisasm -g -c` line information of the same kernel: executed instructions and samples per source line.

  ncu -i X.ncu-rep --page source --csv --kernel-name K --launch-skip 1 --launch-count 1 > SRC.csv
  cuobjdump -xelf all libqsv.so ; nvdisasm -g -c <cubin with K> > DIS.txt
  python tools/sass_lines.py SRC.csv DIS.txt [top]
"""
import collections
import csv
import re
import sys


def main():
    src, dis = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    rows = list(csv.reader(open(src)))
    hdr, data = rows[1], rows[2:]
    ix = {h: i for i, h in enumerate(hdr)}
    cur = ("?", 0)
    lines = []
    for ln in open(dis):
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s", ln):
            lines.append(cur)
    n = min(len(lines), len(data))
    if len(lines) != len(data):
        print(f"warning: {len(lines)} disassembled vs {len(data)} profiled instructions", file=sys.stderr)
    inst, samp, ops, osamp = collections.Counter(), collections.Counter(), collections.Counter(), collections.Counter()
    ti = ts = 0
    for i in range(n):
        e = int(data[i][ix["Instructions Executed"]] or 0)
        s = int(data[i][ix["# Samples"]] or 0)
        inst[lines[i]] += e
        samp[lines[i]] += s
        m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_]+)", data[i][ix["Source"]])
        o = m.group(2) if m else "?"
        ops[o] += e
        osamp[o] += s
        ti += e
        ts += s
    print(f"total warp-instructions {ti}, samples {ts}")
    print("opcodes:", ", ".join(f"{o} {c / ti:.1%}/{osamp[o] / max(ts, 1):.1%}" for o, c in ops.most_common(16)))
    stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    tot = {h: sum(int(r[ix[h]] or 0) for r in data) for h in stalls}
    print("stall samples:", ", ".join(f"{h[6:]} {v / max(ts, 1):.1%}" for h, v in sorted(tot.items(), key=lambda x: -x[1])[:8]))
    for key, e in inst.most_common(top):
        print(f"{key[0]:22s} {key[1]:5d}  inst {e / ti:6.2%}  samples {samp[key] / max(ts, 1):6.2%}")


if __name__ == "__main__":
    main()
