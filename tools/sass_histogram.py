#!/usr/bin/env python
"""Opcode histogram per kernel of the shipped library (cuobjdump -sass, sm_100a cubins): which instructions move the
tiles (UTMALDG / UTMASTG = tensor-map TMA, LDGSTS = cp.async, LDG / STG, LD / ST = generic-address accesses such as the
stores into peer buffers), which exchange lanes (SHFL), spills (LDL / STL),
FP64 work (DFMA / DMUL / DADD) and barriers.

  python tools/sass_histogram.py [quantum-simulations_b200/libqsv.so] > profiles/r2_sass_histogram.txt
"""
import collections
import re
import subprocess
import sys

WATCH = ["UTMALDG", "UTMASTG", "UTMAPF", "UBLKCP", "LDGSTS", "LDG", "STG", "LD", "ST", "LDS", "STS", "SHFL", "LDL", "STL", "DFMA", "DMUL", "DADD",
         "BAR", "SYNCS", "LDC", "LDCU", "BRA"]


def main():
    lib = sys.argv[1] if len(sys.argv) > 1 else "quantum-simulations_b200/libqsv.so"
    sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
    demangle = lambda n: subprocess.run(["cu++filt", n], capture_output=True, text=True).stdout.strip() or n
    kernels = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = kernels.setdefault(m.group(1), collections.Counter())
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m and cur is not None:
            cur[m.group(1)] += 1
    print(f"# {lib}: SASS opcode counts per kernel (static instruction counts; base mnemonic before the first '.')")
    print("# " + " ".join(f"{w:>7}" for w in ["total"] + WATCH) + "  kernel")
    for name, c in kernels.items():
        full = demangle(name)
        m = re.match(r"^(?:void )?(\w+<.*?>)\((?!int\))", full) or re.match(r"^(?:void )?(\w+)\(", full)
        short = m.group(1) if m else full[:80]
        row = [sum(c.values())] + [c.get(w, 0) for w in WATCH]
        print("  " + " ".join(f"{v:>7}" for v in row) + "  " + short)


if __name__ == "__main__":
    main()
