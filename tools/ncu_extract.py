#!/usr/bin/env python
"""Key metrics of an `ncu --set full` report as a small CSV (one column per captured launch), for profiles/.

  python tools/ncu_extract.py gpurun_out/X.ncu-rep profiles/X.csv
"""
import csv
import subprocess
import sys

WANT = ["Kernel Name", "Block Size", "Grid Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "launch__shared_mem_per_block_dynamic", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "l1tex__throughput.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "lts__t_sector_hit_rate.pct"]
STALLS = "smsp__average_warps_issue_stalled_%s_per_issue_active.ratio"
for s in ("wait", "barrier", "short_scoreboard", "long_scoreboard", "mio_throttle", "lg_throttle", "math_pipe_throttle",
          "not_selected", "no_instruction", "dispatch_stall", "branch_resolving", "membar", "imc_miss", "tex_throttle", "sleeping"):
    WANT.append(STALLS % s)


def main():
    rep, out = sys.argv[1], sys.argv[2]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    with open(out, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["metric", "unit"] + [f"launch{i}" for i in range(len(data))])
        for name in WANT:
            if name in hdr:
                i = hdr.index(name)
                w.writerow([name, units[i]] + [r[i] for r in data])


if __name__ == "__main__":
    main()
