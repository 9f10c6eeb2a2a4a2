#!/usr/bin/env python
"""Text summary of an `ncu --set full` report for profiles/: one block per kernel (last captured launch of each name)
with the metrics DESIGN.md argues from.  usage: python tools/ncu_summary.py report.ncu-rep "<what was captured>" > profiles/x.txt"""
import csv
import io
import subprocess
import sys

METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
    "l1tex__throughput.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "launch__grid_size", "launch__registers_per_thread",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "smsp__warps_eligible.avg.per_cycle_active",
]


def main():
    rep, what = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "")
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    ik = hdr.index("Kernel Name")
    last = {}
    count = {}
    for r in rows[2:]:
        last[r[ik]] = r
        count[r[ik]] = count.get(r[ik], 0) + 1
    print(f"ncu --set full --clock-control none: {what}")
    for name, r in last.items():
        print("=====")
        print(f"{name[:110]}   (x{count[name]} captured, last shown)")
        for i, c in enumerate(hdr):
            stall = "issue_stalled" in c and c.endswith("per_issue_active.ratio")
            if c in METRICS or stall:
                try:
                    v = float(r[i].replace(",", ""))
                except ValueError:
                    continue
                if stall and v < 0.2:
                    continue
                print(f"{c:90s} {v:.5g} [{units[i]}]")


if __name__ == "__main__":
    main()
