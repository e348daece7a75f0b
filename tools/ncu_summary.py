#!/usr/bin/env python
"""Summarise `ncu -i X.ncu-rep --page raw --csv` output: the metrics DESIGN.md / bench.py cite.

    ncu -i gpurun_out/prof.ncu-rep --page raw --csv > raw.csv ; python tools/ncu_summary.py raw.csv
"""
import csv
import sys

KEYS = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct",
        "l1tex__t_sector_hit_rate.pct", "l1tex__t_bytes.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__shared_mem_per_block_static",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
        "smsp__warps_eligible.avg.per_cycle_active", "smsp__average_warp_latency_per_inst_issued.ratio",
        "sm__inst_executed_pipe_lsu.sum", "sm__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_alu.sum",
        "smsp__cycles_active.avg", "sm__cycles_elapsed.max"]


def main(path):
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    for i, h in enumerate(hdr):
        stall = ("issue_stalled" in h and h.endswith("per_issue_active.ratio")) or ("warps_issue_stalled" in h and h.endswith("_per_warp_active.pct"))
        if h in KEYS or stall:
            vals = [r[i] for r in data]
            if stall and all(float(v.replace(",", "") or 0) < 0.15 for v in vals):
                continue
            print(f"{h:86s} {units[i]:14s} " + " | ".join(v[:48] for v in vals))


if __name__ == "__main__":
    main(sys.argv[1])
