#!/usr/bin/env python
"""Condenses `ncu -i X.ncu-rep --page raw --csv` into the handful of metrics DESIGN.md quotes.
usage: ncu_summary.py raw.csv"""
import csv
import sys

KEYS = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__occupancy_limit", "launch__grid_size", "launch__block_size",
        "sm__warps_active.avg.per_cycle_active", "smsp__issue_active.avg.per_cycle_active", "smsp__inst_executed.sum ",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__pipe_fmaheavy_cycles_active.avg", "sm__throughput.avg.pct",
        "dram__bytes_read.sum ", "dram__bytes_write.sum ", "dram__bytes_read.sum.per_second", "dram__throughput.avg.pct", "lts__t_sector_hit_rate.pct",
        "l1tex__t_sector_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__average_warps_issue_stalled",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sass__inst_executed_local", "launch__shared_mem_per_block"]
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
for vals in rows[2:]:
    name = vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?"
    print("kernel:", name)
    for h, u, v in zip(hdr, units, vals):
        if any(k in h + " " for k in KEYS) and v not in ("0", "", "n/a"):
            print(f"  {h} [{u}] = {v}")
