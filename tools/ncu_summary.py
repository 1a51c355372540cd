#!/usr/bin/env python3
"""Per-kernel summary of an `ncu --set full` report: the last captured launch of every kernel, one column each.

    ncu -i REPORT.ncu-rep --page raw --csv > raw.csv ; python tools/ncu_summary.py raw.csv > profiles/NAME.csv
"""
import csv
import re
import sys

METRICS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__grid_size", "launch__block_size",
           "launch__registers_per_thread", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
           "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
           "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
           "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
           "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct"]
STALLS = ["wait", "math_pipe_throttle", "short_scoreboard", "long_scoreboard", "barrier", "not_selected", "no_instruction"]
METRICS += [f"smsp__average_warps_issue_stalled_{s}_per_issue_active.ratio" for s in STALLS]

rows = list(csv.reader(open(sys.argv[1])))
hdr, units, body = rows[0], rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
last = {}
for r in body:
    name = re.sub(r"^void |<unnamed>::|\(.*$|\(int\)|\(bool\)", "", r[ix["Kernel Name"]])
    last[name] = r
names = list(last)
w = csv.writer(sys.stdout)
w.writerow(["metric", "unit"] + names)
for m in METRICS:
    if m in ix:
        w.writerow([m, units[ix[m]]] + [last[n][ix[m]] for n in names])
