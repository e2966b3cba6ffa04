#!/usr/bin/env python
"""Condenses `ncu -i <rep> --page raw --csv` into the handful of counters DESIGN.md quotes.
usage: ncu -i X.ncu-rep --page raw --csv | python tools/ncu_summary.py > profiles/<tag>_ncu_full_summary.csv"""
import csv
import re
import sys

KEEP = [r"^Kernel Name$", r"^Grid Size$", r"^Block Size$", r"^gpu__time_duration\.sum$", r"^launch__registers_per_thread$",
        r"^launch__occupancy_limit_", r"^sm__warps_active\.avg\.pct_of_peak_sustained_active$",
        r"^sm__throughput\.avg\.pct", r"^smsp__inst_executed\.sum$", r"^smsp__issue_active\.avg\.(pct|per_cycle_active)",
        r"^sm__inst_executed_pipe_(fp64|alu|fma|lsu|xu|cbu|adu)\.avg\.pct_of_peak_sustained_active$",
        r"^sm__pipe_fp64_cycles_active\.avg\.pct_of_peak_sustained_(active|elapsed)$",
        r"^smsp__sass_thread_inst_executed_op_d(add|mul|fma)_pred_on\.sum(\.per_cycle_elapsed)?$",
        r"^smsp__sass_inst_executed_op_(global|local|shared)_(ld|st)\.sum$",
        r"^dram__bytes_(read|write)\.sum$", r"^dram__throughput\.avg\.pct", r"^lts__t_sector_hit_rate\.pct$",
        r"^l1tex__t_sector_hit_rate\.pct$", r"^sm__cycles_elapsed\.max$", r"^smsp__cycles_active\.avg$",
        r"^smsp__average_warps_issue_stalled_.*_per_issue_active\.ratio$", r"^smsp__inst_executed_op_branch\.sum$"]
rows = list(csv.reader(sys.stdin))
hdr, units = rows[0], rows[1]
w = csv.writer(sys.stdout)
w.writerow(["launch", "metric", "unit", "value"])
for li, vals in enumerate(rows[2:]):
    for i, h in enumerate(hdr):
        if any(re.search(p, h) for p in KEEP):
            w.writerow([li, h, units[i], vals[i]])
