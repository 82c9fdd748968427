#!/usr/bin/env python
"""Condense an .ncu-rep (one kernel, --set full) into the metrics DESIGN.md quotes:
   python tools/ncu_summary.py report.ncu-rep out.csv"""
import csv
import io
import subprocess
import sys

KEEP = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "smsp__thread_inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__icc_request_hit_rate.pct",
        "gcc__cache_requests_type_instruction.sum.pct_of_peak_sustained_elapsed"]
txt = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr, units = rows[0], rows[1]
with open(sys.argv[2], "w", newline="") as f:
    w = csv.writer(f)
    w.writerow(["kernel", "metric", "unit", "value"])
    for vals in rows[2:]:
        d = dict(zip(hdr, vals)); u = dict(zip(hdr, units))
        for k in hdr:
            if k in KEEP or ("issue_stalled" in k and k.endswith("_per_issue_active.ratio")) or \
               (k.startswith("sm__inst_executed_pipe_") and k.endswith(".avg.pct_of_peak_sustained_active")):
                w.writerow([d.get("Kernel Name", ""), k, u[k], d[k]])
print("wrote", sys.argv[2])
