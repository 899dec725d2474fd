#!/usr/bin/env python
"""Prints the handful of raw ncu metrics used in profiles/README.md from an .ncu-rep (first kernel in the report)."""
import csv
import subprocess
import sys

rep = sys.argv[1]
rows = list(csv.reader(subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout.splitlines()))
h, u = rows[0], rows[1]
want = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "launch__registers_per_thread", "launch__cluster_size", "launch__cluster_max_active", "launch__shared_mem_per_block_dynamic",
        "sm__warps_active.avg.per_cycle_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__average_warps_issue_stalled", "lts__t_sector_hit_rate.pct", "sm__inst_executed_pipe_fp64", "smsp__inst_executed.sum "]
for v in rows[2:]:
    for i, n in enumerate(h):
        if any(n.startswith(w) or n == w for w in want) and "pct_of_peak" not in n.replace("smsp__issue_active.avg.pct_of_peak_sustained_active", ""):
            try:
                if n.startswith("smsp__average_warps_issue_stalled") and float(v[i]) < 0.05:
                    continue
            except ValueError:
                pass
            print(f"{n} [{u[i]}] = {v[i]}")
    print("--")
