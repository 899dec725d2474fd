#!/usr/bin/env python
"""Text summary of the first kernel in an .ncu-rep: the raw metrics quoted in profiles/README.md plus the warp-stall
sampling breakdown of the source page.   python tools/ncu_summary.py report.ncu-rep ["header line"] > profiles/rNN_x.txt"""
import collections
import csv
import subprocess
import sys

METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_read.sum.per_second", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "launch__grid_size",
    "launch__block_size", "launch__cluster_size", "launch__cluster_max_active", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static", "smsp__inst_executed.sum",
    "sm__inst_executed.avg.per_cycle_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__cycles_elapsed.avg", "sm__cycles_elapsed.avg.per_second",
]


def page(rep, name):
    out = subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(out.splitlines()))


def main():
    rep = sys.argv[1]
    if len(sys.argv) > 2:
        print("# " + sys.argv[2])
    raw = page(rep, "raw")
    hdr, units, vals = raw[0], raw[1], raw[2]
    col = {n: i for i, n in enumerate(hdr)}
    print(f"{'Kernel Name':85s} {vals[col['Kernel Name']]}")
    for m in METRICS:
        if m in col:
            print(f"{m:85s} {vals[col[m]]} {units[col[m]]}")
    src = page(rep, "source")
    h = src[1]
    stalls = collections.Counter()
    for r in src[2:]:
        for i, n in enumerate(h):
            if n.startswith("stall_") and "Not Issued" not in n and i < len(r) and r[i].isdigit():
                stalls[n[6:]] += int(r[i])
    tot = sum(stalls.values())
    if tot:
        print("warp stall samples (all): " + ", ".join(f"{k} {100.0 * v / tot:.1f}%" for k, v in stalls.most_common(8)))


if __name__ == "__main__":
    main()
