#!/usr/bin/env python
"""DRAM-side rate of the fused sweep by cluster size, one GPU: for each row length V a shard of --gb gigabytes is
filled once per candidate cluster size C (SART_FUSED_CLUSTER=C), solved for a few fixed iterations with the whole-solve
kernel, and the sweep time, HBM GB/s, resident clusters and SMs streaming are printed beside the plan's own choice.

    python tools/cluster_table.py --nvoxel 200000 --gb 40 [--clusters 14,15,16]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package  # noqa: E402
from bench import load_module  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nvoxel", type=int, required=True)
    ap.add_argument("--gb", type=float, default=40.0)
    ap.add_argument("--clusters", default="")
    ap.add_argument("--iterations", type=int, default=5)
    ap.add_argument("--steps", type=int, default=2)
    args = ap.parse_args()
    sart = load_package()
    syn = load_module("synthetic")
    V = args.nvoxel
    P = int(args.gb * 1e9 / (4 * V))
    xs = syn.emission_profile(V)
    cands = [int(c) for c in args.clusters.split(",") if c] or [0]
    if 0 not in cands:
        cands = [0] + cands  # 0 = the plan's own choice
    rows = []
    for C in cands:
        if C:
            os.environ["SART_FUSED_CLUSTER"] = str(C)
        else:
            os.environ.pop("SART_FUSED_CLUSTER", None)
        try:
            with sart.Engine(P, V, device=0, max_iterations=args.iterations, fixed_iterations=True, sweep_mode=sart.SWEEP_FUSED) as eng:
                eng.fill_synthetic(42)
                eng.finalize()
                eng.stage_measurement(eng.forward_project(xs))
                eng.stage_guess(0.9 * xs + 0.01)
                for _ in range(2):
                    eng.solve_resident(True)
                ms = []
                for _ in range(args.steps):
                    eng.solve_resident(True)
                    s = eng.stats()
                    ms.append(s.fused_ms_avg)
                sweep_ms = min(ms)
                rows.append({"nvoxel": V, "rows": P, "forced_cluster": C, "cluster": s.fused_cluster_size, "clusters": s.fused_cluster_count,
                             "sms": s.fused_cluster_size * s.fused_cluster_count, "sweep_ms": sweep_ms,
                             "hbm_gbs": P * V * 4 / (sweep_ms * 1e-3) / 1e9, "between_sweeps_us": 1e3 * s.tail_ms_avg,
                             "whole_solve_kernel": bool(s.solve_launches)})
        except sart.SartError as exc:
            rows.append({"nvoxel": V, "forced_cluster": C, "error": exc.message})
        print("CLUSTER " + json.dumps(rows[-1]), flush=True)


if __name__ == "__main__":
    main()
