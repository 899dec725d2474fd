#!/usr/bin/env python
"""sart_solve_batch (independent frames, --no_guess) against solving the same frames one by one, on one GPU.

    python tools/batch_bench.py [--npixel 200000 --nvoxel 50000 --frames 16 --iterations 10]

Prints one JSON line: seconds of both ways, frame-iterations/s, the speed-up, the batched projection kernels' times per
pass and the largest relative L2 distance between a frame's two solutions. SART_BATCH_FRAMES=4|8 picks the group size.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package  # noqa: E402
from bench import load_module  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--npixel", type=int, default=200000)
    ap.add_argument("--nvoxel", type=int, default=50000)
    ap.add_argument("--frames", type=int, default=16)
    ap.add_argument("--iterations", type=int, default=10)
    args = ap.parse_args()
    sart = load_package()
    syn = load_module("synthetic")
    P, V, F, I = args.npixel, args.nvoxel, args.frames, args.iterations
    with sart.Engine(P, V, device=0, beta_laplace=2e-2, max_iterations=I, fixed_iterations=True) as eng:
        eng.fill_synthetic(42)
        eng.set_laplacian(*syn.laplacian_coo(V))
        eng.finalize()
        frames = np.stack([eng.forward_project(syn.emission_profile(V, t=f / max(1, F - 1))) for f in range(F)])
        eng.solve_batch(frames[:2])  # allocates the group buffers
        t0 = time.perf_counter()
        xb, stb, itb = eng.solve_batch(frames)
        t_batch = time.perf_counter() - t0
        sb = eng.stats()
        eng.solve(frames[0])
        t0 = time.perf_counter()
        singles = [eng.solve(f) for f in frames]
        t_single = time.perf_counter() - t0
        ss = eng.stats()
    worst = max(float(np.linalg.norm(xb[k] - singles[k][0]) / np.linalg.norm(singles[k][0])) for k in range(F))
    out = {"npixel": P, "nvoxel": V, "frames": F, "iterations_per_frame": I, "group": int(os.environ.get("SART_BATCH_FRAMES", "8")),
           "batch_s": t_batch, "one_by_one_s": t_single, "speedup": t_single / t_batch,
           "frame_iterations_per_s_batch": F * I / t_batch, "frame_iterations_per_s_one_by_one": F * I / t_single,
           "ms_per_frame_iteration_batch": 1e3 * t_batch / (F * I), "ms_per_frame_iteration_one_by_one": 1e3 * t_single / (F * I),
           "group_fp_ms": sb.fp_ms_avg, "group_bp_ms": sb.bp_ms_avg, "single_sweep_ms": ss.fused_ms_avg,
           "max_rel_l2_batch_vs_single": worst, "iterations": [int(v) for v in itb]}
    print("BATCH " + json.dumps(out))


if __name__ == "__main__":
    main()
