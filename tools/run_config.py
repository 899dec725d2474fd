#!/usr/bin/env python
"""Runs one of the larger BASELINE.json configurations end to end through the C ABI and prints one JSON line.

    python tools/run_config.py --config c4 [--frames 200]                         # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 ... tools/run_config.py --config c3|c5|c4

  c3  synthetic 1 000 000 x 100 000 fp32 matrix (400 GB), sharded over 8 GPUs (50 GB each)
  c5  synthetic 1 500 000 x 200 000 fp32 matrix (1.2 TB), sharded over 8 GPUs (150 GB each)
  c4  time series on the 200 000 x 50 000 matrix + Laplacian, warm start from the previous solution, default tolerance

These sizes have no CPU oracle (the matrix does not fit a host); what is checked are size-independent properties:
bit-reproducibility, scale covariance, every rank ending with the identical solution, a consistent frame started at its
own solution staying put, and (c4) the per-frame stopping statistics. Parity at these voxel counts is covered by the
row-subsampled tests in tests/test_parity_gpu.py.
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

SHAPES = {"c3": (1_000_000, 100_000), "c5": (1_500_000, 200_000), "c4": (200_000, 50_000), "c2": (200_000, 50_000)}


def rel(a, b):
    return float(np.linalg.norm(a - b) / np.linalg.norm(b))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", required=True, choices=sorted(SHAPES))
    ap.add_argument("--frames", type=int, default=200)
    ap.add_argument("--iterations", type=int, default=5)
    ap.add_argument("--scale", type=float, default=1.0, help="shrink the row count (debugging on fewer GPUs)")
    ap.add_argument("--sweep", default="auto", choices=["auto", "split", "fused"])
    ap.add_argument("--series", action="store_true",
                    help="c4: hand the whole series to sart_solve_series (look-ahead) instead of one resident solve per frame")
    ap.add_argument("--batch", type=int, default=0,
                    help="c2: solve this many independent frames (no warm start, --iterations each) with sart_solve_batch "
                         "and one by one, and report both rates")
    args = ap.parse_args()

    sart = load_package()
    syn = load_module("synthetic")
    cp = load_module("control_plane").ControlPlane.from_env(timeout_s=900)
    P, V = SHAPES[args.config]
    P = int(P * args.scale)
    device = cp.local_rank if cp.world > 1 else 0
    series = args.config == "c4"
    mode = {"auto": 0, "split": 1, "fused": 2}[args.sweep]

    t0 = time.perf_counter()
    series = series and not args.batch
    eng = sart.Engine(P, V, rank=cp.rank, nranks=cp.world, device=device, beta_laplace=2e-2, sweep_mode=mode,
                      max_iterations=2000 if series else args.iterations, fixed_iterations=not series)
    if cp.world > 1:
        eng.comm_init(cp.broadcast_bytes(sart.comm_unique_id() if cp.rank == 0 else None))
    eng.fill_synthetic(42)
    if series and cp.rank == 0:
        eng.set_laplacian(*syn.laplacian_coo(V))
    eng.finalize()
    t_setup = time.perf_counter() - t0
    rho, lam = eng.ray_sums()
    out = {"config": args.config, "npixel": P, "nvoxel": V, "n_gpus": cp.world, "rows_per_gpu": eng.nrows,
           "shard_gb": eng.nrows * V * 4 / 1e9, "matrix_gb": P * V * 4 / 1e9, "setup_s": t_setup,
           "ray_length_mean": float(lam.mean()), "ray_density_min": float(rho.min())}

    if args.batch:
        F = args.batch
        frames = np.stack([eng.forward_project(syn.emission_profile(V, t=f / max(1, F - 1))) for f in range(F)])
        eng.solve_batch(frames[:2])  # warm-up (allocates the group buffers)
        t1 = time.perf_counter()
        xb, stb, itb = eng.solve_batch(frames)
        t_batch = time.perf_counter() - t1
        sb = eng.stats()
        pair_ms = sb.fp_ms_avg + sb.bp_ms_avg  # one batched forward + one batched back projection = an iteration of 4 frames
        t1 = time.perf_counter()
        singles = [eng.solve(f) for f in frames]
        t_single = time.perf_counter() - t1
        single_ms = eng.stats().fused_ms_avg
        worst = max(rel(xb[f], singles[f][0]) for f in range(F))
        out.update({"batch_frames": F, "iterations_per_frame": int(itb[0]), "batch_s": t_batch, "one_by_one_s": t_single,
                    "frame_iterations_per_s_batch": float(np.sum(itb)) / t_batch,
                    "frame_iterations_per_s_one_by_one": float(sum(s[2] for s in singles)) / t_single,
                    "speedup": t_single / t_batch, "group_iteration_ms": pair_ms, "group_fp_ms": sb.fp_ms_avg,
                    "group_bp_ms": sb.bp_ms_avg, "single_sweep_ms": single_ms,
                    "max_rel_l2_batch_vs_single": worst, "x_checksum": float(np.dot(xb[-1], np.arange(1, V + 1)))})
    elif not series:
        xs = syn.emission_profile(V)
        g = eng.forward_project(xs)
        x1, st, it = eng.solve(g)
        s = eng.stats()
        x2 = eng.solve(g)[0]
        x3 = eng.solve(4.0 * g)[0]
        eng.set_max_iterations(2)
        x4 = eng.solve(g, xs)[0]
        sweep_ms = s.fused_ms_avg if s.fused_launches else (s.fp_ms_avg + s.bp_ms_avg)
        out.update({
            "iterations": it, "status": st, "solve_ms": s.solve_ms, "iterations_per_s": it / (s.solve_ms / 1e3),
            "sweep": {1: "split", 2: "fused"}[s.sweep_mode], "ms_per_iteration_sweep": sweep_ms,
            "algorithmic_tb_s_per_gpu": 2 * eng.nrows * V * 4 / (sweep_ms * 1e-3) / 1e12,
            "bit_reproducible": bool(np.array_equal(x1, x2)), "scale_covariant": bool(np.array_equal(x3, 4.0 * x1)),
            "fixed_point_rel_l2": rel(x4, xs), "finite_nonnegative": bool(np.all(np.isfinite(x1)) and np.all(x1 >= 0)),
            "x_checksum": float(np.dot(x1, np.arange(1, V + 1))),
        })
    else:
        F = args.frames
        frames = np.stack([eng.forward_project(syn.emission_profile(V, t=f / max(1, F - 1))) for f in range(F)])
        iters, statuses = [], []
        t1 = time.perf_counter()
        dev_ms = 0.0
        x = None
        if args.series:
            xs_all, sts, its = eng.solve_series(frames)
            ss = eng.stats()
            x, iters, statuses, dev_ms = xs_all[-1], [int(v) for v in its], [int(v) for v in sts], ss.solve_ms
            out.update({"series_call": True, "matrix_passes": ss.sweeps, "frames_handed_their_first_iteration": ss.fp_launches,
                        "sweeps_repeated": ss.bp_launches})
        for f in range(0 if not args.series else F, F):
            eng.stage_measurement(frames[f])
            st, it = eng.solve_resident(f > 0)  # warm start from the previous solution, left on the device
            x = eng.fetch_solution()            # what main() hands to the solution writer
            dev_ms += eng.stats().solve_ms
            iters.append(it)
            statuses.append(st)
        wall = time.perf_counter() - t1
        xt = syn.emission_profile(V, t=1.0)
        out.update({
            "frames": F, "wall_s": wall, "frames_per_s": F / wall, "device_ms_total": dev_ms,
            "iterations_total": int(np.sum(iters)), "iterations_per_s": float(np.sum(iters) / wall),
            "iterations_first_frame": iters[0], "iterations_median_warm": float(np.median(iters[1:])) if F > 1 else None,
            "iterations_max_warm": int(np.max(iters[1:])) if F > 1 else None,
            "all_converged": bool(all(s == 0 for s in statuses)), "sweep": {1: "split", 2: "fused"}[eng.stats().sweep_mode],
            "last_frame_rel_l2_vs_truth": rel(x, xt), "x_checksum": float(np.dot(x, np.arange(1, V + 1))),
        })
    sums = cp.gather(out["x_checksum"])
    out["ranks_agree"] = bool(len(set(sums)) == 1)
    eng.close()
    cp.barrier()
    cp.shutdown()
    if cp.rank == 0:
        print("CONFIG " + json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
