#!/usr/bin/env python
"""Races the engine against the reference's OWN GPU solver on the same box and the same inputs.

The reference GPU path (source/sartsolver_cuda.cpp + source/cuda/sart_kernels.cu + cuBLAS Sgemv/Sdot) is compiled for
sm_100a, unmodified, into oracle/_ref/libsart_ref_gpu.so (oracle/Makefile; its own Makefile stops at sm_86). It needs
the matrix in host memory and spends tens of seconds in its constructor's host-side ray sums, so the race runs on a row
slice of the BASELINE config-2 matrix (same 50 000 voxels, same generator); per-iteration time is linear in the rows.

Prints one JSON line: per-iteration times of both, the ratio, and the relative L2 distance between the two solutions
(engine in SART_SEMANTICS_GPU, i.e. following the GPU reference's initial-guess mask and 1e-7 floor).
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
from oracle.pyoracle import REF_GPU_SO, Reference, Restatement  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=20000)
    ap.add_argument("--nvoxel", type=int, default=50000)
    ap.add_argument("--iterations", type=int, default=10)
    args = ap.parse_args()
    P, V, I = args.rows, args.nvoxel, args.iterations

    sart = load_package()
    syn = load_module("synthetic")
    res = Restatement()
    H = res.fill_synthetic(P, V, 42)
    g = res.forward(H, syn.emission_profile(V))
    lap = syn.laplacian_coo(V)
    beta = 2e-2

    ref = Reference(REF_GPU_SO)
    out = {"rows": P, "nvoxel": V, "slice_gb": P * V * 4 / 1e9, "max_iterations": I}
    # The reference holds sum(fit^2) in fp32 and adds with atomics, so with a vanishing tolerance its convergence value
    # sooner or later ties bit for bit and it stops at a run-dependent iteration. Its time is
    # therefore normalised per matrix sweep: a solve of n iterations makes 2n + 2 sweeps (initial guess, first
    # projection, two per iteration: sartsolver_cuda.cpp:166-189, 231-249). Frame 0 (cuBLAS warm-up) is left out.
    nframes = 6
    frames = np.repeat(g[None, :], nframes, axis=0)
    x_ref, st, it, (t_ctor, t_solve) = ref.solve_frames(H, frames, lap=lap, beta_laplace=beta, max_iterations=I,
                                                        conv_tolerance=1e-300, nranks=1, no_guess=True)
    fs = ref.frame_seconds
    ref_sweeps = int(np.sum(2 * it[1:] + 2))
    ref_ms_per_sweep = float(np.sum(fs[1:])) / ref_sweeps * 1e3
    out["reference_gpu"] = {"ctor_s": t_ctor, "iterations": it.tolist(), "status": st.tolist(),
                            "frame_ms": (fs * 1e3).round(3).tolist(), "ms_per_sweep": ref_ms_per_sweep,
                            "tb_s": P * V * 4 / (ref_ms_per_sweep * 1e-3) / 1e12}
    # compare solutions at the iteration count of the last frame
    I = int(it[-1])
    x_ref_I = x_ref[-1].copy()
    out["iterations_compared"] = I

    ours = {}
    for name, mode in (("split", 1), ("fused", 2)):
        with sart.Engine(P, V, device=0, beta_laplace=beta, max_iterations=I, fixed_iterations=True,
                         semantics=sart.SEMANTICS_GPU, sweep_mode=mode) as eng:
            t0 = time.perf_counter()
            eng.upload_rows(H)
            t_up = time.perf_counter() - t0
            eng.set_laplacian(*lap)
            eng.finalize()
            eng.solve(g)
            x, st, it = eng.solve(g)
            s = eng.stats()
        sweep_ms = s.fused_ms_avg if s.fused_launches else s.fp_ms_avg + s.bp_ms_avg
        ours[name] = {"ms_per_iteration": sweep_ms, "solve_ms": s.solve_ms, "upload_s": t_up,
                      "rel_l2_vs_reference_gpu": float(np.linalg.norm(x - x_ref_I) / np.linalg.norm(x_ref_I))}
    out["engine"] = ours
    for name in ours:
        # the engine's solve of n iterations: initial guess + n fused sweeps + last projection (fused), or
        # initial guess + (n + 1) forward + n back projections (split)
        out[f"speedup_{name}_per_iteration"] = 2 * ref_ms_per_sweep / ours[name]["ms_per_iteration"]
    print("RACE " + json.dumps(out))


if __name__ == "__main__":
    main()
