"""Seeded parity cases shared by tests/golden/make_golden.py (which runs the reference on them) and the tests.

Every case is built from the oracle's synthetic generator so that it can be re-created anywhere from a few
integers; the golden file stores the reference's answers plus checksums of the regenerated inputs.
Edge cases follow what the reference's code paths distinguish (it ships no tests of its own):
saturated (negative) pixels, rows / columns below the ray-length / ray-density thresholds, warm starts,
ragged shapes that do not divide by the rank count or by the vector width, Laplacian on/off, both solvers.
"""
import importlib.util
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def load_synthetic():
    path = os.path.join(ROOT, "mpi-cuda-sartsolver_b200", "synthetic.py")
    spec = importlib.util.spec_from_file_location("sart_b200_synthetic", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


CASES = {
    # name: dict(P, V, seed, solver keyword arguments..., extras)
    "linear_fixed30": dict(P=384, V=160, seed=42, kw=dict(max_iterations=30, conv_tolerance=1e-30)),
    "linear_laplacian_tol_np2": dict(P=512, V=256, seed=7, laplacian=True,
                                     kw=dict(max_iterations=400, conv_tolerance=1e-5, beta_laplace=2e-2, nranks=2)),
    "linear_masks_warmstart": dict(P=300, V=128, seed=11, saturated_every=17, zero_row_every=41, zero_col_every=29,
                                   warm_start=True, laplacian=True,
                                   kw=dict(max_iterations=40, conv_tolerance=1e-30, relaxation=0.8)),
    "linear_masks_guess": dict(P=300, V=128, seed=12, zero_row_every=37, zero_col_every=31,
                               kw=dict(max_iterations=25, conv_tolerance=1e-30)),
    "ragged_np4": dict(P=301, V=77, seed=5, laplacian=True,
                       kw=dict(max_iterations=35, conv_tolerance=1e-30, nranks=4, beta_laplace=1e-2)),
    "log_laplacian_np3": dict(P=320, V=144, seed=9, laplacian=True,
                              kw=dict(max_iterations=30, conv_tolerance=1e-30, logarithmic=True, relaxation=0.7,
                                      nranks=3, beta_laplace=2e-2)),
    # Laplacian rows that do not add up to zero (a constant on the diagonal, as a boundary row with a fixed value has)
    # and a frame far from max(g) = 1: the log solver's penalty beta * L * log(x) then depends on the units of x
    # (source/sartsolver.cpp:292 takes the log of the solution as the caller gave it).
    "log_laplacian_rowsum": dict(P=320, V=144, seed=19, laplacian=True, lap_diag_add=0.5, g_scale=37.5,
                                 kw=dict(max_iterations=20, conv_tolerance=1e-30, logarithmic=True, relaxation=0.8,
                                         beta_laplace=2e-2)),
    "linear_laplacian_rowsum": dict(P=320, V=144, seed=20, laplacian=True, lap_diag_add=0.5, g_scale=37.5,
                                    kw=dict(max_iterations=20, conv_tolerance=1e-30, beta_laplace=2e-2)),
    "log_tol": dict(P=256, V=96, seed=21, kw=dict(max_iterations=500, conv_tolerance=1e-5, logarithmic=True)),
    "max_iterations_exceeded": dict(P=128, V=64, seed=3, kw=dict(max_iterations=3, conv_tolerance=1e-12)),
    "series_warm4": dict(P=256, V=100, seed=13, frames=4, laplacian=True,
                         kw=dict(max_iterations=200, conv_tolerance=1e-4, nranks=2)),
    "series_noguess3": dict(P=256, V=100, seed=14, frames=3, no_guess=True,
                            kw=dict(max_iterations=60, conv_tolerance=1e-4)),
}


def build_case(name, restatement):
    """Returns dict(H, frames [F][P], x0 or None, lap or None, kw, no_guess)."""
    c = CASES[name]
    syn = load_synthetic()
    P, V = c["P"], c["V"]
    H = restatement.fill_synthetic(P, V, c["seed"])
    if c.get("zero_row_every"):
        H[::c["zero_row_every"], :] = 0
    if c.get("zero_col_every"):
        H[:, ::c["zero_col_every"]] = 0
    nframes = c.get("frames", 1)
    frames = np.stack([restatement.forward(H, syn.emission_profile(V, t=f / max(1, nframes - 1) if nframes > 1 else 0.0))
                       for f in range(nframes)])
    frames = frames * c.get("g_scale", 1.0)
    if c.get("saturated_every"):
        frames[:, ::c["saturated_every"]] = -1.0
    x0 = 0.5 * syn.emission_profile(V, t=1.0) + 0.1 if c.get("warm_start") else None
    lap = syn.laplacian_coo(V) if c.get("laplacian") else None
    if lap is not None and c.get("lap_diag_add"):
        idx, val = lap
        val = val.copy()
        val[(idx // np.uint64(V)) == (idx % np.uint64(V))] += np.float32(c["lap_diag_add"])
        lap = (idx, val)
    return dict(H=H, frames=frames, x0=x0, lap=lap, kw=dict(c["kw"]), no_guess=bool(c.get("no_guess")))


def input_checksums(case):
    H, frames = case["H"], case["frames"]
    return np.array([H.astype(np.float64).sum(), float(H[H.shape[0] // 2, H.shape[1] // 3]),
                     frames.sum(), float(frames[-1, frames.shape[1] // 2])])


def rel_l2(a, b):
    return float(np.linalg.norm(np.asarray(a) - np.asarray(b)) / np.linalg.norm(b))
