"""Worker for the multi-GPU parity test: run under torchrun, one rank per GPU.

Every rank owns its row shard (source/main.cpp:67-68), feeds its slice of each frame, and must end with the same
full-length solution as the CPU oracle. Rank 0 alone is given the Laplacian, as in the reference (source/main.cpp:73-74)."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package  # noqa: E402
from bench import load_module  # noqa: E402
from oracle.pyoracle import Restatement  # noqa: E402
from tests.cases import load_synthetic, rel_l2  # noqa: E402


def main():
    sart = load_package()
    cpl = load_module("control_plane")
    cp = cpl.ControlPlane.from_env(timeout_s=300)
    syn = load_synthetic()
    res = Restatement()
    out = {"rank": cp.rank, "world": cp.world, "checks": {}}
    uid = cp.broadcast_bytes(sart.comm_unique_id() if cp.rank == 0 else None)

    def make_engine(P, V, **kw):
        eng = sart.Engine(P, V, rank=cp.rank, nranks=cp.world, device=cp.local_rank, **kw)
        return eng

    first = True
    SPLIT, FUSED = 1, 2
    for name, P, V, kw, use_lap, saturated, mode in [
        ("linear_fixed", 1030, 515, dict(max_iterations=20, fixed_iterations=True), True, False, SPLIT),
        ("linear_fixed_fused", 1030, 515, dict(max_iterations=20, fixed_iterations=True), True, False, FUSED),
        ("linear_tol", 777, 300, dict(max_iterations=400, conv_tolerance=1e-4), False, True, SPLIT),
        ("linear_tol_fused", 777, 300, dict(max_iterations=400, conv_tolerance=1e-4), False, True, FUSED),
        ("wide_fused_cluster2", 300, 20000, dict(max_iterations=6, fixed_iterations=True), True, False, FUSED),
        ("log_fixed", 900, 256, dict(max_iterations=15, fixed_iterations=True, logarithmic=True, relaxation=0.7), True, False, SPLIT),
        ("log_fixed_fused", 900, 256, dict(max_iterations=15, fixed_iterations=True, logarithmic=True, relaxation=0.7), True, False, FUSED),
    ]:
        H = res.fill_synthetic(P, V, 31)
        g = res.forward(H, syn.emission_profile(V))
        if saturated:
            g[::13] = -1.0
        lap = syn.laplacian_coo(V) if use_lap else None
        okw = dict(kw)
        x_ref, st_ref, it_ref = res.solve(H, g, lap=lap, beta_laplace=2e-2, nranks=cp.world, **okw)
        eng = make_engine(P, V, beta_laplace=2e-2, sweep_mode=mode, **kw)
        # NCCL allows one communicator per engine; every engine of this worker needs its own id
        if not first:
            uid = cp.broadcast_bytes(sart.comm_unique_id() if cp.rank == 0 else None)
        first = False
        eng.comm_init(uid)
        off, n = eng.row_offset, eng.nrows
        eng.fill_synthetic(31)
        np.testing.assert_array_equal(eng.download_rows(), H[off:off + n])
        if lap is not None and cp.rank == 0:
            eng.set_laplacian(*lap)
        eng.finalize()
        rho, lam = eng.ray_sums()
        rho_ref, lam_ref = res.ray_sums(H, cp.world)
        x, st, it = eng.solve(g[off:off + n])
        exchange = eng.stats().exchange_mode
        # warm-started second frame
        x2, st2, it2 = eng.solve(1.1 * g[off:off + n], x)
        x2_ref, st2_ref, it2_ref = res.solve(H, 1.1 * g, x0=x_ref, lap=lap, beta_laplace=2e-2, nranks=cp.world, **okw)
        # independent frames through sart_solve_batch (one matrix pass serves all, per-frame sums all-reduced together)
        frames = np.stack([(1.0 + 0.5 * k) * g[off:off + n] for k in range(5)])
        xb, stb, itb = eng.solve_batch(frames)
        batch_exchange = eng.stats().exchange_mode
        singles = [eng.solve(f) for f in frames]
        batch = {"x": max(rel_l2(xb[k], singles[k][0]) for k in range(5)),
                 "same_status": bool(all(int(stb[k]) == singles[k][1] for k in range(5))),
                 "iteration_gap": int(max(abs(int(itb[k]) - singles[k][2]) for k in range(5))),
                 "hash": float(np.sum(xb[4] * np.arange(1, V + 1))), "exchange": int(batch_exchange)}
        # warm-started series through sart_solve_series (look-ahead) against solve-by-solve on the same ranks
        series = None
        if not kw.get("logarithmic"):
            frames = np.stack([(1.0 + 1e-3 * k) * g[off:off + n] for k in range(7)])
            xs, sts, its = eng.solve_series(frames)
            handed = eng.stats().fp_launches
            prev, worst, same = None, 0.0, True
            for k in range(7):
                prev = eng.solve(frames[k]) if prev is None else eng.solve(frames[k], prev[0])
                same = same and int(sts[k]) == prev[1] and abs(int(its[k]) - prev[2]) <= 1
                if int(its[k]) == prev[2]:
                    worst = max(worst, rel_l2(xs[k], prev[0]))
            series = {"x": worst, "consistent": bool(same), "handed": int(handed),
                      "hash": float(np.sum(xs[-1] * np.arange(1, V + 1)))}
        eng.close()
        out["checks"][name] = {
            "series": series,
            "batch": batch,
            "rho": rel_l2(rho, rho_ref), "lam": rel_l2(lam, lam_ref[off:off + n]),
            "x": rel_l2(x, x_ref), "status": [st, st_ref], "iterations": [it, it_ref],
            "x2": rel_l2(x2, x2_ref), "iterations2": [it2, it2_ref],
            "x_hash": float(np.sum(x * np.arange(1, V + 1))), "exchange": exchange, "mode": mode,
        }
    gathered = cp.gather(out)
    cp.barrier()
    cp.shutdown()
    if cp.rank == 0:
        print("RESULT " + json.dumps(gathered))


if __name__ == "__main__":
    main()
