#!/usr/bin/env python
"""Headline benchmark: SART iterations/s on the BASELINE.json workload, with roofline, CPU baseline and end-to-end.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (all N): BASELINE.json configs[1] / configs[3]'s matrix -- synthetic 200 000 pixel x 50 000 voxel fp32
ray-transfer matrix (40 GB) with Laplacian regularisation, row-sharded over the N GPUs exactly as the reference
shards it over MPI ranks (strong scaling: the matrix is fixed, every GPU holds 1/N of its rows).
A step = one frame solved with a warm start and a fixed number of iterations (--iterations, default 10): 11 sweeps of
the shard (each iterate needs its forward projection; all but the last also the back projection), the per-iteration
reductions over clusters and ranks, the updates and the on-device decisions -- by default all inside ONE cooperative
launch of fused_solve_kernel per step, which is what the roofline entry then times (`sweeps_per_launch` sweeps of
algorithmic bytes over the launch's CUDA-event duration).
  value  iterations/s with the frame and the starting solution already in HBM (sart_solve_resident), device time
         from CUDA events on the engine's stream, max over ranks.
  e2e    the same through sart_solve with pinned HOST buffers: the frame and the guess go up, the solution comes
         back, every step; wall clock around the K calls, max over ranks.
The matrix (40 GB) is far larger than L2 (126 MB), so consecutive sweeps cannot hit in cache; no extra flush.
"""
import argparse
import json
import os
import shutil
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
from __graft_entry__ import PKG_DIR, load_package  # noqa: E402

METRIC = "sart_iterations_per_s"
UNIT = "iterations/s"
SEED = 42
BETA = 2e-2  # CLI default of the reference, source/arguments.cpp:127


def load_module(name):
    import importlib.util
    spec = importlib.util.spec_from_file_location("sart_b200_" + name, os.path.join(PKG_DIR, name + ".py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, copy read+write)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md, 6.65 TB/s)"


def profiled_traffic(kernel, sweeps_per_launch=1.0):
    """dram bytes per launch of the dominant kernel from the committed ncu capture, if one was summarised. The
    whole-solve kernel was captured with 4 sweeps per launch: its figure is kept per sweep and scaled to this run."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)
        if kernel == "fused_solve_kernel":
            return int(t["fused_solve_kernel_per_sweep"] * sweeps_per_launch)
        return t.get(kernel)
    except Exception:
        return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 50 ms; only samples inside the timed region count."""
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.rows, self.proc, self.thread = [], None, None
        exe = shutil.which("nvidia-smi")
        if exe:
            try:
                self.proc = subprocess.Popen([exe, "-i", str(device), "--query-gpu=" + self.FIELDS,
                                              "--format=csv,noheader,nounits", "-lms", "50"],
                                             stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
                self.thread = threading.Thread(target=self._pump, daemon=True)
                self.thread.start()
            except Exception:
                self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), line.strip()))

    def summary(self, t0, t1):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for t, line in self.rows:
            if not (t0 - 0.05 <= t <= t1 + 0.05):
                continue
            f = [x.strip() for x in line.split(",")]
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except Exception:
                continue
            for name, val in zip(names, f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "note": "nvidia-smi gave no samples"}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


# ---- the reference's CPU path on a bounded sample -------------------------------------------------------------------

def cpu_baseline(npixel, nvoxel, iterations=4, sample_rows=4096, threads=None, laplacian=True):
    """Times the reference's own CPU solver (oracle/_ref, or the C restatement when that is not built) on the first
    ``sample_rows`` rows of the very same matrix, all voxels, Laplacian on, warm start, ``iterations`` iterations, one
    thread-rank per host core, and extrapolates linearly in the row count (the per-iteration cost is a sweep of the
    matrix: proportional to rows x voxels)."""
    from oracle.pyoracle import Reference, Restatement
    syn = load_module("synthetic")
    res = Restatement()
    cores = threads or os.cpu_count() or 1
    rows = min(sample_rows, npixel)
    H = res.fill_synthetic(rows, nvoxel, SEED)
    xs = syn.emission_profile(nvoxel)
    g = res.forward(H, xs)
    lap = syn.laplacian_coo(nvoxel) if laplacian else None
    x0 = 0.9 * xs + 0.01
    if Reference.available():
        kind = "reference"
        ranks = max(1, min(cores, rows))
        ref = Reference()
        _, st, it, (t_ctor, t_solve) = ref.solve_frames(H, g[None, :], x0=x0, lap=lap, beta_laplace=BETA,
                                                        max_iterations=iterations, conv_tolerance=1e-300, nranks=ranks)
        done = int(it[0])
    else:
        kind, ranks = "port", 1
        os.environ.setdefault("OMP_NUM_THREADS", str(cores))
        sums = res.ray_sums(H)
        t = time.perf_counter()
        _, st, done = res.solve(H, g, x0=x0, lap=lap, beta_laplace=BETA, max_iterations=iterations,
                                fixed_iterations=True, sums=sums)
        t_solve, t_ctor = time.perf_counter() - t, float("nan")
    sample_rate = done / t_solve
    value = sample_rate * rows / npixel
    return {
        "value": value, "unit": UNIT, "cores": ranks if kind == "reference" else cores, "kind": kind,
        "sample": (f"rows 0..{rows - 1} of the same {npixel}x{nvoxel} matrix (all voxels, Laplacian {'on' if laplacian else 'off'}, warm start), "
                   f"{done} iterations in {t_solve:.2f} s on {ranks} thread-rank(s) = {sample_rate:.3f} it/s on the sample, "
                   f"scaled by {rows}/{npixel} rows; constructor sums {t_ctor:.2f} s not counted"),
        "sample_seconds": t_solve,
    }


def run_reference_arm(args):
    vals, last = [], None
    t0 = time.perf_counter()
    for i in range(args.warmup + args.steps):
        last = cpu_baseline(args.npixel, args.nvoxel, iterations=args.cpu_iterations, sample_rows=args.cpu_rows,
                            threads=args.cpu_ranks or None, laplacian=args.laplacian)
        if i >= args.warmup:
            vals.append(last)
        if time.perf_counter() - t0 > 240 and vals:
            break
    value = len(vals) * args.cpu_iterations / sum(v["sample_seconds"] for v in vals) * min(args.cpu_rows, args.npixel) / args.npixel
    cb = dict(last)
    cb["value"] = value
    cb.pop("sample_seconds", None)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": len(vals),
        "warmup": args.warmup, "ms_per_step": 1e3 * sum(v["sample_seconds"] for v in vals) / len(vals),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, 0), "cpu_baseline": cb,
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(args, n):
    gb = args.npixel * args.nvoxel * 4 / 1e9
    return {
        "workload": f"BASELINE {args.workload_name}: synthetic {args.npixel}x{args.nvoxel} fp32 ray-transfer matrix "
                    f"({gb:.1f} GB){' + Laplacian' if args.laplacian else ''}, row-sharded over the GPUs",
        "npixel": args.npixel, "nvoxel": args.nvoxel, "iterations_per_step": args.iterations, "warm_start": True,
        "laplacian": args.laplacian, "beta_laplace": BETA, "seed": SEED,
        "sharding": f"rows/{max(n, 1)} (source/main.cpp:67-68)",
        "l2_policy": ("matrix >> L2 (126 MB); no flush needed between sweeps" if gb > 1.0 else
                      "matrix ~ L2-sized: sweeps partly hit in L2, so the HBM roofline can be exceeded; not flushed"),
    }


# ---- the engine -----------------------------------------------------------------------------------------------------

def run_engine_arm(args, cp):
    import torch
    sart = load_package()
    syn = load_module("synthetic")
    P, V, I = args.npixel, args.nvoxel, args.iterations
    device = cp.local_rank if cp.world > 1 else 0

    eng = sart.Engine(P, V, rank=cp.rank, nranks=cp.world, device=device, beta_laplace=BETA, max_iterations=I,
                      fixed_iterations=True, sweep_mode={"auto": 0, "split": 1, "fused": 2}[args.sweep])
    if cp.world > 1:
        uid = cp.broadcast_bytes(sart.comm_unique_id() if cp.rank == 0 else None)
        eng.comm_init(uid)
    eng.fill_synthetic(SEED)
    if args.laplacian and cp.rank == 0:
        eng.set_laplacian(*syn.laplacian_coo(V))  # the reference reads it on rank 0 only; finalize shares it
    eng.finalize()

    xs = syn.emission_profile(V)
    g_pin = torch.empty(eng.nrows, dtype=torch.float64).pin_memory()
    x_pin = torch.empty(V, dtype=torch.float64).pin_memory()
    g = g_pin.numpy()
    g[:] = eng.forward_project(xs)
    guess = 0.9 * xs + 0.01

    def resident_step():
        st, it = eng.solve_resident(True)
        assert it == I, (st, it)
        return eng.stats()

    # ---- value: inputs resident in HBM -------------------------------------------------------------------------------
    eng.stage_measurement(g)
    eng.stage_guess(guess)
    sampler = ClockSampler(device) if cp.rank == 0 else None  # started before the warm-up so samples flow by t0
    for _ in range(args.warmup):
        resident_step()
    torch.cuda.synchronize()
    cp.barrier()
    w0 = time.perf_counter()
    dev_ms, launches, fp_ms, bp_ms, fu_ms, fp_n, bp_n, fu_n, solve_n = 0.0, 0, 0.0, 0.0, 0.0, 0, 0, 0, 0
    for _ in range(args.steps):
        s = resident_step()
        dev_ms += s.solve_ms
        launches += s.gpu_launches
        fp_ms += s.fp_ms_avg * s.fp_launches
        bp_ms += s.bp_ms_avg * s.bp_launches
        fu_ms += s.fused_ms_avg * s.fused_launches
        fp_n += s.fp_launches
        bp_n += s.bp_launches
        fu_n += s.fused_launches
        solve_n += s.solve_launches
        tail_ms = s.tail_ms_avg
    torch.cuda.synchronize()  # (every step already ended with an event sync on the engine's own stream)
    sweep_mode = {1: "split", 2: "fused"}.get(s.sweep_mode, "?")
    exchange_mode = {0: "none", 1: "nccl", 2: "peer-memory kernel"}.get(s.exchange_mode, "?")
    cp.barrier()
    w1 = time.perf_counter()
    clocks = sampler.summary(w0, w1) if sampler else None
    total_ms = cp.max(dev_ms)
    wall_ms = cp.max((w1 - w0) * 1e3)
    value = args.steps * I / (total_ms / 1e3)

    # ---- e2e: host buffers through sart_solve ---------------------------------------------------------------------------
    x_host = x_pin.numpy()
    for _ in range(max(1, args.warmup // 2)):
        x_host[:] = guess
        eng.solve(g, x_host)
    cp.barrier()
    e0 = time.perf_counter()
    h2d = d2h = 0
    for _ in range(args.steps):
        x_host[:] = guess
        x, st, it = eng.solve(g, x_host)
        s = eng.stats()
        h2d, d2h = s.h2d_bytes, s.d2h_bytes
    cp.barrier()
    e1 = time.perf_counter()
    e2e_s = cp.max(e1 - e0)
    e2e = {"value": args.steps * I / e2e_s, "unit": UNIT, "h2d_bytes_per_step": int(cp.sum(h2d)),
           "d2h_bytes_per_step": int(cp.sum(d2h)), "timing": "host wall clock around K sart_solve calls, max over ranks"}

    # ---- roofline of the dominant kernel (this rank's shard; device events inside the timed region) -------------------
    shard_bytes = eng.nrows * V * 4
    peak, peak_src = measured_peak()
    # algorithmic bytes per launch: one sweep of the shard for a projection kernel; the fused kernel does
    # the work of both sweeps (2 x shard) while reading HBM once -- "hbm_gbs" is its actual DRAM-side rate
    kern = {
        "fp_rows_kernel": {"ms": fp_ms / max(fp_n, 1), "launches": fp_n, "algorithmic_bytes": shard_bytes, "hbm_bytes": shard_bytes},
        "bp_cols_kernel": {"ms": bp_ms / max(bp_n, 1), "launches": bp_n, "algorithmic_bytes": shard_bytes, "hbm_bytes": shard_bytes},
        "fused_sweep_kernel": {"ms": fu_ms / max(fu_n, 1), "launches": fu_n, "algorithmic_bytes": 2 * shard_bytes, "hbm_bytes": shard_bytes},
    }
    if solve_n:
        # the whole solve ran as ONE launch per step (fused_solve_kernel): fu_n counts the sweeps inside those launches
        per = fu_n / solve_n
        kern.pop("fused_sweep_kernel")
        kern["fused_solve_kernel"] = {"ms": fu_ms / solve_n, "launches": solve_n, "sweeps_per_launch": per,
                                      "algorithmic_bytes": 2 * shard_bytes * per, "hbm_bytes": shard_bytes * per}
    for k in kern.values():
        k["gbs"] = k["algorithmic_bytes"] / (k["ms"] * 1e-3) / 1e9 if k["ms"] > 0 else 0.0
        k["hbm_gbs"] = k["hbm_bytes"] / (k["ms"] * 1e-3) / 1e9 if k["ms"] > 0 else 0.0
        k["share_of_step"] = k["ms"] * k["launches"] / dev_ms if dev_ms > 0 else 0.0
    dom = max(kern, key=lambda n: kern[n]["ms"] * kern[n]["launches"])
    roofline = {
        "bound": "hbm", "kernel": dom, "achieved": kern[dom]["gbs"], "peak": peak, "unit": "GB/s",
        "frac": kern[dom]["gbs"] / peak, "traffic": profiled_traffic(dom, kern[dom].get("sweeps_per_launch", 1.0)), "peak_source": peak_src,
        "algorithmic_bytes_per_launch": kern[dom]["algorithmic_bytes"],
        "hbm_bytes_per_launch": kern[dom]["hbm_bytes"], "hbm_achieved": kern[dom]["hbm_gbs"],
        "hbm_frac": kern[dom]["hbm_gbs"] / peak,
        "note": ("fused sweep: algorithmic bytes are the two sweeps of an iteration (2*P*V*4); the kernel reads the shard from "
                 "HBM once and re-reads it from shared memory, so achieved/peak can exceed 1 while hbm_frac cannot"
                 + ("; one launch = all sweeps of a solve (cluster reduction, decision and update between grid barriers inside)"
                    if dom == "fused_solve_kernel" else "")
                 if dom.startswith("fused_") else "one sweep of the shard per launch"),
        "kernels": kern,
        "iteration_roofline_frac": (2.0 * shard_bytes * (value / 1.0) / 1e9) / peak,  # 2*bytes(A) per iteration / peak
    }
    cp_tail = cp.max(tail_ms)
    gathered = cp.gather({"rank": cp.rank, "dev_ms": dev_ms, "dominant_gbs": kern[dom]["gbs"]})

    cb = None
    if cp.rank == 0 and cp.world == 1 and not args.no_cpu_baseline:
        cb = cpu_baseline(P, V, iterations=args.cpu_iterations, sample_rows=args.cpu_rows, threads=args.cpu_ranks or None,
                          laplacian=args.laplacian)
        cb.pop("sample_seconds", None)

    if cp.rank == 0:
        info = sart.device_query(device)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": cp.world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": dict(workload_config(args, cp.world), sweep=sweep_mode, exchange=exchange_mode), "e2e": e2e,
            "between_sweeps_us": 1e3 * cp_tail, "gpu_launches": launches,
            "roofline": roofline, "cpu_baseline": cb, "clocks": clocks,
            "wall_ms_per_step": wall_ms / args.steps, "per_rank": gathered,
            "device": {"name": info.name.decode(), "sms": info.sm_count, "l2_bytes": info.l2_bytes,
                       "total_mem": info.total_mem},
            "host": {"cores": os.cpu_count()},
        }
        print(json.dumps(line), flush=True)
    eng.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--iterations", type=int, default=10, help="SART iterations per step (fixed)")
    ap.add_argument("--npixel", type=int, default=200000)
    ap.add_argument("--nvoxel", type=int, default=50000)
    ap.add_argument("--cpu-rows", type=int, default=16384, help="rows of the matrix the CPU baseline is timed on")
    ap.add_argument("--cpu-iterations", type=int, default=5)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--sweep", default="auto", choices=["auto", "split", "fused"])
    ap.add_argument("--cpu-ranks", type=int, default=0, help="thread-ranks of the CPU baseline (0 = all host cores)")
    ap.add_argument("--workload", default="c2", choices=["c2", "c1"],
                    help="c2: BASELINE configs[1] (the headline, default); c1: configs[0], 8192x4096, 100 iterations, "
                         "no Laplacian, CPU reference on the whole problem with 2 ranks")
    args = ap.parse_args()
    args.laplacian, args.workload_name = True, "configs[1]"
    if args.workload == "c1":
        args.npixel, args.nvoxel, args.iterations = 8192, 4096, 100
        args.cpu_rows, args.cpu_iterations, args.cpu_ranks = 8192, 100, args.cpu_ranks or 2
        args.laplacian, args.workload_name = False, "configs[0]"
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    if args.impl == "reference":
        # CPU only: rank 0 alone works, the other ranks of a torchrun launch leave at once (no rendezvous needed)
        if int(os.environ.get("RANK", "0")) == 0:
            run_reference_arm(args)
        return

    cpl = load_module("control_plane")
    cp = cpl.ControlPlane.from_env()
    if cp.world != max(args.gpus, 1) and cp.rank == 0:
        print(f"bench.py: --gpus {args.gpus} but WORLD_SIZE={cp.world}; using WORLD_SIZE", file=sys.stderr)
    try:
        run_engine_arm(args, cp)
    finally:
        cp.shutdown()


if __name__ == "__main__":
    main()
