#!/usr/bin/env python
"""Headline benchmark: SART iterations/s on the BASELINE.json workloads, with roofline, CPU baseline, end-to-end and a
correctness record of the very run that was timed.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c1|c2|c3|c4|c5] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Default workload (all N): BASELINE.json configs[1] / configs[3]'s matrix -- synthetic 200 000 pixel x 50 000 voxel fp32
ray-transfer matrix (40 GB) with Laplacian regularisation, row-sharded over the N GPUs exactly as the reference shards
it over MPI ranks (strong scaling: the matrix is fixed, every GPU holds 1/N of its rows).
A step = one frame solved with a warm start and a fixed number of iterations (--iterations, default 10): I + 1 forward
and I back projections of the shard, the per-iteration reductions over clusters and ranks, the updates and the on-device
decisions -- by default all inside ONE cooperative launch of fused_solve_kernel per step, which reads the shard from HBM
I + 1 times (each sweep serves both projections).
  value     iterations/s with the frame and the starting solution already in HBM (sart_solve_resident), device time
            from CUDA events on the engine's stream, max over ranks.
  e2e       the same through sart_solve with pinned HOST buffers: the frame and the guess go up, the solution comes
            back, every step; wall clock around the K calls, max over ranks.
  roofline  of the dominant kernel: `achieved` / `frac` are DRAM-side (bytes the launch streams from HBM over its
            CUDA-event duration, against the measured copy peak); `algorithmic_*` bills SURVEY 8(d)'s 2 sweeps per
            iteration ((2 I + 1) shards per launch) and so exceeds 1 for a kernel that reads HBM once per iteration.
  check     outside the timed region: every rank ends with the same solution bits; the same frame through the OTHER
            kernel family and collective (two-kernel sweep, NCCL all-reduce) agrees to <= 1e-6; a consistent frame
            started at its own solution stays there; the first 512 rows of the same matrix agree with the CPU
            reference to <= 1e-5. A failed check makes the run exit non-zero after printing its line.
Other workloads: c1 = configs[0]; c3 / c5 = configs[2] / configs[4] (400 GB / 1.2 TB, as many GPUs as hold them);
c4 = configs[3], the warm-started series (sart_solve_series) on a line-of-sight matrix of the same shape whose frames need
about ten iterations each. With 8 GPUs the default line also carries `extra.c3`: the 400 GB matrix at 50 GB per GPU
beside the same shard alone on one GPU (scaling efficiency at capacity).
The matrices are far larger than L2 (126 MB), so consecutive sweeps cannot hit in cache; no extra flush (c1 excepted).
"""
import argparse
import hashlib
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
READ_PEAK_GBS = 7400.0  # pure-read kernel on this pool's B200 (profiles/r01_membw.txt); the copy peak is read+write
CAPTURE_SHARD_BYTES = 200000 * 50000 * 4  # the shard profiles/traffic.json was captured on

WORKLOADS = {
    # name: BASELINE.json config, shape, iterations per step, Laplacian, CPU-sample rows / iterations / ranks
    "c1": dict(name="configs[0]", npixel=8192, nvoxel=4096, iterations=100, laplacian=False, cpu_rows=8192,
               cpu_iterations=100, cpu_ranks=2),
    "c2": dict(name="configs[1]", npixel=200000, nvoxel=50000, iterations=10, laplacian=True, cpu_rows=16384,
               cpu_iterations=5, cpu_ranks=0),
    "c3": dict(name="configs[2]", npixel=1000000, nvoxel=100000, iterations=5, laplacian=False, cpu_rows=8192,
               cpu_iterations=4, cpu_ranks=0),
    "c4": dict(name="configs[3]", npixel=200000, nvoxel=50000, iterations=0, laplacian=True, cpu_rows=16384,
               cpu_iterations=0, cpu_ranks=0),
    "c5": dict(name="configs[4]", npixel=1500000, nvoxel=200000, iterations=5, laplacian=False, cpu_rows=4096,
               cpu_iterations=4, cpu_ranks=0),
}
SERIES_TOLERANCE = 1e-6   # c4: frames then take about ten iterations (default 1e-5: six)
SERIES_DRIFT = 0.002      # c4: how far the emission profile moves per frame (synthetic.emission_profile's t)


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


def profiled_traffic(kernel, sweeps_per_launch, shard_bytes):
    """dram bytes per launch of the dominant kernel, from the committed ncu capture of the 40 GB shard: kept per sweep and
    scaled to the sweeps of this run's launch and to this rank's shard (the kernel streams the shard once per sweep)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)
        per_sweep = t["fused_solve_kernel_per_sweep"] if kernel == "fused_solve_kernel" else t[kernel]
        return int(per_sweep * sweeps_per_launch * shard_bytes / CAPTURE_SHARD_BYTES)
    except Exception:
        return None


def rel_l2(a, b):
    nb = float(np.linalg.norm(b))
    return float(np.linalg.norm(np.asarray(a) - np.asarray(b)) / nb) if nb > 0 else float("nan")


def bits_hash(x):
    return hashlib.sha256(np.ascontiguousarray(x, dtype=np.float64).tobytes()).hexdigest()[:16]


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

    def window(self, t0, t1):
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for t, line in list(self.rows):
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

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
            self.proc = None


# ---- the reference's CPU path on a bounded sample -------------------------------------------------------------------

def cpu_baseline(npixel, nvoxel, iterations=4, sample_rows=4096, threads=None, laplacian=True):
    """Times the reference's own CPU solver (oracle/_ref, or the C restatement when that is not built) on the first
    ``sample_rows`` rows of the very same matrix, all voxels, warm start, ``iterations`` iterations, one thread-rank per
    host core, and extrapolates linearly in the row count (the per-iteration cost is a sweep of the matrix:
    proportional to rows x voxels)."""
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


def cpu_series_baseline(npixel, nvoxel, frames=3, sample_rows=16384, threads=None):
    """c4: the reference's own CPU solver on a warm-started series of the line-of-sight matrix's first ``sample_rows`` rows
    (same tolerance, Laplacian, drift); iterations/s and frames/s extrapolated linearly in the row count."""
    from oracle.pyoracle import Reference, Restatement
    syn = load_module("synthetic")
    res = Restatement()
    cores = threads or os.cpu_count() or 1
    rows = min(sample_rows, npixel)
    H = res.fill_synthetic(rows, nvoxel, SEED, kind=1)
    truth = [syn.emission_profile(nvoxel, t=SERIES_DRIFT * k) for k in range(frames)]
    g = np.stack([res.forward(H, x) for x in truth])
    lap = syn.laplacian_coo(nvoxel)
    x0 = truth[0].copy()  # the series is timed in its steady state, not on the first frame's long descent
    if Reference.available():
        kind, ranks = "reference", max(1, min(cores, rows))
        ref = Reference()
        _, st, it, (t_ctor, t_solve) = ref.solve_frames(H, g, x0=x0, lap=lap, beta_laplace=BETA, max_iterations=2000,
                                                        conv_tolerance=SERIES_TOLERANCE, nranks=ranks)
    else:
        kind, ranks = "port", 1
        sums = res.ray_sums(H)
        t = time.perf_counter()
        it, x = [], x0
        for k in range(frames):
            x, _, n = res.solve(H, g[k], x0=x, lap=lap, beta_laplace=BETA, conv_tolerance=SERIES_TOLERANCE, sums=sums)
            it.append(n)
        t_solve = time.perf_counter() - t
    total_it = int(np.sum(it))
    scale = rows / npixel
    return {
        "value": total_it / t_solve * scale, "unit": UNIT, "cores": ranks, "kind": kind,
        "frames_per_s": frames / t_solve * scale,
        "sample": (f"rows 0..{rows - 1} of the same {npixel}x{nvoxel} line-of-sight matrix, {frames} warm-started frames, "
                   f"tolerance {SERIES_TOLERANCE:g}, iterations per frame {[int(v) for v in it]}, {t_solve:.2f} s on {ranks} "
                   f"thread-rank(s), scaled by {rows}/{npixel} rows"),
        "sample_seconds": t_solve, "sample_iterations": total_it,
    }


def run_reference_arm(args):
    vals, last = [], None
    t0 = time.perf_counter()
    series = args.workload == "c4"
    for i in range(args.warmup + args.steps):
        if series:
            last = cpu_series_baseline(args.npixel, args.nvoxel, frames=3, sample_rows=args.cpu_rows, threads=args.cpu_ranks or None)
        else:
            last = cpu_baseline(args.npixel, args.nvoxel, iterations=args.cpu_iterations, sample_rows=args.cpu_rows,
                                threads=args.cpu_ranks or None, laplacian=args.laplacian)
        if i >= args.warmup:
            vals.append(last)
        if time.perf_counter() - t0 > 240 and vals:
            break
    secs = sum(v["sample_seconds"] for v in vals)
    its = sum(v["sample_iterations"] for v in vals) if series else len(vals) * args.cpu_iterations
    value = its / secs * min(args.cpu_rows, args.npixel) / args.npixel
    cb = dict(last)
    cb["value"] = value
    cb.pop("sample_seconds", None)
    cb.pop("sample_iterations", None)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": len(vals),
        "warmup": args.warmup, "ms_per_step": 1e3 * secs / len(vals),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, 0), "cpu_baseline": cb,
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(args, n):
    gb = args.npixel * args.nvoxel * 4 / 1e9
    cfg = {
        "workload": f"BASELINE {args.workload_name}: synthetic {args.npixel}x{args.nvoxel} fp32 ray-transfer matrix "
                    f"({gb:.1f} GB){' + Laplacian' if args.laplacian else ''}, row-sharded over the GPUs",
        "npixel": args.npixel, "nvoxel": args.nvoxel, "iterations_per_step": args.iterations, "warm_start": True,
        "laplacian": args.laplacian, "beta_laplace": BETA, "seed": SEED,
        "sharding": f"rows/{max(n, 1)} (source/main.cpp:67-68)",
        "l2_policy": ("matrix >> L2 (126 MB); no flush needed between sweeps" if gb > 1.0 else
                      "matrix ~ L2-sized: sweeps partly hit in L2, so the HBM roofline can be exceeded; not flushed"),
    }
    if args.workload == "c4":
        cfg.update({"workload": f"BASELINE {args.workload_name}: warm-started series of {args.frames} frames on a synthetic "
                                f"{args.npixel}x{args.nvoxel} fp32 line-of-sight matrix ({gb:.1f} GB) + Laplacian, row-sharded over the GPUs",
                    "iterations_per_step": "until |dC| < conv_tolerance", "conv_tolerance": SERIES_TOLERANCE,
                    "frames": args.frames, "drift_per_frame": SERIES_DRIFT, "matrix_kind": "line of sight"})
    return cfg


# ---- the engine -----------------------------------------------------------------------------------------------------

def make_engine(sart, syn, cp, device, P, V, iterations, laplacian, sweep=0, kind=0, env=None, **kw):
    """Engine for this rank's shard of the P x V synthetic matrix, communicator and all. ``env`` = environment variables
    the engine reads while it is set up (exchange / persistence switches), restored afterwards."""
    saved = {}
    for k, v in (env or {}).items():
        saved[k] = os.environ.get(k)
        os.environ[k] = v
    try:
        kw.setdefault("beta_laplace", BETA)
        kw.setdefault("max_iterations", iterations)
        kw.setdefault("fixed_iterations", True)
        eng = sart.Engine(P, V, rank=cp.rank, nranks=cp.world, device=device, sweep_mode=sweep, **kw)
        if cp.world > 1:
            eng.comm_init(cp.broadcast_bytes(sart.comm_unique_id() if cp.rank == 0 else None))
        eng.fill_synthetic(SEED, kind)
        if laplacian and cp.rank == 0:
            eng.set_laplacian(*syn.laplacian_coo(V))  # the reference reads it on rank 0 only; finalize shares it
        eng.finalize()
    finally:
        for k, v in saved.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    return eng


class SingleRank:
    """Stands in for the control plane when one rank builds an engine of its own (no collectives)."""
    rank, world, local_rank = 0, 1, 0

    @staticmethod
    def broadcast_bytes(b):
        return b


def time_resident(eng, cp, torch, steps, warmup, iterations):
    """K resident solves (inputs already staged); device time from the engine's CUDA events, max over ranks."""
    acc = dict(dev_ms=0.0, launches=0, fp_ms=0.0, bp_ms=0.0, fu_ms=0.0, fp_n=0, bp_n=0, fu_n=0, solve_n=0)
    for _ in range(warmup):
        eng.solve_resident(True)
    torch.cuda.synchronize()
    cp.barrier()
    w0 = time.perf_counter()
    s = None
    for _ in range(steps):
        st, it = eng.solve_resident(True)
        assert it == iterations, (st, it)
        s = eng.stats()
        acc["dev_ms"] += s.solve_ms
        acc["launches"] += s.gpu_launches
        acc["fp_ms"] += s.fp_ms_avg * s.fp_launches
        acc["bp_ms"] += s.bp_ms_avg * s.bp_launches
        acc["fu_ms"] += s.fused_ms_avg * s.fused_launches
        acc["fp_n"] += s.fp_launches
        acc["bp_n"] += s.bp_launches
        acc["fu_n"] += s.fused_launches
        acc["solve_n"] += s.solve_launches
    torch.cuda.synchronize()  # (every step already ended with an event sync on the engine's own stream)
    cp.barrier()
    w1 = time.perf_counter()
    acc["t0"], acc["t1"], acc["last"] = w0, w1, s
    return acc


def roofline_of(acc, shard_bytes, iterations, value):
    """Per-kernel rates of this rank's shard from the device events of the timed region."""
    peak, peak_src = measured_peak()
    dev_ms = acc["dev_ms"]
    kern = {
        "fp_rows_kernel": {"ms": acc["fp_ms"] / max(acc["fp_n"], 1), "launches": acc["fp_n"],
                           "algorithmic_bytes": shard_bytes, "hbm_bytes": shard_bytes, "sweeps_per_launch": 1.0},
        "bp_cols_kernel": {"ms": acc["bp_ms"] / max(acc["bp_n"], 1), "launches": acc["bp_n"],
                           "algorithmic_bytes": shard_bytes, "hbm_bytes": shard_bytes, "sweeps_per_launch": 1.0},
        # one fused sweep does a forward and a back projection from one pass over HBM
        "fused_sweep_kernel": {"ms": acc["fu_ms"] / max(acc["fu_n"], 1), "launches": acc["fu_n"],
                               "algorithmic_bytes": 2 * shard_bytes, "hbm_bytes": shard_bytes, "sweeps_per_launch": 1.0},
    }
    if acc["solve_n"]:
        # the whole solve ran as ONE launch per step (fused_solve_kernel): I + 1 sweeps inside it, the last of which
        # only needs its forward projection -- SURVEY 8(d): I + 1 forward + I back projections = (2 I + 1) shards
        per = acc["fu_n"] / acc["solve_n"]
        kern.pop("fused_sweep_kernel")
        kern["fused_solve_kernel"] = {"ms": acc["fu_ms"] / acc["solve_n"], "launches": acc["solve_n"], "sweeps_per_launch": per,
                                      "algorithmic_bytes": (2 * per - 1) * shard_bytes, "hbm_bytes": shard_bytes * per}
    for k in kern.values():
        sec = k["ms"] * 1e-3
        k["algorithmic_gbs"] = k["algorithmic_bytes"] / sec / 1e9 if sec > 0 else 0.0
        k["hbm_gbs"] = k["hbm_bytes"] / sec / 1e9 if sec > 0 else 0.0
        k["share_of_step"] = k["ms"] * k["launches"] / dev_ms if dev_ms > 0 else 0.0
    dom = max(kern, key=lambda n: kern[n]["ms"] * kern[n]["launches"])
    d = kern[dom]
    return {
        "bound": "hbm", "kernel": dom, "achieved": d["hbm_gbs"], "peak": peak, "unit": "GB/s", "frac": d["hbm_gbs"] / peak,
        "traffic": profiled_traffic(dom, d["sweeps_per_launch"], shard_bytes),
        "traffic_source": "ncu dram__bytes_read.sum + dram__bytes_write.sum per sweep of the 40 GB shard (profiles/traffic.json), "
                          "scaled to this launch's sweeps and this rank's shard bytes",
        "peak_source": peak_src, "hbm_bytes_per_launch": d["hbm_bytes"],
        "frac_of_read_peak": d["hbm_gbs"] / READ_PEAK_GBS, "read_peak": READ_PEAK_GBS,
        "read_peak_source": "pure-read streaming kernel on this pool's B200 (profiles/r01_membw.txt); the measured copy peak is read+write",
        "algorithmic_bytes_per_launch": d["algorithmic_bytes"], "algorithmic_achieved": d["algorithmic_gbs"],
        "algorithmic_frac": d["algorithmic_gbs"] / peak,
        "iteration_roofline_frac": (2.0 * shard_bytes * value / 1e9) / peak,  # 2 * bytes(A) per iteration / peak
        "note": ("achieved / frac: bytes streamed from HBM per launch (one shard per sweep) over the launch's CUDA-event duration, "
                 "tails between the sweeps included. algorithmic_*: SURVEY 8(d) bills an iteration as two sweeps of the matrix; "
                 "the fused kernel serves both projections from one pass over HBM (and a second pass over shared memory), so that "
                 "figure exceeds the peak by design" if dom.startswith("fused_") else "one sweep of the shard per launch"),
        "kernels": kern,
    }


def run_checks(args, cp, sart, syn, eng, device, g, guess, xs, x_timed):
    """Correctness of what was just timed (outside the timed region). Returns (record, ok)."""
    P, V, I = args.npixel, args.nvoxel, args.iterations
    rec, ok = {}, True
    # (1) every rank holds the same solution, bit for bit
    hashes = cp.gather(bits_hash(x_timed))
    rec["ranks_agree"] = len(set(hashes)) == 1
    rec["solution_sha256_16"] = hashes[0]
    ok &= rec["ranks_agree"]
    # (2) a consistent frame started at its own solution stays there (no penalty): residual weights vanish
    eng.set_beta_laplace(0.0)
    eng.set_max_iterations(3)
    x_fp = eng.solve(g, xs)[0]
    eng.set_beta_laplace(BETA)
    eng.set_max_iterations(I)
    rec["fixed_point_residual"] = rel_l2(x_fp, xs)
    ok &= rec["fixed_point_residual"] <= 1e-5
    stats = eng.stats()
    eng.close()
    # (3) the same frame through the other kernel family and the other collective
    alt = make_engine(sart, syn, cp, device, P, V, I, args.laplacian, sweep=sart.SWEEP_SPLIT,
                      env={"SART_EXCHANGE": "nccl", "SART_PERSISTENT": "0"})
    x_alt, st_alt, it_alt = alt.solve(g, guess)
    s_alt = alt.stats()
    alt.close()
    rec["rel_l2_vs_split_nccl"] = float(cp.max(rel_l2(x_timed, x_alt)))
    rec["alt_path"] = {"sweep": "split (fp_rows_kernel + bp_cols_kernel)",
                       "exchange": {0: "none", 1: "nccl", 2: "peer-memory kernel"}.get(s_alt.exchange_mode, "?"),
                       "iterations": it_alt}
    ok &= rec["rel_l2_vs_split_nccl"] <= 1e-6 and it_alt == I
    # (4) rank 0: the first rows of the very same matrix against the reference's CPU solver
    if cp.rank == 0:
        try:
            from oracle.pyoracle import Reference, Restatement
            rows, iters = min(512, P), 6
            res = Restatement()
            H = res.fill_synthetic(rows, V, SEED)
            gs = res.forward(H, xs)
            lap = syn.laplacian_coo(V) if args.laplacian else None
            if Reference.available():
                x_ref = Reference().solve(H, gs, lap=lap, beta_laplace=BETA, max_iterations=iters, conv_tolerance=1e-300)[0]
                kind = "reference"
            else:
                x_ref = res.solve(H, gs, lap=lap, beta_laplace=BETA, max_iterations=iters, fixed_iterations=True)[0]
                kind = "port"
            small = make_engine(sart, syn, SingleRank, device, rows, V, iters, args.laplacian,
                                sweep={"auto": 0, "split": 1, "fused": 2}[args.sweep])
            x_small = small.solve(gs)[0]
            mode = small.stats().sweep_mode
            small.close()
            rec["row_subsample_vs_cpu"] = {"rows": rows, "iterations": iters, "rel_l2": rel_l2(x_small, x_ref), "oracle": kind,
                                           "sweep": {1: "split", 2: "fused"}.get(mode, "?")}
            ok &= rec["row_subsample_vs_cpu"]["rel_l2"] <= 1e-5
        except Exception as exc:  # the checker is missing on this machine: say so, do not fail the measurement
            rec["row_subsample_vs_cpu"] = {"unavailable": repr(exc)}
    cp.barrier()
    ok = bool(cp.max(0.0 if ok else 1.0) == 0.0)
    rec["passed"] = ok
    return rec, ok, stats


def c3_record(args, cp, sart, syn, device, torch):
    """BASELINE configs[2] at capacity: 125 000 rows x 100 000 voxels (50 GB) per GPU -- with 8 GPUs the 400 GB matrix --
    five fixed iterations per step, beside the same-sized shard alone on this GPU in the same process (no exchange, no
    peers): how much of the single-GPU rate survives sharding."""
    rows, V, I, steps, warm = args.c3_rows, 100000, 5, 3, 3
    P = rows * cp.world
    xs = syn.emission_profile(V)
    out = {"npixel": P, "nvoxel": V, "matrix_gb": P * V * 4 / 1e9, "shard_gb": rows * V * 4 / 1e9, "iterations_per_step": I,
           "steps": steps, "warmup": warm}
    sampler = ClockSampler(device) if cp.rank == 0 else None

    def timed(engine, plane):
        engine.stage_measurement(engine.forward_project(xs))
        engine.stage_guess(0.9 * xs + 0.01)
        acc = time_resident(engine, plane, torch, steps, warm, I)
        x = engine.fetch_solution()
        s = acc["last"]
        return acc, x, s

    eng = make_engine(sart, syn, cp, device, P, V, I, False)
    acc, x, s = timed(eng, cp)
    eng.close()
    ms_it = cp.max(acc["dev_ms"]) / (steps * I)
    hashes = cp.gather(bits_hash(x))
    out.update({"ms_per_iteration": ms_it, "iterations_per_s": 1e3 / ms_it, "ranks_agree": len(set(hashes)) == 1,
                "hbm_gbs_per_gpu": eng.nrows * V * 4 * (I + 1) / I / (ms_it * 1e-3) / 1e9,
                "between_sweeps_us": 1e3 * cp.max(s.tail_ms_avg), "cluster": [s.fused_cluster_size, s.fused_cluster_count],
                "clocks": sampler.window(acc["t0"], acc["t1"]) if sampler else None})

    class Alone(SingleRank):  # every rank times its own shard-sized matrix on its own GPU, all at the same time
        barrier = staticmethod(cp.barrier)
    one = make_engine(sart, syn, Alone, device, rows, V, I, False)
    acc1, _, _ = timed(one, Alone)
    one.close()
    ms_it1 = cp.max(acc1["dev_ms"]) / (steps * I)
    out.update({"single_shard_ms_per_iteration": ms_it1, "efficiency_vs_single_shard": ms_it1 / ms_it})
    if sampler:
        sampler.stop()
    return out


def run_engine_arm(args, cp):
    import torch
    sart = load_package()
    syn = load_module("synthetic")
    P, V, I = args.npixel, args.nvoxel, args.iterations
    device = cp.local_rank if cp.world > 1 else 0
    sweep = {"auto": 0, "split": 1, "fused": 2}[args.sweep]
    info = sart.device_query(device)
    shard_need = (P // cp.world + 1) * V * 4
    if shard_need + (6 << 30) > info.total_mem:
        raise SystemExit(f"bench.py: a shard of {shard_need / 1e9:.0f} GB does not fit one GPU; workload {args.workload} needs more GPUs")

    eng = make_engine(sart, syn, cp, device, P, V, I, args.laplacian, sweep=sweep)
    xs = syn.emission_profile(V)
    g_pin = torch.empty(eng.nrows, dtype=torch.float64).pin_memory()
    x_pin = torch.empty(V, dtype=torch.float64).pin_memory()
    g = g_pin.numpy()
    g[:] = eng.forward_project(xs)
    guess = 0.9 * xs + 0.01

    # ---- value: inputs resident in HBM -------------------------------------------------------------------------------
    eng.stage_measurement(g)
    eng.stage_guess(guess)
    sampler = ClockSampler(device) if cp.rank == 0 else None  # started before the warm-up so samples flow by t0
    acc = time_resident(eng, cp, torch, args.steps, args.warmup, I)
    s = acc["last"]
    sweep_mode = {1: "split", 2: "fused"}.get(s.sweep_mode, "?")
    exchange_mode = {0: "none", 1: "nccl", 2: "peer-memory kernel"}.get(s.exchange_mode, "?")
    clocks = sampler.window(acc["t0"], acc["t1"]) if sampler else None
    if sampler:
        sampler.stop()
    total_ms = cp.max(acc["dev_ms"])
    wall_ms = cp.max((acc["t1"] - acc["t0"]) * 1e3)
    value = args.steps * I / (total_ms / 1e3)
    tail_us = 1e3 * cp.max(s.tail_ms_avg)
    phases = [float(v) for v in s.tail_phase_us]

    # ---- e2e: host buffers through sart_solve ---------------------------------------------------------------------------
    x_host = x_pin.numpy()
    for _ in range(max(1, args.warmup // 2)):
        x_host[:] = guess
        eng.solve(g, x_host)
    cp.barrier()
    e0 = time.perf_counter()
    h2d = d2h = 0
    x_timed = None
    for _ in range(args.steps):
        x_host[:] = guess
        x_timed, st, it = eng.solve(g, x_host)
        se = eng.stats()
        h2d, d2h = se.h2d_bytes, se.d2h_bytes
    cp.barrier()
    e1 = time.perf_counter()
    e2e_s = cp.max(e1 - e0)
    e2e = {"value": args.steps * I / e2e_s, "unit": UNIT, "h2d_bytes_per_step": int(cp.sum(h2d)),
           "d2h_bytes_per_step": int(cp.sum(d2h)), "timing": "host wall clock around K sart_solve calls, max over ranks"}

    # ---- roofline of the dominant kernel (this rank's shard; device events inside the timed region) -------------------
    shard_bytes = eng.nrows * V * 4
    roofline = roofline_of(acc, shard_bytes, I, value)
    dom = roofline["kernel"]
    gathered = cp.gather({"rank": cp.rank, "dev_ms": acc["dev_ms"], "hbm_gbs": roofline["kernels"][dom]["hbm_gbs"],
                          "tail_phase_us": phases})
    geometry = {"cluster_size": s.fused_cluster_size, "clusters": s.fused_cluster_count,
                "sms_streaming": s.fused_cluster_size * s.fused_cluster_count}

    # ---- correctness of this very configuration -----------------------------------------------------------------------
    check, ok = None, True
    if not args.no_check:
        check, ok, _ = run_checks(args, cp, sart, syn, eng, device, g, guess, xs, x_timed)
    else:
        eng.close()

    extra = {}
    if args.extra_c3 == "on" or (args.extra_c3 == "auto" and cp.world == 8 and args.workload == "c2"):
        try:
            extra["c3"] = c3_record(args, cp, sart, syn, device, torch)
        except Exception as exc:
            extra["c3"] = {"error": repr(exc)}

    cb = None
    if cp.rank == 0 and cp.world == 1 and not args.no_cpu_baseline:
        cb = cpu_baseline(P, V, iterations=args.cpu_iterations, sample_rows=args.cpu_rows, threads=args.cpu_ranks or None,
                          laplacian=args.laplacian)
        cb.pop("sample_seconds", None)

    if cp.rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": cp.world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": dict(workload_config(args, cp.world), sweep=sweep_mode, exchange=exchange_mode), "e2e": e2e,
            "between_sweeps_us": tail_us,
            "between_sweeps_phases_us": dict(zip(["barrier1_wait", "cluster_reduce", "flag_round_trip", "peer_reads_update",
                                                  "barrier2_wait"], phases)),
            "gpu_launches": acc["launches"], "geometry": geometry,
            "roofline": roofline, "check": check, "cpu_baseline": cb, "clocks": clocks,
            "wall_ms_per_step": wall_ms / args.steps, "per_rank": gathered,
            "device": {"name": info.name.decode(), "sms": info.sm_count, "l2_bytes": info.l2_bytes,
                       "total_mem": info.total_mem},
            "host": {"cores": os.cpu_count()},
        }
        if extra:
            line["extra"] = extra
        print(json.dumps(line), flush=True)
    return ok


def run_series_arm(args, cp):
    """BASELINE configs[3]: --frames warm-started frames through sart_solve_series, in chunks of --chunk frames (the
    reference's --max_cached_frames), each frame iterated until the reference's stopping rule fires. One "step" is the
    whole series; value = iterations/s over it, frames/s beside it."""
    import torch
    sart = load_package()
    syn = load_module("synthetic")
    P, V, F = args.npixel, args.nvoxel, args.frames
    device = cp.local_rank if cp.world > 1 else 0
    info = sart.device_query(device)
    eng = make_engine(sart, syn, cp, device, P, V, 2000, True, sweep={"auto": 0, "split": 1, "fused": 2}[args.sweep],
                      kind=sart.SYNTHETIC_LOS, max_iterations=2000, fixed_iterations=False, conv_tolerance=SERIES_TOLERANCE)
    chunk = max(1, min(args.chunk, F))
    g_pin = torch.empty((chunk, eng.nrows), dtype=torch.float64).pin_memory()
    g = g_pin.numpy()

    def frames_of(k0, n):
        for k in range(n):
            g[k] = eng.forward_project(syn.emission_profile(V, t=SERIES_DRIFT * (k0 + k)))
        return g[:n]

    # warm-up: a few frames from the start of the series (also allocates the series buffers), then the timed series
    truth0 = syn.emission_profile(V)
    x_prev = eng.solve_series(frames_of(0, min(4, chunk)), guess=truth0)[0][-1]
    sampler = ClockSampler(device) if cp.rank == 0 else None
    its, sts = [], []
    dev_ms = wall = 0.0
    passes = handed = repeated = launches = 0
    h2d = d2h = 0
    t_first = time.perf_counter()
    x_last, done = None, 0
    x_prev = truth0
    while done < F:
        n = min(chunk, F - done)
        frames = frames_of(done, n)  # generating the frames is not part of the solve
        cp.barrier()
        t0 = time.perf_counter()
        xs_, st_, it_ = eng.solve_series(frames, guess=x_prev)
        cp.barrier()
        wall += time.perf_counter() - t0
        s = eng.stats()
        dev_ms += s.solve_ms
        passes += s.sweeps
        handed += s.fp_launches
        repeated += s.bp_launches
        launches += s.gpu_launches
        h2d += s.h2d_bytes
        d2h += s.d2h_bytes
        its.extend(int(v) for v in it_)
        sts.extend(int(v) for v in st_)
        x_prev = x_last = xs_[-1].copy()
        done += n
    t_last = time.perf_counter()
    clocks = sampler.window(t_first, t_last) if sampler else None
    if sampler:
        sampler.stop()
    total_ms = cp.max(dev_ms)
    wall_s = cp.max(wall)
    total_it = int(np.sum(its))
    truth = syn.emission_profile(V, t=SERIES_DRIFT * (F - 1))
    hashes = cp.gather(bits_hash(x_last))
    shard_bytes = eng.nrows * V * 4
    peak, peak_src = measured_peak()
    hbm_gbs = shard_bytes * passes / (total_ms * 1e-3) / 1e9
    exchange_mode = {0: "none", 1: "nccl", 2: "peer-memory kernel"}.get(s.exchange_mode, "?")
    eng.close()
    cb = None
    if cp.rank == 0 and cp.world == 1 and not args.no_cpu_baseline:
        cb = cpu_series_baseline(P, V, frames=3, sample_rows=args.cpu_rows, threads=args.cpu_ranks or None)
        cb.pop("sample_seconds", None)
        cb.pop("sample_iterations", None)
    ok = len(set(hashes)) == 1 and all(v == 0 for v in sts)
    h2d_all, d2h_all = int(cp.sum(h2d)), int(cp.sum(d2h))  # (collectives: every rank)
    if cp.rank == 0:
        line = {
            "metric": METRIC, "value": total_it / (total_ms / 1e3), "unit": UNIT, "n_gpus": cp.world, "steps": 1,
            "warmup": 1, "ms_per_step": total_ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": dict(workload_config(args, cp.world), sweep="fused", exchange=exchange_mode, chunk=chunk),
            "e2e": {"value": total_it / wall_s, "unit": UNIT, "h2d_bytes_per_step": h2d_all, "d2h_bytes_per_step": d2h_all,
                    "timing": "host wall clock around the sart_solve_series calls (pinned host frames in, solutions out), max over ranks"},
            "series": {"frames": F, "frames_per_s": F / (total_ms / 1e3), "frames_per_s_e2e": F / wall_s,
                       "iterations_total": total_it, "iterations_median": float(np.median(its)),
                       "iterations_min": int(np.min(its)), "iterations_max": int(np.max(its)),
                       "matrix_passes": int(passes), "passes_without_lookahead": int(total_it + F),
                       "frames_handed_their_first_iteration": int(handed), "sweeps_repeated": int(repeated),
                       "all_converged": all(v == 0 for v in sts), "last_frame_rel_l2_vs_truth": rel_l2(x_last, truth),
                       "ms_per_frame": total_ms / F},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": "fused_solve_kernel", "achieved": hbm_gbs, "peak": peak, "unit": "GB/s",
                         "frac": hbm_gbs / peak, "traffic": None, "peak_source": peak_src,
                         "frac_of_read_peak": hbm_gbs / READ_PEAK_GBS,
                         "note": "bytes streamed from HBM (matrix passes x shard) over the device time of the whole series, per-frame set-up and copies included"},
            "check": {"ranks_agree": len(set(hashes)) == 1, "all_converged": all(v == 0 for v in sts), "passed": ok},
            "cpu_baseline": cb, "clocks": clocks,
            "device": {"name": info.name.decode(), "sms": info.sm_count, "l2_bytes": info.l2_bytes, "total_mem": info.total_mem},
            "host": {"cores": os.cpu_count()},
        }
        print(json.dumps(line), flush=True)
    return ok


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS),
                    help="c2: BASELINE configs[1] (the headline, default); c1: configs[0], 8192x4096, 100 iterations, no "
                         "Laplacian, CPU reference on the whole problem with 2 ranks; c3 / c5: configs[2] / configs[4] "
                         "(400 GB / 1.2 TB: as many GPUs as hold the shards); c4: configs[3], the warm-started series")
    ap.add_argument("--iterations", type=int, default=0, help="SART iterations per step (fixed); 0 = the workload's own")
    ap.add_argument("--npixel", type=int, default=0)
    ap.add_argument("--nvoxel", type=int, default=0)
    ap.add_argument("--frames", type=int, default=1000, help="c4: frames of the series")
    ap.add_argument("--chunk", type=int, default=250, help="c4: frames per sart_solve_series call")
    ap.add_argument("--cpu-rows", type=int, default=0, help="rows of the matrix the CPU baseline is timed on")
    ap.add_argument("--cpu-iterations", type=int, default=0)
    ap.add_argument("--cpu-ranks", type=int, default=0, help="thread-ranks of the CPU baseline (0 = all host cores)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-check", action="store_true", help="skip the correctness record (profiling runs)")
    ap.add_argument("--sweep", default="auto", choices=["auto", "split", "fused"])
    ap.add_argument("--extra-c3", default="auto", choices=["auto", "on", "off"],
                    help="add the 50 GB-per-GPU configs[2] record (auto: with 8 GPUs on the default workload)")
    ap.add_argument("--c3-rows", type=int, default=125000, help="rows per GPU of that record")
    args = ap.parse_args()
    w = WORKLOADS[args.workload]
    args.workload_name, args.laplacian = w["name"], w["laplacian"]
    args.npixel = args.npixel or w["npixel"]
    args.nvoxel = args.nvoxel or w["nvoxel"]
    args.iterations = args.iterations or w["iterations"]
    args.cpu_rows = args.cpu_rows or w["cpu_rows"]
    args.cpu_iterations = args.cpu_iterations or w["cpu_iterations"]
    args.cpu_ranks = args.cpu_ranks or w["cpu_ranks"]
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    if args.impl == "reference":
        # CPU only: rank 0 alone works, the other ranks of a torchrun launch leave at once (no rendezvous needed)
        if int(os.environ.get("RANK", "0")) == 0:
            run_reference_arm(args)
        return 0

    cpl = load_module("control_plane")
    cp = cpl.ControlPlane.from_env()
    if cp.world != max(args.gpus, 1) and cp.rank == 0:
        print(f"bench.py: --gpus {args.gpus} but WORLD_SIZE={cp.world}; using WORLD_SIZE", file=sys.stderr)
    try:
        ok = run_series_arm(args, cp) if args.workload == "c4" else run_engine_arm(args, cp)
    finally:
        cp.shutdown()
    if not ok:
        print("bench.py: a correctness check of the timed configuration failed (see \"check\" in the line above)", file=sys.stderr)
        return 3
    return 0


if __name__ == "__main__":
    sys.exit(main())
