"""TEST INFRASTRUCTURE ONLY -- ctypes access to the two CPU checkers.

* ``Restatement``  -> oracle/libsart_oracle.so  (oracle/sart_oracle.c, our C restatement)
* ``Reference``    -> oracle/_ref/libsart_ref.so (the reference's own sartsolver.cpp, compiled in place
  by oracle/Makefile with the thread-per-rank MPI shim)

Only tests/, ``__graft_entry__.smoke()`` and ``bench.py`` (cpu_baseline / ``--impl reference``) may import
this module, and only to check or to time the CPU path. The product never does.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(_HERE, "libsart_oracle.so")
REF_SO = os.path.join(_HERE, "_ref", "libsart_ref.so")
REF_GPU_SO = os.path.join(_HERE, "_ref", "libsart_ref_gpu.so")  # the reference's own CUDA solver, built for sm_100a

SUCCESS = 0
MAX_ITERATIONS_EXCEEDED = -1

_u64 = C.c_uint64
_pf = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_pd = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")


class OracleParams(C.Structure):
    _fields_ = [
        ("ray_density_threshold", C.c_double),
        ("ray_length_threshold", C.c_double),
        ("conv_tolerance", C.c_double),
        ("beta_laplace", C.c_double),
        ("relaxation", C.c_double),
        ("max_iterations", C.c_int),
        ("logarithmic", C.c_int),
        ("gpu_semantics", C.c_int),
        ("fixed_iterations", C.c_int),
        ("nranks", C.c_int),
    ]


def _lap_args(lap):
    if lap is None or len(lap[0]) == 0:
        return None, None, 0, ()
    idx = np.ascontiguousarray(lap[0], dtype=np.uint64)
    val = np.ascontiguousarray(lap[1], dtype=np.float32)
    return idx.ctypes.data_as(C.c_void_p), val.ctypes.data_as(C.c_void_p), len(idx), (idx, val)


class Restatement:
    """oracle/sart_oracle.c through ctypes."""

    def __init__(self, path=ORACLE_SO):
        self.lib = C.CDLL(path)
        L = self.lib
        L.oracle_shard_range.argtypes = [_u64, C.c_int, C.c_int, C.POINTER(_u64), C.POINTER(_u64)]
        L.oracle_ray_sums.argtypes = [_pf, _u64, _u64, C.c_int, _pd, _pd]
        L.oracle_solve.argtypes = [_pf, _u64, _u64, C.c_void_p, C.c_void_p, _u64, _pd, _pd, _pd, _pd, C.c_int,
                                   C.POINTER(OracleParams), C.POINTER(C.c_int), C.c_void_p]
        L.oracle_solve.restype = C.c_int
        L.oracle_fill_synthetic.argtypes = [_pf, _u64, _u64, _u64, _u64]
        L.oracle_fill_los.argtypes = [_pf, _u64, _u64, _u64, _u64]
        L.oracle_los_exponent.argtypes = [_u64]
        L.oracle_los_exponent.restype = C.c_int
        L.oracle_forward.argtypes = [_pf, _u64, _u64, _pd, _pd]
        L.oracle_synth_exponent.argtypes = [_u64]
        L.oracle_synth_exponent.restype = C.c_int

    def shard_range(self, npixel, nranks, rank):
        off, n = _u64(), _u64()
        self.lib.oracle_shard_range(npixel, nranks, rank, C.byref(off), C.byref(n))
        return off.value, n.value

    def fill_synthetic(self, npixel, nvoxel, seed, row0=0, kind=0):
        """kind 0: the dense matrix, 1: the line-of-sight matrix (SART_SYNTHETIC_DENSE / _LOS of the product)."""
        H = np.empty((npixel, nvoxel), dtype=np.float32)
        (self.lib.oracle_fill_los if kind == 1 else self.lib.oracle_fill_synthetic)(H, row0, npixel, nvoxel, seed)
        return H

    def forward(self, H, x):
        out = np.empty(H.shape[0], dtype=np.float64)
        self.lib.oracle_forward(H, H.shape[0], H.shape[1], np.ascontiguousarray(x, dtype=np.float64), out)
        return out

    def ray_sums(self, H, nranks=1):
        P, V = H.shape
        rho, lam = np.empty(V), np.empty(P)
        self.lib.oracle_ray_sums(H, P, V, nranks, rho, lam)
        return rho, lam

    def solve(self, H, g, x0=None, lap=None, ray_density_threshold=1e-6, ray_length_threshold=1e-6,
              conv_tolerance=1e-5, beta_laplace=2e-2, relaxation=1.0, max_iterations=2000, logarithmic=False,
              gpu_semantics=False, fixed_iterations=False, nranks=1, sums=None, history=False):
        """Returns (x, status, iterations[, conv_history])."""
        P, V = H.shape
        rho, lam = sums if sums is not None else self.ray_sums(H, nranks)
        prm = OracleParams(ray_density_threshold, ray_length_threshold, conv_tolerance, beta_laplace, relaxation,
                           max_iterations, int(logarithmic), int(gpu_semantics), int(fixed_iterations), nranks)
        x = np.zeros(V) if x0 is None else np.array(x0, dtype=np.float64, copy=True)
        li, lv, nnz, _keep = _lap_args(lap)
        it = C.c_int(0)
        hist = np.full(max_iterations, np.nan) if history else None
        st = self.lib.oracle_solve(H, P, V, li, lv, nnz, rho, lam, np.ascontiguousarray(g, dtype=np.float64), x,
                                   0 if x0 is None else 1, C.byref(prm), C.byref(it),
                                   hist.ctypes.data_as(C.c_void_p) if history else None)
        return (x, st, it.value, hist) if history else (x, st, it.value)


class Reference:
    """The reference's own SARTSolverMPI / LogSARTSolverMPI object code through oracle/ref_harness.cpp."""

    def __init__(self, path=REF_SO):
        self.lib = C.CDLL(path)
        f = self.lib.ref_sart_solve_frames
        f.argtypes = [_pf, _u64, _u64, C.c_void_p, C.c_void_p, _u64, _pd, _pd, C.c_int, C.c_int, C.c_int, _pd,
                      C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, _pd]
        f.restype = C.c_int

    @staticmethod
    def available(path=REF_SO):
        return os.path.exists(path)

    def solve_frames(self, H, frames, x0=None, lap=None, no_guess=False, ray_density_threshold=1e-6,
                     ray_length_threshold=1e-6, conv_tolerance=1e-5, beta_laplace=2e-2, relaxation=1.0,
                     max_iterations=2000, logarithmic=False, nranks=1):
        """frames [F][P] -> (x [F][V], status [F], iterations [F], (ctor_seconds, solve_seconds)); the wall time of
        each frame's solve is left in ``self.frame_seconds``."""
        P, V = H.shape
        frames = np.ascontiguousarray(np.atleast_2d(frames), dtype=np.float64)
        F = frames.shape[0]
        x = np.zeros((F, V))
        if x0 is not None:
            x[0] = x0
        status = np.zeros(F, dtype=np.int32)
        iters = np.zeros(F, dtype=np.int32)
        secs = np.zeros(2 + F)
        prm = np.array([ray_density_threshold, ray_length_threshold, conv_tolerance, beta_laplace, relaxation])
        li, lv, nnz, _keep = _lap_args(lap)
        rc = self.lib.ref_sart_solve_frames(H, P, V, li, lv, nnz, frames, x, F, 0 if x0 is None else 1,
                                            int(no_guess), prm, max_iterations, int(logarithmic), nranks,
                                            status.ctypes.data_as(C.c_void_p), iters.ctypes.data_as(C.c_void_p), secs)
        if rc != 0:
            raise ValueError("reference would reject these arguments (sartsolver.cpp:63-123)")
        self.frame_seconds = secs[2:].copy()
        return x, status, iters, (secs[0], secs[1])

    def solve(self, H, g, x0=None, **kw):
        x, st, it, secs = self.solve_frames(H, np.asarray(g)[None, :], x0=x0, **kw)
        return x[0], int(st[0]), int(it[0])
