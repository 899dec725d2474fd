"""ctypes view of libsart_b200.so (include/sart_b200.h) for the tests and bench.py.

The product is the shared library; this module only binds its C ABI and offers ``Engine``, a thin object with
the reference solver's vocabulary (``solve(measurement, solution)``, the six setters, status codes:
/root/reference/source/include/sartsolver.hpp:36-51). There is no Python or CPU implementation behind it: if the
library is missing or no CUDA device is present every entry point fails loudly.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libsart_b200.so")

OK = 0
ERR_INVALID_ARGUMENT, ERR_CUDA, ERR_STATE, ERR_NCCL, ERR_NO_DEVICE = 1, 2, 3, 4, 5
SUCCESS = 0
MAX_ITERATIONS_EXCEEDED = -1
SEMANTICS_CPU, SEMANTICS_GPU = 0, 1
SWEEP_AUTO, SWEEP_SPLIT, SWEEP_FUSED = 0, 1, 2
EXCHANGE_NONE, EXCHANGE_NCCL, EXCHANGE_PEER = 0, 1, 2
UNIQUE_ID_BYTES = 128
SYNTHETIC_DENSE, SYNTHETIC_LOS = 0, 1

# every symbol include/sart_b200.h declares (tests check the library exports exactly these)
SYMBOLS = [
    "sart_last_error", "sart_config_default", "sart_shard_range", "sart_device_query", "sart_create", "sart_destroy",
    "sart_comm_unique_id", "sart_comm_init", "sart_upload_rows", "sart_fill_synthetic", "sart_download_rows",
    "sart_set_laplacian_coo", "sart_finalize", "sart_get_ray_sums", "sart_set_ray_density_threshold",
    "sart_set_ray_length_threshold", "sart_set_convolution_tolerance", "sart_set_beta_laplace", "sart_set_relaxation",
    "sart_set_max_iterations", "sart_set_fixed_iterations", "sart_get_config", "sart_solve", "sart_stage_measurement",
    "sart_stage_guess", "sart_solve_resident", "sart_fetch_solution", "sart_forward_project", "sart_get_stats",
    "sart_kernel_forward", "sart_kernel_backward", "sart_back_project_noalloc", "sart_load_rows_file", "sart_solve_batch",
    "sart_solve_series", "sart_fill_synthetic_kind",
]


class Config(C.Structure):
    _fields_ = [
        ("npixel", C.c_uint64), ("nvoxel", C.c_uint64),
        ("rank", C.c_int32), ("nranks", C.c_int32), ("device", C.c_int32), ("logarithmic", C.c_int32),
        ("semantics", C.c_int32), ("max_iterations", C.c_int32), ("fixed_iterations", C.c_int32),
        ("sweep_mode", C.c_int32),
        ("ray_density_threshold", C.c_double), ("ray_length_threshold", C.c_double), ("conv_tolerance", C.c_double),
        ("beta_laplace", C.c_double), ("relaxation", C.c_double),
    ]


class Stats(C.Structure):
    _fields_ = [
        ("status", C.c_int32), ("iterations", C.c_int32), ("sweeps", C.c_int32), ("gpu_launches", C.c_int32),
        ("solve_ms", C.c_double), ("iterate_ms", C.c_double), ("fp_ms_avg", C.c_double), ("bp_ms_avg", C.c_double),
        ("fp_launches", C.c_int32), ("bp_launches", C.c_int32),
        ("matrix_bytes", C.c_uint64), ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64),
        ("fused_ms_avg", C.c_double), ("fused_launches", C.c_int32), ("sweep_mode", C.c_int32),
        ("exchange_mode", C.c_int32), ("solve_launches", C.c_int32), ("tail_ms_avg", C.c_double),
        ("fused_cluster_size", C.c_int32), ("fused_cluster_count", C.c_int32), ("tail_phase_us", C.c_double * 5),
    ]


class DeviceInfo(C.Structure):
    _fields_ = [
        ("name", C.c_char * 256), ("sm_count", C.c_int32), ("cc_major", C.c_int32), ("cc_minor", C.c_int32),
        ("l2_bytes", C.c_int32), ("reserved", C.c_int32), ("total_mem", C.c_uint64), ("free_mem", C.c_uint64),
    ]


class SartError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"[sart_b200 error {code}] {message}")
        self.code = code
        self.message = message


_lib = None


def lib():
    """Loads libsart_b200.so. Raises if it has not been built -- there is nothing to fall back to."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FileNotFoundError(f"{LIB_PATH} is not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                                    "(or make -C mpi-cuda-sartsolver_b200/csrc); this package has no other implementation")
        L = C.CDLL(LIB_PATH)
        L.sart_last_error.restype = C.c_char_p
        _u64, _i32, _vp, _dbl = C.c_uint64, C.c_int32, C.c_void_p, C.c_double
        L.sart_config_default.argtypes = [C.POINTER(Config)]
        L.sart_shard_range.argtypes = [_u64, _i32, _i32, C.POINTER(_u64), C.POINTER(_u64)]
        L.sart_device_query.argtypes = [_i32, C.POINTER(DeviceInfo)]
        L.sart_create.argtypes = [C.POINTER(Config), C.POINTER(_vp)]
        L.sart_destroy.argtypes = [_vp]
        L.sart_comm_unique_id.argtypes = [_vp]
        L.sart_comm_init.argtypes = [_vp, _vp]
        L.sart_upload_rows.argtypes = [_vp, _u64, _u64, _u64, _u64, _vp, _u64]
        L.sart_fill_synthetic.argtypes = [_vp, _u64]
        L.sart_fill_synthetic_kind.argtypes = [_vp, _u64, _i32]
        L.sart_load_rows_file.argtypes = [_vp, C.c_char_p, _u64, _u64, _u64, _u64, _u64, _u64]
        L.sart_download_rows.argtypes = [_vp, _u64, _u64, _vp]
        L.sart_set_laplacian_coo.argtypes = [_vp, _vp, _vp, _u64]
        L.sart_finalize.argtypes = [_vp]
        L.sart_get_ray_sums.argtypes = [_vp, _vp, _vp]
        for name in ("ray_density_threshold", "ray_length_threshold", "convolution_tolerance", "beta_laplace", "relaxation"):
            getattr(L, "sart_set_" + name).argtypes = [_vp, _dbl]
        L.sart_set_max_iterations.argtypes = [_vp, _i32]
        L.sart_set_fixed_iterations.argtypes = [_vp, _i32]
        L.sart_get_config.argtypes = [_vp, C.POINTER(Config)]
        L.sart_solve.argtypes = [_vp, _vp, _vp, _i32, C.POINTER(_i32), C.POINTER(_i32)]
        L.sart_solve_batch.argtypes = [_vp, _i32, _vp, _vp, _vp, _vp]
        L.sart_solve_series.argtypes = [_vp, _i32, _vp, _vp, _vp, _vp, _vp]
        L.sart_stage_measurement.argtypes = [_vp, _vp]
        L.sart_stage_guess.argtypes = [_vp, _vp]
        L.sart_solve_resident.argtypes = [_vp, _i32, C.POINTER(_i32), C.POINTER(_i32)]
        L.sart_fetch_solution.argtypes = [_vp, _vp]
        L.sart_forward_project.argtypes = [_vp, _vp, _vp]
        L.sart_get_stats.argtypes = [_vp, C.POINTER(Stats)]
        L.sart_kernel_forward.argtypes = [_vp, _vp, _vp]
        L.sart_kernel_backward.argtypes = [_vp, _vp, _vp]
        L.sart_back_project_noalloc.argtypes = [_vp, _u64, _u64, _u64, _vp, _vp, _vp]
        for name in SYMBOLS:
            if name != "sart_last_error":
                getattr(L, name).restype = C.c_int
        _lib = L
    return _lib


def check(rc):
    if rc != OK:
        raise SartError(rc, lib().sart_last_error().decode())


def default_config(**overrides):
    cfg = Config()
    check(lib().sart_config_default(C.byref(cfg)))
    for k, v in overrides.items():
        setattr(cfg, k, v)
    return cfg


def shard_range(npixel, nranks, rank):
    off, n = C.c_uint64(), C.c_uint64()
    check(lib().sart_shard_range(npixel, nranks, rank, C.byref(off), C.byref(n)))
    return off.value, n.value


def device_query(device=0):
    info = DeviceInfo()
    check(lib().sart_device_query(device, C.byref(info)))
    return info


def comm_unique_id():
    buf = (C.c_uint8 * UNIQUE_ID_BYTES)()
    check(lib().sart_comm_unique_id(buf))
    return bytes(buf)


def _header_constant(name):
    """The value include/sart_b200.h gives a `#define NAME (A [<< B])` constant: the header is the one place the
    limit of the allocation-free entry point is written down (the library checks its own array against it)."""
    import re
    with open(os.path.join(os.path.dirname(_HERE), "include", "sart_b200.h")) as f:
        m = re.search(r"#define\s+" + name + r"\s+\((\d+)u?(?:\s*<<\s*(\d+))?\)", f.read())
    return int(m.group(1)) << int(m.group(2) or 0)


NOALLOC_MAX_LD = _header_constant("SART_NOALLOC_MAX_LD")
NOALLOC_SCRATCH_BYTES = _header_constant("SART_NOALLOC_SCRATCH_BYTES")


def back_project_noalloc(d_matrix, nrows, ncols, ld, d_weights, d_out, stream=None):
    """sart_back_project_noalloc on raw DEVICE addresses (ints): for callers whose GPU memory is entirely theirs and
    who therefore cannot let a library allocate. Asynchronous on ``stream`` (a cudaStream_t as int, None = default)."""
    check(lib().sart_back_project_noalloc(d_matrix, nrows, ncols, ld, d_weights, d_out, stream))


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


class Engine:
    """One row shard of the ray-transfer matrix on one GPU + the solver over it (one per process/rank)."""

    def __init__(self, npixel, nvoxel, rank=0, nranks=1, device=-1, logarithmic=False, semantics=SEMANTICS_CPU,
                 ray_density_threshold=1e-6, ray_length_threshold=1e-6, conv_tolerance=1e-5, beta_laplace=1e-2,
                 relaxation=1.0, max_iterations=2000, fixed_iterations=False, sweep_mode=SWEEP_AUTO):
        self._h = C.c_void_p()
        cfg = default_config(npixel=npixel, nvoxel=nvoxel, rank=rank, nranks=nranks, device=device,
                             logarithmic=int(logarithmic), semantics=semantics, max_iterations=max_iterations,
                             fixed_iterations=int(fixed_iterations), sweep_mode=sweep_mode,
                             ray_density_threshold=ray_density_threshold,
                             ray_length_threshold=ray_length_threshold, conv_tolerance=conv_tolerance,
                             beta_laplace=beta_laplace, relaxation=relaxation)
        check(lib().sart_create(C.byref(cfg), C.byref(self._h)))
        self.npixel, self.nvoxel, self.rank, self.nranks = npixel, nvoxel, rank, nranks
        self.row_offset, self.nrows = shard_range(npixel, nranks, rank)

    def close(self):
        if self._h:
            lib().sart_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -- set-up ----------------------------------------------------------------------------------------------------
    def comm_init(self, unique_id):
        buf = (C.c_uint8 * UNIQUE_ID_BYTES).from_buffer_copy(unique_id)
        check(lib().sart_comm_init(self._h, buf))

    def upload_rows(self, block, local_row0=0, col0=0):
        block = np.asarray(block, dtype=np.float32)
        if block.ndim != 2 or block.strides[1] != 4:
            block = np.ascontiguousarray(block)
        check(lib().sart_upload_rows(self._h, local_row0, block.shape[0], col0, block.shape[1],
                                     block.ctypes.data_as(C.c_void_p), block.strides[0] // 4))

    def load_rows_file(self, path, file_offset, file_ld, local_row0, nrows, col0, ncols):
        check(lib().sart_load_rows_file(self._h, os.fsencode(path), file_offset, file_ld, local_row0, nrows, col0, ncols))

    def fill_synthetic(self, seed, kind=SYNTHETIC_DENSE):
        check(lib().sart_fill_synthetic_kind(self._h, seed, kind))

    def download_rows(self, local_row0=0, nrows=None):
        nrows = self.nrows - local_row0 if nrows is None else nrows
        out = np.empty((nrows, self.nvoxel), dtype=np.float32)
        check(lib().sart_download_rows(self._h, local_row0, nrows, out.ctypes.data_as(C.c_void_p)))
        return out

    def set_laplacian(self, flat_index, value):
        idx = np.ascontiguousarray(flat_index, dtype=np.uint64)
        val = np.ascontiguousarray(value, dtype=np.float32)
        check(lib().sart_set_laplacian_coo(self._h, idx.ctypes.data_as(C.c_void_p), val.ctypes.data_as(C.c_void_p), len(idx)))

    def finalize(self):
        check(lib().sart_finalize(self._h))

    def ray_sums(self):
        rho, lam = np.empty(self.nvoxel), np.empty(self.nrows)
        check(lib().sart_get_ray_sums(self._h, rho.ctypes.data_as(C.c_void_p), lam.ctypes.data_as(C.c_void_p)))
        return rho, lam

    # -- the reference's setters (sartsolver.hpp:43-48) -----------------------------------------------------------------
    def set_ray_density_threshold(self, v):
        check(lib().sart_set_ray_density_threshold(self._h, v))

    def set_ray_length_threshold(self, v):
        check(lib().sart_set_ray_length_threshold(self._h, v))

    def set_convolution_tolerance(self, v):
        check(lib().sart_set_convolution_tolerance(self._h, v))

    def set_beta_laplace(self, v):
        check(lib().sart_set_beta_laplace(self._h, v))

    def set_relaxation(self, v):
        check(lib().sart_set_relaxation(self._h, v))

    def set_max_iterations(self, v):
        check(lib().sart_set_max_iterations(self._h, v))

    def set_fixed_iterations(self, on):
        check(lib().sart_set_fixed_iterations(self._h, int(on)))

    def config(self):
        cfg = Config()
        check(lib().sart_get_config(self._h, C.byref(cfg)))
        return cfg

    # -- solve ---------------------------------------------------------------------------------------------------------
    def solve(self, measurement, solution=None):
        """``solution=None`` is the reference's empty vector (default initial guess). Returns (x, status, iterations)."""
        g = _f64(measurement)
        if g.shape != (self.nrows,):
            raise ValueError(f"measurement must hold this rank's {self.nrows} pixels")
        if solution is not None and np.shape(solution) != (self.nvoxel,):
            # sartsolver.cpp:139-142
            raise ValueError("Solution vector must be empty or contain nvoxel elements.")
        x = np.zeros(self.nvoxel) if solution is None else np.array(solution, dtype=np.float64, copy=True)
        st, it = C.c_int32(), C.c_int32()
        check(lib().sart_solve(self._h, g.ctypes.data_as(C.c_void_p), x.ctypes.data_as(C.c_void_p),
                               0 if solution is None else 1, C.byref(st), C.byref(it)))
        return x, st.value, it.value

    def solve_batch(self, frames):
        """Independent frames (no warm start), solved four per matrix pass where possible.
        frames [F][nrows] -> (x [F][nvoxel], status [F], iterations [F])."""
        g = np.ascontiguousarray(np.atleast_2d(frames), dtype=np.float64)
        if g.shape[1] != self.nrows:
            raise ValueError(f"each frame must hold this rank's {self.nrows} pixels")
        F = g.shape[0]
        x = np.zeros((F, self.nvoxel))
        st, it = np.zeros(F, dtype=np.int32), np.zeros(F, dtype=np.int32)
        check(lib().sart_solve_batch(self._h, F, g.ctypes.data_as(C.c_void_p), x.ctypes.data_as(C.c_void_p),
                                     st.ctypes.data_as(C.c_void_p), it.ctypes.data_as(C.c_void_p)))
        return x, st, it

    def solve_series(self, frames, guess=None):
        """Warm-started series: frame k starts from the solution of frame k - 1 (frame 0 from ``guess`` or the default
        initial guess). frames [F][nrows] -> (x [F][nvoxel], status [F], iterations [F])."""
        g = np.ascontiguousarray(np.atleast_2d(frames), dtype=np.float64)
        if g.shape[1] != self.nrows:
            raise ValueError(f"each frame must hold this rank's {self.nrows} pixels")
        F = g.shape[0]
        x = np.zeros((F, self.nvoxel))
        st, it = np.zeros(F, dtype=np.int32), np.zeros(F, dtype=np.int32)
        x0 = None if guess is None else _f64(guess)
        check(lib().sart_solve_series(self._h, F, g.ctypes.data_as(C.c_void_p),
                                      None if x0 is None else x0.ctypes.data_as(C.c_void_p), x.ctypes.data_as(C.c_void_p),
                                      st.ctypes.data_as(C.c_void_p), it.ctypes.data_as(C.c_void_p)))
        return x, st, it

    def stage_measurement(self, measurement):
        g = _f64(measurement)
        check(lib().sart_stage_measurement(self._h, g.ctypes.data_as(C.c_void_p)))

    def stage_guess(self, solution):
        x = _f64(solution)
        check(lib().sart_stage_guess(self._h, x.ctypes.data_as(C.c_void_p)))

    def solve_resident(self, has_guess):
        st, it = C.c_int32(), C.c_int32()
        check(lib().sart_solve_resident(self._h, int(has_guess), C.byref(st), C.byref(it)))
        return st.value, it.value

    def fetch_solution(self):
        x = np.empty(self.nvoxel)
        check(lib().sart_fetch_solution(self._h, x.ctypes.data_as(C.c_void_p)))
        return x

    def forward_project(self, x):
        x = _f64(x)
        out = np.empty(self.nrows)
        check(lib().sart_forward_project(self._h, x.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p)))
        return out

    def kernel_forward(self, x):
        x = _f64(x)
        out = np.empty(self.nrows)
        check(lib().sart_kernel_forward(self._h, x.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p)))
        return out

    def kernel_backward(self, w):
        w = _f64(w)
        out = np.empty(self.nvoxel)
        check(lib().sart_kernel_backward(self._h, w.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p)))
        return out

    def stats(self):
        s = Stats()
        check(lib().sart_get_stats(self._h, C.byref(s)))
        return s
