"""CPU tests of the boundary: the C-ABI library loads, exports every symbol include/sart_b200.h declares, its
host-only entry points behave like the reference's, and without a GPU it fails loudly instead of falling back."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "sart_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(sart_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(sart):
    lib = sart.lib()
    declared = _declared_symbols()
    assert sorted(sart.SYMBOLS) == declared
    for name in declared:
        assert getattr(lib, name) is not None


def test_no_oracle_or_torch_dependency_in_the_product(sart):
    """The shared library must not link the oracle, torch or (at link time) NCCL."""
    import subprocess
    out = subprocess.run(["ldd", sart.LIB_PATH], capture_output=True, text=True).stdout
    for forbidden in ("oracle", "torch", "nccl", "cublas", "libcudart"):
        assert forbidden not in out, out
    src = open(os.path.join(ROOT, "mpi-cuda-sartsolver_b200", "__init__.py")).read()
    assert "oracle" not in src and "import torch" not in src


def test_config_defaults_are_the_reference_constructor_defaults(sart):
    cfg = sart.default_config()
    # source/include/sartsolver_cuda.hpp:31-35
    assert (cfg.ray_density_threshold, cfg.ray_length_threshold, cfg.conv_tolerance) == (1e-6, 1e-6, 1e-5)
    assert (cfg.beta_laplace, cfg.relaxation, cfg.max_iterations) == (1e-2, 1.0, 2000)
    assert (cfg.rank, cfg.nranks, cfg.device, cfg.logarithmic, cfg.semantics) == (0, 1, -1, 0, sart.SEMANTICS_CPU)
    assert C.sizeof(sart.Config) == 88 and C.sizeof(sart.Stats) == 160  # sizeof in C (gcc)


def test_shard_range_matches_reference_partition(sart, restatement):
    for npixel, n in ((10, 3), (8192, 2), (200000, 8), (1000003, 8), (5, 5)):
        for r in range(n):
            assert sart.shard_range(npixel, n, r) == restatement.shard_range(npixel, n, r)
    with pytest.raises(sart.SartError):
        sart.shard_range(10, 2, 2)


def test_status_constants_match_reference(sart, reference):
    assert sart.SUCCESS == reference.lib.ref_status_success()
    assert sart.MAX_ITERATIONS_EXCEEDED == reference.lib.ref_status_max_iterations_exceeded()


@pytest.mark.parametrize("field,bad", [
    ("ray_density_threshold", -1.0), ("ray_length_threshold", -1e-9), ("conv_tolerance", 0.0),
    ("beta_laplace", -0.1), ("relaxation", 0.0), ("relaxation", 1.5), ("max_iterations", 0),
])
def test_create_rejects_what_the_reference_setters_reject(sart, field, bad):
    # source/sartsolver.cpp:63-123: same ranges; an error code + message instead of exit(1)
    cfg = sart.default_config(npixel=8, nvoxel=8, **{field: bad})
    h = C.c_void_p()
    rc = sart.lib().sart_create(C.byref(cfg), C.byref(h))
    assert rc == sart.ERR_INVALID_ARGUMENT and not h.value
    assert len(sart.lib().sart_last_error()) > 10


def test_no_gpu_means_loud_failure_not_fallback(sart):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(sart.SartError) as ei:
        sart.Engine(64, 32)
    assert ei.value.code == sart.ERR_NO_DEVICE
    assert "no CPU path" in ei.value.message
