"""Last file pytest collects: the allocation-free back projection (sart_back_project_noalloc).

Its users are hosts that own every byte of GPU memory themselves (a fixed-budget acquisition loop, a framework's
caching allocator) and so cannot let this library call cudaMalloc; here torch plays that host and supplies the device
buffers. The result must agree with the engine's own back projection (sart_kernel_backward, whose partials sit in a
cudaMalloc'ed buffer) on two consecutive calls -- the second call finds the first call's partials still in the
library's built-in scratch array -- and again after a call with a different shape has overwritten part of it.
Because both entry points run the same kernel body, every result is also held against H^T w computed in fp64 by
numpy, the caller's buffers are laid out so that a kernel reading or writing outside its columns shows (non-finite
padding columns in the matrix, a guarded tail behind the output), and the amount of free device memory must not move
across the calls."""
import ctypes

import numpy as np
import pytest

from tests.cases import rel_l2

pytestmark = pytest.mark.gpu

TOL = 1e-5  # the tolerance of the parity tests
GUARD = 64  # doubles behind the output that no call may touch
SENTINEL = -12345.678


def _device_problem(torch, P, V, seed, extra_ld=0):
    """Matrix [P][ld] on the device whose padding columns (V .. ld-1) hold +inf / NaN: they may be read (the kernel works on
    whole 16-byte groups) but nothing of them may reach out[0 .. V-1]."""
    rng = np.random.default_rng(seed)
    H = rng.random((P, V), dtype=np.float32)
    w = rng.standard_normal(P).astype(np.float32)
    ld = (V + 31) // 32 * 32 + extra_ld
    Hd = torch.full((P, ld), float("nan"), dtype=torch.float32, device="cuda")
    Hd[:, V:V + 1] = float("inf")
    Hd[:, :V] = torch.from_numpy(H).cuda()
    return H, w, ld, Hd, torch.from_numpy(w).cuda()


def _free_device_bytes():
    """cudaMemGetInfo of the current device, asked of the runtime torch already loaded (no allocation of its own)."""
    import torch
    free, _total = torch.cuda.mem_get_info()
    return free


def test_noalloc_back_projection_matches_engine_twice_in_a_row(sart):
    import torch
    shapes = [(3001, 2500, 1, 0), (700, 9000, 2, 64)]  # the second one has a leading dimension wider than the padded row
    expected, exact, problems = [], [], []
    for P, V, seed, extra_ld in shapes:
        H, w, ld, Hd, wd = _device_problem(torch, P, V, seed, extra_ld)
        with sart.Engine(P, V, device=0) as eng:
            eng.upload_rows(H)
            expected.append(eng.kernel_backward(w.astype(np.float64)))  # the existing entry point
        exact.append(H.astype(np.float64).T @ w.astype(np.float64))     # and the definition, independent of both
        problems.append((P, V, ld, Hd, wd))
    stream = torch.cuda.current_stream().cuda_stream
    outs = [torch.empty(V + GUARD, dtype=torch.float64, device="cuda") for _, V, _, _ in shapes]
    torch.cuda.synchronize()

    def run(k):
        P, V, ld, Hd, wd = problems[k]
        out = outs[k]
        out[:V] = float("nan")
        out[V:] = SENTINEL
        sart.back_project_noalloc(Hd.data_ptr(), P, V, ld, wd.data_ptr(), out.data_ptr(), stream)
        torch.cuda.synchronize()
        res = out.cpu().numpy()
        assert np.all(res[V:] == SENTINEL), "wrote behind out[ncols - 1]"
        assert np.all(np.isfinite(res[:V])), "padding columns of the matrix leaked into the result"
        return res[:V]

    run(0)  # first call on this device: the library looks its scratch array up (may load the module)
    free_before = _free_device_bytes()
    first, second = run(0), run(0)
    for res in (first, second):
        assert rel_l2(res, expected[0]) <= TOL
        assert rel_l2(res, exact[0]) <= TOL
    np.testing.assert_array_equal(first, second)  # nothing left over from the first call leaks into the second
    third = run(1)                                # another geometry overwrites part of the shared array ...
    assert rel_l2(third, expected[1]) <= TOL and rel_l2(third, exact[1]) <= TOL
    again = run(0)                                # ... and the first problem still comes out right afterwards
    assert rel_l2(again, expected[0]) <= TOL and rel_l2(again, exact[0]) <= TOL
    assert _free_device_bytes() == free_before, "the allocation-free entry point changed the free device memory"


def test_noalloc_back_projection_refuses_what_does_not_fit(sart):
    import torch
    buf = torch.zeros(64, dtype=torch.float32, device="cuda")
    out = torch.zeros(64, dtype=torch.float64, device="cuda")
    assert sart.NOALLOC_MAX_LD * ctypes.sizeof(ctypes.c_float) == sart.NOALLOC_SCRATCH_BYTES  # one row of partials fits
    too_wide = sart.NOALLOC_MAX_LD + 32
    with pytest.raises(sart.SartError) as ei:
        sart.back_project_noalloc(buf.data_ptr(), 1, too_wide, too_wide, buf.data_ptr(), out.data_ptr(), None)
    assert ei.value.code == sart.ERR_INVALID_ARGUMENT and "exceeds the built-in scratch" in ei.value.message
    with pytest.raises(sart.SartError):  # ld not a multiple of 4 floats
        sart.back_project_noalloc(buf.data_ptr(), 2, 6, 6, buf.data_ptr(), out.data_ptr(), None)
    with pytest.raises(sart.SartError):  # null pointer
        sart.back_project_noalloc(None, 2, 8, 8, buf.data_ptr(), out.data_ptr(), None)
