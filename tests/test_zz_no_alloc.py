"""Last file pytest collects: the allocation-free back projection (sart_back_project_noalloc).

Its users are hosts that own every byte of GPU memory themselves (a fixed-budget acquisition loop, a framework's
caching allocator) and so cannot let this library call cudaMalloc; here torch plays that host and supplies the device
buffers. The result must agree with the engine's own back projection (sart_kernel_backward, whose partials sit in a
cudaMalloc'ed buffer) on two consecutive calls -- the second call finds the first call's partials still in the
library's built-in scratch array -- and again after a call with a different shape has overwritten part of it."""
import numpy as np
import pytest

from tests.cases import rel_l2

pytestmark = pytest.mark.gpu

TOL = 1e-5  # the tolerance of the parity tests


def _device_problem(torch, P, V, seed):
    rng = np.random.default_rng(seed)
    H = rng.random((P, V), dtype=np.float32)
    w = rng.standard_normal(P).astype(np.float32)
    ld = (V + 31) // 32 * 32
    Hd = torch.zeros((P, ld), dtype=torch.float32, device="cuda")
    Hd[:, :V] = torch.from_numpy(H).cuda()
    return H, w, ld, Hd, torch.from_numpy(w).cuda()


def test_noalloc_back_projection_matches_engine_twice_in_a_row(sart):
    import torch
    shapes = [(3001, 2500, 1), (700, 9000, 2)]
    expected, problems = [], []
    for P, V, seed in shapes:
        H, w, ld, Hd, wd = _device_problem(torch, P, V, seed)
        with sart.Engine(P, V, device=0) as eng:
            eng.upload_rows(H)
            expected.append(eng.kernel_backward(w.astype(np.float64)))
        problems.append((P, V, ld, Hd, wd))
    stream = torch.cuda.current_stream().cuda_stream

    def run(k):
        P, V, ld, Hd, wd = problems[k]
        out = torch.full((V,), float("nan"), dtype=torch.float64, device="cuda")
        sart.back_project_noalloc(Hd.data_ptr(), P, V, ld, wd.data_ptr(), out.data_ptr(), stream)
        torch.cuda.synchronize()
        return out.cpu().numpy()

    first, second = run(0), run(0)
    assert rel_l2(first, expected[0]) <= TOL
    assert rel_l2(second, expected[0]) <= TOL
    np.testing.assert_array_equal(first, second)  # nothing left over from the first call leaks into the second
    assert rel_l2(run(1), expected[1]) <= TOL      # another geometry overwrites part of the shared array ...
    assert rel_l2(run(0), expected[0]) <= TOL      # ... and the first problem still comes out right afterwards


def test_noalloc_back_projection_refuses_what_does_not_fit(sart):
    import torch
    buf = torch.zeros(64, dtype=torch.float32, device="cuda")
    out = torch.zeros(64, dtype=torch.float64, device="cuda")
    too_wide = sart.NOALLOC_MAX_LD + 32
    with pytest.raises(sart.SartError) as ei:
        sart.back_project_noalloc(buf.data_ptr(), 1, too_wide, too_wide, buf.data_ptr(), out.data_ptr(), None)
    assert ei.value.code == sart.ERR_INVALID_ARGUMENT and "exceeds the built-in scratch" in ei.value.message
    with pytest.raises(sart.SartError):  # ld not a multiple of 4 floats
        sart.back_project_noalloc(buf.data_ptr(), 2, 6, 6, buf.data_ptr(), out.data_ptr(), None)
    with pytest.raises(sart.SartError):  # null pointer
        sart.back_project_noalloc(None, 2, 8, 8, buf.data_ptr(), out.data_ptr(), None)
