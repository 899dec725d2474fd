"""GPU parity tests: the CUDA engine, called through the C ABI, against the oracle.

Tolerance (BASELINE.json north_star): relative L2 <= 1e-5 on the solution after a fixed iteration count, the same
status, and the same stopping iteration +-1 when the convergence test decides. The bound presumes matrices scaled
like physical ray-transfer matrices (row sums of order one), which the synthetic generator provides.
"""
import numpy as np
import pytest

from tests.cases import CASES, build_case, load_synthetic, rel_l2

pytestmark = pytest.mark.gpu

TOL = 1e-5
SPLIT, FUSED = 1, 2  # SART_SWEEP_SPLIT / SART_SWEEP_FUSED
BOTH_MODES = pytest.mark.parametrize("mode", [SPLIT, FUSED], ids=["split", "fused"])


def _engine_for(sart, case, H=None, **extra):
    H = case["H"] if H is None else H
    kw = dict(case["kw"])
    kw.pop("nranks", None)  # same arithmetic on one shard; the rank split only reorders fp64 additions
    kw.setdefault("beta_laplace", 2e-2)  # the golden runs used the CLI default (source/arguments.cpp:127)
    eng = sart.Engine(H.shape[0], H.shape[1], device=0, **kw, **extra)
    eng.upload_rows(H)
    if case["lap"] is not None:
        eng.set_laplacian(*case["lap"])
    eng.finalize()
    return eng


def _run_frames(eng, case):
    xs, sts, its = [], [], []
    guess = case["x0"]
    for f in range(case["frames"].shape[0]):
        x, st, it = eng.solve(case["frames"][f], guess)
        xs.append(x)
        sts.append(st)
        its.append(it)
        guess = None if case["no_guess"] else x  # source/main.cpp:139
    return np.stack(xs), np.array(sts), np.array(its)


@BOTH_MODES
@pytest.mark.parametrize("name", list(CASES))
def test_engine_matches_golden(name, mode, sart, restatement, golden):
    case = build_case(name, restatement)
    fixed = case["kw"]["conv_tolerance"] < 1e-20
    with _engine_for(sart, case, sweep_mode=mode) as eng:
        x, st, it = _run_frames(eng, case)
    gx, gst, git = golden[name + "/x"], golden[name + "/status"], golden[name + "/iterations"]
    if fixed:
        np.testing.assert_array_equal(st, gst)
        np.testing.assert_array_equal(it, git)
        for f in range(x.shape[0]):
            assert rel_l2(x[f], gx[f]) <= TOL
    else:
        # the convergence test decides: stopping iteration within one of the reference's
        assert np.all(np.abs(it - git) <= 1), (it, git)
        if np.array_equal(it, git):
            np.testing.assert_array_equal(st, gst)
            for f in range(x.shape[0]):
                assert rel_l2(x[f], gx[f]) <= TOL


def test_generator_is_bit_identical_to_the_cpu_generator(sart, restatement):
    P, V = 300, 1000  # V not a multiple of 32: exercises the padded leading dimension
    ref = restatement.fill_synthetic(P, V, 42)
    with sart.Engine(P, V, device=0) as eng:
        eng.fill_synthetic(42)
        np.testing.assert_array_equal(eng.download_rows(), ref)
    # a shard of a 3-rank partition holds exactly its rows of the same global matrix
    for rank in range(3):
        with sart.Engine(P, V, rank=rank, nranks=3, device=0) as eng:
            eng.fill_synthetic(42)
            off, n = sart.shard_range(P, 3, rank)
            np.testing.assert_array_equal(eng.download_rows(), ref[off:off + n])


def test_line_of_sight_generator_is_bit_identical_to_the_cpu_generator(sart, restatement):
    P, V = 257, 5000  # 50 x 100 voxel grid; V not a multiple of 32
    ref = restatement.fill_synthetic(P, V, 42, kind=1)
    with sart.Engine(P, V, device=0) as eng:
        eng.fill_synthetic(42, sart.SYNTHETIC_LOS)
        np.testing.assert_array_equal(eng.download_rows(), ref)
    off, n = sart.shard_range(P, 3, 2)
    with sart.Engine(P, V, rank=2, nranks=3, device=0) as eng:
        eng.fill_synthetic(42, sart.SYNTHETIC_LOS)
        np.testing.assert_array_equal(eng.download_rows(), ref[off:off + n])
        with pytest.raises(sart.SartError):
            eng.fill_synthetic(42, 7)


def test_upload_subblocks_roundtrip(sart):
    rng = np.random.default_rng(0)
    P, V = 97, 203
    H = rng.random((P, V), dtype=np.float32)
    with sart.Engine(P, V, device=0) as eng:
        # per "camera" row blocks x per "voxel segment" column blocks, like the reference's files
        for r0, r1 in ((0, 40), (40, 97)):
            for c0, c1 in ((0, 100), (100, 203)):
                eng.upload_rows(H[r0:r1, c0:c1], local_row0=r0, col0=c0)
        np.testing.assert_array_equal(eng.download_rows(), H)
        with pytest.raises(sart.SartError):
            eng.upload_rows(H[:, :10], local_row0=1)


def test_ray_sums_and_forward_projection_fp64(sart, restatement):
    P, V = 777, 333
    H = restatement.fill_synthetic(P, V, 5)
    H[::50] = 0
    rho_ref, lam_ref = restatement.ray_sums(H)
    x = load_synthetic().emission_profile(V)
    with sart.Engine(P, V, device=0) as eng:
        eng.upload_rows(H)
        eng.finalize()
        rho, lam = eng.ray_sums()
        fwd = eng.forward_project(x)
    assert rel_l2(rho, rho_ref) < 1e-13 and rel_l2(lam, lam_ref) < 1e-13
    assert np.all(lam[::50] == 0)
    assert rel_l2(fwd, restatement.forward(H, x)) < 1e-13


@BOTH_MODES
@pytest.mark.parametrize("P,V,iters", [(4096, 2048, 20), (2050, 4100, 12), (33, 7, 15), (1, 40, 5), (700, 20000, 8)])
def test_fixed_iterations_parity_across_shapes(P, V, iters, mode, sart, restatement):
    syn = load_synthetic()
    H = restatement.fill_synthetic(P, V, 99)
    g = restatement.forward(H, syn.emission_profile(V))
    lap = syn.laplacian_coo(V)
    x_ref, st_ref, it_ref = restatement.solve(H, g, lap=lap, max_iterations=iters, fixed_iterations=True)
    with sart.Engine(P, V, device=0, beta_laplace=2e-2, max_iterations=iters, fixed_iterations=True,
                     sweep_mode=mode) as eng:
        eng.upload_rows(H)
        eng.set_laplacian(*lap)
        eng.finalize()
        x, st, it = eng.solve(g)
        s = eng.stats()
    assert (st, it) == (st_ref, it_ref) == (sart.MAX_ITERATIONS_EXCEEDED, iters)
    assert rel_l2(x, x_ref) <= TOL
    assert s.sweep_mode == mode and s.sweeps == 2 * iters + 2
    if mode == FUSED and s.solve_launches:
        # the whole solve ran as one cooperative launch: sweeps 0 .. iters inside it
        assert s.solve_launches == 1 and s.fused_launches == iters + 1 and s.bp_launches == 0 and s.fp_launches == 0
    elif mode == FUSED:
        assert s.fused_launches == min(iters, 64) and s.bp_launches == 0 and s.fp_launches == 1
    else:
        assert s.gpu_launches >= 4 * iters


@pytest.mark.parametrize("P,V,log,tol", [(3000, 4100, 0, 1e-5), (700, 20000, 0, 1e-30), (1500, 9000, 1, 1e-5), (64, 5000, 0, 1e-4),
                                         (1536, 640, 0, 1e-30), (5000, 2000, 0, 1e-5)])
def test_whole_solve_kernel_is_bit_identical_to_one_launch_per_sweep(P, V, log, tol, sart, restatement, monkeypatch):
    """fused_solve_kernel (all sweeps, cluster reduction, decision and update inside one cooperative launch) repeats the
    arithmetic of the per-sweep kernels in the same order: same bits, same status, same stopping iteration."""
    syn = load_synthetic()
    H = restatement.fill_synthetic(P, V, 123)
    g = restatement.forward(H, syn.emission_profile(V))
    if not log:
        g[::17] = -1.0  # saturated pixels
    lap = syn.laplacian_coo(V)
    out = {}
    for persistent in ("1", "0"):
        monkeypatch.setenv("SART_PERSISTENT", persistent)
        with sart.Engine(P, V, device=0, beta_laplace=2e-2, max_iterations=40, conv_tolerance=tol, logarithmic=log,
                         sweep_mode=FUSED) as eng:
            eng.upload_rows(H)
            eng.set_laplacian(*lap)
            eng.finalize()
            cold = eng.solve(g)
            warm = eng.solve(1.5 * g, cold[0])
            out[persistent] = (cold, warm, eng.stats().solve_launches)
    assert out["1"][2] == 1 and out["0"][2] == 0
    for a, b in zip(out["1"][:2], out["0"][:2]):
        assert a[1:] == b[1:]
        assert np.array_equal(a[0], b[0])


def test_gpu_semantics_masked_initial_guess(sart, restatement):
    """The reference GPU path clips saturated pixels out of the initial guess and floors the start at 1e-7 * max(g)
    (sart_kernels.cu:33-34, sartsolver_cuda.cpp:180); the default follows the CPU path, which does neither."""
    syn = load_synthetic()
    P, V = 512, 200
    H = restatement.fill_synthetic(P, V, 17)
    g = restatement.forward(H, syn.emission_profile(V))
    g[::9] = -1.0
    for gpu_sem in (False, True):
        x_ref = restatement.solve(H, g, max_iterations=10, fixed_iterations=True, gpu_semantics=gpu_sem)[0]
        with sart.Engine(P, V, device=0, max_iterations=10, fixed_iterations=True,
                         semantics=sart.SEMANTICS_GPU if gpu_sem else sart.SEMANTICS_CPU) as eng:
            eng.upload_rows(H)
            eng.finalize()
            x = eng.solve(g)[0]
        assert rel_l2(x, x_ref) <= TOL


def test_resident_path_and_setters(sart, restatement):
    syn = load_synthetic()
    P, V = 640, 256
    H = restatement.fill_synthetic(P, V, 3)
    g = restatement.forward(H, syn.emission_profile(V))
    with sart.Engine(P, V, device=0, max_iterations=8, fixed_iterations=True) as eng:
        eng.upload_rows(H)
        eng.finalize()
        x1, st1, it1 = eng.solve(g)
        eng.stage_measurement(g)
        st2, it2 = eng.solve_resident(False)
        x2 = eng.fetch_solution()
        np.testing.assert_array_equal(x1, x2)  # same launches, no atomics: bit-reproducible
        assert (st1, it1) == (st2, it2)
        # warm start from the solution left on the device == warm start from the host copy
        st3, it3 = eng.solve_resident(True)
        x3 = eng.fetch_solution()
        x4 = eng.solve(g, x1)[0]
        np.testing.assert_array_equal(x3, x4)
        # setters validate like the reference and take effect
        with pytest.raises(sart.SartError):
            eng.set_relaxation(0.0)
        eng.set_relaxation(0.5)
        eng.set_max_iterations(5)
        x5 = eng.solve(g)[0]
        x5_ref = restatement.solve(H, g, max_iterations=5, fixed_iterations=True, relaxation=0.5)[0]
        assert rel_l2(x5, x5_ref) <= TOL
        with pytest.raises(ValueError):
            eng.solve(g, np.zeros(V + 1))


@BOTH_MODES
def test_convergence_decision_on_device_matches_oracle(mode, sart, restatement):
    syn = load_synthetic()
    P, V = 1024, 512
    H = restatement.fill_synthetic(P, V, 8)
    g = restatement.forward(H, syn.emission_profile(V))
    x_ref, st_ref, it_ref, hist = restatement.solve(H, g, max_iterations=500, conv_tolerance=1e-4, history=True)
    assert st_ref == 0 and 3 < it_ref < 500
    with sart.Engine(P, V, device=0, max_iterations=500, conv_tolerance=1e-4, sweep_mode=mode) as eng:
        eng.upload_rows(H)
        eng.finalize()
        x, st, it = eng.solve(g)
    assert st == 0 and abs(it - it_ref) <= 1
    if it == it_ref:
        assert rel_l2(x, x_ref) <= TOL
    else:
        assert rel_l2(x, restatement.solve(H, g, max_iterations=it, fixed_iterations=True)[0]) <= TOL


@pytest.mark.parametrize("mode,P,V", [(SPLIT, 512, 50000), (FUSED, 512, 50000), (FUSED, 300, 100000), (FUSED, 150, 150000)],
                         ids=["split-50k", "fused-50k-cluster4", "fused-100k-cluster8", "fused-150k-cluster16"])
def test_same_voxel_count_as_the_bench_workload_on_a_row_subsample(mode, P, V, sart, restatement):
    """BASELINE config 2 has 50 000 voxels; the oracle cannot hold its 200 000 rows, but the generator is keyed by
    global (row, column), so the first rows of that very matrix are a valid reduced problem with the same row
    length, panel count and padding."""
    syn = load_synthetic()
    iters = 6
    H = restatement.fill_synthetic(P, V, 42)
    g = restatement.forward(H, syn.emission_profile(V))
    lap = syn.laplacian_coo(V)
    x_ref = restatement.solve(H, g, lap=lap, max_iterations=iters, fixed_iterations=True)[0]
    with sart.Engine(P, V, device=0, beta_laplace=2e-2, max_iterations=iters, fixed_iterations=True,
                     sweep_mode=mode) as eng:
        eng.fill_synthetic(42)
        eng.set_laplacian(*lap)
        eng.finalize()
        x = eng.solve(g)[0]
        assert eng.stats().sweep_mode == mode
    assert rel_l2(x, x_ref) <= TOL


@BOTH_MODES
def test_full_size_properties_config2(mode, sart):
    """BASELINE config 2 at full size (200 000 x 50 000 fp32 = 40 GB + Laplacian): properties that need no oracle.
    (1) bit-reproducible; (2) scale covariance solve(c g) = c solve(g); (3) a consistent frame g = H x* started at
    x* stays there (residual weights vanish), without the Laplacian."""
    info = sart.device_query(0)
    P, V = 200000, 50000
    if info.free_mem < P * V * 4 + (6 << 30):
        pytest.skip("needs ~46 GB of free device memory")
    syn = load_synthetic()
    xs = syn.emission_profile(V)
    with sart.Engine(P, V, device=0, beta_laplace=2e-2, max_iterations=6, fixed_iterations=True,
                     sweep_mode=mode) as eng:
        eng.fill_synthetic(42)
        eng.set_laplacian(*syn.laplacian_coo(V))
        eng.finalize()
        rho, lam = eng.ray_sums()
        assert 0.5 < lam.mean() < 1.5 and rho.min() > 0
        g = eng.forward_project(xs)
        x1, st, it = eng.solve(g)
        x2 = eng.solve(g)[0]
        np.testing.assert_array_equal(x1, x2)
        assert (st, it) == (sart.MAX_ITERATIONS_EXCEEDED, 6) and np.all(x1 >= 0) and np.all(np.isfinite(x1))
        x3 = eng.solve(8.0 * g)[0]  # power of two: the normalised problem is bit-identical
        np.testing.assert_array_equal(x3, 8.0 * x1)
        x4 = eng.solve(3.7 * g)[0]
        assert rel_l2(x4, 3.7 * x1) <= TOL
        # without the penalty the iteration keeps reducing the data misfit
        eng.set_beta_laplace(0.0)
        r0 = np.linalg.norm(eng.forward_project(eng.solve(g)[0]) - g)
        eng.set_max_iterations(12)
        r1 = np.linalg.norm(eng.forward_project(eng.solve(g)[0]) - g)
        assert r1 < r0
        eng.set_max_iterations(3)
        x5 = eng.solve(g, xs)[0]
        assert rel_l2(x5, xs) <= TOL


@pytest.mark.parametrize("P,V", [(1000, 1024), (257, 1000), (4099, 77), (31, 33), (64, 50000)])
def test_sweep_kernels_known_answers(P, V, sart):
    """The two production kernels alone against fp64 numpy on odd shapes (rows not a multiple of the row group or of
    the 32-row slab step, columns not a multiple of 4 / 32 / the 1024-column panel)."""
    rng = np.random.default_rng(P * 131 + V)
    H = rng.random((P, V), dtype=np.float32)
    H[rng.random(P) < 0.05] = 0
    x = rng.random(V)
    w = rng.standard_normal(P)
    with sart.Engine(P, V, device=0) as eng:
        eng.upload_rows(H)
        fit = eng.kernel_forward(x)
        s = eng.kernel_backward(w)
    H64 = H.astype(np.float64)
    assert rel_l2(fit, H64 @ x.astype(np.float32).astype(np.float64)) < 2e-6
    assert rel_l2(s, H64.T @ w.astype(np.float32).astype(np.float64)) < 2e-6


def test_fused_and_split_sweeps_agree_at_full_size(sart):
    """The two ways of passing over the matrix are the same arithmetic in a different summation order: on the full
    BASELINE config-2 matrix their solutions must agree far inside the parity tolerance."""
    info = sart.device_query(0)
    P, V = 200000, 50000
    if info.free_mem < P * V * 4 + (6 << 30):
        pytest.skip("needs ~46 GB of free device memory")
    syn = load_synthetic()
    xs = syn.emission_profile(V)
    sols = []
    for mode in (SPLIT, FUSED):
        with sart.Engine(P, V, device=0, beta_laplace=2e-2, max_iterations=10, fixed_iterations=True,
                         sweep_mode=mode) as eng:
            eng.fill_synthetic(42)
            eng.set_laplacian(*syn.laplacian_coo(V))
            eng.finalize()
            g = eng.forward_project(xs)
            sols.append(eng.solve(g)[0])
    assert rel_l2(sols[1], sols[0]) <= 1e-6


@BOTH_MODES
def test_baseline_config1_full(mode, sart, restatement):
    """BASELINE configs[0] at full size: 8192 x 4096, single frame, 100 iterations, default initial guess, against the
    oracle after exactly 100 iterations."""
    syn = load_synthetic()
    P, V, iters = 8192, 4096, 100
    H = restatement.fill_synthetic(P, V, 42)
    g = restatement.forward(H, syn.emission_profile(V))
    x_ref, st_ref, it_ref = restatement.solve(H, g, max_iterations=iters, fixed_iterations=True)
    with sart.Engine(P, V, device=0, max_iterations=iters, fixed_iterations=True, sweep_mode=mode) as eng:
        eng.fill_synthetic(42)
        eng.finalize()
        x, st, it = eng.solve(g)
    assert (st, it) == (st_ref, it_ref)
    assert rel_l2(x, x_ref) <= TOL


def test_engine_matches_the_reference_gpu_solver(sart, restatement):
    """Second oracle: the reference's own CUDA solver (sartsolver_cuda.cpp + sart_kernels.cu + cuBLAS), compiled
    unmodified for sm_100a into oracle/_ref/libsart_ref_gpu.so. Its stopping iteration under a vanishing tolerance is
    run-dependent (fp32 sum(fit^2), atomics), so the engine is run for exactly the count the reference reached."""
    import os
    from oracle.pyoracle import REF_GPU_SO, Reference
    if not os.path.exists(REF_GPU_SO):
        pytest.skip("oracle/_ref/libsart_ref_gpu.so not built")
    syn = load_synthetic()
    P, V = 3000, 1500
    H = restatement.fill_synthetic(P, V, 77)
    g = restatement.forward(H, syn.emission_profile(V))
    g[::11] = -1.0  # saturated pixels: the GPU path clips them out of the initial guess
    lap = syn.laplacian_coo(V)
    ref = Reference(REF_GPU_SO)
    x_ref, st_ref, it_ref = ref.solve(H, g, lap=lap, beta_laplace=2e-2, max_iterations=8, conv_tolerance=1e-300)
    assert 1 <= it_ref <= 8
    for mode in (SPLIT, FUSED):
        with sart.Engine(P, V, device=0, beta_laplace=2e-2, max_iterations=it_ref, fixed_iterations=True,
                         semantics=sart.SEMANTICS_GPU, sweep_mode=mode) as eng:
            eng.upload_rows(H)
            eng.set_laplacian(*lap)
            eng.finalize()
            x = eng.solve(g)[0]
        assert rel_l2(x, x_ref) <= TOL


def _rowsum_laplacian(V, add=0.5):
    idx, val = load_synthetic().laplacian_coo(V)
    val = val.copy()
    val[(idx // np.uint64(V)) == (idx % np.uint64(V))] += np.float32(add)
    return idx, val


@BOTH_MODES
@pytest.mark.parametrize("gpu_sem", [False, True], ids=["cpu-semantics", "gpu-semantics"])
def test_log_penalty_units_with_nonzero_laplacian_row_sums(gpu_sem, mode, sart, restatement):
    """The log solver's penalty beta * L * log(x) is not scale-invariant when a Laplacian row does not sum to zero:
    the CPU solver takes log() of the solution in the caller's units (sartsolver.cpp:292), the GPU solver of the
    iterate divided by max(g) (sart_kernels.cu:200). A frame 37.5 times brighter than max(g) = 1 tells them apart."""
    syn = load_synthetic()
    P, V = 400, 180
    H = restatement.fill_synthetic(P, V, 31)
    g = 37.5 * restatement.forward(H, syn.emission_profile(V))
    lap = _rowsum_laplacian(V)
    kw = dict(beta_laplace=2e-2, max_iterations=15, fixed_iterations=True, logarithmic=True, relaxation=0.8)
    x_ref = restatement.solve(H, g, lap=lap, gpu_semantics=gpu_sem, **kw)[0]
    x_other = restatement.solve(H, g, lap=lap, gpu_semantics=not gpu_sem, **kw)[0]
    assert rel_l2(x_other, x_ref) > 1e-2  # the two conventions really differ on this input
    with sart.Engine(P, V, device=0, sweep_mode=mode, semantics=sart.SEMANTICS_GPU if gpu_sem else sart.SEMANTICS_CPU,
                     **kw) as eng:
        eng.upload_rows(H)
        eng.set_laplacian(*lap)
        eng.finalize()
        x = eng.solve(g)[0]
    assert rel_l2(x, x_ref) <= TOL


def test_log_solver_matches_the_reference_gpu_solver_on_nonzero_row_sums(sart, restatement):
    """Pins the GPU-semantics branch of the restatement (and the engine) to the reference's own CUDA log solver where
    the units of the penalty matter. That solver uses __logf / __powf / __expf (about 1e-6 per step), hence the wider
    bound; the unit convention is a 1e-2 effect on this input."""
    import os
    from oracle.pyoracle import REF_GPU_SO, Reference
    if not os.path.exists(REF_GPU_SO):
        pytest.skip("oracle/_ref/libsart_ref_gpu.so not built")
    syn = load_synthetic()
    P, V = 2000, 900
    H = restatement.fill_synthetic(P, V, 78)
    g = 37.5 * restatement.forward(H, syn.emission_profile(V))
    lap = _rowsum_laplacian(V)
    ref = Reference(REF_GPU_SO)
    x_ref, st_ref, it_ref = ref.solve(H, g, lap=lap, beta_laplace=2e-2, max_iterations=6, conv_tolerance=1e-300,
                                      logarithmic=True, relaxation=0.8)
    assert 1 <= it_ref <= 6
    kw = dict(beta_laplace=2e-2, max_iterations=it_ref, fixed_iterations=True, logarithmic=True, relaxation=0.8)
    x_res = restatement.solve(H, g, lap=lap, gpu_semantics=True, **kw)[0]
    assert rel_l2(x_res, x_ref) <= 1e-4
    with sart.Engine(P, V, device=0, semantics=sart.SEMANTICS_GPU, **kw) as eng:
        eng.upload_rows(H)
        eng.set_laplacian(*lap)
        eng.finalize()
        x = eng.solve(g)[0]
    assert rel_l2(x, x_ref) <= 1e-4


def test_file_loader_streams_blocks_through_pinned_staging(sart, tmp_path, monkeypatch):
    """Row blocks x column segments from raw files (what per-camera x per-segment files are) land exactly where
    sart_upload_rows would put them; a 1 MB staging size forces many slabs and the ring to wrap."""
    monkeypatch.setenv("SART_STAGING_MB", "1")
    rng = np.random.default_rng(5)
    P, V = 1500, 777
    H = rng.random((P, V), dtype=np.float32)
    header = 24
    seg = [(0, 300), (300, 777)]           # two voxel segments
    cams = [(0, 900), (900, 1500)]         # two cameras
    with sart.Engine(P, V, device=0) as eng:
        for ci, (r0, r1) in enumerate(cams):
            for si, (c0, c1) in enumerate(seg):
                path = tmp_path / f"cam{ci}_seg{si}.bin"
                with open(path, "wb") as f:
                    f.write(b"\0" * header)
                    f.write(np.ascontiguousarray(H[r0:r1, c0:c1]).tobytes())
                eng.load_rows_file(path, header, c1 - c0, r0, r1 - r0, c0, c1 - c0)
        np.testing.assert_array_equal(eng.download_rows(), H)
        # a block that is a column window of wider file rows (file_ld > ncols)
        whole = tmp_path / "whole.bin"
        H2 = rng.random((P, V), dtype=np.float32)
        whole.write_bytes(H2.tobytes())
        eng.load_rows_file(whole, 0, V, 0, P, 0, V)
        eng.load_rows_file(whole, 100 * 4, V, 10, 50, 100, 200)  # rows 0..49 of the file, columns 100..299 -> rows 10..59
        expect = H2.copy()
        expect[10:60, 100:300] = H2[0:50, 100:300]
        np.testing.assert_array_equal(eng.download_rows(), expect)
        with pytest.raises(sart.SartError):
            eng.load_rows_file(tmp_path / "missing.bin", 0, V, 0, P, 0, V)
        with pytest.raises(sart.SartError):
            eng.load_rows_file(whole, 8, V, 0, P, 0, V)  # would run past the end of the file


def test_call_order_and_argument_errors_are_reported_not_fatal(sart):
    """Every misuse comes back as an error code + message (the reference would print and exit(1))."""
    import ctypes as C
    with sart.Engine(64, 32, device=0) as eng:
        with pytest.raises(sart.SartError) as ei:
            eng.solve(np.ones(64))  # before sart_finalize
        assert ei.value.code == sart.ERR_STATE
        with pytest.raises(sart.SartError):
            eng.upload_rows(np.zeros((65, 32), np.float32))  # more rows than the shard has
        eng.set_laplacian(np.array([32 * 32 + 1], np.uint64), np.array([1.0], np.float32))  # row index 32 of 32 voxels
        with pytest.raises(sart.SartError) as ei:
            eng.finalize()
        assert ei.value.code == sart.ERR_INVALID_ARGUMENT and "out of range" in ei.value.message
        eng.set_laplacian(np.zeros(0, np.uint64), np.zeros(0, np.float32))
        eng.upload_rows(np.ones((64, 32), np.float32))
        eng.finalize()
        with pytest.raises(sart.SartError):
            eng.solve_resident(False)  # nothing staged
        with pytest.raises(ValueError):
            eng.solve(np.ones(63))
        x, st, it = eng.solve(np.zeros(64))  # an all-dark frame: m2 = 0, the metric is NaN, runs to max_iterations
        assert st == sart.MAX_ITERATIONS_EXCEEDED and it == 2000 and np.all(x == 0)
    with pytest.raises(sart.SartError):
        sart.Engine(64, 32, device=99)
    with pytest.raises(sart.SartError) as ei:
        sart.Engine(64, 300000, device=0, sweep_mode=FUSED)  # wider than 16 CTAs can hold
    assert "not available" in ei.value.message


@pytest.mark.parametrize("seed", range(24))
def test_randomised_problems_against_the_oracle(seed, sart, restatement):
    """Shapes, thresholds, masks, relaxation, regularisation, solver kind, sweep kind, semantics and warm start drawn at
    random (reproducibly): the engine must follow the oracle on every combination."""
    rng = np.random.default_rng(1000 + seed)
    syn = load_synthetic()
    P = int(rng.integers(1, 700))
    V = int(rng.choice([1, 3, 4, 5, 31, 32, 33, 100, 255, 256, 257, 1000, 1025, 2500]))
    H = restatement.fill_synthetic(P, V, int(rng.integers(1, 10 ** 6)))
    if rng.random() < 0.5 and P > 4:
        H[rng.random(P) < 0.1] = 0                       # rays that see nothing
    if rng.random() < 0.5 and V > 4:
        H[:, rng.random(V) < 0.1] = 0                    # voxels nobody sees
    if rng.random() < 0.3:
        H[rng.random(P) < 0.2] *= 1e-7                   # rows below the ray-length threshold
    x_true = syn.emission_profile(V) * float(rng.choice([1e-3, 1.0, 1e4]))
    g = restatement.forward(H, x_true)
    saturate = rng.random() < 0.5
    kw = dict(
        ray_density_threshold=float(rng.choice([1e-6, 1e-3])), ray_length_threshold=float(rng.choice([1e-6, 1e-4])),
        relaxation=float(rng.choice([1.0, 0.7, 0.25])), beta_laplace=float(rng.choice([0.0, 1e-2, 5e-2])),
        max_iterations=int(rng.integers(1, 25)), logarithmic=bool(rng.random() < 0.35))
    gpu_sem = bool(rng.random() < 0.3)
    lap = syn.laplacian_coo(V) if rng.random() < 0.6 else None
    x0 = None
    if rng.random() < 0.4:
        x0 = x_true * rng.uniform(0.5, 1.5, V) + 1e-3 * x_true.max()
    # Known limit (precision): the CPU solver's default initial guess adds the -1 markers of saturated pixels
    # into the guess; when that drives it negative the log solver restarts from 1e-100, a value (and a forward projection
    # of it) far below fp32 range, which no fp32 engine -- the reference GPU path included -- can follow.
    if saturate and not (kw["logarithmic"] and x0 is None and not gpu_sem):
        g[rng.random(P) < 0.15] = -1.0                   # saturated pixels
    mode = int(rng.choice([SPLIT, FUSED]))
    x_ref, st_ref, it_ref = restatement.solve(H, g, x0=x0, lap=lap, fixed_iterations=True, gpu_semantics=gpu_sem, **kw)
    with sart.Engine(P, V, device=0, fixed_iterations=True, sweep_mode=mode,
                     semantics=sart.SEMANTICS_GPU if gpu_sem else sart.SEMANTICS_CPU, **kw) as eng:
        eng.upload_rows(H)
        if lap is not None:
            eng.set_laplacian(*lap)
        eng.finalize()
        x, st, it = eng.solve(g, x0)
    assert (st, it) == (st_ref, it_ref)
    if np.all(np.isfinite(x_ref)) and np.linalg.norm(x_ref) > 0:
        assert rel_l2(x, x_ref) <= TOL, dict(P=P, V=V, mode=mode, gpu_sem=gpu_sem, **kw)
    else:
        np.testing.assert_array_equal(np.isfinite(x), np.isfinite(x_ref))


@pytest.mark.parametrize("P,V,tol,log", [(700, 1000, None, False), (512, 50000, None, False), (300, 20000, 1e-4, False),
                                          (257, 77, 1e-4, False), (700, 1000, None, True), (300, 5000, 1e-4, True)])
def test_batch_of_independent_frames_matches_frame_by_frame(P, V, tol, log, sart, restatement):
    """sart_solve_batch (eight frames per matrix pass) against solving every frame on its own and against the oracle: eleven
    frames, so one full group and one of three; frames of a group converge at different iterations when the tolerance
    decides. Both solvers: the log solver's back-projected measurement of every frame is one more batched pass."""
    syn = load_synthetic()
    H = restatement.fill_synthetic(P, V, 321)
    times = (0.0, 0.3, 0.55, 0.8, 1.0, 0.15, 0.65, 0.9, 0.45, 0.05, 0.7)
    frames = np.stack([restatement.forward(H, syn.emission_profile(V, t=t) * (1.0 + 3.0 * t)) for t in times])
    frames[1, ::19] = -1.0
    lap = syn.laplacian_coo(V)
    kw = dict(beta_laplace=2e-2, max_iterations=9, fixed_iterations=True) if tol is None else \
        dict(beta_laplace=2e-2, max_iterations=300, conv_tolerance=tol)
    if log:
        kw.update(logarithmic=True, relaxation=0.8)
    with sart.Engine(P, V, device=0, **kw) as eng:
        eng.upload_rows(H)
        eng.set_laplacian(*lap)
        eng.finalize()
        xb, stb, itb = eng.solve_batch(frames)
        singles = [eng.solve(f) for f in frames]
    for f in range(len(times)):
        xs, sts, its = singles[f]
        assert abs(int(itb[f]) - its) <= 1
        if itb[f] == its:
            assert stb[f] == sts and rel_l2(xb[f], xs) <= 1e-6
        okw = dict(kw)
        okw.pop("fixed_iterations", None)
        x_ref, st_ref, it_ref = restatement.solve(H, frames[f], lap=lap, fixed_iterations=tol is None, **okw)
        assert abs(int(itb[f]) - it_ref) <= 1
        if itb[f] == it_ref:
            assert stb[f] == st_ref and rel_l2(xb[f], x_ref) <= TOL
    if tol is not None:
        assert len(set(itb.tolist())) > 1  # the frames really stopped at different iterations


@pytest.mark.parametrize("P,V,tol,kind", [(3000, 4100, 1e-4, "steady"), (900, 20000, 1e-4, "steady"), (2048, 640, 1e-5, "drift")])
def test_warm_started_series_with_lookahead_matches_frame_by_frame(P, V, tol, kind, sart, restatement, monkeypatch):
    """sart_solve_series against the same frames solved one by one with a warm start (same engine, look-ahead off) and
    against the oracle's sequence: same status, stopping iteration within one, solutions within the parity tolerance --
    while the look-ahead saves a pass over the matrix for every frame that stops where its predecessor did.
    "steady": nearly identical frames (a few iterations each once warm) with one jump that breaks the prediction;
    "drift": a moving profile whose frames need many, changing iteration counts (predictions rarely apply)."""
    syn = load_synthetic()
    H = restatement.fill_synthetic(P, V, 77)
    F = 10
    if kind == "steady":
        g0 = restatement.forward(H, syn.emission_profile(V))
        frames = np.stack([g0 * (1.0 + 1e-3 * k) for k in range(F)])
    else:
        frames = np.stack([restatement.forward(H, syn.emission_profile(V, t=0.02 * k)) * (1.0 + 0.1 * k) for k in range(F)])
    frames[:, ::23] = -1.0  # saturated pixels in every frame
    frames[6] *= 3.0       # a jump: this frame needs more iterations than its predecessor
    lap = syn.laplacian_coo(V)
    kw = dict(beta_laplace=2e-2, max_iterations=200, conv_tolerance=tol, sweep_mode=FUSED)
    out = {}
    for look in ("1", "0"):
        monkeypatch.setenv("SART_SERIES_LOOKAHEAD", look)
        with sart.Engine(P, V, device=0, **kw) as eng:
            eng.upload_rows(H)
            eng.set_laplacian(*lap)
            eng.finalize()
            xs, sts, its = eng.solve_series(frames)
            s = eng.stats()
            out[look] = (xs, sts, its, s.sweeps, s.fp_launches, s.bp_launches)
            if look == "0":
                x1 = [eng.solve(frames[0])]
                for k in range(1, F):
                    x1.append(eng.solve(frames[k], x1[-1][0]))
    xs, sts, its, sweeps, handed, repeated = out["1"]
    xs0, sts0, its0, sweeps0, handed0, repeated0 = out["0"]
    assert handed0 == 0 and repeated0 == 0
    for k in range(F):  # look-ahead off == plain warm-started solves, bit for bit
        assert np.array_equal(xs0[k], x1[k][0]) and (sts0[k], its0[k]) == x1[k][1:]
    x_prev = None
    for k in range(F):
        assert sts[k] == sts0[k] and abs(int(its[k]) - int(its0[k])) <= 1
        if its[k] == its0[k]:
            assert rel_l2(xs[k], xs0[k]) <= TOL
        okw = dict(lap=lap, beta_laplace=2e-2, max_iterations=200, conv_tolerance=tol)
        x_ref, st_ref, it_ref = restatement.solve(H, frames[k], x0=x_prev, **okw)
        x_prev = x_ref
        assert sts[k] == st_ref and abs(int(its[k]) - it_ref) <= 1
        if its[k] == it_ref:
            assert rel_l2(xs[k], x_ref) <= TOL
    # passes over the matrix: initial guess + one per iterate of every frame, minus the handed-over first iterations
    assert sweeps == 1 + int(np.sum(its + 1)) - handed + repeated
    if kind == "steady":
        assert handed >= 2 and repeated >= 1, (its.tolist(), handed, repeated)
    assert sweeps0 == 1 + int(np.sum(its0 + 1))


@pytest.mark.parametrize("P,V", [(2048, 4096), (600, 20000)], ids=["cluster1", "cluster2"])
def test_precision_envelope_against_mean_ray_length(P, V, sart, restatement, capsys):
    """Where the 1e-5 parity bound stops holding. The reference's default initial guess (Eq. 4) is not normalised by the ray
    length, so on a matrix whose rows sum to lambda it overshoots by about lambda and the first update cancels it: any
    engine that keeps the iterate in fp32 loses ~lambda * 2^-24 there. The matrix is scaled by powers
    of two (exact in every product), 30 iterations from the default guess; both sweep kinds are held to the same envelope
    -- the fused sweep sums a row in an fp32 tree (lanes, warps, CTAs of the cluster), the two-kernel sweep in fp64."""
    syn = load_synthetic()
    H1 = restatement.fill_synthetic(P, V, 5)
    g = restatement.forward(H1, syn.emission_profile(V))
    rows = []
    for scale in (1.0, 8.0, 64.0):
        H = H1 * np.float32(scale)
        gs = g * scale  # the same emission seen through the scaled matrix
        x_ref = restatement.solve(H, gs, max_iterations=30, fixed_iterations=True, beta_laplace=0.0)[0]
        errs = {}
        for mode, name in ((SPLIT, "split"), (FUSED, "fused")):
            with sart.Engine(P, V, device=0, max_iterations=30, fixed_iterations=True, sweep_mode=mode) as eng:
                eng.upload_rows(H)
                eng.finalize()
                errs[name] = rel_l2(eng.solve(gs)[0], x_ref)
        rows.append((scale, float(H.sum(axis=1).mean()), errs["split"], errs["fused"]))
    with capsys.disabled():
        for scale, lam, es, ef in rows:
            print(f"\n  precision envelope {P}x{V}: mean ray length {lam:7.2f}  rel-L2 vs oracle: split {es:.2e}  fused {ef:.2e}", end="")
    for scale, lam, es, ef in rows:
        bound = max(TOL, 4e-6 * lam)  # 1e-5 at physical scale; beyond it the fp32 iterate sets the floor
        assert es <= bound and ef <= bound, rows
        assert ef <= 4 * es + 1e-7, rows  # the fp32 row tree of the fused sweep is not what limits the accuracy
