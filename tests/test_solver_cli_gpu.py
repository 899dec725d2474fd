"""The `sartsolver` stand-in (host/sart_b200_solver.cpp) end to end on the GPU: the reference's command line, inputs as
one ray-transfer file per camera x voxel segment and one image file per camera (any order on the command line), the
matrix streamed through the pinned staging ring, frame loop with warm start / --no_guess -- against the golden vectors."""
import os
import subprocess

import numpy as np
import pytest

from tests.cases import build_case, rel_l2
from tests.rawfiles import read_solution, write_img, write_lap, write_rtm

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "mpi-cuda-sartsolver_b200", "host", "sart_b200_solver")


def _write_inputs(tmp_path, case):
    H, frames = case["H"], case["frames"]
    P, V = H.shape
    pcut, vcut = P // 3, V // 2 + 1
    files = []
    for ci, (r0, r1) in enumerate(((0, pcut), (pcut, P))):        # two cameras
        for si, (c0, c1) in enumerate(((0, vcut), (vcut, V))):    # two voxel segments
            p = tmp_path / f"cam{ci}_seg{si}.rtm"
            write_rtm(p, H[r0:r1, c0:c1], r0, c0)
            files.append(str(p))
        p = tmp_path / f"cam{ci}.img"
        write_img(p, frames[:, r0:r1], np.arange(frames.shape[0], dtype=float), r0)
        files.append(str(p))
    return files[::-1]  # the reference accepts its input files in any order


@pytest.mark.parametrize("name", ["series_warm4", "series_noguess3", "linear_masks_guess"])
def test_solver_cli_matches_golden(name, restatement, golden, tmp_path):
    if not os.path.exists(EXE):
        pytest.fail(f"{EXE} is not built (python -c 'import __graft_entry__ as g; g.build()')")
    case = build_case(name, restatement)
    kw = case["kw"]
    files = _write_inputs(tmp_path, case)
    out_path = tmp_path / "solution.bin"
    # --use_cpu: the golden vectors are the reference's CPU solver; without it the driver follows the GPU solver where the
    # two differ (saturated pixels in the initial guess, the 1e-7 floor), as the reference's own default run does
    argv = [EXE, "--use_cpu", "-o", str(out_path), "-m", str(kw["max_iterations"]), "-c", repr(kw["conv_tolerance"]),
            "-R", str(kw.get("relaxation", 1.0))]
    if case["lap"] is not None:
        lap = tmp_path / "laplacian.lap"
        write_lap(lap, case["lap"][0], case["lap"][1], case["H"].shape[1])
        argv += ["-l", str(lap), "-b", str(kw.get("beta_laplace", 2e-2))]
    if case["no_guess"]:
        argv.append("--no_guess")
    env = dict(os.environ, SART_STAGING_MB="1")
    out = subprocess.run(argv + files, capture_output=True, text=True, timeout=300, env=env)
    assert out.returncode == 0, out.stderr
    nframes = case["frames"].shape[0]
    assert out.stdout.count("Processed in:") == nframes
    records = read_solution(out_path)
    assert len(records) == nframes
    gx, gst, git = golden[name + "/x"], golden[name + "/status"], golden[name + "/iterations"]
    fixed = kw["conv_tolerance"] < 1e-20
    same_history = True
    for f, (t, status, iterations, x) in enumerate(records):
        assert t == float(f)
        if fixed:
            assert (status, iterations) == (gst[f], git[f])
            assert rel_l2(x, gx[f]) <= 1e-5
        else:
            assert abs(iterations - git[f]) <= 1
            same_history = same_history and iterations == git[f]
            if same_history:  # a warm-started series only stays comparable while every earlier frame stopped alike
                assert status == gst[f] and rel_l2(x, gx[f]) <= 1e-5


def test_solver_cli_rejects_like_the_reference(tmp_path, restatement):
    case = build_case("max_iterations_exceeded", restatement)
    files = _write_inputs(tmp_path, case)
    out = subprocess.run([EXE, "-R", "2", *files], capture_output=True, text=True)
    assert out.returncode == 1 and "relaxation must be within (0, 1] interval" in out.stderr
    out = subprocess.run([EXE, files[0], files[1]], capture_output=True, text=True)  # an image file and one rtm block only
    assert out.returncode == 1


def test_solver_cli_default_follows_the_gpu_solver_and_use_cpu_the_cpu_solver(tmp_path, restatement):
    """The reference's default command line runs its CUDA solver, --use_cpu its CPU solver (source/main.cpp:88-97); where
    the two differ observably -- here: saturated pixels in the default initial guess -- the driver follows the one the same
    command line would have run."""
    case = build_case("linear_masks_guess", restatement)
    case["frames"] = case["frames"].copy()
    case["frames"][:, ::13] = -1.0  # saturated pixels, no warm start
    files = _write_inputs(tmp_path, case)
    H, g = case["H"], case["frames"][0]
    sols = {}
    for flag in ("--use_cpu", None):
        out_path = tmp_path / f"solution_{flag}.bin"
        argv = [EXE] + ([flag] if flag else []) + ["-o", str(out_path), "-m", "3", "-c", "1e-30"]
        out = subprocess.run(argv + files, capture_output=True, text=True, timeout=300)
        assert out.returncode == 0, out.stderr
        sols[flag] = read_solution(out_path)[0][3]
        x_ref = restatement.solve(H, g, max_iterations=3, fixed_iterations=True, beta_laplace=0.0, gpu_semantics=flag is None)[0]
        assert rel_l2(sols[flag], x_ref) <= 1e-5
    assert rel_l2(sols[None], sols["--use_cpu"]) > 1e-4


def _write_hdf5_inputs(tmp_path, case, rtm_name="with_reflections"):
    """The same problem as HDF5 files in the reference's schemas: two cameras (pixels picked out of a larger sensor by a
    frame mask) x two voxel segments (the second one of camB stored sparse), one image file per camera."""
    from tests.h5w import write_image_h5, write_rtm_h5
    H, frames = case["H"], case["frames"]
    P, V = H.shape
    pcut, vcut = P // 3, V // 2 + 1
    rng = np.random.default_rng(123)
    files = []
    times = np.arange(frames.shape[0], dtype=float) * 0.5 + 10.0
    for cam, (r0, r1) in (("camA", (0, pcut)), ("camB", (pcut, P))):
        n = r1 - r0
        width = 16
        height = n // width + 3
        positions = np.sort(rng.choice(height * width, size=n, replace=False))
        mask = np.zeros(height * width, dtype=np.int32)
        mask[positions] = 1
        for seg, (c0, c1) in enumerate(((0, vcut), (vcut, V))):
            p = tmp_path / f"{cam}_seg{seg}.h5"
            write_rtm_h5(p, H[r0:r1, c0:c1], cam, rtm_name, mask.reshape(height, width), np.arange(c0, c1), (V, 1, 1),
                         sparse=(cam == "camB" and seg == 1))
            files.append(str(p))
        sensor = np.full((frames.shape[0], height * width), -999.0)
        sensor[:, positions] = frames[:, r0:r1]
        p = tmp_path / f"{cam}_image.h5"
        write_image_h5(p, cam, times, sensor.reshape(frames.shape[0], height, width))
        files.append(str(p))
    return files[::-1], times


@pytest.mark.parametrize("name", ["series_warm4", "linear_masks_guess"])
def test_solver_cli_on_hdf5_files_matches_golden(name, restatement, golden, tmp_path):
    from tests.h5w import h5ls, write_laplacian_h5
    case = build_case(name, restatement)
    kw = case["kw"]
    files, times = _write_hdf5_inputs(tmp_path, case)
    out_path = tmp_path / "solution.h5"
    argv = [EXE, "--use_cpu", "-o", str(out_path), "-m", str(kw["max_iterations"]), "-c", repr(kw["conv_tolerance"]),
            "-R", str(kw.get("relaxation", 1.0)), "-n", "with_reflections"]
    V = case["H"].shape[1]
    if case["lap"] is not None:
        lap = tmp_path / "laplacian.h5"
        write_laplacian_h5(lap, case["lap"][0], case["lap"][1], V)
        argv += ["-l", str(lap), "-b", str(kw.get("beta_laplace", 2e-2))]
    out = subprocess.run(argv + ["--max_cached_solutions", "3"] + files, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    o = h5ls(out_path)
    nframes = case["frames"].shape[0]
    x = np.array(o["solution/value"]["values"]).reshape(nframes, V)
    assert o["solution/value"]["dims"] == [nframes, V]
    np.testing.assert_allclose(o["solution/time"]["values"], times)
    np.testing.assert_allclose(o["solution/time_camA"]["values"], times)
    np.testing.assert_allclose(o["solution/time_camB"]["values"], times)
    gx, gst, git = golden[name + "/x"], golden[name + "/status"], golden[name + "/iterations"]
    status, iterations = o["solution/status"]["values"], o["solution/iterations"]["values"]
    fixed = kw["conv_tolerance"] < 1e-20
    same_history = True
    for f in range(nframes):
        if fixed:
            assert (status[f], iterations[f]) == (gst[f], git[f]) and rel_l2(x[f], gx[f]) <= 1e-5
        else:
            assert abs(iterations[f] - git[f]) <= 1
            same_history = same_history and iterations[f] == git[f]
            if same_history:
                assert status[f] == gst[f] and rel_l2(x[f], gx[f]) <= 1e-5
    # the voxel grid travels from the RTM files to the output (source/main.cpp:125, 143)
    vm = o["voxel_map"]["attributes"]
    assert (vm["nx"], vm["ny"], vm["nz"]) == (V, 1, 1) and vm["coordinate_system"] == "cylindrical"
    assert (vm["xmin"], vm["xmax"], vm["ymin"], vm["ymax"]) == (1.5, 4.5, 0, 1)
    assert o["voxel_map/i"]["values"] == list(range(V)) and o["voxel_map/value"]["values"] == list(range(V))


def test_solver_cli_hdf5_consistency_checks(tmp_path, restatement):
    from tests.h5w import write_image_h5, write_rtm_h5
    case = build_case("max_iterations_exceeded", restatement)
    files, _ = _write_hdf5_inputs(tmp_path, case)
    # an image whose wavelength is too far from the RTM's (source/hdf5files.cpp:302-316)
    bad = [f for f in files if f.endswith("camA_image.h5")][0]
    frames = np.zeros((1, 3, 16))
    write_image_h5(bad, "camA", [0.0], frames, wavelength=900.0)
    out = subprocess.run([EXE, "-o", str(tmp_path / "o.h5"), *files], capture_output=True, text=True)
    assert out.returncode == 1 and "wavelength" in out.stderr
    out = subprocess.run([EXE, "-o", str(tmp_path / "o.h5"), "-n", "no_such_rtm", *files], capture_output=True, text=True)
    assert out.returncode == 1 and "no_such_rtm" in out.stderr
