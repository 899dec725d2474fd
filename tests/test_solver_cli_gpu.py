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
    argv = [EXE, "-o", str(out_path), "-m", str(kw["max_iterations"]), "-c", repr(kw["conv_tolerance"]),
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
    out = subprocess.run([EXE, "--use_cpu", *files], capture_output=True, text=True)
    assert out.returncode == 1 and "no CPU solver" in out.stderr
    out = subprocess.run([EXE, files[0], files[1]], capture_output=True, text=True)  # an image file and one rtm block only
    assert out.returncode == 1
