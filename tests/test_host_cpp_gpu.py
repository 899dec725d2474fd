"""The C++ host side (SARTSolverB200 adaptor + the raw-file frame-loop driver) end to end on the GPU, against the
golden vectors of the reference: the same frame loop, warm start and --no_guess handling as source/main.cpp:127-140."""
import os
import struct
import subprocess

import numpy as np
import pytest

from tests.cases import CASES, build_case, rel_l2

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DRIVER = os.path.join(ROOT, "mpi-cuda-sartsolver_b200", "host", "sart_b200_frames")
# the same driver compiled against the reference's own RayTransferMatrix / LaplacianMatrix headers (oracle/Makefile; test
# only, built where /root/reference exists and shipped to the GPU box with oracle/_ref/)
DRIVER_REF_CONTAINERS = os.path.join(ROOT, "oracle", "_ref", "sart_b200_frames_refcontainers")


def _write_problem(path, case):
    H, frames, kw = case["H"], case["frames"], case["kw"]
    lap = case["lap"] if case["lap"] is not None else (np.zeros(0, np.uint64), np.zeros(0, np.float32))
    flags = (1 if case["x0"] is not None else 0) | (2 if case["no_guess"] else 0) | (4 if kw.get("logarithmic") else 0)
    with open(path, "wb") as f:
        f.write(struct.pack("<6Q", H.shape[0], H.shape[1], len(lap[0]), frames.shape[0], flags, kw["max_iterations"]))
        f.write(struct.pack("<5d", 1e-6, 1e-6, kw["conv_tolerance"], kw.get("beta_laplace", 2e-2), kw.get("relaxation", 1.0)))
        f.write(np.ascontiguousarray(H, np.float32).tobytes())
        f.write(np.ascontiguousarray(lap[0], np.uint64).tobytes())
        f.write(np.ascontiguousarray(lap[1], np.float32).tobytes())
        f.write(np.ascontiguousarray(frames, np.float64).tobytes())
        if case["x0"] is not None:
            f.write(np.ascontiguousarray(case["x0"], np.float64).tobytes())


@pytest.mark.parametrize("name", ["linear_masks_warmstart", "series_warm4", "series_noguess3", "log_laplacian_np3"])
def test_cpp_frame_loop_matches_golden(name, restatement, golden, tmp_path):
    if not os.path.exists(DRIVER):
        pytest.fail(f"{DRIVER} is not built (python -c 'import __graft_entry__ as g; g.build()')")
    case = build_case(name, restatement)
    prob, sol = tmp_path / "problem.bin", tmp_path / "solution.bin"
    _write_problem(prob, case)
    out = subprocess.run([DRIVER, str(prob), str(sol)], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    assert out.stdout.count("Processed in:") == case["frames"].shape[0]
    V = case["H"].shape[1]
    raw = open(sol, "rb").read()
    rec = 4 + 8 * V
    gx, gst, git = golden[name + "/x"], golden[name + "/status"], golden[name + "/iterations"]
    fixed = case["kw"]["conv_tolerance"] < 1e-20
    for f in range(case["frames"].shape[0]):
        status = struct.unpack_from("<i", raw, f * rec)[0]
        x = np.frombuffer(raw, np.float64, V, f * rec + 4)
        if fixed:
            assert status == gst[f]
            assert rel_l2(x, gx[f]) <= 1e-5
        else:
            iters = int(out.stdout.splitlines()[f].split("(")[1].split()[0])
            assert abs(iters - git[f]) <= 1
            if iters == git[f]:
                assert status == gst[f] and rel_l2(x, gx[f]) <= 1e-5


@pytest.mark.parametrize("name", ["linear_masks_warmstart", "series_warm4"])
def test_template_constructor_takes_the_reference_containers_as_they_are(name, restatement, tmp_path):
    """SARTSolverB200's constructor is a template over the matrix containers; this driver instantiates it with the
    reference's real classes (source/include/raytransfer.hpp:15-37, laplacian.hpp:15-38) and must give the very same bytes
    as the build with the look-alike containers."""
    if not os.path.exists(DRIVER_REF_CONTAINERS):
        pytest.skip("oracle/_ref/sart_b200_frames_refcontainers not built (needs /root/reference at build time)")
    case = build_case(name, restatement)
    prob = tmp_path / "problem.bin"
    _write_problem(prob, case)
    sols = []
    for exe in (DRIVER, DRIVER_REF_CONTAINERS):
        sol = tmp_path / (os.path.basename(exe) + ".bin")
        out = subprocess.run([exe, str(prob), str(sol)], capture_output=True, text=True, timeout=300)
        assert out.returncode == 0, out.stderr
        sols.append(open(sol, "rb").read())
    assert sols[0] == sols[1] and len(sols[0]) == case["frames"].shape[0] * (4 + 8 * case["H"].shape[1])


def test_cpp_adaptor_rejects_bad_arguments_like_the_reference(tmp_path, restatement):
    """A wrong-sized warm start must end the process with the reference's message (sartsolver_cuda.cpp:209-212)."""
    case = build_case("max_iterations_exceeded", restatement)
    case["x0"] = np.ones(case["H"].shape[1])
    prob = tmp_path / "p.bin"
    _write_problem(prob, case)
    raw = bytearray(open(prob, "rb").read())
    struct.pack_into("<d", raw, 48 + 3 * 8, -1.0)  # beta_laplace < 0
    open(prob, "wb").write(raw)
    out = subprocess.run([DRIVER, str(prob), str(tmp_path / "s.bin")], capture_output=True, text=True, timeout=300)
    assert out.returncode == 1 and "beta_laplace must be non-negative" in out.stderr
