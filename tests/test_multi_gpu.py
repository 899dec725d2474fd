"""Row-sharded solve on several GPUs of one box (one process per GPU, NCCL all-reduce between the engines) against the
CPU oracle run with the same rank split. Skipped on single-GPU boxes; run it with `gpurun --gpus 2`."""
import json
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _gpu_count():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_sharded_solve_matches_oracle(world):
    if _gpu_count() < world:
        pytest.skip(f"needs {world} GPUs")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "multi_gpu_worker.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-4000:]
    line = [l for l in out.stdout.splitlines() if l.startswith("RESULT ")][0]
    ranks = json.loads(line[7:])
    assert len(ranks) == world
    for name in ranks[0]["checks"]:
        hashes = set()
        for r in ranks:
            c = r["checks"][name]
            assert c["rho"] < 1e-13 and c["lam"] < 1e-13, (name, c)
            assert c["status"][0] == c["status"][1], (name, c)
            assert abs(c["iterations"][0] - c["iterations"][1]) <= 1, (name, c)
            if c["iterations"][0] == c["iterations"][1]:
                assert c["x"] <= 1e-5, (name, c)
            if c["iterations2"][0] == c["iterations2"][1]:
                assert c["x2"] <= 1e-5, (name, c)
            hashes.add(c["x_hash"])
            if c["batch"] is not None:  # sart_solve_batch against frame-by-frame solves on the same ranks
                b = c["batch"]
                assert b["iteration_gap"] <= 1, (name, b)
                import os as _os2  # the group's reduction travels over peer memory too, unless NCCL was asked for
                assert b["exchange"] == (1 if _os2.environ.get("SART_EXCHANGE") == "nccl" else 2), (name, b)
                if b["iteration_gap"] == 0:
                    assert b["same_status"] and b["x"] <= 1e-5, (name, b)
            if c["series"] is not None:  # sart_solve_series against warm-started solves on the same ranks
                assert c["series"]["consistent"] and c["series"]["x"] <= 1e-5, (name, c["series"])
            # the fused sweep reduces over peer memory (one kernel), the two-kernel sweep over NCCL
            import os as _os
            if c["mode"] == 2 and _os.environ.get("SART_EXCHANGE") != "nccl":
                assert c["exchange"] == 2, (name, c)
            else:
                assert c["exchange"] == 1, (name, c)
        assert len(hashes) == 1, f"{name}: ranks ended with different solutions"
        batch_hashes = {r["checks"][name]["batch"]["hash"] for r in ranks if r["checks"][name]["batch"] is not None}
        assert len(batch_hashes) <= 1, f"{name}: ranks ended with different batch solutions"
        series_hashes = {r["checks"][name]["series"]["hash"] for r in ranks if r["checks"][name]["series"] is not None}
        assert len(series_hashes) <= 1, f"{name}: ranks ended with different series solutions"


def test_cpp_solver_driver_on_two_gpus(tmp_path):
    """The C++ driver as two processes (rank / world size from the environment, communicator id through a file), each
    loading its own rows of HDF5 inputs, against the golden vectors of the reference."""
    if _gpu_count() < 2:
        pytest.skip("needs 2 GPUs")
    import numpy as np
    sys.path.insert(0, ROOT)
    from oracle.pyoracle import Restatement
    from tests.cases import build_case, rel_l2
    from tests.h5w import h5ls, write_laplacian_h5
    from tests.test_solver_cli_gpu import EXE, _write_hdf5_inputs
    name = "series_warm4"
    case = build_case(name, Restatement())
    golden = np.load(os.path.join(ROOT, "tests", "golden", "golden_v1.npz"))
    kw = case["kw"]
    files, times = _write_hdf5_inputs(tmp_path, case)
    V = case["H"].shape[1]
    lap = tmp_path / "laplacian.h5"
    write_laplacian_h5(lap, case["lap"][0], case["lap"][1], V)
    out_path = tmp_path / "solution.h5"
    argv = [EXE, "--use_cpu", "-o", str(out_path), "-m", str(kw["max_iterations"]), "-c", repr(kw["conv_tolerance"]), "-l", str(lap),
            "-b", str(kw.get("beta_laplace", 2e-2)), *files]
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE="2", LOCAL_RANK=str(rank),
                   SART_COMM_ID_FILE=str(tmp_path / "comm_id"))
        procs.append(subprocess.Popen(argv, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    outs = [p.communicate(timeout=600) for p in procs]
    for p, (so, se) in zip(procs, outs):
        assert p.returncode == 0, se[-2000:]
    assert outs[0][0].count("Processed in:") == 4 and outs[1][0].count("Processed in:") == 0
    o = h5ls(out_path)
    x = np.array(o["solution/value"]["values"]).reshape(4, V)
    its = o["solution/iterations"]["values"]
    git = golden[name + "/iterations"]
    same = True
    for f in range(4):
        assert abs(its[f] - git[f]) <= 1
        same = same and its[f] == git[f]
        if same:
            assert rel_l2(x[f], golden[name + "/x"][f]) <= 1e-5
