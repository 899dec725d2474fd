"""CPU tests of the one-process-per-GPU plumbing (world size 2 over gloo): the communicator id reaches every rank,
timings combine as max / sum over ranks, and frames are sliced by the reference's row partition."""
import os
import socket
import subprocess
import sys
import textwrap

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = textwrap.dedent("""
    import json, os, sys
    sys.path.insert(0, %r)
    import numpy as np
    from bench import load_module
    from __graft_entry__ import load_package
    cpl = load_module("control_plane")
    sart = load_package()
    cp = cpl.ControlPlane.from_env(timeout_s=120)
    uid = cp.broadcast_bytes(bytes(range(128)) if cp.rank == 0 else None)
    npixel = 1001
    off, n = cpl.shard_range(npixel, cp.world, cp.rank)
    assert (off, n) == sart.shard_range(npixel, cp.world, cp.rank)
    frame = np.arange(npixel, dtype=np.float64)
    local = frame[off:off + n]
    out = {
        "rank": cp.rank, "world": cp.world, "uid_ok": uid == bytes(range(128)),
        "max": cp.max(10.0 + cp.rank), "sum": cp.sum(local.sum()), "ranges": cp.gather((off, n)),
    }
    cp.barrier()
    cp.shutdown()
    print("RESULT " + json.dumps(out))
""") % ROOT


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_world_size_two_gloo():
    port = _free_port()
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE="2", LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, "-c", WORKER], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.PIPE, text=True))
    results = []
    for p in procs:
        out, err = p.communicate(timeout=300)
        assert p.returncode == 0, err[-2000:]
        import json
        results.append(json.loads([l for l in out.splitlines() if l.startswith("RESULT ")][0][7:]))
    for r in results:
        assert r["world"] == 2 and r["uid_ok"]
        assert r["max"] == 11.0
        assert r["sum"] == float(np.arange(1001).sum())
        assert r["ranges"] == [[0, 501], [501, 500]]


def test_single_process_control_plane_needs_no_rendezvous():
    sys.path.insert(0, ROOT)
    from bench import load_module
    cpl = load_module("control_plane")
    cp = cpl.ControlPlane(0, 1, 0)
    assert cp.broadcast_bytes(b"abc") == b"abc" and cp.max(3) == 3.0 and cp.sum(2) == 2.0 and cp.gather(5) == [5]
    cp.barrier()
    covered = 0
    for r in range(7):
        off, n = cpl.shard_range(100, 7, r)
        assert off == covered
        covered += n
    assert covered == 100
