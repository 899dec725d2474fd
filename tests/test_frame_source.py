"""Frame selection and multi-camera synchronisation of the driver (host/frame_source.cpp), the restatement of the
reference's CompositeImage time logic (source/image.cpp:53-196), on hand-checkable timelines. CPU only."""
import json
import os
import subprocess

import numpy as np
import pytest

from tests.rawfiles import write_img

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "mpi-cuda-sartsolver_b200", "host", "sart_b200_timesync")


def timesync(tmp_path, time_range, *timelines):
    if not os.path.exists(EXE):
        subprocess.check_call(["make", "-C", os.path.dirname(EXE), "sart_b200_timesync"])
    paths, off = [], 0
    for c, times in enumerate(timelines):
        p = tmp_path / f"cam{c}.img"
        write_img(p, np.zeros((len(times), 3)), times, off)
        off += 3
        paths.append(str(p))
    out = subprocess.run([EXE, time_range, *paths], capture_output=True, text=True)
    if out.returncode:
        return out.returncode, out.stderr.strip()
    return 0, [json.loads(l) for l in out.stdout.splitlines()]


def test_single_camera_all_frames_and_interval(tmp_path):
    rc, fr = timesync(tmp_path, "", [0.0, 1.0, 2.0, 3.0, 4.0])
    assert rc == 0 and [f["indices"] for f in fr] == [[0], [1], [2], [3], [4]]
    assert np.allclose([f["time"] for f in fr], [0, 1, 2, 3, 4])
    rc, fr = timesync(tmp_path, "1:3", [0.0, 1.0, 2.0, 3.0, 4.0])  # both limits inclusive (source/image.cpp:90)
    assert rc == 0 and [f["indices"] for f in fr] == [[1], [2], [3]]
    rc, fr = timesync(tmp_path, "0.5:1.5,2.5:4", [0.0, 1.0, 2.0, 3.0, 4.0])
    assert rc == 0 and [f["indices"] for f in fr] == [[1], [3], [4]]


def test_two_cameras_aligned_and_slightly_shifted(tmp_path):
    t = [0.0, 1.0, 2.0, 3.0]
    rc, fr = timesync(tmp_path, "", t, t)
    assert rc == 0 and [f["indices"] for f in fr] == [[0, 0], [1, 1], [2, 2], [3, 3]]
    rc, fr = timesync(tmp_path, "", t, [x + 0.1 for x in t])  # within the default threshold (= step = 1 s)
    assert rc == 0 and [f["indices"] for f in fr] == [[0, 0], [1, 1], [2, 2], [3, 3]]
    assert np.allclose([f["camera_time"][1] - f["camera_time"][0] for f in fr], 0.1)


def test_threshold_drops_unsynchronised_frames(tmp_path):
    t = [0.0, 1.0, 2.0, 3.0]
    rc, err = timesync(tmp_path, "0:10:1:0.2", t, [x + 0.4 for x in t])
    assert rc == 1 and "No composite images can be created" in err
    # only the frames of camera 1 that come within 0.2 s of a grid point survive
    rc, fr = timesync(tmp_path, "0:10:1:0.2", t, [0.05, 1.4, 2.1, 3.45])
    assert rc == 0 and [f["indices"] for f in fr] == [[0, 0], [2, 2]]


def test_slower_camera_sets_the_step(tmp_path):
    # step 0 -> the largest of the cameras' smallest spacings (2 s): the fast camera contributes every other frame
    rc, fr = timesync(tmp_path, "", [0.0, 1.0, 2.0, 3.0, 4.0], [0.0, 2.0, 4.0])
    assert rc == 0 and [f["indices"] for f in fr] == [[0, 0], [2, 1], [4, 2]]
    assert np.allclose([f["time"] for f in fr], [0, 2, 4])


def test_unsorted_timeline_is_rejected(tmp_path):
    rc, err = timesync(tmp_path, "", [0.0, 2.0, 1.0])
    assert rc == 1 and "not sorted by time" in err


@pytest.mark.parametrize("seed", range(40))
def test_random_timelines_select_what_the_reference_selects(tmp_path, seed):
    """The driver's synchronisation (a search per grid point) against the restatement of the reference's own routine
    (oracle/timesync_oracle.py, source/image.cpp:83-196: it scatters every frame onto its grid point and the two
    neighbours): same composite frames, same camera frames in each, same time stamps -- jittered cameras of different
    rates, dropped frames, explicit and default step / threshold, several intervals."""
    from oracle.timesync_oracle import composite_frames
    rng = np.random.default_rng(100 + seed)
    ncam = int(rng.integers(1, 4))
    timelines = []
    for _ in range(ncam):
        period = float(rng.choice([0.5, 1.0, 1.0, 2.0]))
        n = int(rng.integers(3, 30))
        t = 0.25 * rng.random() + period * np.arange(n) + rng.normal(0, 0.08 * period, n) * (rng.random() < 0.7)
        t = np.sort(np.round(np.abs(t), 6))
        t = t[rng.random(len(t)) > 0.15] if len(t) > 4 else t  # dropped frames
        t = np.unique(t)
        timelines.append([float(v) for v in t])
    choice = seed % 4
    if choice == 0:
        spec, intervals = "", [(0.0, float("inf"), 0.0, 0.0)]
    elif choice == 1:
        spec, intervals = "0:100:1:0.3", [(0.0, 100.0, 1.0, 0.3)]
    elif choice == 2:
        spec, intervals = "1:6,7.5:40:0.5", [(1.0, 6.0, 0.0, 0.0), (7.5, 40.0, 0.5, 0.0)]
    else:
        spec, intervals = "0:50:2:1.7", [(0.0, 50.0, 2.0, 1.7)]  # (the parser refuses a threshold above the step)
    want_idx, want_cam_t, want_t = composite_frames(timelines, intervals)
    rc, got = timesync(tmp_path, spec, *timelines)
    if not want_idx:
        assert rc == 1 and "No composite images can be created" in got
        return
    assert rc == 0, got
    assert [f["indices"] for f in got] == want_idx
    assert np.allclose([f["time"] for f in got], want_t, rtol=0, atol=1e-9)
    assert np.allclose([f["camera_time"] for f in got], want_cam_t, rtol=0, atol=1e-9)


def test_chunk_reads_fetch_runs_of_frames_in_one_go(tmp_path):
    """The frame cache of the reference (source/image.cpp:268-331) as the driver has it: the frames of a chunk come in with
    one read per camera and run of consecutive frame indices, and hold exactly what frame-by-frame reads give -- whole
    frames and a rank's pixel window alike."""
    rng = np.random.default_rng(7)
    t0 = [0.0, 1.0, 2.0, 3.0, 4.0, 5.0, 6.0, 7.0]
    t1 = [0.0, 0.5, 1.0, 1.5, 2.0, 2.5, 3.0, 3.5, 4.0, 4.5, 5.0, 5.5, 6.0, 6.5, 7.0]  # twice as fast: every other frame is used
    paths, off = [], 0
    for c, times in enumerate((t0, t1)):
        p = tmp_path / f"cam{c}.img"
        write_img(p, rng.random((len(times), 5)), times, off)
        off += 5
        paths.append(str(p))
    if not os.path.exists(EXE):
        subprocess.check_call(["make", "-C", os.path.dirname(EXE), "sart_b200_timesync"])
    for window in (None, "3:4"):
        env = dict(os.environ, SART_TIMESYNC_CHUNK="4")
        if window:
            env["SART_TIMESYNC_WINDOW"] = window
        out = subprocess.run([EXE, "", *paths], capture_output=True, text=True, env=env)
        assert out.returncode == 0, out.stderr
        lines = [json.loads(l) for l in out.stdout.splitlines()]
        frames, summary = lines[:-1], lines[-1]
        assert [f["indices"] for f in frames] == [[k, 2 * k] for k in range(8)]
        assert summary["equal"] == 1 and summary["checksum"] > 0
        # camera 0: one run per chunk of four frames; camera 1 uses every other frame: no runs, one read per frame
        assert summary["chunk_reads"] == 2 * 1 + 2 * 4 and summary["frame_reads"] == 16
