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
