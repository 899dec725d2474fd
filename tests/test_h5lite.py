"""h5lite (host/h5lite.cpp), the native HDF5 subset reader/writer. Not verifiable against libhdf5 on this image; checked
against the one third-party HDF5 file present (scipy's MATLAB v7.3 test file), against an independent pure-Python
writer for chunked / compressed data and variable-length strings, and writer -> reader round trips. CPU only."""
import glob
import os

import numpy as np
import pytest

from tests.h5w import H5Writer, h5ls, write_chunked_h5


def test_reads_the_third_party_file_on_the_image():
    import scipy.io.matlab
    hits = glob.glob(os.path.join(os.path.dirname(scipy.io.matlab.__file__), "tests", "data", "testhdf5_7.4_GLNX86.mat"))
    if not hits:
        pytest.skip("scipy's MATLAB v7.3 test file is not installed")
    objs = h5ls(hits[0])  # written by MATLAB 7.4's HDF5 library in 2008: 512-byte user block, version-2 layout message
    d = objs["testdouble"]
    assert (d["kind"], d["class"], d["size"], d["dims"], d["layout"]) == ("dataset", "float", 8, [9, 1], 1)
    np.testing.assert_allclose(d["values"], np.linspace(0, 2 * np.pi, 9), rtol=1e-15)
    assert d["attributes"]["MATLAB_class"] == "double"
    assert d["contiguous_offset"] == 4096


def test_writer_reader_round_trip(tmp_path):
    p = tmp_path / "rt.h5"
    rng = np.random.default_rng(1)
    value = rng.standard_normal((5, 37))
    frames = rng.random((3, 4, 6))
    w = H5Writer(p)
    w.group("solution")
    w.group("solution/deep")
    w.dataset("solution/value", value)
    w.dataset("solution/status", np.array([0, -1, 0, 0, -1], dtype=np.int32))
    w.dataset("solution/deep/frames", frames)
    w.dataset("solution/idx", np.array([0, 2 ** 40, 7], dtype=np.uint64))
    w.dataset("solution/f32", value.astype(np.float32))
    for k in range(40):  # more links than the default 8 per symbol node
        w.dataset(f"solution/deep/d{k:02d}", np.full(3, float(k)))
    w.attr("solution", "nvoxel", 37)
    w.attr("solution", "beta", 0.02)
    w.attr("solution", "name", "with_reflections")
    w.attr("solution/value", "flag", np.int32(-7))
    w.close()
    o = h5ls(p)
    np.testing.assert_array_equal(np.array(o["solution/value"]["values"]).reshape(5, 37), value)
    np.testing.assert_array_equal(np.array(o["solution/deep/frames"]["values"]).reshape(3, 4, 6), frames)
    assert o["solution/status"]["values"] == [0, -1, 0, 0, -1] and o["solution/status"]["signed"] == 1
    assert o["solution/idx"]["values"] == [0, 2 ** 40, 7]
    np.testing.assert_array_equal(np.array(o["solution/f32"]["values"], dtype=np.float32).reshape(5, 37), value.astype(np.float32))
    assert o["solution"]["attributes"] == {"beta": 0.02, "name": "with_reflections", "nvoxel": 37}
    assert o["solution/value"]["attributes"] == {"flag": -7}
    assert all(o[f"solution/deep/d{k:02d}"]["values"] == [float(k)] * 3 for k in range(40))


@pytest.mark.parametrize("deflate,shuffle,chunk", [(True, True, (16, 32)), (True, False, (7, 5)), (False, False, (64, 64)), (False, True, (10, 100))])
def test_reads_chunked_filtered_datasets_and_vlen_strings(tmp_path, deflate, shuffle, chunk):
    p = tmp_path / "chunked.h5"
    a = np.random.default_rng(2).random((50, 70), dtype=np.float32)
    write_chunked_h5(p, a, chunk, deflate=deflate, shuffle=shuffle, vlen_attr=("camera_name", "divertor_view_3"))
    o = h5ls(p)
    d = o["data"]
    assert d["layout"] == 2 and d["dims"] == [50, 70] and d["contiguous_offset"] == -1
    assert d["filters"] == ([2] if shuffle else []) + ([1] if deflate else [])
    np.testing.assert_array_equal(np.array(d["values"], dtype=np.float32).reshape(50, 70), a)
    assert o["/"]["attributes"]["camera_name"] == "divertor_view_3"  # variable-length string from the global heap


def test_rejects_what_it_cannot_read(tmp_path):
    import subprocess
    from tests.h5w import HOST
    p = tmp_path / "not.h5"
    p.write_bytes(b"definitely not hdf5" * 10)
    out = subprocess.run([os.path.join(HOST, "sart_b200_h5ls"), str(p)], capture_output=True, text=True)
    assert out.returncode == 1 and "is not an HDF5 file" in out.stderr
    v2 = bytearray(bytes([0x89]) + b"HDF\r\n\x1a\n" + bytes([9]) + b"\0" * 100)  # a superblock version that does not exist
    p.write_bytes(bytes(v2))
    out = subprocess.run([os.path.join(HOST, "sart_b200_h5ls"), str(p)], capture_output=True, text=True)
    assert out.returncode == 1 and "not supported" in out.stderr


def test_extendible_datasets_grow_flush_by_flush(tmp_path):
    """The solution file's shape (source/solution.cpp:60-165): datasets with an unlimited first dimension, extended every
    few frames. After EVERY flush the file on disk must be a complete file holding exactly the rows flushed so far --
    through a chunk index of one, two and three levels (64 children per node), one-row chunks for the 2-D dataset and
    a partly filled last chunk for the 1-D ones -- next to ordinary datasets and attributes written at lay-out time."""
    p = tmp_path / "grow.h5"
    rng = np.random.default_rng(3)
    nvox = 11
    w = H5Writer(p)
    w.group("solution")
    w.attr("solution", "nvoxel", nvox)
    w.dataset("solution/fixed", np.arange(6, dtype=np.float64).reshape(2, 3))
    w.extendible("solution/value", np.float64, (nvox,), 1)
    w.extendible("solution/time", np.float64, (), 7)
    w.extendible("solution/status", np.int32, (), 7)
    value = rng.standard_normal((3, nvox))
    time_ = np.arange(3, dtype=np.float64)
    status = np.array([0, -1, 0], dtype=np.int32)
    w.append("solution/value", value)   # rows buffered before the file exists
    w.append("solution/time", time_)
    w.append("solution/status", status)
    w.layout()

    def check():
        o = h5ls(p)
        n = len(time_)
        assert o["solution/value"]["dims"] == [n, nvox] and o["solution/value"]["layout"] == 2
        np.testing.assert_array_equal(np.array(o["solution/value"]["values"]).reshape(n, nvox), value)
        np.testing.assert_array_equal(o["solution/time"]["values"], time_)
        assert o["solution/status"]["values"] == status.tolist()
        assert o["solution/fixed"]["values"] == [0.0, 1.0, 2.0, 3.0, 4.0, 5.0] and o["solution"]["attributes"] == {"nvoxel": nvox}
        assert os.path.getsize(p) >= 96

    import os
    check()
    for batch in (2, 5, 60, 1, 70, 4100):  # 3 -> 5 -> 10 -> 70 (two index levels) -> 71 -> 141 -> 4241 rows (three levels)
        v = rng.standard_normal((batch, nvox))
        t = len(time_) + np.arange(batch, dtype=np.float64)
        s = rng.integers(-1, 1, batch).astype(np.int32)
        w.append("solution/value", v)
        w.append("solution/time", t)
        w.append("solution/status", s)
        before = h5ls(p, values=False)["solution/value"]["dims"][0]
        assert before == len(time_)  # appended rows do not show before the flush
        w.flush()
        value, time_, status = np.vstack([value, v]), np.concatenate([time_, t]), np.concatenate([status, s])
        check()
    w.close()
    check()


def test_reads_libver_latest_structures(tmp_path):
    """Superblock 2, version-2 object headers (one with a continuation block), new-style groups with link messages, layout
    version 4 with contiguous / compact / single-chunk / implicit / fixed-array storage: what h5py writes with
    libver="latest" for small groups, here from an independent pure-Python writer (tests/h5w.py)."""
    from tests.h5w import write_latest_h5
    p = tmp_path / "latest.h5"
    a = np.random.default_rng(4).standard_normal((7, 10))
    small = write_latest_h5(p, a, attr_value=2.5)
    o = h5ls(p)
    assert o["g"]["kind"] == "group"
    for name in ("contiguous", "single", "implicit", "fixed_array"):
        d = o["g/" + name]
        assert d["dims"] == [7, 10] and d["class"] == "float" and d["size"] == 8, (name, d)
        np.testing.assert_array_equal(np.array(d["values"]).reshape(7, 10), a, err_msg=name)
    np.testing.assert_array_equal(np.array(o["g/compact"]["values"]).reshape(2, 3), small)
    assert o["g/single"]["attributes"] == {"scale": 2.5}
    assert o["g/contiguous"]["layout"] == 1 and o["g/fixed_array"]["layout"] == 2
