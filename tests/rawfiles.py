"""Writers for the flat stand-in files of mpi-cuda-sartsolver_b200/host/raw_files.hpp (tests only)."""
import struct

import numpy as np


def write_rtm(path, block, pixel_offset, voxel_offset):
    block = np.ascontiguousarray(block, dtype=np.float32)
    with open(path, "wb") as f:
        f.write(b"SARTRTM1")
        f.write(struct.pack("<4Q", pixel_offset, block.shape[0], voxel_offset, block.shape[1]))
        f.write(block.tobytes())


def write_img(path, frames, times, pixel_offset):
    frames = np.ascontiguousarray(frames, dtype=np.float64)
    with open(path, "wb") as f:
        f.write(b"SARTIMG1")
        f.write(struct.pack("<3Q", pixel_offset, frames.shape[1], frames.shape[0]))
        f.write(np.ascontiguousarray(times, dtype=np.float64).tobytes())
        f.write(frames.tobytes())


def write_lap(path, flat_index, value, nvoxel):
    flat_index = np.asarray(flat_index, dtype=np.uint64)
    with open(path, "wb") as f:
        f.write(b"SARTLAP1")
        f.write(struct.pack("<2Q", nvoxel, len(flat_index)))
        f.write((flat_index // np.uint64(nvoxel)).astype(np.uint64).tobytes())
        f.write((flat_index % np.uint64(nvoxel)).astype(np.uint64).tobytes())
        f.write(np.ascontiguousarray(value, dtype=np.float32).tobytes())


def read_solution(path):
    raw = open(path, "rb").read()
    assert raw[:8] == b"SARTSOL1"
    nvoxel, nframes = struct.unpack_from("<2Q", raw, 8)
    rec = 16 + 8 * nvoxel
    out = []
    for k in range(nframes):
        base = 24 + k * rec
        t, status, iterations = struct.unpack_from("<dii", raw, base)
        out.append((t, status, iterations, np.frombuffer(raw, np.float64, nvoxel, base + 16)))
    return out
