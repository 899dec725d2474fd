"""Host-side synthetic workload pieces shared by the tests and bench.py (numpy only, no device code).

The ray-transfer matrix itself is generated on the device (``sart_fill_synthetic``); what lives here is
the small stuff a caller of the reference would read from files: an emission profile to project, and a
Laplacian regularisation matrix in the reference's COO convention (flattened index ``i * nvoxel + j``
sorted ascending, fp32 values: /root/reference/source/include/laplacian.hpp:15-19,
/root/reference/source/laplacian.cpp:67-82).
"""
import numpy as np


def voxel_grid_shape(nvoxel):
    """Factor nvoxel = nx * nz with nx <= nz as square as possible (a 2-D R x Z grid)."""
    nx = int(np.floor(np.sqrt(nvoxel)))
    while nx > 1 and nvoxel % nx:
        nx -= 1
    return nx, nvoxel // nx


def emission_profile(nvoxel, t=0.0, dtype=np.float64):
    """Smooth, strictly positive emission on the voxel grid: two Gaussian blobs on a small pedestal.

    ``t`` in [0, 1] moves the blobs slowly (time series / warm-start workloads)."""
    nx, nz = voxel_grid_shape(nvoxel)
    u = (np.arange(nx) + 0.5) / nx
    v = (np.arange(nz) + 0.5) / nz
    U, W = np.meshgrid(u, v, indexing="ij")
    c1 = (0.35 + 0.10 * t, 0.40 + 0.05 * t)
    c2 = (0.70 - 0.05 * t, 0.65 - 0.10 * t)
    prof = (0.05 + 1.00 * np.exp(-((U - c1[0]) ** 2 + (W - c1[1]) ** 2) / (2 * 0.08 ** 2))
            + 0.60 * np.exp(-((U - c2[0]) ** 2 + (W - c2[1]) ** 2) / (2 * 0.12 ** 2)))
    return prof.reshape(-1).astype(dtype)


def laplacian_coo(nvoxel):
    """5-point graph Laplacian on the nx x nz voxel grid: diagonal = number of neighbours, off-diagonals -1.

    Returns (flat_index uint64 [nnz], value float32 [nnz]) sorted by flat_index = i * nvoxel + j."""
    nx, nz = voxel_grid_shape(nvoxel)
    cell = np.arange(nvoxel, dtype=np.int64).reshape(nx, nz)
    rows, cols, vals = [], [], []
    degree = np.zeros((nx, nz), dtype=np.int64)
    for di, dj in ((1, 0), (-1, 0), (0, 1), (0, -1)):
        src = cell[max(0, -di):nx - max(0, di), max(0, -dj):nz - max(0, dj)]
        dst = cell[max(0, di):nx - max(0, -di), max(0, dj):nz - max(0, -dj)]
        rows.append(src.reshape(-1))
        cols.append(dst.reshape(-1))
        vals.append(np.full(src.size, -1.0, dtype=np.float32))
        degree[max(0, -di):nx - max(0, di), max(0, -dj):nz - max(0, dj)] += 1
    rows.append(cell.reshape(-1))
    cols.append(cell.reshape(-1))
    vals.append(degree.reshape(-1).astype(np.float32))
    rows, cols, vals = np.concatenate(rows), np.concatenate(cols), np.concatenate(vals)
    flat = rows.astype(np.uint64) * np.uint64(nvoxel) + cols.astype(np.uint64)
    order = np.argsort(flat, kind="stable")
    return flat[order], vals[order]
