// TEST INFRASTRUCTURE ONLY -- the members of the reference's RayTransferMatrix / LaplacianMatrix whose own bodies live
// in HDF5-dependent translation units (declared: /root/reference/source/include/raytransfer.hpp:24-36,
// /root/reference/source/include/laplacian.hpp:26-36), filled from caller memory instead of files. Linked with the
// frame-loop driver built against the reference's REAL headers (oracle/Makefile: _ref/sart_b200_frames_refcontainers) to
// show that SARTSolverB200's template constructor takes those containers as they are. Never part of the product.
#include "ref_containers.hpp"

#include "laplacian.hpp"
#include "raytransfer.hpp"

namespace {
const float* g_matrix = nullptr;
const size_t* g_lap_index = nullptr;
const float* g_lap_value = nullptr;
size_t g_lap_nnz = 0;
}  // namespace

void ref_containers_set_sources(const float* matrix, const size_t* lap_index, const float* lap_value, size_t lap_nnz) {
    g_matrix = matrix;
    g_lap_index = lap_index;
    g_lap_value = lap_value;
    g_lap_nnz = lap_nnz;
}

RayTransferMatrix::RayTransferMatrix(size_t npixel, size_t nvoxel, size_t offset_pixel)
    : offset_pix(offset_pixel), npix(npixel), nvox(nvoxel) {}

const std::vector<float>& RayTransferMatrix::matrix() const { return mat; }

int RayTransferMatrix::read_hdf5(const std::map<std::string, std::vector<std::string>>&, const std::string&) {
    const float* first = g_matrix + offset_pix * nvox;  // the source holds ALL rows; this rank keeps its own
    mat.assign(first, first + npix * nvox);
    return 0;
}

LaplacianMatrix::LaplacianMatrix(size_t nvoxel) : nvox(nvoxel) {}

const std::vector<float>& LaplacianMatrix::value() const { return val; }

const std::vector<size_t>& LaplacianMatrix::index() const { return index1d; }

int LaplacianMatrix::read_hdf5(const std::string&) {
    index1d.assign(g_lap_index, g_lap_index + g_lap_nnz);
    val.assign(g_lap_value, g_lap_value + g_lap_nnz);
    return 0;
}
