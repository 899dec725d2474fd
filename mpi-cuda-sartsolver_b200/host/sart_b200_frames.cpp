// Frame-loop driver over raw binary files: the part of the reference's main() that surrounds the hot path
// (source/main.cpp:88-110 solver set-up, :127-140 frame loop with warm start / --no_guess), with the HDF5 readers
// replaced by one flat little-endian file so that it runs where libhdf5 does not exist. Used by the tests to drive
// the C++ adaptor end to end; single rank.
//
//   sart_b200_frames <problem.bin> <solution.bin>
//
// problem.bin : u64 npixel, nvoxel, nnz, nframes, flags (1 has_guess | 2 no_guess | 4 logarithmic), max_iterations;
//               f64 ray_density_threshold, ray_length_threshold, conv_tolerance, beta_laplace, relaxation;
//               f32 H[npixel*nvoxel]; u64 lap_index[nnz]; f32 lap_value[nnz]; f64 frames[nframes*npixel];
//               f64 guess[nvoxel] when has_guess
// solution.bin: per frame i32 status, f64 solution[nvoxel]
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <iostream>
#include <vector>

#include "sartsolver_b200.hpp"

#ifdef SART_USE_REFERENCE_CONTAINERS
// Test build (oracle/Makefile, where the reference's sources exist): the reference's OWN container classes, included from
// where they lie; their HDF5-backed members are defined by oracle/ref_containers.cpp to fill from memory.
#include "laplacian.hpp"
#include "raytransfer.hpp"
#include "ref_containers.hpp"
#endif

namespace {

#ifndef SART_USE_REFERENCE_CONTAINERS
// Stand-ins with the member names of the reference's containers (source/include/raytransfer.hpp:15-37,
// source/include/laplacian.hpp:15-38): what SARTSolverB200's template constructor expects.
class RayTransferMatrix {
    size_t offset_pix, npix, nvox;
    std::vector<float> mat;

public:
    RayTransferMatrix(size_t npixel, size_t nvoxel, size_t offset_pixel = 0)
        : offset_pix(offset_pixel), npix(npixel), nvox(nvoxel), mat(npixel * nvoxel) {}
    std::vector<float>& storage() { return mat; }
    const std::vector<float>& matrix() const { return mat; }
    size_t offset_pixel() const { return offset_pix; }
    size_t npixel() const { return npix; }
    size_t nvoxel() const { return nvox; }
};

class LaplacianMatrix {
public:
    std::vector<size_t> index1d;
    std::vector<float> val;
    size_t size() const { return val.size(); }
    const std::vector<float>& value() const { return val; }
    const std::vector<size_t>& index() const { return index1d; }
};
#endif

template <typename T>
bool read_n(FILE* f, T* dst, size_t n) {
    return n == 0 || std::fread(dst, sizeof(T), n, f) == n;
}

}  // namespace

int main(int argc, char** argv) {
    if (argc != 3) {
        std::cerr << "usage: sart_b200_frames <problem.bin> <solution.bin>" << std::endl;
        return 2;
    }
    FILE* f = std::fopen(argv[1], "rb");
    if (!f) {
        std::cerr << "cannot open " << argv[1] << std::endl;
        return 1;
    }
    uint64_t hdr[6];
    double prm[5];
    if (!read_n(f, hdr, 6) || !read_n(f, prm, 5)) return 1;
    const uint64_t npixel = hdr[0], nvoxel = hdr[1], nnz = hdr[2], nframes = hdr[3], flags = hdr[4];
    const int max_iterations = (int)hdr[5];
    std::vector<double> frames(nframes * npixel), solution_vec;
#ifdef SART_USE_REFERENCE_CONTAINERS
    std::vector<float> matrix_src(npixel * nvoxel), lap_val(nnz);
    std::vector<size_t> lap_idx(nnz);
    bool ok = read_n(f, matrix_src.data(), npixel * nvoxel) && read_n(f, lap_idx.data(), nnz) &&
              read_n(f, lap_val.data(), nnz) && read_n(f, frames.data(), nframes * npixel);
    ref_containers_set_sources(matrix_src.data(), lap_idx.data(), lap_val.data(), nnz);
    RayTransferMatrix raytransfer(npixel, nvoxel, 0);  // source/main.cpp:76-77
    raytransfer.read_hdf5({}, "");
    LaplacianMatrix laplacian(nnz ? nvoxel : 0);       // source/main.cpp:73-74
    if (nnz) laplacian.read_hdf5("");
#else
    RayTransferMatrix raytransfer(npixel, nvoxel, 0);
    LaplacianMatrix laplacian;
    laplacian.index1d.resize(nnz);
    laplacian.val.resize(nnz);
    bool ok = read_n(f, raytransfer.storage().data(), npixel * nvoxel) && read_n(f, laplacian.index1d.data(), nnz) &&
              read_n(f, laplacian.val.data(), nnz) && read_n(f, frames.data(), nframes * npixel);
#endif
    if (ok && (flags & 1)) {
        solution_vec.resize(nvoxel);
        ok = read_n(f, solution_vec.data(), nvoxel);
    }
    std::fclose(f);
    if (!ok) {
        std::cerr << "short read on " << argv[1] << std::endl;
        return 1;
    }

    // source/main.cpp:88-110
    SARTSolverB200 solver(raytransfer, laplacian, /*rank=*/0, /*nranks=*/1, npixel, nullptr, (flags & 4) != 0);
    solver.set_ray_density_threshold(prm[0]);
    solver.set_ray_length_threshold(prm[1]);
    solver.set_convolution_tolerance(prm[2]);
    solver.set_beta_laplace(prm[3]);
    solver.set_relaxation(prm[4]);
    solver.set_max_iterations(max_iterations);

    FILE* out = std::fopen(argv[2], "wb");
    if (!out) {
        std::cerr << "cannot open " << argv[2] << std::endl;
        return 1;
    }
    // source/main.cpp:127-140
    std::vector<double> frame(npixel);
    for (uint64_t k = 0; k < nframes; ++k) {
        frame.assign(frames.begin() + k * npixel, frames.begin() + (k + 1) * npixel);
        const auto last = std::chrono::steady_clock::now();
        const int32_t status = solver.solve(solution_vec, frame);
        const std::chrono::duration<double, std::milli> duration(std::chrono::steady_clock::now() - last);
        std::cout << "Processed in: " << duration.count() << " ms (" << solver.last_stats().iterations << " iterations)"
                  << std::endl;
        std::fwrite(&status, sizeof(status), 1, out);
        std::fwrite(solution_vec.data(), sizeof(double), nvoxel, out);
        if (flags & 2) solution_vec.clear();
    }
    std::fclose(out);
    return 0;
}
