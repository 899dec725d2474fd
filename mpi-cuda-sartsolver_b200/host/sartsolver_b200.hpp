// C++ host side above the C ABI: the reference's solver interface on top of libsart_b200.so.
//
// SARTSolverB200 has the public surface of the reference's BaseSARTSolverMPICuda family
// (source/include/sartsolver.hpp:33-51, source/include/sartsolver_cuda.hpp:29-35): the same status constants, the
// same six getters / setters with the same range checks and messages, and
//     int solve(std::vector<double>& solution, const std::vector<double>& measurement) const;
// with the same conventions: an empty `solution` asks for the default initial guess and is resized to nvoxel, a
// non-empty one must hold nvoxel values and is the warm start; `measurement` is this rank's slice of the frame
// (negative = saturated pixel); on return `solution` holds the un-normalised result. Like the reference, errors print
// to std::cerr and std::exit(1) (source/sartsolver_cuda.cpp:45-75) -- the C ABI underneath returns codes instead.
//
// The constructor is a template over the matrix containers so that the reference's own RayTransferMatrix /
// LaplacianMatrix (source/include/raytransfer.hpp, laplacian.hpp) can be passed as they are; any type with
// matrix().data(), npixel(), nvoxel(), offset_pixel() resp. index().data(), value().data(), size() works.
// No CUDA, NCCL or torch header is needed to compile against this file.
#pragma once

#include <cstddef>
#include <cstdint>
#include <vector>

#include "../../include/sart_b200.h"

class SARTSolverB200 {
public:
    static const int SUCCESS = SART_STATUS_SUCCESS;
    static const int MAX_ITERATIONS_EXCEEDED = SART_STATUS_MAX_ITERATIONS_EXCEEDED;

    // rank / nranks / npixel_total: the MPI rank, communicator size and total row count of source/main.cpp:63-68.
    // comm_id: the 128 bytes of sart_comm_unique_id() made on rank 0 and broadcast by the caller (e.g. MPI_Bcast);
    // ignored when nranks == 1.
    template <class RayTransfer, class Laplacian>
    SARTSolverB200(const RayTransfer& rtm, const Laplacian& lm, int rank, int nranks, std::size_t npixel_total,
                   const std::uint8_t* comm_id = nullptr, bool logarithmic = false,
                   double ray_density_threshold = 1.e-6, double ray_length_threshold = 1.e-6,
                   double convolution_tolerance = 1.e-5, double beta_laplace = 1.e-2, double relaxation = 1.,
                   int max_iterations = 2000, int device = -1)
        : SARTSolverB200(rtm.matrix().data(), rtm.npixel(), rtm.nvoxel(), rtm.offset_pixel(), npixel_total,
                         lm.size() ? lm.index().data() : nullptr, lm.size() ? lm.value().data() : nullptr, lm.size(),
                         rank, nranks, comm_id, logarithmic, ray_density_threshold, ray_length_threshold,
                         convolution_tolerance, beta_laplace, relaxation, max_iterations, device) {}

    // The same from raw arrays: matrix is this rank's row block, row-major [npixel_local][nvoxel].
    SARTSolverB200(const float* matrix, std::size_t npixel_local, std::size_t nvoxel, std::size_t offset_pixel,
                   std::size_t npixel_total, const std::size_t* laplacian_index, const float* laplacian_value,
                   std::size_t laplacian_size, int rank, int nranks, const std::uint8_t* comm_id, bool logarithmic,
                   double ray_density_threshold, double ray_length_threshold, double convolution_tolerance,
                   double beta_laplace, double relaxation, int max_iterations, int device);

    ~SARTSolverB200();
    SARTSolverB200(const SARTSolverB200&) = delete;
    SARTSolverB200& operator=(const SARTSolverB200&) = delete;

    double get_ray_density_threshold() const;
    double get_ray_length_threshold() const;
    double get_convolution_tolerance() const;
    double get_beta_laplace() const;
    double get_relaxation() const;
    int get_max_iterations() const;

    void set_ray_density_threshold(double value);
    void set_ray_length_threshold(double value);
    void set_convolution_tolerance(double value);
    void set_beta_laplace(double value);
    void set_relaxation(double value);
    void set_max_iterations(int value);

    int solve(std::vector<double>& solution, const std::vector<double>& measurement) const;

    // Not in the reference: a cached chunk of frames in one call (the loop around solve(), source/main.cpp:125-140).
    // frames = [n][npixel_local] row-major, solutions = [n][nvoxel] (resized); returns the n status codes.
    //   solve_series: frame k starts from the solution of frame k - 1; `solution` is the start of frame 0 (empty = default
    //                 initial guess) and holds the last frame's solution on return, as after n calls of solve().
    //   solve_independent: no frame starts from another one (--no_guess).
    // Same answers as solve() frame by frame within the parity tolerance, fewer passes over the matrix.
    std::vector<int> solve_series(std::vector<double>& solution, const std::vector<double>& frames,
                                  std::vector<double>& solutions) const;
    std::vector<int> solve_independent(const std::vector<double>& frames, std::vector<double>& solutions) const;

    // not in the reference: what the last solve did (iterations, device times, bytes moved)
    sart_stats last_stats() const;
    std::size_t nvoxel() const { return nvox; }
    std::size_t npixel_local() const { return npix; }

private:
    sart_engine* engine = nullptr;
    std::size_t npix = 0, nvox = 0;
};
