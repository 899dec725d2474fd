// See sartsolver_b200.hpp. Error behaviour follows the reference: message on std::cerr, std::exit(1)
// (source/sartsolver.cpp:63-67, source/sartsolver_cuda.cpp:45-54).
#include "sartsolver_b200.hpp"

#include <cstdlib>
#include <iostream>

namespace {

void must(int rc) {
    if (rc != SART_OK) {
        std::cerr << sart_last_error() << std::endl;
        std::exit(1);
    }
}

}  // namespace

SARTSolverB200::SARTSolverB200(const float* matrix, std::size_t npixel_local, std::size_t nvoxel,
                               std::size_t offset_pixel, std::size_t npixel_total, const std::size_t* laplacian_index,
                               const float* laplacian_value, std::size_t laplacian_size, int rank, int nranks,
                               const std::uint8_t* comm_id, bool logarithmic, double ray_density_threshold,
                               double ray_length_threshold, double convolution_tolerance, double beta_laplace,
                               double relaxation, int max_iterations, int device)
    : npix(npixel_local), nvox(nvoxel) {
    sart_config cfg;
    must(sart_config_default(&cfg));
    cfg.npixel = npixel_total;
    cfg.nvoxel = nvoxel;
    cfg.rank = rank;
    cfg.nranks = nranks;
    cfg.device = device;
    cfg.logarithmic = logarithmic ? 1 : 0;
    cfg.ray_density_threshold = ray_density_threshold;
    cfg.ray_length_threshold = ray_length_threshold;
    cfg.conv_tolerance = convolution_tolerance;
    cfg.beta_laplace = beta_laplace;
    cfg.relaxation = relaxation;
    cfg.max_iterations = max_iterations;

    // the caller's partition must be the reference's (source/main.cpp:67-68): the engine shards by the same rule
    std::uint64_t off = 0, n = 0;
    must(sart_shard_range(npixel_total, nranks, rank, &off, &n));
    if (off != offset_pixel || n != npixel_local) {
        std::cerr << "Row block [" << offset_pixel << ", +" << npixel_local << ") is not rank " << rank << "'s share of "
                  << npixel_total << " pixels over " << nranks << " ranks." << std::endl;
        std::exit(1);
    }

    must(sart_create(&cfg, &engine));
    if (nranks > 1) {
        if (!comm_id) {
            std::cerr << "A communicator id (sart_comm_unique_id on rank 0, broadcast to all) is required when nranks > 1."
                      << std::endl;
            std::exit(1);
        }
        must(sart_comm_init(engine, comm_id));
    }
    must(sart_upload_rows(engine, 0, npixel_local, 0, nvoxel, matrix, nvoxel));
    static_assert(sizeof(std::size_t) == sizeof(std::uint64_t), "the reference's size_t Laplacian index is 64-bit");
    must(sart_set_laplacian_coo(engine, reinterpret_cast<const std::uint64_t*>(laplacian_index), laplacian_value,
                                laplacian_size));
    must(sart_finalize(engine));
}

SARTSolverB200::~SARTSolverB200() {
    if (engine) sart_destroy(engine);
}

namespace {
sart_config config_of(const sart_engine* e) {
    sart_config c;
    must(sart_get_config(e, &c));
    return c;
}
}  // namespace

double SARTSolverB200::get_ray_density_threshold() const { return config_of(engine).ray_density_threshold; }
double SARTSolverB200::get_ray_length_threshold() const { return config_of(engine).ray_length_threshold; }
double SARTSolverB200::get_convolution_tolerance() const { return config_of(engine).conv_tolerance; }
double SARTSolverB200::get_beta_laplace() const { return config_of(engine).beta_laplace; }
double SARTSolverB200::get_relaxation() const { return config_of(engine).relaxation; }
int SARTSolverB200::get_max_iterations() const { return config_of(engine).max_iterations; }

void SARTSolverB200::set_ray_density_threshold(double value) { must(sart_set_ray_density_threshold(engine, value)); }
void SARTSolverB200::set_ray_length_threshold(double value) { must(sart_set_ray_length_threshold(engine, value)); }
void SARTSolverB200::set_convolution_tolerance(double value) { must(sart_set_convolution_tolerance(engine, value)); }
void SARTSolverB200::set_beta_laplace(double value) { must(sart_set_beta_laplace(engine, value)); }
void SARTSolverB200::set_relaxation(double value) { must(sart_set_relaxation(engine, value)); }
void SARTSolverB200::set_max_iterations(int value) { must(sart_set_max_iterations(engine, value)); }

int SARTSolverB200::solve(std::vector<double>& solution, const std::vector<double>& measurement) const {
    if (!solution.empty() && solution.size() != nvox) {  // source/sartsolver_cuda.cpp:209-212
        std::cerr << "Solution vector must be empty or contain nvoxel elements." << std::endl;
        std::exit(1);
    }
    if (measurement.size() != npix) {
        std::cerr << "Measurement vector must contain this rank's npixel elements." << std::endl;
        std::exit(1);
    }
    const int has_guess = solution.empty() ? 0 : 1;
    if (!has_guess) solution.resize(nvox);  // source/sartsolver_cuda.cpp:166-167
    std::int32_t status = 0, iterations = 0;
    must(sart_solve(engine, measurement.data(), solution.data(), has_guess, &status, &iterations));
    return status;
}

namespace {
std::size_t frames_in(const std::vector<double>& frames, std::size_t npix) {
    if (npix == 0 || frames.size() % npix != 0) {
        std::cerr << "Frames vector must contain a whole number of this rank's npixel-element frames." << std::endl;
        std::exit(1);
    }
    return frames.size() / npix;
}
}  // namespace

std::vector<int> SARTSolverB200::solve_series(std::vector<double>& solution, const std::vector<double>& frames,
                                              std::vector<double>& solutions) const {
    if (!solution.empty() && solution.size() != nvox) {  // source/sartsolver_cuda.cpp:209-212
        std::cerr << "Solution vector must be empty or contain nvoxel elements." << std::endl;
        std::exit(1);
    }
    const std::size_t n = frames_in(frames, npix);
    solutions.resize(n * nvox);
    std::vector<std::int32_t> status(n, 0);
    must(sart_solve_series(engine, (std::int32_t)n, frames.data(), solution.empty() ? nullptr : solution.data(),
                           solutions.data(), status.data(), nullptr));
    if (n) solution.assign(solutions.end() - (std::ptrdiff_t)nvox, solutions.end());
    return std::vector<int>(status.begin(), status.end());
}

std::vector<int> SARTSolverB200::solve_independent(const std::vector<double>& frames, std::vector<double>& solutions) const {
    const std::size_t n = frames_in(frames, npix);
    solutions.resize(n * nvox);
    std::vector<std::int32_t> status(n, 0);
    must(sart_solve_batch(engine, (std::int32_t)n, frames.data(), solutions.data(), status.data(), nullptr));
    return std::vector<int>(status.begin(), status.end());
}

sart_stats SARTSolverB200::last_stats() const {
    sart_stats s;
    must(sart_get_stats(engine, &s));
    return s;
}
