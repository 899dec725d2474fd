// TEST INFRASTRUCTURE ONLY -- harness that runs the UNMODIFIED reference CPU solver as the oracle.
//
// This TU is compiled together with /root/reference/source/sartsolver.cpp (in place, never copied)
// into oracle/_ref/libsart_ref.so by oracle/Makefile. It supplies
//   * the member functions of RayTransferMatrix / LaplacianMatrix whose reference bodies live in
//     HDF5-dependent TUs (declarations: source/include/raytransfer.hpp:24-36,
//     source/include/laplacian.hpp:26-36) -- here they are filled from caller memory instead of files;
//   * MPI_Allreduce for the thread-per-rank shim declared in oracle/mpi_shim/mpi.h;
//   * a C entry point that plays the role of main()'s rank set-up and frame loop
//     (source/main.cpp:63-110, 127-140): row partition, Laplacian on rank 0 only for the linear
//     solver, setters, warm start between frames unless no_guess.
// The solver arithmetic itself (constructor sums, SARTSolverMPI::solve, LogSARTSolverMPI::solve)
// is the reference's own object code.
//
// Only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline / --impl reference) may load
// the resulting library, and only as the checker or the timed CPU baseline.

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "sartsolver.hpp"
#ifdef REF_WITH_CUDA
// second build of this harness (oracle/_ref/libsart_ref_gpu.so): the reference's own GPU solver, i.e. its
// source/sartsolver_cuda.cpp + source/cuda/sart_kernels.cu compiled for sm_100a in place, as the kernel to beat
#include "sartsolver_cuda.hpp"
#endif

namespace {

struct Team {
    int n = 1;
    std::vector<const void*> send;
    std::mutex mtx;
    std::condition_variable cv;
    int waiting = 0;
    unsigned long generation = 0;

    void barrier() {
        std::unique_lock<std::mutex> lock(mtx);
        const unsigned long gen = generation;
        if (++waiting == n) {
            waiting = 0;
            ++generation;
            cv.notify_all();
        } else {
            cv.wait(lock, [&] { return gen != generation; });
        }
    }
};

struct RankContext {
    Team* team = nullptr;
    int rank = 0;
    // data sources for the stand-in read_hdf5 members
    const float* matrix_src = nullptr;  // global matrix, row-major [npixel_total][nvoxel]
    const uint64_t* lap_idx = nullptr;
    const float* lap_val = nullptr;
    size_t lap_nnz = 0;
    long scalar_allreduce_calls = 0;  // MPI_Allreduce calls with count == 1 (see iterations below)
};

thread_local RankContext t_ctx;

template <typename T>
void reduce_in_rank_order(const Team& team, void* recvbuf, int count, MPI_Op op) {
    T* out = static_cast<T*>(recvbuf);
    for (int i = 0; i < count; ++i) {
        T acc = static_cast<const T*>(team.send[0])[i];
        for (int r = 1; r < team.n; ++r) {
            const T v = static_cast<const T*>(team.send[r])[i];
            acc = (op == MPI_MAX) ? std::max(acc, v) : acc + v;
        }
        out[i] = acc;
    }
}

}  // namespace

int MPI_Allreduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype datatype, MPI_Op op, MPI_Comm) {
    Team& team = *t_ctx.team;
    if (count == 1) ++t_ctx.scalar_allreduce_calls;
    team.send[t_ctx.rank] = sendbuf;
    team.barrier();
    if (datatype == MPI_DOUBLE) reduce_in_rank_order<double>(team, recvbuf, count, op);
    else reduce_in_rank_order<float>(team, recvbuf, count, op);
    team.barrier();
    return MPI_SUCCESS;
}

// ---- stand-ins for the HDF5-backed members (see header comment) -----------------------------------------------------

RayTransferMatrix::RayTransferMatrix(size_t npixel, size_t nvoxel, size_t offset_pixel)
    : offset_pix(offset_pixel), npix(npixel), nvox(nvoxel) {}

const std::vector<float>& RayTransferMatrix::matrix() const { return mat; }

int RayTransferMatrix::read_hdf5(const std::map<std::string, std::vector<std::string>>&, const std::string&) {
    const float* first = t_ctx.matrix_src + offset_pix * nvox;
    mat.assign(first, first + npix * nvox);
    return 0;
}

LaplacianMatrix::LaplacianMatrix(size_t nvoxel) : nvox(nvoxel) {}

const std::vector<float>& LaplacianMatrix::value() const { return val; }

const std::vector<size_t>& LaplacianMatrix::index() const { return index1d; }

int LaplacianMatrix::read_hdf5(const std::string&) {
    index1d.assign(t_ctx.lap_idx, t_ctx.lap_idx + t_ctx.lap_nnz);
    val.assign(t_ctx.lap_val, t_ctx.lap_val + t_ctx.lap_nnz);
    return 0;
}

// ---- C entry point --------------------------------------------------------------------------------------------------

extern "C" {

// params = {ray_density_threshold, ray_length_threshold, conv_tolerance, beta_laplace, relaxation}
// g        [nframes][npixel]   measurements (double), full (un-sharded) frames
// x        [nframes][nvoxel]   out: solution per frame; in: x[0] is the initial guess when has_guess != 0
// status, iterations [nframes] out. iterations = number of passes through the reference's iteration loop,
//          recovered from the shim (each pass makes exactly one scalar MPI_Allreduce: sartsolver.cpp:222 / :329,
//          plus one before the loop: :164 / :268).
// seconds  [2 + nframes] out: {constructor wall time, total solve wall time, solve wall time of each frame} as seen
//          by rank 0.
// Returns 0, or -1 for arguments the reference would std::exit(1) on (sartsolver.cpp:63-123).
int ref_sart_solve_frames(const float* H, uint64_t npixel, uint64_t nvoxel, const uint64_t* lap_idx,
                          const float* lap_val, uint64_t lap_nnz, const double* g, double* x, int nframes,
                          int has_guess, int no_guess, const double* params, int max_iterations, int logarithmic,
                          int nranks, int* status, int* iterations, double* seconds) {
    if (params[0] < 0 || params[1] < 0 || !(params[2] > 0) || params[3] < 0 || !(params[4] > 0) ||
        params[4] > 1.0 || max_iterations <= 0 || nranks < 1 || (uint64_t)nranks > npixel || nframes < 1)
        return -1;

    Team team;
    team.n = nranks;
    team.send.assign(nranks, nullptr);
    double t_ctor = 0, t_solve = 0;

    auto rank_main = [&](int rank) {
        t_ctx = RankContext();
        t_ctx.team = &team;
        t_ctx.rank = rank;
        t_ctx.matrix_src = H;
        t_ctx.lap_idx = lap_idx;
        t_ctx.lap_val = lap_val;
        t_ctx.lap_nnz = lap_nnz;

        // row partition exactly as source/main.cpp:67-68
        const size_t numproc = (size_t)nranks;
        const size_t offset_pixel = (size_t)rank * (npixel / numproc) + std::min((size_t)rank, (size_t)(npixel % numproc));
        const size_t npixel_local = ((size_t)rank < npixel % numproc) ? npixel / numproc + 1 : npixel / numproc;

        // source/main.cpp:73-74: only the root rank holds the Laplacian when the solver is linear
        LaplacianMatrix laplacian((lap_nnz == 0 || (rank && !logarithmic)) ? 0 : nvoxel);
        if (laplacian.nvoxel() > 0) laplacian.read_hdf5("");

        RayTransferMatrix raytransfer(npixel_local, nvoxel, offset_pixel);
        raytransfer.read_hdf5({}, "");

        team.barrier();
        auto t0 = std::chrono::steady_clock::now();
        BaseSARTSolverMPI* solver;
#ifdef REF_WITH_CUDA
        if (logarithmic) solver = new LogSARTSolverMPICuda(raytransfer, laplacian, rank);
        else solver = new SARTSolverMPICuda(raytransfer, laplacian, rank);
#else
        if (logarithmic) solver = new LogSARTSolverMPI(raytransfer, laplacian);
        else solver = new SARTSolverMPI(raytransfer, laplacian);
#endif
        solver->set_ray_density_threshold(params[0]);
        solver->set_ray_length_threshold(params[1]);
        solver->set_convolution_tolerance(params[2]);
        solver->set_beta_laplace(params[3]);
        solver->set_relaxation(params[4]);
        solver->set_max_iterations(max_iterations);
        team.barrier();
        auto t1 = std::chrono::steady_clock::now();

        std::vector<double> solution_vec, frame(npixel_local);
        if (has_guess) solution_vec.assign(x, x + nvoxel);
        double solve_seconds = 0;
        for (int f = 0; f < nframes; ++f) {
            const double* gf = g + (size_t)f * npixel + offset_pixel;
            frame.assign(gf, gf + npixel_local);
            t_ctx.scalar_allreduce_calls = 0;
            team.barrier();
            auto s0 = std::chrono::steady_clock::now();
            const int st = solver->solve(solution_vec, frame);
            team.barrier();
            auto s1 = std::chrono::steady_clock::now();
            solve_seconds += std::chrono::duration<double>(s1 - s0).count();
            if (rank == 0 && seconds) seconds[2 + f] = std::chrono::duration<double>(s1 - s0).count();
            if (rank == 0) {
                std::copy(solution_vec.begin(), solution_vec.end(), x + (size_t)f * nvoxel);
                status[f] = st;
#ifdef REF_WITH_CUDA
                // GPU path: two scalar reductions before the loop (max, sum g^2: sartsolver_cuda.cpp:149, 156), one per pass
                iterations[f] = (int)(t_ctx.scalar_allreduce_calls - 2);
#else
                iterations[f] = (int)(t_ctx.scalar_allreduce_calls - 1);
#endif
            }
            if (no_guess) solution_vec.clear();  // source/main.cpp:139
        }
        if (rank == 0) {
            t_ctor = std::chrono::duration<double>(t1 - t0).count();
            t_solve = solve_seconds;
        }
        delete solver;
    };

    std::vector<std::thread> threads;
    for (int r = 1; r < nranks; ++r) threads.emplace_back(rank_main, r);
    rank_main(0);
    for (auto& t : threads) t.join();
    if (seconds) {
        seconds[0] = t_ctor;
        seconds[1] = t_solve;
    }
    return 0;
}

// status constants of the reference (source/sartsolver.cpp:16-17), for the tests to assert on
int ref_status_success(void) { return BaseSARTSolverMPI::SUCCESS; }
int ref_status_max_iterations_exceeded(void) { return BaseSARTSolverMPI::MAX_ITERATIONS_EXCEEDED; }

}  // extern "C"
