/* sart_b200.h -- C ABI of libsart_b200.so: a B200 (sm_100a) constrained-SART engine that is a drop-in for the
 * solver path of vsnever/mpi-cuda-sartsolver (SARTSolverMPICuda / LogSARTSolverMPICuda).
 *
 * One process drives one GPU and owns one row shard of the ray-transfer matrix, exactly like one MPI rank of the
 * reference (source/main.cpp:63-104). Plain pointers and sizes only: no C++ or torch types cross this boundary, no
 * exceptions, no exit(). Every function returns SART_OK (0) or an error code; the text is in sart_last_error()
 * (thread-local). File:line citations name the reference interface each entry point replaces
 * (paths relative to the reference repository root).
 *
 * Call order:  sart_create -> [sart_comm_init if nranks > 1] -> sart_upload_rows* | sart_fill_synthetic
 *              -> [sart_set_laplacian_coo] -> sart_finalize -> sart_solve* -> sart_destroy
 * With nranks > 1 the calls marked COLLECTIVE must be made by every rank, in the same order.
 */
#ifndef SART_B200_H
#define SART_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* error codes */
#define SART_OK 0
#define SART_ERR_INVALID_ARGUMENT 1 /* what the reference answers with a message + std::exit(1) */
#define SART_ERR_CUDA 2
#define SART_ERR_STATE 3 /* call order violated */
#define SART_ERR_NCCL 4
#define SART_ERR_NO_DEVICE 5

/* solve status: BaseSARTSolverMPI::SUCCESS / MAX_ITERATIONS_EXCEEDED, source/sartsolver.cpp:16-17 */
#define SART_STATUS_SUCCESS 0
#define SART_STATUS_MAX_ITERATIONS_EXCEEDED (-1)

/* Where the reference's CPU and GPU solvers differ observably (initial guess over saturated pixels,
 * source/sartsolver.cpp:153 vs source/cuda/sart_kernels.cu:33-34; 1e-7 floor on the starting solution,
 * source/sartsolver_cuda.cpp:180) the engine follows one or the other. Default: the CPU solver. */
#define SART_SEMANTICS_CPU 0
#define SART_SEMANTICS_GPU 1

/* How one iteration passes over the matrix (no reference counterpart; results agree within the parity tolerance).
 * SPLIT: forward projection and back projection as two kernels, two reads of the shard from HBM.
 * FUSED: one cluster kernel that keeps each row in shared memory between its forward and back projection, one read
 *        of the shard from HBM; needs nvoxel <= 229 376 (16 CTAs x 3584 float4 columns). AUTO picks FUSED for long rows. */
#define SART_SWEEP_AUTO 0
#define SART_SWEEP_SPLIT 1
#define SART_SWEEP_FUSED 2

/* How the per-iteration sum over the row shards travels between the GPUs of one box (replaces the reference's
 * D2H -> MPI_Allreduce -> H2D, source/sartsolver_cuda.cpp:242-244): PEER = one kernel per rank that adds all ranks'
 * partial vectors straight out of each other's memory over NVLink (CUDA IPC mappings made in sart_comm_init), takes the
 * stop/continue decision and updates the solution; NCCL = ncclAllReduce + separate kernels (always used by the
 * two-kernel sweep, and whenever the peers cannot be mapped or SART_EXCHANGE=nccl is set in the environment). */
#define SART_EXCHANGE_NONE 0
#define SART_EXCHANGE_NCCL 1
#define SART_EXCHANGE_PEER 2

#define SART_UNIQUE_ID_BYTES 128

typedef struct sart_engine sart_engine;

/* Construction arguments: BaseSARTSolverMPICuda::BaseSARTSolverMPICuda(rtm, lm, rank, thresholds..., comm),
 * source/include/sartsolver_cuda.hpp:31-35, plus the matrix extents the reference takes from RayTransferMatrix. */
typedef struct sart_config {
    uint64_t npixel;  /* TOTAL number of matrix rows over all ranks */
    uint64_t nvoxel;  /* number of matrix columns (every rank holds all of them, source/main.cpp:76) */
    int32_t rank;     /* this process's rank; it owns the rows sart_shard_range() gives */
    int32_t nranks;   /* number of ranks (= GPUs) */
    int32_t device;   /* CUDA ordinal, or -1 for rank % device_count (source/sartsolver_cuda.cpp:96-98) */
    int32_t logarithmic;      /* 0: SARTSolverMPICuda, 1: LogSARTSolverMPICuda */
    int32_t semantics;        /* SART_SEMANTICS_* */
    int32_t max_iterations;   /* default 2000 */
    int32_t fixed_iterations; /* 1: never test convergence, always run max_iterations (benchmark switch; not in the reference) */
    int32_t sweep_mode;       /* SART_SWEEP_*: how an iteration passes over the matrix (default AUTO) */
    double ray_density_threshold; /* default 1e-6 */
    double ray_length_threshold;  /* default 1e-6 */
    double conv_tolerance;        /* default 1e-5 */
    double beta_laplace;          /* default 1e-2 (the constructor default, source/include/sartsolver.hpp:55) */
    double relaxation;            /* default 1 */
} sart_config;

/* Per-solve measurements kept by the engine (device time from CUDA events on the engine's own stream). */
typedef struct sart_stats {
    int32_t status;          /* of the last solve */
    int32_t iterations;      /* passes through the reference's iteration loop the last solve made */
    int32_t sweeps;          /* full passes over the shard: iterations + 1 forward and iterations back projections, + 1 for an initial guess */
    int32_t gpu_launches;    /* kernels of this library launched by the last solve */
    double solve_ms;         /* device time of the whole solve, copies in and out included */
    double iterate_ms;       /* device time of the iteration loop only (first sweep .. last update) */
    double fp_ms_avg;        /* mean duration of one forward-projection launch */
    double bp_ms_avg;        /* mean duration of one back-projection launch */
    int32_t fp_launches;     /* launches behind fp_ms_avg */
    int32_t bp_launches;     /* launches behind bp_ms_avg */
    uint64_t matrix_bytes;   /* npixel_local * nvoxel * 4: algorithmic bytes of ONE sweep of this shard */
    uint64_t h2d_bytes;      /* host->device bytes moved by the last solve */
    uint64_t d2h_bytes;      /* device->host bytes moved by the last solve */
    double fused_ms_avg;     /* mean duration of one fused forward+back projection sweep (SART_SWEEP_FUSED); when the
                                solve ran as one launch (solve_launches == 1): that launch / fused_launches, i.e.
                                including what happens between the sweeps */
    int32_t fused_launches;  /* fused sweeps behind fused_ms_avg (one launch each unless solve_launches == 1) */
    int32_t sweep_mode;      /* SART_SWEEP_SPLIT or SART_SWEEP_FUSED: what the engine runs */
    int32_t exchange_mode;   /* SART_EXCHANGE_*: how the per-iteration reduction over the ranks travels */
    int32_t solve_launches;  /* 1: all sweeps of the solve ran inside ONE cooperative launch (fused_solve_kernel), 0: a launch per sweep */
    double tail_ms_avg;      /* fused sweep: mean device time between two sweeps (reduction over clusters and ranks, decision, update) */
    int32_t fused_cluster_size;  /* CTAs per thread-block cluster of the fused sweep (0: two-kernel sweep) */
    int32_t fused_cluster_count; /* clusters resident at once = its grid / cluster size */
    double tail_phase_us[5]; /* whole-solve kernel, mean per sweep as CTA 0 saw it: wait at the first grid barrier (the
                                slowest cluster finishing its rows), sum over the clusters' partials, flag round trip
                                with the other ranks, peer reads + update, second grid barrier */
} sart_stats;

typedef struct sart_device_info {
    char name[256];
    int32_t sm_count;
    int32_t cc_major, cc_minor;
    int32_t l2_bytes;
    int32_t reserved;
    uint64_t total_mem, free_mem;
} sart_device_info;

const char* sart_last_error(void);

/* Defaults of the reference constructor (source/include/sartsolver_cuda.hpp:31-35), rank 0 of 1, device -1. */
int sart_config_default(sart_config* cfg);

/* The reference's row partition: offset_pixel / npixel_local of source/main.cpp:67-68. */
int sart_shard_range(uint64_t npixel, int32_t nranks, int32_t rank, uint64_t* row_offset, uint64_t* nrows);

int sart_device_query(int32_t device, sart_device_info* info);

/* Replaces the constructor's device selection and allocations, source/sartsolver_cuda.cpp:78-126. Validates the
 * six solver parameters like the setters below. Allocates the zero-filled shard [nrows][ld] fp32, ld = nvoxel
 * rounded up to 32 floats so that every row starts on a 128-byte boundary. */
int sart_create(const sart_config* cfg, sart_engine** out);
int sart_destroy(sart_engine* e);

/* Replaces MPI_Init / MPI_COMM_WORLD of source/main.cpp:63-65 for the engine's own reductions: rank 0 makes an
 * id, the caller ships the 128 bytes to every rank by any means, then every rank calls sart_comm_init
 * (COLLECTIVE; NCCL over NVLink, loaded with dlopen so a single-GPU caller needs no NCCL at all). */
int sart_comm_unique_id(uint8_t id[SART_UNIQUE_ID_BYTES]);
int sart_comm_init(sart_engine* e, const uint8_t id[SART_UNIQUE_ID_BYTES]);

/* Replaces RayTransferMatrix::read_hdf5's row-at-a-time fill of the host matrix + the constructor's one
 * cudaMemcpy (source/raytransfer.cpp:92-110, source/sartsolver_cuda.cpp:104-106): copies the block
 * rows [local_row0, +nrows) x columns [col0, +ncols) from src (row-major, leading dimension src_ld floats, pinned
 * or pageable) straight into the shard. Arbitrary sub-blocks, because files come per camera x voxel segment.
 * Returns after the copy has completed. */
int sart_upload_rows(sart_engine* e, uint64_t local_row0, uint64_t nrows, uint64_t col0, uint64_t ncols,
                     const float* src, uint64_t src_ld);
/* The same from a file, without a host-resident copy of the matrix: the block rows [local_row0, +nrows) x columns
 * [col0, +ncols) is streamed from a raw little-endian fp32 array (row-major, file_ld floats per row, first element at
 * byte file_offset of `path`) through a ring of three pinned staging buffers (SART_STAGING_MB each, default 16): the
 * pread of slab k+1 overlaps the asynchronous copy of slab k, and host memory stays O(staging) however large the shard
 * is (the reference holds the whole rank-local matrix in host RAM: source/include/raytransfer.hpp:20, and copies it in
 * one blocking cudaMemcpy: source/sartsolver_cuda.cpp:104-106). Returns when the block is on the device. */
int sart_load_rows_file(sart_engine* e, const char* path, uint64_t file_offset, uint64_t file_ld, uint64_t local_row0,
                        uint64_t nrows, uint64_t col0, uint64_t ncols);
/* Test/bench helpers without a reference counterpart: counter-based synthetic matrix keyed by GLOBAL row and
 * column (shard-invariant), and read-back of shard rows (dst leading dimension nvoxel). */
int sart_fill_synthetic(sart_engine* e, uint64_t seed);
/* The same with a choice of matrix: DENSE = the fill above (every element non-zero, 1 in 64 a strong "direct" one: the
 * reflected-radiation matrices of BASELINE configs[1..4]); LOS = one straight ray per row through the nx x nz voxel grid
 * over a weak dense background, which conditions the problem like a real tomography set-up -- warm-started frames then
 * take tens of iterations and converge to the emission they were projected from, which the dense kind does not. */
#define SART_SYNTHETIC_DENSE 0
#define SART_SYNTHETIC_LOS 1
int sart_fill_synthetic_kind(sart_engine* e, uint64_t seed, int32_t kind);
int sart_download_rows(sart_engine* e, uint64_t local_row0, uint64_t nrows, float* dst);

/* Replaces LaplacianMatrix (source/include/laplacian.hpp:15-19) as the solver sees it: COO entries with the
 * flattened index i * nvoxel + j and fp32 values, any order (the reference sorts on load, source/laplacian.cpp:67-82).
 * nnz == 0 clears it. With nranks > 1 the reference gives the matrix to rank 0 only (source/main.cpp:73-74);
 * here sart_finalize broadcasts rank 0's copy so that every GPU can apply the penalty. */
int sart_set_laplacian_coo(sart_engine* e, const uint64_t* flat_index, const float* value, uint64_t nnz);

/* Replaces the base-class constructor's two host sweeps, source/sartsolver.cpp:35-56: ray_length (row sums of the
 * local rows) and ray_density (column sums over ALL ranks' rows), accumulated in double on the device.
 * COLLECTIVE. Must be called again after the matrix or the Laplacian changes. */
int sart_finalize(sart_engine* e);
int sart_get_ray_sums(sart_engine* e, double* ray_density /* [nvoxel] or NULL */, double* ray_length /* [nrows] or NULL */);

/* Setters / getters with the reference's range checks, source/sartsolver.cpp:61-123
 * (thresholds >= 0, tolerance > 0, beta >= 0, relaxation in (0, 1], max_iterations > 0). */
int sart_set_ray_density_threshold(sart_engine* e, double v);
int sart_set_ray_length_threshold(sart_engine* e, double v);
int sart_set_convolution_tolerance(sart_engine* e, double v);
int sart_set_beta_laplace(sart_engine* e, double v);
int sart_set_relaxation(sart_engine* e, double v);
int sart_set_max_iterations(sart_engine* e, int32_t v);
int sart_set_fixed_iterations(sart_engine* e, int32_t on);
int sart_get_config(const sart_engine* e, sart_config* out);

/* Replaces SARTSolverMPICuda::solve / LogSARTSolverMPICuda::solve, source/sartsolver_cuda.cpp:197-274, 277-354
 * (declared source/include/sartsolver.hpp:50-51). COLLECTIVE.
 *   measurement  [nrows]  this rank's slice of the frame (host, double); negative = saturated pixel
 *   solution     [nvoxel] host, double; in: starting solution when has_guess != 0; out: the solution
 *   has_guess    0 = the reference's "solution.empty()": start from the default initial guess
 *   status       SART_STATUS_*;  iterations: passes through the iteration loop (the reference does not report it) */
int sart_solve(sart_engine* e, const double* measurement, double* solution, int32_t has_guess, int32_t* status,
               int32_t* iterations);

/* A warm-started time series in one call: the frame loop of source/main.cpp:125-140 (without --no_guess), where every
 * frame starts from the previous frame's solution. COLLECTIVE.
 *   measurements [nframes][nrows] (host, double), guess [nvoxel] or NULL (start of frame 0; NULL = default initial
 *   guess), solutions [nframes][nvoxel] (host, double, out), status / iterations [nframes] or NULL.
 * Every frame gets what the sequence sart_solve(frame 0), sart_solve(frame 1, has_guess = 1), ... gives, within the
 * parity tolerance. Knowing the next frame in advance lets the engine save a pass over the matrix per frame: the sweep
 * predicted to be a frame's last (the previous frame's iteration count, once two consecutive frames agreed to within two
 * iterations) back-projects the next frame's measurement, so a
 * frame that stops as predicted hands its successor the first iteration (linear solver, SART_SEMANTICS_CPU, whole-solve
 * kernel; otherwise the frames are simply solved one after the other). After the call sart_get_stats reports for the
 * whole series: sweeps = passes over the matrix actually made, fp_launches = frames that received their first iteration
 * from their predecessor, bp_launches = sweeps that had to be repeated because a frame did not stop where predicted. */
int sart_solve_series(sart_engine* e, int32_t nframes, const double* measurements, const double* guess, double* solutions,
                      int32_t* status, int32_t* iterations);

/* Several frames that do NOT start from each other's solutions -- the reference's --no_guess loop, source/main.cpp:131-140
 * with solution_vec.clear() after every frame -- in one call. Such frames are independent right-hand sides of the same
 * matrix, so the engine solves them eight at a time (SART_BATCH_FRAMES=4: four): every pass over the matrix (forward
 * projection, back projection) serves the whole group, 2 passes per iteration for eight frames instead of one fused pass per frame (linear solver,
 * log solver, any number of ranks; a last single frame is solved on its own, same results). COLLECTIVE.
 *   measurements [nframes][nrows], solutions [nframes][nvoxel] (host, double); status, iterations [nframes] or NULL.
 * Every frame gets exactly what sart_solve(has_guess = 0) would give it, within the parity tolerance. */
int sart_solve_batch(sart_engine* e, int32_t nframes, const double* measurements, double* solutions, int32_t* status,
                     int32_t* iterations);

/* The same solve split at the host<->device copies, for callers that keep frames and solutions on the device
 * (time series with warm start: the previous solution never leaves HBM). sart_solve == stage_measurement
 * [+ stage_guess] + solve_resident + fetch_solution. */
int sart_stage_measurement(sart_engine* e, const double* measurement);
int sart_stage_guess(sart_engine* e, const double* solution);
int sart_solve_resident(sart_engine* e, int32_t has_guess, int32_t* status, int32_t* iterations); /* COLLECTIVE */
int sart_fetch_solution(sart_engine* e, double* solution);

/* Synthetic frames without a reference counterpart: out[i] = sum_j H[i][j] * x[j] over the local rows, fp64
 * accumulation on the device (what a camera would measure for emission x). */
int sart_forward_project(sart_engine* e, const double* x, double* out);

/* The two production sweep kernels on their own, for known-answer tests of the kernels (no reference counterpart as
 * entry points; the arithmetic is cublasSgemv's, source/sartsolver_cuda.cpp:248-249, and the main loop of
 * PropagateKernel, source/cuda/sart_kernels.cu:86-101):
 *   sart_kernel_forward   fit[i] = sum_j H[i][j] * (float)x[j]     over the local rows (fp32 kernel, fp64 host I/O)
 *   sart_kernel_backward  s[j]   = sum_i H[i][j] * (float)w[i]     over the local rows, NOT reduced over ranks
 * They use the engine's per-solve work buffers: do not interleave them with a solve on another thread. */
int sart_kernel_forward(sart_engine* e, const double* x, double* fit);
int sart_kernel_backward(sart_engine* e, const double* w, double* s);

/* Back projection without any device allocation -- this entry point has no counterpart in the reference.
 * For hosts that reserved all GPU memory for their own arena (acquisition loops with a fixed budget, frameworks on a
 * caching allocator, capped partitions): there a library-side cudaMalloc can fail, and it stalls every stream of the
 * device when it does not. Such a caller cannot create an engine (sart_create allocates); it passes its own DEVICE
 * buffers and gets the same column reduction the engine's back projection performs:
 *     d_out[j] = sum_i d_matrix[i * ld + j] * d_weights[i]      0 <= i < nrows, 0 <= j < ncols   (fp64 result)
 * d_matrix: row-major fp32, 16-byte aligned, leading dimension ld floats, ld % 4 == 0, ld >= ncols (the padding
 * columns are read and ignored); d_weights fp32 [nrows]; d_out fp64 [ncols]; cuda_stream: a cudaStream_t (NULL = default
 * stream) on the CURRENT device, to which all three buffers must belong. The two launches are asynchronous on it.
 * Scratch: a 16 MiB array built into the library, one per GPU, resident on every GPU that loads the library until
 * exit. Limit: ld <= SART_NOALLOC_MAX_LD; larger is refused (SART_ERR_INVALID_ARGUMENT). The array is shared by all
 * calls and keeps the previous call's contents -- nothing in it needs clearing, a call reads only what it wrote -- but
 * calls that overlap on one GPU race on it: serialise them (one stream, or events). The first call on each device looks
 * up the array's address; make that call before starting any stream capture. */
#define SART_NOALLOC_SCRATCH_BYTES (16u << 20)
#define SART_NOALLOC_MAX_LD (4u << 20)
int sart_back_project_noalloc(const float* d_matrix, uint64_t nrows, uint64_t ncols, uint64_t ld,
                              const float* d_weights, double* d_out, void* cuda_stream);

int sart_get_stats(const sart_engine* e, sart_stats* out);

#ifdef __cplusplus
}
#endif
#endif /* SART_B200_H */
