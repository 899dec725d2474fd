/* TEST INFRASTRUCTURE ONLY -- CPU restatement (plain C, double precision) of the reference's
 * constrained-SART solve path. See sart_oracle.c for the per-function reference citations.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use it, and only as
 * the checker; the product (libsart_b200.so) never links or loads it.
 */
#ifndef SART_ORACLE_H
#define SART_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORACLE_SUCCESS 0
#define ORACLE_MAX_ITERATIONS_EXCEEDED (-1)

typedef struct {
    double ray_density_threshold; /* h_d */
    double ray_length_threshold;  /* h_l */
    double conv_tolerance;        /* C_tol */
    double beta_laplace;
    double relaxation;
    int max_iterations;
    int logarithmic;         /* 0: SARTSolverMPI, 1: LogSARTSolverMPI */
    int gpu_semantics;       /* 0: CPU path (primary oracle); 1: the two observable deviations of the reference GPU path */
    int fixed_iterations;    /* 1: never test convergence (bench/parity switch; not in the reference) */
    int nranks;              /* emulate the row partition: partial sums per rank block, added in rank order */
} oracle_params;

void oracle_shard_range(uint64_t npixel, int nranks, int rank, uint64_t* offset, uint64_t* nrows);

void oracle_ray_sums(const float* H, uint64_t npixel, uint64_t nvoxel, int nranks, double* ray_density,
                     double* ray_length);

int oracle_solve(const float* H, uint64_t npixel, uint64_t nvoxel, const uint64_t* lap_idx, const float* lap_val,
                 uint64_t lap_nnz, const double* ray_density, const double* ray_length, const double* g, double* x,
                 int has_guess, const oracle_params* prm, int* iterations, double* conv_history);

/* synthetic inputs (mirrors the device generator in the product; compared bit-for-bit by the tests) */
uint64_t oracle_splitmix64(uint64_t z);
int oracle_synth_exponent(uint64_t nvoxel);
float oracle_synth_element(uint64_t seed, uint64_t flat_index, int exponent);
void oracle_fill_synthetic(float* H, uint64_t row0, uint64_t nrows, uint64_t nvoxel, uint64_t seed);
void oracle_grid_shape(uint64_t nvoxel, uint64_t* nx, uint64_t* nz);
int oracle_los_exponent(uint64_t nvoxel);
void oracle_fill_los(float* H, uint64_t row0, uint64_t nrows, uint64_t nvoxel, uint64_t seed);
void oracle_forward(const float* H, uint64_t npixel, uint64_t nvoxel, const double* x, double* out);

#ifdef __cplusplus
}
#endif
#endif
