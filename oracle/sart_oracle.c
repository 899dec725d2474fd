/* TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's constrained-SART solve path.
 *
 * Plain C, double accumulation over the fp32 matrix, exactly as the reference's CPU solver does
 * (/root/reference/source/sartsolver.cpp). It is an independent second statement of the algorithm:
 * the first is the reference's own object code (oracle/_ref/libsart_ref.so, built by oracle/Makefile
 * from the reference sources where they lie). tests/test_oracle.py pins this file against that
 * library and against the golden vectors under tests/golden/ that the library produced.
 *
 * Loop nests are re-arranged to walk the row-major matrix along rows (the reference walks columns,
 * sartsolver.cpp:149-157, 191-205), but every per-voxel sum still adds its terms in ascending pixel
 * order and every per-pixel sum in ascending voxel order, so with nranks == 1 the results are
 * bit-identical to the reference's, and with nranks > 1 they reproduce the rank-ordered
 * MPI_Allreduce of the thread shim.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use this file, and only
 * as the checker or the timed baseline; the product never links, loads or calls it.
 */
#include "sart_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#define EPSILON_LOG_CPU 1.e-100 /* sartsolver.cpp:14 */
#define EPSILON_GPU 1.e-7       /* sartsolver_cuda.cpp:15-17, in units of max(g) */

/* source/main.cpp:67-68 -- contiguous row blocks, the first npixel % nranks ranks get one extra row */
void oracle_shard_range(uint64_t npixel, int nranks, int rank, uint64_t* offset, uint64_t* nrows) {
    const uint64_t n = (uint64_t)nranks, r = (uint64_t)rank;
    const uint64_t base = npixel / n, rem = npixel % n;
    *offset = r * base + (r < rem ? r : rem);
    *nrows = (r < rem) ? base + 1 : base;
}

/* sartsolver.cpp:35-56 -- ray_density_j = sum_i H_ij over ALL rows (rank partials added in rank order,
 * :47), ray_length_i = sum_j H_ij; both accumulated in double. */
void oracle_ray_sums(const float* H, uint64_t npixel, uint64_t nvoxel, int nranks, double* ray_density,
                     double* ray_length) {
    double* part = (double*)malloc(nvoxel * sizeof(double));
    memset(ray_density, 0, nvoxel * sizeof(double));
    for (int r = 0; r < nranks; ++r) {
        uint64_t off, n;
        oracle_shard_range(npixel, nranks, r, &off, &n);
        memset(part, 0, nvoxel * sizeof(double));
        for (uint64_t i = off; i < off + n; ++i) {
            const float* row = H + i * nvoxel;
            for (uint64_t j = 0; j < nvoxel; ++j) part[j] += (double)row[j];
        }
        for (uint64_t j = 0; j < nvoxel; ++j) ray_density[j] += part[j];
    }
    free(part);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < (int64_t)npixel; ++i) {
        const float* row = H + (uint64_t)i * nvoxel;
        double res = 0;
        for (uint64_t j = 0; j < nvoxel; ++j) res += (double)row[j];
        ray_length[i] = res;
    }
}

/* fitted_i = sum_j H_ij x_j (sartsolver.cpp:166-173, 213-220) and the rank-ordered sum of squares (:219-222) */
static double forward_and_square(const float* H, uint64_t npixel, uint64_t nvoxel, int nranks, const double* x,
                                 double* fitted) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < (int64_t)npixel; ++i) {
        const float* row = H + (uint64_t)i * nvoxel;
        double res = 0;
        for (uint64_t j = 0; j < nvoxel; ++j) res += (double)row[j] * x[j];
        fitted[i] = res;
    }
    double total = 0;
    for (int r = 0; r < nranks; ++r) {
        uint64_t off, n;
        oracle_shard_range(npixel, nranks, r, &off, &n);
        double loc = 0;
        for (uint64_t i = off; i < off + n; ++i) loc += fitted[i] * fitted[i];
        total = (r == 0) ? loc : total + loc;
    }
    return total;
}

/* Column sums of rows [off, off+n) weighted per row: acc_j = sum_i (H_ij / div_i) * w_i in ascending i.
 * div == NULL means no division (initial guess, sartsolver.cpp:152-154); rows with skip_i != 0 are left out
 * (:196). Threads split the voxel range so the per-voxel order of additions is unchanged. */
static void weighted_column_sums(const float* H, uint64_t nvoxel, uint64_t off, uint64_t n, const double* div,
                                 const double* w, const unsigned char* skip, double* acc) {
    const int64_t chunk = 2048;
    const int64_t nchunks = ((int64_t)nvoxel + chunk - 1) / chunk;
#pragma omp parallel for schedule(static)
    for (int64_t c = 0; c < nchunks; ++c) {
        const uint64_t j0 = (uint64_t)(c * chunk);
        const uint64_t j1 = (j0 + (uint64_t)chunk < nvoxel) ? j0 + (uint64_t)chunk : nvoxel;
        for (uint64_t j = j0; j < j1; ++j) acc[j] = 0;
        for (uint64_t i = off; i < off + n; ++i) {
            if (skip && skip[i]) continue;
            const float* row = H + i * nvoxel;
            const double wi = w[i];
            if (div) {
                const double d = div[i];
                for (uint64_t j = j0; j < j1; ++j) acc[j] += ((double)row[j] / d) * wi;
            } else {
                for (uint64_t j = j0; j < j1; ++j) acc[j] += (double)row[j] * wi;
            }
        }
    }
}

/* SARTSolverMPI::solve (sartsolver.cpp:133-232) and LogSARTSolverMPI::solve (:235-339).
 * x: in = initial guess when has_guess, out = solution. Returns the reference's status code. */
int oracle_solve(const float* H, uint64_t npixel, uint64_t nvoxel, const uint64_t* lap_idx, const float* lap_val,
                 uint64_t lap_nnz, const double* ray_density, const double* ray_length, const double* g, double* x,
                 int has_guess, const oracle_params* prm, int* iterations, double* conv_history) {
    const int nranks = prm->nranks > 0 ? prm->nranks : 1;
    const double h_d = prm->ray_density_threshold, h_l = prm->ray_length_threshold;
    const double beta = prm->beta_laplace, relax = prm->relaxation;
    double* fitted = (double*)malloc(npixel * sizeof(double));
    double* resid = (double*)malloc(npixel * sizeof(double));
    double* part = (double*)malloc(2 * nvoxel * sizeof(double));
    double* total = (double*)malloc(2 * nvoxel * sizeof(double));
    double* pen = (double*)malloc(nvoxel * sizeof(double));
    unsigned char* skip = (unsigned char*)malloc(npixel);
    int status = ORACLE_MAX_ITERATIONS_EXCEEDED;
    int iters_done = 0;

    double norm = g[0];
    for (uint64_t i = 1; i < npixel; ++i) norm = g[i] > norm ? g[i] : norm; /* sartsolver_cuda.cpp:147-149 */
    const double eps_log = prm->gpu_semantics ? EPSILON_GPU * norm : EPSILON_LOG_CPU;

    if (!has_guess) {
        /* sartsolver.cpp:144-159. CPU path: negative g enters the sum; GPU path: clipped to 0 (sart_kernels.cu:33-34) */
        double* gw = resid;
        for (uint64_t i = 0; i < npixel; ++i) gw[i] = (prm->gpu_semantics && !(g[i] > 0)) ? 0.0 : g[i];
        for (uint64_t j = 0; j < nvoxel; ++j) x[j] = 0;
        for (int r = 0; r < nranks; ++r) {
            uint64_t off, n;
            oracle_shard_range(npixel, nranks, r, &off, &n);
            weighted_column_sums(H, nvoxel, off, n, NULL, gw, NULL, part);
            for (uint64_t j = 0; j < nvoxel; ++j) {
                const double loc = (ray_density[j] > h_d) ? part[j] / ray_density[j] : 0.0;
                x[j] = (r == 0) ? loc : x[j] + loc;
            }
        }
    }
    if (prm->gpu_semantics) { /* sartsolver_cuda.cpp:180 (both solvers, warm starts too) */
        for (uint64_t j = 0; j < nvoxel; ++j)
            if (x[j] < EPSILON_GPU * norm) x[j] = EPSILON_GPU * norm;
    } else if (prm->logarithmic) { /* sartsolver.cpp:263 */
        for (uint64_t j = 0; j < nvoxel; ++j)
            if (x[j] < EPSILON_LOG_CPU) x[j] = EPSILON_LOG_CPU;
    }

    /* sartsolver.cpp:161-164 -- saturated (negative) pixels do not count; rank partials in rank order */
    double m2 = 0;
    for (int r = 0; r < nranks; ++r) {
        uint64_t off, n;
        oracle_shard_range(npixel, nranks, r, &off, &n);
        double loc = 0;
        for (uint64_t i = off; i < off + n; ++i) loc += (g[i] > 0) ? g[i] * g[i] : 0;
        m2 = (r == 0) ? loc : m2 + loc;
    }

    forward_and_square(H, npixel, nvoxel, nranks, x, fitted); /* :166-173 */

    for (uint64_t i = 0; i < npixel; ++i) skip[i] = !((ray_length[i] > h_l) && (g[i] >= 0)); /* :196 */

    double conv_prev = 0;
    for (int iter = 0; iter < prm->max_iterations; ++iter) {
        /* :183-189 / :287-293 -- penalty uses the solution BEFORE this iteration's update */
        for (uint64_t j = 0; j < nvoxel; ++j) pen[j] = 0;
        for (uint64_t k = 0; k < lap_nnz; ++k) {
            const uint64_t iv = lap_idx[k] / nvoxel, jv = lap_idx[k] % nvoxel;
            /* CPU solver: log of the solution in caller units (:292). GPU solver: __logf of the iterate it keeps
             * divided by max(g) (sart_kernels.cu:200 on sartsolver_cuda.cpp:177's normalised vector). */
            const double s = prm->logarithmic ? log(prm->gpu_semantics ? x[jv] / norm : x[jv]) : x[jv];
            pen[iv] += beta * (double)lap_val[k] * s;
        }

        if (!prm->logarithmic) {
            for (uint64_t i = 0; i < npixel; ++i) resid[i] = g[i] - fitted[i];
            for (int r = 0; r < nranks; ++r) {
                uint64_t off, n;
                oracle_shard_range(npixel, nranks, r, &off, &n);
                weighted_column_sums(H, nvoxel, off, n, ray_length, resid, skip, part); /* :191-200 */
                for (uint64_t j = 0; j < nvoxel; ++j) {
                    double diff_j = 0;
                    if (ray_density[j] > h_d) diff_j = part[j] * (relax / ray_density[j]); /* :201 */
                    /* :204 + main.cpp:73 -- only rank 0 holds the Laplacian, so it is subtracted once */
                    const double loc = diff_j - (r == 0 ? pen[j] : 0.0);
                    total[j] = (r == 0) ? loc : total[j] + loc; /* :206 */
                }
            }
            for (uint64_t j = 0; j < nvoxel; ++j) { /* :208-209 */
                x[j] += total[j];
                if (signbit(x[j])) x[j] = 0;
            }
        } else {
            for (int r = 0; r < nranks; ++r) { /* :295-312 */
                uint64_t off, n;
                oracle_shard_range(npixel, nranks, r, &off, &n);
                weighted_column_sums(H, nvoxel, off, n, ray_length, g, skip, part);
                weighted_column_sums(H, nvoxel, off, n, ray_length, fitted, skip, part + nvoxel);
                for (uint64_t j = 0; j < nvoxel; ++j) {
                    const int on = ray_density[j] > h_d;
                    const double o = on ? part[j] : 0.0, f = on ? part[nvoxel + j] : 0.0;
                    total[j] = (r == 0) ? o : total[j] + o;
                    total[nvoxel + j] = (r == 0) ? f : total[nvoxel + j] + f;
                }
            }
            for (uint64_t j = 0; j < nvoxel; ++j) /* :314-316 */
                x[j] *= pow((total[j] + eps_log) / (total[nvoxel + j] + eps_log), relax) * exp(-pen[j]);
        }

        const double f2 = forward_and_square(H, npixel, nvoxel, nranks, x, fitted); /* :211-222 */
        const double conv = (m2 - f2) / m2;                                          /* :224 */
        if (conv_history) conv_history[iter] = conv;
        iters_done = iter + 1;
        if (!prm->fixed_iterations && iter && (fabs(conv - conv_prev) < prm->conv_tolerance)) { /* :226 */
            status = ORACLE_SUCCESS;
            break;
        }
        conv_prev = conv;
    }

    if (iterations) *iterations = iters_done;
    free(fitted);
    free(resid);
    free(part);
    free(total);
    free(pen);
    free(skip);
    return status;
}

/* ---- synthetic ray-transfer matrix (counter-based, shard-invariant; mirrors the product's device generator) ---------
 * One splitmix64 draw per element, keyed by the GLOBAL flattened index i*nvoxel + j:
 *   background ("reflections"): 16-bit, never zero           bits 63..48 | 1
 *   line of sight (1 element in 64): extra 23-bit term       bits 28..6  when bits 5..0 == 0
 * The integer q < 2^24 is exact in fp32 and is scaled by one power of two chosen so that the mean row sum is ~1,
 * the scale of a physical ray-transfer matrix (row sums are ray lengths in metres). */
uint64_t oracle_splitmix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

int oracle_synth_exponent(uint64_t nvoxel) {
    /* mean q = 32768 + 2^22/64 = 98304 */
    return (int)lround(log2((double)nvoxel * 98304.0));
}

float oracle_synth_element(uint64_t seed, uint64_t flat_index, int exponent) {
    const uint64_t h = oracle_splitmix64(seed ^ (flat_index * 0xD1342543DE82EF95ULL));
    uint32_t q = (uint32_t)(h >> 48) | 1u;
    if ((h & 63u) == 0) q += (uint32_t)((h >> 6) & 0x7FFFFFu);
    return ldexpf((float)q, -exponent);
}

void oracle_fill_synthetic(float* H, uint64_t row0, uint64_t nrows, uint64_t nvoxel, uint64_t seed) {
    const int e = oracle_synth_exponent(nvoxel);
#pragma omp parallel for schedule(static)
    for (int64_t r = 0; r < (int64_t)nrows; ++r) {
        const uint64_t i = row0 + (uint64_t)r;
        float* row = H + (uint64_t)r * nvoxel;
        for (uint64_t j = 0; j < nvoxel; ++j) row[j] = oracle_synth_element(seed, i * nvoxel + j, e);
    }
}

/* The "line of sight" kind of the product's generator (SART_SYNTHETIC_LOS, kernels.cuh): one straight ray per row through
 * the nx x nz voxel grid over a weak 8-bit background, integers only. */
void oracle_grid_shape(uint64_t nvoxel, uint64_t* nx, uint64_t* nz) {
    uint64_t n = (uint64_t)floor(sqrt((double)nvoxel));
    while (n > 1 && nvoxel % n) --n;
    *nx = n;
    *nz = nvoxel / n;
}

int oracle_los_exponent(uint64_t nvoxel) {
    return (int)lround(log2(1024.0 * 2430.0 * sqrt((double)nvoxel) + 128.0 * (double)nvoxel));
}

void oracle_fill_los(float* H, uint64_t row0, uint64_t nrows, uint64_t nvoxel, uint64_t seed) {
    uint64_t nx, nz;
    oracle_grid_shape(nvoxel, &nx, &nz);
    const int e = oracle_los_exponent(nvoxel);
#pragma omp parallel for schedule(static)
    for (int64_t r = 0; r < (int64_t)nrows; ++r) {
        const uint64_t i = row0 + (uint64_t)r;
        const uint64_t h = oracle_splitmix64((seed ^ 0xA5A5A5A55A5A5A5AULL) ^ (i * 0xD1342543DE82EF95ULL));
        const uint64_t h2 = oracle_splitmix64(h);
        const int64_t a = (int64_t)(h & 2047u) - 1023;
        const int64_t mag = 1024 - (a < 0 ? -a : a);
        const int64_t b = ((h >> 11) & 1u) ? mag : -mag;
        const int64_t u0 = (int64_t)((h2 & 0xFFFFFFu) % nx), w0 = (int64_t)(((h2 >> 24) & 0xFFFFFFu) % nz);
        const int64_t c = a * u0 + b * w0;
        float* row = H + (uint64_t)r * nvoxel;
        for (uint64_t j = 0; j < nvoxel; ++j) {
            int64_t d = a * (int64_t)(j / nz) + b * (int64_t)(j % nz) - c;
            d = d < 0 ? -d : d;
            const uint32_t direct = d < 2048 ? (uint32_t)(2048 - d) >> 1 : 0u;
            const uint32_t bg = (uint32_t)(oracle_splitmix64((seed + 1) ^ ((i * nvoxel + j) * 0xD1342543DE82EF95ULL)) >> 56) | 1u;
            row[j] = ldexpf((float)(direct * 1024u + bg), -e);
        }
    }
}

void oracle_forward(const float* H, uint64_t npixel, uint64_t nvoxel, const double* x, double* out) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < (int64_t)npixel; ++i) {
        const float* row = H + (uint64_t)i * nvoxel;
        double res = 0;
        for (uint64_t j = 0; j < nvoxel; ++j) res += (double)row[j] * x[j];
        out[i] = res;
    }
}
