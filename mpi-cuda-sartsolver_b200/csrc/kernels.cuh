// Device code of the B200 constrained-SART engine (sm_100a).
//
// One SART iteration of the reference is (source/sartsolver_cuda.cpp:231-262, CPU twin source/sartsolver.cpp:180-229)
//   penalty -> PropagateKernel (column reduction over H) -> host MPI_Allreduce -> UpdateSolutionKernel
//   -> cublasSgemv (row reduction over H) -> cublasSdot -> host MPI_Allreduce -> host convergence test.
// Here the same arithmetic is regrouped into one "sweep" per solution iterate x_t:
//   forward projection  fit = H x_t  with the ray-length-normalised residual w and sum(fit^2) in its epilogue,
//   decision            convergence test for x_t on the device (after an all-reduce of the scalar across row shards),
//   back projection     s   = H^T w  (deterministic: per-slab partials, fixed-order reduction, no atomics),
//   [all-reduce of s across row shards],
//   update              x_{t+1} = max(x_t + alpha/rho * s - beta L x_t, 0)
// so the only host involvement is a status poll every few sweeps. The matrix is streamed with 128-bit
// L1-bypassing loads, fp32 FMAs, fp64 for every cross-thread / cross-CTA sum and for all scalars.
#pragma once

#include <cstdint>
#include <cuda_runtime.h>

namespace sart {

constexpr unsigned FULL_MASK = 0xffffffffu;

// Device-resident control block of one solve; two copies, ping-ponged by sweep parity so that a kernel never
// reads the copy another block of the same launch writes.
struct IterState {
    double conv_prev;  // convergence value of the previous reference iteration (sartsolver.cpp:179, 228)
    int k;             // solution updates applied so far == index of the sweep that evaluates the current iterate
    int done;          // 1: the solve is over, every later launch returns at once
    int status;        // SART_STATUS_* once done
    int iterations;    // passes through the reference's iteration loop
    int final_buf;     // which of the two solution buffers holds the answer
    int pad;
};

// What one frame of a warm-started series leaves for the next one (sart_solve_series): read by the next launch of the
// whole-solve kernel and by its set-up kernels, so that the host can enqueue a whole chunk of frames without waiting
// for any of them.
struct SeriesCtl {
    int handed;              // 1: the previous frame's last sweep already did this frame's first iteration
    int pflip;               // solution-buffer flip of the next launch when handed
    int it_prev, it_prev2;   // iteration counts of the last two frames (-1: none yet)
    unsigned long long epoch;  // exchange epoch after the last launch (several ranks)
    unsigned long long passes; // sweeps run so far in this series
    int frames_handed, sweeps_repeated;
    int timed_out, pad;
};

struct FrameScalars {
    double gmax;  // max_i g_i over all ranks (sartsolver_cuda.cpp:147-149)
    double gsq;   // sum over g_i > 0 of g_i^2 over all ranks (sartsolver.cpp:161-164)
    double norm;  // gmax, or 1 when gmax <= 0
    double m2;    // gsq / norm^2 (sartsolver_cuda.cpp:157)
};

__device__ __forceinline__ float4 ldg_stream(const float4* p) {
    float4 r;
    asm("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];"
        : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
        : "l"(p));
    return r;
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL_MASK, v, o);
    return v;
}

__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(FULL_MASK, v, o));
    return v;
}

// Exchange regions mapped between the ranks of one box (CUDA IPC): flags are kFlagStride bytes apart, written with
// system-scope release stores into every peer's region and polled with acquire loads in one's own.
constexpr int kMaxPeers = 16;
constexpr int kFlagStride = 128;
constexpr int kStatSlots = 1024;  // frames per statistics exchange (a longer chunk goes in several rounds)

__device__ __forceinline__ void st_release_sys_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

// ---------------------------------------------------------------------------------------------------------------------
// Forward projection: fit_i = sum_j H_ij x_j over the local rows  (replaces cublasSgemv, sartsolver_cuda.cpp:188-189,
// 248-249; CPU twin sartsolver.cpp:166-173, 213-220), with in its epilogue
//   w_i     = [lambda_i > h_l][g_i >= 0] (g_i - fit_i) / lambda_i        (prologue of PropagateKernel, sart_kernels.cu:75-82)
//   or, log solver:  [..][..] fit_i / lambda_i                            (LogPropagateKernel, sart_kernels.cu:126-140)
//   sum_i fit_i^2 over ALL rows, one fp64 partial per CTA                 (replaces cublasSdot, sartsolver_cuda.cpp:251-253)
// A CTA owns R full rows at a time (persistent, strided over row groups). Every thread keeps R fp32 accumulators,
// loads one float4 of x per column step and R float4 of H against it, UNROLL steps in flight.
// ---------------------------------------------------------------------------------------------------------------------
struct FpArgs {
    const float* H;      // [P][ld]
    uint64_t ld4;        // ld / 4
    uint64_t P;
    const float* x;      // [ld], zero beyond nvoxel
    const float* g;      // [P] normalised measurement
    const float* rinv;   // [P] (lambda_i > h_l) ? 1 / lambda_i : 0
    float* fit;          // [P]
    float* w;            // [P]
    double* f2part;      // [gridDim.x]
    const IterState* st; // may be null
    uint32_t nrg;        // number of row groups = ceil(P / R)
    int log_mode;
};

template <int R, int NT, int UNROLL>
__global__ void __launch_bounds__(NT, 2) fp_rows_kernel(const FpArgs a) {
    if (a.st && a.st->done) return;
    constexpr int NW = NT / 32;
    __shared__ double red[R][NW];
    __shared__ double f2s[R];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float4* __restrict__ H4 = reinterpret_cast<const float4*>(a.H);
    const float4* __restrict__ x4 = reinterpret_cast<const float4*>(a.x);
    const uint32_t ld4 = (uint32_t)a.ld4;
    double f2acc = 0.0;

    for (uint32_t rg = blockIdx.x; rg < a.nrg; rg += gridDim.x) {
        const uint64_t row0 = (uint64_t)rg * R;
        const int nvalid = (int)min((uint64_t)R, a.P - row0);
        const float4* __restrict__ base = H4 + row0 * a.ld4;
        uint32_t roff[R];
#pragma unroll
        for (int r = 0; r < R; ++r) roff[r] = (r < nvalid ? r : 0) * ld4;  // rows past the end re-read row0, result unused

        float acc[R];
#pragma unroll
        for (int r = 0; r < R; ++r) acc[r] = 0.f;

        uint32_t c = tid;
        for (; c + (UNROLL - 1) * NT < ld4; c += UNROLL * NT) {
            float4 xv[UNROLL];
            float4 h[UNROLL][R];
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) xv[u] = __ldg(x4 + c + u * NT);
#pragma unroll
            for (int u = 0; u < UNROLL; ++u)
#pragma unroll
                for (int r = 0; r < R; ++r) h[u][r] = ldg_stream(base + roff[r] + c + u * NT);
#pragma unroll
            for (int u = 0; u < UNROLL; ++u)
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    acc[r] = fmaf(h[u][r].x, xv[u].x, acc[r]);
                    acc[r] = fmaf(h[u][r].y, xv[u].y, acc[r]);
                    acc[r] = fmaf(h[u][r].z, xv[u].z, acc[r]);
                    acc[r] = fmaf(h[u][r].w, xv[u].w, acc[r]);
                }
        }
        for (; c < ld4; c += NT) {
            const float4 xv = __ldg(x4 + c);
            float4 h[R];
#pragma unroll
            for (int r = 0; r < R; ++r) h[r] = ldg_stream(base + roff[r] + c);
#pragma unroll
            for (int r = 0; r < R; ++r) {
                acc[r] = fmaf(h[r].x, xv.x, acc[r]);
                acc[r] = fmaf(h[r].y, xv.y, acc[r]);
                acc[r] = fmaf(h[r].z, xv.z, acc[r]);
                acc[r] = fmaf(h[r].w, xv.w, acc[r]);
            }
        }

        // cross-thread sums in fp64, fixed order
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const double s = warp_sum((double)acc[r]);
            if (lane == 0) red[r][warp] = s;
        }
        __syncthreads();
        if (tid < nvalid) {
            double s = 0.0;
#pragma unroll
            for (int k = 0; k < NW; ++k) s += red[tid][k];
            const uint64_t i = row0 + tid;
            const float fitv = (float)s;
            f2acc += s * s;
            const float gi = a.g[i];
            const float ri = a.rinv[i];
            a.fit[i] = fitv;
            float wi = 0.f;
            if (gi >= 0.f) wi = a.log_mode ? fitv * ri : (gi - fitv) * ri;
            a.w[i] = wi;
        }
        __syncthreads();
    }

    if (tid < R) f2s[tid] = f2acc;
    __syncthreads();
    if (tid == 0) {
        double s = 0.0;
#pragma unroll
        for (int r = 0; r < R; ++r) s += f2s[r];
        a.f2part[blockIdx.x] = s;
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Back projection: s_j = sum_i H_ij w_i over the rows of one slab  (the main loop of PropagateKernel,
// sart_kernels.cu:86-101, of InitialGuessKernel :40-54 and of LogPropagateKernel :146-170; CPU twins
// sartsolver.cpp:149-157, 191-200, 295-310). grid = (column panels of NT float4, row slabs). A thread owns four
// adjacent columns, walks down its slab with U rows in flight; the 32 row weights of a step are loaded once per
// warp and broadcast by shuffle. Each (slab, column) partial is written exactly once: no atomics, no memset,
// bit-reproducible. The slabs are added in fixed order by reduce_partials_kernel.
// ---------------------------------------------------------------------------------------------------------------------
struct BpArgs {
    const float* H;
    uint64_t ld4;
    uint64_t P;
    const float* w;          // [P]
    float* part;             // [nslabs][ld]
    uint32_t rows_per_slab;
    const IterState* st;     // may be null
};

template <int NT, int U>
__device__ __forceinline__ void bp_cols_body(const BpArgs& a, float* __restrict__ part) {
    const int tid = threadIdx.x, lane = tid & 31;
    const uint64_t c4 = (uint64_t)blockIdx.x * NT + tid;
    const bool active = c4 < a.ld4;
    const float4* __restrict__ base = reinterpret_cast<const float4*>(a.H) + (active ? c4 : a.ld4 - 1);
    const uint64_t r0 = (uint64_t)blockIdx.y * a.rows_per_slab;
    const uint64_t r1 = min(a.P, r0 + a.rows_per_slab);

    float4 acc0 = make_float4(0.f, 0.f, 0.f, 0.f), acc1 = acc0;
    for (uint64_t r = r0; r < r1; r += 32) {
        const uint64_t rl = r + lane;
        const float wl = (rl < r1) ? __ldg(a.w + rl) : 0.f;
        const float4* __restrict__ p = base + r * a.ld4;
        if (r + 32 <= r1) {
#pragma unroll
            for (int k = 0; k < 32; k += U) {
                float4 h[U];
#pragma unroll
                for (int u = 0; u < U; ++u) h[u] = ldg_stream(p + (uint64_t)(k + u) * a.ld4);
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const float wv = __shfl_sync(FULL_MASK, wl, k + u);
                    float4& acc = (u & 1) ? acc1 : acc0;
                    acc.x = fmaf(h[u].x, wv, acc.x);
                    acc.y = fmaf(h[u].y, wv, acc.y);
                    acc.z = fmaf(h[u].z, wv, acc.z);
                    acc.w = fmaf(h[u].w, wv, acc.w);
                }
            }
        } else {
            const int n = (int)(r1 - r);
            for (int k = 0; k < n; ++k) {
                const float4 h = ldg_stream(p + (uint64_t)k * a.ld4);
                const float wv = __shfl_sync(FULL_MASK, wl, k);
                acc0.x = fmaf(h.x, wv, acc0.x);
                acc0.y = fmaf(h.y, wv, acc0.y);
                acc0.z = fmaf(h.z, wv, acc0.z);
                acc0.w = fmaf(h.w, wv, acc0.w);
            }
        }
    }
    if (active) {
        float4 o;
        o.x = acc0.x + acc1.x;
        o.y = acc0.y + acc1.y;
        o.z = acc0.z + acc1.z;
        o.w = acc0.w + acc1.w;
        reinterpret_cast<float4*>(part)[(uint64_t)blockIdx.y * a.ld4 + c4] = o;
    }
}

template <int NT, int U>
__global__ void __launch_bounds__(NT) bp_cols_kernel(const BpArgs a) {
    if (a.st && a.st->done) return;
    bp_cols_body<NT, U>(a, a.part);
}

// ---------------------------------------------------------------------------------------------------------------------
// F independent frames per pass over the matrix (sart_solve_batch). Frames solved without a warm start
// share nothing but H, so one sweep can serve F of them: the forward projection becomes [P x V] x [V x F], the back
// projection [V x P] x [P x F] -- still CUDA-core fp32 (TF32 tensor cores would lose the 1e-5 parity) and still
// HBM-bound at F = 4: 4 FMAs per matrix element against 0.5 flop/byte of headroom. Same thread mappings, summation
// orders and epilogues as the single-frame kernels above, with an extra frame index on x, g, w, the partials and the
// control blocks. A frame that is done keeps being multiplied (its results are ignored): the pass costs the same.
// ---------------------------------------------------------------------------------------------------------------------
struct FpBatchArgs {
    const float* H;
    uint64_t ld4, P;
    const float* x;       // [F][ld]
    const float* g;       // [F][P]
    const float* rinv;    // [P]
    float* w;             // [F][P]
    double* f2part;       // [F][gridDim.x]
    const IterState* st;  // [F] or null
    uint32_t nrg;
    int nframes;          // frames in use (<= F); the others are skipped in the epilogue
    int log_mode;         // log solver: w = fit / lambda instead of (g - fit) / lambda
};

template <int R, int NT, int F>
__global__ void __launch_bounds__(NT, 2) fp_rows_batch_kernel(const FpBatchArgs a) {
    if (a.st) {
        bool all_done = true;
        for (int f = 0; f < a.nframes; ++f) all_done = all_done && a.st[f].done;
        if (all_done) return;
    }
    constexpr int NW = NT / 32;
    __shared__ double red[R][F][NW];
    __shared__ double f2s[R][F];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float4* __restrict__ H4 = reinterpret_cast<const float4*>(a.H);
    const float4* __restrict__ x4 = reinterpret_cast<const float4*>(a.x);
    const uint32_t ld4 = (uint32_t)a.ld4;
    double f2acc = 0.0;  // thread (r, f) = tid: row slot r of frame f

    for (uint32_t rg = blockIdx.x; rg < a.nrg; rg += gridDim.x) {
        const uint64_t row0 = (uint64_t)rg * R;
        const int nvalid = (int)min((uint64_t)R, a.P - row0);
        const float4* __restrict__ base = H4 + row0 * a.ld4;
        uint32_t roff[R];
#pragma unroll
        for (int r = 0; r < R; ++r) roff[r] = (r < nvalid ? r : 0) * ld4;

        float acc[R][F];
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int f = 0; f < F; ++f) acc[r][f] = 0.f;

        for (uint32_t c = tid; c < ld4; c += NT) {
            float4 xv[F], h[R];
#pragma unroll
            for (int r = 0; r < R; ++r) h[r] = ldg_stream(base + roff[r] + c);
#pragma unroll
            for (int f = 0; f < F; ++f) xv[f] = __ldg(x4 + (uint64_t)f * ld4 + c);
#pragma unroll
            for (int r = 0; r < R; ++r)
#pragma unroll
                for (int f = 0; f < F; ++f) {
                    acc[r][f] = fmaf(h[r].x, xv[f].x, acc[r][f]);
                    acc[r][f] = fmaf(h[r].y, xv[f].y, acc[r][f]);
                    acc[r][f] = fmaf(h[r].z, xv[f].z, acc[r][f]);
                    acc[r][f] = fmaf(h[r].w, xv[f].w, acc[r][f]);
                }
        }
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int f = 0; f < F; ++f) {
                const double s = warp_sum((double)acc[r][f]);
                if (lane == 0) red[r][f][warp] = s;
            }
        __syncthreads();
        if (tid < R * F) {
            const int r = tid / F, f = tid % F;
            if (r < nvalid && f < a.nframes) {
                double s = 0.0;
#pragma unroll
                for (int k = 0; k < NW; ++k) s += red[r][f][k];
                const uint64_t i = row0 + r;
                const float fitv = (float)s;
                f2acc += s * s;
                const float gi = a.g[(uint64_t)f * a.P + i];
                a.w[(uint64_t)f * a.P + i] = (gi >= 0.f) ? (a.log_mode ? fitv : gi - fitv) * a.rinv[i] : 0.f;
            }
        }
        __syncthreads();
    }
    if (tid < R * F) f2s[tid / F][tid % F] = f2acc;
    __syncthreads();
    if (tid < a.nframes) {
        double s = 0.0;
#pragma unroll
        for (int r = 0; r < R; ++r) s += f2s[r][tid];
        a.f2part[(uint64_t)tid * gridDim.x + blockIdx.x] = s;
    }
}

// The same for wide groups (F = 8), where re-reading x dominates: above, a CTA walks ONE row group and every thread
// fetches F float4 of x per column step -- with F = 8 that is twice the bytes of the matrix itself coming out of L2.
// Here every WARP owns a row group of R rows (R * F = 32 accumulators per thread, lane l reads columns l, l + 32, ...)
// and the warps of a CTA walk the columns side by side, so a column step's x values are fetched from L2 once per CTA
// and found in L1 by the other warps. A row's dot products never leave the warp: shuffle tree, lane r * F + f finishes
// (row r, frame f).
template <int R, int NT, int F, int MINB>
__global__ void __launch_bounds__(NT, MINB) fp_rows_batch_warp_kernel(const FpBatchArgs a) {
    static_assert((R * F) % 32 == 0 && 32 % F == 0, "whole (row, frame) slots per lane in the epilogue");
    constexpr int PER_LANE = R * F / 32;  // lane l finishes (row l / F + q * 32 / F, frame l % F), q < PER_LANE
    if (a.st) {
        bool all_done = true;
        for (int f = 0; f < a.nframes; ++f) all_done = all_done && a.st[f].done;
        if (all_done) return;
    }
    constexpr int NW = NT / 32;
    __shared__ double f2s[NW][F];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float4* __restrict__ H4 = reinterpret_cast<const float4*>(a.H);
    const float4* __restrict__ x4 = reinterpret_cast<const float4*>(a.x);
    const uint32_t ld4 = (uint32_t)a.ld4;
    const int my_r = lane / F, my_f = lane % F;
    double f2acc = 0.0;  // of frame my_f, this lane's row slots

    for (uint32_t rg = blockIdx.x * NW + warp; rg < a.nrg; rg += gridDim.x * NW) {
        const uint64_t row0 = (uint64_t)rg * R;
        const int nvalid = (int)min((uint64_t)R, a.P - row0);
        const float4* __restrict__ base = H4 + row0 * a.ld4;
        uint32_t roff[R];
#pragma unroll
        for (int r = 0; r < R; ++r) roff[r] = (r < nvalid ? r : 0) * ld4;
        float acc[R][F];
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int f = 0; f < F; ++f) acc[r][f] = 0.f;
#pragma unroll 1
        for (uint32_t c = lane; c < ld4; c += 32) {
            float4 xv[F], h[R];
#pragma unroll
            for (int r = 0; r < R; ++r) h[r] = ldg_stream(base + roff[r] + c);
#pragma unroll
            for (int f = 0; f < F; ++f) xv[f] = __ldg(x4 + (uint64_t)f * ld4 + c);
#pragma unroll
            for (int r = 0; r < R; ++r)
#pragma unroll
                for (int f = 0; f < F; ++f) {
                    acc[r][f] = fmaf(h[r].x, xv[f].x, acc[r][f]);
                    acc[r][f] = fmaf(h[r].y, xv[f].y, acc[r][f]);
                    acc[r][f] = fmaf(h[r].z, xv[f].z, acc[r][f]);
                    acc[r][f] = fmaf(h[r].w, xv[f].w, acc[r][f]);
                }
        }
        double mine[PER_LANE];
#pragma unroll
        for (int q = 0; q < PER_LANE; ++q) mine[q] = 0.0;
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int f = 0; f < F; ++f) {
                const double s = warp_sum((double)acc[r][f]);
                if (lane == (r * F + f) % 32) mine[(r * F + f) / 32] = s;
            }
#pragma unroll
        for (int q = 0; q < PER_LANE; ++q) {
            const int r = my_r + q * (32 / F);
            if (r < nvalid && my_f < a.nframes) {
                const uint64_t i = row0 + r;
                const float fitv = (float)mine[q];
                f2acc += mine[q] * mine[q];
                const float gi = a.g[(uint64_t)my_f * a.P + i];
                a.w[(uint64_t)my_f * a.P + i] = (gi >= 0.f) ? (a.log_mode ? fitv : gi - fitv) * a.rinv[i] : 0.f;
            }
        }
    }
    // sum(fit^2) per frame: over the row slots of a warp (lanes f, f + F, ...), then over the warps in order
#pragma unroll
    for (int o = F; o < 32; o <<= 1) f2acc += __shfl_xor_sync(FULL_MASK, f2acc, o);
    if (lane < F) f2s[warp][lane] = f2acc;
    __syncthreads();
    if ((int)threadIdx.x < a.nframes) {
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < NW; ++k) s += f2s[k][threadIdx.x];
        a.f2part[(uint64_t)threadIdx.x * gridDim.x + blockIdx.x] = s;
    }
}

struct BpBatchArgs {
    const float* H;
    uint64_t ld4, P;
    const float* w;   // [F][P]
    float* part;      // [F][nslabs][ld]
    uint32_t rows_per_slab;
    const IterState* st;  // [F] or null
    int nframes;
};

template <int NT, int U, int F>
__global__ void __launch_bounds__(NT) bp_cols_batch_kernel(const BpBatchArgs a) {
    if (a.st) {
        bool all_done = true;
        for (int f = 0; f < a.nframes; ++f) all_done = all_done && a.st[f].done;
        if (all_done) return;
    }
    const int tid = threadIdx.x, lane = tid & 31;
    const uint64_t c4 = (uint64_t)blockIdx.x * NT + tid;
    const bool active = c4 < a.ld4;
    const float4* __restrict__ base = reinterpret_cast<const float4*>(a.H) + (active ? c4 : a.ld4 - 1);
    const uint64_t r0 = (uint64_t)blockIdx.y * a.rows_per_slab;
    const uint64_t r1 = min(a.P, r0 + a.rows_per_slab);

    float4 acc[F];
#pragma unroll
    for (int f = 0; f < F; ++f) acc[f] = make_float4(0.f, 0.f, 0.f, 0.f);
    auto fma_row = [&](const float4& h, const float (&wl)[F], int src_lane) {
#pragma unroll
        for (int f = 0; f < F; ++f) {
            const float wv = __shfl_sync(FULL_MASK, wl[f], src_lane);
            acc[f].x = fmaf(h.x, wv, acc[f].x);
            acc[f].y = fmaf(h.y, wv, acc[f].y);
            acc[f].z = fmaf(h.z, wv, acc[f].z);
            acc[f].w = fmaf(h.w, wv, acc[f].w);
        }
    };
    for (uint64_t r = r0; r < r1; r += 32) {
        const uint64_t rl = r + lane;
        float wl[F];
#pragma unroll
        for (int f = 0; f < F; ++f) wl[f] = (rl < r1 && f < a.nframes) ? __ldg(a.w + (uint64_t)f * a.P + rl) : 0.f;
        const float4* __restrict__ p = base + r * a.ld4;
        if (r + 32 <= r1) {
#pragma unroll
            for (int k = 0; k < 32; k += U) {
                float4 h[U];  // U unconditional 128-bit loads in flight before the first FMA
#pragma unroll
                for (int u = 0; u < U; ++u) h[u] = ldg_stream(p + (uint64_t)(k + u) * a.ld4);
#pragma unroll
                for (int u = 0; u < U; ++u) fma_row(h[u], wl, k + u);
            }
        } else {
            const int n = (int)(r1 - r);
            for (int k = 0; k < n; ++k) fma_row(ldg_stream(p + (uint64_t)k * a.ld4), wl, k);
        }
    }
    if (active) {
        const uint64_t nslabs = gridDim.y;
#pragma unroll
        for (int f = 0; f < F; ++f)
            if (f < a.nframes) reinterpret_cast<float4*>(a.part)[((uint64_t)f * nslabs + blockIdx.y) * a.ld4 + c4] = acc[f];
    }
}

// Wide groups (F = 8): the F weights of a row come out of shared memory (two broadcast 128-bit loads) instead of F
// shuffles per row and thread; the weights of the next 32 rows are fetched while the current 32 are multiplied.
template <int NT, int U, int F>
__global__ void __launch_bounds__(NT) bp_cols_batch_smem_kernel(const BpBatchArgs a) {
    static_assert(NT >= 32 * F && F % 4 == 0, "one thread per (row, frame) of a 32-row block");
    if (a.st) {
        bool all_done = true;
        for (int f = 0; f < a.nframes; ++f) all_done = all_done && a.st[f].done;
        if (all_done) return;
    }
    __shared__ __align__(16) float ws[2][32][F];
    const int tid = threadIdx.x;
    const uint64_t c4 = (uint64_t)blockIdx.x * NT + tid;
    const bool active = c4 < a.ld4;
    const float4* __restrict__ base = reinterpret_cast<const float4*>(a.H) + (active ? c4 : a.ld4 - 1);
    const uint64_t r0 = (uint64_t)blockIdx.y * a.rows_per_slab;
    const uint64_t r1 = min(a.P, r0 + a.rows_per_slab);
    const int wrow = tid & 31, wf = tid >> 5;  // this thread's (row, frame) of a block's weights
    auto stage = [&](uint64_t r, int buf) {
        if (wf < F) ws[buf][wrow][wf] = (r + wrow < r1 && wf < a.nframes) ? __ldg(a.w + (uint64_t)wf * a.P + r + wrow) : 0.f;
    };
    float4 acc[F];
#pragma unroll
    for (int f = 0; f < F; ++f) acc[f] = make_float4(0.f, 0.f, 0.f, 0.f);
    auto fma_row = [&](const float4& h, const float* wrow_ptr) {
#pragma unroll
        for (int f4 = 0; f4 < F / 4; ++f4) {
            const float4 wv = *reinterpret_cast<const float4*>(wrow_ptr + 4 * f4);
            const float wf4[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                float4& ac = acc[4 * f4 + j];
                ac.x = fmaf(h.x, wf4[j], ac.x);
                ac.y = fmaf(h.y, wf4[j], ac.y);
                ac.z = fmaf(h.z, wf4[j], ac.z);
                ac.w = fmaf(h.w, wf4[j], ac.w);
            }
        }
    };
    int buf = 0;
    if (r0 < r1) stage(r0, 0);
    __syncthreads();
    for (uint64_t r = r0; r < r1; r += 32) {
        if (r + 32 < r1) stage(r + 32, buf ^ 1);
        const float4* __restrict__ p = base + r * a.ld4;
        if (r + 32 <= r1) {
#pragma unroll
            for (int k = 0; k < 32; k += U) {
                float4 h[U];  // U unconditional 128-bit loads in flight before the first FMA
#pragma unroll
                for (int u = 0; u < U; ++u) h[u] = ldg_stream(p + (uint64_t)(k + u) * a.ld4);
#pragma unroll
                for (int u = 0; u < U; ++u) fma_row(h[u], &ws[buf][k + u][0]);
            }
        } else {
            const int n = (int)(r1 - r);
            for (int k = 0; k < n; ++k) fma_row(ldg_stream(p + (uint64_t)k * a.ld4), &ws[buf][k][0]);
        }
        __syncthreads();  // everybody is done with ws[buf]; ws[buf ^ 1] is complete
        buf ^= 1;
    }
    if (active) {
        const uint64_t nslabs = gridDim.y;
#pragma unroll
        for (int f = 0; f < F; ++f)
            if (f < a.nframes) reinterpret_cast<float4*>(a.part)[((uint64_t)f * nslabs + blockIdx.y) * a.ld4 + c4] = acc[f];
    }
}

// s_j = sum over slabs (fixed order, fp64) -> sred[j]
// When f2part is given, warp 0 of block 0 also adds the nf2 partials of sum(fit^2) into sred[V] (one launch less
// where the scalar travels with the vector).
__global__ void reduce_partials_kernel(const float* __restrict__ part, uint32_t nslabs, uint64_t ld, uint64_t V,
                                       double* __restrict__ sred, const IterState* st,
                                       const double* __restrict__ f2part = nullptr, uint32_t nf2 = 0) {
    if (st && st->done) return;
    if (f2part && blockIdx.x == 0 && threadIdx.x < 32) {
        double f = 0.0;
        for (uint32_t k = threadIdx.x; k < nf2; k += 32) f += f2part[k];
        f = warp_sum(f);
        if (threadIdx.x == 0) sred[V] = f;
    }
    const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= V) return;
    double s = 0.0;
    for (uint32_t k = 0; k < nslabs; ++k) s += (double)part[(uint64_t)k * ld + j];
    sred[j] = s;
}

// The same for the frames of a group (blockIdx.y = frame): partials part_stride apart, sums sred_stride apart, the frames'
// sum(fit^2) partials f2_stride apart. One launch instead of one per frame.
__global__ void reduce_partials_batch_kernel(const float* __restrict__ part, uint32_t nslabs, uint64_t ld, uint64_t V,
                                             uint64_t part_stride, double* __restrict__ sred, uint64_t sred_stride,
                                             const IterState* st, const double* __restrict__ f2part = nullptr, uint32_t nf2 = 0,
                                             uint32_t f2_stride = 0) {
    const uint32_t fr = blockIdx.y;
    if (st && st[fr].done) return;
    part += (uint64_t)fr * part_stride;
    sred += (uint64_t)fr * sred_stride;
    if (f2part && blockIdx.x == 0 && threadIdx.x < 32) {
        double f = 0.0;
        for (uint32_t k = threadIdx.x; k < nf2; k += 32) f += f2part[(uint64_t)fr * f2_stride + k];
        f = warp_sum(f);
        if (threadIdx.x == 0) sred[V] = f;
    }
    const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= V) return;
    double s = 0.0;
    for (uint32_t k = 0; k < nslabs; ++k) s += (double)part[(uint64_t)k * ld + j];
    sred[j] = s;
}

// sum(fit^2) of this shard from the forward projection's per-CTA partials (one warp, fixed order); only launched
// when several ranks have to all-reduce the scalar before the decision.
__global__ void f2_sum_kernel(const double* __restrict__ f2part, uint32_t n, double* __restrict__ out, const IterState* st) {
    if (st && st->done) return;
    double f = 0.0;
    for (uint32_t k = threadIdx.x; k < n; k += 32) f += f2part[k];
    f = warp_sum(f);
    if (threadIdx.x == 0) *out = f;
}

// ... of every frame of a group: CTA f sums frame f's partials (stride apart) into out[f]
__global__ void f2_sum_batch_kernel(const double* __restrict__ f2part, uint32_t n, uint32_t stride, double* __restrict__ out,
                                    const IterState* st) {
    const uint32_t fr = blockIdx.x;
    if (st && st[fr].done) return;
    double f = 0.0;
    for (uint32_t k = threadIdx.x; k < n; k += 32) f += f2part[(uint64_t)fr * stride + k];
    f = warp_sum(f);
    if (threadIdx.x == 0) out[fr] = f;
}

// ---------------------------------------------------------------------------------------------------------------------
// Decision for the iterate x_t the forward projection just evaluated (the host-side test of the reference,
// sartsolver_cuda.cpp:251-261, CPU twin sartsolver.cpp:222-228), taken on the device BEFORE the back projection so
// that a converged solve does not pay for one more matrix sweep. One warp. Reads state[t & 1], writes
// state[(t + 1) & 1]; every later kernel of the sweep (and all later sweeps) looks at the new state.
// ---------------------------------------------------------------------------------------------------------------------
struct DecideArgs {
    const double* f2part;  // per-CTA partials of sum(fit^2), or the all-reduced scalar (n == 1)
    uint32_t n;
    const FrameScalars* fs;
    const IterState* st_in;
    IterState* st_out;
    double tol;
    int max_iter, fixed, t;
};

__device__ __forceinline__ void decide_body(const DecideArgs& a) {
    const IterState s = *a.st_in;
    if (s.done) {
        if (threadIdx.x == 0) *a.st_out = s;
        return;
    }
    double f2 = 0.0;
    for (uint32_t k = threadIdx.x; k < a.n; k += 32) f2 += a.f2part[k];
    f2 = warp_sum(f2);
    if (threadIdx.x != 0) return;
    IterState o = s;
    bool stop = false;
    if (a.t >= 1) {
        // sweep t evaluated x_t, the result of reference iteration iter = t - 1
        const double m2 = a.fs->m2;
        const double conv = (m2 - f2) / m2;
        const int iter = a.t - 1;
        if (!a.fixed && iter >= 1 && fabs(conv - s.conv_prev) < a.tol) {
            stop = true;
            o.status = 0;
        } else if (a.t >= a.max_iter) {
            stop = true;
            o.status = -1;
        }
        o.conv_prev = conv;
    }
    if (stop) {
        o.done = 1;
        o.iterations = a.t;
        o.final_buf = a.t & 1;
    } else {
        o.k = a.t + 1;
    }
    *a.st_out = o;
}

__global__ void decide_kernel(const DecideArgs a) { decide_body(a); }

// the frames of a group (sart_solve_batch) in one launch: CTA f decides frame f; f2_stride apart lie the frames' partials
__global__ void decide_batch_kernel(DecideArgs a, uint32_t f2_stride) {
    const uint32_t f = blockIdx.x;
    a.f2part += (uint64_t)f * f2_stride;
    a.fs += f;
    a.st_in += f;
    a.st_out += f;
    decide_body(a);
}

// ---------------------------------------------------------------------------------------------------------------------
// Update: x_{t+1} from x_t and the (all-reduced) back projection.
// Replaces GradPenaltyKernel (sart_kernels.cu:179-189), the epilogue of PropagateKernel (:103-109) and
// UpdateSolutionKernel (:205-212) / UpdateLogSolutionKernel (:215-224); CPU twins sartsolver.cpp:183-189, 201-209
// and :287-293, 314-316. The Laplacian is CSR (rows sorted by column: the order the reference's sorted COO adds in);
// the penalty uses the iterate BEFORE the update, which lives in the other buffer.
// ---------------------------------------------------------------------------------------------------------------------
struct UpdArgs {
    const double* sred;         // [V] back-projected sums, already reduced over ranks
    const double* xd_cur;       // [V] current iterate (normalised by norm), fp64 master copy
    double* xd_next;            // [V]
    float* xf_next;             // [ld] fp32 copy the next forward projection reads
    const double* ray_density;  // [V]
    const double* obs;          // [V] log solver: back-projected measurement (iteration-invariant)
    const int* lap_rowptr;      // [V + 1] or null
    const int* lap_col;
    const float* lap_val;
    const FrameScalars* fs;
    const IterState* st;        // state after this sweep's decision
    uint64_t V;
    double beta, relax, h_d;
    double eps_norm, eps_orig;  // log solver's epsilon: in normalised units (GPU semantics) or caller units (CPU, 1e-100)
    int log_mode;
    int log_caller_units;       // log solver: the penalty takes log(x) of the solution in the caller's units (CPU solver,
                                // sartsolver.cpp:292) instead of the normalised one (GPU solver, sart_kernels.cu:200)
    double out_scale;           // 1, or the factor into the next frame's normalisation (series hand-over, fused.cuh)
};

// x_{t+1,j} from x_t and the reduced back projection s_j of voxel j. x_t is read through L2 (ld.global.cg): in the
// whole-solve kernel it was written by other SMs earlier in the same launch.
// The iterate's buffers and the output scale are passed apart from the constant fields, so that the whole-solve kernel can
// hand in its kernel parameter block as it is (no per-sweep copy of the struct in registers).
__device__ __forceinline__ void apply_update_at(const UpdArgs& a, const double* __restrict__ xd_cur, double* __restrict__ xd_next,
                                                float* __restrict__ xf_next, double out_scale, uint64_t j, double sj) {
    double pen = 0.0;
    if (a.lap_rowptr) {
        const int b = __ldg(a.lap_rowptr + j), e = __ldg(a.lap_rowptr + j + 1);
        // The iterate is kept as x / max(g). The linear penalty is scale-covariant; the log solver's is not: the CPU
        // solver takes log() of the solution in the caller's units, so a Laplacian row whose entries do not add up to
        // zero sees the normalisation (beta * log(norm) * rowsum). The GPU solver takes it of the normalised iterate.
        const double unit = (a.log_mode && a.log_caller_units) ? a.fs->norm : 1.0;
        // Entries of a row in batches: the column indices and values first (read-only for the life of the engine: through
        // L1, where a CTA of the whole-solve kernel finds its own voxels' rows again every sweep -- the matrix streams
        // past L1), then the iterate's values (written by other SMs earlier in the launch: through L2), then the sum in
        // the stored order. Three dependent memory round trips per batch instead of two per entry.
        constexpr int LAPB = 2;
        for (int k0 = b; k0 < e; k0 += LAPB) {
            int col[LAPB];
            float val[LAPB];
            double xv[LAPB];
#pragma unroll
            for (int i = 0; i < LAPB; ++i)
                if (k0 + i < e) {
                    col[i] = __ldg(a.lap_col + k0 + i);
                    val[i] = __ldg(a.lap_val + k0 + i);
                }
#pragma unroll
            for (int i = 0; i < LAPB; ++i)
                if (k0 + i < e) xv[i] = __ldcg(xd_cur + col[i]);
#pragma unroll
            for (int i = 0; i < LAPB; ++i)
                if (k0 + i < e) pen += a.beta * (double)val[i] * (a.log_mode ? log(xv[i] * unit) : xv[i]);
        }
    }
    const double rho = __ldg(a.ray_density + j);
    const bool on = rho > a.h_d;
    double xn;
    if (!a.log_mode) {
        const double diff = on ? sj * (a.relax / rho) : 0.0;
        xn = __ldcg(xd_cur + j) + (diff - pen);
        if (signbit(xn)) xn = 0.0;
    } else {
        const double fitj = on ? sj : 0.0;
        const double eps = a.eps_norm + a.eps_orig / a.fs->norm;
        xn = __ldcg(xd_cur + j) * pow((a.obs[j] + eps) / (fitj + eps), a.relax) * exp(-pen);
    }
    xn *= out_scale;
    xd_next[j] = xn;
    xf_next[j] = (float)xn;
}

__device__ __forceinline__ void apply_update(const UpdArgs& a, uint64_t j, double sj) {
    apply_update_at(a, a.xd_cur, a.xd_next, a.xf_next, a.out_scale, j, sj);
}

__global__ void update_kernel(const UpdArgs a) {
    if (a.st->done) return;
    const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= a.V) return;
    apply_update(a, j, a.sred[j]);
}

// Frame f of a group: the frames' vectors lie one after the other ([V + 2] reduced sums, [V] iterates and obs, [ld] fp32
// copies), their scalars and control blocks in arrays.
__device__ __forceinline__ UpdArgs frame_view(UpdArgs u, uint32_t f, uint64_t ld) {
    u.sred += (uint64_t)f * (u.V + 2);
    u.xd_cur += (uint64_t)f * u.V;
    u.xd_next += (uint64_t)f * u.V;
    u.xf_next += (uint64_t)f * ld;
    u.obs += (uint64_t)f * u.V;
    u.fs += f;
    if (u.st) u.st += f;
    return u;
}

// update of every frame of a group in one launch (blockIdx.y = frame); sums already reduced over slabs and ranks
__global__ void update_batch_kernel(const UpdArgs a, uint64_t ld) {
    const UpdArgs u = frame_view(a, blockIdx.y, ld);
    if (u.st->done) return;
    const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= u.V) return;
    apply_update(u, j, u.sred[j]);
}

// one rank: the slab partials of the back projection summed (the order of reduce_partials_kernel) and the update applied
// in the same launch, for every frame of the group
__global__ void reduce_update_batch_kernel(const UpdArgs a, uint64_t ld, const float* __restrict__ part, uint32_t nslabs,
                                           uint64_t part_stride) {
    const UpdArgs u = frame_view(a, blockIdx.y, ld);
    if (u.st->done) return;
    const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= u.V) return;
    part += (uint64_t)blockIdx.y * part_stride;
    double s = 0.0;
    for (uint32_t k = 0; k < nslabs; ++k) s += (double)part[(uint64_t)k * ld + j];
    apply_update(u, j, s);
}

// ---------------------------------------------------------------------------------------------------------------------
// Cross-GPU reduction + decision + update in ONE kernel over NVLink peer memory (replaces the D2H -> MPI_Allreduce ->
// H2D round trip of the reference, sartsolver_cuda.cpp:242-244 and :253-255, and this engine's own
// reduce -> ncclAllReduce -> decide -> update launch chain when the ranks can map each other's memory).
// Every rank owns an exchange region that all ranks of the box map through CUDA IPC:
//     [nranks flags, 128 B apart][buffer 0: V + 1 doubles][buffer 1: V + 1 doubles]
// Sweep number `epoch` uses buffer epoch & 1. Before this kernel the rank's own reduced back projection s[V] and its
// sum(fit^2) have been written to its buffer (reduce_partials_kernel). Then
//   1. block 0 publishes `epoch` into flag[rank] of EVERY rank's region (release, system scope);
//   2. every block waits until all of its own region's flags have reached `epoch` (acquire, system scope): all
//      partial vectors are then complete and visible;
//   3. every thread adds the nranks values of its voxel in rank order (peer loads over NVLink, L1-bypassing), so all
//      ranks obtain the bit-identical sum, takes the same stop/continue decision from the summed sum(fit^2), and
//      applies the update for its voxel.
// Buffer reuse needs no second barrier: a rank overwrites buffer b two sweeps later, after it has passed step 2 of
// the sweep in between, which every peer only signals once it has finished reading buffer b.
// A peer that never arrives (crashed rank) ends the wait after ~4 s with status SART_STATUS_EXCHANGE_TIMEOUT.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int kStatusExchangeTimeout = -3;

struct XchgArgs {
    char* const* peers;  // [nranks] base address of every rank's exchange region as mapped in THIS process
    uint32_t nranks, rank;
    unsigned long long epoch;
    uint64_t buf_off;  // byte offset of this epoch's buffer inside a region
    const IterState* st_in;
    IterState* st_out;
    double tol;
    int max_iter, fixed, t;
};

__device__ __forceinline__ void xchg_decide_update_body(const XchgArgs& x, const UpdArgs& u) {
    const IterState s = *x.st_in;
    const bool writer = (blockIdx.x == 0 && threadIdx.x == 0);
    if (s.done) {
        if (writer) *x.st_out = s;
        return;
    }
    __shared__ int timed_out;
    __shared__ double f2r[kMaxPeers];
    if (threadIdx.x == 0) timed_out = 0;
    __syncthreads();
    if (blockIdx.x == 0 && threadIdx.x < x.nranks) {
        __threadfence_system();
        st_release_sys_u64(reinterpret_cast<unsigned long long*>(x.peers[threadIdx.x] + x.rank * kFlagStride), x.epoch);
    }
    if (threadIdx.x < x.nranks) {
        const unsigned long long* flag =
            reinterpret_cast<const unsigned long long*>(x.peers[x.rank] + threadIdx.x * kFlagStride);
        const long long t0 = clock64();
        while (ld_acquire_sys_u64(flag) < x.epoch) {
            if (clock64() - t0 > 8000000000LL) {  // ~4 s at 2 GHz
                timed_out = 1;
                break;
            }
        }
        // this rank's sum(fit^2), straight from the owner's buffer
        f2r[threadIdx.x] = __ldcv(reinterpret_cast<const double*>(x.peers[threadIdx.x] + x.buf_off) + u.V);
    }
    __syncthreads();

    IterState o = s;
    bool stop = false;
    if (timed_out) {
        stop = true;
        o.status = kStatusExchangeTimeout;
    } else if (x.t >= 1) {
        double f2 = 0.0;
        for (uint32_t r = 0; r < x.nranks; ++r) f2 += f2r[r];
        const double m2 = u.fs->m2;
        const double conv = (m2 - f2) / m2;
        const int iter = x.t - 1;
        if (!x.fixed && iter >= 1 && fabs(conv - s.conv_prev) < x.tol) {
            stop = true;
            o.status = 0;
        } else if (x.t >= x.max_iter) {
            stop = true;
            o.status = -1;
        }
        o.conv_prev = conv;
    }
    if (stop) {
        o.done = 1;
        o.iterations = x.t;
        o.final_buf = x.t & 1;
    } else {
        o.k = x.t + 1;
    }
    if (writer) *x.st_out = o;
    const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (stop || j >= u.V) return;

    double part[kMaxPeers];
#pragma unroll
    for (int r = 0; r < kMaxPeers; ++r)
        if (r < (int)x.nranks) part[r] = __ldcv(reinterpret_cast<const double*>(x.peers[r] + x.buf_off) + j);
    double sj = 0.0;
#pragma unroll
    for (int r = 0; r < kMaxPeers; ++r)
        if (r < (int)x.nranks) sj += part[r];
    apply_update(u, j, sj);
}

__global__ void __launch_bounds__(256) xchg_decide_update_kernel(const XchgArgs x, const UpdArgs u) { xchg_decide_update_body(x, u); }

// The frames of a group (blockIdx.y = frame): frame f's partial vector and sum(fit^2) lie at buffer + f (V + 1) doubles.
// One flag round trip serves all frames: CTA (0, f) of every frame still running publishes the epoch (the same value,
// so the repeats are harmless), and because the ranks take identical decisions the same frames run on every rank.
__global__ void __launch_bounds__(256) xchg_decide_update_batch_kernel(XchgArgs x, const UpdArgs u, uint64_t ld) {
    const uint32_t f = blockIdx.y;
    x.st_in += f;
    x.st_out += f;
    x.buf_off += (uint64_t)f * (u.V + 1) * sizeof(double);
    xchg_decide_update_body(x, frame_view(u, f, ld));
}

// ---------------------------------------------------------------------------------------------------------------------
// Per-frame prologue / epilogue (replaces pre_iteration_setup, sartsolver_cuda.cpp:138-194, and :264-265).
// ---------------------------------------------------------------------------------------------------------------------
// Frame statistics (max g and sum over g > 0 of g^2: sartsolver_cuda.cpp:147-157, sartsolver.cpp:161-164) in two stages
// and for any number of frames at once -- a whole chunk of a series, the frames of a batch, or one frame:
//   frame_stats_partial_kernel  grid (nb, nframes): CTA (b, f) reduces slice b of frame f (four accumulators per thread,
//                               warp tree, warps in order: a fixed summation order);
//   frame_stats_final_kernel    one CTA: per frame the nb partials in order, then -- with several ranks -- the ranks' values
//                               in rank order straight out of each other's exchange regions (one flag round trip for all
//                               the frames together, like xchg_decide_update_kernel below), so that a frame costs no
//                               ncclAllReduce at all. Without mapped peers the caller all-reduces the two arrays instead.
// They replace one single-CTA pass per frame plus two (series: three) scalar all-reduces per frame.
struct StatPair {
    double gmax, gsq;
};

__global__ void __launch_bounds__(1024) frame_stats_partial_kernel(const double* __restrict__ g, uint64_t P, uint64_t frame_stride,
                                                                    StatPair* __restrict__ part) {
    __shared__ double smax[32], ssum[32];
    const uint32_t nb = gridDim.x, b = blockIdx.x, f = blockIdx.y;
    const uint64_t span = (P + nb - 1) / nb, i0 = min(P, (uint64_t)b * span), i1 = min(P, i0 + span);
    const double* __restrict__ gf = g + (uint64_t)f * frame_stride;
    double mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY}, sm[4] = {0.0, 0.0, 0.0, 0.0};
    const uint64_t stride = blockDim.x;
    uint64_t i = i0 + threadIdx.x;
    for (; i + 3 * stride < i1; i += 4 * stride) {
        double v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) v[u] = gf[i + u * stride];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            mx[u] = fmax(mx[u], v[u]);
            sm[u] += (v[u] > 0) ? v[u] * v[u] : 0.0;
        }
    }
    for (; i < i1; i += stride) {
        const double v = gf[i];
        mx[0] = fmax(mx[0], v);
        sm[0] += (v > 0) ? v * v : 0.0;
    }
    double m = warp_max(fmax(fmax(mx[0], mx[1]), fmax(mx[2], mx[3])));
    double s = warp_sum((sm[0] + sm[1]) + (sm[2] + sm[3]));
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    if (lane == 0) {
        smax[warp] = m;
        ssum[warp] = s;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        m = smax[0];
        s = ssum[0];
        for (int k = 1; k < nw; ++k) {
            m = fmax(m, smax[k]);
            s += ssum[k];
        }
        part[(uint64_t)f * nb + b] = {m, s};
    }
}

struct StatXchg {
    char* const* peers;  // [nranks] exchange regions as mapped here; null: this rank's values only
    uint32_t nranks, rank;
    unsigned long long epoch;  // of this exchange; slots alternate with its parity
    uint64_t flag_off;         // byte offset of the statistics flags [nranks] (kFlagStride apart) inside a region
    uint64_t slot_off[2];      // byte offsets of the two slot arrays [kStatSlots] StatPair
};

__global__ void __launch_bounds__(1024) frame_stats_final_kernel(const StatPair* __restrict__ part, uint32_t nb, uint32_t nframes,
                                                                  double* __restrict__ gmax, double* __restrict__ gsq, const StatXchg x) {
    // this rank's value of frame f = threadIdx.x (nframes <= kStatSlots <= blockDim.x)
    const uint32_t f = threadIdx.x;
    double m = -INFINITY, s = 0.0;
    if (f < nframes) {
        const StatPair* p = part + (uint64_t)f * nb;
        for (uint32_t b = 0; b < nb; ++b) {
            m = fmax(m, p[b].gmax);
            s += p[b].gsq;
        }
    }
    if (x.peers) {
        StatPair* mine = reinterpret_cast<StatPair*>(x.peers[x.rank] + x.slot_off[x.epoch & 1]);
        if (f < nframes) mine[f] = {m, s};
        __threadfence_system();
        __syncthreads();
        if (f < x.nranks) {
            st_release_sys_u64(reinterpret_cast<unsigned long long*>(x.peers[f] + x.flag_off + (uint64_t)x.rank * kFlagStride), x.epoch);
            const unsigned long long* flag =
                reinterpret_cast<const unsigned long long*>(x.peers[x.rank] + x.flag_off + (uint64_t)f * kFlagStride);
            const long long t0 = clock64();
            while (ld_acquire_sys_u64(flag) < x.epoch)
                if (clock64() - t0 > 8000000000LL) break;  // a dead peer: the solve's own exchange reports it
        }
        __syncthreads();
        if (f < nframes) {
            m = -INFINITY;
            s = 0.0;
            for (uint32_t r = 0; r < x.nranks; ++r) {  // rank order: every rank gets the same bits
                const StatPair* theirs = reinterpret_cast<const StatPair*>(x.peers[r] + x.slot_off[x.epoch & 1]);
                const double pm = __ldcv(&theirs[f].gmax), ps = __ldcv(&theirs[f].gsq);
                m = fmax(m, pm);
                s += ps;
            }
        }
    }
    if (f < nframes) {
        gmax[f] = m;
        gsq[f] = s;
    }
}

// weights_mode: 0 none, 1 initial guess CPU semantics (w = g), 2 initial guess GPU semantics (w = max(g, 0)),
//               3 log solver's back-projected measurement (w = [g >= 0] g / lambda)
// stat_max / stat_sum: this frame's max g and sum over g > 0 of g^2 over all ranks (frame_stats_final_kernel)
__global__ void frame_prepare_kernel(const double* __restrict__ gd, uint64_t P, FrameScalars* fs,
                                     const double* __restrict__ stat_max, const double* __restrict__ stat_sum,
                                     const double* __restrict__ ray_length, double h_l, float* __restrict__ g,
                                     float* __restrict__ rinv, float* __restrict__ w, int weights_mode) {
    const double gmax = *stat_max;
    const double norm = (gmax > 0) ? gmax : 1.0;
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) {
        fs->gmax = gmax;
        fs->gsq = *stat_sum;
        fs->norm = norm;
        fs->m2 = *stat_sum / (norm * norm);
    }
    if (i >= P) return;
    const float gi = (float)(gd[i] / norm);
    const double lam = ray_length[i];
    const float ri = (lam > h_l) ? (float)(1.0 / lam) : 0.f;
    g[i] = gi;
    rinv[i] = ri;
    if (weights_mode == 1) w[i] = gi;
    else if (weights_mode == 2) w[i] = (gi > 0.f) ? gi : 0.f;
    else if (weights_mode == 3) w[i] = (gi >= 0.f) ? gi * ri : 0.f;
}

// Series look-ahead (sart_solve_series): the NEXT frame's measurement in the CURRENT frame's normalisation, which the
// last sweep of the current frame back-projects instead of its own measurement, and the factor that converts the update
// it produces into the next frame's normalisation. Saturated pixels stay negative (the norm is positive).
__global__ void series_lookahead_kernel(const double* __restrict__ g_next, uint64_t P, const FrameScalars* cur,
                                        const double* __restrict__ next_gmax, float* __restrict__ g_spec, double* __restrict__ scale) {
    const double norm = cur->norm;
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) *scale = norm / ((*next_gmax > 0) ? *next_gmax : 1.0);
    if (i < P) g_spec[i] = (float)(g_next[i] / norm);
}

// Starting solution. from_guess: x = caller's solution / norm (sartsolver_cuda.cpp:176-178); else the default
// initial guess x_j = [rho_j > h_d] s_j / rho_j from the back-projected measurement (Eq. 4; sartsolver.cpp:149-157).
// The lower clamp of the solver, if any, is max(floor_norm, floor_orig / norm): GPU semantics 1e-7 in normalised
// units (sartsolver_cuda.cpp:180), CPU log solver 1e-100 in the caller's units (sartsolver.cpp:263).
__global__ void start_solution_kernel(const double* __restrict__ sred, const double* __restrict__ xio,
                                      const double* __restrict__ ray_density, double h_d, const FrameScalars* fs,
                                      double floor_norm, double floor_orig, int from_guess, double* __restrict__ xd,
                                      float* __restrict__ xf, uint64_t V, const SeriesCtl* series = nullptr) {
    if (series && series->handed) return;  // the previous frame of the series has already produced this frame's x_1
    const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= V) return;
    double x;
    if (from_guess) x = xio[j] / fs->norm;
    else {
        const double rho = ray_density[j];
        x = (rho > h_d) ? sred[j] / rho : 0.0;
    }
    const double floor_value = fmax(floor_norm, floor_orig / fs->norm);
    if (floor_value > 0 && x < floor_value) x = floor_value;
    xd[j] = x;
    xf[j] = (float)x;
}

__global__ void mask_obs_kernel(const double* __restrict__ sred, const double* __restrict__ ray_density, double h_d,
                                double* __restrict__ obs, uint64_t V) {
    const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j < V) obs[j] = (ray_density[j] > h_d) ? sred[j] : 0.0;
}

__global__ void init_state_kernel(IterState* st0, IterState* st1) {
    IterState s;
    s.conv_prev = 0.0;
    s.k = 0;
    s.done = 0;
    s.status = -1;
    s.iterations = 0;
    s.final_buf = 0;
    s.pad = 0;
    *st0 = s;
    *st1 = s;
}

// xio_j = norm * x_final_j (sartsolver_cuda.cpp:264-265)
__global__ void finish_solution_kernel(const double* __restrict__ xd0, const double* __restrict__ xd1,
                                       const IterState* st, const FrameScalars* fs, double* __restrict__ xio,
                                       uint64_t V) {
    const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= V) return;
    const double* src = st->final_buf ? xd1 : xd0;
    xio[j] = fs->norm * src[j];
}

// ---------------------------------------------------------------------------------------------------------------------
// Once-per-matrix sums in fp64 (replaces the two host loops of the base constructor, sartsolver.cpp:38-56) and the
// fp64 forward projection used to synthesise frames. Not on the per-iteration path.
// ---------------------------------------------------------------------------------------------------------------------
// out_i = sum_j H_ij * (x ? x_j : 1), one warp per row
__global__ void forward_f64_kernel(const float* __restrict__ H, uint64_t ld4, uint64_t P, uint64_t V,
                                   const double* __restrict__ x, double* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const uint64_t row = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= P) return;
    const float4* __restrict__ p = reinterpret_cast<const float4*>(H) + row * ld4;
    double acc = 0.0;
    for (uint64_t c = lane; c < ld4; c += 32) {
        const float4 h = ldg_stream(p + c);
        const uint64_t j = 4 * c;
        if (x) {
            // columns past nvoxel hold zeros in H; x is only nvoxel long
            acc += (double)h.x * (j + 0 < V ? x[j + 0] : 0.0);
            acc += (double)h.y * (j + 1 < V ? x[j + 1] : 0.0);
            acc += (double)h.z * (j + 2 < V ? x[j + 2] : 0.0);
            acc += (double)h.w * (j + 3 < V ? x[j + 3] : 0.0);
        } else {
            acc += (double)h.x + (double)h.y + (double)h.z + (double)h.w;
        }
    }
    acc = warp_sum(acc);
    if (lane == 0) out[row] = acc;
}

// part[slab][j] = sum over the slab's rows of H_ij, thread per column
__global__ void col_sums_f64_kernel(const float* __restrict__ H, uint64_t ld, uint64_t P, uint64_t V,
                                    uint32_t rows_per_slab, double* __restrict__ part) {
    const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= V) return;
    const uint64_t r0 = (uint64_t)blockIdx.y * rows_per_slab;
    const uint64_t r1 = min(P, r0 + rows_per_slab);
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
    const float* __restrict__ p = H + j;
    uint64_t r = r0;
    for (; r + 4 <= r1; r += 4) {
        const float h0 = __ldg(p + (r + 0) * ld), h1 = __ldg(p + (r + 1) * ld);
        const float h2 = __ldg(p + (r + 2) * ld), h3 = __ldg(p + (r + 3) * ld);
        a0 += (double)h0;
        a1 += (double)h1;
        a2 += (double)h2;
        a3 += (double)h3;
    }
    for (; r < r1; ++r) a0 += (double)__ldg(p + r * ld);
    part[(uint64_t)blockIdx.y * V + j] = (a0 + a1) + (a2 + a3);
}

__global__ void reduce_f64_partials_kernel(const double* __restrict__ part, uint32_t nslabs, uint64_t V,
                                           double* __restrict__ out) {
    const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= V) return;
    double s = 0.0;
    for (uint32_t k = 0; k < nslabs; ++k) s += part[(uint64_t)k * V + j];
    out[j] = s;
}

// ---------------------------------------------------------------------------------------------------------------------
// Synthetic ray-transfer matrix: one splitmix64 draw per element keyed by the GLOBAL flattened index, an exact
// integer < 2^24 scaled by one power of two (bit-identical on any rank count and to the CPU generator the tests use).
// ---------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t splitmix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

__device__ __forceinline__ float synth_element(uint64_t seed, uint64_t flat, float scale) {
    const uint64_t h = splitmix64(seed ^ (flat * 0xD1342543DE82EF95ULL));
    uint32_t q = (uint32_t)(h >> 48) | 1u;
    if ((h & 63u) == 0) q += (uint32_t)((h >> 6) & 0x7FFFFFu);
    return (float)q * scale;
}

__global__ void fill_synthetic_kernel(float* __restrict__ H, uint64_t ld4, uint64_t P, uint64_t V, uint64_t row_offset,
                                      uint64_t seed, float scale) {
    for (uint64_t row = blockIdx.y; row < P; row += gridDim.y) {
        const uint64_t flat0 = (row_offset + row) * V;
        float4* __restrict__ p = reinterpret_cast<float4*>(H) + row * ld4;
        for (uint64_t c = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; c < ld4; c += (uint64_t)gridDim.x * blockDim.x) {
            const uint64_t j = 4 * c;
            float4 v;
            v.x = (j + 0 < V) ? synth_element(seed, flat0 + j + 0, scale) : 0.f;
            v.y = (j + 1 < V) ? synth_element(seed, flat0 + j + 1, scale) : 0.f;
            v.z = (j + 2 < V) ? synth_element(seed, flat0 + j + 2, scale) : 0.f;
            v.w = (j + 3 < V) ? synth_element(seed, flat0 + j + 3, scale) : 0.f;
            p[c] = v;
        }
    }
}

// "Line of sight" matrix (SART_SYNTHETIC_LOS): every row is one ray through the nx x nz voxel grid -- a band a few
// voxels wide around a straight line, drawn per row from the hash of the GLOBAL row index -- over a weak dense
// background ("reflections", 8 bits per element). Integer arithmetic only, so the CPU generator of the tests produces
// the same bits: with (u, w) = (j / nz, j % nz) the voxel of column j,
//     direct = max(0, 2048 - |a u + b w - c|) / 2        |a| + |b| = 1024, the line passes through a voxel (u0, w0)
//     q      = direct * 1024 + (bg | 1)                   q < 2^21, exact in fp32
// and H_ij = q * 2^-e. Unlike the dense kind this one behaves like a tomography problem: a warm-started frame needs
// some ten iterations at a tolerance of 1e-6 and ends within a few per cent of the emission it was projected from.
struct LosRay {
    long long a, b, c;
};
__device__ __forceinline__ LosRay los_ray(uint64_t seed, uint64_t row, uint32_t nx, uint32_t nz) {
    const uint64_t h = splitmix64((seed ^ 0xA5A5A5A55A5A5A5AULL) ^ (row * 0xD1342543DE82EF95ULL));
    const uint64_t h2 = splitmix64(h);
    LosRay r;
    r.a = (long long)(h & 2047u) - 1023;
    const long long mag = 1024 - (r.a < 0 ? -r.a : r.a);
    r.b = ((h >> 11) & 1u) ? mag : -mag;
    const long long u0 = (long long)((h2 & 0xFFFFFFu) % nx), w0 = (long long)(((h2 >> 24) & 0xFFFFFFu) % nz);
    r.c = r.a * u0 + r.b * w0;
    return r;
}
__device__ __forceinline__ float los_element(const LosRay& r, uint64_t seed, uint64_t flat, uint32_t j, uint32_t nz, float scale) {
    const long long u = j / nz, w = j % nz;
    long long d = r.a * u + r.b * w - r.c;
    d = d < 0 ? -d : d;
    const uint32_t direct = d < 2048 ? (uint32_t)(2048 - d) >> 1 : 0u;
    const uint32_t bg = (uint32_t)(splitmix64((seed + 1) ^ (flat * 0xD1342543DE82EF95ULL)) >> 56) | 1u;
    return (float)(direct * 1024u + bg) * scale;
}

__global__ void fill_los_kernel(float* __restrict__ H, uint64_t ld4, uint64_t P, uint64_t V, uint64_t row_offset,
                                uint64_t seed, float scale, uint32_t nx, uint32_t nz) {
    for (uint64_t row = blockIdx.y; row < P; row += gridDim.y) {
        const uint64_t flat0 = (row_offset + row) * V;
        const LosRay ray = los_ray(seed, row_offset + row, nx, nz);
        float4* __restrict__ p = reinterpret_cast<float4*>(H) + row * ld4;
        for (uint64_t c = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; c < ld4; c += (uint64_t)gridDim.x * blockDim.x) {
            const uint32_t j = (uint32_t)(4 * c);
            float4 v;
            v.x = (j + 0 < V) ? los_element(ray, seed, flat0 + j + 0, j + 0, nz, scale) : 0.f;
            v.y = (j + 1 < V) ? los_element(ray, seed, flat0 + j + 1, j + 1, nz, scale) : 0.f;
            v.z = (j + 2 < V) ? los_element(ray, seed, flat0 + j + 2, j + 2, nz, scale) : 0.f;
            v.w = (j + 3 < V) ? los_element(ray, seed, flat0 + j + 3, j + 3, nz, scale) : 0.f;
            p[c] = v;
        }
    }
}

}  // namespace sart
