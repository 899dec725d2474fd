// Forward projection of a group of frames (sart_solve_batch) with the matrix staged through a shared-memory ring by
// bulk-async copies, like the fused sweep (fused.cuh), instead of through the register file.
//
// fp_rows_batch_warp_kernel (kernels.cuh) keeps 8 rows x 8 frames of accumulators per thread and needs eight 128-bit
// global loads of the matrix in flight per thread on top of them: 168 registers, 12 warps per SM, and ncu shows it
// latency-bound by occupancy at ~5.6 TB/s. Here the loads in flight cost no registers: a producer thread streams
// [32 rows x 256 columns] tiles of the matrix and the matching [F frames x 256 columns] tile of x into a ring of stages,
// one TMA tensor copy each (cp.async.bulk.tensor.2d over tensor maps of the shard and of the iterate's buffer; the
// first version issued one 1 KB cp.async.bulk per row and was bound by the copy engine's rate of ~25 small copies per
// microsecond and SM: 13.6 ms per pass); eight consumer warps -- four row groups of 8 rows x two column halves of 128
// columns -- multiply a stage out of shared memory alone (per lane and stage: F + 8 128-bit shared loads, 32 F FMAs).
// x comes out of L2 once per stage and serves the stage's 32 rows. Nine warps: at most three per SM sub-partition, so a
// thread may use 168 registers (a 13-warp version was capped at 128 and spilled). Elements outside the tensors (rows
// behind the shard, columns behind ld) arrive as zeros.
//
//   producer (one lane): per stage: wait empty[s] -> expect_tx(full[s], stage bytes) -> the two tensor copies
//   consumer warp (g, q): per stage: wait full[s]; acc[r][f] += H[8 g + r][q] . x[f][q]; arrive empty[s]
//   per row block of 32 rows: shuffle-tree sums inside the warp (lane (r F + f) % 32 ends up with row r, frame f), the
//   halves added in order through shared memory, epilogue (w, sum(fit^2)) by the warps of half 0 -- the epilogue of
//   fp_rows_batch_warp_kernel.
#pragma once

#include <cuda.h>  // CUtensorMap (types only: the encoder is looked up at run time, the library does not link libcuda)

#include "fused.cuh"

namespace sart {

constexpr int kRingRowsPerGroup = 8, kRingRowGroups = 4, kRingHalves = 2;
constexpr int kRingRows = kRingRowsPerGroup * kRingRowGroups;            // rows of a stage
constexpr int kRingCols4 = 32 * kRingHalves;                           // float4 columns of a stage
constexpr int kRingConsumerWarps = kRingRowGroups * kRingHalves;       // 8
constexpr int kRingThreads = 32 * (kRingConsumerWarps + 1);              // + the producer warp
constexpr int kRingRowBytes = kRingCols4 * 16;                           // 1 KB of a row (or of a frame's x) per stage
static_assert(kRingRows <= 256 && kRingCols4 * 4 <= 256, "a box of a tensor map has at most 256 elements per dimension");

// [rows x 256 columns] box of a row-major fp32 tensor into shared memory; coordinates in elements (column, row)
__device__ __forceinline__ void tensor_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t col, uint32_t row, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst),
                 "l"(reinterpret_cast<uint64_t>(map)), "r"(col), "r"(row), "r"(bar)
                 : "memory");
}

template <int F>
__host__ __device__ constexpr uint32_t ring_stage_bytes() {
    return (uint32_t)(kRingRows + F) * kRingRowBytes;
}
// stages that fit: the ring, two mbarriers per stage, the static arrays of the kernel (cross-half sums, sum(fit^2))
template <int F>
inline uint32_t ring_stages(int max_smem) {
    const int reserve = 2 * kRingRowGroups * kRingHalves * 64 * 8 + 1024;
    int s = (max_smem - reserve) / (int)(ring_stage_bytes<F>() + 16);
    return (uint32_t)(s > 8 ? 8 : s);
}
template <int F>
inline uint32_t ring_smem_bytes(uint32_t stages) {
    return stages * (ring_stage_bytes<F>() + 16);
}

// a.nrg = number of row blocks (kRingRows rows each); a.f2part is [F][gridDim.x]
template <int F>
__global__ void __launch_bounds__(kRingThreads, 1)
    fp_rows_batch_ring_kernel(const FpBatchArgs a, uint32_t S, const __grid_constant__ CUtensorMap map_h,
                              const __grid_constant__ CUtensorMap map_x) {
    constexpr int R = kRingRowsPerGroup, G = kRingRowGroups, Q = kRingHalves, NWC = kRingConsumerWarps;
    static_assert((R * F) % 32 == 0 && 32 % F == 0, "whole (row, frame) slots per lane in the epilogue");
    constexpr int PER_LANE = R * F / 32;
    if (a.st) {
        bool all_done = true;
        for (int f = 0; f < a.nframes; ++f) all_done = all_done && a.st[f].done;
        if (all_done) return;
    }
    extern __shared__ __align__(128) unsigned char ring[];
    __shared__ double red[2][G][Q][R * F];  // a row block's sums per column half (two row blocks in flight)
    __shared__ double f2s[G][F];
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr uint32_t stage_bytes = ring_stage_bytes<F>();
    uint32_t ring_u32;  // through an asm statement, so that it is kept in a register instead of being re-derived per stage
    asm volatile("mov.u32 %0, %1;" : "=r"(ring_u32) : "r"(smem_u32(ring)));
    const uint32_t full_u32 = ring_u32 + S * stage_bytes, empty_u32 = full_u32 + 8 * S;
    if (tid == 0) {
        for (uint32_t s = 0; s < S; ++s) {
            mbar_init(full_u32 + 8 * s, 1);
            mbar_init(empty_u32 + 8 * s, NWC);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const uint32_t ld4 = (uint32_t)a.ld4;
    const uint32_t nch = (ld4 + kRingCols4 - 1) / kRingCols4;
    uint32_t stage = 0, parity = 0;

    if (warp == NWC) {
        // ---- producer ---------------------------------------------------------------------------------------------------
        if (lane == 0) {
            parity = 1;  // a fresh "empty" barrier lets the first S waits through
            for (uint32_t rb = blockIdx.x; rb < a.nrg; rb += gridDim.x) {
                for (uint32_t ch = 0; ch < nch; ++ch) {
                    const uint32_t dst = ring_u32 + stage * stage_bytes, bar = full_u32 + 8 * stage;
                    mbar_wait(empty_u32 + 8 * stage, parity);
                    mbar_expect_tx(bar, stage_bytes);  // whole boxes are counted, zero-filled parts included
                    tensor_load_2d(dst, &map_h, ch * (kRingCols4 * 4), rb * kRingRows, bar);
                    tensor_load_2d(dst + kRingRows * kRingRowBytes, &map_x, ch * (kRingCols4 * 4), 0, bar);
                    if (++stage == S) {
                        stage = 0;
                        parity ^= 1;
                    }
                }
            }
        }
        return;  // (the consumers meet at named barrier 1 only)
    }

    // ---- consumers ------------------------------------------------------------------------------------------------------
    const uint32_t g = warp / Q, q = warp % Q;
    const int my_r = lane / F, my_f = lane % F;
    const uint32_t h_off = (g * R * kRingCols4 + q * 32 + lane) * 16;   // this lane's float4 of row 8 g in a stage
    const uint32_t x_off = (kRingRows * kRingCols4 + q * 32 + lane) * 16;  // ... of frame 0's x
    double f2acc = 0.0;  // of frame my_f, this lane's row slots (warps of half 0)
    uint32_t nblock = 0;
    for (uint32_t rb = blockIdx.x; rb < a.nrg; rb += gridDim.x, ++nblock) {
        const uint64_t row0 = (uint64_t)rb * kRingRows + g * R;
        const int nvalid = (row0 < a.P) ? (int)min((uint64_t)R, a.P - row0) : 0;
        float acc[R][F];
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int f = 0; f < F; ++f) acc[r][f] = 0.f;
#pragma unroll 1
        for (uint32_t ch = 0; ch < nch; ++ch) {
            mbar_wait(full_u32 + 8 * stage, parity);
            if (ch * kRingCols4 + q * 32 + lane < ld4) {
                const uint32_t base = ring_u32 + stage * stage_bytes;
                float4 xv[F];
#pragma unroll
                for (int f = 0; f < F; ++f) xv[f] = lds128(base + x_off + f * kRingRowBytes);
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const float4 h = lds128(base + h_off + r * kRingRowBytes);
#pragma unroll
                    for (int f = 0; f < F; ++f) {
                        acc[r][f] = fmaf(h.x, xv[f].x, acc[r][f]);
                        acc[r][f] = fmaf(h.y, xv[f].y, acc[r][f]);
                        acc[r][f] = fmaf(h.z, xv[f].z, acc[r][f]);
                        acc[r][f] = fmaf(h.w, xv[f].w, acc[r][f]);
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(empty_u32 + 8 * stage);
            if (++stage == S) {
                stage = 0;
                parity ^= 1;
            }
        }
        // sums over the warp's 128 columns: lane (r F + f) % 32 keeps (row r, frame f)
        double mine[PER_LANE];
#pragma unroll
        for (int k = 0; k < PER_LANE; ++k) mine[k] = 0.0;
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int f = 0; f < F; ++f) {
                const double s = warp_sum((double)acc[r][f]);
                if ((int)lane == (r * F + f) % 32) mine[(r * F + f) / 32] = s;
            }
        double(*slot)[Q][R * F] = red[nblock & 1];
#pragma unroll
        for (int k = 0; k < PER_LANE; ++k) slot[g][q][k * 32 + lane] = mine[k];
        // all consumer warps: the block's halves are complete (and the other slot of `red` is free again, see below)
        asm volatile("bar.sync 1, %0;" ::"n"(32 * NWC) : "memory");
        if (q == 0) {
#pragma unroll
            for (int k = 0; k < PER_LANE; ++k) {
                double s = 0.0;
#pragma unroll
                for (int qq = 0; qq < Q; ++qq) s += slot[g][qq][k * 32 + lane];
                const int r = my_r + k * (32 / F);
                if (r < nvalid && my_f < a.nframes) {
                    const uint64_t i = row0 + r;
                    const float fitv = (float)s;
                    f2acc += s * s;
                    const float gi = a.g[(uint64_t)my_f * a.P + i];
                    a.w[(uint64_t)my_f * a.P + i] = (gi >= 0.f) ? (a.log_mode ? fitv : gi - fitv) * a.rinv[i] : 0.f;
                }
            }
        }
        // slot reuse: block n + 2 writes red[n & 1] after the barrier of block n + 1, which the half-0 warps only
        // reach once they have read block n's sums
    }
    // sum(fit^2) per frame: over the row slots of a warp (lanes f, f + F, ...), then over the row groups in order
#pragma unroll
    for (int o = F; o < 32; o <<= 1) f2acc += __shfl_xor_sync(FULL_MASK, f2acc, o);
    if (q == 0 && lane < (uint32_t)F) f2s[g][lane] = f2acc;
    asm volatile("bar.sync 1, %0;" ::"n"(32 * NWC) : "memory");
    if ((int)tid < a.nframes) {
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < G; ++k) s += f2s[k][tid];
        a.f2part[(uint64_t)tid * gridDim.x + blockIdx.x] = s;
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Back projection of a group of frames through the same kind of ring: CTA (panel, slab) owns 256 columns and the rows of
// one slab; a stage is a [64 rows x 256 columns] tile of the matrix (one tensor copy) plus the 64 x F weights of those
// rows, which a second helper warp fetches with ordinary loads (w is [F][P] with an arbitrary P: no 16-byte-aligned row
// pitch for a tensor map, and it is 1 / 6000 of the traffic) and stores row-major [row][F], so that a row's F weights
// are two broadcast 128-bit shared loads. Eight consumer warps = two column halves x four row quarters of a stage; a
// lane owns one float4 column x F frames of accumulators (the mapping of bp_cols_batch_smem_kernel); at the end of the
// slab the quarters are added in order through the (by then idle) ring and quarter 0 writes the slab's partial.
// Rows behind the shard arrive as zeros (matrix) and are given zero weights.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int kBpRingRows = 64, kBpRingHalves = 2, kBpRingQuarters = 4;
constexpr int kBpRingCols4 = 32 * kBpRingHalves;                          // float4 columns of a panel
constexpr int kBpRingConsumerWarps = kBpRingHalves * kBpRingQuarters;     // 8
constexpr int kBpRingThreads = 32 * (kBpRingConsumerWarps + 2);           // + tensor-copy warp + weights warp
constexpr int kBpRingRowBytes = kBpRingCols4 * 16;                        // 1 KB
static_assert(kBpRingCols4 == kRingCols4, "both ring kernels use tensor maps with boxes of the same width");

template <int F>
__host__ __device__ constexpr uint32_t bp_ring_stage_bytes() {
    return (uint32_t)kBpRingRows * kBpRingRowBytes + (uint32_t)kBpRingRows * F * 4;
}
template <int F>
inline uint32_t bp_ring_stages(int max_smem) {
    int s = (max_smem - 1024) / (int)(bp_ring_stage_bytes<F>() + 16);
    return (uint32_t)(s > 8 ? 8 : s);
}
template <int F>
inline uint32_t bp_ring_smem_bytes(uint32_t stages) {
    return stages * (bp_ring_stage_bytes<F>() + 16);
}

// grid (panels of 256 columns, slabs); a.rows_per_slab is a multiple of kBpRingRows
template <int F>
__global__ void __launch_bounds__(kBpRingThreads, 1)
    bp_cols_batch_ring_kernel(const BpBatchArgs a, uint32_t S, const __grid_constant__ CUtensorMap map_h) {
    static_assert(F == 8, "a row's weights are two float4");
    constexpr int H = kBpRingHalves, Q = kBpRingQuarters, NWC = kBpRingConsumerWarps, RQ = kBpRingRows / Q;  // 16 rows per quarter
    static_assert((Q - 1) * H * 32 * F * 16 <= kBpRingRows * kBpRingRowBytes, "the quarters' accumulators fit one stage's tile");
    if (a.st) {
        bool all_done = true;
        for (int f = 0; f < a.nframes; ++f) all_done = all_done && a.st[f].done;
        if (all_done) return;
    }
    extern __shared__ __align__(128) unsigned char ring[];
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr uint32_t stage_bytes = bp_ring_stage_bytes<F>(), w_off = kBpRingRows * kBpRingRowBytes;
    uint32_t ring_u32;
    asm volatile("mov.u32 %0, %1;" : "=r"(ring_u32) : "r"(smem_u32(ring)));
    const uint32_t full_u32 = ring_u32 + S * stage_bytes, empty_u32 = full_u32 + 8 * S;
    if (tid == 0) {
        for (uint32_t s = 0; s < S; ++s) {
            mbar_init(full_u32 + 8 * s, 2);  // the tensor copy's expect_tx and the weights warp
            mbar_init(empty_u32 + 8 * s, NWC);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const uint64_t r0 = (uint64_t)blockIdx.y * a.rows_per_slab;
    const uint64_t r1 = min(a.P, r0 + a.rows_per_slab);
    const uint32_t nst = (r1 > r0) ? (uint32_t)((r1 - r0 + kBpRingRows - 1) / kBpRingRows) : 0;
    uint32_t stage = 0, parity = 0;

    if (warp == NWC) {
        // ---- tensor copies ----------------------------------------------------------------------------------------------
        if (lane == 0) {
            parity = 1;
            for (uint32_t st = 0; st < nst; ++st) {
                const uint32_t bar = full_u32 + 8 * stage;
                mbar_wait(empty_u32 + 8 * stage, parity);
                mbar_expect_tx(bar, kBpRingRows * kBpRingRowBytes);
                tensor_load_2d(ring_u32 + stage * stage_bytes, &map_h, blockIdx.x * (kBpRingCols4 * 4),
                               (uint32_t)(r0 + (uint64_t)st * kBpRingRows), bar);
                if (++stage == S) {
                    stage = 0;
                    parity ^= 1;
                }
            }
        }
        return;
    }
    if (warp == NWC + 1) {
        // ---- weights: lane l fetches rows l and l + 32 of the stage, all frames, and stores them [row][F] -----------------
        parity = 1;
        for (uint32_t st = 0; st < nst; ++st) {
            const uint64_t row = r0 + (uint64_t)st * kBpRingRows + lane;
            float wl[2][F];
#pragma unroll
            for (int k = 0; k < 2; ++k)
#pragma unroll
                for (int f = 0; f < F; ++f)
                    wl[k][f] = (row + 32 * k < r1 && f < a.nframes) ? __ldcg(a.w + (uint64_t)f * a.P + row + 32 * k) : 0.f;
            mbar_wait(empty_u32 + 8 * stage, parity);
            const uint32_t dst = ring_u32 + stage * stage_bytes + w_off + lane * (F * 4);
#pragma unroll
            for (int k = 0; k < 2; ++k)
#pragma unroll
                for (int f4 = 0; f4 < F; f4 += 4)
                    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(dst + k * 32 * (F * 4) + f4 * 4), "f"(wl[k][f4]),
                                 "f"(wl[k][f4 + 1]), "f"(wl[k][f4 + 2]), "f"(wl[k][f4 + 3])
                                 : "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(full_u32 + 8 * stage);  // release: the consumers see the weights
            if (++stage == S) {
                stage = 0;
                parity ^= 1;
            }
        }
        return;
    }

    // ---- consumers ------------------------------------------------------------------------------------------------------
    const uint32_t h = warp % H, rq = warp / H;
    const uint32_t col_off = (h * 32 + lane) * 16;  // this lane's float4 inside a row of the tile
    float4 acc[F];
#pragma unroll
    for (int f = 0; f < F; ++f) acc[f] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
    for (uint32_t st = 0; st < nst; ++st) {
        mbar_wait(full_u32 + 8 * stage, parity);
        const uint32_t base = ring_u32 + stage * stage_bytes;
#pragma unroll
        for (int r4 = 0; r4 < RQ; r4 += 4) {
            float4 hv[4], wlo[4], whi[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t row = rq * RQ + r4 + k;
                hv[k] = lds128(base + row * kBpRingRowBytes + col_off);
                wlo[k] = lds128(base + w_off + row * (F * 4));
                whi[k] = lds128(base + w_off + row * (F * 4) + 16);
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float wv[F] = {wlo[k].x, wlo[k].y, wlo[k].z, wlo[k].w, whi[k].x, whi[k].y, whi[k].z, whi[k].w};
#pragma unroll
                for (int f = 0; f < F; ++f) {
                    acc[f].x = fmaf(hv[k].x, wv[f], acc[f].x);
                    acc[f].y = fmaf(hv[k].y, wv[f], acc[f].y);
                    acc[f].z = fmaf(hv[k].z, wv[f], acc[f].z);
                    acc[f].w = fmaf(hv[k].w, wv[f], acc[f].w);
                }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(empty_u32 + 8 * stage);
        if (++stage == S) {
            stage = 0;
            parity ^= 1;
        }
    }
    // the four row quarters of every column, added in order: through the first stage's tile, which nobody fills any more
    asm volatile("bar.sync 1, %0;" ::"n"(32 * NWC) : "memory");  // every consumer has left its last stage
    if (rq > 0) {
#pragma unroll
        for (int f = 0; f < F; ++f) {
            const uint32_t dst = ring_u32 + ((((rq - 1) * H + h) * F + f) * 32 + lane) * 16;
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "f"(acc[f].x), "f"(acc[f].y), "f"(acc[f].z), "f"(acc[f].w)
                         : "memory");
        }
    }
    asm volatile("bar.sync 1, %0;" ::"n"(32 * NWC) : "memory");
    const uint64_t c4 = (uint64_t)blockIdx.x * kBpRingCols4 + h * 32 + lane;
    if (rq == 0 && c4 < a.ld4) {
        const uint64_t nslabs = gridDim.y;
#pragma unroll
        for (int f = 0; f < F; ++f) {
            float4 s4 = acc[f];
#pragma unroll
            for (int k = 0; k < Q - 1; ++k) {
                const float4 v = lds128(ring_u32 + (((k * H + h) * F + f) * 32 + lane) * 16);
                s4.x += v.x;
                s4.y += v.y;
                s4.z += v.z;
                s4.w += v.w;
            }
            if (f < a.nframes) reinterpret_cast<float4*>(a.part)[((uint64_t)f * nslabs + blockIdx.y) * a.ld4 + c4] = s4;
        }
    }
}

}  // namespace sart
