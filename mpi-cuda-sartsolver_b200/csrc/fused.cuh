// Fused sweep: forward projection AND back projection from ONE pass over the matrix in HBM.
//
// The two sweeps of an iteration read the same matrix; what separates them is only that the back projection of row
// i needs w_i, which needs the complete dot product of row i. A row of 50 000 - 200 000 voxels (200 - 800 KB) does not
// fit one SM, but it fits a thread-block CLUSTER: the C CTAs of a cluster split the columns into C panels, each CTA
// streams its panel of the cluster's current row into shared memory with one bulk-async copy (TMA, mbarrier
// completion), multiplies it with its slice of x (held in registers), the per-CTA partial dot products are
// exchanged through distributed shared memory, every thread forms the same fit_i / w_i in the same order, and the row
// panel -- still in shared memory -- is accumulated into the CTA's slice of s = H^T w (held in registers).
// HBM is read once per iteration instead of twice; shared memory is read twice instead.
//
//   producer warp (one lane): per row: wait empty[stage] -> arm the row's dot barrier (expect_tx 4*C) -> expect_tx
//       (full[stage]) -> cp.async.bulk of the panel (50 KB at 50 000 voxels, 4 CTAs)
//   consumer warps, software-pipelined so that the exchange latency of row t hides behind row t+1:
//       FP(0); for t: { FP(t+1); FIN(t) }
//       FP(t):  wait full[stage] ; partial dot from smem x registers ; warp sum ; lane 0 stores it into the CTA's
//               warp-partial slot and arrives on the CTA-local barrier wbar[slot]
//       FIN(t): wait dot[slot] ; add the C per-CTA sums in rank order (one or a few broadcast 128-bit loads: the same
//               additions in every thread of the cluster, so fit_i is identical everywhere) ; w_i ;
//               s += row * w_i from the same smem stage ; arrive empty[stage]
//   reducer warp: per row: wait wbar[slot] ; fixed-order sum of the CTA's warp partials ; lanes q < C send it into
//       CTA q's slot [t mod 2S][rank] with st.async ... mbarrier::complete_tx (DSMEM). One message per CTA pair and row:
//       an earlier version let every warp send to every CTA (C x NW messages per row and CTA), which cost 2 % at C = 4
//       and 20 % at C = 16. No fence anywhere: a plain remote store + mbarrier.arrive.release.cluster is a
//       MEMBAR.ALL.GPU per row per warp and made the first version 2x slower than two separate sweeps.
// No __syncthreads in the loop: warps only meet at mbarriers. Slot reuse (pipeline depth D): a CTA reaches FP(t) only
// after its own FIN(t-S), which needed every peer's row t-S, which those peers produced after their FIN(t-S-D-1); with
// 2S slots and S >= D+1 nothing for row t can land before every CTA has consumed row t-2S, and the producer re-arms a
// slot when it re-issues the stage. At the end each CTA writes its slice of the cluster's partial s; the clusters' partials are added in
// fixed order by reduce_partials_kernel, exactly like the slabs of the two-kernel path. Replaces cublasSgemv +
// cublasSdot + PropagateKernel's prologue and main loop (source/sartsolver_cuda.cpp:239-253,
// source/cuda/sart_kernels.cu:63-101) in one launch.
#pragma once

#include "kernels.cuh"

namespace sart {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t cluster_nctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t cluster_id_x() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t cluster_count_x() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%nclusterid.x;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(bar), "r"(parity)
            : "memory");
    } while (!ok);
}
__device__ __forceinline__ uint32_t map_to_cta(uint32_t local_addr, uint32_t cta) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(cta));
    return r;
}
// Asynchronous store into another CTA's shared memory that completes 4 bytes of transaction on that CTA's mbarrier:
// the waiter on the barrier sees the value, and no release fence is needed on the sending side. (A plain remote store
// followed by mbarrier.arrive.release.cluster costs a MEMBAR.ALL.GPU per row per warp, which waits for every
// outstanding global load of the warp -- measured 2 us per row.)
__device__ __forceinline__ void st_async_remote_f32(uint32_t cluster_addr, float v, uint32_t cluster_bar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.f32 [%0], %1, [%2];" ::"r"(cluster_addr),
                 "f"(v), "r"(cluster_bar)
                 : "memory");
}
__device__ __forceinline__ void bulk_load(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}

__device__ __forceinline__ float warp_sum_f32(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL_MASK, v, o);
    return v;
}

struct FusedArgs {
    const float* H;        // [P][ld]
    uint64_t ld4;          // ld / 4
    uint64_t P;
    const float* x;        // [ld] solution, zero beyond nvoxel (the whole-solve kernel takes it from SolveArgs::xf)
    const float* g;        // [P] normalised measurement
    const float* rinv;     // [P] (lambda_i > h_l) ? 1 / lambda_i : 0
    float* part;           // [nclusters][ld] per-cluster partial back projection
    double* f2part;        // [nclusters] per-cluster sum(fit^2)
    const IterState* st;   // may be null
    uint32_t Wc;           // float4 columns per CTA panel (last panel may be narrower)
    uint32_t stages;       // S >= 2
    uint32_t stage_bytes;  // bytes of one stage, multiple of 128
    int log_mode;
};

__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ float lds32(uint32_t addr) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
    return v;
}

// ---- whole solve in one launch --------------------------------------------------------------------------------------
// With PERSIST the kernel does not return after the sweep: every CTA of the grid is resident (the grid is sized by
// cudaOccupancyMaxActiveClusters and launched cooperatively), so the sweeps of one solve are separated by two grid
// barriers instead of three kernel boundaries, and what the separate kernels did in between runs in the same CTAs:
//     sweep t (both projections)                                        -> per-cluster partials of s and sum(fit^2)
//     grid barrier
//     tail: CTA b owns voxels [b * chunk, (b + 1) * chunk): adds the clusters' partials in fixed order (fp64), with
//           several ranks publishes that to the peers and adds the ranks' values in rank order (NVLink peer memory,
//           one flag per CTA and rank: a CTA starts as soon as ITS chunk is complete everywhere), takes the
//           stop / continue decision (every CTA of every rank computes the same one from the same numbers in the same
//           order) and writes x_{t+1} for its voxels
//     grid barrier                                                       -> x_{t+1} visible to every CTA
// The arithmetic and its order are those of reduce_partials_kernel, decide_kernel, update_kernel and
// xchg_decide_update_kernel, so both ways of running a solve give bit-identical results.
// Everything a later sweep reads that an earlier tail wrote goes through L2 (ld.global.cg / .cv): L1 is not coherent
// within a launch. A barrier or a peer that does not arrive within ~4 s ends the solve with SART_STATUS_EXCHANGE_TIMEOUT.
struct SolveArgs {
    IterState* st[2];      // [0] is read on entry; the final state is written to both copies
    float* xf[2];          // fp32 iterate: sweep t reads xf[(t + pflip) & 1], its tail writes the other one
    double* xd[2];         // fp64 master copy, same parity rule
    UpdArgs upd;           // constant fields of the update (xd_cur / xd_next / xf_next / sred / st are set per sweep)
    double tol;
    int max_iter, fixed;
    unsigned long long* gbar;  // grid barrier counter, zero at launch
    uint64_t V, ld;
    char* const* peers;    // [nranks] exchange regions as mapped here, null with one rank
    uint32_t nranks, rank;
    unsigned long long epoch0;  // sweep t uses epoch0 + t + 1
    uint64_t buf_off[2];   // byte offsets of the two exchange buffers inside a region
    uint64_t flag_off;     // byte offset of the per-CTA flags [nranks][gridDim.x] inside a region
    // Warm-started series (sart_solve_series), series != null: the launch takes its first sweep (1 when the previous
    // frame's last sweep already produced this frame's first update), the buffer flip (buffers are indexed by
    // (t + pflip) & 1) and the exchange epoch from the control block the previous frame left, and predicts its own last
    // sweep t_spec (the previous frame's iteration count, once two consecutive frames agreed to within two iterations, and
    // only when a next frame exists). At sweep t_spec the back projection uses g_spec -- the NEXT frame's measurement in
    // this frame's normalisation -- instead of g: if the solve stops there, as predicted, that sweep has at the same time
    // evaluated this frame's last iterate and done the next frame's first iteration (*out_scale converts it to the next
    // frame's normalisation); if it does not stop, the sweep is run again with this frame's own measurement.
    SeriesCtl* series;
    int has_next;
    const float* g_spec;
    const double* out_scale;  // device scalar: this frame's norm / the next frame's norm
    int* out_status;          // this frame's slots in the series' result arrays (device), or null
    int* out_iterations;
    unsigned long long* exchanges;  // number of tails run (= exchanges over the ranks), written at the end
    unsigned long long* tail_ns;  // [3], zero at launch: ns CTA 0 spent between the end of a sweep (first grid barrier passed) and the start of the next, how often, scratch
    // [kTailPhases], zero at launch: ns CTA 0 spent in each part of the tails (summed over the tails tail_ns[1] counts):
    // waiting at the first grid barrier (= how much later the slowest cluster finished its rows), adding the clusters'
    // partials, the flag round trip with the other ranks, reading the peers' records + update, the second grid barrier
    unsigned long long* phase_ns;
    int prefetch;  // 1: the producer warps run ahead into the next sweep while the tail is being computed
};
constexpr int kTailPhases = 5;

__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

constexpr long long kSpinLimit = 8000000000LL;            // ~4 s at 2 GHz
constexpr unsigned long long kBarrierPoison = 1ull << 40;  // added to the barrier counter by a CTA that gives up

// true: every CTA arrived; false: this or another CTA timed out (the counter is poisoned so nobody waits again)
// (*target: advanced by thread 0 before the call; only thread 0 reads it, behind the first __syncthreads)
__device__ __forceinline__ bool grid_barrier(unsigned long long* ctr, const unsigned long long* target_ptr, int* s_abort) {
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned long long target = *target_ptr;
        __threadfence();
        atomicAdd(ctr, 1ull);
        const long long t0 = clock64();
        for (;;) {
            unsigned long long v;
            asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(ctr) : "memory");
            if (v >= kBarrierPoison) {
                *s_abort = 1;
                break;
            }
            if (v >= target) break;
            if (clock64() - t0 > kSpinLimit) {
                atomicAdd(ctr, kBarrierPoison);
                *s_abort = 1;
                break;
            }
        }
        __threadfence();
    }
    __syncthreads();
    return *s_abort == 0;
}

// What one sweep of a launch leaves for the next one, and what the tails decide about the sweeps to come. It lives in shared
// memory: as registers these ~20 values would stay occupied through the row loop, where the kernel sits at its 128-register
// cap, and the compiler then re-materialises loop invariants (thread and cluster ids, kernel parameters) inside that loop --
// 7 % more instructions per row, which wide rows (fewer SMs streaming) feel one to one. Written by one thread of the role
// that owns a field, read by everybody behind a CTA-wide barrier.
struct SweepCarry {
    double conv_prev;
    int t, t_spec, pflip, t_first, repeated, ahead_prev;
    uint32_t pr_stage, pr_slot, pr_parity;                                  // producer's ring / slot position
    uint32_t rd_slot, rd_parity;                                            // reducer's
    uint32_t fp_stage, fp_parity, fp_slot, fin_stage, fin_slot, fin_parity;  // consumers'
};

// warps 0..NWC-1 consumers, warp NWC TMA producer, warp NWC+1 reducer: 16 warps in all, so that every SM sub-partition
// holds 4 and a thread may use 128 registers (a 17th warp caps the kernel at 96 and spills).
// D = software-pipeline depth of the consumers: FP runs D rows ahead of FIN, so the exchange latency of a row (reducer
// wake-up, DSMEM hop, barrier) is spread over D rows of work. Requires S >= D + 1 (ring) and 2S >= S + D + 1 (slot
// reuse, above).
// shared memory: [S stages][2S x NWC warp partials][2S x 16 per-CTA sums][full[S], empty[S], wbar[2S], dot[2S]]
// All ring / slot / barrier positions advance incrementally (no integer division in the row loop) and carry over
// from one sweep of a PERSIST launch to the next.
template <int NT, int K, int D, bool PERSIST>
__device__ __forceinline__ void fused_body(const FusedArgs& a, const SolveArgs* sv) {
    constexpr uint32_t NWC = NT / 32;
    constexpr uint32_t CSTRIDE = 16;  // floats per slot of per-CTA sums (cluster size <= 16), 64-byte aligned
    // 128-bit shared loads issued back to back before their FMAs (their latency overlaps); bounded by the register budget
    constexpr int G = (K >= 7) ? 1 : (K < 4 ? K : 4);
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ int s_abort;
    __shared__ double s_f2, s_f2r[kMaxPeers];
    // What only the tails between the sweeps need lives in shared memory, not in registers that would stay occupied through
    // the row loop: the grid-barrier target and the exchange epoch (thread 0 advances them before the barrier that opens a
    // tail; everybody reads them behind it), and the series block as the launch found it.
    __shared__ unsigned long long s_bar_target, s_tails, s_epoch0;
    __shared__ SeriesCtl s_ctl;
    __shared__ SweepCarry s_c;

    double conv_prev0 = 0.0;
    if constexpr (PERSIST) {
        const IterState s0 = *sv->st[0];
        if (s0.done) return;
        conv_prev0 = s0.conv_prev;
    } else if (a.st && a.st->done) {
        return;
    }
    const uint32_t S = a.stages;
    const uint32_t rank = cluster_ctarank(), C = cluster_nctarank();
    const uint32_t cid = cluster_id_x(), ncl = cluster_count_x();
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    // The two bases every row needs go through an asm statement: the compiler may then not re-derive them inside the row
    // loop (S2R CgaCtaId + parameter loads + multiply, several times per row) when registers are tight; it keeps them.
    uint32_t ring_u32, wpart_u32;
    asm volatile("mov.u32 %0, %1;" : "=r"(ring_u32) : "r"(smem_u32(smem)));
    asm volatile("mov.u32 %0, %1;" : "=r"(wpart_u32) : "r"(ring_u32 + S * a.stage_bytes));  // [2S][NWC] floats
    const uint32_t cdot_u32 = wpart_u32 + 2 * S * NWC * 4;       // [2S][CSTRIDE] floats
    const uint32_t full_u32 = cdot_u32 + 2 * S * CSTRIDE * 4;
    const uint32_t empty_u32 = full_u32 + 8 * S, wbar_u32 = empty_u32 + 8 * S, dot_u32 = wbar_u32 + 16 * S;

    const uint32_t c0 = rank * a.Wc;
    const uint32_t wcols = ((uint64_t)c0 < a.ld4) ? (uint32_t)min((uint64_t)a.Wc, a.ld4 - c0) : 0u;
    // What the bulk copies and the peers never write must read as zero, so that the row loop needs no column guards:
    // the tail of every stage behind the panel (stage_bytes = K * NT float4) and the per-CTA sum slots of ranks >= C.
    for (uint32_t s = 0; s < S; ++s)
        for (uint32_t o = (wcols + tid) * 16; o < a.stage_bytes; o += blockDim.x * 16)
            asm volatile("st.shared.v4.f32 [%0], {%1, %1, %1, %1};" ::"r"(ring_u32 + s * a.stage_bytes + o), "f"(0.f) : "memory");
    for (uint32_t i = tid; i < 2 * S * CSTRIDE; i += blockDim.x)
        asm volatile("st.shared.f32 [%0], %1;" ::"r"(cdot_u32 + 4 * i), "f"(0.f) : "memory");
    if (tid == 0) {
        s_abort = 0;
        s_bar_target = 0;
        s_tails = 0;
        if constexpr (PERSIST) {
            s_epoch0 = sv->epoch0;
            SeriesCtl c = {};
            if (sv->series) {  // every CTA reads the block long before CTA 0 rewrites it at the end of the solve
                c = *sv->series;
                s_epoch0 = c.epoch;
            }
            s_ctl = c;
        }
        {
            SweepCarry k = {};
            k.conv_prev = conv_prev0;
            k.t_spec = -1;
            k.pr_parity = 1;
            if constexpr (PERSIST) {
                if (sv->series) {
                    if (s_ctl.handed) {
                        k.t_first = 1;
                        k.pflip = s_ctl.pflip;
                    }
                    const int d = s_ctl.it_prev - s_ctl.it_prev2;
                    if (sv->has_next && s_ctl.it_prev >= 2 && s_ctl.it_prev2 >= 2 && d >= -2 && d <= 2) k.t_spec = s_ctl.it_prev;
                }
            }
            k.t = k.t_first;
            s_c = k;
        }
        for (uint32_t s = 0; s < S; ++s) {
            mbar_init(full_u32 + 8 * s, 1);
            mbar_init(empty_u32 + 8 * s, NWC);
        }
        for (uint32_t s = 0; s < 2 * S; ++s) {
            mbar_init(wbar_u32 + 8 * s, NWC);
            mbar_init(dot_u32 + 8 * s, 1);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    cluster_sync_all();

    const uint32_t n = (a.P > cid) ? (uint32_t)((a.P - cid + ncl - 1) / ncl) : 0;
    // The ring / slot positions of the three roles run on from sweep to sweep (s_c, each role loads its own at the start of
    // a sweep and stores them at the end). s_c.ahead_prev: rows of the coming sweep whose copies the producer already issued
    // while the previous tail was running -- the matrix does not depend on the iterate, so the producer simply keeps going
    // into the next sweep's first rows as the consumers release the last stages of this one, and the reduction / exchange /
    // update between two sweeps overlaps with refilling the whole shared-memory ring (S row panels per SM) from HBM.
    // Another sweep can only follow while t < max_iter (the decision of sweep max_iter is always "stop"), and none is
    // expected after the sweep a series predicts to be the frame's last.
    auto runs_ahead = [&](int t, int t_spec) {
        if constexpr (PERSIST) return sv->prefetch && t < sv->max_iter && t != t_spec;
        else return false;
    };
    for (;;) {
        if (warp == NWC) {
            // ---- producer --------------------------------------------------------------------------------------------
            if (lane == 0) {
                uint32_t pr_stage = s_c.pr_stage, pr_slot = s_c.pr_slot, pr_parity = s_c.pr_parity;
                const bool ahead_prev = s_c.ahead_prev != 0;
                const uint32_t bytes = wcols * 16;
                const char* src0 = reinterpret_cast<const char*>(a.H) + ((uint64_t)cid * a.ld4 + c0) * 16;
                const uint64_t src_step = (uint64_t)ncl * a.ld4 * 16;
                auto issue = [&](const char* src) {
                    mbar_wait(empty_u32 + 8 * pr_stage, pr_parity);
                    mbar_expect_tx(dot_u32 + 8 * pr_slot, 4 * C);  // one float from every CTA of the cluster
                    mbar_expect_tx(full_u32 + 8 * pr_stage, bytes);
                    if (bytes) bulk_load(ring_u32 + pr_stage * a.stage_bytes, src, bytes, full_u32 + 8 * pr_stage);
                    if (++pr_stage == S) {
                        pr_stage = 0;
                        pr_parity ^= 1;
                    }
                    if (++pr_slot == 2 * S) pr_slot = 0;
                };
                const uint32_t pre = ahead_prev ? min(S, n) : 0u;
                const char* src = src0 + (uint64_t)pre * src_step;
                for (uint32_t r = pre; r < n; ++r, src += src_step) issue(src);
                if constexpr (PERSIST) {
                    const uint32_t npre = runs_ahead(s_c.t, s_c.t_spec) ? min(S, n) : 0u;
                    src = src0;
                    for (uint32_t r = 0; r < npre; ++r, src += src_step) issue(src);  // the next sweep's first rows
                }
                s_c.pr_stage = pr_stage;
                s_c.pr_slot = pr_slot;
                s_c.pr_parity = pr_parity;
            }
            __syncwarp();
        } else if (warp == NWC + 1) {
            // ---- reducer: CTA-level sum of the warp partials, one message per peer (and frame) ---------------------------
            const uint32_t peer = (lane < C) ? lane : 0;
            const uint32_t peer_cdot0 = map_to_cta(cdot_u32 + rank * 4, peer);
            const uint32_t peer_dot0 = map_to_cta(dot_u32, peer);
            uint32_t rd_slot = s_c.rd_slot, rd_parity = s_c.rd_parity;
            for (uint32_t r = 0; r < n; ++r) {
                mbar_wait(wbar_u32 + 8 * rd_slot, rd_parity);
                float p = (lane < NWC) ? lds32(wpart_u32 + (rd_slot * NWC + lane) * 4) : 0.f;
                p = warp_sum_f32(p);  // fixed tree over the NWC partials
                if (lane < C) st_async_remote_f32(peer_cdot0 + rd_slot * (CSTRIDE * 4), p, peer_dot0 + 8 * rd_slot);
                if (++rd_slot == 2 * S) {
                    rd_slot = 0;
                    rd_parity ^= 1;
                }
            }
            __syncwarp();
            if (lane == 0) {
                s_c.rd_slot = rd_slot;
                s_c.rd_parity = rd_parity;
            }
        } else {
            // ---- consumers -------------------------------------------------------------------------------------------
            const float* xcur = a.x;
            const float* gsrc = a.g;
            if constexpr (PERSIST) {
                const int t = s_c.t;
                xcur = sv->xf[(t + s_c.pflip) & 1];
                if (t == s_c.t_spec) gsrc = sv->g_spec;
            }
            uint32_t fp_stage = s_c.fp_stage, fp_parity = s_c.fp_parity, fp_slot = s_c.fp_slot;
            uint32_t fin_stage = s_c.fin_stage, fin_slot = s_c.fin_slot, fin_parity = s_c.fin_parity;
            const int mine = (wcols > tid) ? (int)((wcols - tid + NT - 1) / NT) : 0;
            float4 xr[K], sacc[K];
            const float4* x4 = reinterpret_cast<const float4*>(xcur) + c0 + tid;
#pragma unroll
            for (int m = 0; m < K; ++m) {
                // a PERSIST launch reads what the previous tail wrote: through L2, never the read-only path
                xr[m] = (m < mine) ? (PERSIST ? __ldcg(x4 + m * NT) : __ldg(x4 + m * NT)) : make_float4(0.f, 0.f, 0.f, 0.f);
                sacc[m] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
            double f2acc = 0.0;
            // Measurements and inverse ray lengths of the cluster's rows are fetched 32 rows at a time: lane l holds row
            // 32j + l of batch j, the batch after the current one is already in flight, and each row's values are
            // broadcast with a shuffle. A per-row load, however early it is issued, ends up waited for within the row
            // (its register is touched when the pipeline queue shifts), which put one exposed L2/HBM latency on every
            // row and capped short rows at ~0.75 us each.
            float g_cur = 0.f, g_nxt = 0.f, r_cur = 0.f, r_nxt = 0.f;
            auto load_batch = [&](uint32_t t0) {
                const uint32_t row = t0 + lane;
                const uint64_t off = (uint64_t)cid + (uint64_t)row * ncl;
                g_nxt = (row < n) ? __ldg(gsrc + off) : 0.f;
                r_nxt = (row < n) ? __ldg(a.rinv + off) : 0.f;
            };
            const uint32_t c4 = (C + 3) / 4;  // 128-bit loads that cover the C per-CTA sums

            auto forward = [&]() {
                mbar_wait(full_u32 + 8 * fp_stage, fp_parity);
                const uint32_t base = ring_u32 + fp_stage * a.stage_bytes + tid * 16;
                float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int m0 = 0; m0 < K; m0 += G) {
                    float4 h[G];
#pragma unroll
                    for (int j = 0; j < G; ++j)
                        if (m0 + j < K) h[j] = lds128(base + (m0 + j) * (NT * 16));
#pragma unroll
                    for (int j = 0; j < G; ++j)
                        if (m0 + j < K) {
                            acc[0] = fmaf(h[j].x, xr[m0 + j].x, acc[0]);
                            acc[1] = fmaf(h[j].y, xr[m0 + j].y, acc[1]);
                            acc[2] = fmaf(h[j].z, xr[m0 + j].z, acc[2]);
                            acc[3] = fmaf(h[j].w, xr[m0 + j].w, acc[3]);
                        }
                }
                const float tot = warp_sum_f32((acc[0] + acc[1]) + (acc[2] + acc[3]));
                if (lane == 0) {
                    asm volatile("st.shared.f32 [%0], %1;" ::"r"(wpart_u32 + (fp_slot * NWC + warp) * 4), "f"(tot) : "memory");
                    mbar_arrive(wbar_u32 + 8 * fp_slot);  // release: the reducer sees the partials
                }
                if (++fp_stage == S) {
                    fp_stage = 0;
                    fp_parity ^= 1;
                }
                if (++fp_slot == 2 * S) fp_slot = 0;
            };

            auto finish = [&](float gi, float ri) {
                mbar_wait(dot_u32 + 8 * fin_slot, fin_parity);
                // the C per-CTA sums in rank order: the same additions in every thread of the cluster
                float fitv = 0.f;
                const uint32_t cd = cdot_u32 + fin_slot * (CSTRIDE * 4);
                for (uint32_t q = 0; q < c4; ++q) {  // slots of ranks >= C hold zeros
                    const float4 v = lds128(cd + q * 16);
                    fitv = (((fitv + v.x) + v.y) + v.z) + v.w;
                }
                float wi = 0.f;
                if (gi >= 0.f) wi = a.log_mode ? fitv * ri : (gi - fitv) * ri;
                f2acc += (double)fitv * (double)fitv;
                const uint32_t base = ring_u32 + fin_stage * a.stage_bytes + tid * 16;
#pragma unroll
                for (int m0 = 0; m0 < K; m0 += G) {
                    float4 h[G];
#pragma unroll
                    for (int j = 0; j < G; ++j)
                        if (m0 + j < K) h[j] = lds128(base + (m0 + j) * (NT * 16));
#pragma unroll
                    for (int j = 0; j < G; ++j)
                        if (m0 + j < K) {
                            sacc[m0 + j].x = fmaf(h[j].x, wi, sacc[m0 + j].x);
                            sacc[m0 + j].y = fmaf(h[j].y, wi, sacc[m0 + j].y);
                            sacc[m0 + j].z = fmaf(h[j].z, wi, sacc[m0 + j].z);
                            sacc[m0 + j].w = fmaf(h[j].w, wi, sacc[m0 + j].w);
                        }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(empty_u32 + 8 * fin_stage);
                if (++fin_stage == S) fin_stage = 0;
                if (++fin_slot == 2 * S) {
                    fin_slot = 0;
                    fin_parity ^= 1;
                }
            };

            // prologue: the first D rows go through FP; then per row: FP(r + D) ; FIN(r)
            load_batch(0);
#pragma unroll
            for (int d = 0; d < D; ++d)
                if ((uint32_t)d < n) forward();
            for (uint32_t r = 0; r < n; ++r) {
                if ((r & 31) == 0) {
                    r_cur = r_nxt;
                    g_cur = g_nxt;
                    load_batch(r + 32);
                }
                const float gi = __shfl_sync(FULL_MASK, g_cur, r & 31);
                const float ri = __shfl_sync(FULL_MASK, r_cur, r & 31);
                if (r + D < n) forward();
                finish(gi, ri);
            }

            float4* __restrict__ out4 = reinterpret_cast<float4*>(a.part) + (uint64_t)cid * a.ld4 + c0 + tid;
#pragma unroll
            for (int m = 0; m < K; ++m)
                if (m < mine) out4[m * NT] = sacc[m];
            if (rank == 0 && tid == 0) a.f2part[cid] = f2acc;
            if (tid == 0) {  // (every consumer thread holds the same positions)
                s_c.fp_stage = fp_stage;
                s_c.fp_parity = fp_parity;
                s_c.fp_slot = fp_slot;
                s_c.fin_stage = fin_stage;
                s_c.fin_slot = fin_slot;
                s_c.fin_parity = fin_parity;
            }
        }
        if constexpr (!PERSIST) {
            break;
        } else {
            // ---- tail of sweep t ----------------------------------------------------------------------------------------
            const uint32_t nblk = gridDim.x, blk = blockIdx.x;
            const bool timer = (blk == 0 && tid == 0);
            unsigned long long tp = timer ? global_timer_ns() : 0ull;  // CTA 0 has finished its own rows
            auto phase_done = [&](int k) {
                if (timer) {
                    const unsigned long long now = global_timer_ns();
                    sv->phase_ns[k] += now - tp;
                    tp = now;
                }
            };
            if (tid == 0) {
                s_bar_target += nblk;
                s_tails += 1;  // one exchange per tail, repeated sweeps included
            }
            bool ok = grid_barrier(sv->gbar, &s_bar_target, &s_abort);
            if (timer) sv->tail_ns[2] = global_timer_ns();  // the slowest cluster has finished its rows
            phase_done(0);
            // the carry as the sweep left it; thread 0 rewrites it at the end of the tail, behind the __syncthreads below
            const int t = s_c.t, pflip = s_c.pflip, t_first = s_c.t_first;
            int t_spec = s_c.t_spec;
            bool repeated = s_c.repeated != 0;
            double conv_prev = s_c.conv_prev;
            const uint64_t V = sv->V, chunk = (V + nblk - 1) / nblk;
            const uint64_t j0 = min(V, (uint64_t)blk * chunk), j1 = min(V, j0 + chunk);
            // this rank's partial of voxel j: the clusters' partials in fixed order (reduce_partials_kernel)
            auto local_sum = [&](uint64_t j) {
                // whole batches of LB unconditional L2 loads in flight, then the rest one by one; added in cluster order
                double sj = 0.0;
                constexpr uint32_t LB = 8;
                const uint64_t stride = sv->ld;
                const float* p = a.part + j;
                uint32_t k = 0;
                for (; k + LB <= ncl; k += LB, p += LB * stride) {
                    float v[LB];
#pragma unroll
                    for (uint32_t i = 0; i < LB; ++i) v[i] = __ldcg(p + i * stride);
#pragma unroll
                    for (uint32_t i = 0; i < LB; ++i) sj += (double)v[i];
                }
                for (; k < ncl; ++k, p += stride) sj += (double)__ldcg(p);
                return sj;
            };
            if (ok && warp == 0) {
                double f = 0.0;
                for (uint32_t k = lane; k < ncl; k += 32) f += __ldcg(a.f2part + k);
                f = warp_sum(f);
                if (lane == 0) s_f2 = f;
            }
            const bool multi = sv->nranks > 1;
            const unsigned long long tails = s_tails, epoch = s_epoch0 + tails;
            const uint64_t rec = (uint64_t)blk * (chunk + 1);  // this CTA's record in an exchange buffer: [sum(fit^2)][chunk]
            if (ok && multi) {
                double* mine = reinterpret_cast<double*>(sv->peers[sv->rank] + sv->buf_off[epoch & 1]) + rec;
                for (uint64_t j = j0 + tid; j < j1; j += blockDim.x) mine[1 + (j - j0)] = local_sum(j);
                __syncthreads();  // s_f2, and this CTA's record complete
                if (tid == 0) mine[0] = s_f2;
                __threadfence_system();
                __syncthreads();
                phase_done(1);
                if (tid < sv->nranks) {
                    const uint64_t fl = sv->flag_off + ((uint64_t)sv->rank * nblk + blk) * 8;
                    st_release_sys_u64(reinterpret_cast<unsigned long long*>(sv->peers[tid] + fl), epoch);
                    const unsigned long long* flag = reinterpret_cast<const unsigned long long*>(
                        sv->peers[sv->rank] + sv->flag_off + ((uint64_t)tid * nblk + blk) * 8);
                    const long long t0 = clock64();
                    while (ld_acquire_sys_u64(flag) < epoch) {
                        if (clock64() - t0 > kSpinLimit) {
                            s_abort = 1;
                            break;
                        }
                    }
                    s_f2r[tid] = __ldcv(reinterpret_cast<const double*>(sv->peers[tid] + sv->buf_off[epoch & 1]) + rec);
                }
            }
            __syncthreads();
            if (multi) phase_done(2);
            if (s_abort) ok = false;
            // decision for x_t (decide_kernel / xchg_decide_update_kernel): identical in every CTA of every rank
            bool stop = false;
            int status = 0;
            if (!ok) {
                atomicAdd(sv->gbar, kBarrierPoison);  // nobody waits for this CTA again
                stop = true;
                status = kStatusExchangeTimeout;
            } else if (repeated) {
                repeated = false;  // sweep t run again after a wrong prediction: "continue" was decided the first time
            } else if (t >= 1) {
                double f2 = s_f2;
                if (multi) {
                    f2 = 0.0;
                    for (uint32_t r = 0; r < sv->nranks; ++r) f2 += s_f2r[r];
                }
                const double m2 = sv->upd.fs->m2;
                const double conv = (m2 - f2) / m2;
                if (!sv->fixed && t - 1 >= 1 && fabs(conv - conv_prev) < sv->tol) {
                    stop = true;
                    status = 0;
                } else if (t >= sv->max_iter) {
                    stop = true;
                    status = -1;
                }
                conv_prev = conv;
            }
            const bool spec = ok && (t == t_spec);
            const bool hand_over = spec && stop;  // predicted stop: this sweep's back projection is the next frame's first
            const bool redo = spec && !stop;      // wrong prediction: the sweep has to be run again with this frame's own g
            if (stop && (blk == 0 || !ok) && tid == 0) {
                IterState o;
                o.conv_prev = conv_prev;
                o.k = t;
                o.done = 1;
                o.status = status;
                o.iterations = t;
                o.final_buf = (t + pflip) & 1;
                o.pad = hand_over ? 1 : 0;
                *sv->st[0] = o;
                *sv->st[1] = o;
                *sv->exchanges = tails;
                if (sv->series) {
                    const SeriesCtl ctl = s_ctl;
                    SeriesCtl c = ctl;
                    c.handed = hand_over ? 1 : 0;
                    c.pflip = ((t + 1 + pflip) & 1) ^ 1;  // the next launch's sweep 1 reads what the hand-over writes
                    c.it_prev2 = ctl.it_prev;
                    c.it_prev = t;
                    c.epoch = epoch;
                    c.passes = ctl.passes + tails;
                    c.frames_handed = ctl.frames_handed + (hand_over ? 1 : 0);
                    c.sweeps_repeated = ctl.sweeps_repeated + ((int)tails > t - t_first + 1 ? 1 : 0);
                    c.timed_out = ctl.timed_out | (ok ? 0 : 1);
                    *sv->series = c;
                    if (sv->out_status) *sv->out_status = status;
                    if (sv->out_iterations) *sv->out_iterations = t;
                }
            }
            // Leaving the loop with copies of the next sweep's first rows still in flight: they must have landed before the
            // CTA may exit (its shared memory is their destination).
            auto drain = [&]() {
                if (tid == 0) {
                    uint32_t st = s_c.fp_stage, par = s_c.fp_parity;  // (thread 0 stored them itself)
                    const uint32_t npre = runs_ahead(t, t_spec) ? min(S, n) : 0u;
                    for (uint32_t i = 0; i < npre; ++i) {
                        mbar_wait(full_u32 + 8 * st, par);
                        if (++st == S) {
                            st = 0;
                            par ^= 1;
                        }
                    }
                }
            };
            if (stop && !hand_over) {
                drain();
                break;
            }
            if (!redo) {
                const double* xd_cur = sv->xd[(t + pflip) & 1];
                double* xd_next = sv->xd[(t + 1 + pflip) & 1];
                float* xf_next = sv->xf[(t + 1 + pflip) & 1];
                const double out_scale = hand_over ? *sv->out_scale : 1.0;
                for (uint64_t j = j0 + tid; j < j1; j += blockDim.x) {
                    double sj;
                    if (multi) {
                        // all ranks' values of this voxel requested before the first one is used: one NVLink round trip
                        // instead of one per rank; added in rank order, so every rank gets the same bits
                        sj = 0.0;
                        for (int r0 = 0; r0 < (int)sv->nranks; r0 += 8) {  // (8 at a time: one batch on a box of 8 GPUs)
                            double pr[8];
#pragma unroll
                            for (int r = 0; r < 8; ++r)
                                if (r0 + r < (int)sv->nranks)
                                    pr[r] = __ldcv(reinterpret_cast<const double*>(sv->peers[r0 + r] + sv->buf_off[epoch & 1]) + rec + 1 + (j - j0));
#pragma unroll
                            for (int r = 0; r < 8; ++r)
                                if (r0 + r < (int)sv->nranks) sj += pr[r];
                        }
                    } else {
                        const unsigned long long d0 = (timer && j == j0) ? global_timer_ns() : 0ull;
                        sj = local_sum(j);
                        if (timer && j == j0) {  // diagnostics (SART_TAIL_DEBUG): thread 0's own sum over the clusters
                            sv->phase_ns[5] += global_timer_ns() - d0;
                            sv->phase_ns[7] += d0 - tp;  // what came before it in this phase (sum(fit^2), decision)
                        }
                    }
                    const unsigned long long d1 = (timer && j == j0) ? global_timer_ns() : 0ull;
                    apply_update_at(sv->upd, xd_cur, xd_next, xf_next, out_scale, j, sj);
                    if (timer && j == j0) sv->phase_ns[6] += global_timer_ns() - d1;  // ... and its own update
                }
            }
            phase_done(multi ? 3 : 1);
            if (hand_over) {
                drain();
                break;
            }
            if (tid == 0) {
                s_bar_target += nblk;
                // the next sweep: t + 1, or -- after a wrong prediction, once the partials of this sweep have been read by
                // everyone (barrier below) -- sweep t again with this frame's own measurement
                s_c.ahead_prev = runs_ahead(t, t_spec) ? 1 : 0;
                s_c.conv_prev = conv_prev;
                s_c.t = redo ? t : t + 1;
                s_c.t_spec = redo ? -1 : t_spec;
                s_c.repeated = redo ? 1 : (repeated ? 1 : 0);
            }
            if (!grid_barrier(sv->gbar, &s_bar_target, &s_abort)) {
                drain();
                if (tid == 0) {
                    IterState o;
                    o.conv_prev = conv_prev;
                    o.k = t;
                    o.done = 1;
                    o.status = kStatusExchangeTimeout;
                    o.iterations = t;
                    o.final_buf = (t + pflip) & 1;
                    o.pad = 0;
                    *sv->st[0] = o;
                    *sv->st[1] = o;
                    *sv->exchanges = s_tails;
                    if (sv->series) {
                        sv->series->timed_out = 1;
                        if (sv->out_status) *sv->out_status = kStatusExchangeTimeout;
                    }
                }
                break;
            }
            phase_done(4);
            if (timer) {
                sv->tail_ns[0] += global_timer_ns() - sv->tail_ns[2];
                sv->tail_ns[1] += 1;
            }
        }
    }
    cluster_sync_all();
}

template <int NT, int K, int D>
__global__ void __launch_bounds__(NT + 64, 1) fused_sweep_kernel(const FusedArgs a) {
    fused_body<NT, K, D, false>(a, nullptr);
}

// the whole solve: sweeps 0 .. stop in one cooperative launch (one frame)
template <int NT, int K, int D>
__global__ void __launch_bounds__(NT + 64, 1) fused_solve_kernel(const FusedArgs a, const SolveArgs sv) {
    fused_body<NT, K, D, true>(a, &sv);
}

}  // namespace sart
