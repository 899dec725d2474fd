// libsart_b200.so -- host side of the engine and the C ABI declared in include/sart_b200.h.
//
// What the reference spreads over BaseSARTSolverMPICuda's constructor, pre_iteration_setup and the two solve()
// bodies (source/sartsolver_cuda.cpp:78-354) lives here as one engine object per process (= per GPU = per row
// shard). All device memory is allocated once per engine (the reference allocates and frees five vectors per
// frame, sartsolver_cuda.cpp:160-191, 223, 267-271); a solve is a fixed sequence of launches on one stream with the
// convergence decision taken on the device, polled by the host once per batch of sweeps.
// There is no CPU fallback: without a CUDA device sart_create fails.
#include <fcntl.h>
#include <unistd.h>

#include <algorithm>
#include <cerrno>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <mutex>
#include <numeric>
#include <string>
#include <vector>

#include "../../include/sart_b200.h"
#include "kernels.cuh"
#include "fused.cuh"
#include "batch_ring.cuh"
#include "nccl_dyn.h"

using namespace sart;

namespace {

thread_local std::string g_error;

int fail(int code, const std::string& msg) {
    g_error = msg;
    return code;
}

#define CUDA_TRY(expr)                                                                                       \
    do {                                                                                                     \
        cudaError_t err__ = (expr);                                                                          \
        if (err__ != cudaSuccess)                                                                            \
            return fail(SART_ERR_CUDA, std::string("CUDA error: ") + cudaGetErrorString(err__) + " at " +    \
                                           __FILE__ + ":" + std::to_string(__LINE__) + " (" #expr ")");      \
    } while (0)

#define NCCL_TRY(api, expr)                                                                                  \
    do {                                                                                                     \
        ncclResult_t r__ = (expr);                                                                           \
        if (r__ != ncclSuccess)                                                                              \
            return fail(SART_ERR_NCCL, std::string("NCCL error: ") + (api)->GetErrorString(r__) + " at " +   \
                                           __FILE__ + ":" + std::to_string(__LINE__));                       \
    } while (0)

#define SART_TRY(expr)          \
    do {                        \
        int rc__ = (expr);      \
        if (rc__) return rc__;  \
    } while (0)

constexpr int kProfiledSweeps = 64;  // sweeps per solve whose kernels are bracketed by events
constexpr int kSolveCtlWords = 16;   // control words of the whole-solve kernel (barrier counter, timers, counters)
constexpr int kBpThreads = 256;
constexpr int kFpThreads = 256;

int env_int(const char* name, int dflt) {
    const char* v = std::getenv(name);
    return (v && *v) ? std::atoi(v) : dflt;
}

inline uint64_t ceil_div(uint64_t a, uint64_t b) { return (a + b - 1) / b; }

// True when a CUDA profiler / tools library has been injected into this process (Nsight Compute, Nsight Systems,
// compute-sanitizer: they all come in through CUDA_INJECTION64_PATH or leave NV_COMPUTE_PROFILER_* / NV_NSIGHT_* /
// NSYS_* variables in the environment of the process they launch). Looked up once.
bool profiler_attached() {
    static const bool attached = [] {
        const int forced = env_int("SART_ASSUME_PROFILER", -1);
        if (forced >= 0) return forced != 0;
        static const char* const prefixes[] = {"CUDA_INJECTION64_PATH=", "CUDA_INJECTION32_PATH=", "NV_COMPUTE_PROFILER_",
                                               "NV_NSIGHT_", "NSYS_PROFILING_SESSION_ID=", "NV_SANITIZER_", "NVTX_INJECTION64_PATH="};
        for (char** v = environ; v && *v; ++v)
            for (const char* p : prefixes)
                if (std::strncmp(*v, p, std::strlen(p)) == 0) return true;
        return false;
    }();
    return attached;
}

// messages of the reference's setters, source/sartsolver.cpp:63-123
int check_params(const sart_config& c) {
    if (!(c.ray_density_threshold >= 0)) return fail(SART_ERR_INVALID_ARGUMENT, "Ray density threshold must be non-negative.");
    if (!(c.ray_length_threshold >= 0)) return fail(SART_ERR_INVALID_ARGUMENT, "Ray length threshold must be non-negative.");
    if (!(c.conv_tolerance > 0)) return fail(SART_ERR_INVALID_ARGUMENT, "Convolution tolerance must be positive.");
    if (!(c.beta_laplace >= 0)) return fail(SART_ERR_INVALID_ARGUMENT, "Attribute beta_laplace must be non-negative.");
    if (!(c.relaxation > 0) || c.relaxation > 1.) return fail(SART_ERR_INVALID_ARGUMENT, "Attribute relaxation must be within (0, 1] interval.");
    if (c.max_iterations <= 0) return fail(SART_ERR_INVALID_ARGUMENT, "Attribute max_iterations must be positive.");
    return SART_OK;
}

}  // namespace

struct sart_engine {
    sart_config cfg{};
    int device = 0, sm_count = 0;
    cudaStream_t stream = nullptr;
    uint64_t row_offset = 0, P = 0, V = 0, ld = 0, ld4 = 0;

    float* dH = nullptr;
    double *d_rho = nullptr, *d_lambda = nullptr;
    bool finalized = false;

    std::vector<uint64_t> lap_idx;
    std::vector<float> lap_val;
    int *d_lap_rowptr = nullptr, *d_lap_col = nullptr;
    float* d_lap_val = nullptr;
    uint64_t lap_nnz = 0;

    double *d_gd = nullptr, *d_xio = nullptr, *d_obs = nullptr, *d_sred = nullptr, *d_f2part = nullptr, *d_f2 = nullptr;
    double* d_xd[2] = {nullptr, nullptr};
    float* d_xf[2] = {nullptr, nullptr};
    float *d_g = nullptr, *d_rinv = nullptr, *d_fit = nullptr, *d_w = nullptr, *d_part = nullptr;
    FrameScalars* d_fs = nullptr;
    IterState* d_st[2] = {nullptr, nullptr};
    IterState* h_st = nullptr;  // pinned, 2 slots

    // launch geometry
    uint32_t bp_panels = 0, bp_slabs = 0, bp_rows_per_slab = 0;
    uint32_t fp_grid = 0, fp_nrg = 0;
    int fp_variant = 0, bp_variant = 0, fp_rows = 8;
    int batch = 1;
    // fused sweep (cluster kernel): geometry, 0 clusters = not available for this shape / device
    bool fused = false;
    uint32_t fu_cluster = 0, fu_wc = 0, fu_stages = 0, fu_stage_bytes = 0, fu_smem = 0, fu_nclusters = 0;
    void (*fu_kernel)(const sart::FusedArgs) = nullptr;
    uint32_t fu_threads = 0;
    // whole solve in one cooperative launch (fused_solve_kernel); falls back to one launch per sweep when unavailable
    bool persistent = false;
    void (*fs_kernel)(const sart::FusedArgs, const sart::SolveArgs) = nullptr;
    uint32_t fs_nclusters = 0;
    unsigned long long* d_solve_ctl = nullptr;  // [0] grid barrier counter, [1..3] tail timer, [4] tails run, [8..12] tail phases
    unsigned long long* h_solve_ctl = nullptr;  // pinned copy of the control words after a solve
    int last_prologue_sweeps = 0;               // matrix passes of the last frame's set-up (initial guess, log "obs")
    // warm-started series (sart_solve_series): look-ahead buffers for the next frame, allocated on first use
    double* d_series_in = nullptr;    // [capacity][P] the chunk's measurements, uploaded once
    float* d_gspec = nullptr;
    double* d_spec_scale = nullptr;
    SeriesCtl* d_series_ctl = nullptr;
    double* d_series_out = nullptr;   // [capacity][V] solutions of a chunk, copied to the caller at the end
    int* d_series_res = nullptr;      // [2][capacity] status, iterations
    int32_t series_capacity = 0;
    uint64_t xchg_cta_flag_off = 0;             // per-CTA flags of the in-kernel exchange inside a region
    // frame statistics (max g, sum g^2) of up to stat_capacity frames: per-CTA partials, reduced values, exchange over peers
    StatPair* d_stat_part = nullptr;
    double *d_stat_max = nullptr, *d_stat_sum = nullptr;
    int32_t stat_capacity = 0;
    uint32_t stat_blocks = 1;
    uint64_t xchg_stat_flag_off = 0, xchg_stat_slot_off[2] = {0, 0};
    unsigned long long stat_epoch = 0;

    cudaEvent_t ev_solve0 = nullptr, ev_solve1 = nullptr, ev_it0 = nullptr;
    cudaEvent_t ev_batch[2] = {nullptr, nullptr};
    cudaEvent_t ev_prof[kProfiledSweeps][4];  // before/after FP, before/after BP
    bool profiling = true;

    ncclComm_t comm = nullptr;
    // peer-memory exchange (one region per rank, mapped by every rank through CUDA IPC); see xchg_decide_update_kernel
    char* d_xchg = nullptr;             // this rank's region
    char** d_peers = nullptr;           // device array [nranks]: every rank's region as mapped here
    std::vector<void*> peer_mappings;   // what cudaIpcOpenMemHandle returned (to close)
    uint64_t xchg_buf_off[2] = {0, 0};
    unsigned long long xchg_epoch = 0;
    bool peer_exchange = false;
    // several frames per matrix pass (sart_solve_batch): geometry + buffers for kGroup frames, allocated on first use
    bool group_planned = false, group_ok = false;
    uint32_t gp_fp_grid = 0, gp_bp_slabs = 0, gp_bp_rows_per_slab = 0;
    int group_frames = 8, gp_fp_rows = 8, gp_fp_threads = 384;  // frames per matrix pass; rows per thread / threads per CTA of the batched forward kernel
    uint32_t gp_fp_ring_stages = 0, gp_fp_ring_smem = 0;         // > 0: the shared-memory-ring forward kernel (batch_ring.cuh) runs
    CUtensorMap gp_map_h, gp_map_x[2];                           // its tensor maps: the shard, the two buffers of the group's iterates
    uint32_t gp_bp_ring_stages = 0, gp_bp_ring_smem = 0;         // > 0: the ring back-projection kernel runs
    CUtensorMap gp_map_h_bp;                                     // the shard in boxes of kBpRingRows rows
    double *d_pgd = nullptr, *d_pxio = nullptr, *d_psred = nullptr, *d_pf2part = nullptr, *d_pf2sum = nullptr, *d_pobs = nullptr;
    double* d_pxd[2] = {nullptr, nullptr};   // [parity] -> [kGroup][V]
    float* d_pxf[2] = {nullptr, nullptr};    // [parity] -> [kGroup][ld]
    float *d_pg = nullptr, *d_pw = nullptr, *d_ppart = nullptr;
    FrameScalars* d_pfs = nullptr;           // [kGroup]
    IterState* d_pst[2] = {nullptr, nullptr};  // [parity] -> [kGroup]
    IterState* h_pst = nullptr;              // pinned, 2 slots x kGroup frames
    // pinned staging ring of the file loader (allocated on first use)
    static constexpr int kStagingBuffers = 3;
    char* h_staging[kStagingBuffers] = {nullptr, nullptr, nullptr};
    cudaEvent_t ev_staging[kStagingBuffers] = {nullptr, nullptr, nullptr};
    uint64_t staging_bytes = 0;
    bool measurement_staged = false, guess_staged = false;
    int launches = 0;
    sart_stats stats{};

    sart_engine() {
        for (auto& s : ev_prof) s[0] = s[1] = s[2] = s[3] = nullptr;
    }
};

namespace {

template <typename T>
int dev_alloc(T** p, uint64_t count, bool zero = true) {
    CUDA_TRY(cudaMalloc(reinterpret_cast<void**>(p), std::max<uint64_t>(count, 1) * sizeof(T)));
    if (zero) CUDA_TRY(cudaMemset(*p, 0, std::max<uint64_t>(count, 1) * sizeof(T)));
    return SART_OK;
}

// ---- launchers ------------------------------------------------------------------------------------------------------

typedef void (*FpKernel)(const FpArgs);
typedef void (*BpKernel)(const BpArgs);

struct FpVariant {
    FpKernel fn;
    int rows;
    const char* name;
};
struct BpVariant {
    BpKernel fn;
    int threads;
    const char* name;
};

const FpVariant kFpVariants[] = {
    {fp_rows_kernel<8, kFpThreads, 2>, 8, "R8xU2"},
    {fp_rows_kernel<4, kFpThreads, 4>, 4, "R4xU4"},
    {fp_rows_kernel<8, kFpThreads, 1>, 8, "R8xU1"},
    {fp_rows_kernel<4, kFpThreads, 2>, 4, "R4xU2"},
};
const BpVariant kBpVariants[] = {
    {bp_cols_kernel<kBpThreads, 8>, kBpThreads, "T256xU8"},
    {bp_cols_kernel<kBpThreads, 16>, kBpThreads, "T256xU16"},
    {bp_cols_kernel<128, 8>, 128, "T128xU8"},
    {bp_cols_kernel<128, 16>, 128, "T128xU16"},
};
constexpr int kNumFpVariants = sizeof(kFpVariants) / sizeof(kFpVariants[0]);
constexpr int kNumBpVariants = sizeof(kBpVariants) / sizeof(kBpVariants[0]);

int plan_launches(sart_engine* e) {
    e->fp_variant = std::min(std::max(env_int("SART_FP_VARIANT", 0), 0), kNumFpVariants - 1);
    e->bp_variant = std::min(std::max(env_int("SART_BP_VARIANT", 0), 0), kNumBpVariants - 1);
    const FpVariant& fv = kFpVariants[e->fp_variant];
    const BpVariant& bv = kBpVariants[e->bp_variant];

    int fp_occ = 0, bp_occ = 0;
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&fp_occ, fv.fn, kFpThreads, 0));
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bp_occ, bv.fn, bv.threads, 0));
    fp_occ = std::max(fp_occ, 1);
    bp_occ = std::max(bp_occ, 1);

    e->fp_rows = fv.rows;
    e->fp_nrg = (uint32_t)ceil_div(e->P, fv.rows);
    const int fp_ctas_per_sm = std::min(env_int("SART_FP_CTAS_PER_SM", fp_occ), 32);  // d_f2part holds 32 per SM
    e->fp_grid = (uint32_t)std::min<uint64_t>(e->fp_nrg, (uint64_t)e->sm_count * std::max(fp_ctas_per_sm, 1));

    // back projection: one wave of (panels x slabs) CTAs, slabs a multiple of 32 rows
    e->bp_panels = (uint32_t)ceil_div(e->ld4, bv.threads);
    const uint64_t target = (uint64_t)e->sm_count * std::max(env_int("SART_BP_CTAS_PER_SM", bp_occ), 1);
    uint64_t slabs = std::max<uint64_t>(1, target / e->bp_panels);
    const int forced = env_int("SART_BP_SLABS", 0);
    if (forced > 0) slabs = (uint64_t)forced;
    slabs = std::min<uint64_t>(slabs, ceil_div(e->P, 32));
    e->bp_rows_per_slab = (uint32_t)(ceil_div(ceil_div(e->P, slabs), 32) * 32);
    e->bp_slabs = (uint32_t)ceil_div(e->P, e->bp_rows_per_slab);

    // sweeps per status poll: enough work per poll that the host round trip hides behind the next batch
    // (from the global shape, so that every rank enqueues -- and therefore joins collectives -- in identical batches)
    const double rows_per_rank = std::ceil((double)e->cfg.npixel / (double)e->cfg.nranks);
    const double sweep_us = 2.0 * rows_per_rank * (double)e->ld * 4.0 / 6.0e6;  // at ~6 TB/s
    e->batch = std::min(16, std::max(1, (int)std::ceil(400.0 / std::max(sweep_us, 1.0))));
    e->batch = std::max(1, env_int("SART_BATCH", e->batch));
    return SART_OK;
}

typedef void (*FusedKernel)(const FusedArgs);

// 14 consumer warps + producer + reducer = 16 warps (see fused.cuh). K = float4 columns a consumer thread owns (its
// slices of x and s live in registers): the smallest K that covers the panel is instantiated.
constexpr int kFusedKMax = 8;
constexpr int kFusedThreads = 448;  // consumer threads
// pipeline depth: 3 rows ahead when the ring holds 8 stages (short rows), else 1 (long rows, HBM-bound anyway)
inline int fused_depth(uint32_t stages) {
    const int want = env_int("SART_FUSED_DEPTH", stages >= 8 ? 3 : 1);
    return (want >= 3 && stages >= 4) ? 3 : 1;  // the ring must hold the rows in flight plus one being loaded
}
template <int K>
FusedKernel pick_fused_kernel(int k, int depth) {
    if constexpr (K > kFusedKMax) return nullptr;
    else if (k <= K) return depth == 3 ? fused_sweep_kernel<kFusedThreads, K, 3> : fused_sweep_kernel<kFusedThreads, K, 1>;
    else return pick_fused_kernel<K + 1>(k, depth);
}
typedef void (*SolveKernel)(const FusedArgs, const SolveArgs);
template <int K>
SolveKernel pick_solve_kernel(int k, int depth) {
    if constexpr (K > kFusedKMax) return nullptr;
    else if (k <= K) return depth == 3 ? fused_solve_kernel<kFusedThreads, K, 3> : fused_solve_kernel<kFusedThreads, K, 1>;
    else return pick_solve_kernel<K + 1>(k, depth);
}
// Geometry of the fused sweep. A cluster of C CTAs splits a row into C column panels; each CTA keeps its panel of x and
// s in registers (<= K float4 per thread) and as many shared-memory stages of one row panel as fit. Every cluster size
// from the smallest that can hold the row up to 16 is a candidate -- not only powers of two: how many clusters of C
// the device keeps resident depends on how C packs into its GPCs (148 SMs in GPCs of 16/18/20: clusters of 4 use 132
// SMs, of 16 only 112), so the plan asks cudaOccupancyMaxActiveClusters for each candidate, for the sweep kernel and
// the whole-solve kernel alike, and takes the one that keeps the most SMs streaming (ties: the smaller cluster, whose
// rows need fewer exchange messages). SART_FUSED_CLUSTER forces a size. The choice depends on the shape and the
// device model only, so every rank of a job plans the same grid.
// Leaves e->fused false when the shape or the device cannot run it.
struct FusedCandidate {
    uint32_t C = 0, wc = 0, S = 0, stage_bytes = 0, smem = 0;
    int k = 0, nclusters = 0, nclusters_solve = 0;
    FusedKernel fk = nullptr;
    SolveKernel sk = nullptr;
    std::string why;  // set when the candidate cannot run
};

FusedCandidate probe_fused(sart_engine* e, uint32_t C, int max_smem) {
    FusedCandidate c;
    c.C = C;
    const uint32_t nt = kFusedThreads, nw = nt / 32, extra_warps = 2;  // + producer + reducer
    c.wc = (uint32_t)ceil_div(e->ld4, C);
    c.k = (int)ceil_div(c.wc, nt);
    // a stage holds K x NT float4 per row panel: the kernel reads all of it and zero-fills what the copy leaves out
    c.stage_bytes = (uint32_t)c.k * nt * 16;
    const uint32_t s_cap = (uint32_t)std::max(2, env_int("SART_FUSED_STAGES", 8));
    for (uint32_t s = s_cap; s >= 2; --s) {
        // stages + warp partials [2S][NWC] + per-CTA sums [2S][16] + barriers full[S], empty[S], wbar[2S], dot[2S]
        const uint64_t need = (uint64_t)s * c.stage_bytes + (uint64_t)2 * s * nw * 4 + (uint64_t)2 * s * 16 * 4 + (uint64_t)6 * s * 8;
        if (need + 512 <= (uint64_t)max_smem) {  // + the kernel's static shared variables (384 bytes)
            c.S = s;
            c.smem = (uint32_t)need;
            break;
        }
    }
    if (!c.S) {
        c.why = "a row panel does not fit twice into shared memory";
        return c;
    }
    c.fk = pick_fused_kernel<1>(c.k, fused_depth(c.S));
    if (!c.fk) {
        c.why = "no kernel instance for this panel width";
        return c;
    }
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = C;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.gridDim = dim3(C * (uint32_t)e->sm_count, 1, 1);
    cfg.blockDim = dim3(nt + 32 * extra_warps, 1, 1);
    cfg.dynamicSmemBytes = c.smem;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaError_t err = cudaFuncSetAttribute(c.fk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c.smem);
    if (err == cudaSuccess && C > 8) err = cudaFuncSetAttribute(c.fk, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    if (err == cudaSuccess) err = cudaOccupancyMaxActiveClusters(&c.nclusters, c.fk, &cfg);
    if (err != cudaSuccess || c.nclusters <= 0) {
        (void)cudaGetLastError();
        c.why = std::string("the device cannot co-schedule a cluster of ") + std::to_string(C) + " CTAs (" +
                (err != cudaSuccess ? cudaGetErrorString(err) : "0 active clusters") + ")";
        c.nclusters = 0;
        return c;
    }
    // the whole-solve kernel needs more registers for its tail: on narrow panels fewer of its CTAs fit an SM
    c.sk = pick_solve_kernel<1>(c.k, fused_depth(c.S));
    if (c.sk && cudaFuncSetAttribute(c.sk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c.smem) == cudaSuccess &&
        (C <= 8 || cudaFuncSetAttribute(c.sk, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess) &&
        cudaOccupancyMaxActiveClusters(&c.nclusters_solve, c.sk, &cfg) == cudaSuccess && c.nclusters_solve > 0) {
        // both kernels run the smaller grid: same clusters, same partial sums, bit-identical results
    } else {
        c.sk = nullptr;
        c.nclusters_solve = 0;
    }
    (void)cudaGetLastError();
    return c;
}

int plan_fused(sart_engine* e) {
    e->fused = false;
    const int mode = e->cfg.sweep_mode;
    if (mode == SART_SWEEP_SPLIT) return SART_OK;
    const uint64_t max_wc = (uint64_t)kFusedThreads * kFusedKMax;
    const uint32_t kMaxCluster = 16;
    auto unavailable = [&](const std::string& why) {
        if (mode == SART_SWEEP_FUSED) return fail(SART_ERR_INVALID_ARGUMENT, "SART_SWEEP_FUSED is not available: " + why);
        return (int)SART_OK;
    };
    const uint64_t c_need = ceil_div(e->ld4, max_wc);
    if (c_need > kMaxCluster) return unavailable("nvoxel exceeds what 16 CTAs can hold in registers");
    const uint32_t c_min = (uint32_t)c_need;
    // long rows only, unless forced: short rows make the per-row exchange dominate
    const uint32_t min_wc = (uint32_t)env_int("SART_FUSED_MIN_WC", 1024);
    if (mode == SART_SWEEP_AUTO && ceil_div(e->ld4, c_min) < min_wc) return SART_OK;
    int max_smem = 0;
    CUDA_TRY(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, e->device));

    const int forced_c = env_int("SART_FUSED_CLUSTER", 0);
    FusedCandidate best;
    std::string first_reason;
    auto sms_used = [&](const FusedCandidate& c) {
        const int n = c.nclusters_solve > 0 ? std::min(c.nclusters, c.nclusters_solve) : c.nclusters;
        return (uint64_t)c.C * std::min<uint64_t>((uint64_t)n, e->P);
    };
    if (forced_c > 0 && (uint32_t)forced_c >= c_min && (uint32_t)forced_c <= kMaxCluster) {
        best = probe_fused(e, (uint32_t)forced_c, max_smem);
        first_reason = best.why;
    } else {
        // panels narrower than half the widest one (or than the short-row limit) are not worth their exchange traffic
        const uint32_t c_max = std::min<uint32_t>(kMaxCluster, std::max(c_min, 2 * c_min - 1));
        for (uint32_t C = c_min; C <= c_max; ++C) {
            if (C > c_min && ceil_div(e->ld4, C) < min_wc) break;
            FusedCandidate c = probe_fused(e, C, max_smem);
            if (!c.nclusters) {
                if (first_reason.empty()) first_reason = c.why;
                continue;
            }
            if (!best.nclusters || sms_used(c) > sms_used(best)) best = c;
        }
    }
    if (!best.nclusters) return unavailable(first_reason.empty() ? "no cluster size fits this shape" : first_reason);

    // the probes of other candidates may have left another shared-memory size on the same kernel instance
    CUDA_TRY(cudaFuncSetAttribute(best.fk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)best.smem));
    if (best.sk) CUDA_TRY(cudaFuncSetAttribute(best.sk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)best.smem));
    int nclusters = best.nclusters;
    const int forced_n = env_int("SART_FUSED_CLUSTERS", 0);
    if (forced_n > 0) nclusters = std::min(nclusters, forced_n);
    e->fu_kernel = best.fk;
    e->fu_threads = kFusedThreads + 64;
    e->fu_cluster = best.C;
    e->fu_wc = best.wc;
    e->fu_stages = best.S;
    e->fu_stage_bytes = best.stage_bytes;
    e->fu_smem = best.smem;
    e->fu_nclusters = (uint32_t)std::min<uint64_t>((uint64_t)nclusters, e->P);
    e->fused = true;

    // Whole solve in one cooperative launch (fused_solve_kernel), when the device can co-schedule its grid.
    e->persistent = false;
    int coop = 0;
    CUDA_TRY(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, e->device));
    if (best.sk) {
        e->fu_nclusters = std::min(e->fu_nclusters, (uint32_t)best.nclusters_solve);
        e->fs_kernel = best.sk;
        e->fs_nclusters = e->fu_nclusters;
        e->persistent = coop && env_int("SART_PERSISTENT", 1) != 0;
        if (e->persistent && !e->d_solve_ctl) {
            SART_TRY(dev_alloc(&e->d_solve_ctl, kSolveCtlWords));
            CUDA_TRY(cudaMallocHost(reinterpret_cast<void**>(&e->h_solve_ctl), kSolveCtlWords * sizeof(unsigned long long)));
        }
    }
    (void)cudaGetLastError();
    return SART_OK;
}

int launch_fused(sart_engine* e, const float* x, const IterState* st) {
    FusedArgs a;
    a.H = e->dH;
    a.ld4 = e->ld4;
    a.P = e->P;
    a.x = x;
    a.g = e->d_g;
    a.rinv = e->d_rinv;
    a.part = e->d_part;
    a.f2part = e->d_f2part;
    a.st = st;
    a.Wc = e->fu_wc;
    a.stages = e->fu_stages;
    a.stage_bytes = e->fu_stage_bytes;
    a.log_mode = e->cfg.logarithmic;
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = e->fu_cluster;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.gridDim = dim3(e->fu_cluster * e->fu_nclusters, 1, 1);
    cfg.blockDim = dim3(e->fu_threads, 1, 1);
    cfg.dynamicSmemBytes = e->fu_smem;
    cfg.stream = e->stream;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    CUDA_TRY(cudaLaunchKernelEx(&cfg, e->fu_kernel, a));
    ++e->launches;
    return SART_OK;
}

int launch_fp(sart_engine* e, const float* x, const IterState* st) {
    FpArgs a;
    a.H = e->dH;
    a.ld4 = e->ld4;
    a.P = e->P;
    a.x = x;
    a.g = e->d_g;
    a.rinv = e->d_rinv;
    a.fit = e->d_fit;
    a.w = e->d_w;
    a.f2part = e->d_f2part;
    a.st = st;
    a.nrg = e->fp_nrg;
    a.log_mode = e->cfg.logarithmic;
    kFpVariants[e->fp_variant].fn<<<e->fp_grid, kFpThreads, 0, e->stream>>>(a);
    CUDA_TRY(cudaGetLastError());
    ++e->launches;
    return SART_OK;
}

int launch_bp(sart_engine* e, const IterState* st) {
    BpArgs a;
    a.H = e->dH;
    a.ld4 = e->ld4;
    a.P = e->P;
    a.w = e->d_w;
    a.part = e->d_part;
    a.rows_per_slab = e->bp_rows_per_slab;
    a.st = st;
    const BpVariant& bv = kBpVariants[e->bp_variant];
    bv.fn<<<dim3(e->bp_panels, e->bp_slabs), bv.threads, 0, e->stream>>>(a);
    CUDA_TRY(cudaGetLastError());
    ++e->launches;
    return SART_OK;
}

int launch_reduce(sart_engine* e, const IterState* st, uint32_t nslabs = 0, double* dst = nullptr,
                  const double* f2part = nullptr, uint32_t nf2 = 0) {
    const uint32_t blocks = (uint32_t)ceil_div(e->V, 256);
    reduce_partials_kernel<<<blocks, 256, 0, e->stream>>>(e->d_part, nslabs ? nslabs : e->bp_slabs, e->ld, e->V,
                                                          dst ? dst : e->d_sred, st, f2part, nf2);
    CUDA_TRY(cudaGetLastError());
    ++e->launches;
    return SART_OK;
}

int allreduce_f64(sart_engine* e, double* buf, uint64_t count, ncclRedOp_t op) {
    if (e->cfg.nranks <= 1) return SART_OK;
    NcclApi* api = nccl_api();
    if (!api || !e->comm) return fail(SART_ERR_STATE, "nranks > 1 but sart_comm_init has not been called");
    NCCL_TRY(api, api->AllReduce(buf, buf, count, ncclDouble, op, e->comm, e->stream));
    return SART_OK;
}

// One sweep for the iterate x_t: forward projection, decision, and -- unless t is the last possible sweep, where
// the decision can only be "stop" -- back projection, reduction and update. Kernels after the decision test the
// freshly written state, so nothing but the forward projection runs once the solve is over.
UpdArgs make_update_args(sart_engine* e, int t, const IterState* st) {
    UpdArgs u;
    u.sred = e->d_sred;
    u.xd_cur = e->d_xd[t & 1];
    u.xd_next = e->d_xd[(t + 1) & 1];
    u.xf_next = e->d_xf[(t + 1) & 1];
    u.ray_density = e->d_rho;
    u.obs = e->d_obs;
    u.lap_rowptr = e->lap_nnz ? e->d_lap_rowptr : nullptr;
    u.lap_col = e->d_lap_col;
    u.lap_val = e->d_lap_val;
    u.fs = e->d_fs;
    u.st = st;
    u.V = e->V;
    u.beta = e->cfg.beta_laplace;
    u.relax = e->cfg.relaxation;
    u.h_d = e->cfg.ray_density_threshold;
    u.eps_norm = (e->cfg.semantics == SART_SEMANTICS_GPU) ? 1.e-7 : 0.0;
    u.eps_orig = (e->cfg.semantics == SART_SEMANTICS_GPU) ? 0.0 : 1.e-100;
    u.log_mode = e->cfg.logarithmic;
    u.log_caller_units = (e->cfg.semantics == SART_SEMANTICS_GPU) ? 0 : 1;
    u.out_scale = 1.0;
    return u;
}

int launch_update(sart_engine* e, int t, const IterState* st) {
    update_kernel<<<(uint32_t)ceil_div(e->V, 256), 256, 0, e->stream>>>(make_update_args(e, t, st));
    CUDA_TRY(cudaGetLastError());
    ++e->launches;
    return SART_OK;
}

// reduction over ranks + decision + update in one kernel over peer memory (fused sweep, several ranks, IPC available)
int launch_xchg_update(sart_engine* e, int t, const IterState* st_in, IterState* st_out) {
    XchgArgs x;
    x.peers = e->d_peers;
    x.nranks = (uint32_t)e->cfg.nranks;
    x.rank = (uint32_t)e->cfg.rank;
    x.epoch = e->xchg_epoch;
    x.buf_off = e->xchg_buf_off[e->xchg_epoch & 1];
    x.st_in = st_in;
    x.st_out = st_out;
    x.tol = e->cfg.conv_tolerance;
    x.max_iter = e->cfg.max_iterations;
    x.fixed = e->cfg.fixed_iterations;
    x.t = t;
    xchg_decide_update_kernel<<<(uint32_t)ceil_div(e->V, 256), 256, 0, e->stream>>>(x, make_update_args(e, t, nullptr));
    CUDA_TRY(cudaGetLastError());
    ++e->launches;
    return SART_OK;
}

// The whole solve (sweeps 0 .. stop) in ONE cooperative launch of fused_solve_kernel: reduction over clusters and ranks,
// decision and update run inside the kernel between two grid barriers. cudaErrorCooperativeLaunchTooLarge or an
// unsupported attribute combination is reported as "not launched" so that the caller can fall back.
struct SeriesLaunch {  // one frame of sart_solve_series (see SolveArgs)
    SeriesCtl* ctl = nullptr;
    int has_next = 0;
    int* out_status = nullptr;
    int* out_iterations = nullptr;
};

int launch_solve(sart_engine* e, bool* launched, const SeriesLaunch& series = SeriesLaunch()) {
    *launched = false;
    FusedArgs a;
    a.H = e->dH;
    a.ld4 = e->ld4;
    a.P = e->P;
    a.x = nullptr;
    a.g = e->d_g;
    a.rinv = e->d_rinv;
    a.part = e->d_part;
    a.f2part = e->d_f2part;
    a.st = nullptr;
    a.Wc = e->fu_wc;
    a.stages = e->fu_stages;
    a.stage_bytes = e->fu_stage_bytes;
    a.log_mode = e->cfg.logarithmic;
    SolveArgs sv;
    for (int k = 0; k < 2; ++k) {
        sv.st[k] = e->d_st[k];
        sv.xf[k] = e->d_xf[k];
        sv.xd[k] = e->d_xd[k];
    }
    sv.upd = make_update_args(e, 0, nullptr);
    sv.tol = e->cfg.conv_tolerance;
    sv.max_iter = e->cfg.max_iterations;
    sv.fixed = e->cfg.fixed_iterations;
    sv.gbar = e->d_solve_ctl;
    sv.tail_ns = e->d_solve_ctl + 1;
    sv.V = e->V;
    sv.ld = e->ld;
    const bool multi = e->cfg.nranks > 1;
    sv.peers = multi ? e->d_peers : nullptr;
    sv.nranks = (uint32_t)e->cfg.nranks;
    sv.rank = (uint32_t)e->cfg.rank;
    sv.epoch0 = e->xchg_epoch;
    sv.buf_off[0] = e->xchg_buf_off[0];
    sv.buf_off[1] = e->xchg_buf_off[1];
    sv.flag_off = e->xchg_cta_flag_off;
    sv.series = series.ctl;
    sv.has_next = series.has_next;
    sv.out_status = series.out_status;
    sv.out_iterations = series.out_iterations;
    sv.g_spec = e->d_gspec;
    sv.out_scale = e->d_spec_scale;
    sv.exchanges = e->d_solve_ctl + 4;
    sv.phase_ns = e->d_solve_ctl + 8;
    static_assert(8 + kTailPhases <= kSolveCtlWords, "control words");
    sv.prefetch = env_int("SART_PREFETCH", 1) != 0 ? 1 : 0;
    CUDA_TRY(cudaMemsetAsync(e->d_solve_ctl, 0, kSolveCtlWords * sizeof(unsigned long long), e->stream));
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = e->fu_cluster;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeCooperative;
    attr[1].val.cooperative = 1;
    cfg.gridDim = dim3(e->fu_cluster * e->fs_nclusters, 1, 1);
    cfg.blockDim = dim3(e->fu_threads, 1, 1);
    cfg.dynamicSmemBytes = e->fu_smem;
    cfg.stream = e->stream;
    cfg.attrs = attr;
    // Without the cooperative attribute the grid is the same and still fits the device at once
    // (cudaOccupancyMaxActiveClusters), but nothing checks that the GPU is not shared; a CTA that never becomes
    // resident ends the solve through the barrier timeout instead of a launch error. The attribute is dropped
    //   * on request (SART_COOPERATIVE=0),
    //   * when a profiler has injected itself into the process (Nsight Compute 2025.2 cannot replay a cooperative
    //     cluster launch: "LaunchFailed", and the process dies) -- so `ncu python ...` over any caller just works,
    //   * for a second attempt when the driver refuses the cooperative launch.
    // If that fails as well the caller falls back to one launch per sweep.
    const bool coop = env_int("SART_COOPERATIVE", 1) != 0 && !profiler_attached();
    cfg.numAttrs = coop ? 2 : 1;
    cudaError_t err = cudaLaunchKernelEx(&cfg, e->fs_kernel, a, sv);
    if (err != cudaSuccess && coop) {
        (void)cudaGetLastError();
        cfg.numAttrs = 1;
        err = cudaLaunchKernelEx(&cfg, e->fs_kernel, a, sv);
    }
    if (err != cudaSuccess) {
        // not launched: nothing has run, the error is not sticky (a sticky one shows up again in the fallback path)
        (void)cudaGetLastError();
        return SART_OK;
    }
    ++e->launches;
    *launched = true;
    return SART_OK;
}

int launch_decide(sart_engine* e, int t, const double* f2, uint32_t n, const IterState* st_in, IterState* st_out,
                  const FrameScalars* fs = nullptr) {
    DecideArgs d;
    d.f2part = f2;
    d.n = n;
    d.fs = fs ? fs : e->d_fs;
    d.st_in = st_in;
    d.st_out = st_out;
    d.tol = e->cfg.conv_tolerance;
    d.max_iter = e->cfg.max_iterations;
    d.fixed = e->cfg.fixed_iterations;
    d.t = t;
    decide_kernel<<<1, 32, 0, e->stream>>>(d);
    CUDA_TRY(cudaGetLastError());
    ++e->launches;
    return SART_OK;
}

// One sweep for the iterate x_t.
// SPLIT: forward projection, decision, and -- unless t is the last possible sweep, where the decision can only be
//   "stop" -- back projection, reduction and update; kernels after the decision test the freshly written state, so
//   nothing but the forward projection runs once the solve is over.
// FUSED: one cluster kernel does both projections from a single pass over HBM, then reduction (the scalar sum(fit^2)
//   rides along in the same all-reduce), decision, update. The last possible sweep only needs sum(fit^2): it runs the
//   plain forward projection.
int enqueue_sweep(sart_engine* e, int t) {
    const IterState* st_in = e->d_st[t & 1];
    IterState* st_out = e->d_st[(t + 1) & 1];
    const bool prof = e->profiling && t < kProfiledSweeps;
    const bool last = t >= e->cfg.max_iterations;
    const bool multi = e->cfg.nranks > 1;

    if (e->fused && !last) {
        if (prof) CUDA_TRY(cudaEventRecord(e->ev_prof[t][2], e->stream));
        SART_TRY(launch_fused(e, e->d_xf[t & 1], st_in));
        if (prof) CUDA_TRY(cudaEventRecord(e->ev_prof[t][3], e->stream));
        if (multi && e->peer_exchange) {
            ++e->xchg_epoch;  // same count on every rank: batches are enqueued identically (plan_launches)
            double* buf = reinterpret_cast<double*>(e->d_xchg + e->xchg_buf_off[e->xchg_epoch & 1]);
            SART_TRY(launch_reduce(e, st_in, e->fu_nclusters, buf, e->d_f2part, e->fu_nclusters));
            return launch_xchg_update(e, t, st_in, st_out);
        }
        SART_TRY(launch_reduce(e, st_in, e->fu_nclusters));
        const double* f2 = e->d_f2part;
        uint32_t nf2 = e->fu_nclusters;
        if (multi) {
            f2_sum_kernel<<<1, 32, 0, e->stream>>>(e->d_f2part, e->fu_nclusters, e->d_sred + e->V, st_in);
            CUDA_TRY(cudaGetLastError());
            ++e->launches;
            SART_TRY(allreduce_f64(e, e->d_sred, e->V + 1, ncclSum));
            f2 = e->d_sred + e->V;
            nf2 = 1;
        }
        SART_TRY(launch_decide(e, t, f2, nf2, st_in, st_out));
        return launch_update(e, t, st_out);
    }

    if (prof) CUDA_TRY(cudaEventRecord(e->ev_prof[t][0], e->stream));
    SART_TRY(launch_fp(e, e->d_xf[t & 1], st_in));
    if (prof) CUDA_TRY(cudaEventRecord(e->ev_prof[t][1], e->stream));
    const double* f2 = e->d_f2part;
    uint32_t nf2 = e->fp_grid;
    if (multi) {
        f2_sum_kernel<<<1, 32, 0, e->stream>>>(e->d_f2part, e->fp_grid, e->d_f2, st_in);
        CUDA_TRY(cudaGetLastError());
        ++e->launches;
        SART_TRY(allreduce_f64(e, e->d_f2, 1, ncclSum));
        f2 = e->d_f2;
        nf2 = 1;
    }
    SART_TRY(launch_decide(e, t, f2, nf2, st_in, st_out));
    if (last) return SART_OK;

    if (prof) CUDA_TRY(cudaEventRecord(e->ev_prof[t][2], e->stream));
    SART_TRY(launch_bp(e, st_out));
    if (prof) CUDA_TRY(cudaEventRecord(e->ev_prof[t][3], e->stream));
    SART_TRY(launch_reduce(e, st_out));
    SART_TRY(allreduce_f64(e, e->d_sred, e->V, ncclSum));
    return launch_update(e, t, st_out);
}

// column reduction of the staged weights d_w into d_sred (reduced over ranks): initial guess / log-solver "obs"
int backproject_weights(sart_engine* e) {
    SART_TRY(launch_bp(e, nullptr));
    SART_TRY(launch_reduce(e, nullptr));
    SART_TRY(allreduce_f64(e, e->d_sred, e->V, ncclSum));
    return SART_OK;
}

// Statistics of nframes frames (gd: [nframes][frame_stride] raw measurements on the device) into d_stat_max / d_stat_sum
// [first .. first + nframes): max g and sum over g > 0 of g^2 over ALL ranks. Two kernels for all the frames together; with
// the peers mapped the second one also does the exchange, otherwise the two arrays are all-reduced.
int compute_frame_stats(sart_engine* e, const double* gd, uint64_t frame_stride, int32_t nframes, int32_t first = 0) {
    if (first + nframes > e->stat_capacity) {
        if (first != 0) return fail(SART_ERR_STATE, "statistics buffer too small (internal error)");
        for (void* p : {(void*)e->d_stat_part, (void*)e->d_stat_max, (void*)e->d_stat_sum})
            if (p) CUDA_TRY(cudaFree(p));
        e->d_stat_part = nullptr;
        e->d_stat_max = e->d_stat_sum = nullptr;
        e->stat_capacity = 0;
        // enough CTAs to stream a frame at memory speed, few enough that the second stage stays trivial
        e->stat_blocks = (uint32_t)std::min<uint64_t>(64, std::max<uint64_t>(1, ceil_div(e->P, 8192)));
        const int32_t cap = std::max<int32_t>(nframes, 4);
        SART_TRY(dev_alloc(&e->d_stat_part, (uint64_t)cap * e->stat_blocks, false));
        SART_TRY(dev_alloc(&e->d_stat_max, (uint64_t)cap));
        SART_TRY(dev_alloc(&e->d_stat_sum, (uint64_t)cap));
        e->stat_capacity = cap;
    }
    const bool multi = e->cfg.nranks > 1;
    for (int32_t f0 = 0; f0 < nframes; f0 += kStatSlots) {
        const int32_t n = std::min<int32_t>(kStatSlots, nframes - f0);
        StatPair* part = e->d_stat_part + (uint64_t)(first + f0) * e->stat_blocks;
        frame_stats_partial_kernel<<<dim3(e->stat_blocks, (uint32_t)n), 1024, 0, e->stream>>>(gd + (uint64_t)f0 * frame_stride, e->P,
                                                                                              frame_stride, part);
        CUDA_TRY(cudaGetLastError());
        StatXchg x = {};
        if (multi && e->peer_exchange) {
            x.peers = e->d_peers;
            x.nranks = (uint32_t)e->cfg.nranks;
            x.rank = (uint32_t)e->cfg.rank;
            x.epoch = ++e->stat_epoch;  // every rank makes the same calls in the same order
            x.flag_off = e->xchg_stat_flag_off;
            x.slot_off[0] = e->xchg_stat_slot_off[0];
            x.slot_off[1] = e->xchg_stat_slot_off[1];
        }
        frame_stats_final_kernel<<<1, 1024, 0, e->stream>>>(part, e->stat_blocks, (uint32_t)n, e->d_stat_max + first + f0,
                                                            e->d_stat_sum + first + f0, x);
        CUDA_TRY(cudaGetLastError());
        e->launches += 2;
        if (multi && !e->peer_exchange) {
            SART_TRY(allreduce_f64(e, e->d_stat_max + first + f0, (uint64_t)n, ncclMax));
            SART_TRY(allreduce_f64(e, e->d_stat_sum + first + f0, (uint64_t)n, ncclSum));
        }
    }
    return SART_OK;
}

// stat: index of the frame's statistics in d_stat_max / d_stat_sum
int launch_prepare(sart_engine* e, int weights_mode, const double* gd, int32_t stat = 0) {
    frame_prepare_kernel<<<(uint32_t)ceil_div(e->P, 256), 256, 0, e->stream>>>(
        gd, e->P, e->d_fs, e->d_stat_max + stat, e->d_stat_sum + stat, e->d_lambda, e->cfg.ray_length_threshold, e->d_g, e->d_rinv,
        e->d_w, weights_mode);
    CUDA_TRY(cudaGetLastError());
    ++e->launches;
    return SART_OK;
}

// What follows the whole-solve kernel: final state to the host, un-normalised solution, statistics.
int finish_persistent_solve(sart_engine* e, int sweeps_before, uint32_t vblocks) {
    if (e->profiling) CUDA_TRY(cudaEventRecord(e->ev_prof[0][3], e->stream));
    CUDA_TRY(cudaMemcpyAsync(&e->h_st[0], e->d_st[0], sizeof(IterState), cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(cudaMemcpyAsync(e->h_solve_ctl, e->d_solve_ctl, kSolveCtlWords * sizeof(unsigned long long), cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(cudaEventRecord(e->ev_batch[0], e->stream));
    CUDA_TRY(cudaEventSynchronize(e->ev_batch[0]));
    const IterState fin = e->h_st[0];
    const unsigned long long* tail = e->h_solve_ctl + 1;  // tail time (ns), tails timed, scratch, tails run
    const unsigned long long* phase = e->h_solve_ctl + 8;
    if (!fin.done) return fail(SART_ERR_STATE, "whole-solve kernel ended without a decision (internal error)");
    if (fin.status == kStatusExchangeTimeout)
        return fail(SART_ERR_NCCL, "in-kernel barrier or peer-memory exchange timed out: a CTA or a rank of the job did not arrive");
    if (e->cfg.nranks > 1) e->xchg_epoch += tail[3];  // one exchange per sweep run, every rank alike

    finish_solution_kernel<<<vblocks, 256, 0, e->stream>>>(e->d_xd[0], e->d_xd[1], e->d_st[0], e->d_fs, e->d_xio, e->V);
    CUDA_TRY(cudaGetLastError());
    ++e->launches;

    const int nsweeps = (int)tail[3];  // sweeps run inside the one launch (iterations + 1 for a solve of its own)
    e->stats.status = fin.status;
    e->stats.iterations = fin.iterations;
    e->stats.sweeps = sweeps_before + 2 * fin.iterations + 1;
    float ms = 0.f;
    CUDA_TRY(cudaEventElapsedTime(&ms, e->ev_it0, e->ev_batch[0]));
    e->stats.iterate_ms = ms;
    e->stats.fp_ms_avg = e->stats.bp_ms_avg = 0;
    e->stats.fp_launches = e->stats.bp_launches = 0;
    e->stats.fused_ms_avg = 0;
    e->stats.fused_launches = nsweeps;
    e->stats.solve_launches = 1;
    e->stats.tail_ms_avg = tail[1] ? 1e-6 * (double)tail[0] / (double)tail[1] : 0;
    for (int k = 0; k < kTailPhases; ++k) e->stats.tail_phase_us[k] = tail[1] ? 1e-3 * (double)phase[k] / (double)tail[1] : 0;
    if (env_int("SART_TAIL_DEBUG", 0) && tail[1])
        std::fprintf(stderr, "sart tail debug (us per sweep, thread 0 of CTA 0): own cluster sum %.2f, own update %.2f, before them %.2f\n",
                     1e-3 * (double)phase[5] / (double)tail[1], 1e-3 * (double)phase[6] / (double)tail[1],
                     1e-3 * (double)phase[7] / (double)tail[1]);
    e->stats.sweep_mode = SART_SWEEP_FUSED;
    e->stats.exchange_mode = e->cfg.nranks <= 1 ? SART_EXCHANGE_NONE : SART_EXCHANGE_PEER;
    if (e->profiling) {
        CUDA_TRY(cudaEventElapsedTime(&ms, e->ev_prof[0][2], e->ev_prof[0][3]));
        e->stats.fused_ms_avg = ms / nsweeps;  // per sweep, in-kernel tails included
    }
    return SART_OK;
}

// The whole solve from device-resident inputs (d_gd, and d_xio when has_guess) to device-resident output (d_xio).
// Per-frame set-up from the staged measurement d_gd (and d_xio when has_guess): statistics, normalisation, starting
// solution into buffer 0, fresh control blocks. *sweeps counts the matrix passes it took (initial guess, log "obs").
// gd: the frame's raw measurement on the device (default: the staging buffer d_gd).
// stat >= 0: the frame's statistics are already in d_stat_max / d_stat_sum [stat] (a series computes them for the whole chunk).
int frame_prologue(sart_engine* e, int has_guess, int* sweeps, const SeriesCtl* series = nullptr, const double* gd = nullptr,
                   int32_t stat = -1) {
    if (!gd) gd = e->d_gd;
    const uint32_t vblocks = (uint32_t)ceil_div(e->V, 256);
    const bool gpu_sem = e->cfg.semantics == SART_SEMANTICS_GPU;
    const bool log_mode = e->cfg.logarithmic != 0;

    if (stat < 0) {
        SART_TRY(compute_frame_stats(e, gd, e->P, 1));
        stat = 0;
    }

    *sweeps = 0;
    if (log_mode) {
        SART_TRY(launch_prepare(e, 3, gd, stat));
        SART_TRY(backproject_weights(e));
        mask_obs_kernel<<<vblocks, 256, 0, e->stream>>>(e->d_sred, e->d_rho, e->cfg.ray_density_threshold, e->d_obs, e->V);
        CUDA_TRY(cudaGetLastError());
        ++e->launches;
        ++*sweeps;
    }
    SART_TRY(launch_prepare(e, has_guess ? 0 : (gpu_sem ? 2 : 1), gd, stat));
    if (!has_guess) {
        SART_TRY(backproject_weights(e));
        ++*sweeps;
    }
    // lower clamp of the starting solution: GPU semantics 1e-7 (normalised units) for both solvers
    // (sartsolver_cuda.cpp:180); CPU semantics 1e-100 (caller units) for the log solver only (sartsolver.cpp:263)
    const double floor_norm = gpu_sem ? 1.e-7 : 0.0;
    const double floor_orig = (!gpu_sem && log_mode) ? 1.e-100 : 0.0;
    start_solution_kernel<<<vblocks, 256, 0, e->stream>>>(e->d_sred, e->d_xio, e->d_rho, e->cfg.ray_density_threshold,
                                                          e->d_fs, floor_norm, floor_orig, has_guess, e->d_xd[0],
                                                          e->d_xf[0], e->V, series);
    CUDA_TRY(cudaGetLastError());
    ++e->launches;
    init_state_kernel<<<1, 1, 0, e->stream>>>(e->d_st[0], e->d_st[1]);
    CUDA_TRY(cudaGetLastError());
    ++e->launches;
    return SART_OK;
}

int run_solve(sart_engine* e, int has_guess) {
    const uint32_t vblocks = (uint32_t)ceil_div(e->V, 256);
    int sweeps = 0;
    SART_TRY(frame_prologue(e, has_guess, &sweeps));
    e->last_prologue_sweeps = sweeps;
    CUDA_TRY(cudaEventRecord(e->ev_it0, e->stream));

    if (e->fused && e->persistent && (e->cfg.nranks <= 1 || e->peer_exchange)) {
        bool launched = false;
        if (e->profiling) CUDA_TRY(cudaEventRecord(e->ev_prof[0][2], e->stream));
        SART_TRY(launch_solve(e, &launched));
        if (launched) return finish_persistent_solve(e, sweeps, vblocks);
        e->persistent = false;  // this device / driver refuses the launch: one launch per sweep from now on
    }

    // Sweeps t = 0 .. max_iterations. Batches are enqueued one ahead of the status poll, so the device never
    // waits for the host; launches that arrive after the device has decided to stop return immediately.
    const int t_max = e->cfg.max_iterations;
    int t = 0, slot = 0, done_slot = -1;
    std::deque<int> in_flight;
    while (true) {
        if (t <= t_max) {
            const int nb = std::min(e->batch, t_max - t + 1);
            for (int i = 0; i < nb; ++i) SART_TRY(enqueue_sweep(e, t + i));
            t += nb;
            CUDA_TRY(cudaMemcpyAsync(&e->h_st[slot], e->d_st[t & 1], sizeof(IterState), cudaMemcpyDeviceToHost, e->stream));
            CUDA_TRY(cudaEventRecord(e->ev_batch[slot], e->stream));
            in_flight.push_back(slot);
            slot ^= 1;
        }
        if (in_flight.size() == 2 || t > t_max) {
            if (in_flight.empty()) return fail(SART_ERR_STATE, "solve ended without a decision (internal error)");
            const int s = in_flight.front();
            in_flight.pop_front();
            CUDA_TRY(cudaEventSynchronize(e->ev_batch[s]));
            if (e->h_st[s].done) {
                done_slot = s;
                break;
            }
        }
    }
    const IterState fin = e->h_st[done_slot];
    if (fin.status == kStatusExchangeTimeout)
        return fail(SART_ERR_NCCL, "peer-memory exchange timed out: a rank of the job did not reach the reduction");

    finish_solution_kernel<<<vblocks, 256, 0, e->stream>>>(e->d_xd[0], e->d_xd[1], e->d_st[t & 1], e->d_fs, e->d_xio, e->V);
    CUDA_TRY(cudaGetLastError());
    ++e->launches;

    e->stats.status = fin.status;
    e->stats.iterations = fin.iterations;
    e->stats.sweeps = sweeps + 2 * fin.iterations + 1;
    float ms = 0.f;
    CUDA_TRY(cudaEventElapsedTime(&ms, e->ev_it0, e->ev_batch[done_slot]));
    e->stats.iterate_ms = ms;
    e->stats.fp_ms_avg = e->stats.bp_ms_avg = 0;
    e->stats.fp_launches = e->stats.bp_launches = 0;
    e->stats.fused_ms_avg = 0;
    e->stats.fused_launches = 0;
    e->stats.solve_launches = 0;
    e->stats.tail_ms_avg = 0;
    e->stats.sweep_mode = e->fused ? SART_SWEEP_FUSED : SART_SWEEP_SPLIT;
    e->stats.exchange_mode = e->cfg.nranks <= 1 ? SART_EXCHANGE_NONE
                             : (e->fused && e->peer_exchange) ? SART_EXCHANGE_PEER : SART_EXCHANGE_NCCL;
    if (e->profiling) {
        // every profiled sweep up to the deciding one has completed (they precede ev_batch[done_slot]).
        // Events [0],[1] bracket a forward projection, [2],[3] a back projection or a fused sweep.
        const int nsweeps = std::min(kProfiledSweeps, fin.iterations + 1);  // sweeps 0 .. iterations
        double fp = 0, second = 0;
        int nfp = 0, nsecond = 0;
        for (int k = 0; k < nsweeps; ++k) {
            const bool last = k >= e->cfg.max_iterations;
            const bool fused_sweep = e->fused && !last;
            float a = 0.f;
            if (!fused_sweep) {
                CUDA_TRY(cudaEventElapsedTime(&a, e->ev_prof[k][0], e->ev_prof[k][1]));
                fp += a;
                ++nfp;
            }
            // the deciding sweep of the split path ran its forward projection only; a fused sweep always ran whole
            if (fused_sweep || (!last && k < fin.iterations)) {
                CUDA_TRY(cudaEventElapsedTime(&a, e->ev_prof[k][2], e->ev_prof[k][3]));
                second += a;
                ++nsecond;
            }
        }
        // what follows a fused sweep before the next one may start: reduction over clusters and ranks, decision, update
        double tail = 0;
        int ntail = 0;
        for (int k = 0; e->fused && k + 1 < nsweeps && k + 1 < e->cfg.max_iterations; ++k) {
            float a = 0.f;
            CUDA_TRY(cudaEventElapsedTime(&a, e->ev_prof[k][3], e->ev_prof[k + 1][2]));
            tail += a;
            ++ntail;
        }
        e->stats.tail_ms_avg = ntail ? tail / ntail : 0;
        e->stats.fp_ms_avg = nfp ? fp / nfp : 0;
        e->stats.fp_launches = nfp;
        if (e->fused) {
            e->stats.fused_ms_avg = nsecond ? second / nsecond : 0;
            e->stats.fused_launches = nsecond;
        } else {
            e->stats.bp_ms_avg = nsecond ? second / nsecond : 0;
            e->stats.bp_launches = nsecond;
        }
    }
    return SART_OK;
}

// ---- several frames per matrix pass ---------------------------------------------------------------------------------------
// Frames solved without a warm start are independent right-hand sides of the same matrix, so kGroup of them ride on
// every pass over it (fp_rows_batch_kernel / bp_cols_batch_kernel): 2 passes per iteration for the whole group instead of
// one fused pass per frame. Linear solver; with several ranks the per-frame sums travel in one NCCL all-reduce per
// pass. The log solver rides along (its back-projected measurement of every frame comes from one more batched pass at
// set-up). A last single frame is solved on its own by the caller.
// Group size: 8 frames per pass (forward kernel: 4 rows x 8 frames of accumulators per thread) or 4 (8 rows x 4 frames);
// SART_BATCH_FRAMES picks, buffers are sized for kGroup = the larger one.
constexpr int kGroup = 8;

// Tensor map of a row-major fp32 array [rows][ld] for boxes of [box_rows][kRingCols4 * 4 columns] (batch_ring.cuh). The
// encoder is a driver function; it is looked up through the runtime, so the library still does not link libcuda.
bool encode_rows_map(CUtensorMap* map, const float* base, uint64_t rows, uint64_t ld, uint32_t box_rows) {
    using EncodeTiled = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                     const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                     CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static EncodeTiled encode = [] {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult found = cudaDriverEntryPointSymbolNotFound;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &found) != cudaSuccess ||
            found != cudaDriverEntryPointSuccess)
            fn = nullptr;
        (void)cudaGetLastError();
        return reinterpret_cast<EncodeTiled>(fn);
    }();
    if (!encode || (reinterpret_cast<uintptr_t>(base) & 15) || (ld * sizeof(float)) % 16 || ld >= (1ull << 32) || rows >= (1ull << 32))
        return false;
    const cuuint64_t dims[2] = {ld, rows};
    const cuuint64_t strides[1] = {ld * sizeof(float)};
    const cuuint32_t box[2] = {(cuuint32_t)(kRingCols4 * 4), box_rows};
    const cuuint32_t steps[2] = {1, 1};
    return encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, steps,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

constexpr const char* kBatchFpDefault = "ring";
constexpr const char* kBatchBpDefault = "ring";  // back projection of eight-frame groups: "ring" or "smem" (kernels.cuh)  // forward kernel of eight-frame groups: "ring" (batch_ring.cuh) or "warp"
constexpr int kFp8Threads = 384, kFp8Rows = 8;  // eight frames: 8 rows x 8 frames of accumulators per thread, one CTA per SM
template <int F>
struct GroupKernels {
    static constexpr int kFpRows = (F >= 8) ? kFp8Rows : 32 / F;
    static constexpr int kFpThreadsOf = (F >= 8) ? kFp8Threads : kFpThreads;
    // eight frames: a row group per warp (x shared through L1) and weights from shared memory; four: a row group per CTA
    // and weights by shuffle (kernels.cuh)
    static void (*fp())(const FpBatchArgs) {
        if (F >= 8) return fp_rows_batch_warp_kernel<kFp8Rows, kFp8Threads, F, 1>;
        return fp_rows_batch_kernel<32 / F, kFpThreads, F>;
    }
    static void (*bp())(const BpBatchArgs) {
        if (F >= 8) return bp_cols_batch_smem_kernel<kBpThreads, 8, F>;
        return bp_cols_batch_kernel<kBpThreads, 8, F>;
    }
};

int plan_group(sart_engine* e) {
    if (e->group_planned) return SART_OK;
    e->group_planned = true;
    e->group_ok = false;
    int fp_occ = 0, bp_occ = 0;
    e->group_frames = env_int("SART_BATCH_FRAMES", 8) >= 8 ? 8 : 4;
    e->gp_fp_rows = e->group_frames == 8 ? GroupKernels<8>::kFpRows : GroupKernels<4>::kFpRows;
    e->gp_fp_threads = e->group_frames == 8 ? GroupKernels<8>::kFpThreadsOf : GroupKernels<4>::kFpThreadsOf;
    if (e->group_frames == 8) {
        CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&fp_occ, GroupKernels<8>::fp(), kFp8Threads, 0));
        CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bp_occ, GroupKernels<8>::bp(), kBpThreads, 0));
    } else {
        CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&fp_occ, GroupKernels<4>::fp(), kFpThreads, 0));
        CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bp_occ, GroupKernels<4>::bp(), kBpThreads, 0));
    }
    const uint64_t nrg = ceil_div(e->P, (uint64_t)e->gp_fp_rows);
    const uint64_t groups_per_cta = e->group_frames == 8 ? kFp8Threads / 32 : 1;
    e->gp_fp_grid = (uint32_t)std::min<uint64_t>(ceil_div(nrg, groups_per_cta), (uint64_t)e->sm_count * std::max(fp_occ, 1));
    // eight frames: the forward kernel that stages the matrix through a shared-memory ring (SART_BATCH_FP=warp keeps the
    // register-file one), when the device grants the shared memory for at least two stages
    e->gp_fp_ring_stages = 0;
    const uint32_t warp_grid = e->gp_fp_grid;
    const int warp_rows = e->gp_fp_rows, warp_threads = e->gp_fp_threads;
    const char* fp_kind = std::getenv("SART_BATCH_FP");
    if (e->group_frames == 8 && std::string(fp_kind ? fp_kind : kBatchFpDefault) == "ring") {
        int max_smem = 0;
        CUDA_TRY(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, e->device));
        const uint32_t stages = ring_stages<8>(max_smem);
        if (stages >= 2 &&
            cudaFuncSetAttribute(fp_rows_batch_ring_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)ring_smem_bytes<8>(stages)) == cudaSuccess) {
            e->gp_fp_ring_stages = stages;
            e->gp_fp_ring_smem = ring_smem_bytes<8>(stages);
            e->gp_fp_rows = kRingRows;
            e->gp_fp_threads = kRingThreads;
            e->gp_fp_grid = (uint32_t)std::min<uint64_t>(ceil_div(e->P, (uint64_t)kRingRows), (uint64_t)e->sm_count);
        }
        (void)cudaGetLastError();
    }
    // back projection: one wave of (panels x slabs) CTAs at this kernel's own occupancy, slabs a multiple of 32 rows
    const uint64_t panels = ceil_div(e->ld4, kBpThreads);
    uint64_t slabs = std::max<uint64_t>(1, (uint64_t)e->sm_count * std::max(bp_occ, 1) / panels);
    slabs = std::min<uint64_t>(slabs, ceil_div(e->P, 32));
    e->gp_bp_rows_per_slab = (uint32_t)(ceil_div(ceil_div(e->P, slabs), 32) * 32);
    e->gp_bp_slabs = (uint32_t)ceil_div(e->P, e->gp_bp_rows_per_slab);

    // the same for the back projection (SART_BATCH_BP): panels of 256 columns x slabs of rows, about four waves of CTAs
    e->gp_bp_ring_stages = 0;
    const uint32_t smem_slabs = e->gp_bp_slabs, smem_rows_per_slab = e->gp_bp_rows_per_slab;
    const char* bp_kind = std::getenv("SART_BATCH_BP");
    if (e->group_frames == 8 && std::string(bp_kind ? bp_kind : kBatchBpDefault) == "ring") {
        int max_smem = 0;
        CUDA_TRY(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, e->device));
        const uint32_t stages = bp_ring_stages<8>(max_smem);
        if (stages >= 2 &&
            cudaFuncSetAttribute(bp_cols_batch_ring_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)bp_ring_smem_bytes<8>(stages)) == cudaSuccess) {
            const uint64_t ring_panels = ceil_div(e->ld4, (uint64_t)kBpRingCols4);
            uint64_t ring_slabs = std::max<uint64_t>(1, (uint64_t)e->sm_count * 4 / ring_panels);
            ring_slabs = std::min<uint64_t>(ring_slabs, ceil_div(e->P, (uint64_t)kBpRingRows));
            e->gp_bp_rows_per_slab = (uint32_t)(ceil_div(ceil_div(e->P, ring_slabs), (uint64_t)kBpRingRows) * kBpRingRows);
            e->gp_bp_slabs = (uint32_t)ceil_div(e->P, e->gp_bp_rows_per_slab);
            e->gp_bp_ring_stages = stages;
            e->gp_bp_ring_smem = bp_ring_smem_bytes<8>(stages);
        }
        (void)cudaGetLastError();
    }

    const uint64_t G = kGroup;
    SART_TRY(dev_alloc(&e->d_pgd, G * e->P));
    SART_TRY(dev_alloc(&e->d_pg, G * e->P));
    SART_TRY(dev_alloc(&e->d_pw, G * e->P));
    SART_TRY(dev_alloc(&e->d_pxio, G * e->V));
    SART_TRY(dev_alloc(&e->d_pobs, G * e->V));
    SART_TRY(dev_alloc(&e->d_psred, G * (e->V + 2)));
    SART_TRY(dev_alloc(&e->d_pf2part, G * (uint64_t)std::max(e->gp_fp_grid, warp_grid)));
    SART_TRY(dev_alloc(&e->d_pf2sum, G));
    SART_TRY(dev_alloc(&e->d_ppart, G * (uint64_t)std::max(e->gp_bp_slabs, smem_slabs) * e->ld));
    SART_TRY(dev_alloc(&e->d_pfs, G));
    for (int k = 0; k < 2; ++k) {
        SART_TRY(dev_alloc(&e->d_pxd[k], G * e->V));
        SART_TRY(dev_alloc(&e->d_pxf[k], G * e->ld));
        SART_TRY(dev_alloc(&e->d_pst[k], G));
    }
    CUDA_TRY(cudaMallocHost(reinterpret_cast<void**>(&e->h_pst), 2 * G * sizeof(IterState)));
    if (e->gp_fp_ring_stages &&
        !(encode_rows_map(&e->gp_map_h, e->dH, e->P, e->ld, kRingRows) && encode_rows_map(&e->gp_map_x[0], e->d_pxf[0], G, e->ld, (uint32_t)G) &&
          encode_rows_map(&e->gp_map_x[1], e->d_pxf[1], G, e->ld, (uint32_t)G))) {
        e->gp_fp_ring_stages = 0;  // no tensor maps on this driver: the register-file kernel
        e->gp_fp_grid = warp_grid;
        e->gp_fp_rows = warp_rows;
        e->gp_fp_threads = warp_threads;
    }
    if (e->gp_bp_ring_stages && !encode_rows_map(&e->gp_map_h_bp, e->dH, e->P, e->ld, kBpRingRows)) {
        e->gp_bp_ring_stages = 0;
        e->gp_bp_slabs = smem_slabs;
        e->gp_bp_rows_per_slab = smem_rows_per_slab;
    }
    e->group_ok = true;
    return SART_OK;
}

// nf (2 .. kGroup) frames from d_pgd to d_pxio, default initial guess, no warm start. fin[f] = final state of frame f.
int run_solve_group(sart_engine* e, int nf, IterState* fin) {
    const uint32_t vblocks = (uint32_t)ceil_div(e->V, 256), pblocks = (uint32_t)ceil_div(e->P, 256);
    const bool gpu_sem = e->cfg.semantics == SART_SEMANTICS_GPU;
    const uint64_t P = e->P, V = e->V, ld = e->ld;
    const uint64_t part_stride = (uint64_t)e->gp_bp_slabs * ld;

    const bool multi = e->cfg.nranks > 1;
    const dim3 vgrid(vblocks, (uint32_t)nf);  // one CTA row per frame: the per-frame steps of all frames in one launch
    auto launch_bp = [&](const IterState* st) -> int {  // d_pw -> slab partials, every frame in one pass
        BpBatchArgs b;
        b.H = e->dH;
        b.ld4 = e->ld4;
        b.P = P;
        b.w = e->d_pw;
        b.part = e->d_ppart;
        b.rows_per_slab = e->gp_bp_rows_per_slab;
        b.st = st;
        b.nframes = nf;
        if (e->gp_bp_ring_stages)
            bp_cols_batch_ring_kernel<8><<<dim3((uint32_t)ceil_div(e->ld4, (uint64_t)kBpRingCols4), e->gp_bp_slabs), kBpRingThreads,
                                           e->gp_bp_ring_smem, e->stream>>>(b, e->gp_bp_ring_stages, e->gp_map_h_bp);
        else
            (e->group_frames == 8 ? GroupKernels<8>::bp() : GroupKernels<4>::bp())
                <<<dim3((uint32_t)ceil_div(e->ld4, kBpThreads), e->gp_bp_slabs), kBpThreads, 0, e->stream>>>(b);
        CUDA_TRY(cudaGetLastError());
        ++e->launches;
        return SART_OK;
    };
    // slab partials -> dst[f * dst_stride + j] (with_f2: the frame's sum(fit^2) behind its vector, at [V])
    auto reduce_all = [&](const IterState* st, double* dst, uint64_t dst_stride, bool with_f2) -> int {
        reduce_partials_batch_kernel<<<vgrid, 256, 0, e->stream>>>(e->d_ppart, e->gp_bp_slabs, ld, V, part_stride, dst, dst_stride, st,
                                                                    with_f2 ? e->d_pf2part : nullptr, e->gp_fp_grid, e->gp_fp_grid);
        CUDA_TRY(cudaGetLastError());
        ++e->launches;
        return SART_OK;
    };
    auto back_project = [&](const IterState* st) -> int {  // d_pw -> d_psred[f], reduced over the row shards
        SART_TRY(launch_bp(st));
        SART_TRY(reduce_all(st, e->d_psred, V + 2, false));
        return allreduce_f64(e, e->d_psred, (uint64_t)nf * (V + 2), ncclSum);  // all frames at once
    };

    // per frame: statistics, normalisation, weights of the default initial guess; one pass back-projects all of them
    SART_TRY(compute_frame_stats(e, e->d_pgd, P, nf));  // all frames of the group together
    const bool log_mode = e->cfg.logarithmic != 0;
    auto prepare = [&](int weights_mode) -> int {
        for (int f = 0; f < nf; ++f) {
            frame_prepare_kernel<<<pblocks, 256, 0, e->stream>>>(e->d_pgd + f * P, P, e->d_pfs + f, e->d_stat_max + f,
                                                                 e->d_stat_sum + f, e->d_lambda, e->cfg.ray_length_threshold,
                                                                 e->d_pg + f * P, e->d_rinv, e->d_pw + f * P, weights_mode);
            CUDA_TRY(cudaGetLastError());
            ++e->launches;
        }
        return SART_OK;
    };
    if (log_mode) {  // the log solver's iteration-invariant back-projected measurement of every frame, in one pass
        SART_TRY(prepare(3));
        SART_TRY(back_project(nullptr));
        for (int f = 0; f < nf; ++f) {
            mask_obs_kernel<<<vblocks, 256, 0, e->stream>>>(e->d_psred + f * (V + 2), e->d_rho, e->cfg.ray_density_threshold,
                                                            e->d_pobs + f * V, V);
            CUDA_TRY(cudaGetLastError());
            ++e->launches;
        }
    }
    SART_TRY(prepare(gpu_sem ? 2 : 1));
    SART_TRY(back_project(nullptr));
    for (int f = 0; f < nf; ++f) {
        // lower clamp of the starting solution as in frame_prologue: GPU semantics 1e-7 (normalised), CPU log solver 1e-100
        start_solution_kernel<<<vblocks, 256, 0, e->stream>>>(e->d_psred + f * (V + 2), nullptr, e->d_rho,
                                                              e->cfg.ray_density_threshold, e->d_pfs + f, gpu_sem ? 1.e-7 : 0.0,
                                                              (!gpu_sem && log_mode) ? 1.e-100 : 0.0, 0, e->d_pxd[0] + f * V, e->d_pxf[0] + f * ld, V);
        init_state_kernel<<<1, 1, 0, e->stream>>>(e->d_pst[0] + f, e->d_pst[1] + f);
        CUDA_TRY(cudaGetLastError());
        e->launches += 2;
    }
    CUDA_TRY(cudaEventRecord(e->ev_it0, e->stream));

    // sweep t, like the two-kernel path of enqueue_sweep: projection, decision per frame, and -- unless t is the last
    // possible sweep -- back projection, reduction and update per frame
    auto enqueue = [&](int t) -> int {
        const IterState* st_in = e->d_pst[t & 1];
        IterState* st_out = e->d_pst[(t + 1) & 1];
        const bool prof = e->profiling && t < kProfiledSweeps;
        FpBatchArgs a;
        a.H = e->dH;
        a.ld4 = e->ld4;
        a.P = P;
        a.x = e->d_pxf[t & 1];
        a.g = e->d_pg;
        a.rinv = e->d_rinv;
        a.w = e->d_pw;
        a.f2part = e->d_pf2part;
        a.st = st_in;
        a.nrg = (uint32_t)ceil_div(P, (uint64_t)e->gp_fp_rows);
        a.nframes = nf;
        a.log_mode = e->cfg.logarithmic;
        if (prof) CUDA_TRY(cudaEventRecord(e->ev_prof[t][0], e->stream));
        if (e->gp_fp_ring_stages)
            fp_rows_batch_ring_kernel<8><<<e->gp_fp_grid, kRingThreads, e->gp_fp_ring_smem, e->stream>>>(a, e->gp_fp_ring_stages, e->gp_map_h,
                                                                                                           e->gp_map_x[t & 1]);
        else
            (e->group_frames == 8 ? GroupKernels<8>::fp() : GroupKernels<4>::fp())<<<e->gp_fp_grid, e->gp_fp_threads, 0, e->stream>>>(a);
        CUDA_TRY(cudaGetLastError());
        if (prof) CUDA_TRY(cudaEventRecord(e->ev_prof[t][1], e->stream));
        ++e->launches;
        const bool last = t >= e->cfg.max_iterations;
        UpdArgs u = make_update_args(e, t, st_out);  // base pointers: the kernels step to their frame
        u.sred = e->d_psred;
        u.xd_cur = e->d_pxd[t & 1];
        u.xd_next = e->d_pxd[(t + 1) & 1];
        u.xf_next = e->d_pxf[(t + 1) & 1];
        u.fs = e->d_pfs;
        u.obs = e->d_pobs;
        if (multi && e->peer_exchange && !last) {
            // back projection, then reduction over the ranks + decision + update of every frame in one kernel over peer
            // memory (the frames' partial vectors lie one after the other in this rank's exchange buffer)
            if (prof) CUDA_TRY(cudaEventRecord(e->ev_prof[t][2], e->stream));
            SART_TRY(launch_bp(st_in));
            if (prof) CUDA_TRY(cudaEventRecord(e->ev_prof[t][3], e->stream));
            ++e->xchg_epoch;  // same count on every rank
            SART_TRY(reduce_all(st_in, reinterpret_cast<double*>(e->d_xchg + e->xchg_buf_off[e->xchg_epoch & 1]), V + 1, true));
            XchgArgs x;
            x.peers = e->d_peers;
            x.nranks = (uint32_t)e->cfg.nranks;
            x.rank = (uint32_t)e->cfg.rank;
            x.epoch = e->xchg_epoch;
            x.buf_off = e->xchg_buf_off[e->xchg_epoch & 1];
            x.st_in = st_in;
            x.st_out = st_out;
            x.tol = e->cfg.conv_tolerance;
            x.max_iter = e->cfg.max_iterations;
            x.fixed = e->cfg.fixed_iterations;
            x.t = t;
            u.st = nullptr;
            xchg_decide_update_batch_kernel<<<vgrid, 256, 0, e->stream>>>(x, u, ld);
            CUDA_TRY(cudaGetLastError());
            ++e->launches;
            return SART_OK;
        }
        DecideArgs d;
        d.f2part = e->d_pf2part;
        d.n = e->gp_fp_grid;
        d.fs = e->d_pfs;
        d.st_in = st_in;
        d.st_out = st_out;
        d.tol = e->cfg.conv_tolerance;
        d.max_iter = e->cfg.max_iterations;
        d.fixed = e->cfg.fixed_iterations;
        d.t = t;
        uint32_t f2_stride = e->gp_fp_grid;
        if (multi) {  // sum(fit^2) of every frame over the row shards before the decisions
            f2_sum_batch_kernel<<<nf, 32, 0, e->stream>>>(e->d_pf2part, e->gp_fp_grid, e->gp_fp_grid, e->d_pf2sum, st_in);
            CUDA_TRY(cudaGetLastError());
            ++e->launches;
            SART_TRY(allreduce_f64(e, e->d_pf2sum, (uint64_t)nf, ncclSum));
            d.f2part = e->d_pf2sum;
            d.n = 1;
            f2_stride = 1;
        }
        decide_batch_kernel<<<nf, 32, 0, e->stream>>>(d, f2_stride);
        CUDA_TRY(cudaGetLastError());
        ++e->launches;
        if (last) return SART_OK;
        if (prof) CUDA_TRY(cudaEventRecord(e->ev_prof[t][2], e->stream));
        SART_TRY(launch_bp(st_out));
        if (prof) CUDA_TRY(cudaEventRecord(e->ev_prof[t][3], e->stream));
        if (!multi) {  // slab sums and update of every frame in one launch
            reduce_update_batch_kernel<<<vgrid, 256, 0, e->stream>>>(u, ld, e->d_ppart, e->gp_bp_slabs, part_stride);
        } else {
            SART_TRY(reduce_all(st_out, e->d_psred, V + 2, false));
            SART_TRY(allreduce_f64(e, e->d_psred, (uint64_t)nf * (V + 2), ncclSum));
            update_batch_kernel<<<vgrid, 256, 0, e->stream>>>(u, ld);
        }
        CUDA_TRY(cudaGetLastError());
        ++e->launches;
        return SART_OK;
    };

    const int t_max = e->cfg.max_iterations;
    int t = 0, slot = 0, done_slot = -1;
    std::deque<int> in_flight;
    while (true) {
        if (t <= t_max) {
            const int nb = std::min(e->batch, t_max - t + 1);
            for (int i = 0; i < nb; ++i) SART_TRY(enqueue(t + i));
            t += nb;
            CUDA_TRY(cudaMemcpyAsync(&e->h_pst[kGroup * slot], e->d_pst[t & 1], nf * sizeof(IterState), cudaMemcpyDeviceToHost, e->stream));
            CUDA_TRY(cudaEventRecord(e->ev_batch[slot], e->stream));
            in_flight.push_back(slot);
            slot ^= 1;
        }
        if (in_flight.size() == 2 || t > t_max) {
            if (in_flight.empty()) return fail(SART_ERR_STATE, "group solve ended without a decision (internal error)");
            const int sl = in_flight.front();
            in_flight.pop_front();
            CUDA_TRY(cudaEventSynchronize(e->ev_batch[sl]));
            bool all = true;
            for (int f = 0; f < nf; ++f) all = all && e->h_pst[kGroup * sl + f].done;
            if (all) {
                done_slot = sl;
                break;
            }
        }
    }
    int its = 0;
    for (int f = 0; f < nf; ++f) {
        fin[f] = e->h_pst[kGroup * done_slot + f];
        its = std::max(its, fin[f].iterations);
        finish_solution_kernel<<<vblocks, 256, 0, e->stream>>>(e->d_pxd[0] + f * V, e->d_pxd[1] + f * V, e->d_pst[t & 1] + f,
                                                               e->d_pfs + f, e->d_pxio + f * V, V);
        CUDA_TRY(cudaGetLastError());
        ++e->launches;
    }
    float ms = 0.f;
    CUDA_TRY(cudaEventElapsedTime(&ms, e->ev_it0, e->ev_batch[done_slot]));
    e->stats.iterate_ms = ms;
    e->stats.fused_ms_avg = e->stats.tail_ms_avg = 0;
    e->stats.fused_launches = e->stats.solve_launches = 0;
    e->stats.fp_ms_avg = e->stats.bp_ms_avg = 0;
    e->stats.fp_launches = e->stats.bp_launches = 0;
    if (e->profiling) {
        const int n = std::min(kProfiledSweeps, its + 1);
        double fp = 0, bp = 0;
        int nbp = 0;
        for (int k = 0; k < n; ++k) {
            float a = 0.f;
            CUDA_TRY(cudaEventElapsedTime(&a, e->ev_prof[k][0], e->ev_prof[k][1]));
            fp += a;
            if (k < its && k < e->cfg.max_iterations) {
                CUDA_TRY(cudaEventElapsedTime(&a, e->ev_prof[k][2], e->ev_prof[k][3]));
                bp += a;
                ++nbp;
            }
        }
        e->stats.fp_ms_avg = n ? fp / n : 0;
        e->stats.fp_launches = n;
        e->stats.bp_ms_avg = nbp ? bp / nbp : 0;
        e->stats.bp_launches = nbp;
    }
    return SART_OK;
}

int build_laplacian(sart_engine* e) {
    // with several ranks the reference hands the matrix to rank 0 only (source/main.cpp:73-74): share rank 0's copy
    if (e->cfg.nranks > 1) {
        NcclApi* api = nccl_api();
        if (!api || !e->comm) return fail(SART_ERR_STATE, "nranks > 1 but sart_comm_init has not been called");
        unsigned long long* d_n = nullptr;
        SART_TRY(dev_alloc(&d_n, 1));
        unsigned long long n = e->lap_idx.size();
        CUDA_TRY(cudaMemcpyAsync(d_n, &n, sizeof(n), cudaMemcpyHostToDevice, e->stream));
        NCCL_TRY(api, api->Broadcast(d_n, d_n, 1, ncclUint64, 0, e->comm, e->stream));
        CUDA_TRY(cudaMemcpyAsync(&n, d_n, sizeof(n), cudaMemcpyDeviceToHost, e->stream));
        CUDA_TRY(cudaStreamSynchronize(e->stream));
        CUDA_TRY(cudaFree(d_n));
        if (n) {
            unsigned long long* d_i = nullptr;
            float* d_v = nullptr;
            SART_TRY(dev_alloc(&d_i, n, false));
            SART_TRY(dev_alloc(&d_v, n, false));
            if (e->cfg.rank == 0) {
                CUDA_TRY(cudaMemcpyAsync(d_i, e->lap_idx.data(), n * 8, cudaMemcpyHostToDevice, e->stream));
                CUDA_TRY(cudaMemcpyAsync(d_v, e->lap_val.data(), n * 4, cudaMemcpyHostToDevice, e->stream));
            }
            NCCL_TRY(api, api->Broadcast(d_i, d_i, n, ncclUint64, 0, e->comm, e->stream));
            NCCL_TRY(api, api->Broadcast(d_v, d_v, n, ncclFloat, 0, e->comm, e->stream));
            e->lap_idx.resize(n);
            e->lap_val.resize(n);
            CUDA_TRY(cudaMemcpyAsync(e->lap_idx.data(), d_i, n * 8, cudaMemcpyDeviceToHost, e->stream));
            CUDA_TRY(cudaMemcpyAsync(e->lap_val.data(), d_v, n * 4, cudaMemcpyDeviceToHost, e->stream));
            CUDA_TRY(cudaStreamSynchronize(e->stream));
            CUDA_TRY(cudaFree(d_i));
            CUDA_TRY(cudaFree(d_v));
        } else {
            e->lap_idx.clear();
            e->lap_val.clear();
        }
    }

    if (e->d_lap_rowptr) cudaFree(e->d_lap_rowptr);
    if (e->d_lap_col) cudaFree(e->d_lap_col);
    if (e->d_lap_val) cudaFree(e->d_lap_val);
    e->d_lap_rowptr = e->d_lap_col = nullptr;
    e->d_lap_val = nullptr;
    e->lap_nnz = e->lap_idx.size();
    if (!e->lap_nnz) return SART_OK;

    const uint64_t n = e->lap_nnz, V = e->V;
    if (n > 0x7fffffffULL) return fail(SART_ERR_INVALID_ARGUMENT, "Laplacian has more than 2^31 entries");
    std::vector<uint32_t> order(n);
    std::iota(order.begin(), order.end(), 0u);
    if (!std::is_sorted(e->lap_idx.begin(), e->lap_idx.end()))
        std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return e->lap_idx[a] < e->lap_idx[b]; });
    std::vector<int> rowptr(V + 1, 0), col(n);
    std::vector<float> val(n);
    for (uint64_t k = 0; k < n; ++k) {
        const uint64_t flat = e->lap_idx[order[k]];
        const uint64_t i = flat / V, j = flat % V;
        if (i >= V) return fail(SART_ERR_INVALID_ARGUMENT, "Laplacian index out of range for nvoxel");
        ++rowptr[i + 1];
        col[k] = (int)j;
        val[k] = e->lap_val[order[k]];
    }
    for (uint64_t i = 0; i < V; ++i) rowptr[i + 1] += rowptr[i];
    SART_TRY(dev_alloc(&e->d_lap_rowptr, V + 1, false));
    SART_TRY(dev_alloc(&e->d_lap_col, n, false));
    SART_TRY(dev_alloc(&e->d_lap_val, n, false));
    CUDA_TRY(cudaMemcpy(e->d_lap_rowptr, rowptr.data(), (V + 1) * sizeof(int), cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemcpy(e->d_lap_col, col.data(), n * sizeof(int), cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemcpy(e->d_lap_val, val.data(), n * sizeof(float), cudaMemcpyHostToDevice));
    return SART_OK;
}

int require(sart_engine* e, bool finalized) {
    if (!e) return fail(SART_ERR_INVALID_ARGUMENT, "engine is NULL");
    CUDA_TRY(cudaSetDevice(e->device));
    if (finalized && !e->finalized) return fail(SART_ERR_STATE, "sart_finalize has not been called");
    return SART_OK;
}

// Maps every rank's exchange region into this process (CUDA IPC over NVLink). All ranks of one box only; when the
// mapping is not possible (SART_EXCHANGE=nccl, more than 16 ranks, IPC refused) the engine keeps the NCCL all-reduce.
// COLLECTIVE: every rank takes the same decision, because the outcome is all-reduced (MIN) before it is adopted.
int setup_peer_exchange(sart_engine* e) {
    NcclApi* api = nccl_api();
    const char* mode = std::getenv("SART_EXCHANGE");
    int ok = !(mode && std::string(mode) == "nccl") && e->cfg.nranks <= kMaxPeers;
    // [one flag per rank, 128 B apart][one flag per rank and CTA of the whole-solve kernel][statistics: one flag per rank,
    // two slot arrays][buffer 0][buffer 1]
    // a buffer holds V + 1 doubles (per-sweep kernels) or one record [sum(fit^2)][chunk] per CTA (whole-solve kernel)
    const uint64_t nblk = std::max<uint64_t>((uint64_t)e->fu_cluster * (e->persistent ? e->fs_nclusters : e->fu_nclusters), 1);
    const uint64_t rank_flags_bytes = ceil_div((uint64_t)e->cfg.nranks * kFlagStride, 256) * 256;
    // The layout must be the same on every rank (a rank addresses its peers' regions with its own offsets), but the grid
    // is not always: ranks whose shards differ by a row can plan one cluster more or less on tiny problems. So the
    // regions are laid out for the largest grid any rank could run, not for this rank's own.
    constexpr uint64_t kLayoutCtas = 4096;  // >= 16 CTAs per cluster x resident clusters on any device
    if (nblk > kLayoutCtas) ok = 0;
    const uint64_t cta_flags_bytes = ceil_div((uint64_t)e->cfg.nranks * kLayoutCtas * 8, 256) * 256;
    const uint64_t stat_slots_bytes = ceil_div((uint64_t)kStatSlots * sizeof(StatPair), 256) * 256;
    const uint64_t flags_bytes = rank_flags_bytes + cta_flags_bytes + rank_flags_bytes + 2 * stat_slots_bytes;
    // ... or, in sart_solve_batch, the vectors of the frames of a group one after the other
    const uint64_t buf_bytes = ceil_div(std::max<uint64_t>(e->V + 2 * kLayoutCtas + 2, (uint64_t)kGroup * (e->V + 1)) * sizeof(double), 256) * 256;
    e->xchg_cta_flag_off = rank_flags_bytes;
    e->xchg_stat_flag_off = rank_flags_bytes + cta_flags_bytes;
    e->xchg_stat_slot_off[0] = e->xchg_stat_flag_off + rank_flags_bytes;
    e->xchg_stat_slot_off[1] = e->xchg_stat_slot_off[0] + stat_slots_bytes;
    e->xchg_buf_off[0] = flags_bytes;
    e->xchg_buf_off[1] = flags_bytes + buf_bytes;

    cudaIpcMemHandle_t mine;
    std::memset(&mine, 0, sizeof(mine));
    if (ok) {
        if (cudaMalloc(reinterpret_cast<void**>(&e->d_xchg), flags_bytes + 2 * buf_bytes) != cudaSuccess ||
            cudaMemsetAsync(e->d_xchg, 0, flags_bytes + 2 * buf_bytes, e->stream) != cudaSuccess ||
            cudaIpcGetMemHandle(&mine, e->d_xchg) != cudaSuccess) {
            (void)cudaGetLastError();
            ok = 0;
        }
    }
    // handles of all ranks (the zeroed flags are ordered before the handle leaves this rank: same stream)
    const int n = e->cfg.nranks;
    char *d_mine = nullptr, *d_all = nullptr;
    SART_TRY(dev_alloc(&d_mine, sizeof(mine), false));
    SART_TRY(dev_alloc(&d_all, sizeof(mine) * n, false));
    CUDA_TRY(cudaMemcpyAsync(d_mine, &mine, sizeof(mine), cudaMemcpyHostToDevice, e->stream));
    NCCL_TRY(api, api->AllGather(d_mine, d_all, sizeof(mine), ncclChar, e->comm, e->stream));
    std::vector<cudaIpcMemHandle_t> all(n);
    CUDA_TRY(cudaMemcpyAsync(all.data(), d_all, sizeof(mine) * n, cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    CUDA_TRY(cudaFree(d_mine));
    CUDA_TRY(cudaFree(d_all));

    std::vector<char*> peers(n, nullptr);
    e->peer_mappings.assign(n, nullptr);
    for (int r = 0; r < n && ok; ++r) {
        if (r == e->cfg.rank) {
            peers[r] = e->d_xchg;
            continue;
        }
        void* p = nullptr;
        if (cudaIpcOpenMemHandle(&p, all[r], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
            (void)cudaGetLastError();
            ok = 0;
            break;
        }
        e->peer_mappings[r] = p;
        peers[r] = static_cast<char*>(p);
    }
    // adopt the peer path only if every rank managed to map every region, and the whole-solve kernel only if every
    // rank can run it with the same grid (the voxel chunks of the in-kernel exchange are per CTA)
    double* d_ok = nullptr;
    SART_TRY(dev_alloc(&d_ok, 4, false));
    const double mine_ok[4] = {ok ? 1.0 : 0.0, (e->fused && e->persistent) ? 1.0 : 0.0, (double)nblk, -(double)nblk};
    double all_ok[4] = {0, 0, 0, 0};
    CUDA_TRY(cudaMemcpyAsync(d_ok, mine_ok, sizeof(mine_ok), cudaMemcpyHostToDevice, e->stream));
    NCCL_TRY(api, api->AllReduce(d_ok, d_ok, 4, ncclDouble, ncclMin, e->comm, e->stream));
    CUDA_TRY(cudaMemcpyAsync(all_ok, d_ok, sizeof(all_ok), cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    CUDA_TRY(cudaFree(d_ok));
    if (all_ok[1] < 0.5 || all_ok[2] != -all_ok[3]) e->persistent = false;
    if (all_ok[0] < 0.5) {
        e->peer_exchange = false;
        e->persistent = false;  // the whole-solve kernel has no NCCL variant
        return SART_OK;
    }
    SART_TRY(dev_alloc(&e->d_peers, (uint64_t)n, false));
    CUDA_TRY(cudaMemcpy(e->d_peers, peers.data(), sizeof(char*) * n, cudaMemcpyHostToDevice));
    e->xchg_epoch = 0;
    e->stat_epoch = 0;
    e->peer_exchange = true;
    return SART_OK;
}

void free_engine(sart_engine* e) {
    if (!e) return;
    cudaSetDevice(e->device);
    if (e->stream) cudaStreamSynchronize(e->stream);
    if (e->comm) {
        NcclApi* api = nccl_api();
        if (api) api->CommDestroy(e->comm);
    }
    for (void* p : e->peer_mappings)
        if (p) cudaIpcCloseMemHandle(p);
    if (e->d_xchg) cudaFree(e->d_xchg);
    if (e->d_peers) cudaFree(e->d_peers);
    if (e->d_solve_ctl) cudaFree(e->d_solve_ctl);
    if (e->h_solve_ctl) cudaFreeHost(e->h_solve_ctl);
    void* series_ptrs[] = {e->d_series_in, e->d_gspec, e->d_spec_scale, e->d_series_ctl, e->d_series_out, e->d_series_res,
                           e->d_stat_part, e->d_stat_max, e->d_stat_sum};
    for (void* p : series_ptrs)
        if (p) cudaFree(p);
    void* pair_ptrs[] = {e->d_pobs, e->d_pgd, e->d_pxio, e->d_psred, e->d_pf2part, e->d_pf2sum, e->d_pxd[0], e->d_pxd[1], e->d_pxf[0], e->d_pxf[1],
                         e->d_pg, e->d_pw, e->d_ppart, e->d_pfs, e->d_pst[0], e->d_pst[1]};
    for (void* p : pair_ptrs)
        if (p) cudaFree(p);
    if (e->h_pst) cudaFreeHost(e->h_pst);
    void* ptrs[] = {e->dH, e->d_rho, e->d_lambda, e->d_lap_rowptr, e->d_lap_col, e->d_lap_val, e->d_gd, e->d_xio,
                    e->d_obs, e->d_sred, e->d_f2part, e->d_f2, e->d_xd[0], e->d_xd[1], e->d_xf[0], e->d_xf[1], e->d_g,
                    e->d_rinv, e->d_fit, e->d_w, e->d_part, e->d_fs, e->d_st[0], e->d_st[1]};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    if (e->h_st) cudaFreeHost(e->h_st);
    for (int k = 0; k < sart_engine::kStagingBuffers; ++k) {
        if (e->h_staging[k]) cudaFreeHost(e->h_staging[k]);
        if (e->ev_staging[k]) cudaEventDestroy(e->ev_staging[k]);
    }
    cudaEvent_t evs[] = {e->ev_solve0, e->ev_solve1, e->ev_it0, e->ev_batch[0], e->ev_batch[1]};
    for (cudaEvent_t ev : evs)
        if (ev) cudaEventDestroy(ev);
    for (auto& s : e->ev_prof)
        for (int k = 0; k < 4; ++k)
            if (s[k]) cudaEventDestroy(s[k]);
    if (e->stream) cudaStreamDestroy(e->stream);
    delete e;
}

int create_impl(const sart_config* cfg, sart_engine* e) {
    e->cfg = *cfg;
    int count = 0;
    cudaError_t err = cudaGetDeviceCount(&count);
    if (err != cudaSuccess || count == 0)
        return fail(SART_ERR_NO_DEVICE, std::string("no CUDA device available (") +
                                            (err != cudaSuccess ? cudaGetErrorString(err) : "device count 0") +
                                            "); this engine has no CPU path");
    e->device = (cfg->device >= 0) ? cfg->device : cfg->rank % count;  // sartsolver_cuda.cpp:96-98
    if (e->device >= count) return fail(SART_ERR_INVALID_ARGUMENT, "device ordinal out of range");
    CUDA_TRY(cudaSetDevice(e->device));
    CUDA_TRY(cudaDeviceGetAttribute(&e->sm_count, cudaDevAttrMultiProcessorCount, e->device));
    CUDA_TRY(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking));

    SART_TRY(sart_shard_range(cfg->npixel, cfg->nranks, cfg->rank, &e->row_offset, &e->P));
    e->V = cfg->nvoxel;
    e->ld = ceil_div(e->V, 32) * 32;
    e->ld4 = e->ld / 4;
    if (e->P == 0) return fail(SART_ERR_INVALID_ARGUMENT, "this rank owns no rows (nranks > npixel)");
    if (e->ld4 * 8 >= (1ull << 32)) return fail(SART_ERR_INVALID_ARGUMENT, "nvoxel too large");
    e->cfg.device = e->device;

    SART_TRY(plan_launches(e));
    SART_TRY(plan_fused(e));

    size_t free_b = 0, total_b = 0;
    CUDA_TRY(cudaMemGetInfo(&free_b, &total_b));
    const uint64_t need = e->P * e->ld * 4;
    if (need + (512ull << 20) > free_b)
        return fail(SART_ERR_CUDA, "matrix shard of " + std::to_string(need >> 20) + " MiB does not fit in the " +
                                       std::to_string(free_b >> 20) + " MiB free on device " + std::to_string(e->device));
    SART_TRY(dev_alloc(&e->dH, e->P * e->ld));
    SART_TRY(dev_alloc(&e->d_rho, e->V));
    SART_TRY(dev_alloc(&e->d_lambda, e->P));
    SART_TRY(dev_alloc(&e->d_gd, e->P));
    SART_TRY(dev_alloc(&e->d_xio, e->V));
    SART_TRY(dev_alloc(&e->d_obs, e->V));
    SART_TRY(dev_alloc(&e->d_sred, e->V + 2));
    SART_TRY(dev_alloc(&e->d_f2part, (uint64_t)e->sm_count * 32));
    SART_TRY(dev_alloc(&e->d_f2, 1));
    for (int k = 0; k < 2; ++k) {
        SART_TRY(dev_alloc(&e->d_xd[k], e->V));
        SART_TRY(dev_alloc(&e->d_xf[k], e->ld));
        SART_TRY(dev_alloc(&e->d_st[k], 1));
    }
    SART_TRY(dev_alloc(&e->d_g, e->P));
    SART_TRY(dev_alloc(&e->d_rinv, e->P));
    SART_TRY(dev_alloc(&e->d_fit, e->P));
    SART_TRY(dev_alloc(&e->d_w, e->P));
    SART_TRY(dev_alloc(&e->d_part, (uint64_t)std::max(e->bp_slabs, e->fu_nclusters) * e->ld));
    SART_TRY(dev_alloc(&e->d_fs, 1));
    CUDA_TRY(cudaMallocHost(reinterpret_cast<void**>(&e->h_st), 2 * sizeof(IterState)));

    CUDA_TRY(cudaEventCreate(&e->ev_solve0));
    CUDA_TRY(cudaEventCreate(&e->ev_solve1));
    CUDA_TRY(cudaEventCreate(&e->ev_it0));
    CUDA_TRY(cudaEventCreate(&e->ev_batch[0]));
    CUDA_TRY(cudaEventCreate(&e->ev_batch[1]));
    e->profiling = env_int("SART_PROFILE_KERNELS", 1) != 0;
    if (e->profiling)
        for (auto& s : e->ev_prof)
            for (int k = 0; k < 4; ++k) CUDA_TRY(cudaEventCreate(&s[k]));
    e->stats.matrix_bytes = e->P * e->V * 4;
    CUDA_TRY(cudaDeviceSynchronize());
    return SART_OK;
}

}  // namespace

// ---- C ABI ----------------------------------------------------------------------------------------------------------

extern "C" {

const char* sart_last_error(void) { return g_error.c_str(); }

int sart_config_default(sart_config* cfg) {
    if (!cfg) return fail(SART_ERR_INVALID_ARGUMENT, "cfg is NULL");
    std::memset(cfg, 0, sizeof(*cfg));
    cfg->nranks = 1;
    cfg->device = -1;
    cfg->semantics = SART_SEMANTICS_CPU;
    cfg->max_iterations = 2000;
    cfg->ray_density_threshold = 1.e-6;
    cfg->ray_length_threshold = 1.e-6;
    cfg->conv_tolerance = 1.e-5;
    cfg->beta_laplace = 1.e-2;
    cfg->relaxation = 1.;
    return SART_OK;
}

int sart_shard_range(uint64_t npixel, int32_t nranks, int32_t rank, uint64_t* row_offset, uint64_t* nrows) {
    if (nranks < 1 || rank < 0 || rank >= nranks || !row_offset || !nrows)
        return fail(SART_ERR_INVALID_ARGUMENT, "sart_shard_range: need 0 <= rank < nranks and non-NULL outputs");
    const uint64_t n = (uint64_t)nranks, r = (uint64_t)rank;
    const uint64_t base = npixel / n, rem = npixel % n;
    *row_offset = r * base + std::min(r, rem);
    *nrows = (r < rem) ? base + 1 : base;
    return SART_OK;
}

int sart_device_query(int32_t device, sart_device_info* info) {
    if (!info) return fail(SART_ERR_INVALID_ARGUMENT, "info is NULL");
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || device < 0 || device >= count)
        return fail(SART_ERR_NO_DEVICE, "no such CUDA device");
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    std::memset(info, 0, sizeof(*info));
    std::snprintf(info->name, sizeof(info->name), "%s", prop.name);
    info->sm_count = prop.multiProcessorCount;
    info->cc_major = prop.major;
    info->cc_minor = prop.minor;
    info->l2_bytes = prop.l2CacheSize;
    CUDA_TRY(cudaSetDevice(device));
    size_t f = 0, t = 0;
    CUDA_TRY(cudaMemGetInfo(&f, &t));
    info->free_mem = f;
    info->total_mem = t;
    return SART_OK;
}

int sart_create(const sart_config* cfg, sart_engine** out) {
    if (!cfg || !out) return fail(SART_ERR_INVALID_ARGUMENT, "cfg or out is NULL");
    *out = nullptr;
    if (cfg->npixel == 0 || cfg->nvoxel == 0) return fail(SART_ERR_INVALID_ARGUMENT, "npixel and nvoxel must be positive");
    if (cfg->nranks < 1 || cfg->rank < 0 || cfg->rank >= cfg->nranks) return fail(SART_ERR_INVALID_ARGUMENT, "need 0 <= rank < nranks");
    if (cfg->semantics != SART_SEMANTICS_CPU && cfg->semantics != SART_SEMANTICS_GPU) return fail(SART_ERR_INVALID_ARGUMENT, "unknown semantics");
    if (cfg->sweep_mode < SART_SWEEP_AUTO || cfg->sweep_mode > SART_SWEEP_FUSED) return fail(SART_ERR_INVALID_ARGUMENT, "unknown sweep_mode");
    SART_TRY(check_params(*cfg));
    sart_engine* e = new sart_engine();
    const int rc = create_impl(cfg, e);
    if (rc) {
        const std::string keep = g_error;
        free_engine(e);
        g_error = keep;
        return rc;
    }
    *out = e;
    return SART_OK;
}

int sart_destroy(sart_engine* e) {
    free_engine(e);
    return SART_OK;
}

int sart_comm_unique_id(uint8_t id[SART_UNIQUE_ID_BYTES]) {
    static_assert(sizeof(ncclUniqueId) == SART_UNIQUE_ID_BYTES, "ncclUniqueId size");
    NcclApi* api = nccl_api();
    if (!api) return fail(SART_ERR_NCCL, nccl_storage().error);
    ncclUniqueId uid;
    NCCL_TRY(api, api->GetUniqueId(&uid));
    std::memcpy(id, &uid, sizeof(uid));
    return SART_OK;
}

int sart_comm_init(sart_engine* e, const uint8_t id[SART_UNIQUE_ID_BYTES]) {
    SART_TRY(require(e, false));
    if (e->cfg.nranks <= 1) return SART_OK;
    if (e->comm) return fail(SART_ERR_STATE, "communicator already initialised");
    NcclApi* api = nccl_api();
    if (!api) return fail(SART_ERR_NCCL, nccl_storage().error);
    ncclUniqueId uid;
    std::memcpy(&uid, id, sizeof(uid));
    NCCL_TRY(api, api->CommInitRank(&e->comm, e->cfg.nranks, uid, e->cfg.rank));
    return setup_peer_exchange(e);
}

int sart_upload_rows(sart_engine* e, uint64_t local_row0, uint64_t nrows, uint64_t col0, uint64_t ncols,
                     const float* src, uint64_t src_ld) {
    SART_TRY(require(e, false));
    if (!src || local_row0 + nrows > e->P || col0 + ncols > e->V || src_ld < ncols)
        return fail(SART_ERR_INVALID_ARGUMENT, "sart_upload_rows: block outside the shard or src_ld < ncols");
    if (nrows == 0 || ncols == 0) return SART_OK;
    CUDA_TRY(cudaMemcpy2DAsync(e->dH + local_row0 * e->ld + col0, e->ld * sizeof(float), src, src_ld * sizeof(float),
                               ncols * sizeof(float), nrows, cudaMemcpyHostToDevice, e->stream));
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    e->finalized = false;
    return SART_OK;
}

int sart_load_rows_file(sart_engine* e, const char* path, uint64_t file_offset, uint64_t file_ld, uint64_t local_row0,
                        uint64_t nrows, uint64_t col0, uint64_t ncols) {
    SART_TRY(require(e, false));
    if (!path || local_row0 + nrows > e->P || col0 + ncols > e->V || file_ld < ncols)
        return fail(SART_ERR_INVALID_ARGUMENT, "sart_load_rows_file: block outside the shard or file_ld < ncols");
    if (nrows == 0 || ncols == 0) return SART_OK;
    const int fd = ::open(path, O_RDONLY);
    if (fd < 0) return fail(SART_ERR_INVALID_ARGUMENT, std::string("cannot open ") + path + ": " + std::strerror(errno));

    // ring of pinned buffers: the read of slab k+1 (host thread, pread) overlaps the copy of slab k (DMA engine)
    const uint64_t want = (uint64_t)std::max(1, env_int("SART_STAGING_MB", 16)) << 20;
    const uint64_t row_bytes = ncols * sizeof(float);
    const uint64_t slab_rows = std::max<uint64_t>(1, want / row_bytes);
    const uint64_t need = slab_rows * row_bytes;
    int rc = SART_OK;
    if (e->staging_bytes < need) {
        for (int k = 0; k < sart_engine::kStagingBuffers; ++k) {
            if (e->h_staging[k]) cudaFreeHost(e->h_staging[k]);
            e->h_staging[k] = nullptr;
            if (cudaHostAlloc(reinterpret_cast<void**>(&e->h_staging[k]), need, cudaHostAllocDefault) != cudaSuccess) {
                ::close(fd);
                e->staging_bytes = 0;
                return fail(SART_ERR_CUDA, "cannot allocate pinned staging memory");
            }
            if (!e->ev_staging[k] && cudaEventCreateWithFlags(&e->ev_staging[k], cudaEventDisableTiming) != cudaSuccess) {
                ::close(fd);
                return fail(SART_ERR_CUDA, "cannot create staging event");
            }
        }
        e->staging_bytes = need;
    }
    bool in_use[sart_engine::kStagingBuffers] = {false, false, false};
    int k = 0;
    for (uint64_t r = 0; r < nrows && rc == SART_OK; r += slab_rows, k = (k + 1) % sart_engine::kStagingBuffers) {
        const uint64_t n = std::min(slab_rows, nrows - r);
        if (in_use[k] && cudaEventSynchronize(e->ev_staging[k]) != cudaSuccess) {
            rc = fail(SART_ERR_CUDA, "staging event failed");
            break;
        }
        char* dst = e->h_staging[k];
        // rows of the block are contiguous in the file only when the block spans whole file rows
        const bool contiguous = (ncols == file_ld);
        const uint64_t chunks = contiguous ? 1 : n;
        const uint64_t chunk_bytes = contiguous ? n * row_bytes : row_bytes;
        for (uint64_t c = 0; c < chunks && rc == SART_OK; ++c) {
            uint64_t done = 0;
            const uint64_t off = file_offset + ((r + c) * file_ld) * sizeof(float);
            while (done < chunk_bytes) {
                const ssize_t got = ::pread(fd, dst + c * row_bytes + done, chunk_bytes - done, (off_t)(off + done));
                if (got <= 0) {
                    rc = fail(SART_ERR_INVALID_ARGUMENT, std::string("short read on ") + path +
                                                             (got < 0 ? std::string(": ") + std::strerror(errno) : ""));
                    break;
                }
                done += (uint64_t)got;
            }
        }
        if (rc != SART_OK) break;
        if (cudaMemcpy2DAsync(e->dH + (local_row0 + r) * e->ld + col0, e->ld * sizeof(float), dst, row_bytes, row_bytes, n,
                              cudaMemcpyHostToDevice, e->stream) != cudaSuccess ||
            cudaEventRecord(e->ev_staging[k], e->stream) != cudaSuccess) {
            rc = fail(SART_ERR_CUDA, std::string("copy into the shard failed: ") + cudaGetErrorString(cudaGetLastError()));
            break;
        }
        in_use[k] = true;
    }
    ::close(fd);
    cudaError_t err = cudaStreamSynchronize(e->stream);
    if (rc == SART_OK && err != cudaSuccess) rc = fail(SART_ERR_CUDA, std::string("CUDA error: ") + cudaGetErrorString(err));
    e->finalized = false;
    return rc;
}

int sart_fill_synthetic_kind(sart_engine* e, uint64_t seed, int32_t kind) {
    SART_TRY(require(e, false));
    const dim3 grid((uint32_t)std::min<uint64_t>(ceil_div(e->ld4, 256), 64), (uint32_t)std::min<uint64_t>(e->P, 32768));
    if (kind == SART_SYNTHETIC_DENSE) {
        const int exponent = (int)std::lround(std::log2((double)e->V * 98304.0));
        const float scale = std::ldexp(1.0f, -exponent);
        fill_synthetic_kernel<<<grid, 256, 0, e->stream>>>(e->dH, e->ld4, e->P, e->V, e->row_offset, seed, scale);
    } else if (kind == SART_SYNTHETIC_LOS) {
        // the voxel grid nx x nz: nvoxel factored as squarely as possible (the tests' voxel_grid_shape)
        uint64_t nx = (uint64_t)std::floor(std::sqrt((double)e->V));
        while (nx > 1 && e->V % nx) --nx;
        const uint64_t nz = e->V / nx;
        // mean row sum of q: a ray crosses about 2430 sqrt(V) units of "direct" weight, plus 128 per background element
        const int exponent = (int)std::lround(std::log2(1024.0 * 2430.0 * std::sqrt((double)e->V) + 128.0 * (double)e->V));
        const float scale = std::ldexp(1.0f, -exponent);
        fill_los_kernel<<<grid, 256, 0, e->stream>>>(e->dH, e->ld4, e->P, e->V, e->row_offset, seed, scale, (uint32_t)nx, (uint32_t)nz);
    } else {
        return fail(SART_ERR_INVALID_ARGUMENT, "sart_fill_synthetic_kind: unknown kind");
    }
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    e->finalized = false;
    return SART_OK;
}

int sart_fill_synthetic(sart_engine* e, uint64_t seed) { return sart_fill_synthetic_kind(e, seed, SART_SYNTHETIC_DENSE); }

int sart_download_rows(sart_engine* e, uint64_t local_row0, uint64_t nrows, float* dst) {
    SART_TRY(require(e, false));
    if (!dst || local_row0 + nrows > e->P) return fail(SART_ERR_INVALID_ARGUMENT, "sart_download_rows: rows outside the shard");
    if (nrows == 0) return SART_OK;
    CUDA_TRY(cudaMemcpy2DAsync(dst, e->V * sizeof(float), e->dH + local_row0 * e->ld, e->ld * sizeof(float),
                               e->V * sizeof(float), nrows, cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    return SART_OK;
}

int sart_set_laplacian_coo(sart_engine* e, const uint64_t* flat_index, const float* value, uint64_t nnz) {
    SART_TRY(require(e, false));
    if (nnz && (!flat_index || !value)) return fail(SART_ERR_INVALID_ARGUMENT, "Laplacian arrays are NULL");
    e->lap_idx.assign(flat_index, flat_index + nnz);
    e->lap_val.assign(value, value + nnz);
    e->finalized = false;
    return SART_OK;
}

int sart_finalize(sart_engine* e) {
    SART_TRY(require(e, false));
    // ray_length_i: row sums of the local rows
    forward_f64_kernel<<<(uint32_t)ceil_div(e->P, 8), 256, 0, e->stream>>>(e->dH, e->ld4, e->P, e->V, nullptr, e->d_lambda);
    CUDA_TRY(cudaGetLastError());
    // ray_density_j: column sums, slab partials then fixed-order sum, then over ranks
    const uint32_t cblocks = (uint32_t)ceil_div(e->V, 256);
    uint64_t slabs = std::max<uint64_t>(1, (uint64_t)e->sm_count * 8 / cblocks);
    slabs = std::min<uint64_t>(slabs, e->P);
    const uint32_t rows_per_slab = (uint32_t)ceil_div(e->P, slabs);
    slabs = ceil_div(e->P, rows_per_slab);
    double* d_part = nullptr;
    SART_TRY(dev_alloc(&d_part, slabs * e->V, false));
    const int rc = [&]() -> int {  // so that the partials are freed on every way out
        col_sums_f64_kernel<<<dim3(cblocks, (uint32_t)slabs), 256, 0, e->stream>>>(e->dH, e->ld, e->P, e->V, rows_per_slab, d_part);
        CUDA_TRY(cudaGetLastError());
        reduce_f64_partials_kernel<<<cblocks, 256, 0, e->stream>>>(d_part, (uint32_t)slabs, e->V, e->d_rho);
        CUDA_TRY(cudaGetLastError());
        SART_TRY(allreduce_f64(e, e->d_rho, e->V, ncclSum));
        CUDA_TRY(cudaStreamSynchronize(e->stream));
        return SART_OK;
    }();
    const cudaError_t free_err = cudaFree(d_part);
    SART_TRY(rc);
    CUDA_TRY(free_err);
    SART_TRY(build_laplacian(e));
    e->finalized = true;
    return SART_OK;
}

int sart_get_ray_sums(sart_engine* e, double* ray_density, double* ray_length) {
    SART_TRY(require(e, true));
    if (ray_density) CUDA_TRY(cudaMemcpy(ray_density, e->d_rho, e->V * sizeof(double), cudaMemcpyDeviceToHost));
    if (ray_length) CUDA_TRY(cudaMemcpy(ray_length, e->d_lambda, e->P * sizeof(double), cudaMemcpyDeviceToHost));
    return SART_OK;
}

#define SART_SETTER(name, field, type)                      \
    int name(sart_engine* e, type v) {                      \
        if (!e) return fail(SART_ERR_INVALID_ARGUMENT, "engine is NULL"); \
        sart_config c = e->cfg;                             \
        c.field = v;                                        \
        SART_TRY(check_params(c));                          \
        e->cfg.field = v;                                   \
        return SART_OK;                                     \
    }
SART_SETTER(sart_set_ray_density_threshold, ray_density_threshold, double)
SART_SETTER(sart_set_ray_length_threshold, ray_length_threshold, double)
SART_SETTER(sart_set_convolution_tolerance, conv_tolerance, double)
SART_SETTER(sart_set_beta_laplace, beta_laplace, double)
SART_SETTER(sart_set_relaxation, relaxation, double)
SART_SETTER(sart_set_max_iterations, max_iterations, int32_t)
#undef SART_SETTER

int sart_set_fixed_iterations(sart_engine* e, int32_t on) {
    if (!e) return fail(SART_ERR_INVALID_ARGUMENT, "engine is NULL");
    e->cfg.fixed_iterations = on ? 1 : 0;
    return SART_OK;
}

int sart_get_config(const sart_engine* e, sart_config* out) {
    if (!e || !out) return fail(SART_ERR_INVALID_ARGUMENT, "engine or out is NULL");
    *out = e->cfg;
    return SART_OK;
}

int sart_stage_measurement(sart_engine* e, const double* measurement) {
    SART_TRY(require(e, true));
    if (!measurement) return fail(SART_ERR_INVALID_ARGUMENT, "measurement is NULL");
    CUDA_TRY(cudaMemcpyAsync(e->d_gd, measurement, e->P * sizeof(double), cudaMemcpyHostToDevice, e->stream));
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    e->measurement_staged = true;
    return SART_OK;
}

int sart_stage_guess(sart_engine* e, const double* solution) {
    SART_TRY(require(e, true));
    if (!solution) return fail(SART_ERR_INVALID_ARGUMENT, "solution is NULL");
    CUDA_TRY(cudaMemcpyAsync(e->d_xio, solution, e->V * sizeof(double), cudaMemcpyHostToDevice, e->stream));
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    e->guess_staged = true;
    return SART_OK;
}

int sart_solve_resident(sart_engine* e, int32_t has_guess, int32_t* status, int32_t* iterations) {
    SART_TRY(require(e, true));
    if (!e->measurement_staged) return fail(SART_ERR_STATE, "no measurement on the device (sart_stage_measurement)");
    if (has_guess && !e->guess_staged) return fail(SART_ERR_STATE, "no starting solution on the device");
    e->launches = 0;
    CUDA_TRY(cudaEventRecord(e->ev_solve0, e->stream));
    SART_TRY(run_solve(e, has_guess));
    CUDA_TRY(cudaEventRecord(e->ev_solve1, e->stream));
    CUDA_TRY(cudaEventSynchronize(e->ev_solve1));
    float ms = 0.f;
    CUDA_TRY(cudaEventElapsedTime(&ms, e->ev_solve0, e->ev_solve1));
    e->stats.solve_ms = ms;
    e->stats.gpu_launches = e->launches;
    e->stats.h2d_bytes = e->stats.d2h_bytes = 0;
    e->guess_staged = true;  // the solution stays on the device: the next frame may warm-start from it
    if (status) *status = e->stats.status;
    if (iterations) *iterations = e->stats.iterations;
    return SART_OK;
}

int sart_fetch_solution(sart_engine* e, double* solution) {
    SART_TRY(require(e, true));
    if (!solution) return fail(SART_ERR_INVALID_ARGUMENT, "solution is NULL");
    CUDA_TRY(cudaMemcpyAsync(solution, e->d_xio, e->V * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    return SART_OK;
}

int sart_solve(sart_engine* e, const double* measurement, double* solution, int32_t has_guess, int32_t* status,
               int32_t* iterations) {
    SART_TRY(require(e, true));
    if (!measurement || !solution) return fail(SART_ERR_INVALID_ARGUMENT, "measurement or solution is NULL");
    e->launches = 0;
    CUDA_TRY(cudaEventRecord(e->ev_solve0, e->stream));
    CUDA_TRY(cudaMemcpyAsync(e->d_gd, measurement, e->P * sizeof(double), cudaMemcpyHostToDevice, e->stream));
    e->measurement_staged = true;
    uint64_t h2d = e->P * sizeof(double);
    if (has_guess) {
        CUDA_TRY(cudaMemcpyAsync(e->d_xio, solution, e->V * sizeof(double), cudaMemcpyHostToDevice, e->stream));
        h2d += e->V * sizeof(double);
    }
    SART_TRY(run_solve(e, has_guess));
    CUDA_TRY(cudaMemcpyAsync(solution, e->d_xio, e->V * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(cudaEventRecord(e->ev_solve1, e->stream));
    CUDA_TRY(cudaEventSynchronize(e->ev_solve1));
    float ms = 0.f;
    CUDA_TRY(cudaEventElapsedTime(&ms, e->ev_solve0, e->ev_solve1));
    e->stats.solve_ms = ms;
    e->stats.gpu_launches = e->launches;
    e->stats.h2d_bytes = h2d;
    e->stats.d2h_bytes = e->V * sizeof(double) + 2 * sizeof(IterState);
    e->guess_staged = true;
    if (status) *status = e->stats.status;
    if (iterations) *iterations = e->stats.iterations;
    return SART_OK;
}

// A warm-started series in one call (the frame loop of source/main.cpp:125-140 without --no_guess). Frame by frame as
// sart_solve would do it, except that -- when the whole-solve kernel runs (linear solver, CPU semantics) -- the sweep
// that is predicted to be a frame's last one back-projects the NEXT frame's measurement: that sweep evaluates the
// frame's final iterate anyway, and its forward projection is, up to the ratio of the two normalisations, the next
// frame's first one. A frame that stops as predicted therefore hands the next frame its first update and saves it one
// pass over the matrix (2 instead of 3 for a frame that needs two iterations); a frame that does not stop there repeats
// that sweep with its own measurement (one pass lost). What a frame leaves for the next one (handed over or not, buffer
// flip, iteration counts for the prediction, exchange epoch) stays on the device in a SeriesCtl block, results collect
// in device buffers, and the host enqueues the whole chunk without waiting for any frame: one synchronisation per call.
static int solve_series_impl(sart_engine* e, int32_t nframes, const double* measurements, const double* guess, double* solutions,
                             int32_t* status, int32_t* iterations) {
    const uint32_t vblocks = (uint32_t)ceil_div(e->V, 256), pblocks = (uint32_t)ceil_div(e->P, 256);
    const bool whole = e->fused && e->persistent && (e->cfg.nranks <= 1 || e->peer_exchange);
    const bool chained = whole && !e->cfg.logarithmic && e->cfg.semantics == SART_SEMANTICS_CPU &&
                         env_int("SART_SERIES_LOOKAHEAD", 1) != 0;
    e->launches = 0;
    CUDA_TRY(cudaEventRecord(e->ev_solve0, e->stream));
    uint64_t h2d = 0, d2h = 0;
    if (nframes > 0 && guess) {
        CUDA_TRY(cudaMemcpyAsync(e->d_xio, guess, e->V * sizeof(double), cudaMemcpyHostToDevice, e->stream));
        h2d += e->V * sizeof(double);
    }

    if (!chained) {  // the plain frame loop
        int total_sweeps = 0, total_fused = 0;
        for (int32_t f = 0; f < nframes; ++f) {
            CUDA_TRY(cudaMemcpyAsync(e->d_gd, measurements + (uint64_t)f * e->P, e->P * sizeof(double), cudaMemcpyHostToDevice, e->stream));
            h2d += e->P * sizeof(double);
            SART_TRY(run_solve(e, (f > 0 || guess) ? 1 : 0));
            // passes over the matrix: set-up + one per iterate with the fused sweep, two per iteration without
            const int it = e->stats.iterations;
            total_sweeps += e->last_prologue_sweeps + (e->stats.sweep_mode == SART_SWEEP_FUSED ? it + 1 : 2 * it + 1);
            total_fused += e->stats.fused_launches;
            if (status) status[f] = e->stats.status;
            if (iterations) iterations[f] = e->stats.iterations;
            CUDA_TRY(cudaMemcpyAsync(solutions + (uint64_t)f * e->V, e->d_xio, e->V * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
            d2h += e->V * sizeof(double) + 2 * sizeof(IterState);
        }
        CUDA_TRY(cudaEventRecord(e->ev_solve1, e->stream));
        CUDA_TRY(cudaEventSynchronize(e->ev_solve1));
        e->stats.sweeps = total_sweeps;
        e->stats.fused_launches = total_fused;
        e->stats.fp_launches = e->stats.bp_launches = 0;
    } else {
        if (!e->d_gspec) {
            SART_TRY(dev_alloc(&e->d_gspec, e->P));
            SART_TRY(dev_alloc(&e->d_spec_scale, 1));
            SART_TRY(dev_alloc(&e->d_series_ctl, 1));
        }
        if (nframes > e->series_capacity) {
            if (e->d_series_in) CUDA_TRY(cudaFree(e->d_series_in));
            if (e->d_series_out) CUDA_TRY(cudaFree(e->d_series_out));
            if (e->d_series_res) CUDA_TRY(cudaFree(e->d_series_res));
            e->d_series_in = e->d_series_out = nullptr;
            e->d_series_res = nullptr;
            e->series_capacity = 0;
            SART_TRY(dev_alloc(&e->d_series_in, (uint64_t)nframes * e->P, false));
            SART_TRY(dev_alloc(&e->d_series_out, (uint64_t)nframes * e->V, false));
            SART_TRY(dev_alloc(&e->d_series_res, 2 * (uint64_t)nframes, false));
            e->series_capacity = nframes;
        }
        SeriesCtl init = {};
        init.it_prev = init.it_prev2 = -1;
        init.epoch = e->xchg_epoch;
        CUDA_TRY(cudaMemcpyAsync(e->d_series_ctl, &init, sizeof(init), cudaMemcpyHostToDevice, e->stream));
        // the whole chunk of measurements in one copy: nothing the host does per frame waits for the device afterwards
        CUDA_TRY(cudaMemcpyAsync(e->d_series_in, measurements, (uint64_t)nframes * e->P * sizeof(double), cudaMemcpyHostToDevice, e->stream));
        h2d += (uint64_t)nframes * e->P * sizeof(double);
        // statistics of every frame of the chunk in one go: no collective is left inside the frame loop
        SART_TRY(compute_frame_stats(e, e->d_series_in, e->P, nframes));
        int setup_sweeps = 0;
        for (int32_t f = 0; f < nframes; ++f) {
            const bool more = f + 1 < nframes;
            int sweeps = 0;
            // set-up of frame f; its starting solution is left alone on the device when frame f - 1 handed x_1 over
            const double* gd = e->d_series_in + (uint64_t)f * e->P;
            SART_TRY(frame_prologue(e, (f > 0 || guess) ? 1 : 0, &sweeps, e->d_series_ctl, gd, f));
            setup_sweeps += sweeps;
            if (more) {  // look-ahead: the next frame's measurement in this frame's normalisation
                const double* gd_next = gd + e->P;
                series_lookahead_kernel<<<pblocks, 256, 0, e->stream>>>(gd_next, e->P, e->d_fs, e->d_stat_max + f + 1, e->d_gspec,
                                                                        e->d_spec_scale);
                CUDA_TRY(cudaGetLastError());
                ++e->launches;
            }
            SeriesLaunch sl;
            sl.ctl = e->d_series_ctl;
            sl.has_next = more ? 1 : 0;
            sl.out_status = e->d_series_res + f;
            sl.out_iterations = e->d_series_res + nframes + f;
            bool launched = false;
            SART_TRY(launch_solve(e, &launched, sl));
            if (!launched) return fail(SART_ERR_CUDA, "the device refused the cooperative launch of the whole-solve kernel in the middle of a series");
            finish_solution_kernel<<<vblocks, 256, 0, e->stream>>>(e->d_xd[0], e->d_xd[1], e->d_st[0], e->d_fs, e->d_xio, e->V);
            CUDA_TRY(cudaGetLastError());
            ++e->launches;
            CUDA_TRY(cudaMemcpyAsync(e->d_series_out + (uint64_t)f * e->V, e->d_xio, e->V * sizeof(double), cudaMemcpyDeviceToDevice, e->stream));
        }
        // one synchronisation for the whole chunk
        SeriesCtl fin = {};
        std::vector<int> res(2 * (size_t)nframes);
        if (nframes > 0) {
            CUDA_TRY(cudaMemcpyAsync(solutions, e->d_series_out, (uint64_t)nframes * e->V * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
            CUDA_TRY(cudaMemcpyAsync(res.data(), e->d_series_res, res.size() * sizeof(int), cudaMemcpyDeviceToHost, e->stream));
        }
        CUDA_TRY(cudaMemcpyAsync(&fin, e->d_series_ctl, sizeof(fin), cudaMemcpyDeviceToHost, e->stream));
        CUDA_TRY(cudaEventRecord(e->ev_solve1, e->stream));
        CUDA_TRY(cudaEventSynchronize(e->ev_solve1));
        d2h += (uint64_t)nframes * (e->V * sizeof(double) + 2 * sizeof(int)) + sizeof(fin);
        e->xchg_epoch = fin.epoch;
        if (fin.timed_out)
            return fail(SART_ERR_NCCL, "in-kernel barrier or peer-memory exchange timed out during the series: a CTA or a rank of the job did not arrive");
        for (int32_t f = 0; f < nframes; ++f) {
            if (status) status[f] = res[f];
            if (iterations) iterations[f] = res[nframes + f];
        }
        e->stats.status = nframes ? res[nframes - 1] : 0;
        e->stats.iterations = nframes ? res[2 * nframes - 1] : 0;
        e->stats.sweeps = setup_sweeps + (int)fin.passes;
        e->stats.fused_launches = (int)fin.passes;
        e->stats.solve_launches = nframes;
        e->stats.fp_launches = fin.frames_handed;
        e->stats.bp_launches = fin.sweeps_repeated;
        e->stats.fp_ms_avg = e->stats.bp_ms_avg = e->stats.tail_ms_avg = 0;
        e->stats.sweep_mode = SART_SWEEP_FUSED;
        e->stats.exchange_mode = e->cfg.nranks <= 1 ? SART_EXCHANGE_NONE : SART_EXCHANGE_PEER;
    }
    float ms = 0.f;
    CUDA_TRY(cudaEventElapsedTime(&ms, e->ev_solve0, e->ev_solve1));
    e->stats.solve_ms = ms;
    if (chained) e->stats.fused_ms_avg = e->stats.fused_launches ? ms / e->stats.fused_launches : 0;  // per pass, everything included
    e->stats.gpu_launches = e->launches;
    e->stats.h2d_bytes = h2d;
    e->stats.d2h_bytes = d2h;
    if (!chained) e->measurement_staged = nframes > 0;
    e->guess_staged = nframes > 0;
    return SART_OK;
}

int sart_solve_series(sart_engine* e, int32_t nframes, const double* measurements, const double* guess, double* solutions,
                      int32_t* status, int32_t* iterations) {
    SART_TRY(require(e, true));
    if (nframes < 0 || (nframes > 0 && (!measurements || !solutions)))
        return fail(SART_ERR_INVALID_ARGUMENT, "sart_solve_series: NULL buffers or negative frame count");
    return solve_series_impl(e, nframes, measurements, guess, solutions, status, iterations);
}

int sart_solve_batch(sart_engine* e, int32_t nframes, const double* measurements, double* solutions, int32_t* status,
                     int32_t* iterations) {
    SART_TRY(require(e, true));
    if (nframes < 0 || (nframes > 0 && (!measurements || !solutions)))
        return fail(SART_ERR_INVALID_ARGUMENT, "sart_solve_batch: NULL buffers or negative frame count");
    SART_TRY(plan_group(e));
    int32_t f = 0;
    int total_launches = 0;
    double total_ms = 0;
    uint64_t h2d = 0, d2h = 0;
    while (f < nframes) {
        const int nf = std::min<int32_t>(e->group_frames, nframes - f);
        if (e->group_ok && nf >= 2) {
            e->launches = 0;
            CUDA_TRY(cudaEventRecord(e->ev_solve0, e->stream));
            CUDA_TRY(cudaMemcpyAsync(e->d_pgd, measurements + (uint64_t)f * e->P, (uint64_t)nf * e->P * sizeof(double),
                                     cudaMemcpyHostToDevice, e->stream));
            IterState fin[kGroup];
            SART_TRY(run_solve_group(e, nf, fin));
            CUDA_TRY(cudaMemcpyAsync(solutions + (uint64_t)f * e->V, e->d_pxio, (uint64_t)nf * e->V * sizeof(double),
                                     cudaMemcpyDeviceToHost, e->stream));
            CUDA_TRY(cudaEventRecord(e->ev_solve1, e->stream));
            CUDA_TRY(cudaEventSynchronize(e->ev_solve1));
            float ms = 0.f;
            CUDA_TRY(cudaEventElapsedTime(&ms, e->ev_solve0, e->ev_solve1));
            total_ms += ms;
            total_launches += e->launches;
            h2d += (uint64_t)nf * e->P * sizeof(double);
            d2h += (uint64_t)nf * e->V * sizeof(double);
            for (int b = 0; b < nf; ++b) {
                if (status) status[f + b] = fin[b].status;
                if (iterations) iterations[f + b] = fin[b].iterations;
            }
            e->stats.status = fin[nf - 1].status;
            e->stats.iterations = fin[nf - 1].iterations;
            e->stats.sweep_mode = SART_SWEEP_SPLIT;
            e->stats.exchange_mode = e->cfg.nranks <= 1 ? SART_EXCHANGE_NONE : (e->peer_exchange ? SART_EXCHANGE_PEER : SART_EXCHANGE_NCCL);
            f += nf;
        } else {
            int32_t st = 0, it = 0;
            SART_TRY(sart_solve(e, measurements + (uint64_t)f * e->P, solutions + (uint64_t)f * e->V, 0, &st, &it));
            if (status) status[f] = st;
            if (iterations) iterations[f] = it;
            total_ms += e->stats.solve_ms;
            total_launches += e->stats.gpu_launches;
            h2d += e->stats.h2d_bytes;
            d2h += e->stats.d2h_bytes;
            f += 1;
        }
    }
    e->stats.solve_ms = total_ms;
    e->stats.gpu_launches = total_launches;
    e->stats.h2d_bytes = h2d;
    e->stats.d2h_bytes = d2h;
    e->guess_staged = false;
    return SART_OK;
}

int sart_forward_project(sart_engine* e, const double* x, double* out) {
    SART_TRY(require(e, false));
    if (!x || !out) return fail(SART_ERR_INVALID_ARGUMENT, "x or out is NULL");
    double *d_x = nullptr, *d_o = nullptr;
    SART_TRY(dev_alloc(&d_x, e->V, false));
    SART_TRY(dev_alloc(&d_o, e->P, false));
    CUDA_TRY(cudaMemcpyAsync(d_x, x, e->V * sizeof(double), cudaMemcpyHostToDevice, e->stream));
    forward_f64_kernel<<<(uint32_t)ceil_div(e->P, 8), 256, 0, e->stream>>>(e->dH, e->ld4, e->P, e->V, d_x, d_o);
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpyAsync(out, d_o, e->P * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    CUDA_TRY(cudaFree(d_x));
    CUDA_TRY(cudaFree(d_o));
    return SART_OK;
}

int sart_kernel_forward(sart_engine* e, const double* x, double* fit) {
    SART_TRY(require(e, false));
    if (!x || !fit) return fail(SART_ERR_INVALID_ARGUMENT, "x or fit is NULL");
    std::vector<float> xf(e->ld, 0.f), out(e->P);
    for (uint64_t j = 0; j < e->V; ++j) xf[j] = (float)x[j];
    CUDA_TRY(cudaMemcpyAsync(e->d_xf[0], xf.data(), e->ld * sizeof(float), cudaMemcpyHostToDevice, e->stream));
    SART_TRY(launch_fp(e, e->d_xf[0], nullptr));
    CUDA_TRY(cudaMemcpyAsync(out.data(), e->d_fit, e->P * sizeof(float), cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    for (uint64_t i = 0; i < e->P; ++i) fit[i] = out[i];
    return SART_OK;
}

int sart_kernel_backward(sart_engine* e, const double* w, double* s) {
    SART_TRY(require(e, false));
    if (!w || !s) return fail(SART_ERR_INVALID_ARGUMENT, "w or s is NULL");
    std::vector<float> wf(e->P);
    for (uint64_t i = 0; i < e->P; ++i) wf[i] = (float)w[i];
    CUDA_TRY(cudaMemcpyAsync(e->d_w, wf.data(), e->P * sizeof(float), cudaMemcpyHostToDevice, e->stream));
    SART_TRY(launch_bp(e, nullptr));
    SART_TRY(launch_reduce(e, nullptr));
    CUDA_TRY(cudaMemcpyAsync(s, e->d_sred, e->V * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    return SART_OK;
}

int sart_get_stats(const sart_engine* e, sart_stats* out) {
    if (!e || !out) return fail(SART_ERR_INVALID_ARGUMENT, "engine or out is NULL");
    *out = e->stats;
    out->matrix_bytes = e->P * e->V * 4;
    out->fused_cluster_size = e->fused ? (int32_t)e->fu_cluster : 0;
    out->fused_cluster_count = e->fused ? (int32_t)e->fu_nclusters : 0;
    return SART_OK;
}

}  // extern "C"

// =====================================================================================================================
// Back projection for callers that must not allocate device memory.
//
// Who that is: hosts that gave the whole GPU to their own arena at start-up -- a data-acquisition loop that sized every
// buffer up front, a framework sitting on a caching allocator, a partition with a hard memory cap -- and that call the
// back projection from inside that loop. There, a cudaMalloc inside a library can fail outright (nothing is left outside
// the arena) and, when it succeeds, it serialises the whole device, which shows up as a hitch in every other stream of
// the loop. sart_create() allocates, so such callers cannot use an engine; they own the matrix, the weights and the
// output themselves and only need the kernel.
//
// The slab partials that bp_cols_kernel keeps in a cudaMalloc'ed buffer live here in a fixed array inside the library
// image instead. Consequences a caller has to know:
//   * The array costs SART_NOALLOC_SCRATCH_BYTES (16 MiB) on every GPU that loads the library, until the process exits.
//   * It bounds the problem: the padded row length ld may not exceed SART_NOALLOC_MAX_LD floats; anything larger is
//     refused with SART_ERR_INVALID_ARGUMENT, nothing is launched.
//   * It is ONE array per GPU, shared by every call and never cleared. No element has to start at zero: the reduction
//     reads only the slabs x ld partials that the first kernel of the same call has just written, and stale values
//     from an earlier call beyond that range are never read. Two calls that overlap on the same GPU (different streams
//     or host threads without ordering between them) write the same partials and race; order them yourself.
//   * The first call on each device looks the array's device address up (cudaGetSymbolAddress with that device
//     current; the name of a __device__ array is only a host-side handle and must never be handed to a kernel). That
//     first call may make the runtime load the module, so make it before any stream capture begins.
// The function calls no cudaMalloc / cudaFree / cudaMallocAsync and no library routine that allocates on the device.
// =====================================================================================================================
#define SART_NOALLOC_SCRATCH_FLOATS (SART_NOALLOC_SCRATCH_BYTES / sizeof(float))
static_assert(SART_NOALLOC_SCRATCH_FLOATS == SART_NOALLOC_MAX_LD,
              "the documented limit (SART_NOALLOC_MAX_LD) is one row of partials in the built-in array");

static __device__ __align__(256) float g_noalloc_scratch[SART_NOALLOC_SCRATCH_FLOATS];

namespace {

constexpr int kNoallocMaxDevices = 64;
float* g_noalloc_ptr[kNoallocMaxDevices];  // device address of g_noalloc_scratch on each device, null until looked up
int g_noalloc_sms[kNoallocMaxDevices];
std::mutex g_noalloc_mutex;

// bp_cols_kernel with its partials in the library's own array (address passed in: see the note on the first call)
template <int NT, int U>
__global__ void __launch_bounds__(NT) bp_cols_scratch_kernel(const BpArgs a, float* scratch) {
    bp_cols_body<NT, U>(a, scratch);
}

}  // namespace

extern "C" int sart_back_project_noalloc(const float* d_matrix, uint64_t nrows, uint64_t ncols, uint64_t ld,
                                         const float* d_weights, double* d_out, void* cuda_stream) {
    if (!d_matrix || !d_weights || !d_out) return fail(SART_ERR_INVALID_ARGUMENT, "sart_back_project_noalloc: NULL device pointer");
    if (nrows == 0 || ncols == 0 || ld < ncols || ld % 4 != 0 || (reinterpret_cast<uintptr_t>(d_matrix) & 15))
        return fail(SART_ERR_INVALID_ARGUMENT,
                    "sart_back_project_noalloc: need nrows, ncols > 0, ld >= ncols, ld a multiple of 4 floats and a 16-byte aligned matrix");
    if (ld > SART_NOALLOC_SCRATCH_FLOATS)
        return fail(SART_ERR_INVALID_ARGUMENT, "sart_back_project_noalloc: row length ld = " + std::to_string(ld) +
                                                   " floats exceeds the built-in scratch (at most " +
                                                   std::to_string(SART_NOALLOC_SCRATCH_FLOATS) +
                                                   "); use an engine (sart_create) for wider matrices");
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kNoallocMaxDevices) return fail(SART_ERR_INVALID_ARGUMENT, "sart_back_project_noalloc: device ordinal too large");
    float* scratch = nullptr;
    int sms = 0;
    {
        std::lock_guard<std::mutex> lock(g_noalloc_mutex);
        if (!g_noalloc_ptr[dev]) {  // first call on this device, which is the current one
            float* ptr = nullptr;
            CUDA_TRY(cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch));
            CUDA_TRY(cudaDeviceGetAttribute(&g_noalloc_sms[dev], cudaDevAttrMultiProcessorCount, dev));
            g_noalloc_ptr[dev] = ptr;
        }
        scratch = g_noalloc_ptr[dev];
        sms = g_noalloc_sms[dev];
    }

    // one wave of (1024-column panels) x (row slabs of a multiple of 32 rows), as many slabs as the array can hold
    const uint64_t ld4 = ld / 4;
    const uint64_t panels = ceil_div(ld4, kBpThreads);
    uint64_t slabs = std::max<uint64_t>(1, (uint64_t)sms * 4 / panels);
    slabs = std::min<uint64_t>(slabs, SART_NOALLOC_SCRATCH_FLOATS / ld);
    slabs = std::min<uint64_t>(slabs, ceil_div(nrows, 32));
    const uint64_t rows_per_slab = ceil_div(ceil_div(nrows, slabs), 32) * 32;
    slabs = ceil_div(nrows, rows_per_slab);
    if (panels > 0x7fffffffULL || slabs > 65535 || rows_per_slab > 0xffffffffULL)
        return fail(SART_ERR_INVALID_ARGUMENT, "sart_back_project_noalloc: problem too large for one launch");

    cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
    BpArgs a;
    a.H = d_matrix;
    a.ld4 = ld4;
    a.P = nrows;
    a.w = d_weights;
    a.part = nullptr;
    a.rows_per_slab = (uint32_t)rows_per_slab;
    a.st = nullptr;
    bp_cols_scratch_kernel<kBpThreads, 8><<<dim3((uint32_t)panels, (uint32_t)slabs), kBpThreads, 0, stream>>>(a, scratch);
    CUDA_TRY(cudaGetLastError());
    reduce_partials_kernel<<<(uint32_t)ceil_div(ncols, 256), 256, 0, stream>>>(scratch, (uint32_t)slabs, ld, ncols, d_out, nullptr);
    CUDA_TRY(cudaGetLastError());
    return SART_OK;
}
