// lwc_api.cu -- the C ABI (include/lwc_b200.h) over the sm_100a scoring kernels.
// Host side: context + streams, shape validation, kernel routing, and the chunked
// host-memory pipeline (H2D on a copy stream overlapped with kernels on the compute stream).
// There is no CPU fallback anywhere in this file: without a CUDA device every entry point
// that computes returns 503/500.
#include <cuda.h>
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>

#include "../../include/lwc_b200.h"
#include "score_generic.cuh"
#include "score_small.cuh"
#include "score_stream.cuh"
#include "score_tc.cuh"
#include "pool.cuh"
#include "vote_kernel.cuh"

#define LWC_VERSION_STRING "lwc_b200 0.1.0 (sm_100a)"

// Working memory of lwc_pool_consensus_static: part of the device image, so no allocation call is ever
// needed for it.  In host code this name is only a handle for cudaGetSymbolAddress -- it is NOT a
// device address and must never be handed to a kernel as one.
static __device__ __align__(256) unsigned char g_static_scratch[LWC_STATIC_SCRATCH_BYTES];

namespace {

std::mutex g_err_mu;
std::string g_create_err = "";

struct Slot {
    void *buf[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    size_t cap[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    cudaEvent_t h2d_done = nullptr; // inputs of this slot are on the device
    cudaEvent_t all_done = nullptr; // kernel + D2H of this slot finished (slot reusable)
    bool used = false;
};

} // namespace

struct lwc_ctx {
    int device = 0;
    int sm_count = 0;
    int chunk_requests = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;
    std::mutex mu;
    std::string err;
    uint64_t launches = 0;
    Slot slots[2];
    int small_kw = 4; // warps per request in the small-pool kernel (LWC_SMALL_KW)
    int small_kw_set = 0;
    int small_u = 0;  // k-chunks in flight per lane, 0 = per-shape default (LWC_SMALL_U)
    int small_impl = 0;   // 0 = persistent streaming kernel, 1 = one-CTA-per-request kernel (LWC_SMALL_IMPL=direct)
    int force_stream = 0; // LWC_SMALL_IMPL=stream: streaming kernel for every small shape
    int stream_ncw = 0;   // consumer warps of the streaming kernel, 0 = per-shape default (LWC_STREAM_NCW)
    int stream_kc = 0;    // upper bound on columns per stage, 0 = as large as fits (LWC_STREAM_KC)
    int stream_slots = 0; // ring depth override (LWC_STREAM_SLOTS)
    int stream_nf = 0;    // finaliser warps, 0 = 1 (LWC_STREAM_NF)
    int max_smem_optin = 0;
    ncclComm_t comm = nullptr; // attached by lwc_comm_init for the row-sharded pool
    int nranks = 1, rank = 0;
    // fused peer-to-peer gather of the pool path (lwc_pool_p2p_export / _attach)
    bool p2p_ready = false;
    long long p2p_n_own = 0;
    int p2p_D = 0;
    void *p2p_mapped[64][2] = {{nullptr}}; // cudaIpcOpenMemHandle results per peer: ehat_all, t_all
    void *p2p_dev = nullptr;               // device arrays: peer_ehat[nranks], peer_t[nranks], shard_done[nranks]
    void *pool_buf[12] = {nullptr};
    size_t pool_cap[12] = {0};
    int no_p2p = 0;    // LWC_POOL_NO_P2P=1: keep the NCCL all-gather even when peer mappings are attached
    int tc_ko = 0;     // operand slabs per stage of the tcgen05 kernel (LWC_TC_KO)
    int tc_stages = 0; // its ring depth (LWC_TC_STAGES)
};

namespace {

int fail(lwc_ctx *ctx, int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (ctx) {
        ctx->err = buf;
    } else {
        std::lock_guard<std::mutex> lk(g_err_mu);
        g_create_err = buf;
    }
    return code;
}

#define CUDA_TRY(ctx, expr)                                                                              \
    do {                                                                                                 \
        cudaError_t _e = (expr);                                                                         \
        if (_e != cudaSuccess)                                                                           \
            return fail((ctx), LWC_ERR_INTERNAL, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                        __FILE__, __LINE__);                                                             \
    } while (0)

enum Route { ROUTE_INVALID = 0, ROUTE_GENERIC, ROUTE_SMALL, ROUTE_TC };

struct Plan {
    Route route = ROUTE_INVALID;
    int np = 0;      // padded pool size
    bool x3 = true;  // small route: compensated 3xTF32
    const char *name = "";
};

Plan plan_for(const lwc_batch_desc &d, const void *E) {
    Plan pl;
    if (d.R < 0 || d.N < 1 || d.N > LWC_MAX_CHOICES || d.M < 0 || d.M > LWC_MAX_VOTERS) return pl;
    if (d.similarity != LWC_SIM_IDENTITY && d.similarity != LWC_SIM_COSINE) return pl;
    if (d.precision < LWC_PREC_FP32 || d.precision > LWC_PREC_BF16) return pl;
    if (d.memory != LWC_MEM_HOST && d.memory != LWC_MEM_DEVICE) return pl;
    if (d.similarity == LWC_SIM_COSINE && d.D < 1) return pl;
    const int np = d.N <= 32 ? 32 : (d.N <= 64 ? 64 : 128);
    if (d.similarity == LWC_SIM_IDENTITY) {
        pl.route = ROUTE_GENERIC;
        pl.np = np;
        pl.name = "tally";
        return pl;
    }
    const bool aligned = (d.D % 4 == 0) && ((reinterpret_cast<uintptr_t>(E) & 15u) == 0);
    if (d.N <= 32 && aligned) {
        pl.route = ROUTE_SMALL;
        pl.np = d.N <= 16 ? 16 : 32;
        pl.x3 = (d.precision == LWC_PREC_FP32 || d.precision == LWC_PREC_TF32X3);
        pl.name = pl.np == 16 ? (pl.x3 ? "small16_tf32x3" : "small16_tf32") : (pl.x3 ? "small32_tf32x3" : "small32_tf32");
        return pl;
    }
    // large pools: the similarity on tcgen05 tensor cores -- single-pass TF32 when the caller accepts
    // TF32-grade inputs, the compensated three-MMA scheme for fp32-grade results
    if (d.N > 32 && aligned && d.D % 32 == 0) {
        pl.route = ROUTE_TC;
        pl.np = 128;
        pl.x3 = (d.precision == LWC_PREC_FP32 || d.precision == LWC_PREC_TF32X3);
        pl.name = pl.x3 ? "tc128_tf32x3" : "tc128_tf32";
        return pl;
    }
    pl.route = ROUTE_GENERIC;
    pl.np = np;
    pl.name = np == 32 ? "generic32_fp32" : (np == 64 ? "generic64_fp32" : "generic128_fp32");
    return pl;
}

template <int NP>
cudaError_t launch_generic(const lwc::ScoreParams &p, cudaStream_t st) {
    using C = lwc::GenericCfg<NP>;
    static bool configured[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (!configured[dev & 63]) {
        cudaError_t e = cudaFuncSetAttribute(lwc::score_generic_kernel<NP>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)C::smem_bytes());
        if (e != cudaSuccess) return e;
        configured[dev & 63] = true;
    }
    lwc::score_generic_kernel<NP><<<p.R, lwc::kGenThreads, C::smem_bytes(), st>>>(p);
    return cudaGetLastError();
}

template <int NT, bool X3, int KW, int U>
cudaError_t launch_small_one(const lwc::ScoreParams &p, cudaStream_t st) {
    using C = lwc::SmallCfg<NT, X3, KW, U>;
    static bool configured[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (!configured[dev & 63]) {
        cudaError_t e = cudaFuncSetAttribute(lwc::score_small_kernel<NT, X3, KW, U>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::smem_bytes());
        if (e != cudaSuccess) return e;
        configured[dev & 63] = true;
    }
    lwc::score_small_kernel<NT, X3, KW, U><<<p.R, KW * 32, C::smem_bytes(), st>>>(p);
    return cudaGetLastError();
}

template <int NT, bool X3>
cudaError_t launch_small(const lwc::ScoreParams &p, cudaStream_t st, int kw, int u) {
    if (u == 0) u = (NT == 16) ? 4 : 2;
    const int key = kw * 10 + u;
    switch (key) {
        case 22: return launch_small_one<NT, X3, 2, 2>(p, st);
        case 24: return launch_small_one<NT, X3, 2, 4>(p, st);
        case 42: return launch_small_one<NT, X3, 4, 2>(p, st);
        case 44: return launch_small_one<NT, X3, 4, 4>(p, st);
        case 82: return launch_small_one<NT, X3, 8, 2>(p, st);
        case 84: return launch_small_one<NT, X3, 8, 4>(p, st);
        default: return launch_small_one<NT, X3, 4, 2>(p, st);
    }
}

// cuTensorMapEncodeTiled is a driver entry point; fetch it through the runtime so that the
// library does not link libcuda directly.
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void *sym = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(sym);
    }
    return fn;
}

// E viewed as a rank-4 fp32 tensor {32 (k_in), N (row), D/32 (k_out), R}: one box {32, 128, ko, 1} lands as
// ko K-major SWIZZLE_128B operand slabs of [128 rows x 128 B] (score_tc.cuh).
bool make_tensor_map_tc(CUtensorMap *tm, const float *E, int R, int N, int D, int ko) {
    EncodeTiledFn fn = encode_tiled_fn();
    if (!fn) return false;
    cuuint64_t dims[4] = {32, (cuuint64_t)N, (cuuint64_t)(D / 32), (cuuint64_t)R};
    cuuint64_t strides[3] = {(cuuint64_t)D * 4, 128, (cuuint64_t)N * D * 4};
    cuuint32_t box[4] = {32, 128, (cuuint32_t)ko, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    return fn(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float *>(E), dims, strides, box, estr,
              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template <bool X3>
cudaError_t launch_tc(lwc_ctx *ctx, const lwc::ScoreParams &p, cudaStream_t st) {
    lwc::TcGeom geo;
    geo.ko = ctx->tc_ko > 0 ? ctx->tc_ko : (X3 ? 2 : 4); // 64 KB per stage either way (profiles/r1_tma_probe2.txt)
    const int nslab = p.D / 32;
    if (geo.ko > nslab) geo.ko = nslab;
    geo.stages = lwc::kTcMaxStages;
    while (geo.stages > 1 && lwc::tc_smem_bytes(geo.ko, geo.stages, X3) > (size_t)ctx->max_smem_optin) --geo.stages;
    if (ctx->tc_stages > 0 && ctx->tc_stages < geo.stages) geo.stages = ctx->tc_stages;
    if (lwc::tc_smem_bytes(geo.ko, geo.stages, X3) > (size_t)ctx->max_smem_optin) return cudaErrorInvalidConfiguration;
    static bool configured[64] = {false};
    if (!configured[ctx->device & 63]) {
        cudaError_t e = cudaFuncSetAttribute(lwc::score_tc_kernel<X3>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             ctx->max_smem_optin);
        if (e != cudaSuccess) return e;
        configured[ctx->device & 63] = true;
    }
    CUtensorMap tm;
    if (!make_tensor_map_tc(&tm, p.E, p.R, p.N, p.D, geo.ko)) return cudaErrorInvalidValue;
    const int grid = p.R < ctx->sm_count ? p.R : ctx->sm_count;
    lwc::score_tc_kernel<X3><<<grid, X3 ? lwc::kTcThreadsX3 : lwc::kTcThreads, lwc::tc_smem_bytes(geo.ko, geo.stages, X3), st>>>(
        tm, p, geo);
    return cudaGetLastError();
}

template <int NT, bool X3, int NCW, int NF>
cudaError_t launch_stream_one(lwc_ctx *ctx, const lwc::ScoreParams &p, cudaStream_t st) {
    using C = lwc::StreamCfg<NT, X3, NCW, NF>;
    const size_t avail = (size_t)ctx->max_smem_optin - 128 - C::fixed_bytes();
    lwc::StreamGeom geo;
    // stage width: 3 KB per row copy is where the bulk-copy engine reaches HBM speed
    // (profiles/r1_tma_probe.txt); shrink it until at least 3 stages fit.
    const int dpad = (p.D + 15) / 16 * 16;
    int ks = ctx->stream_kc >= 16 ? ctx->stream_kc : 768;
    if (ks > dpad) ks = dpad;
    while (ks > 64 && avail / C::stage_bytes(ks) < 3) ks = ((ks / 2) + 15) / 16 * 16;
    geo.ks = ks;
    geo.stages = (int)(avail / C::stage_bytes(ks));
    if (geo.stages > C::MAXS) geo.stages = C::MAXS;
    if (ctx->stream_slots >= 1 && ctx->stream_slots < geo.stages) geo.stages = ctx->stream_slots;
    if (geo.stages < 1) return cudaErrorInvalidConfiguration;
    static bool configured[64] = {false};
    if (!configured[ctx->device & 63]) {
        cudaError_t e = cudaFuncSetAttribute(lwc::score_stream_kernel<NT, X3, NCW, NF>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, ctx->max_smem_optin);
        if (e != cudaSuccess) return e;
        configured[ctx->device & 63] = true;
    }
    const int grid = p.R < ctx->sm_count ? p.R : ctx->sm_count;
    lwc::score_stream_kernel<NT, X3, NCW, NF><<<grid, C::THREADS, C::smem_bytes(geo.ks, geo.stages), st>>>(p, geo);
    return cudaGetLastError();
}

template <int NT, bool X3>
cudaError_t launch_stream(lwc_ctx *ctx, const lwc::ScoreParams &p, cudaStream_t st) {
    const int ncw = ctx->stream_ncw ? ctx->stream_ncw : (NT == 16 ? 12 : 8);
    const int nf = ctx->stream_nf ? ctx->stream_nf : (NT == 16 ? 4 : 2);
    switch (ncw * 10 + nf) {
        case 41: return launch_stream_one<NT, X3, 4, 1>(ctx, p, st);
        case 42: return launch_stream_one<NT, X3, 4, 2>(ctx, p, st);
        case 44: return launch_stream_one<NT, X3, 4, 4>(ctx, p, st);
        case 81: return launch_stream_one<NT, X3, 8, 1>(ctx, p, st);
        case 82: return launch_stream_one<NT, X3, 8, 2>(ctx, p, st);
        case 83: return launch_stream_one<NT, X3, 8, 3>(ctx, p, st);
        case 84: return launch_stream_one<NT, X3, 8, 4>(ctx, p, st);
        case 86: return launch_stream_one<NT, X3, 8, 6>(ctx, p, st);
        case 122: return launch_stream_one<NT, X3, 12, 2>(ctx, p, st);
        case 124: return launch_stream_one<NT, X3, 12, 4>(ctx, p, st);
        default: return launch_stream_one<NT, X3, 8, 2>(ctx, p, st);
    }
}

int launch_plan(lwc_ctx *ctx, const Plan &pl, const lwc::ScoreParams &p, cudaStream_t st) {
    if (p.R == 0) return LWC_OK;
    cudaError_t e = cudaSuccess;
    // 32-candidate pools with the compensated 3xTF32 Gram are bound by the legacy mma.sync pipe, not by
    // HBM; there the one-CTA-per-request kernel (more warps in flight) measured 83 % vs 80 % of HBM peak.
    const bool stream_route = ctx->small_impl == 0 && !(pl.np == 32 && pl.x3 && !ctx->force_stream);
    if (pl.route == ROUTE_TC) {
        e = pl.x3 ? launch_tc<true>(ctx, p, st) : launch_tc<false>(ctx, p, st);
    } else if (pl.route == ROUTE_SMALL && stream_route) {
        if (pl.np == 16)
            e = pl.x3 ? launch_stream<16, true>(ctx, p, st) : launch_stream<16, false>(ctx, p, st);
        else
            e = pl.x3 ? launch_stream<32, true>(ctx, p, st) : launch_stream<32, false>(ctx, p, st);
    } else if (pl.route == ROUTE_SMALL) {
        if (pl.np == 16)
            e = pl.x3 ? launch_small<16, true>(p, st, ctx->small_kw, ctx->small_u)
                      : launch_small<16, false>(p, st, ctx->small_kw, ctx->small_u);
        else // 32-row pools: 2 warps per request measured best (1.157 vs 1.185 ms at KW=4, cfg3 slice)
            e = pl.x3 ? launch_small<32, true>(p, st, ctx->small_kw_set ? ctx->small_kw : 2, ctx->small_u)
                      : launch_small<32, false>(p, st, ctx->small_kw_set ? ctx->small_kw : 2, ctx->small_u);
    } else {
        if (pl.np == 32)
            e = launch_generic<32>(p, st);
        else if (pl.np == 64)
            e = launch_generic<64>(p, st);
        else
            e = launch_generic<128>(p, st);
    }
    if (e != cudaSuccess) return fail(ctx, LWC_ERR_INTERNAL, "kernel launch (%s) failed: %s", pl.name, cudaGetErrorString(e));
    ctx->launches++;
    return LWC_OK;
}

int ensure(lwc_ctx *ctx, Slot &s, int which, size_t bytes) {
    if (bytes <= s.cap[which]) return LWC_OK;
    if (s.buf[which]) CUDA_TRY(ctx, cudaFree(s.buf[which]));
    s.buf[which] = nullptr;
    s.cap[which] = 0;
    size_t want = bytes + bytes / 8 + 256;
    CUDA_TRY(ctx, cudaMalloc(&s.buf[which], want));
    s.cap[which] = want;
    return LWC_OK;
}

int score_host(lwc_ctx *ctx, const lwc_batch_desc &d, const Plan &pl, const float *E, const float *V, const float *w,
               float *conf, float *cw, float *agree, int32_t *winner, int32_t *rank) {
    const bool cosine = d.similarity == LWC_SIM_COSINE;
    const size_t e_req = cosine ? (size_t)d.N * d.D * sizeof(float) : 0;
    const size_t v_req = (size_t)d.M * d.N * sizeof(float);
    int chunk = ctx->chunk_requests;
    if (chunk <= 0) {
        // ~32 MiB of embeddings per chunk: large enough for PCIe efficiency, small enough to overlap
        size_t per = e_req + v_req + 64;
        chunk = (int)((32u << 20) / per);
        if (chunk < 1) chunk = 1;
    }
    if (chunk > d.R) chunk = d.R;
    // the shared weight vector (w_per_request == 0) is uploaded once into slot 0's w buffer
    const float *dw_shared = nullptr;
    if (!d.w_per_request && d.M > 0) {
        int rc = ensure(ctx, ctx->slots[0], 7, (size_t)d.M * sizeof(float));
        if (rc) return rc;
        CUDA_TRY(ctx, cudaMemcpyAsync(ctx->slots[0].buf[7], w, (size_t)d.M * sizeof(float), cudaMemcpyHostToDevice,
                                      ctx->stream));
        dw_shared = static_cast<const float *>(ctx->slots[0].buf[7]);
    }
    int ci = 0;
    for (int r0 = 0; r0 < d.R; r0 += chunk, ++ci) {
        const int rc_n = (d.R - r0 < chunk) ? d.R - r0 : chunk;
        Slot &s = ctx->slots[ci & 1];
        int rc;
        // a slot may be refilled only after its previous kernel and D2H finished
        if (s.used) CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->copy_stream, s.all_done, 0));
        // (re)allocation happens only while growing; cudaFree synchronises, which is safe here
        if ((rc = ensure(ctx, s, 0, e_req * rc_n))) return rc;
        if ((rc = ensure(ctx, s, 1, v_req * rc_n))) return rc;
        if ((rc = ensure(ctx, s, 2, (size_t)d.M * rc_n * sizeof(float)))) return rc;
        if ((rc = ensure(ctx, s, 3, (size_t)d.N * rc_n * sizeof(float) * 2))) return rc;             // conf | cw
        if ((rc = ensure(ctx, s, 4, (size_t)d.M * rc_n * sizeof(float)))) return rc;                 // agree
        if ((rc = ensure(ctx, s, 5, (size_t)rc_n * sizeof(int32_t) * ((size_t)d.N + 1)))) return rc; // winner | rank
        if (cosine)
            CUDA_TRY(ctx, cudaMemcpyAsync(s.buf[0], E + (size_t)r0 * d.N * d.D, e_req * rc_n, cudaMemcpyHostToDevice,
                                          ctx->copy_stream));
        if (v_req)
            CUDA_TRY(ctx, cudaMemcpyAsync(s.buf[1], V + (size_t)r0 * d.M * d.N, v_req * rc_n, cudaMemcpyHostToDevice,
                                          ctx->copy_stream));
        if (d.w_per_request && d.M > 0)
            CUDA_TRY(ctx, cudaMemcpyAsync(s.buf[2], w + (size_t)r0 * d.M, (size_t)d.M * rc_n * sizeof(float),
                                          cudaMemcpyHostToDevice, ctx->copy_stream));
        CUDA_TRY(ctx, cudaEventRecord(s.h2d_done, ctx->copy_stream));
        CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->stream, s.h2d_done, 0));

        lwc::ScoreParams p;
        p.E = static_cast<const float *>(s.buf[0]);
        p.V = static_cast<const float *>(s.buf[1]);
        p.w = d.w_per_request ? static_cast<const float *>(s.buf[2]) : dw_shared;
        float *dconf = static_cast<float *>(s.buf[3]);
        float *dcw = dconf + (size_t)d.N * rc_n;
        int32_t *dwin = static_cast<int32_t *>(s.buf[5]);
        int32_t *drank = dwin + rc_n;
        p.conf = conf ? dconf : nullptr;
        p.cw = cw ? dcw : nullptr;
        p.agree = agree ? static_cast<float *>(s.buf[4]) : nullptr;
        p.winner = winner ? dwin : nullptr;
        p.rank = rank ? drank : nullptr;
        p.R = rc_n, p.N = d.N, p.D = d.D, p.M = d.M;
        p.w_per_request = d.w_per_request;
        p.cosine = cosine ? 1 : 0;
        if ((rc = launch_plan(ctx, pl, p, ctx->stream))) return rc;

        if (conf)
            CUDA_TRY(ctx, cudaMemcpyAsync(conf + (size_t)r0 * d.N, dconf, (size_t)d.N * rc_n * sizeof(float),
                                          cudaMemcpyDeviceToHost, ctx->stream));
        if (cw)
            CUDA_TRY(ctx, cudaMemcpyAsync(cw + (size_t)r0 * d.N, dcw, (size_t)d.N * rc_n * sizeof(float),
                                          cudaMemcpyDeviceToHost, ctx->stream));
        if (agree && d.M > 0)
            CUDA_TRY(ctx, cudaMemcpyAsync(agree + (size_t)r0 * d.M, s.buf[4], (size_t)d.M * rc_n * sizeof(float),
                                          cudaMemcpyDeviceToHost, ctx->stream));
        if (winner)
            CUDA_TRY(ctx, cudaMemcpyAsync(winner + r0, dwin, (size_t)rc_n * sizeof(int32_t), cudaMemcpyDeviceToHost,
                                          ctx->stream));
        if (rank)
            CUDA_TRY(ctx, cudaMemcpyAsync(rank + (size_t)r0 * d.N, drank, (size_t)d.N * rc_n * sizeof(int32_t),
                                          cudaMemcpyDeviceToHost, ctx->stream));
        CUDA_TRY(ctx, cudaEventRecord(s.all_done, ctx->stream));
        s.used = true;
    }
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return LWC_OK;
}


// ---------------------------------------------------------------- NCCL, loaded on demand
struct NcclApi {
    void *handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    bool ok = false;
};

NcclApi &nccl_api() {
    static NcclApi api;
    static bool tried = false;
    if (!tried) {
        tried = true;
        for (const char *name : {"libnccl.so.2", "libnccl.so"}) {
            api.handle = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
            if (api.handle) break;
        }
        if (api.handle) {
            api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(dlsym(api.handle, "ncclGetUniqueId"));
            api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(dlsym(api.handle, "ncclCommInitRank"));
            api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(dlsym(api.handle, "ncclCommDestroy"));
            api.AllGather = reinterpret_cast<decltype(api.AllGather)>(dlsym(api.handle, "ncclAllGather"));
            api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(dlsym(api.handle, "ncclAllReduce"));
            api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(dlsym(api.handle, "ncclGetErrorString"));
            api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.AllGather && api.AllReduce && api.GetErrorString;
        }
    }
    return api;
}

#define NCCL_TRY(ctx, expr)                                                                                      \
    do {                                                                                                         \
        ncclResult_t _r = (expr);                                                                                \
        if (_r != ncclSuccess)                                                                                   \
            return fail((ctx), LWC_ERR_INTERNAL, "%s failed: %s (%s:%d)", #expr, nccl_api().GetErrorString(_r), \
                        __FILE__, __LINE__);                                                                     \
    } while (0)

int pool_ensure(lwc_ctx *ctx, int which, size_t bytes) {
    if (bytes <= ctx->pool_cap[which]) return LWC_OK;
    if (ctx->pool_buf[which]) CUDA_TRY(ctx, cudaFree(ctx->pool_buf[which]));
    ctx->pool_buf[which] = nullptr;
    ctx->pool_cap[which] = 0;
    CUDA_TRY(ctx, cudaMalloc(&ctx->pool_buf[which], bytes + 256));
    ctx->pool_cap[which] = bytes + 256;
    return LWC_OK;
}

bool make_tensor_map_bf16_2d(CUtensorMap *tm, const void *base, long long rows, int D, int box_rows) {
    EncodeTiledFn fn = encode_tiled_fn();
    if (!fn) return false;
    cuuint64_t dims[2] = {(cuuint64_t)D, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)D * 2};
    cuuint32_t box[2] = {(cuuint32_t)lwc::kPoolKS, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    return fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(base), dims, strides, box, estr,
              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// Device-side core of the pool path.  Everything it touches beyond inputs/outputs is passed in `sc`.
struct PoolScratch {
    __nv_bfloat16 *ehat_all; // [n_total][D]
    float *t_all;            // [n_total]
    float *ypart;            // [nch][n_own]
    float *cw;               // [n_own]
    float *conf;             // [n_own]
    float *blk_sum, *blk_max; // [nblk] (+ [nranks] when sharded)
    long long *blk_arg;
    float *blk_agree;        // [nblk][M]
    float *scal;             // {sum, max} + gathered copies
    long long *arg;          // {argmax} + gathered copies
    float *agree;            // [M]
};

int pool_run(lwc_ctx *ctx, const PoolScratch &sc, const float *dE, const float *dV, const float *dw, long long n_own, int D,
             int M, cudaStream_t st) {
    const int nranks = ctx->comm ? ctx->nranks : 1, rank = ctx->comm ? ctx->rank : 0;
    const long long n_total = n_own * nranks, row0 = n_own * rank;
    NcclApi &nc = nccl_api();
    // 1. normalise own rows (bf16) + tally vector
    {
        const int warps = 8;
        const long long blocks = (n_own + warps - 1) / warps;
        lwc::pool_prep_kernel<<<(unsigned)blocks, warps * 32, 0, st>>>(dE, dV, dw, sc.ehat_all + (size_t)row0 * D,
                                                                     sc.t_all + row0, n_own, D, M);
        CUDA_TRY(ctx, cudaGetLastError());
        ctx->launches++;
    }
    // 2. the exchange step.  Fused route: the tensor pass itself pulls the peers' shards over NVLink, so
    //    only a stream-ordered cross-rank barrier is needed here (every rank's prep has finished).
    //    Otherwise: NCCL all-gather of normalised rows and tally vector (in place).
    const bool fused = nranks > 1 && ctx->p2p_ready && ctx->p2p_n_own == n_own && ctx->p2p_D == D &&
                       n_own % lwc::kPoolTN == 0 && !ctx->no_p2p;
    if (fused) {
        NCCL_TRY(ctx, nc.AllReduce(sc.scal + 4, sc.scal + 4, 1, ncclFloat32, ncclSum, ctx->comm, st));
        CUDA_TRY(ctx, cudaMemsetAsync(static_cast<char *>(ctx->p2p_dev) + 2 * 64 * sizeof(void *), 0, 64 * sizeof(unsigned), st));
    } else if (nranks > 1) {
        NCCL_TRY(ctx, nc.AllGather(sc.ehat_all + (size_t)row0 * D, sc.ehat_all, (size_t)n_own * D * 2, ncclUint8, ctx->comm, st));
        NCCL_TRY(ctx, nc.AllGather(sc.t_all + row0, sc.t_all, (size_t)n_own, ncclFloat32, ctx->comm, st));
    }
    // 3. all-pairs pass on tensor cores
    const int nI = (int)((n_own + lwc::kPoolTM - 1) / lwc::kPoolTM);
    const int njb = (int)((n_total + lwc::kPoolTN - 1) / lwc::kPoolTN);
    // column shards (fused gather: one per rank) x chunks per shard; at most 64 chunks in all
    const int nshards = fused ? nranks : 1;
    const int jbs = fused ? (int)(n_own / lwc::kPoolTN) : njb; // 256-column blocks per shard
    int cps = ((8 * ctx->sm_count + nI - 1) / nI + nshards - 1) / nshards;
    if (cps > jbs) cps = jbs;
    if (cps * nshards > 64) cps = 64 / nshards;
    if (cps < 1) cps = 1;
    const int jbpc = (jbs + cps - 1) / cps;
    cps = (jbs + jbpc - 1) / jbpc;
    const int nch = cps * nshards;
    {
        CUtensorMap tmA, tmB;
        if (!make_tensor_map_bf16_2d(&tmA, sc.ehat_all, n_total, D, lwc::kPoolTM) ||
            !make_tensor_map_bf16_2d(&tmB, sc.ehat_all, n_total, D, lwc::kPoolTN))
            return fail(ctx, LWC_ERR_INTERNAL, "cuTensorMapEncodeTiled failed for the pool operand (D=%d)", D);
        static bool configured[64] = {false};
        if (!configured[ctx->device & 63]) {
            CUDA_TRY(ctx, cudaFuncSetAttribute(lwc::pool_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                               (int)lwc::pool_gemm_smem_bytes()));
            configured[ctx->device & 63] = true;
        }
        lwc::PoolGemmParams gp;
        gp.t_all = sc.t_all, gp.ypart = sc.ypart;
        gp.n_own = n_own, gp.n_total = n_total, gp.row0 = row0;
        gp.D = D, gp.nch = nch, gp.jb_per_chunk = jbpc;
        gp.nshards = nshards, gp.rank = fused ? rank : 0, gp.cps = cps, gp.jb_per_shard = jbs;
        gp.peer_ehat = nullptr, gp.peer_t = nullptr, gp.ehat_all = sc.ehat_all, gp.t_all_w = sc.t_all, gp.shard_done = nullptr;
        if (fused) {
            char *pd = static_cast<char *>(ctx->p2p_dev);
            gp.peer_ehat = reinterpret_cast<const __nv_bfloat16 *const *>(pd);
            gp.peer_t = reinterpret_cast<const float *const *>(pd + 64 * sizeof(void *));
            gp.shard_done = reinterpret_cast<unsigned int *>(pd + 2 * 64 * sizeof(void *));
        }
        const int nunits = nI * nch;
        const int grid = nunits < ctx->sm_count ? nunits : ctx->sm_count;
        lwc::pool_gemm_kernel<<<grid, lwc::kPoolThreads, lwc::pool_gemm_smem_bytes(), st>>>(tmA, tmB, gp);
        CUDA_TRY(ctx, cudaGetLastError());
        ctx->launches++;
    }
    // 4. cw, sum, (max, argmax): per block, then folded in a fixed order
    const int nblk = (int)((n_own + lwc::kPoolRedPerBlock - 1) / lwc::kPoolRedPerBlock);
    lwc::pool_reduce_kernel<<<nblk, lwc::kPoolRedThreads, 0, st>>>(sc.ypart, nch, n_own, row0, sc.cw, sc.blk_sum, sc.blk_max,
                                                                  sc.blk_arg);
    lwc::pool_fold_kernel<<<1, 32, 0, st>>>(sc.blk_sum, sc.blk_max, sc.blk_arg, nblk, sc.scal, sc.scal + 1, sc.arg);
    CUDA_TRY(ctx, cudaGetLastError());
    ctx->launches += 2;
    if (nranks > 1) { // three scalars per rank, folded identically on every rank
        float *gs = sc.scal + 8, *gm = gs + nranks;
        long long *ga = sc.arg + 8;
        NCCL_TRY(ctx, nc.AllGather(sc.scal, gs, 1, ncclFloat32, ctx->comm, st));
        NCCL_TRY(ctx, nc.AllGather(sc.scal + 1, gm, 1, ncclFloat32, ctx->comm, st));
        NCCL_TRY(ctx, nc.AllGather(sc.arg, ga, 1, ncclInt64, ctx->comm, st));
        lwc::pool_fold_kernel<<<1, 32, 0, st>>>(gs, gm, ga, nranks, sc.scal, sc.scal + 1, sc.arg);
        CUDA_TRY(ctx, cudaGetLastError());
        ctx->launches++;
    }
    // 5. confidences and voter agreement
    lwc::pool_finish_kernel<<<nblk, lwc::kPoolRedThreads, 0, st>>>(sc.cw, sc.scal, dV, M, n_own, sc.conf, sc.blk_agree);
    ctx->launches++;
    if (M > 0) {
        lwc::pool_fold_agree_kernel<<<(M + 127) / 128, 128, 0, st>>>(sc.blk_agree, nblk, M, dV, n_own, sc.agree);
        ctx->launches++;
        if (nranks > 1) NCCL_TRY(ctx, nc.AllReduce(sc.agree, sc.agree, (size_t)M, ncclFloat32, ncclSum, ctx->comm, st));
    }
    CUDA_TRY(ctx, cudaGetLastError());
    return LWC_OK;
}

} // namespace

extern "C" {

const char *lwc_version(void) { return LWC_VERSION_STRING; }

const char *lwc_last_error_message(const lwc_ctx *ctx) {
    if (ctx) return ctx->err.c_str();
    std::lock_guard<std::mutex> lk(g_err_mu);
    static thread_local std::string copy;
    copy = g_create_err;
    return copy.c_str();
}

int32_t lwc_ctx_create(const lwc_config *cfg, lwc_ctx **out) {
    if (!out) return fail(nullptr, LWC_ERR_INVALID, "lwc_ctx_create: out is NULL");
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(nullptr, LWC_ERR_UNAVAILABLE, "no CUDA device: %s (this library has no CPU path)",
                    e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    const int dev = cfg ? cfg->device : 0;
    if (dev < 0 || dev >= ndev) return fail(nullptr, LWC_ERR_INVALID, "device %d out of range (0..%d)", dev, ndev - 1);
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, dev)) != cudaSuccess)
        return fail(nullptr, LWC_ERR_INTERNAL, "cudaGetDeviceProperties: %s", cudaGetErrorString(e));
    if (prop.major != 10)
        return fail(nullptr, LWC_ERR_UNAVAILABLE, "device %d is sm_%d%d; this library carries sm_100a code only", dev,
                    prop.major, prop.minor);
    lwc_ctx *ctx = new lwc_ctx();
    ctx->device = dev;
    ctx->sm_count = prop.multiProcessorCount;
    ctx->chunk_requests = cfg ? cfg->chunk_requests : 0;
    if (const char *s = getenv("LWC_SMALL_KW")) {
        ctx->small_kw = atoi(s);
        ctx->small_kw_set = 1;
    }
    if (const char *s = getenv("LWC_SMALL_U")) ctx->small_u = atoi(s);
    if (const char *s = getenv("LWC_SMALL_IMPL")) {
        ctx->small_impl = (strcmp(s, "direct") == 0) ? 1 : 0;
        ctx->force_stream = (strcmp(s, "stream") == 0) ? 1 : 0;
    }
    if (const char *s = getenv("LWC_STREAM_NCW")) ctx->stream_ncw = atoi(s);
    if (const char *s = getenv("LWC_STREAM_KC")) ctx->stream_kc = atoi(s);
    if (const char *s = getenv("LWC_STREAM_SLOTS")) ctx->stream_slots = atoi(s);
    if (const char *s = getenv("LWC_STREAM_NF")) ctx->stream_nf = atoi(s);
    if (const char *s = getenv("LWC_POOL_NO_P2P")) ctx->no_p2p = atoi(s);
    if (const char *s = getenv("LWC_TC_KO")) ctx->tc_ko = atoi(s);
    if (const char *s = getenv("LWC_TC_STAGES")) ctx->tc_stages = atoi(s);
    if (ctx->stream_ncw != 0 && ctx->stream_ncw != 4 && ctx->stream_ncw != 8 && ctx->stream_ncw != 12) ctx->stream_ncw = 0;
    if (ctx->stream_kc < 0 || ctx->stream_kc % 16) ctx->stream_kc = 0;
    cudaDeviceGetAttribute(&ctx->max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    if (ctx->small_kw != 2 && ctx->small_kw != 4 && ctx->small_kw != 8) ctx->small_kw = 4;
    if (ctx->small_u != 0 && ctx->small_u != 2 && ctx->small_u != 4) ctx->small_u = 0;
    bool ok = cudaSetDevice(dev) == cudaSuccess &&
              cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) == cudaSuccess &&
              cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking) == cudaSuccess;
    for (int i = 0; ok && i < 2; ++i)
        ok = cudaEventCreateWithFlags(&ctx->slots[i].h2d_done, cudaEventDisableTiming) == cudaSuccess &&
             cudaEventCreateWithFlags(&ctx->slots[i].all_done, cudaEventDisableTiming) == cudaSuccess;
    if (!ok) {
        int rc = fail(nullptr, LWC_ERR_INTERNAL, "context setup failed: %s", cudaGetErrorString(cudaGetLastError()));
        delete ctx;
        return rc;
    }
    *out = ctx;
    return LWC_OK;
}

void lwc_ctx_destroy(lwc_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    if (ctx->copy_stream) cudaStreamSynchronize(ctx->copy_stream);
    for (auto &m : ctx->p2p_mapped)
        for (auto &q : m)
            if (q) cudaIpcCloseMemHandle(q);
    if (ctx->p2p_dev) cudaFree(ctx->p2p_dev);
    for (auto &b : ctx->pool_buf)
        if (b) cudaFree(b);
    for (auto &s : ctx->slots) {
        for (auto &b : s.buf)
            if (b) cudaFree(b);
        if (s.h2d_done) cudaEventDestroy(s.h2d_done);
        if (s.all_done) cudaEventDestroy(s.all_done);
    }
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    delete ctx;
}

int32_t lwc_host_alloc(lwc_ctx *ctx, size_t bytes, void **out) {
    if (!ctx || !out) return fail(ctx, LWC_ERR_INVALID, "lwc_host_alloc: NULL argument");
    std::lock_guard<std::mutex> lk(ctx->mu);
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    CUDA_TRY(ctx, cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault));
    return LWC_OK;
}

int32_t lwc_host_free(lwc_ctx *ctx, void *p) {
    if (!ctx) return LWC_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (p) CUDA_TRY(ctx, cudaFreeHost(p));
    return LWC_OK;
}

int32_t lwc_device_alloc(lwc_ctx *ctx, size_t bytes, void **out) {
    if (!ctx || !out) return fail(ctx, LWC_ERR_INVALID, "lwc_device_alloc: NULL argument");
    std::lock_guard<std::mutex> lk(ctx->mu);
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    CUDA_TRY(ctx, cudaMalloc(out, bytes ? bytes : 1));
    return LWC_OK;
}

int32_t lwc_device_free(lwc_ctx *ctx, void *p) {
    if (!ctx) return LWC_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    if (p) CUDA_TRY(ctx, cudaFree(p));
    return LWC_OK;
}

int32_t lwc_memcpy(lwc_ctx *ctx, void *dst, const void *src, size_t bytes, int32_t kind) {
    if (!ctx || (bytes && (!dst || !src))) return fail(ctx, LWC_ERR_INVALID, "lwc_memcpy: NULL argument");
    if (kind != 0 && kind != 1) return fail(ctx, LWC_ERR_INVALID, "lwc_memcpy: kind must be 0 (H2D) or 1 (D2H)");
    std::lock_guard<std::mutex> lk(ctx->mu);
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    CUDA_TRY(ctx, cudaMemcpyAsync(dst, src, bytes, kind == 0 ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost,
                                  ctx->stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return LWC_OK;
}

void *lwc_ctx_stream(lwc_ctx *ctx) { return ctx ? static_cast<void *>(ctx->stream) : nullptr; }

int32_t lwc_stream_synchronize(lwc_ctx *ctx, void *stream) {
    if (!ctx) return LWC_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    CUDA_TRY(ctx, cudaStreamSynchronize(stream ? static_cast<cudaStream_t>(stream) : ctx->stream));
    return LWC_OK;
}

const char *lwc_kernel_for(const lwc_ctx *ctx, const lwc_batch_desc *desc) {
    (void)ctx;
    if (!desc) return "";
    // alignment is unknown here: assume the documented 16-byte alignment of E
    return plan_for(*desc, nullptr).name;
}

uint64_t lwc_launch_count(const lwc_ctx *ctx) { return ctx ? ctx->launches : 0; }

int32_t lwc_score_batch(lwc_ctx *ctx, const lwc_batch_desc *desc, const float *E, const float *V, const float *w,
                        float *conf, float *cw, float *agree, int32_t *winner, int32_t *rank, void *stream) {
    if (!ctx) return LWC_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (!desc) return fail(ctx, LWC_ERR_INVALID, "lwc_score_batch: desc is NULL");
    const lwc_batch_desc d = *desc;
    const Plan pl = plan_for(d, E);
    if (pl.route == ROUTE_INVALID)
        return fail(ctx, LWC_ERR_INVALID,
                    "lwc_score_batch: invalid descriptor (R=%d N=%d D=%d M=%d sim=%d prec=%d mem=%d); need R>=0, "
                    "1<=N<=%d, 0<=M<=%d, D>=1 for cosine",
                    d.R, d.N, d.D, d.M, d.similarity, d.precision, d.memory, LWC_MAX_CHOICES, LWC_MAX_VOTERS);
    if (d.R == 0) return LWC_OK;
    if (d.similarity == LWC_SIM_COSINE && !E) return fail(ctx, LWC_ERR_INVALID, "lwc_score_batch: E is NULL");
    if (d.M > 0 && (!V || !w)) return fail(ctx, LWC_ERR_INVALID, "lwc_score_batch: V or w is NULL with M > 0");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    if (d.memory == LWC_MEM_HOST) return score_host(ctx, d, pl, E, V, w, conf, cw, agree, winner, rank);

    lwc::ScoreParams p;
    p.E = E, p.V = V, p.w = w;
    p.conf = conf, p.cw = cw, p.agree = agree, p.winner = winner, p.rank = rank;
    p.R = d.R, p.N = d.N, p.D = d.D, p.M = d.M;
    p.w_per_request = d.w_per_request;
    p.cosine = d.similarity == LWC_SIM_COSINE ? 1 : 0;
    return launch_plan(ctx, pl, p, stream ? static_cast<cudaStream_t>(stream) : ctx->stream);
}

int32_t lwc_merge_error_codes(const uint16_t *codes, int32_t n_voters) {
    // client.rs:386-409: only while every voter so far has errored does the merge continue.
    bool have = false;
    uint16_t merged = 0;
    for (int32_t m = 0; m < n_voters; ++m) {
        const uint16_t e = codes[m];
        if (e == 0) return 0; // a voter without an error: not an all-failed request
        if (!have) {
            merged = e;
            have = true;
        } else if (e != merged) {
            const bool both_4xx = e >= 400 && e < 500 && merged >= 400 && merged < 500;
            merged = both_4xx ? 400 : 500;
        }
    }
    return have ? merged : 500; // AllVotesFailed(None) -> 500 (score/completions/error.rs:41)
}

int32_t lwc_vote_from_logprobs(lwc_ctx *ctx, int32_t n_rows, int32_t n_top, int32_t n_choices, const float *logprob,
                               const int32_t *leaf, float *V, int32_t memory, void *stream) {
    if (!ctx) return LWC_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (n_rows < 0 || n_top < 0 || n_choices < 1)
        return fail(ctx, LWC_ERR_INVALID, "lwc_vote_from_logprobs: bad shape (rows=%d top=%d choices=%d)", n_rows, n_top,
                    n_choices);
    if (n_rows == 0) return LWC_OK;
    if (!V || (n_top > 0 && (!logprob || !leaf))) return fail(ctx, LWC_ERR_INVALID, "lwc_vote_from_logprobs: NULL argument");
    if (memory != LWC_MEM_HOST && memory != LWC_MEM_DEVICE)
        return fail(ctx, LWC_ERR_INVALID, "lwc_vote_from_logprobs: memory must be LWC_MEM_HOST or LWC_MEM_DEVICE");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = (memory == LWC_MEM_DEVICE && stream) ? static_cast<cudaStream_t>(stream) : ctx->stream;
    const float *dlp = logprob;
    const int32_t *dlf = leaf;
    float *dV = V;
    const size_t in_bytes = (size_t)n_rows * n_top * 4, out_bytes = (size_t)n_rows * n_choices * 4;
    if (memory == LWC_MEM_HOST) {
        Slot &s = ctx->slots[0];
        int rc;
        if ((rc = ensure(ctx, s, 0, in_bytes + 4))) return rc;
        if ((rc = ensure(ctx, s, 1, in_bytes + 4))) return rc;
        if ((rc = ensure(ctx, s, 3, out_bytes))) return rc;
        if (in_bytes) {
            CUDA_TRY(ctx, cudaMemcpyAsync(s.buf[0], logprob, in_bytes, cudaMemcpyHostToDevice, st));
            CUDA_TRY(ctx, cudaMemcpyAsync(s.buf[1], leaf, in_bytes, cudaMemcpyHostToDevice, st));
        }
        dlp = static_cast<const float *>(s.buf[0]);
        dlf = static_cast<const int32_t *>(s.buf[1]);
        dV = static_cast<float *>(s.buf[3]);
    }
    const int warps = 4;
    lwc::vote_from_logprobs_kernel<<<(n_rows + warps - 1) / warps, warps * 32, 0, st>>>(dlp, dlf, dV, n_rows, n_top,
                                                                                     n_choices);
    CUDA_TRY(ctx, cudaGetLastError());
    ctx->launches++;
    if (memory == LWC_MEM_HOST) {
        CUDA_TRY(ctx, cudaMemcpyAsync(V, dV, out_bytes, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(ctx, cudaStreamSynchronize(st));
    }
    return LWC_OK;
}

int32_t lwc_comm_unique_id(void *id128) {
    if (!id128) return LWC_ERR_INVALID;
    NcclApi &nc = nccl_api();
    if (!nc.ok) return fail(nullptr, LWC_ERR_UNAVAILABLE, "NCCL (libnccl.so.2) could not be loaded");
    ncclUniqueId id;
    if (nc.GetUniqueId(&id) != ncclSuccess) return fail(nullptr, LWC_ERR_INTERNAL, "ncclGetUniqueId failed");
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    memcpy(id128, &id, 128);
    return LWC_OK;
}

int32_t lwc_comm_init(lwc_ctx *ctx, int32_t nranks, int32_t rank, const void *id128) {
    if (!ctx || !id128) return LWC_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (nranks < 1 || rank < 0 || rank >= nranks) return fail(ctx, LWC_ERR_INVALID, "lwc_comm_init: bad rank %d of %d", rank, nranks);
    NcclApi &nc = nccl_api();
    if (!nc.ok) return fail(ctx, LWC_ERR_UNAVAILABLE, "NCCL (libnccl.so.2) could not be loaded");
    if (ctx->comm) return fail(ctx, LWC_ERR_INVALID, "lwc_comm_init: a communicator is already attached");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    ncclUniqueId id;
    memcpy(&id, id128, 128);
    NCCL_TRY(ctx, nc.CommInitRank(&ctx->comm, nranks, id, rank));
    ctx->nranks = nranks;
    ctx->rank = rank;
    return LWC_OK;
}

int32_t lwc_comm_destroy(lwc_ctx *ctx) {
    if (!ctx) return LWC_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (ctx->comm) {
        cudaSetDevice(ctx->device);
        cudaStreamSynchronize(ctx->stream);
        nccl_api().CommDestroy(ctx->comm);
        ctx->comm = nullptr;
        ctx->nranks = 1;
        ctx->rank = 0;
    }
    return LWC_OK;
}

int32_t lwc_pool_p2p_export(lwc_ctx *ctx, int64_t n_own, int32_t D, void *handles_out) {
    if (!ctx || !handles_out) return LWC_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (!ctx->comm) return fail(ctx, LWC_ERR_INVALID, "lwc_pool_p2p_export: attach a communicator first (lwc_comm_init)");
    if (n_own < 1 || D < 8 || D % 8 != 0 || n_own * ctx->nranks > 0x7fffff00LL)
        return fail(ctx, LWC_ERR_INVALID, "lwc_pool_p2p_export: invalid shape (N=%lld D=%d)", (long long)n_own, D);
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    const long long n_total = n_own * ctx->nranks;
    int rc;
    // the buffers peers will read from must exist (and stay put) before their handles are handed out
    if ((rc = pool_ensure(ctx, 0, (size_t)n_total * D * 2))) return rc;
    if ((rc = pool_ensure(ctx, 1, (size_t)n_total * 4))) return rc;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
    cudaIpcMemHandle_t h[2];
    CUDA_TRY(ctx, cudaIpcGetMemHandle(&h[0], ctx->pool_buf[0]));
    CUDA_TRY(ctx, cudaIpcGetMemHandle(&h[1], ctx->pool_buf[1]));
    memcpy(handles_out, h, 128);
    ctx->p2p_ready = false;
    ctx->p2p_n_own = n_own;
    ctx->p2p_D = D;
    return LWC_OK;
}

int32_t lwc_pool_p2p_attach(lwc_ctx *ctx, const void *all_handles) {
    if (!ctx || !all_handles) return LWC_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (!ctx->comm || ctx->p2p_n_own == 0)
        return fail(ctx, LWC_ERR_INVALID, "lwc_pool_p2p_attach: call lwc_comm_init and lwc_pool_p2p_export first");
    if (ctx->nranks > 64) return fail(ctx, LWC_ERR_INVALID, "lwc_pool_p2p_attach: at most 64 ranks");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    const char *hp = static_cast<const char *>(all_handles);
    void *host_ptrs[2 * 64] = {nullptr};
    for (int r = 0; r < ctx->nranks; ++r) {
        for (int k = 0; k < 2; ++k) {
            void *ptr = ctx->pool_buf[k];
            if (r != ctx->rank) {
                if (ctx->p2p_mapped[r][k]) {
                    cudaIpcCloseMemHandle(ctx->p2p_mapped[r][k]);
                    ctx->p2p_mapped[r][k] = nullptr;
                }
                cudaIpcMemHandle_t h;
                memcpy(&h, hp + (size_t)r * 128 + k * 64, 64);
                CUDA_TRY(ctx, cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
                ctx->p2p_mapped[r][k] = ptr;
            }
            host_ptrs[k * 64 + r] = ptr;
        }
    }
    if (!ctx->p2p_dev) CUDA_TRY(ctx, cudaMalloc(&ctx->p2p_dev, 2 * 64 * sizeof(void *) + 64 * sizeof(unsigned)));
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->p2p_dev, host_ptrs, sizeof host_ptrs, cudaMemcpyHostToDevice, ctx->stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->p2p_ready = true;
    return LWC_OK;
}

int32_t lwc_pool_consensus(lwc_ctx *ctx, const lwc_pool_desc *desc, const float *E, const float *V, const float *w,
                           float *conf, float *cw, float *agree, int64_t *winner, void *stream) {
    if (!ctx) return LWC_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (!desc) return fail(ctx, LWC_ERR_INVALID, "lwc_pool_consensus: desc is NULL");
    const lwc_pool_desc d = *desc;
    const int nranks = ctx->comm ? ctx->nranks : 1;
    if (d.N < 1 || d.D < 8 || d.D % 8 != 0 || d.M < 0 || d.M > LWC_MAX_VOTERS || d.N * nranks > 0x7fffff00LL)
        return fail(ctx, LWC_ERR_INVALID, "lwc_pool_consensus: invalid shape (N=%lld D=%d M=%d); need N>=1, D%%8==0, 0<=M<=%d",
                    (long long)d.N, d.D, d.M, LWC_MAX_VOTERS);
    if (d.precision != LWC_PREC_BF16)
        return fail(ctx, LWC_ERR_INVALID, "lwc_pool_consensus: only LWC_PREC_BF16 is implemented for the pool path");
    if (d.memory != LWC_MEM_HOST && d.memory != LWC_MEM_DEVICE) return fail(ctx, LWC_ERR_INVALID, "lwc_pool_consensus: bad memory kind");
    if (!E || (d.M > 0 && (!V || !w))) return fail(ctx, LWC_ERR_INVALID, "lwc_pool_consensus: NULL input");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = (d.memory == LWC_MEM_DEVICE && stream) ? static_cast<cudaStream_t>(stream) : ctx->stream;
    const long long n_own = d.N, n_total = d.N * nranks;
    const int nblk = (int)((n_own + lwc::kPoolRedPerBlock - 1) / lwc::kPoolRedPerBlock);
    int rc;
    // scratch: 0 ehat_all, 1 t_all, 2 ypart, 3 cw, 4 conf, 5 blk_sum|blk_max, 6 blk_arg, 7 blk_agree, 8 scal, 9 arg, 10 agree, 11 staged inputs
    if ((rc = pool_ensure(ctx, 0, (size_t)n_total * d.D * 2))) return rc;
    if ((rc = pool_ensure(ctx, 1, (size_t)n_total * 4))) return rc;
    if ((rc = pool_ensure(ctx, 2, (size_t)64 * n_own * 4))) return rc;
    if ((rc = pool_ensure(ctx, 3, (size_t)n_own * 4))) return rc;
    if ((rc = pool_ensure(ctx, 4, (size_t)n_own * 4))) return rc;
    if ((rc = pool_ensure(ctx, 5, (size_t)2 * (nblk + 16) * 4))) return rc;
    if ((rc = pool_ensure(ctx, 6, (size_t)(nblk + 16) * 8))) return rc;
    if ((rc = pool_ensure(ctx, 7, (size_t)nblk * (d.M > 0 ? d.M : 1) * 4))) return rc;
    if ((rc = pool_ensure(ctx, 8, (size_t)(8 + 2 * nranks) * 4))) return rc;
    if ((rc = pool_ensure(ctx, 9, (size_t)(8 + nranks) * 8))) return rc;
    if ((rc = pool_ensure(ctx, 10, (size_t)(d.M > 0 ? d.M : 1) * 4))) return rc;
    PoolScratch sc;
    sc.ehat_all = static_cast<__nv_bfloat16 *>(ctx->pool_buf[0]);
    sc.t_all = static_cast<float *>(ctx->pool_buf[1]);
    sc.ypart = static_cast<float *>(ctx->pool_buf[2]);
    sc.cw = static_cast<float *>(ctx->pool_buf[3]);
    sc.conf = static_cast<float *>(ctx->pool_buf[4]);
    sc.blk_sum = static_cast<float *>(ctx->pool_buf[5]);
    sc.blk_max = sc.blk_sum + nblk + 16;
    sc.blk_arg = static_cast<long long *>(ctx->pool_buf[6]);
    sc.blk_agree = static_cast<float *>(ctx->pool_buf[7]);
    sc.scal = static_cast<float *>(ctx->pool_buf[8]);
    sc.arg = static_cast<long long *>(ctx->pool_buf[9]);
    sc.agree = static_cast<float *>(ctx->pool_buf[10]);
    const float *dE = E, *dV = V, *dw = w;
    const size_t e_bytes = (size_t)n_own * d.D * 4, v_bytes = (size_t)d.M * n_own * 4, w_bytes = (size_t)d.M * 4;
    if (d.memory == LWC_MEM_HOST) {
        const size_t v_off = (e_bytes + 255) & ~(size_t)255, w_off = (v_off + v_bytes + 255) & ~(size_t)255;
        if ((rc = pool_ensure(ctx, 11, w_off + w_bytes + 256))) return rc;
        char *base = static_cast<char *>(ctx->pool_buf[11]);
        CUDA_TRY(ctx, cudaMemcpyAsync(base, E, e_bytes, cudaMemcpyHostToDevice, st));
        if (v_bytes) CUDA_TRY(ctx, cudaMemcpyAsync(base + v_off, V, v_bytes, cudaMemcpyHostToDevice, st));
        if (w_bytes) CUDA_TRY(ctx, cudaMemcpyAsync(base + w_off, w, w_bytes, cudaMemcpyHostToDevice, st));
        dE = reinterpret_cast<const float *>(base);
        dV = reinterpret_cast<const float *>(base + v_off);
        dw = reinterpret_cast<const float *>(base + w_off);
    }
    if ((rc = pool_run(ctx, sc, dE, dV, dw, n_own, d.D, d.M, st))) return rc;
    const cudaMemcpyKind kind = d.memory == LWC_MEM_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice;
    if (conf) CUDA_TRY(ctx, cudaMemcpyAsync(conf, sc.conf, (size_t)n_own * 4, kind, st));
    if (cw) CUDA_TRY(ctx, cudaMemcpyAsync(cw, sc.cw, (size_t)n_own * 4, kind, st));
    if (agree && d.M > 0) CUDA_TRY(ctx, cudaMemcpyAsync(agree, sc.agree, (size_t)d.M * 4, kind, st));
    float scal_h[2] = {0.f, 0.f};
    long long arg_h = -1;
    CUDA_TRY(ctx, cudaMemcpyAsync(scal_h, sc.scal, 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(&arg_h, sc.arg, 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    if (winner) {
        const int64_t wi = scal_h[0] > 0.f ? (int64_t)arg_h : -1; // winner is always a host scalar
        if (d.memory == LWC_MEM_HOST)
            *winner = wi;
        else
            CUDA_TRY(ctx, cudaMemcpy(winner, &wi, 8, cudaMemcpyHostToDevice));
    }
    (void)n_total;
    return LWC_OK;
}

int32_t lwc_stage_f64(lwc_ctx *ctx, const double *src, size_t n, float *dst) {
    if (n && (!src || !dst)) return fail(ctx, LWC_ERR_INVALID, "lwc_stage_f64: NULL argument");
    for (size_t i = 0; i < n; ++i) dst[i] = static_cast<float>(src[i]);
    return LWC_OK;
}

} // extern "C"

namespace {

std::mutex g_static_mu;
unsigned char *g_static_ptr[64] = {nullptr}; // device address of g_static_scratch, per device ordinal

// Resolves (once per device, with that device current) where g_static_scratch lives on ctx's device.
int static_scratch_address(lwc_ctx *ctx, unsigned char **out) {
    std::lock_guard<std::mutex> lk(g_static_mu);
    const int slot = ctx->device & 63;
    if (!g_static_ptr[slot]) {
        CUDA_TRY(ctx, cudaSetDevice(ctx->device));
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || !p)
            return fail(ctx, LWC_ERR_INTERNAL, "cudaGetSymbolAddress(static scratch) failed on device %d: %s", ctx->device,
                        cudaGetErrorString(e));
        g_static_ptr[slot] = static_cast<unsigned char *>(p);
    }
    *out = g_static_ptr[slot];
    return LWC_OK;
}

__global__ void pool_winner_kernel(const float *__restrict__ scal, const long long *__restrict__ arg,
                                   long long *__restrict__ winner) {
    if (threadIdx.x == 0 && blockIdx.x == 0) *winner = scal[0] > 0.f ? arg[0] : -1;
}

} // namespace

extern "C" int32_t lwc_pool_consensus_static(lwc_ctx *ctx, const lwc_pool_desc *desc, const float *E, const float *V,
                                             const float *w, float *conf, float *cw, float *agree, int64_t *winner,
                                             void *stream) {
    if (!ctx) return LWC_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (!desc) return fail(ctx, LWC_ERR_INVALID, "lwc_pool_consensus_static: desc is NULL");
    const lwc_pool_desc d = *desc;
    if (d.memory != LWC_MEM_DEVICE)
        return fail(ctx, LWC_ERR_INVALID, "lwc_pool_consensus_static: every pointer must be device memory (LWC_MEM_DEVICE); "
                                          "staging host buffers would need an allocation");
    if (ctx->comm) return fail(ctx, LWC_ERR_INVALID, "lwc_pool_consensus_static: single-GPU only, but a communicator is attached");
    if (d.N < 1 || d.D < 8 || d.D % 8 != 0 || d.M < 0 || d.M > LWC_MAX_VOTERS)
        return fail(ctx, LWC_ERR_INVALID, "lwc_pool_consensus_static: invalid shape (N=%lld D=%d M=%d); need N>=1, D%%8==0, 0<=M<=%d",
                    (long long)d.N, d.D, d.M, LWC_MAX_VOTERS);
    if (d.N > LWC_STATIC_POOL_MAX_ROWS || d.N * (long long)d.D > LWC_STATIC_POOL_MAX_ELEMS)
        return fail(ctx, LWC_ERR_INVALID,
                    "lwc_pool_consensus_static: pool of %lld x %d exceeds the fixed scratch (at most %d rows and %d elements, "
                    "%u bytes per GPU); use lwc_pool_consensus",
                    (long long)d.N, d.D, LWC_STATIC_POOL_MAX_ROWS, LWC_STATIC_POOL_MAX_ELEMS, (unsigned)LWC_STATIC_SCRATCH_BYTES);
    if (d.precision != LWC_PREC_BF16)
        return fail(ctx, LWC_ERR_INVALID, "lwc_pool_consensus_static: only LWC_PREC_BF16 is implemented for the pool path");
    if (!E || (d.M > 0 && (!V || !w))) return fail(ctx, LWC_ERR_INVALID, "lwc_pool_consensus_static: NULL input");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    unsigned char *base = nullptr;
    int rc = static_scratch_address(ctx, &base);
    if (rc) return rc;
    cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : ctx->stream;

    // carve the fixed array (256-byte aligned pieces); sizes follow pool_run's needs
    const long long n = d.N;
    const int nblk = (int)((n + lwc::kPoolRedPerBlock - 1) / lwc::kPoolRedPerBlock);
    const int m1 = d.M > 0 ? d.M : 1;
    size_t off = 0;
    auto take = [&off](size_t bytes) {
        const size_t at = off;
        off = (off + bytes + 255) & ~(size_t)255;
        return at;
    };
    const size_t o_small = take(1024);                 // scalars first: the only part that must start at zero
    const size_t o_ehat = take((size_t)n * d.D * 2);
    const size_t o_t = take((size_t)n * 4);
    const size_t o_ypart = take((size_t)64 * n * 4);
    const size_t o_cw = take((size_t)n * 4);
    const size_t o_conf = take((size_t)n * 4);
    const size_t o_bsum = take((size_t)2 * (nblk + 16) * 4);
    const size_t o_barg = take((size_t)(nblk + 16) * 8);
    const size_t o_bagree = take((size_t)nblk * m1 * 4);
    const size_t o_agree = take((size_t)m1 * 4);
    if (off > (size_t)LWC_STATIC_SCRATCH_BYTES)
        return fail(ctx, LWC_ERR_INVALID, "lwc_pool_consensus_static: needs %zu bytes of scratch, the fixed array has %u", off,
                    (unsigned)LWC_STATIC_SCRATCH_BYTES);
    PoolScratch sc;
    sc.scal = reinterpret_cast<float *>(base + o_small);            // 8 floats
    sc.arg = reinterpret_cast<long long *>(base + o_small + 512);   // 8 int64
    sc.ehat_all = reinterpret_cast<__nv_bfloat16 *>(base + o_ehat);
    sc.t_all = reinterpret_cast<float *>(base + o_t);
    sc.ypart = reinterpret_cast<float *>(base + o_ypart);
    sc.cw = reinterpret_cast<float *>(base + o_cw);
    sc.conf = reinterpret_cast<float *>(base + o_conf);
    sc.blk_sum = reinterpret_cast<float *>(base + o_bsum);
    sc.blk_max = sc.blk_sum + nblk + 16;
    sc.blk_arg = reinterpret_cast<long long *>(base + o_barg);
    sc.blk_agree = reinterpret_cast<float *>(base + o_bagree);
    sc.agree = reinterpret_cast<float *>(base + o_agree);
    // the array still holds the previous call's data: zero the reduction words, the rest is overwritten before use
    CUDA_TRY(ctx, cudaMemsetAsync(base + o_small, 0, 1024, st));
    if ((rc = pool_run(ctx, sc, E, V, w, n, d.D, d.M, st))) return rc;
    if (conf) CUDA_TRY(ctx, cudaMemcpyAsync(conf, sc.conf, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
    if (cw) CUDA_TRY(ctx, cudaMemcpyAsync(cw, sc.cw, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
    if (agree && d.M > 0) CUDA_TRY(ctx, cudaMemcpyAsync(agree, sc.agree, (size_t)d.M * 4, cudaMemcpyDeviceToDevice, st));
    if (winner) {
        pool_winner_kernel<<<1, 32, 0, st>>>(sc.scal, sc.arg, reinterpret_cast<long long *>(winner));
        CUDA_TRY(ctx, cudaGetLastError());
        ctx->launches++;
    }
    return LWC_OK;
}
