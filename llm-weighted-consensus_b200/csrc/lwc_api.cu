// lwc_api.cu -- the C ABI (include/lwc_b200.h) over the sm_100a scoring kernels: context and streams,
// shape validation, kernel routing, and the chunked host-memory pipeline (H2D on a copy stream
// overlapped with kernels on the compute stream).  The single-pool path lives in lwc_pool_api.cu.
// There is no CPU fallback anywhere in this library: without a CUDA device every entry point that
// computes returns 503/500.
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>
#include <algorithm>

#include "lwc_internal.hpp"
#include "score_generic.cuh"
#include "score_small.cuh"
#include "score_stream.cuh"
#include "score_tc.cuh"
#include "score_tc16.cuh"
#include "score_tc16a.cuh"
#include "convert.cuh"
#include "vote_kernel.cuh"

#define LWC_VERSION_STRING "lwc_b200 0.1.0 (sm_100a)"

namespace lwc_host {

static std::mutex g_err_mu;
static std::string g_create_err = "";

int fail(lwc_ctx *ctx, int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (ctx) {
        std::lock_guard<std::mutex> lk(ctx->err_mu);
        ctx->err = buf;
    } else {
        std::lock_guard<std::mutex> lk(g_err_mu);
        g_create_err = buf;
    }
    return code;
}

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

std::string last_create_error() {
    std::lock_guard<std::mutex> lk(g_err_mu);
    return g_create_err;
}

} // namespace lwc_host

using lwc_host::fail;
using lwc_host::Slot;
using lwc_host::EncodeTiledFn;
using lwc_host::encode_tiled_fn;

namespace {

enum Route { ROUTE_INVALID = 0, ROUTE_GENERIC, ROUTE_SMALL, ROUTE_TC };

struct Plan {
    Route route = ROUTE_INVALID;
    int np = 0;      // padded pool size
    bool x3 = true;  // small route: compensated 3xTF32
    bool pair = false; // tcgen05 fp32-grade route: two requests of <= 64 candidates per tile
    size_t req_bytes = 0; // N * D * 4: embeddings of one request
    int n = 0;            // candidates
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
    pl.req_bytes = (size_t)d.N * d.D * 4;
    pl.n = d.N;
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
        pl.name = pl.x3 ? "tc128_f16x2" : "tc128_tf32"; // LWC_TC_X3=tf32 runs the fp32-grade route on tf32 pieces ("tc128_tf32x3")
        // pools of at most 64 candidates: two requests per 128-row tile (score_tc16.cuh, PAIR); LWC_TC_PAIR=0 keeps one
        pl.pair = d.N <= 64;
        if (pl.pair) pl.name = pl.x3 ? "tc64x2_f16x2" : "tc64x2_tf32";
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
    static std::atomic<bool> configured[64]; // zero-initialised; set once per device (contexts on other threads may race here)
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

template <int NT, bool X3, int KW, int U, bool F16>
cudaError_t launch_small_one(const lwc::ScoreParams &p, cudaStream_t st) {
    using C = lwc::SmallCfg<NT, X3, KW, U>;
    static std::atomic<bool> configured[64]; // zero-initialised; set once per device (contexts on other threads may race here)
    int dev = 0;
    cudaGetDevice(&dev);
    if (!configured[dev & 63]) {
        cudaError_t e = cudaFuncSetAttribute(lwc::score_small_kernel<NT, X3, KW, U, F16>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::smem_bytes());
        if (e != cudaSuccess) return e;
        configured[dev & 63] = true;
    }
    lwc::score_small_kernel<NT, X3, KW, U, F16><<<p.R, KW * 32, C::smem_bytes(), st>>>(p);
    return cudaGetLastError();
}

// the cfg3 route: 32-row fp32-grade pools, fp16 pieces, loads pipelined through PS shared-memory stages
template <int PS, int NB = 4>
cudaError_t launch_small_pipe32(const lwc::ScoreParams &p, cudaStream_t st) {
    constexpr size_t smem = lwc::small_smem_bytes<32, true, 2, 2, PS>();
    static std::atomic<bool> configured[64];
    int dev = 0;
    cudaGetDevice(&dev);
    if (!configured[dev & 63]) {
        cudaError_t e = cudaFuncSetAttribute(lwc::score_small_kernel<32, true, 2, 2, true, PS, NB>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        configured[dev & 63] = true;
    }
    lwc::score_small_kernel<32, true, 2, 2, true, PS, NB><<<p.R, 64, smem, st>>>(p);
    return cudaGetLastError();
}

// f16: fp32-grade route on fp16 pieces (default for X3; LWC_SMALL_X3=tf32 keeps tf32 pieces)
template <int NT, bool X3>
cudaError_t launch_small(const lwc::ScoreParams &p, cudaStream_t st, int kw, int u, bool f16) {
    if (u == 0) u = (NT == 16) ? 4 : 2;
    constexpr bool F = X3; // the fp16-piece instantiations exist for the fp32-grade route only
    const int key = kw * 10 + u;
    if (F && f16) {
        switch (key) {
            case 22: return launch_small_one<NT, X3, 2, 2, F>(p, st);
            case 24: return launch_small_one<NT, X3, 2, 4, F>(p, st);
            case 42: return launch_small_one<NT, X3, 4, 2, F>(p, st);
            case 44: return launch_small_one<NT, X3, 4, 4, F>(p, st);
            case 82: return launch_small_one<NT, X3, 8, 2, F>(p, st);
            case 84: return launch_small_one<NT, X3, 8, 4, F>(p, st);
            default: return launch_small_one<NT, X3, 4, 2, F>(p, st);
        }
    }
    switch (key) {
        case 22: return launch_small_one<NT, X3, 2, 2, false>(p, st);
        case 24: return launch_small_one<NT, X3, 2, 4, false>(p, st);
        case 42: return launch_small_one<NT, X3, 4, 2, false>(p, st);
        case 44: return launch_small_one<NT, X3, 4, 4, false>(p, st);
        case 82: return launch_small_one<NT, X3, 8, 2, false>(p, st);
        case 84: return launch_small_one<NT, X3, 8, 4, false>(p, st);
        default: return launch_small_one<NT, X3, 4, 2, false>(p, st);
    }
}

// E viewed as a rank-4 fp32 tensor {32 (k_in), N (row), D/32 (k_out), R}: one box {32, 128, ko, 1} lands as
// ko K-major SWIZZLE_128B operand slabs of [128 rows x 128 B] (score_tc.cuh).
// Pools of at most 64 candidates, two requests to a tile: the view {32 (k_in), N (row), R, D/32 (k_out)}; one box
// {32, 64, 2, ko} lands as ko slabs whose rows 0..63 are request 2g's and 64..127 request 2g+1's (zero-filled past N and R).
bool make_tensor_map_tc_pair(CUtensorMap *tm, const float *E, int R, int N, int D, int ko) {
    EncodeTiledFn fn = encode_tiled_fn();
    if (!fn) return false;
    cuuint64_t dims[4] = {32, (cuuint64_t)N, (cuuint64_t)R, (cuuint64_t)(D / 32)};
    cuuint64_t strides[3] = {(cuuint64_t)D * 4, (cuuint64_t)N * D * 4, 128};
    cuuint32_t box[4] = {32, 64, 2, (cuuint32_t)ko};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    return fn(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float *>(E), dims, strides, box, estr,
              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

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

template <bool X3, bool PAIR = false>
cudaError_t launch_tc(lwc_ctx *ctx, const lwc::ScoreParams &p, cudaStream_t st, const int *fix_flags = nullptr) {
    lwc::TcGeom geo;
    geo.ko = ctx->tc_ko > 0 ? ctx->tc_ko : (X3 ? 2 : 4); // 64 KB per stage either way (profiles/r1_tma_probe2.txt)
    const int nslab = p.D / 32;
    if (geo.ko > nslab) geo.ko = nslab;
    geo.stages = lwc::kTcMaxStages;
    while (geo.stages > 1 && lwc::tc_smem_bytes(geo.ko, geo.stages, X3) > (size_t)ctx->max_smem_optin) --geo.stages;
    if (ctx->tc_stages > 0 && ctx->tc_stages < geo.stages) geo.stages = ctx->tc_stages;
    if (lwc::tc_smem_bytes(geo.ko, geo.stages, X3) > (size_t)ctx->max_smem_optin) return cudaErrorInvalidConfiguration;
    static std::atomic<bool> configured[64]; // zero-initialised; set once per device (contexts on other threads may race here)
    if (!configured[ctx->device & 63]) {
        cudaError_t e = cudaFuncSetAttribute(lwc::score_tc_kernel<X3, PAIR>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             ctx->max_smem_optin);
        if (e != cudaSuccess) return e;
        configured[ctx->device & 63] = true;
    }
    CUtensorMap tm;
    // fp32-grade route: one box per slab (the slabs of a stage are interleaved with their remainder slabs)
    if (!(PAIR ? make_tensor_map_tc_pair(&tm, p.E, p.R, p.N, p.D, geo.ko) : make_tensor_map_tc(&tm, p.E, p.R, p.N, p.D, X3 ? 1 : geo.ko)))
        return cudaErrorInvalidValue;
    const int ntiles = PAIR ? (p.R + 1) / 2 : p.R;
    const int grid = ntiles < ctx->sm_count ? ntiles : ctx->sm_count;
    lwc::score_tc_kernel<X3, PAIR><<<grid, X3 ? lwc::kTcThreadsX3 : lwc::kTcThreads, lwc::tc_smem_bytes(geo.ko, geo.stages, X3), st>>>(
        tm, p, geo, fix_flags);
    return cudaGetLastError();
}

cudaError_t ensure_fix_flags(lwc_ctx *ctx, int R);

// fp32-grade route at kind::f16 rate (score_tc16.cuh) + its fix-up pass: the requests whose norms fall outside
// what fp16 pieces represent exactly enough are marked by the first kernel and re-scored by the tf32 kernel.
cudaError_t launch_tc16(lwc_ctx *ctx, const lwc::ScoreParams &p, cudaStream_t st, bool pair) {
    lwc::TcGeom geo;
    geo.ko = 2;
    geo.stages = lwc::kT16MaxStages;
    while (geo.stages > 2 && lwc::tc16_smem_bytes(geo.stages) > (size_t)ctx->max_smem_optin) --geo.stages;
    if (ctx->tc_stages > 1 && ctx->tc_stages < geo.stages) geo.stages = ctx->tc_stages;
    if (lwc::tc16_smem_bytes(geo.stages) > (size_t)ctx->max_smem_optin) return cudaErrorInvalidConfiguration;
    static std::atomic<bool> configured[64];
    if (!configured[ctx->device & 63]) {
        cudaError_t e = cudaFuncSetAttribute(lwc::score_tc16_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, ctx->max_smem_optin);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(lwc::score_tc16_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, ctx->max_smem_optin);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(lwc::score_tc16a_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, ctx->max_smem_optin);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(lwc::score_tc16a_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, ctx->max_smem_optin);
        if (e != cudaSuccess) return e;
        configured[ctx->device & 63] = true;
    }
    {
        cudaError_t e = ensure_fix_flags(ctx, p.R);
        if (e != cudaSuccess) return e;
    }
    CUtensorMap tm;
    if (!(pair ? make_tensor_map_tc_pair(&tm, p.E, p.R, p.N, p.D, 2) : make_tensor_map_tc(&tm, p.E, p.R, p.N, p.D, 2)))
        return cudaErrorInvalidValue;
    const int ntiles = pair ? (p.R + 1) / 2 : p.R;
    const int grid = ntiles < ctx->sm_count ? ntiles : ctx->sm_count;
    // LWC_TC16_A=tmem: the A operand in tensor memory (score_tc16a.cuh); smem: both operands in shared memory
    if (ctx->tc16_a_tmem) {
        if (pair)
            lwc::score_tc16a_kernel<true><<<grid, lwc::kT16aThreads, lwc::tc16_smem_bytes(geo.stages), st>>>(tm, p, geo, static_cast<int *>(ctx->tc_flags));
        else
            lwc::score_tc16a_kernel<false><<<grid, lwc::kT16aThreads, lwc::tc16_smem_bytes(geo.stages), st>>>(tm, p, geo, static_cast<int *>(ctx->tc_flags));
    } else if (pair)
        lwc::score_tc16_kernel<true><<<grid, lwc::kT16Threads, lwc::tc16_smem_bytes(geo.stages), st>>>(tm, p, geo, static_cast<int *>(ctx->tc_flags));
    else
        lwc::score_tc16_kernel<false><<<grid, lwc::kT16Threads, lwc::tc16_smem_bytes(geo.stages), st>>>(tm, p, geo, static_cast<int *>(ctx->tc_flags));
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    ctx->launches++;
    return launch_tc<true>(ctx, p, st, static_cast<const int *>(ctx->tc_flags));
}

// per-request fix-up marks of the fp16-piece tcgen05 route (score_tc16.cuh): one word per request
cudaError_t ensure_fix_flags(lwc_ctx *ctx, int R) {
    if ((size_t)R * sizeof(int) <= ctx->tc_flags_cap) return cudaSuccess;
    if (ctx->tc_flags) {
        cudaError_t e = cudaFree(ctx->tc_flags);
        if (e != cudaSuccess) return e;
    }
    ctx->tc_flags = nullptr, ctx->tc_flags_cap = 0;
    const size_t want = (size_t)R * sizeof(int) * 2 + 256;
    cudaError_t e = cudaMalloc(&ctx->tc_flags, want);
    if (e == cudaSuccess) ctx->tc_flags_cap = want;
    return e;
}

// list != nullptr: one launch over a list of same-shaped batches (score_stream_list_kernel); p.R = all requests of the list
// GROUP: 8 / N requests of at most 4 candidates share one 8-row tile of the half-tile form (score_stream.cuh)
template <int NT, bool X3, int NCW, int NF, int NB = NT / 8, bool GROUP = false>
cudaError_t launch_stream_one(lwc_ctx *ctx, const lwc::ScoreParams &p, cudaStream_t st, const lwc::ScoreList *list = nullptr) {
    using C = lwc::StreamCfg<NT, X3, NCW, NF>;
    lwc::StreamGeom geo;
    // LWC_STREAM_PARKS=2: a second park buffer, so that the consumers hand request q over while a finaliser is still
    // draining q - 1.  Built because ncu at 16 x 384 showed 18 % of all stall samples on that one wait; measured: no
    // change at any width (256..1024) -- the wait is where the consumers idle, not why.  Default 1.
    geo.parks = ctx->stream_parks == 2 ? 2 : 1;
    const size_t avail = (size_t)ctx->max_smem_optin - 128 - C::fixed_bytes(geo.parks);
    // stage width: 3 KB per row copy is where the bulk-copy engine reaches HBM speed
    // (profiles/r1_tma_probe.txt); shrink it until at least 3 stages fit.
    const int dpad = (p.D + 15) / 16 * 16;
    constexpr int rows = NB * 8; // rows per stage
    // (half-tile form: 6 KB per row copy -- its stages hold half the rows, and a stage hand-off costs the consumers the
    // same whatever it carries: 0.84 -> 0.98 of HBM at 8 candidates x 3072, profiles/r2_pool_sizes.jsonl)
    int ks = ctx->stream_kc >= 16 ? ctx->stream_kc : (NB == 1 ? 1536 : 768);
    if (ks > dpad) ks = dpad;
    while (ks > 64 && avail / C::stage_bytes(ks, rows) < 3) ks = ((ks / 2) + 15) / 16 * 16;
    geo.ks = ks;
    geo.stages = (int)(avail / C::stage_bytes(ks, rows));
    if (geo.stages > C::MAXS) geo.stages = C::MAXS;
    if (ctx->stream_slots >= 1 && ctx->stream_slots < geo.stages) geo.stages = ctx->stream_slots;
    if (geo.stages < 1) return cudaErrorInvalidConfiguration;
    static std::atomic<bool> configured[64]; // zero-initialised; set once per device (contexts on other threads may race here)
    if (!configured[ctx->device & 63]) {
        cudaError_t e = cudaFuncSetAttribute(lwc::score_stream_kernel<NT, X3, NCW, NF, NB, GROUP>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, ctx->max_smem_optin);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(lwc::score_stream_list_kernel<NT, X3, NCW, NF, NB, GROUP>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     ctx->max_smem_optin);
        if (e != cudaSuccess) return e;
        configured[ctx->device & 63] = true;
    }
    const int per_tile = GROUP ? 8 / p.N : 1;
    const int ntiles = (p.R + per_tile - 1) / per_tile;
    const int grid = ntiles < ctx->sm_count ? ntiles : ctx->sm_count;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(C::THREADS);
    cfg.dynamicSmemBytes = C::smem_bytes(geo.ks, geo.stages, rows, geo.parks);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; // pairs with griddepcontrol.* in the kernel
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if (list) return cudaLaunchKernelEx(&cfg, lwc::score_stream_list_kernel<NT, X3, NCW, NF, NB, GROUP>, p, geo, *list);
    return cudaLaunchKernelEx(&cfg, lwc::score_stream_kernel<NT, X3, NCW, NF, NB, GROUP>, p, geo);
}

template <int NT, bool X3>
cudaError_t launch_stream(lwc_ctx *ctx, const lwc::ScoreParams &p, cudaStream_t st) {
    const int ncw = ctx->stream_ncw ? ctx->stream_ncw : (NT == 16 ? 12 : 8);
    const int nf = ctx->stream_nf ? ctx->stream_nf : (NT == 16 ? 4 : 2);
    // pools of at most 8 candidates: the half-tile form (score_stream.cuh, NB = 1); LWC_STREAM_HALF=0 keeps the full tile
    // at most 4 candidates: 8 / N requests per tile of that form; LWC_STREAM_GROUP=0 keeps one
    if (NT == 16 && p.N <= 4 && !ctx->stream_no_half && !ctx->stream_no_group && ncw == 12 && nf == 4)
        return launch_stream_one<16, X3, 12, 4, 1, true>(ctx, p, st);
    if (NT == 16 && p.N <= 8 && !ctx->stream_no_half && ncw == 12 && nf == 4) return launch_stream_one<16, X3, 12, 4, 1>(ctx, p, st);
    if (NT == 32 && p.N <= 24 && !ctx->stream_no_half && ncw == 8 && nf == 2) return launch_stream_one<32, X3, 8, 2, 3>(ctx, p, st);
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
        case 126: return launch_stream_one<NT, X3, 12, 6>(ctx, p, st);
        case 128: return launch_stream_one<NT, X3, 12, 8>(ctx, p, st);
        default: return launch_stream_one<NT, X3, 8, 2>(ctx, p, st);
    }
}

// Small pools (N <= 32): which kernel, on which pieces.
//   ring kernel (score_stream): tf32 pieces.  16-row pools run it at 0.97 of HBM.  An fp16-piece form with a fix-up pass
//   was built and measured for 32-row pools (cfg3 slice 1.231 ms against 1.127 ms for the direct kernel below: its eight
//   consumer warps became issue-bound on the piece arithmetic) and dropped.
//   direct kernel (score_small): fp32-grade 32-row pools by default, and everything under LWC_SMALL_IMPL=direct;
//   fp16 pieces with an inline tf32 fallback (LWC_SMALL_X3=tf32: tf32 pieces only).
constexpr size_t kDirectBelowBytes = (size_t)56 << 10, kDirectBelowBytesTf32 = (size_t)40 << 10, kDirectBelowBytes8 = (size_t)20 << 10,
                 kDirectBelowBytes32 = (size_t)72 << 10;
struct SmallRoute {
    bool stream; // persistent ring kernel, else one CTA per request
    bool f16;    // fp32-grade Gram on fp16 pieces
};
SmallRoute small_route(const lwc_ctx *ctx, const Plan &pl) {
    const bool f16_ok = pl.x3 && !ctx->small_x3_tf32;
    SmallRoute r;
    if (ctx->small_impl == 1)
        r.stream = false;
    else if (ctx->force_stream)
        r.stream = true;
    else if (pl.np == 32 && pl.x3)
        r.stream = false; // 32-row fp32-grade pools: the direct kernel is faster on either kind of pieces
    else
        // Small requests: the ring kernel pays ~1.1 us per request whatever its size (every consumer warp meets every
        // stage, parks, a finaliser drains), the direct kernel a CTA launch.  Measured crossover (profiles/
        // r2_width_sweep.jsonl): 16 x 384 runs at 0.78 of HBM direct against 0.49 on the ring, 16 x 768 at 0.86 / 0.82,
        // 16 x 1536 (cfg2) at 0.93 / 1.00.  LWC_SMALL_IMPL=stream keeps the ring for everything.
        // (at most 4 candidates: 8 / N requests share a tile and the ring is kept down to 5 KB per tile -- 3 x 768 runs at
        // 0.50 of HBM there, 0.24 one CTA per request; 3 x 384 at 0.17 one CTA per request)
        // Pools of at most 8 candidates keep the ring down to 20 KB (its half-tile form does half the work per byte;
        // 8 x 1024 runs at 0.74 there against 0.61 direct), single-pass TF32 down to 40 KB (16 x 768: 1.00 / 0.96).
        // 9..12 candidates, fp32-grade: the ring does a full 16-row tile's work for them at any width (10 x 1536: 0.63 of
        // HBM, compute-bound like 16 x 1536) -- direct at every size (0.75-0.77; 12: 0.80-0.85 against 0.75-0.84).
        r.stream = !(pl.x3 && pl.n >= 9 && pl.n <= 12) && pl.req_bytes * (pl.n >= 1 && pl.n <= 4 && !ctx->stream_no_group ? 4 * (8 / pl.n) : 1) > (pl.n <= 8 ? kDirectBelowBytes8 : (pl.np == 32 ? kDirectBelowBytes32 : (pl.x3 ? kDirectBelowBytes : kDirectBelowBytesTf32)));
    r.f16 = f16_ok && !r.stream;
    return r;
}

// One launch for a run of same-shaped batches on the ring kernel (default consumer / finaliser counts only).
// Returns false when this shape does not take that route: the caller then launches batch by batch.
bool list_route(const lwc_ctx *ctx, const Plan &pl) {
    return pl.route == ROUTE_SMALL && small_route(ctx, pl).stream && ctx->stream_ncw == 0 && ctx->stream_nf == 0;
}
int launch_list(lwc_ctx *ctx, const Plan &pl, const lwc::ScoreParams &p, const lwc::ScoreList &list, cudaStream_t st) {
    cudaError_t e;
    if (pl.np == 16 && p.N <= 4 && !ctx->stream_no_half && !ctx->stream_no_group && list.R_each % (8 / p.N) == 0)
        e = pl.x3 ? launch_stream_one<16, true, 12, 4, 1, true>(ctx, p, st, &list) : launch_stream_one<16, false, 12, 4, 1, true>(ctx, p, st, &list);
    else if (pl.np == 16 && p.N <= 8 && !ctx->stream_no_half)
        e = pl.x3 ? launch_stream_one<16, true, 12, 4, 1>(ctx, p, st, &list) : launch_stream_one<16, false, 12, 4, 1>(ctx, p, st, &list);
    else if (pl.np == 16)
        e = pl.x3 ? launch_stream_one<16, true, 12, 4>(ctx, p, st, &list) : launch_stream_one<16, false, 12, 4>(ctx, p, st, &list);
    else if (p.N <= 24 && !ctx->stream_no_half)
        e = pl.x3 ? launch_stream_one<32, true, 8, 2, 3>(ctx, p, st, &list) : launch_stream_one<32, false, 8, 2, 3>(ctx, p, st, &list);
    else
        e = pl.x3 ? launch_stream_one<32, true, 8, 2>(ctx, p, st, &list) : launch_stream_one<32, false, 8, 2>(ctx, p, st, &list);
    if (e != cudaSuccess) return fail(ctx, LWC_ERR_INTERNAL, "kernel launch (%s, list of %d) failed: %s", pl.name, list.n, cudaGetErrorString(e));
    ctx->launches++;
    return LWC_OK;
}

int launch_plan(lwc_ctx *ctx, const Plan &pl, const lwc::ScoreParams &p, cudaStream_t st) {
    if (p.R == 0) return LWC_OK;
    cudaError_t e = cudaSuccess;
    const SmallRoute sr = small_route(ctx, pl);
    if (pl.route == ROUTE_TC) {
        const bool pair = pl.pair && !ctx->tc_no_pair;
        e = !pl.x3 ? (pair ? launch_tc<false, true>(ctx, p, st) : launch_tc<false>(ctx, p, st)) : (ctx->tc_x3_tf32 ? launch_tc<true>(ctx, p, st) : launch_tc16(ctx, p, st, pair));
    } else if (pl.route == ROUTE_SMALL && sr.stream) {
        if (pl.np == 16)
            e = pl.x3 ? launch_stream<16, true>(ctx, p, st) : launch_stream<16, false>(ctx, p, st);
        else
            e = pl.x3 ? launch_stream<32, true>(ctx, p, st) : launch_stream<32, false>(ctx, p, st);
    } else if (pl.route == ROUTE_SMALL) {
        if (pl.np == 16)
            // (2 warps per request: 16 x 384 at 0.78 of HBM against 0.60 with 4)
            e = pl.x3 ? launch_small<16, true>(p, st, ctx->small_kw_set ? ctx->small_kw : 2, ctx->small_u, sr.f16)
                      : launch_small<16, false>(p, st, ctx->small_kw_set ? ctx->small_kw : 2, ctx->small_u, false);
        else if (pl.x3 && sr.f16 && !ctx->small_kw_set && ctx->small_u == 0 && ctx->small_pipe != 0) // the cfg3 route
            e = ctx->small_pipe != 3 ? launch_small_pipe32<4>(p, st)
                                     : (p.N <= 24 && !ctx->stream_no_half ? launch_small_pipe32<3, 3>(p, st) : launch_small_pipe32<3>(p, st));
        else // 32-row pools: 2 warps per request measured best (1.157 vs 1.185 ms at KW=4, cfg3 slice)
            e = pl.x3 ? launch_small<32, true>(p, st, ctx->small_kw_set ? ctx->small_kw : 2, ctx->small_u, sr.f16)
                      : launch_small<32, false>(p, st, ctx->small_kw_set ? ctx->small_kw : 2, ctx->small_u, false);
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

size_t elem_size(int e_type) { return e_type == LWC_ELEM_F64 ? 8 : (e_type == LWC_ELEM_BF16 ? 2 : 4); }

// raw (f64 / bf16) -> fp32 on `st`; both device pointers 16-byte aligned (cudaMalloc)
int launch_convert(lwc_ctx *ctx, int e_type, const void *raw, float *dst, size_t n, cudaStream_t st) {
    if (n == 0) return LWC_OK;
    const size_t per_block = 256 * (e_type == LWC_ELEM_F64 ? 4 : 8);
    size_t blocks = (n + per_block - 1) / per_block;
    const size_t cap = (size_t)ctx->sm_count * 8; // grid-stride: eight resident CTAs per SM keep the loads in flight
    if (blocks > cap) blocks = cap;
    if (e_type == LWC_ELEM_F64)
        lwc::convert_f64_f32_kernel<<<(unsigned)blocks, 256, 0, st>>>(static_cast<const double *>(raw), dst, n);
    else
        lwc::expand_bf16_f32_kernel<<<(unsigned)blocks, 256, 0, st>>>(static_cast<const uint16_t *>(raw), dst, n);
    CUDA_TRY(ctx, cudaGetLastError());
    ctx->launches++;
    return LWC_OK;
}

// Host-memory call for a batch of a few requests (BASELINE configs[0] is ONE request): latency, not bandwidth.  The
// chunked pipeline below issues three H2D and five D2H copies over two streams with events between them.  Here V and w
// are packed into a page-locked block the kernel reads over PCIe, the kernel writes every output into the same block
// (posted PCIe writes), and E is read where it lies when the caller's buffer is page-locked, else copied once: one
// launch, one synchronise, at most one copy.  Measured on B200, 1 x 16 x 1536 x 4 from pinned buffers, median wall
// clock around the call: 46-50 us chunked pipeline, 41.6 all-HBM packed (mode 0), 32.1 outputs over PCIe (1),
// 22.8-24.6 V/w too (2), 19.2 E in place (4, default); the CPU port takes 48.5 us (profiles/r2_cfg1_latency.jsonl).
constexpr size_t kSmallCallBytes = (size_t)256 << 10;

int score_host_small(lwc_ctx *ctx, const lwc_batch_desc &d, const Plan &pl, const float *E, const float *V, const float *w,
                     float *conf, float *cw, float *agree, int32_t *winner, int32_t *rank) {
    const bool cosine = d.similarity == LWC_SIM_COSINE;
    auto up = [](size_t x) { return (x + 255) & ~(size_t)255; };
    const size_t R = (size_t)d.R, N = (size_t)d.N, M = (size_t)d.M;
    const size_t e_b = cosine ? R * N * d.D * 4 : 0, v_b = R * M * N * 4, w_b = (d.w_per_request ? R : 1) * M * 4;
    const size_t o_e = 0, o_v = up(e_b), o_w = o_v + up(v_b), o_out = o_w + up(w_b);
    const size_t c_b = R * N * 4, a_b = R * M * 4, wi_b = R * 4;
    const size_t o_conf = o_out, o_cw = o_conf + up(c_b), o_ag = o_cw + up(c_b), o_win = o_ag + up(a_b), o_rank = o_win + up(wi_b);
    const size_t total = o_rank + up(c_b);
    if (total > ctx->small_cap) {
        if (ctx->small_dev) CUDA_TRY(ctx, cudaFree(ctx->small_dev));
        if (ctx->small_pin) CUDA_TRY(ctx, cudaFreeHost(ctx->small_pin));
        ctx->small_dev = ctx->small_pin = nullptr, ctx->small_cap = 0;
        const size_t want = total * 2 + 4096;
        CUDA_TRY(ctx, cudaMalloc(&ctx->small_dev, want));
        CUDA_TRY(ctx, cudaHostAlloc(&ctx->small_pin, want, cudaHostAllocDefault));
        ctx->small_cap = want;
    }
    char *dev = static_cast<char *>(ctx->small_dev), *pin = static_cast<char *>(ctx->small_pin);
    cudaStream_t st = ctx->stream;
    // small_zero_copy: 0 = every buffer in HBM (3 copies); 1 = outputs written into the pinned block; 2 = V and w read
    // from it too (E copied); 3 = E through the block as well; 4 = E read in place when page-locked, else as 2.
    const int zc = ctx->small_zero_copy;
    char *in_e = zc >= 3 ? pin : dev, *in_vw = zc >= 2 ? pin : dev, *outb = zc >= 1 ? pin : dev;
    const float *e_direct = nullptr; // mode 4: the caller's E is page-locked -- the kernel reads it where it lies
    if (cosine && zc >= 4) {
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, E) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer)
            e_direct = static_cast<const float *>(at.devicePointer);
        else
            in_e = dev;
        (void)cudaGetLastError();
    }
    if (cosine && !e_direct) {
        if (in_e == pin) memcpy(pin + o_e, E, e_b);
        else CUDA_TRY(ctx, cudaMemcpyAsync(dev + o_e, E, e_b, cudaMemcpyHostToDevice, st));
    }
    if (M > 0) {
        memcpy(pin + o_v, V, v_b);
        memcpy(pin + o_w, w, w_b);
        if (zc < 2) CUDA_TRY(ctx, cudaMemcpyAsync(dev + o_v, pin + o_v, o_out - o_v, cudaMemcpyHostToDevice, st));
    }
    lwc::ScoreParams p;
    p.E = e_direct ? e_direct : reinterpret_cast<const float *>(in_e + o_e);
    p.V = reinterpret_cast<const float *>(in_vw + o_v);
    p.w = reinterpret_cast<const float *>(in_vw + o_w);
    p.conf = conf ? reinterpret_cast<float *>(outb + o_conf) : nullptr;
    p.cw = cw ? reinterpret_cast<float *>(outb + o_cw) : nullptr;
    p.agree = agree ? reinterpret_cast<float *>(outb + o_ag) : nullptr;
    p.winner = winner ? reinterpret_cast<int32_t *>(outb + o_win) : nullptr;
    p.rank = rank ? reinterpret_cast<int32_t *>(outb + o_rank) : nullptr;
    p.R = d.R, p.N = d.N, p.D = d.D, p.M = d.M;
    p.w_per_request = d.w_per_request;
    p.cosine = cosine ? 1 : 0;
    p.inputs_resident = 0;
    int rc = launch_plan(ctx, pl, p, st);
    if (rc) return rc;
    if (zc < 1) CUDA_TRY(ctx, cudaMemcpyAsync(pin + o_out, dev + o_out, total - o_out, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    if (conf) memcpy(conf, pin + o_conf, c_b);
    if (cw) memcpy(cw, pin + o_cw, c_b);
    if (agree && M > 0) memcpy(agree, pin + o_ag, a_b);
    if (winner) memcpy(winner, pin + o_win, wi_b);
    if (rank) memcpy(rank, pin + o_rank, c_b);
    return LWC_OK;
}

int score_host(lwc_ctx *ctx, const lwc_batch_desc &d, const Plan &pl, int e_type, const void *E, const float *V, const float *w,
               float *conf, float *cw, float *agree, int32_t *winner, int32_t *rank) {
    const bool cosine = d.similarity == LWC_SIM_COSINE;
    if (e_type == LWC_ELEM_F32 && ctx->chunk_requests <= 0 &&
        (size_t)d.R * ((cosine ? (size_t)d.N * d.D * 4 : 0) + (size_t)d.M * d.N * 4 + (size_t)d.N * 16 + (size_t)d.M * 8 + 1024) <= kSmallCallBytes)
        return score_host_small(ctx, d, pl, static_cast<const float *>(E), V, w, conf, cw, agree, winner, rank);
    const size_t e_req = cosine ? (size_t)d.N * d.D * sizeof(float) : 0;         // fp32 tile the kernels read
    const size_t e_raw = cosine ? (size_t)d.N * d.D * elem_size(e_type) : 0;     // what crosses PCIe
    const bool typed = e_type != LWC_ELEM_F32;
    const size_t v_req = (size_t)d.M * d.N * sizeof(float);
    int chunk = ctx->chunk_requests;
    if (chunk <= 0) {
        // ~96 MiB of embeddings per chunk: PCIe efficiency peaks there (measured: 548 k req/s at 32 MiB,
        // 555 k at 96-192 MiB on cfg2) and four chunks still overlap copies with kernels
        size_t per = e_raw + v_req + 64;
        chunk = (int)((96u << 20) / per);
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
        if (typed && (rc = ensure(ctx, s, 6, e_raw * rc_n))) return rc; // f64 / bf16 as it arrived
        if ((rc = ensure(ctx, s, 1, v_req * rc_n))) return rc;
        if ((rc = ensure(ctx, s, 2, (size_t)d.M * rc_n * sizeof(float)))) return rc;
        if ((rc = ensure(ctx, s, 3, (size_t)d.N * rc_n * sizeof(float) * 2))) return rc;             // conf | cw
        if ((rc = ensure(ctx, s, 4, (size_t)d.M * rc_n * sizeof(float)))) return rc;                 // agree
        if ((rc = ensure(ctx, s, 5, (size_t)rc_n * sizeof(int32_t) * ((size_t)d.N + 1)))) return rc; // winner | rank
        if (cosine)
            CUDA_TRY(ctx, cudaMemcpyAsync(typed ? s.buf[6] : s.buf[0], static_cast<const char *>(E) + (size_t)r0 * e_raw, e_raw * rc_n,
                                          cudaMemcpyHostToDevice, ctx->copy_stream));
        if (v_req)
            CUDA_TRY(ctx, cudaMemcpyAsync(s.buf[1], V + (size_t)r0 * d.M * d.N, v_req * rc_n, cudaMemcpyHostToDevice,
                                          ctx->copy_stream));
        if (d.w_per_request && d.M > 0)
            CUDA_TRY(ctx, cudaMemcpyAsync(s.buf[2], w + (size_t)r0 * d.M, (size_t)d.M * rc_n * sizeof(float),
                                          cudaMemcpyHostToDevice, ctx->copy_stream));
        CUDA_TRY(ctx, cudaEventRecord(s.h2d_done, ctx->copy_stream));
        CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->stream, s.h2d_done, 0));
        if (typed && cosine &&
            (rc = launch_convert(ctx, e_type, s.buf[6], static_cast<float *>(s.buf[0]), (size_t)rc_n * d.N * d.D, ctx->stream)))
            return rc;

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
        p.inputs_resident = 0; // staged inputs arrive through the copy stream: keep full stream ordering
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


} // namespace

extern "C" {


const char *lwc_version(void) { return LWC_VERSION_STRING; }

const char *lwc_last_error_message(const lwc_ctx *ctx) {
    // The text is copied under the lock that failing calls hold while they write it, into storage owned by the
    // asking thread: the pointer stays valid until this thread asks again, whatever other threads do meanwhile.
    static thread_local std::string copy;
    if (ctx) {
        std::lock_guard<std::mutex> lk(ctx->err_mu);
        copy = ctx->err;
    } else {
        copy = lwc_host::last_create_error();
    }
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
    ctx->flags = cfg ? cfg->flags : 0;
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
    if (const char *s = getenv("LWC_STREAM_PARKS")) ctx->stream_parks = (atoi(s) == 1 || atoi(s) == 2) ? atoi(s) : 0;
    if (const char *s = getenv("LWC_STREAM_NF")) ctx->stream_nf = atoi(s);
    if (const char *s = getenv("LWC_POOL_NO_P2P")) ctx->no_p2p = atoi(s);
    if (const char *s = getenv("LWC_CHUNK_REQUESTS")) ctx->chunk_requests = atoi(s);
    if (const char *s = getenv("LWC_LIST_FUSION")) ctx->no_list_fusion = atoi(s) == 0 ? 1 : 0;
    if (const char *s = getenv("LWC_TC_KO")) ctx->tc_ko = atoi(s);
    if (const char *s = getenv("LWC_TC_X3")) ctx->tc_x3_tf32 = strcmp(s, "tf32") == 0 ? 1 : 0;
    if (const char *s = getenv("LWC_TC_PAIR")) ctx->tc_no_pair = atoi(s) == 0 ? 1 : 0;
    if (const char *s = getenv("LWC_TC16_A")) ctx->tc16_a_tmem = strcmp(s, "tmem") == 0 ? 1 : 0;
    if (const char *s = getenv("LWC_STREAM_HALF")) ctx->stream_no_half = atoi(s) == 0 ? 1 : 0;
    if (const char *s = getenv("LWC_STREAM_GROUP")) ctx->stream_no_group = atoi(s) == 0 ? 1 : 0;
    if (const char *s = getenv("LWC_SMALL_X3")) ctx->small_x3_tf32 = strcmp(s, "tf32") == 0 ? 1 : 0;
    if (const char *s = getenv("LWC_SMALL_PIPE")) ctx->small_pipe = atoi(s);
    if (const char *s = getenv("LWC_SMALL_ZERO_COPY")) ctx->small_zero_copy = atoi(s);
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
    if (ctx->typed_buf) cudaFree(ctx->typed_buf);
    if (ctx->small_dev) cudaFree(ctx->small_dev);
    if (ctx->small_pin) cudaFreeHost(ctx->small_pin);
    if (ctx->tc_flags) cudaFree(ctx->tc_flags);
    for (auto &b : ctx->table_buf)
        if (b) cudaFree(b);
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
    const Plan pl = plan_for(*desc, nullptr);
    if (ctx && ctx->tc_x3_tf32 && pl.route == ROUTE_TC && pl.x3) return "tc128_tf32x3";
    if (ctx && ctx->tc_no_pair && pl.pair) return pl.x3 ? "tc128_f16x2" : "tc128_tf32";
    if (ctx && pl.route == ROUTE_SMALL && pl.x3) {
        const SmallRoute sr = small_route(ctx, pl);
        if (sr.f16) return pl.np == 16 ? "small16_f16x2" : "small32_f16x2";
    }
    return pl.name;
}

uint64_t lwc_launch_count(const lwc_ctx *ctx) { return ctx ? ctx->launches : 0; }

static int32_t score_typed(lwc_ctx *ctx, const char *fn, const lwc_batch_desc *desc, int e_type, const void *E, const float *V,
                           const float *w, float *conf, float *cw, float *agree, int32_t *winner, int32_t *rank, void *stream) {
    if (!ctx) return LWC_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (!desc) return fail(ctx, LWC_ERR_INVALID, "%s: desc is NULL", fn);
    lwc_batch_desc d = *desc;
    if (e_type != LWC_ELEM_F32 && e_type != LWC_ELEM_F64 && e_type != LWC_ELEM_BF16)
        return fail(ctx, LWC_ERR_INVALID, "%s: element type %d is not LWC_ELEM_F32 / _F64 / _BF16", fn, e_type);
    // bf16 inputs are exact in tf32 (8 of its 10 mantissa bits): the error-compensating second product would be zeros
    if (e_type == LWC_ELEM_BF16 && (d.precision == LWC_PREC_FP32 || d.precision == LWC_PREC_TF32X3 || d.precision == LWC_PREC_BF16))
        d.precision = LWC_PREC_TF32;
    const bool typed = e_type != LWC_ELEM_F32;
    // typed embeddings are converted into a 16-byte aligned fp32 buffer of the library's own: plan for that
    const Plan pl = plan_for(d, typed ? nullptr : E);
    if (pl.route == ROUTE_INVALID)
        return fail(ctx, LWC_ERR_INVALID,
                    "%s: invalid descriptor (R=%d N=%d D=%d M=%d sim=%d prec=%d mem=%d); need R>=0, "
                    "1<=N<=%d, 0<=M<=%d, D>=1 for cosine",
                    fn, d.R, d.N, d.D, d.M, d.similarity, d.precision, d.memory, LWC_MAX_CHOICES, LWC_MAX_VOTERS);
    if (d.R == 0) return LWC_OK;
    const bool cosine = d.similarity == LWC_SIM_COSINE;
    if (cosine && !E) return fail(ctx, LWC_ERR_INVALID, "%s: E is NULL", fn);
    if (typed && cosine && (reinterpret_cast<uintptr_t>(E) & 15u))
        return fail(ctx, LWC_ERR_INVALID, "%s: E must be 16-byte aligned", fn);
    if (d.M > 0 && (!V || !w)) return fail(ctx, LWC_ERR_INVALID, "%s: V or w is NULL with M > 0", fn);
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    if (d.memory == LWC_MEM_HOST) return score_host(ctx, d, pl, e_type, E, V, w, conf, cw, agree, winner, rank);

    cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : ctx->stream;
    lwc::ScoreParams p;
    p.E = static_cast<const float *>(E), p.V = V, p.w = w;
    p.conf = conf, p.cw = cw, p.agree = agree, p.winner = winner, p.rank = rank;
    p.R = d.R, p.N = d.N, p.D = d.D, p.M = d.M;
    p.w_per_request = d.w_per_request;
    p.cosine = cosine ? 1 : 0;
    p.inputs_resident = (ctx->flags & LWC_CTX_INPUTS_RESIDENT) ? 1 : 0;
    if (ctx->weights_in_flight) { // lwc_table_weights(LWC_MEM_DEVICE) queued a producer of `w` on the context's stream:
        p.inputs_resident = 0;    // this launch must wait for it before it reads anything
        ctx->weights_in_flight = 0;
    }
    if (typed && cosine) {
        // device-resident f64 / bf16: converted into the context's conversion buffer on the same stream
        const size_t n = (size_t)d.R * d.N * d.D;
        if (n * sizeof(float) > ctx->typed_cap) {
            if (ctx->typed_buf) CUDA_TRY(ctx, cudaFree(ctx->typed_buf));
            ctx->typed_buf = nullptr, ctx->typed_cap = 0;
            CUDA_TRY(ctx, cudaMalloc(&ctx->typed_buf, n * sizeof(float) + 256));
            ctx->typed_cap = n * sizeof(float) + 256;
        }
        int rc = launch_convert(ctx, e_type, E, static_cast<float *>(ctx->typed_buf), n, st);
        if (rc) return rc;
        p.E = static_cast<const float *>(ctx->typed_buf);
        p.inputs_resident = 0; // E was just produced on this stream: the score kernel's reads must wait for it
    }
    return launch_plan(ctx, pl, p, st);
}

int32_t lwc_score_batch(lwc_ctx *ctx, const lwc_batch_desc *desc, const float *E, const float *V, const float *w,
                        float *conf, float *cw, float *agree, int32_t *winner, int32_t *rank, void *stream) {
    return score_typed(ctx, "lwc_score_batch", desc, LWC_ELEM_F32, E, V, w, conf, cw, agree, winner, rank, stream);
}

int32_t lwc_score_batch_typed(lwc_ctx *ctx, const lwc_batch_desc *desc, int32_t e_type, const void *E, const float *V,
                              const float *w, float *conf, float *cw, float *agree, int32_t *winner, int32_t *rank,
                              void *stream) {
    return score_typed(ctx, "lwc_score_batch_typed", desc, e_type, E, V, w, conf, cw, agree, winner, rank, stream);
}

int32_t lwc_score_batches(lwc_ctx *ctx, int32_t n_batches, const lwc_batch_desc *descs, const float *const *E,
                          const float *const *V, const float *const *w, float *const *conf, float *const *cw,
                          float *const *agree, int32_t *const *winner, int32_t *const *rank, void *stream) {
    if (!ctx) return LWC_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (n_batches < 0 || (n_batches > 0 && !descs)) return fail(ctx, LWC_ERR_INVALID, "lwc_score_batches: bad batch list");
    // validate everything first: either all batches are enqueued or none is
    for (int32_t b = 0; b < n_batches; ++b) {
        const lwc_batch_desc &d = descs[b];
        if (d.memory != LWC_MEM_DEVICE)
            return fail(ctx, LWC_ERR_INVALID, "lwc_score_batches: batch %d is not LWC_MEM_DEVICE (host batches are pipelined by lwc_score_batch)", b);
        if (plan_for(d, E ? E[b] : nullptr).route == ROUTE_INVALID)
            return fail(ctx, LWC_ERR_INVALID, "lwc_score_batches: batch %d has an invalid descriptor (R=%d N=%d D=%d M=%d)", b, d.R, d.N,
                        d.D, d.M);
        if (d.R > 0 && d.similarity == LWC_SIM_COSINE && (!E || !E[b]))
            return fail(ctx, LWC_ERR_INVALID, "lwc_score_batches: batch %d: E is NULL", b);
        if (d.R > 0 && d.M > 0 && (!V || !w || !V[b] || !w[b]))
            return fail(ctx, LWC_ERR_INVALID, "lwc_score_batches: batch %d: V or w is NULL with M > 0", b);
    }
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : ctx->stream;
    auto same_shape = [&](int32_t a, int32_t b) {
        const lwc_batch_desc &x = descs[a], &y = descs[b];
        return x.R == y.R && x.N == y.N && x.D == y.D && x.M == y.M && x.similarity == y.similarity && x.precision == y.precision &&
               x.w_per_request == y.w_per_request;
    };
    for (int32_t b = 0; b < n_batches;) {
        const lwc_batch_desc &d = descs[b];
        if (d.R == 0) {
            ++b;
            continue;
        }
        const Plan pl = plan_for(d, E ? E[b] : nullptr);
        lwc::ScoreParams p;
        p.E = E ? E[b] : nullptr, p.V = V ? V[b] : nullptr, p.w = w ? w[b] : nullptr;
        p.conf = conf ? conf[b] : nullptr, p.cw = cw ? cw[b] : nullptr, p.agree = agree ? agree[b] : nullptr;
        p.winner = winner ? winner[b] : nullptr, p.rank = rank ? rank[b] : nullptr;
        p.R = d.R, p.N = d.N, p.D = d.D, p.M = d.M;
        p.w_per_request = d.w_per_request;
        p.cosine = d.similarity == LWC_SIM_COSINE ? 1 : 0;
        p.inputs_resident = (ctx->flags & LWC_CTX_INPUTS_RESIDENT) ? 1 : 0;
        if (ctx->weights_in_flight) { // see score_typed
            p.inputs_resident = 0;
            ctx->weights_in_flight = 0;
        }
        // A run of same-shaped batches on the ring kernel becomes ONE launch (up to kMaxList batches): the persistent
        // CTAs walk the whole run, so nothing is lost between batches (launch latency, the first bytes' latency, the
        // last requests' tail) and nothing depends on the host keeping ahead of the GPU.  Results are the bits of
        // single launches: a request's arithmetic does not depend on where in a launch it sits.
        int32_t run = 1;
        if (list_route(ctx, pl) && d.similarity == LWC_SIM_COSINE && !ctx->no_list_fusion)
            while (b + run < n_batches && run < lwc::kMaxList && same_shape(b, b + run) &&
                   plan_for(descs[b + run], E[b + run]).route == pl.route && (long long)(run + 1) * d.R < 0x7fffffffLL)
                ++run;
        if (run > 1) {
            lwc::ScoreList list;
            list.n = run, list.R_each = d.R;
            for (int32_t k = 0; k < run; ++k) {
                list.E[k] = E[b + k], list.V[k] = V ? V[b + k] : nullptr, list.w[k] = w ? w[b + k] : nullptr;
                list.conf[k] = conf ? conf[b + k] : nullptr, list.cw[k] = cw ? cw[b + k] : nullptr;
                list.agree[k] = agree ? agree[b + k] : nullptr, list.winner[k] = winner ? winner[b + k] : nullptr;
                list.rank[k] = rank ? rank[b + k] : nullptr;
            }
            p.R = run * d.R;
            const int rc = launch_list(ctx, pl, p, list, st);
            if (rc) return rc;
        } else {
            const int rc = launch_plan(ctx, pl, p, st);
            if (rc) return rc;
        }
        b += run;
    }
    return LWC_OK;
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
    const size_t tile = lwc::vote_tile_smem(n_top, n_choices);
    const size_t tile4 = lwc::vote_tile4_smem(n_top, n_choices);
    const bool vec4 = n_top >= 4 && n_top % 4 == 0 && tile4 <= 48 * 1024 &&
                      ((reinterpret_cast<uintptr_t>(dlp) | reinterpret_cast<uintptr_t>(dlf)) & 15u) == 0;
    const unsigned tiles = (unsigned)((n_rows + lwc::kVoteRows - 1) / lwc::kVoteRows);
    const size_t ring = lwc::vote_ring_smem(n_top, n_choices);
    static const bool no_ring = getenv("LWC_VOTE_IMPL") && !strcmp(getenv("LWC_VOTE_IMPL"), "tile"); // A/B knob
    if (vec4 && !no_ring && ring <= (size_t)ctx->max_smem_optin && tiles >= 2u * (unsigned)ctx->sm_count) {
        // many tiles: the persistent ring kernel (inputs by bulk async copies, several tiles in flight per CTA)
        static std::atomic<bool> configured[64];
        if (!configured[ctx->device & 63]) {
            CUDA_TRY(ctx, cudaFuncSetAttribute(lwc::vote_ring_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ctx->max_smem_optin));
            configured[ctx->device & 63] = true;
        }
        int per_sm = (int)((size_t)ctx->max_smem_optin / (ring + 1024));
        if (per_sm < 1) per_sm = 1;
        if (per_sm > 4) per_sm = 4;
        unsigned grid = (unsigned)(ctx->sm_count * per_sm);
        if (grid > tiles) grid = tiles;
        lwc::vote_ring_kernel<<<grid, lwc::kVoteRows, ring, st>>>(dlp, dlf, dV, n_rows, n_top, n_choices);
    } else if (vec4) {
        lwc::vote_tile4_kernel<<<tiles, lwc::kVoteRows, tile4, st>>>(dlp, dlf, dV, n_rows, n_top, n_choices,
                                                                    lwc::vote_tile4_pitch4(n_top));
    } else if (tile <= 48 * 1024) {
        lwc::vote_tile_kernel<<<tiles, lwc::kVoteRows, tile, st>>>(dlp, dlf, dV, n_rows, n_top, n_choices);
    } else {
        const int warps = 4;
        lwc::vote_from_logprobs_kernel<<<(n_rows + warps - 1) / warps, warps * 32, 0, st>>>(dlp, dlf, dV, n_rows, n_top,
                                                                                         n_choices);
    }
    CUDA_TRY(ctx, cudaGetLastError());
    ctx->launches++;
    if (memory == LWC_MEM_HOST) {
        CUDA_TRY(ctx, cudaMemcpyAsync(V, dV, out_bytes, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(ctx, cudaStreamSynchronize(st));
    }
    return LWC_OK;
}

int32_t lwc_stage_f64(lwc_ctx *ctx, const double *src, size_t n, float *dst) {
    if (n && (!src || !dst)) return fail(ctx, LWC_ERR_INVALID, "lwc_stage_f64: NULL argument");
    // one embedding (a few thousand values) converts inline -- this is the per-row call of the request packers, so
    // nothing but the loop runs; a whole batch is split over host threads so that the conversion keeps up with the
    // PCIe copy that follows it
    const size_t kPerThread = (size_t)1 << 20;
    if (n <= kPerThread) {
        lwc_host::convert_f64_f32(src, n, dst);
        return LWC_OK;
    }
    unsigned nt = (unsigned)((n + kPerThread - 1) / kPerThread);
    static const unsigned hw = std::thread::hardware_concurrency();
    if (nt > (hw ? hw : 1u)) nt = hw ? hw : 1u;
    if (nt > 16u) nt = 16u;
    auto convert = [src, dst](size_t a, size_t b) { lwc_host::convert_f64_f32(src + a, b - a, dst + a); };
    if (nt <= 1) {
        convert(0, n);
        return LWC_OK;
    }
    std::vector<std::thread> pool;
    pool.reserve(nt - 1);
    const size_t per = (n + nt - 1) / nt;
    for (unsigned t = 1; t < nt; ++t) pool.emplace_back(convert, std::min(n, t * per), std::min(n, (t + 1) * per));
    convert(0, std::min(n, per));
    for (auto &th : pool) th.join();
    return LWC_OK;
}

int32_t lwc_stage_f64_bf16(lwc_ctx *ctx, const double *src, size_t n, uint16_t *dst) {
    if (n && (!src || !dst)) return fail(ctx, LWC_ERR_INVALID, "lwc_stage_f64_bf16: NULL argument");
    // f64 -> fp32 (nearest even) -> bf16 (nearest even); NaN stays NaN.  The double rounding differs from a direct
    // f64 -> bf16 rounding only on exact fp32 ties, far below the format's 2^-9 resolution.
    lwc_host::convert_f64_bf16(src, n, dst);
    return LWC_OK;
}

} // extern "C"
