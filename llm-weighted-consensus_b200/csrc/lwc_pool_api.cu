// lwc_pool_api.cu -- C ABI of the single-pool path (BASELINE.json config 5): lwc_pool_consensus, the NCCL
// communicator hooks, the fused peer-to-peer gather set-up, and lwc_pool_consensus_static.
#include <dlfcn.h>

#include <cstdlib>
#include <cstring>

#include "lwc_internal.hpp"
#include "pool.cuh"

// Working memory of lwc_pool_consensus_static: part of the device image, so no allocation call is ever
// needed for it.  In host code this name is only a handle for cudaGetSymbolAddress -- it is NOT a
// device address and must never be handed to a kernel as one.
static __device__ __align__(256) unsigned char g_static_scratch[LWC_STATIC_SCRATCH_BYTES];

using lwc_host::fail;
using lwc_host::EncodeTiledFn;
using lwc_host::encode_tiled_fn;

namespace {

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

// Drops the fused-gather state: this rank's mappings of its peers' buffers are closed and later calls use the
// NCCL all-gather until the caller exports and attaches again.
void p2p_detach(lwc_ctx *ctx) {
    for (auto &m : ctx->p2p_mapped)
        for (auto &q : m)
            if (q) {
                cudaIpcCloseMemHandle(q);
                q = nullptr;
            }
    ctx->p2p_ready = false;
    ctx->p2p_n_own = 0;
    ctx->p2p_D = 0;
}

int pool_ensure(lwc_ctx *ctx, int which, size_t bytes) {
    if (bytes <= ctx->pool_cap[which]) return LWC_OK;
    // Slots 0 and 1 are the buffers lwc_pool_p2p_export handed to the peers.  Growing one frees the memory their
    // mappings point at, so the fused route ends here -- on every rank at the same call, because the ranks of a
    // communicator call lwc_pool_consensus with the same shape (the peers have long finished reading: the previous
    // call's closing collectives follow every rank's tensor pass).
    if (which < 2 && ctx->pool_buf[which] && ctx->p2p_n_own != 0) p2p_detach(ctx);
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
    static NcclApi no_nccl;                              // never called: every use below is under nranks > 1
    NcclApi &nc = nranks > 1 ? nccl_api() : no_nccl;     // a single-GPU call must not dlopen NCCL
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
        static std::atomic<bool> configured[64]; // zero-initialised; set once per device (contexts on other threads may race here)
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
    NcclApi &nc = nccl_api();
    // device block: peer_ehat[64], peer_t[64], shard_done[64], one agreement word
    const size_t flag_off = 2 * 64 * sizeof(void *) + 64 * sizeof(unsigned);
    if (!ctx->p2p_dev) CUDA_TRY(ctx, cudaMalloc(&ctx->p2p_dev, flag_off + 64));
    const char *hp = static_cast<const char *>(all_handles);
    void *host_ptrs[2 * 64] = {nullptr};
    int local_ok = ctx->no_p2p ? 0 : 1, local_rc = LWC_OK;
    for (int r = 0; r < ctx->nranks && local_ok; ++r) {
        for (int k = 0; k < 2 && local_ok; ++k) {
            void *ptr = ctx->pool_buf[k];
            if (r != ctx->rank) {
                if (ctx->p2p_mapped[r][k]) {
                    cudaIpcCloseMemHandle(ctx->p2p_mapped[r][k]);
                    ctx->p2p_mapped[r][k] = nullptr;
                }
                cudaIpcMemHandle_t h;
                memcpy(&h, hp + (size_t)r * 128 + k * 64, 64);
                const cudaError_t e = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
                if (e != cudaSuccess) {
                    cudaGetLastError();
                    local_ok = 0;
                    local_rc = fail(ctx, LWC_ERR_INTERNAL, "lwc_pool_p2p_attach: cudaIpcOpenMemHandle(rank %d) failed: %s", r,
                                    cudaGetErrorString(e));
                    break;
                }
                ctx->p2p_mapped[r][k] = ptr;
            }
            host_ptrs[k * 64 + r] = ptr;
        }
    }
    // The route is chosen per call from this state, and the two routes issue different collectives: every rank must
    // hold the same answer.  So the ranks agree here (min over "my mappings are open and LWC_POOL_NO_P2P is unset");
    // a rank whose attach failed still takes part, then reports its error.
    int *flag_d = reinterpret_cast<int *>(static_cast<char *>(ctx->p2p_dev) + flag_off);
    int agreed = 0;
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->p2p_dev, host_ptrs, sizeof host_ptrs, cudaMemcpyHostToDevice, ctx->stream));
    CUDA_TRY(ctx, cudaMemcpyAsync(flag_d, &local_ok, sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    NCCL_TRY(ctx, nc.AllReduce(flag_d, flag_d, 1, ncclInt32, ncclMin, ctx->comm, ctx->stream));
    CUDA_TRY(ctx, cudaMemcpyAsync(&agreed, flag_d, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->p2p_ready = agreed != 0;
    if (!ctx->p2p_ready) { // some rank cannot use the fused route: nobody does
        const long long n_own = ctx->p2p_n_own;
        const int D = ctx->p2p_D;
        p2p_detach(ctx);
        ctx->p2p_n_own = n_own, ctx->p2p_D = D; // the export stays valid for another attach
    }
    return local_rc;
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

