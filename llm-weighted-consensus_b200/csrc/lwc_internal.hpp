// lwc_internal.hpp -- shared host-side pieces of the C ABI implementation (not installed).
#pragma once

#include <cuda.h>
#include <cuda_runtime.h>
#include <nccl.h>

#include <cstdarg>
#include <cstdio>
#include <atomic>
#include <mutex>
#include <string>

#include "../../include/lwc_b200.h"

namespace lwc_host {

struct Slot {
    void *buf[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    size_t cap[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    cudaEvent_t h2d_done = nullptr; // inputs of this slot are on the device
    cudaEvent_t all_done = nullptr; // kernel + D2H of this slot finished (slot reusable)
    bool used = false;
};

} // namespace lwc_host

struct lwc_ctx {
    int device = 0;
    int sm_count = 0;
    int chunk_requests = 0;
    uint64_t flags = 0; // LWC_CTX_*
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;
    std::mutex mu;
    mutable std::mutex err_mu; // guards `err` alone: lwc_last_error_message may run while another thread's call fails
    std::string err;
    uint64_t launches = 0;
    lwc_host::Slot slots[2];
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
    // resident training table (lwc_table_load): embeddings, inverse norms, per-LLM scores, scratch
    void *table_buf[5] = {nullptr};
    size_t table_cap[5] = {0};
    long long table_rows = 0;
    int table_D = 0, table_M = 0;
    int no_list_fusion = 0;    // LWC_LIST_FUSION=0: lwc_score_batches launches batch by batch even where one launch could take a run
    int weights_in_flight = 0; // lwc_table_weights(LWC_MEM_DEVICE) has queued a kernel that writes a caller's `w`: the next
                               // lwc_score_batch must not let its input reads run ahead of the stream (LWC_CTX_INPUTS_RESIDENT)
    void *small_dev = nullptr, *small_pin = nullptr; // device block + pinned bounce of the few-request host path
    size_t small_cap = 0;
    int small_zero_copy = 4; // LWC_SMALL_ZERO_COPY: which of the few-request path's buffers the kernel reaches over PCIe (0 = none)
    void *typed_buf = nullptr; // fp32 copy of device-resident f64 / bf16 embeddings (lwc_score_batch_typed)
    size_t typed_cap = 0;
    void *tc_flags = nullptr;  // per-request fix-up marks of the fp16-piece tcgen05 route (score_tc16.cuh)
    size_t tc_flags_cap = 0;
    int small_pipe = 3;        // LWC_SMALL_PIPE=0|3|4: shared-memory load pipeline stages of the 32-row fp16-piece kernel (0: register loads)
    int small_x3_tf32 = 0;     // LWC_SMALL_X3=tf32: the one-CTA-per-request fp32-grade kernel on tf32 pieces
    int stream_parks = 0;      // LWC_STREAM_PARKS=1|2: park buffers of the ring kernel (0: by request size)
    int stream_no_group = 0;   // LWC_STREAM_GROUP=0: pools of <= 4 candidates one to an 8-row tile of the ring kernel
    int stream_no_half = 0;    // LWC_STREAM_HALF=0: pools of <= 8 candidates on the full 16-row tile of the ring kernel
    int tc16_a_tmem = 1;       // LWC_TC16_A=tmem|smem: fp32-grade tcgen05 route with the A operand in tensor memory (score_tc16a.cuh)
    int tc_no_pair = 0;        // LWC_TC_PAIR=0: pools of <= 64 candidates one to a tile (score_tc16_kernel<false>)
    int tc_x3_tf32 = 0;        // LWC_TC_X3=tf32: fp32-grade large pools on tf32 pieces (score_tc_kernel<true>) only
    int tc_ko = 0;     // operand slabs per stage of the tcgen05 kernel (LWC_TC_KO)
    int tc_stages = 0; // its ring depth (LWC_TC_STAGES)
};

namespace lwc_host {

// Records the failure text (ResponseError.message, src/error.rs:11) and returns `code`.
int fail(lwc_ctx *ctx, int code, const char *fmt, ...);

#define CUDA_TRY(ctx, expr)                                                                                         \
    do {                                                                                                            \
        cudaError_t _e = (expr);                                                                                    \
        if (_e != cudaSuccess)                                                                                      \
            return lwc_host::fail((ctx), LWC_ERR_INTERNAL, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                                  __FILE__, __LINE__);                                                              \
    } while (0)

// cuTensorMapEncodeTiled is a driver entry point; fetched through the runtime so that the library does
// not link libcuda directly.
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled_fn();

// host_convert.cpp: vectorised element conversion of embedding rows
void convert_f64_f32(const double *src, size_t n, float *dst);
void convert_f64_bf16(const double *src, size_t n, uint16_t *dst);

} // namespace lwc_host
