// lwc_table_api.cu -- C ABI of the training-table weight lookup (BUILD-DEFINED, see table.cuh).
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "lwc_internal.hpp"
#include "table.cuh"

using lwc_host::fail;

namespace {

int table_ensure(lwc_ctx *ctx, int which, size_t bytes) {
    if (bytes <= ctx->table_cap[which]) return LWC_OK;
    if (ctx->table_buf[which]) CUDA_TRY(ctx, cudaFree(ctx->table_buf[which]));
    ctx->table_buf[which] = nullptr;
    ctx->table_cap[which] = 0;
    CUDA_TRY(ctx, cudaMalloc(&ctx->table_buf[which], bytes + 256));
    ctx->table_cap[which] = bytes + 256;
    return LWC_OK;
}

// Host -> device upload of a large, possibly pageable array.  cudaMemcpy from pageable memory goes through the
// driver's single staging buffer (measured: 11 GB/s, 0.55 s for a 6.1 GB table); here host threads copy 32 MiB
// pieces into two pinned buffers while the previous piece crosses PCIe.  Memory that is already pinned
// (lwc_host_alloc, cudaHostRegister) is copied directly.
int upload_staged(lwc_ctx *ctx, void *dst, const void *src, size_t bytes) {
    cudaPointerAttributes at{};
    if (cudaPointerGetAttributes(&at, src) == cudaSuccess && at.type == cudaMemoryTypeHost) {
        CUDA_TRY(ctx, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
        return LWC_OK;
    }
    cudaGetLastError();
    const size_t piece = (size_t)32 << 20;
    if (bytes <= piece) {
        CUDA_TRY(ctx, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
        CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream)); // pageable source: the caller may reuse it after the call
        return LWC_OK;
    }
    void *pin[2] = {nullptr, nullptr};
    cudaEvent_t done[2] = {nullptr, nullptr};
    int rc = LWC_OK;
    for (int i = 0; i < 2 && !rc; ++i)
        if (cudaHostAlloc(&pin[i], piece, cudaHostAllocDefault) != cudaSuccess ||
            cudaEventCreateWithFlags(&done[i], cudaEventDisableTiming) != cudaSuccess)
            rc = fail(ctx, LWC_ERR_INTERNAL, "lwc_table_load: pinned staging allocation failed: %s", cudaGetErrorString(cudaGetLastError()));
    const unsigned hw = std::thread::hardware_concurrency();
    const unsigned nt = hw >= 8 ? 4 : (hw >= 2 ? 2 : 1);
    size_t off = 0;
    for (int k = 0; !rc && off < bytes; ++k, off += piece) {
        const int b = k & 1;
        const size_t n = bytes - off < piece ? bytes - off : piece;
        if (k >= 2 && cudaEventSynchronize(done[b]) != cudaSuccess) rc = fail(ctx, LWC_ERR_INTERNAL, "lwc_table_load: staging wait failed");
        if (rc) break;
        std::vector<std::thread> th;
        const size_t per = (n + nt - 1) / nt;
        const char *s = static_cast<const char *>(src) + off;
        char *d = static_cast<char *>(pin[b]);
        for (unsigned t = 1; t < nt; ++t)
            if (t * per < n) th.emplace_back([=] { memcpy(d + t * per, s + t * per, std::min(per, n - t * per)); });
        memcpy(d, s, std::min(per, n));
        for (auto &x : th) x.join();
        if (cudaMemcpyAsync(static_cast<char *>(dst) + off, pin[b], n, cudaMemcpyHostToDevice, ctx->stream) != cudaSuccess ||
            cudaEventRecord(done[b], ctx->stream) != cudaSuccess)
            rc = fail(ctx, LWC_ERR_INTERNAL, "lwc_table_load: staged copy failed: %s", cudaGetErrorString(cudaGetLastError()));
    }
    if (cudaStreamSynchronize(ctx->stream) != cudaSuccess && !rc) rc = fail(ctx, LWC_ERR_INTERNAL, "lwc_table_load: upload failed");
    for (int i = 0; i < 2; ++i) {
        if (done[i]) cudaEventDestroy(done[i]);
        if (pin[i]) cudaFreeHost(pin[i]);
    }
    return rc;
}

} // namespace

extern "C" {

int32_t lwc_table_load(lwc_ctx *ctx, const float *T, const float *Q, int64_t rows, int32_t D, int32_t M) {
    if (!ctx) return LWC_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (!T || !Q || rows < 1 || rows > 0x7fffffffLL || D < 1 || M < 1 || M > LWC_MAX_VOTERS)
        return fail(ctx, LWC_ERR_INVALID, "lwc_table_load: invalid table (rows=%lld D=%d M=%d)", (long long)rows, D, M);
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    ctx->table_rows = 0; // no table until this load has succeeded
    int rc;
    if ((rc = table_ensure(ctx, 0, (size_t)rows * D * 4))) return rc;
    if ((rc = table_ensure(ctx, 1, (size_t)rows * 4))) return rc;
    if ((rc = table_ensure(ctx, 2, (size_t)rows * M * 4))) return rc;
    if ((rc = upload_staged(ctx, ctx->table_buf[0], T, (size_t)rows * D * 4))) return rc;
    if ((rc = upload_staged(ctx, ctx->table_buf[2], Q, (size_t)rows * M * 4))) return rc;
    const int warps = 8;
    lwc::table_norm_kernel<<<(unsigned)((rows + warps - 1) / warps), warps * 32, 0, ctx->stream>>>(
        static_cast<const float *>(ctx->table_buf[0]), static_cast<float *>(ctx->table_buf[1]), rows, D);
    CUDA_TRY(ctx, cudaGetLastError());
    ctx->launches++;
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->table_rows = rows;
    ctx->table_D = D;
    ctx->table_M = M;
    return LWC_OK;
}

int32_t lwc_table_weights(lwc_ctx *ctx, const float *q, int32_t n_prompts, int32_t top, const float *base,
                          const float *minw, const float *maxw, float *weights, int32_t *top_idx, float *top_sim,
                          int32_t memory) {
    if (!ctx) return LWC_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (ctx->table_rows == 0) return fail(ctx, LWC_ERR_INVALID, "lwc_table_weights: no table loaded (lwc_table_load)");
    // `top` >= 1 as WeightTrainingTable::validate demands (src/score/model/mod.rs:292-297)
    if (memory != LWC_MEM_HOST && memory != LWC_MEM_DEVICE)
        return fail(ctx, LWC_ERR_INVALID, "lwc_table_weights: memory must be LWC_MEM_HOST or LWC_MEM_DEVICE");
    const bool dev = memory == LWC_MEM_DEVICE;
    if (!q || !base || !minw || !maxw || !weights || n_prompts < 0 || top < 1 || top > lwc::kTableMaxTop)
        return fail(ctx, LWC_ERR_INVALID, "lwc_table_weights: invalid argument (prompts=%d top=%d, top must be 1..%d)",
                    n_prompts, top, lwc::kTableMaxTop);
    const int M = ctx->table_M, D = ctx->table_D;
    const long long Tn = ctx->table_rows;
    for (int m = 0; m < M; ++m) // WeightTrainingTable::validate, src/score/llm/mod.rs:671-687
        if (!(minw[m] > 0.f) || !(base[m] >= minw[m]) || !(base[m] <= maxw[m]))
            return fail(ctx, LWC_ERR_INVALID, "lwc_table_weights: LLM %d needs 0 < min <= base <= max (got %g, %g, %g)", m,
                        minw[m], base[m], maxw[m]);
    if (n_prompts == 0) return LWC_OK;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const int QB = lwc::kTableQB;
    int rc;
    // scratch: [3] scores [QB][Tn] | per-chunk candidates (value, row) [QB][chunks][top]; [4] q | idx | sim | limits | weights for one pass
    const long long chunks = (Tn + lwc::kTopkChunk - 1) / lwc::kTopkChunk;
    const size_t o_cv = ((size_t)QB * Tn * 4 + 255) & ~(size_t)255, cand_bytes = (size_t)QB * chunks * top * 4;
    if ((rc = table_ensure(ctx, 3, o_cv + 4 * cand_bytes))) return rc; // two (value, row) list buffers
    char *s3 = static_cast<char *>(ctx->table_buf[3]);
    const size_t o_q = 0, o_idx = (size_t)QB * D * 4, o_sim = o_idx + (size_t)QB * top * 4, o_lim = o_sim + (size_t)QB * top * 4,
                 o_w = o_lim + (size_t)3 * M * 4, total = o_w + (size_t)QB * M * 4;
    if ((rc = table_ensure(ctx, 4, total))) return rc;
    char *sb = static_cast<char *>(ctx->table_buf[4]);
    CUDA_TRY(ctx, cudaMemcpyAsync(sb + o_lim, base, (size_t)M * 4, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(sb + o_lim + (size_t)M * 4, minw, (size_t)M * 4, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(sb + o_lim + (size_t)2 * M * 4, maxw, (size_t)M * 4, cudaMemcpyHostToDevice, st));
    // Route: the bulk-copy ring when the rows are 16-byte multiples and a ring of RS-row slots fits beside the
    // prompt tile; else the direct-load kernel.  LWC_TABLE_IMPL=direct forces the latter (tests).
    int RS = 0;
    const char *impl = getenv("LWC_TABLE_IMPL");
    if (D % 4 == 0 && !(impl && !strcmp(impl, "direct")))
        for (int rs = 4; rs >= 1 && !RS; rs >>= 1)
            if (lwc::table_ring_smem(rs, D, QB) <= (size_t)ctx->max_smem_optin) RS = rs;
    const size_t smem = RS ? lwc::table_ring_smem(RS, D, QB) : ((size_t)QB * D + QB) * 4;
    if (smem > (size_t)ctx->max_smem_optin)
        return fail(ctx, LWC_ERR_INVALID, "lwc_table_weights: D=%d too large for the prompt tile in shared memory", D);
    const void *fn = RS == 4   ? (const void *)lwc::table_scores_ring_kernel<4>
                     : RS == 2 ? (const void *)lwc::table_scores_ring_kernel<2>
                     : RS == 1 ? (const void *)lwc::table_scores_ring_kernel<1>
                               : (const void *)lwc::table_scores_kernel;
    if (smem > 48 * 1024) CUDA_TRY(ctx, cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    for (int b0 = 0; b0 < n_prompts; b0 += QB) {
        const int nq = n_prompts - b0 < QB ? n_prompts - b0 : QB;
        if (!dev) CUDA_TRY(ctx, cudaMemcpyAsync(sb + o_q, q + (size_t)b0 * D, (size_t)nq * D * 4, cudaMemcpyHostToDevice, st));
        const float *Td = static_cast<const float *>(ctx->table_buf[0]), *rn = static_cast<const float *>(ctx->table_buf[1]);
        const float *qd = dev ? q + (size_t)b0 * D : reinterpret_cast<const float *>(sb + o_q);
        // device memory: results go straight to the caller's buffers; host memory: through the pass's scratch
        int32_t *idx_o = dev && top_idx ? top_idx + (size_t)b0 * top : reinterpret_cast<int32_t *>(sb + o_idx);
        float *sim_o = dev && top_sim ? top_sim + (size_t)b0 * top : reinterpret_cast<float *>(sb + o_sim);
        float *w_o = dev ? weights + (size_t)b0 * M : reinterpret_cast<float *>(sb + o_w);
        float *scd = reinterpret_cast<float *>(s3);
        if (RS) {
            const long long groups = (Tn + RS - 1) / RS;
            const unsigned grid = (unsigned)(groups < ctx->sm_count ? groups : ctx->sm_count);
            const unsigned thr = lwc::kTableRingThreads;
            const size_t sm = lwc::table_ring_smem(RS, D, nq);
            if (RS == 4) lwc::table_scores_ring_kernel<4><<<grid, thr, sm, st>>>(Td, rn, qd, scd, Tn, D, nq);
            else if (RS == 2) lwc::table_scores_ring_kernel<2><<<grid, thr, sm, st>>>(Td, rn, qd, scd, Tn, D, nq);
            else lwc::table_scores_ring_kernel<1><<<grid, thr, sm, st>>>(Td, rn, qd, scd, Tn, D, nq);
        } else {
            long long blocks = (Tn + 8 * lwc::kTableRows - 1) / (8 * lwc::kTableRows);
            if (blocks > 8LL * ctx->sm_count) blocks = 8LL * ctx->sm_count;
            lwc::table_scores_kernel<<<(unsigned)blocks, 256, ((size_t)nq * D + nq) * 4, st>>>(Td, rn, qd, scd, Tn, D, nq);
        }
        { // top-k reduction tree: scores -> per-chunk lists -> ... -> idx / sim
            const float *iv = scd;
            const int32_t *ii = nullptr;
            long long n = Tn;
            int level = 0;
            for (;; ++level) {
                const long long ch = (n + lwc::kTopkChunk - 1) / lwc::kTopkChunk;
                const bool last = ch == 1;
                float *ov = reinterpret_cast<float *>(s3 + o_cv + (size_t)(level & 1) * 2 * cand_bytes);
                int32_t *oi = reinterpret_cast<int32_t *>(s3 + o_cv + (size_t)(level & 1) * 2 * cand_bytes + cand_bytes);
                int32_t *fi = last ? idx_o : nullptr;
                float *fs = last ? sim_o : nullptr;
                const dim3 grid((unsigned)ch, nq);
                if (ii) lwc::table_topk_kernel<true><<<grid, lwc::kTopkThreads, 0, st>>>(iv, ii, n, top, ov, oi, fi, fs);
                else lwc::table_topk_kernel<false><<<grid, lwc::kTopkThreads, 0, st>>>(iv, ii, n, top, ov, oi, fi, fs);
                ctx->launches++;
                if (last) break;
                iv = ov, ii = oi, n = ch * top;
            }
        }
        lwc::table_weights_kernel<<<(nq + 3) / 4, 128, 0, st>>>(
            idx_o, sim_o, static_cast<const float *>(ctx->table_buf[2]), reinterpret_cast<const float *>(sb + o_lim),
            reinterpret_cast<const float *>(sb + o_lim + (size_t)M * 4), reinterpret_cast<const float *>(sb + o_lim + (size_t)2 * M * 4),
            w_o, nq, top, M);
        CUDA_TRY(ctx, cudaGetLastError());
        ctx->launches += 2;
        if (!dev) { // stream order protects the scratch: these copies precede the next pass's upload and kernels
            CUDA_TRY(ctx, cudaMemcpyAsync(weights + (size_t)b0 * M, w_o, (size_t)nq * M * 4, cudaMemcpyDeviceToHost, st));
            if (top_idx) CUDA_TRY(ctx, cudaMemcpyAsync(top_idx + (size_t)b0 * top, idx_o, (size_t)nq * top * 4, cudaMemcpyDeviceToHost, st));
            if (top_sim) CUDA_TRY(ctx, cudaMemcpyAsync(top_sim + (size_t)b0 * top, sim_o, (size_t)nq * top * 4, cudaMemcpyDeviceToHost, st));
        }
    }
    if (!dev) CUDA_TRY(ctx, cudaStreamSynchronize(st)); // host results are valid on return; device results in stream order
    if (dev) ctx->weights_in_flight = 1; // see lwc_score_batch: the next launch keeps full stream order
    return LWC_OK;
}

int32_t lwc_table_merge(int32_t n_parts, int32_t n_prompts, int32_t top, int32_t M, const int64_t *row_offset,
                        const int32_t *idx, const float *sim, const float *q_rows, const float *base, const float *minw,
                        const float *maxw, float *weights, int64_t *top_row, float *top_sim) {
    // Host only: n_parts * top candidates per prompt.  Same arithmetic, in the same order, as table_weights_kernel
    // (fmaf per neighbour in merged order; the mean of rho through the same butterfly as the kernel's warp_sum), so a
    // table split over ranks gives the bits of the unsplit table.
    if (n_parts < 1 || n_prompts < 0 || top < 1 || top > lwc::kTableMaxTop || M < 1 || M > LWC_MAX_VOTERS || !row_offset ||
        !idx || !sim || !q_rows || !base || !minw || !maxw || !weights)
        return fail(nullptr, LWC_ERR_INVALID, "lwc_table_merge: invalid argument (parts=%d prompts=%d top=%d M=%d)", n_parts,
                    n_prompts, top, M);
    for (int m = 0; m < M; ++m)
        if (!(minw[m] > 0.f) || !(base[m] >= minw[m]) || !(base[m] <= maxw[m]))
            return fail(nullptr, LWC_ERR_INVALID, "lwc_table_merge: LLM %d needs 0 < min <= base <= max", m);
    struct Cand {
        float sim;
        int64_t row;
        const float *q;
    };
    std::vector<Cand> cand;
    std::vector<float> rho(M);
    for (int p = 0; p < n_prompts; ++p) {
        cand.clear();
        for (int r = 0; r < n_parts; ++r)
            for (int k = 0; k < top; ++k) {
                const size_t at = ((size_t)r * n_prompts + p) * top + k;
                if (idx[at] >= 0) cand.push_back({sim[at], row_offset[r] + idx[at], q_rows + at * M});
            }
        std::sort(cand.begin(), cand.end(),
                  [](const Cand &a, const Cand &b) { return a.sim > b.sim || (a.sim == b.sim && a.row < b.row); });
        const int n = (int)cand.size() < top ? (int)cand.size() : top;
        float ssum = 0.f;
        for (int k = 0; k < n; ++k) ssum += fmaxf(cand[k].sim, 0.f);
        float lane_part[32] = {0.f};
        for (int m = 0; m < M; ++m) {
            float r = 0.f;
            if (ssum > 0.f) {
                for (int k = 0; k < n; ++k) r = fmaf(fmaxf(cand[k].sim, 0.f), cand[k].q[m], r);
                r /= ssum;
            }
            rho[m] = r;
        }
        for (int lane = 0; lane < 32; ++lane) // the kernel: lane l adds rho[l], rho[32 + l], ... in that order
            for (int j = 0; j * 32 + lane < M; ++j) lane_part[lane] += rho[j * 32 + lane];
        for (int o = 16; o > 0; o >>= 1) { // warp_sum's xor butterfly
            float next[32];
            for (int l = 0; l < 32; ++l) next[l] = lane_part[l] + lane_part[l ^ o];
            std::memcpy(lane_part, next, sizeof(next));
        }
        const float mean = lane_part[0] / (float)M;
        for (int m = 0; m < M; ++m) {
            float wv = base[m];
            if (ssum > 0.f && mean > 0.f) wv = base[m] * (rho[m] / mean);
            weights[(size_t)p * M + m] = fminf(fmaxf(wv, minw[m]), maxw[m]);
        }
        for (int k = 0; k < top; ++k) {
            if (top_row) top_row[(size_t)p * top + k] = k < n ? cand[k].row : -1;
            if (top_sim) top_sim[(size_t)p * top + k] = k < n ? cand[k].sim : 0.f;
        }
    }
    return LWC_OK;
}

} // extern "C"
