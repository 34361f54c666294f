// vote_kernel.cuh -- numeric tail of get_vote (src/score/completions/client.rs:1765-1791) on device.
// For each voter: the top-logprob alternatives whose character is a sibling leaf contribute
// p = exp(logprob) to vote[leaf]; the row is then divided by the probability sum
// (client.rs:1789-1791).  The string/regex part of get_vote (client.rs:1669-1762) stays on the host
// and produces (logprob, leaf) pairs; leaf < 0 marks an alternative that is not a sibling leaf.
// Rows are written straight into the [R][M][N] vote buffer the scorer reads.
//
// vote_tile_kernel (the route for ordinary shapes): HBM-bound, 8 bytes in per alternative and 4 out per choice.
// A CTA stages kVoteRows rows through shared memory with coalesced 128-bit-wide traffic on both sides; each
// thread then owns one row: one expf per alternative, accumulated into its row of the output tile in
// alternative order (deterministic), then the row is divided by its probability sum.
// vote_from_logprobs_kernel (one warp per row) remains for shapes whose tile does not fit 48 KB.
#pragma once

#include "score_common.cuh"

namespace lwc {

__global__ void __launch_bounds__(128) vote_from_logprobs_kernel(const float *__restrict__ logprob,
                                                                  const int32_t *__restrict__ leaf, float *__restrict__ V,
                                                                  int n_rows, int n_top, int n_choices) {
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= n_rows) return;
    const float *lp = logprob + (size_t)row * n_top;
    const int32_t *lf = leaf + (size_t)row * n_top;
    float *out = V + (size_t)row * n_choices;
    // probability sum over the admissible alternatives (lanes stride the <= 20 alternatives)
    float s = 0.f;
    for (int k = lane; k < n_top; k += 32) {
        const int l = lf[k];
        const float x = lp[k];
        if (l >= 0 && l < n_choices && !isnan(x)) s += expf(x);
    }
    s = warp_sum(s);
    if (!(s > 0.f)) { // the reference treats this as unreachable (client.rs:1786-1788): mark the voter absent
        for (int j = lane; j < n_choices; j += 32) out[j] = __int_as_float(0x7fc00000);
        return;
    }
    // each lane owns choices j = lane, lane+32, ...: accumulate in alternative order (deterministic)
    for (int j = lane; j < n_choices; j += 32) {
        float v = 0.f;
        for (int k = 0; k < n_top; ++k) {
            const float x = lp[k];
            if (lf[k] == j && !isnan(x)) v += expf(x);
        }
        out[j] = v / s;
    }
}

constexpr int kVoteRows = 128; // rows per CTA of vote_tile_kernel (= threads)

inline size_t vote_tile_smem(int n_top, int n_choices) {
    return (size_t)kVoteRows * ((size_t)n_top * 8 + (size_t)(n_choices + 1) * 4);
}

__global__ void __launch_bounds__(kVoteRows) vote_tile_kernel(const float *__restrict__ logprob,
                                                              const int32_t *__restrict__ leaf, float *__restrict__ V,
                                                              int n_rows, int n_top, int n_choices) {
    extern __shared__ __align__(16) unsigned char vsm[];
    float *lp_s = reinterpret_cast<float *>(vsm);                                   // [rows][n_top]
    int32_t *lf_s = reinterpret_cast<int32_t *>(lp_s + (size_t)kVoteRows * n_top);  // [rows][n_top]
    float *out_s = reinterpret_cast<float *>(lf_s + (size_t)kVoteRows * n_top);     // [rows][n_choices + 1]
    const int pitch = n_choices + 1; // odd or even, consecutive rows start in different banks
    const long long row0 = (long long)blockIdx.x * kVoteRows;
    const int rows = (int)(n_rows - row0 < kVoteRows ? n_rows - row0 : kVoteRows);
    const int n_in = rows * n_top;
    const float *lp = logprob + (size_t)row0 * n_top;
    const int32_t *lf = leaf + (size_t)row0 * n_top;
    for (int i = threadIdx.x; i < n_in; i += kVoteRows) {
        lp_s[i] = lp[i];
        lf_s[i] = lf[i];
    }
    for (int i = threadIdx.x; i < rows * pitch; i += kVoteRows) out_s[i] = 0.f;
    __syncthreads();
    if ((int)threadIdx.x < rows) {
        const float *x = lp_s + (size_t)threadIdx.x * n_top;
        const int32_t *l = lf_s + (size_t)threadIdx.x * n_top;
        float *o = out_s + (size_t)threadIdx.x * pitch;
        float s = 0.f;
        for (int k = 0; k < n_top; ++k) {
            const int j = l[k];
            const float v = x[k];
            if (j >= 0 && j < n_choices && !isnan(v)) {
                const float pk = expf(v);
                o[j] += pk; // client.rs:1765-1784, in alternative order
                s += pk;
            }
        }
        // normalise (client.rs:1789-1791) and write the row out from this thread: consecutive threads own
        // consecutive rows, so a warp's stores cover one contiguous span (no index arithmetic per element)
        const bool ok = s > 0.f; // else unreachable!() in the reference (client.rs:1786-1788): mark the voter absent
        const float inv = ok ? 1.0f / s : 0.f;
        const float nanv = __int_as_float(0x7fc00000);
        float *out = V + (size_t)(row0 + threadIdx.x) * n_choices;
        if ((n_choices & 3) == 0 && (reinterpret_cast<uintptr_t>(V) & 15u) == 0) {
            for (int j = 0; j < n_choices; j += 4) {
                float4 r4;
                r4.x = ok ? o[j] * inv : nanv;
                r4.y = ok ? o[j + 1] * inv : nanv;
                r4.z = ok ? o[j + 2] * inv : nanv;
                r4.w = ok ? o[j + 3] * inv : nanv;
                *reinterpret_cast<float4 *>(out + j) = r4;
            }
        } else {
            for (int j = 0; j < n_choices; ++j) out[j] = ok ? o[j] * inv : nanv;
        }
    }
}

} // namespace lwc
