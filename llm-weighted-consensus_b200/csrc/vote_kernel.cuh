// vote_kernel.cuh -- numeric tail of get_vote (src/score/completions/client.rs:1765-1791) on device.
// For each voter: the top-logprob alternatives whose character is a sibling leaf contribute
// p = exp(logprob) to vote[leaf]; the row is then divided by the probability sum
// (client.rs:1789-1791).  The string/regex part of get_vote (client.rs:1669-1762) stays on the host
// and produces (logprob, leaf) pairs; leaf < 0 marks an alternative that is not a sibling leaf.
// Rows are written straight into the [R][M][N] vote buffer the scorer reads.
//
// vote_tile_kernel (the route for ordinary shapes): HBM-bound, 8 bytes in per alternative and 4 out per choice.
// A CTA stages kVoteRows rows through shared memory with coalesced traffic on both sides (tiles transposed so
// that every phase is bank-conflict free); each thread then owns one row: one expf per alternative, accumulated
// into its row of the output tile in alternative order (deterministic), then the row is divided by its
// probability sum on the way out.
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

constexpr int kVoteRows = 128;            // rows per CTA of vote_tile_kernel (= threads)
constexpr int kVotePitch = kVoteRows + 1; // tiles are stored TRANSPOSED, [column][row] with an odd pitch

inline size_t vote_tile_smem(int n_top, int n_choices) {
    return (size_t)kVotePitch * ((size_t)n_top * 8 + (size_t)n_choices * 4) + (size_t)kVoteRows * 4;
}

// Shared-memory layout: every tile is [column][row] with pitch 129 words, so that
//   * the compute phase (thread = row, all threads at the same column k, or at data-dependent columns j of
//     the output tile) touches bank (column + row) mod 32 -- conflict-free whatever j each thread picks;
//   * the staging phases (consecutive threads = consecutive global words = consecutive columns of one or
//     two rows) are conflict-free (inputs) or two-way (outputs) as well.
// Round 1 kept the tiles row-major: pitch 20 gave 4-way conflicts on every input read and random-bank
// read-modify-writes on the output row, and the kernel sat at 0.56 of HBM on the shared-memory pipe.
__global__ void __launch_bounds__(kVoteRows) vote_tile_kernel(const float *__restrict__ logprob,
                                                              const int32_t *__restrict__ leaf, float *__restrict__ V,
                                                              int n_rows, int n_top, int n_choices) {
    extern __shared__ __align__(16) unsigned char vsm[];
    float *lp_s = reinterpret_cast<float *>(vsm);                                   // [n_top][129]
    int32_t *lf_s = reinterpret_cast<int32_t *>(lp_s + (size_t)kVotePitch * n_top);  // [n_top][129]
    float *out_s = reinterpret_cast<float *>(lf_s + (size_t)kVotePitch * n_top);     // [n_choices][129]
    float *inv_s = out_s + (size_t)kVotePitch * n_choices;                           // [128] 1 / probability sum, or NaN
    const long long row0 = (long long)blockIdx.x * kVoteRows;
    const int rows = (int)(n_rows - row0 < kVoteRows ? n_rows - row0 : kVoteRows);
    const int tid = threadIdx.x;

    // ---- stage in: coalesced global reads, word i of the CTA's span -> (row i / n_top, column i % n_top)
    if (n_top > 0) {
        const int n_in = rows * n_top;
        const float *lp = logprob + (size_t)row0 * n_top;
        const int32_t *lf = leaf + (size_t)row0 * n_top;
        int r = tid / n_top, k = tid - r * n_top;
        const int dr = kVoteRows / n_top, dk = kVoteRows - dr * n_top;
#pragma unroll 4
        for (int i = tid; i < n_in; i += kVoteRows) {
            lp_s[k * kVotePitch + r] = __ldg(lp + i);
            lf_s[k * kVotePitch + r] = __ldg(lf + i);
            r += dr;
            k += dk;
            if (k >= n_top) {
                k -= n_top;
                ++r;
            }
        }
    }
    for (int j = 0; j < n_choices; ++j) out_s[j * kVotePitch + tid] = 0.f; // own row of the output tile
    __syncthreads();

    // ---- thread = row: one exp per admissible alternative, accumulated in alternative order (client.rs:1765-1784)
    if (tid < rows) {
        float s = 0.f;
#pragma unroll 4
        for (int k = 0; k < n_top; ++k) {
            const int j = lf_s[k * kVotePitch + tid];
            const float v = lp_s[k * kVotePitch + tid];
            if (j >= 0 && j < n_choices && !isnan(v)) {
                const float pk = expf(v);
                out_s[j * kVotePitch + tid] += pk;
                s += pk;
            }
        }
        // client.rs:1789-1791; a zero sum is unreachable!() in the reference (1786-1788): the voter is marked absent
        inv_s[tid] = s > 0.f ? 1.0f / s : __int_as_float(0x7fc00000);
    }
    __syncthreads();

    // ---- stage out: coalesced global writes, word i -> (row i / n_choices, column i % n_choices)
    {
        const int n_out = rows * n_choices;
        float *out = V + (size_t)row0 * n_choices;
        int r = tid / n_choices, j = tid - r * n_choices;
        const int dr = kVoteRows / n_choices, dj = kVoteRows - dr * n_choices;
#pragma unroll 4
        for (int i = tid; i < n_out; i += kVoteRows) {
            const float inv = inv_s[r];
            out[i] = isnan(inv) ? inv : out_s[j * kVotePitch + r] * inv;
            r += dr;
            j += dj;
            if (j >= n_choices) {
                j -= n_choices;
                ++r;
            }
        }
    }
}

} // namespace lwc
