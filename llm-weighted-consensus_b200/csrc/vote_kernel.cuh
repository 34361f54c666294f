// vote_kernel.cuh -- numeric tail of get_vote (src/score/completions/client.rs:1765-1791) on device.
// For each voter: the top-logprob alternatives whose character is a sibling leaf contribute
// p = exp(logprob) to vote[leaf]; the row is then divided by the probability sum
// (client.rs:1789-1791).  The string/regex part of get_vote (client.rs:1669-1762) stays on the host
// and produces (logprob, leaf) pairs; leaf < 0 marks an alternative that is not a sibling leaf.
// One warp per vote row, rows written straight into the [R][M][N] vote buffer the scorer reads.
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

} // namespace lwc
