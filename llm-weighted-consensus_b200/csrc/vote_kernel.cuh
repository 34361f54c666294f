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

#include "score_stream.cuh" // mbarrier / bulk-copy helpers

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
constexpr int kVotePitch = kVoteRows + 1; // input tiles are stored TRANSPOSED, [column][row], with an odd pitch

inline size_t vote_tile_smem(int n_top, int n_choices) {
    return (size_t)kVotePitch * (size_t)n_top * 8 + (size_t)kVoteRows * (size_t)n_choices * 4;
}

// Shared-memory layout (round 1 kept the tiles row-major: pitch 20 gave 4-way conflicts on every input read and
// random-bank read-modify-writes on the output row; the kernel sat at 0.56 of HBM on the shared-memory pipe):
//   inputs  [k][129]: staging writes (consecutive threads = consecutive global words = consecutive k of one or two
//           rows) land in bank (k + row) mod 32 and the compute reads (thread = row, all threads at the same k)
//           too -- both conflict-free;
//   output  [j][128]: thread `row` only ever touches bank row mod 32, WHATEVER choice j its alternative selects,
//           so the data-dependent read-modify-writes are conflict-free as well.  Each thread reads its own column
//           back at the end and stores its finished row with 128-bit stores (consecutive threads own
//           consecutive rows: a warp covers one contiguous 2 KB span).
__global__ void __launch_bounds__(kVoteRows) vote_tile_kernel(const float *__restrict__ logprob,
                                                              const int32_t *__restrict__ leaf, float *__restrict__ V,
                                                              int n_rows, int n_top, int n_choices) {
    extern __shared__ __align__(16) unsigned char vsm[];
    float *lp_s = reinterpret_cast<float *>(vsm);                                   // [n_top][129]
    int32_t *lf_s = reinterpret_cast<int32_t *>(lp_s + (size_t)kVotePitch * n_top);  // [n_top][129]
    float *out_s = reinterpret_cast<float *>(lf_s + (size_t)kVotePitch * n_top);     // [n_choices][128]
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
    for (int j = 0; j < n_choices; ++j) out_s[j * kVoteRows + tid] = 0.f; // own column of the output tile
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
                out_s[j * kVoteRows + tid] += pk;
                s += pk;
            }
        }
        // normalise (client.rs:1789-1791) and write the row out; a zero sum is unreachable!() in the reference
        // (client.rs:1786-1788): the voter is marked absent
        const bool ok = s > 0.f;
        const float inv = ok ? 1.0f / s : 0.f;
        const float nanv = __int_as_float(0x7fc00000);
        float *out = V + (size_t)(row0 + tid) * n_choices;
        if ((n_choices & 3) == 0 && (reinterpret_cast<uintptr_t>(V) & 15u) == 0) {
            for (int j = 0; j < n_choices; j += 4) {
                float4 r4;
                r4.x = ok ? out_s[j * kVoteRows + tid] * inv : nanv;
                r4.y = ok ? out_s[(j + 1) * kVoteRows + tid] * inv : nanv;
                r4.z = ok ? out_s[(j + 2) * kVoteRows + tid] * inv : nanv;
                r4.w = ok ? out_s[(j + 3) * kVoteRows + tid] * inv : nanv;
                *reinterpret_cast<float4 *>(out + j) = r4;
            }
        } else {
            for (int j = 0; j < n_choices; ++j) out[j] = ok ? out_s[j * kVoteRows + tid] * inv : nanv;
        }
    }
}

// exp(x) for a logprob: 2^(x log2 e) with the product carried in two pieces (t + r), so that the result keeps fp32
// accuracy for any x a logprob can take (the plain ex2.approx(x * log2e) loses |x| * 6e-8: 1e-6 at x = -20).
// Six instructions against ~15 for expf's full range handling; subnormal results flush to zero.
__device__ __forceinline__ float exp_logprob(float x) {
    const float kL2eHi = 1.44269502162933349609375f, kL2eLo = 1.925963033500011e-8f, kLn2 = 0.693147182464599609375f;
    const float t = x * kL2eHi;
    float r = fmaf(x, kL2eHi, -t);
    r = fmaf(x, kL2eLo, r);
    float p;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(p) : "f"(t));
    return fmaf(p * kLn2, r, p);
}

// vote_tile4_kernel: the tile kernel for n_top % 4 == 0 and 16-byte aligned inputs (the reference's shapes:
// top_logprobs <= 20, llm/mod.rs:455-467).  Round 2's profile of the generic tile kernel showed it bound by
// instruction issue (~1,000 instructions per row, most of them index arithmetic of the element-wise staging and the
// full-range expf), not by shared-memory conflicts.  Here:
//   * inputs are staged ROW-MAJOR with 128-bit copies (coalesced on both sides, ~20 instructions per row);
//   * a thread reads its row back with 128-bit shared loads -- for a pitch of 4 x odd words (n_top = 20: 5) the eight
//     lanes of a quarter warp hit eight distinct 4-bank groups, conflict-free;
//   * every alternative is handled branch-free: an inadmissible one (leaf < 0 or >= n_choices, logprob NaN) adds 0 to
//     a spare column; exp is exp_logprob above;
//   * the output tile is [j][128] (+1 spare column): bank = thread, whatever column an alternative selects.
__global__ void __launch_bounds__(kVoteRows) vote_tile4_kernel(const float *__restrict__ logprob,
                                                               const int32_t *__restrict__ leaf, float *__restrict__ V,
                                                               int n_rows, int n_top, int n_choices, int pitch4) {
    extern __shared__ __align__(16) unsigned char vsm4[];
    const int pitch = pitch4 * 4;                                                   // words per staged row, 4 x odd
    float4 *lp_s = reinterpret_cast<float4 *>(vsm4);                                // [128][pitch4]
    int4 *lf_s = reinterpret_cast<int4 *>(lp_s + (size_t)kVoteRows * pitch4);       // [128][pitch4]
    float *out_s = reinterpret_cast<float *>(lf_s + (size_t)kVoteRows * pitch4);    // [n_choices + 1][128]
    const long long row0 = (long long)blockIdx.x * kVoteRows;
    const int rows = (int)(n_rows - row0 < kVoteRows ? n_rows - row0 : kVoteRows);
    const int tid = threadIdx.x;
    const int q4 = n_top >> 2; // 16-byte pieces per row
    {
        const float4 *lp = reinterpret_cast<const float4 *>(logprob + (size_t)row0 * n_top);
        const int4 *lf = reinterpret_cast<const int4 *>(leaf + (size_t)row0 * n_top);
        const int n4 = rows * q4;
        int r = tid / q4, k = tid - r * q4;
        const int dr = kVoteRows / q4, dk = kVoteRows - dr * q4;
#pragma unroll 5
        for (int i = tid; i < n4; i += kVoteRows) {
            lp_s[r * pitch4 + k] = __ldg(lp + i);
            lf_s[r * pitch4 + k] = __ldg(lf + i);
            r += dr;
            k += dk;
            if (k >= q4) {
                k -= q4;
                ++r;
            }
        }
    }
    for (int j = 0; j <= n_choices; ++j) out_s[j * kVoteRows + tid] = 0.f;
    __syncthreads();
    (void)pitch;
    if (tid < rows) {
        float s = 0.f;
        const unsigned nc = (unsigned)n_choices;
#pragma unroll 5
        for (int k = 0; k < q4; ++k) {
            const float4 x = lp_s[tid * pitch4 + k];
            const int4 l = lf_s[tid * pitch4 + k];
            const float xs[4] = {x.x, x.y, x.z, x.w};
            const int ls[4] = {l.x, l.y, l.z, l.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const bool ok = (unsigned)ls[e] < nc && xs[e] == xs[e];
                const float pk = ok ? exp_logprob(xs[e]) : 0.f;
                const int j = ok ? ls[e] : n_choices; // spare column
                out_s[j * kVoteRows + tid] += pk;     // client.rs:1765-1784, in alternative order
                s += pk;
            }
        }
        const bool ok = s > 0.f; // else unreachable!() in the reference (client.rs:1786-1788): mark the voter absent
        const float inv = ok ? 1.0f / s : 0.f;
        const float nanv = __int_as_float(0x7fc00000);
        float *out = V + (size_t)(row0 + tid) * n_choices;
        if ((n_choices & 3) == 0 && (reinterpret_cast<uintptr_t>(V) & 15u) == 0) {
            for (int j = 0; j < n_choices; j += 4) {
                float4 r4;
                r4.x = ok ? out_s[j * kVoteRows + tid] * inv : nanv;
                r4.y = ok ? out_s[(j + 1) * kVoteRows + tid] * inv : nanv;
                r4.z = ok ? out_s[(j + 2) * kVoteRows + tid] * inv : nanv;
                r4.w = ok ? out_s[(j + 3) * kVoteRows + tid] * inv : nanv;
                *reinterpret_cast<float4 *>(out + j) = r4;
            }
        } else {
            for (int j = 0; j < n_choices; ++j) out[j] = ok ? out_s[j * kVoteRows + tid] * inv : nanv;
        }
    }
}

inline int vote_tile4_pitch4(int n_top) { // 16-byte pieces per staged row: odd, so that 128-bit row reads do not conflict
    const int q4 = n_top >> 2;
    return q4 | 1;
}
inline size_t vote_tile4_smem(int n_top, int n_choices) {
    return (size_t)kVoteRows * vote_tile4_pitch4(n_top) * 32 + (size_t)kVoteRows * (size_t)(n_choices + 1) * 4;
}

// vote_ring_kernel: persistent form of vote_tile4_kernel.  Round 2's ncu capture of the tile kernel showed neither
// DRAM (65 %) nor the LSU pipe (71 %) saturated: every CTA ran load -> compute -> store phases in sequence, and the
// 7 CTAs an SM holds overlapped them only by chance.  Here a CTA keeps a ring of kVoteRingStages input tiles in
// flight: a tile's (logprob, leaf) rows are two CONTIGUOUS blocks of global memory (128 rows x n_top x 4 B), so one
// thread fetches each with a single bulk async copy (cp.async.bulk, completion on the stage's mbarrier) -- no
// staging instructions, no LSU wavefronts, and the loads of tile k+2 are in flight while tile k is computed.  The
// compute phase is vote_tile4_kernel's: thread = row, 128-bit shared reads of its own row (pitch n_top words,
// conflict-free when n_top / 4 is odd, two-way otherwise), branch-free alternatives, output tile [j][128].
constexpr int kVoteRingStages = 3;

inline size_t vote_ring_smem(int n_top, int n_choices) {
    return (size_t)kVoteRingStages * kVoteRows * (size_t)n_top * 8 + (size_t)kVoteRows * (size_t)(n_choices + 1) * 4 + 64;
}

__global__ void __launch_bounds__(kVoteRows) vote_ring_kernel(const float *__restrict__ logprob, const int32_t *__restrict__ leaf,
                                                              float *__restrict__ V, int n_rows, int n_top, int n_choices) {
    extern __shared__ __align__(128) unsigned char vring[];
    const int tile_words = kVoteRows * n_top;                                   // words per array and stage
    float *lp_s = reinterpret_cast<float *>(vring);                              // [stages][128][n_top]
    int32_t *lf_s = reinterpret_cast<int32_t *>(lp_s + (size_t)kVoteRingStages * tile_words);
    float *out_s = reinterpret_cast<float *>(lf_s + (size_t)kVoteRingStages * tile_words); // [n_choices + 1][128]
    uint64_t *bars = reinterpret_cast<uint64_t *>(out_s + (size_t)kVoteRows * (n_choices + 1));
    const uint32_t full0 = smem_u32(bars);
    const int tid = threadIdx.x;
    const int ntiles = (n_rows + kVoteRows - 1) / kVoteRows;
    const int q4 = n_top >> 2;

    if (tid == 0) {
        for (int s = 0; s < kVoteRingStages; ++s) mbar_init(full0 + 8 * s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    auto issue = [&](int tile, int stage) { // thread 0: both arrays of one tile, two bulk copies
        const long long row0 = (long long)tile * kVoteRows;
        const int rows = (int)(n_rows - row0 < kVoteRows ? n_rows - row0 : kVoteRows);
        const uint32_t bytes = (uint32_t)rows * (uint32_t)n_top * 4u;
        mbar_arrive_expect_tx(full0 + 8 * stage, 2 * bytes);
        bulk_g2s(smem_u32(lp_s + (size_t)stage * tile_words), logprob + (size_t)row0 * n_top, bytes, full0 + 8 * stage);
        bulk_g2s(smem_u32(lf_s + (size_t)stage * tile_words), leaf + (size_t)row0 * n_top, bytes, full0 + 8 * stage);
    };
    if (tid == 0)
        for (int k = 0; k < kVoteRingStages; ++k) {
            const int tile = blockIdx.x + k * gridDim.x;
            if (tile < ntiles) issue(tile, k);
        }
    int stage = 0;
    uint32_t phase = 0;
    const unsigned nc = (unsigned)n_choices;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const long long row0 = (long long)tile * kVoteRows;
        const int rows = (int)(n_rows - row0 < kVoteRows ? n_rows - row0 : kVoteRows);
        for (int j = 0; j <= n_choices; ++j) out_s[j * kVoteRows + tid] = 0.f; // own column only: no barrier needed
        mbar_wait(full0 + 8 * stage, phase);
        if (tid < rows) {
            const float4 *x4 = reinterpret_cast<const float4 *>(lp_s + (size_t)stage * tile_words + (size_t)tid * n_top);
            const int4 *l4 = reinterpret_cast<const int4 *>(lf_s + (size_t)stage * tile_words + (size_t)tid * n_top);
            float s = 0.f;
#pragma unroll 5
            for (int k = 0; k < q4; ++k) {
                const float4 x = x4[k];
                const int4 l = l4[k];
                const float xs[4] = {x.x, x.y, x.z, x.w};
                const int ls[4] = {l.x, l.y, l.z, l.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const bool ok = (unsigned)ls[e] < nc && xs[e] == xs[e];
                    const float pk = ok ? exp_logprob(xs[e]) : 0.f;
                    const int j = ok ? ls[e] : n_choices; // spare column
                    out_s[j * kVoteRows + tid] += pk;     // client.rs:1765-1784, in alternative order
                    s += pk;
                }
            }
            const bool ok = s > 0.f; // else unreachable!() in the reference (client.rs:1786-1788): mark the voter absent
            const float inv = ok ? 1.0f / s : 0.f;
            const float nanv = __int_as_float(0x7fc00000);
            float *out = V + (size_t)(row0 + tid) * n_choices;
            if ((n_choices & 3) == 0 && (reinterpret_cast<uintptr_t>(V) & 15u) == 0) {
                for (int j = 0; j < n_choices; j += 4) {
                    float4 r4;
                    r4.x = ok ? out_s[j * kVoteRows + tid] * inv : nanv;
                    r4.y = ok ? out_s[(j + 1) * kVoteRows + tid] * inv : nanv;
                    r4.z = ok ? out_s[(j + 2) * kVoteRows + tid] * inv : nanv;
                    r4.w = ok ? out_s[(j + 3) * kVoteRows + tid] * inv : nanv;
                    *reinterpret_cast<float4 *>(out + j) = r4;
                }
            } else {
                for (int j = 0; j < n_choices; ++j) out[j] = ok ? out_s[j * kVoteRows + tid] * inv : nanv;
            }
        }
        __syncthreads(); // every thread has read the stage: it may be refilled
        if (tid == 0) {
            const int next = tile + kVoteRingStages * gridDim.x;
            if (next < ntiles) issue(next, stage);
        }
        if (++stage == kVoteRingStages) {
            stage = 0;
            phase ^= 1u;
        }
    }
}

} // namespace lwc
