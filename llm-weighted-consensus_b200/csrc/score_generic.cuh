// score_generic.cuh -- shape-agnostic scoring kernel: any N <= 128, any D (no alignment
// requirement), exact fp32 FMA Gram.  One CTA of 256 threads per request.  It is the route for
// LWC_SIM_IDENTITY (the pure reference tally: E is never read) and for shapes the specialised
// kernels do not take (D % 4 != 0, unaligned E, 32 < N <= 128 until the tcgen05 route applies).
#pragma once

#include "score_common.cuh"

namespace lwc {

constexpr int kGenThreads = 256;
constexpr int kGenKC = 32; // k-chunk staged in shared memory

template <int NP> // padded pool size: 32, 64 or 128
struct GenericCfg {
    static constexpr int TS = NP / 16; // per-thread sub-tile edge (16x16 thread grid)
    static constexpr int SP = NP + 1;  // pitch of S
    static constexpr int EP = kGenKC + 1;
    static constexpr size_t smem_bytes() {
        return ((size_t)NP * SP + (size_t)NP * EP + 4 * NP + 64) * sizeof(float);
    }
};

// CTA-wide tail (thread j < N owns choice j). S: raw Gram in shared memory (pitch SP).
// scratch: >= 4*NP + 64 floats of shared memory.
template <int NP>
__device__ __forceinline__ void block_tally_and_rank(const ScoreParams &p, int r, const float *S, int SP,
                                                     float *scratch) {
    const int N = p.N, M = p.M;
    const int j = threadIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float *Vr = p.V + (size_t)r * M * N;
    const float *wr = p.w + (p.w_per_request ? (size_t)r * M : 0);
    const bool act = j < N;
    float *u_s = scratch;          // [NP] t_i * rn_i
    float *cf_s = scratch + NP;    // [NP] confidences
    float *part = scratch + 2 * NP; // [32] per-warp partials
    int *ipart = reinterpret_cast<int *>(scratch + 2 * NP + 32);

    float t = 0.f;
    if (act) {
        for (int m = 0; m < M; ++m) {
            const float v0 = __ldg(Vr + (size_t)m * N);
            if (isnan(v0)) continue;
            float wm = __ldg(wr + m);
            if (isnan(wm)) wm = 0.f;
            t = fmaf(__ldg(Vr + (size_t)m * N + j), wm, t);
        }
    }
    float c = t;
    if (p.cosine) {
        float rn = 0.f;
        if (act) {
            const float d = S[j * SP + j];
            rn = d > 0.f ? 1.0f / sqrtf(d) : 0.f;
            u_s[j] = t * rn;
        }
        __syncthreads();
        if (act) {
            float acc = 0.f;
            for (int i = 0; i < N; ++i) acc = fmaf(u_s[i], S[i * SP + j], acc);
            c = acc * rn;
        }
    }
    if (!act) c = 0.f;
    // block sum in a fixed order: warp shuffle, then warp 0 adds the per-warp partials
    float ws = warp_sum(c);
    if (lane == 0) part[warp] = ws;
    __syncthreads();
    float sum = 0.f;
#pragma unroll
    for (int k = 0; k < kGenThreads / 32; ++k) sum += part[k];
    const bool positive = sum > 0.f;
    const float cf = (positive && act) ? c / sum : 0.f;
    if (act) {
        cf_s[j] = cf;
        if (p.cw) p.cw[(size_t)r * N + j] = c;
        if (p.conf) p.conf[(size_t)r * N + j] = cf;
    }
    __syncthreads();
    if (p.winner) {
        float bv = act ? cf : -1.f;
        int bi = act ? j : 0x7fffffff;
        warp_argmax(bv, bi);
        if (lane == 0) {
            part[8 + warp] = bv;
            ipart[warp] = bi;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            float best = part[8];
            int besti = ipart[0];
            for (int k = 1; k < kGenThreads / 32; ++k)
                if (part[8 + k] > best || (part[8 + k] == best && ipart[k] < besti)) {
                    best = part[8 + k];
                    besti = ipart[k];
                }
            p.winner[r] = positive ? besti : -1;
        }
    }
    if (p.rank && act) {
        int pos = 0;
        for (int i = 0; i < N; ++i) {
            const float ci = cf_s[i];
            pos += (ci > cf || (ci == cf && i < j)) ? 1 : 0;
        }
        p.rank[(size_t)r * N + pos] = j;
    }
    if (p.agree) {
        // one warp per voter, round-robin
        for (int m = warp; m < M; m += kGenThreads / 32) {
            const float v0 = __ldg(Vr + (size_t)m * N);
            float a;
            if (isnan(v0)) {
                a = __int_as_float(0x7fc00000);
            } else {
                float s = 0.f;
                for (int i = lane; i < N; i += 32) s = fmaf(cf_s[i], __ldg(Vr + (size_t)m * N + i), s);
                a = warp_sum(s);
            }
            if (lane == 0) p.agree[(size_t)r * M + m] = a;
        }
    }
}

template <int NP>
__global__ void __launch_bounds__(kGenThreads) score_generic_kernel(const ScoreParams p) {
    using C = GenericCfg<NP>;
    constexpr int TS = C::TS, SP = C::SP, EP = C::EP;
    extern __shared__ __align__(16) float smem[];
    float *S = smem;                    // [NP][SP]
    float *Es = S + (size_t)NP * SP;    // [NP][EP]
    float *scratch = Es + (size_t)NP * EP;

    const int r = blockIdx.x;
    const int N = p.N, D = p.D;
    if (p.cosine) {
        const float *Er = p.E + (size_t)r * N * D;
        const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
        float acc[TS][TS];
#pragma unroll
        for (int a = 0; a < TS; ++a)
#pragma unroll
            for (int b = 0; b < TS; ++b) acc[a][b] = 0.f;
        for (int k0 = 0; k0 < D; k0 += kGenKC) {
            // stage rows [0,NP) x k [k0,k0+32): a warp reads 32 consecutive floats of one row
            for (int idx = threadIdx.x; idx < NP * kGenKC; idx += kGenThreads) {
                const int row = idx >> 5, kk = idx & 31;
                float v = 0.f;
                if (row < N && k0 + kk < D) v = __ldg(Er + (size_t)row * D + k0 + kk);
                Es[row * EP + kk] = v;
            }
            __syncthreads();
#pragma unroll 8
            for (int kk = 0; kk < kGenKC; ++kk) {
                float ra[TS], rb[TS];
#pragma unroll
                for (int a = 0; a < TS; ++a) ra[a] = Es[(ty + 16 * a) * EP + kk];
#pragma unroll
                for (int b = 0; b < TS; ++b) rb[b] = Es[(tx + 16 * b) * EP + kk];
#pragma unroll
                for (int a = 0; a < TS; ++a)
#pragma unroll
                    for (int b = 0; b < TS; ++b) acc[a][b] = fmaf(ra[a], rb[b], acc[a][b]);
            }
            __syncthreads();
        }
#pragma unroll
        for (int a = 0; a < TS; ++a)
#pragma unroll
            for (int b = 0; b < TS; ++b) S[(ty + 16 * a) * SP + tx + 16 * b] = acc[a][b];
        __syncthreads();
    }
    block_tally_and_rank<NP>(p, r, S, SP, scratch);
}

} // namespace lwc
