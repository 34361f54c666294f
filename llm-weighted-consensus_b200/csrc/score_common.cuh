// score_common.cuh -- shared device-side pieces of the scoring kernels (sm_100a).
//
// Everything after the similarity matrix is the reference's tally, restated per request:
//   t_i   = sum_m w_m V_mi            src/score/completions/client.rs:410-415
//   sum   = sum_j cw_j                client.rs:422-424
//   conf  = cw / sum if sum > 0       client.rs:428-437
//   agree = sum_i conf_i V_mi         client.rs:438-449
// with cw = t (similarity identity, the reference) or cw_j = sum_i t_i S_ij (cosine,
// build-defined).  winner / rank are build-defined (8.0 step 6).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace lwc {

struct ScoreParams {
    const float *E;  // [R][N][D]
    const float *V;  // [R][M][N]
    const float *w;  // [R][M] or [M]
    float *conf;     // [R][N]
    float *cw;       // [R][N]
    float *agree;    // [R][M]
    int32_t *winner; // [R]
    int32_t *rank;   // [R][N]
    int R, N, D, M;
    int w_per_request;
    int cosine;
    int inputs_resident; // LWC_CTX_INPUTS_RESIDENT: input reads need not wait for the previous grid (PDL)
};

// Programmatic dependent launch: let the next grid of the stream be scheduled as soon as this one's CTAs
// retire, and (wait) block until the previous grid has completed and its writes are visible.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait_prior_grid() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

constexpr unsigned kFullMask = 0xffffffffu;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFullMask, v, o);
    return v;
}

// (value, index) arg-max with ties resolved to the lowest index.
// Sum V per-lane values across the warp; afterwards lane l holds the total of value l >> (5 - log2 V).
template <int V>
__device__ __forceinline__ float warp_transpose_sum(float (&v)[V], int lane) {
    static_assert(V == 8 || V == 16 || V == 32, "V");
    int off = 16;
#pragma unroll
    for (int n = V / 2; n >= 1; n >>= 1, off >>= 1) {
        const bool up = (lane & off) != 0;
#pragma unroll
        for (int i = 0; i < n; ++i) {
            const float send = up ? v[i] : v[i + n];
            const float keep = up ? v[i + n] : v[i];
            v[i] = keep + __shfl_xor_sync(kFullMask, send, off);
        }
    }
    float r = v[0];
    for (; off >= 1; off >>= 1) r += __shfl_xor_sync(kFullMask, r, off);
    return r;
}

__device__ __forceinline__ void warp_argmax(float &v, int &idx) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        float ov = __shfl_xor_sync(kFullMask, v, o);
        int oi = __shfl_xor_sync(kFullMask, idx, o);
        if (ov > v || (ov == v && oi < idx)) {
            v = ov;
            idx = oi;
        }
    }
}

// 128-bit streaming load: read-only path, no L1 allocation (every byte of E is used once).
__device__ __forceinline__ float4 ld_stream_f4(const float *p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p));
    return r;
}

// Warp-wide tail of the path for one request whose candidates fit one warp (N <= NTP <= 32):
// lane j owns choice j.  S is the raw Gram matrix E E^T in shared memory (pitch `pitch`, rows
// and columns >= N all zero), ignored when !p.cosine.  The first MU voters' rows/weights may
// arrive prefetched in registers (pv[m] = V[r][m][lane], pw[m] = w[m]); the rest are read from
// global memory.  Loops run to the compile-time pool size NTP so that their shuffles, shared
// loads and FMAs pipeline; padding lanes carry zeros.
template <int MU, int NTP>
__device__ __forceinline__ void warp_tally_and_rank_pf(const ScoreParams &p, int r, const float *S, int pitch,
                                                       int lane, const float *pv, const float *pw) {
    const int N = p.N, M = p.M;
    const float *Vr = p.V + (size_t)r * M * N;
    const float *wr = p.w + (p.w_per_request ? (size_t)r * M : 0);
    const bool act = lane < N;

    float t = 0.f;
#pragma unroll
    for (int m = 0; m < MU; ++m) {
        if (m < M) {
            const float v0 = __shfl_sync(kFullMask, pv[m], 0);
            float wm = pw[m];
            if (isnan(wm)) wm = 0.f;
            if (!isnan(v0) && act) t = fmaf(pv[m], wm, t);
        }
    }
    // the voters that were not prefetched: eight voters' loads in flight at a time (a load-test-load chain per
    // voter costs ~1.5 k clk each, during which this warp's CTA streams nothing)
    for (int m0 = MU; m0 < M; m0 += 8) {
        float v0[8], wv[8], vl[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int m = m0 + u;
            const bool ok = m < M;
            v0[u] = ok ? __ldg(Vr + (size_t)m * N) : __int_as_float(0x7fc00000);
            wv[u] = ok ? __ldg(wr + m) : 0.f;
            vl[u] = (ok && act) ? __ldg(Vr + (size_t)m * N + lane) : 0.f;
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            if (isnan(v0[u])) continue; // voter without a vote (client.rs:410) -- warp-uniform
            const float wm = isnan(wv[u]) ? 0.f : wv[u]; // weight None -> 0 (client.rs:412-413)
            if (act) t = fmaf(vl[u], wm, t);
        }
    }

    float c = t;
    if (p.cosine) {
        const int jl = lane < NTP ? lane : 0;
        const float d = act ? S[jl * pitch + jl] : 0.f;
        const float rn = d > 0.f ? 1.0f / sqrtf(d) : 0.f; // zero-norm row: E^_i = 0
        const float u = t * rn;                            // 0 on padding lanes
        float col[NTP];
#pragma unroll
        for (int i = 0; i < NTP; ++i) col[i] = S[i * pitch + jl];
        float acc = 0.f;
#pragma unroll
        for (int i = 0; i < NTP; ++i) acc = fmaf(__shfl_sync(kFullMask, u, i), col[i], acc);
        c = acc * rn;
    }
    if (!act) c = 0.f;
    const float sum = warp_sum(c);
    const bool positive = sum > 0.f;
    const float cf = (positive && act) ? c / sum : 0.f;

    if (act) {
        if (p.cw) p.cw[(size_t)r * N + lane] = c;
        if (p.conf) p.conf[(size_t)r * N + lane] = cf;
    }
    if (p.winner) {
        float bv = act ? cf : -1.f;
        int bi = act ? lane : 0x7fffffff;
        warp_argmax(bv, bi);
        if (lane == 0) p.winner[r] = positive ? bi : -1;
    }
    if (p.rank) {
        int pos = 0;
#pragma unroll
        for (int i = 0; i < NTP; ++i) {
            const float ci = __shfl_sync(kFullMask, cf, i);
            pos += (i < N && (ci > cf || (ci == cf && i < lane))) ? 1 : 0;
        }
        if (act) p.rank[(size_t)r * N + pos] = lane;
    }
    if (p.agree) {
        float part[MU > 0 ? MU : 1];
#pragma unroll
        for (int m = 0; m < MU; ++m) part[m] = (act && m < M) ? cf * pv[m] : 0.f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
            for (int m = 0; m < MU; ++m) part[m] += __shfl_xor_sync(kFullMask, part[m], o);
        }
#pragma unroll
        for (int m = 0; m < MU; ++m) {
            if (m < M) {
                const float v0 = __shfl_sync(kFullMask, pv[m], 0);
                if (lane == 0) p.agree[(size_t)r * M + m] = isnan(v0) ? __int_as_float(0x7fc00000) : part[m];
            }
        }
        for (int m0 = MU; m0 < M; m0 += 8) {
            float v0[8], vl[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int m = m0 + u;
                v0[u] = m < M ? __ldg(Vr + (size_t)m * N) : 0.f;
                vl[u] = (m < M && act) ? __ldg(Vr + (size_t)m * N + lane) : 0.f;
            }
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int m = m0 + u;
                if (m >= M) break;
                const float a = isnan(v0[u]) ? __int_as_float(0x7fc00000) : warp_sum(act ? cf * vl[u] : 0.f);
                if (lane == 0) p.agree[(size_t)r * M + m] = a;
            }
        }
    }
}

// The same tail for a GROUP of G = 8 / N requests of at most 4 candidates each whose rows share one 8-row Gram tile:
// lane l < G * N owns choice l % N of request r0 + l / N.  S holds the 8 x 8 Gram matrix of the stacked rows; only its
// diagonal N x N blocks mean anything (the others are products across requests) and every sum below runs over the
// lane's own block -- in the order warp_sum's butterfly adds four leading lanes, (v0 + v2) + (v1 + v3), so that a request
// gets the bits it would get alone in a tile.  n_valid: requests of the group that exist (the last group of a batch may
// be short; rows of absent requests hold stale data that is never selected into a sum).  The first MU voters' entries
// arrive prefetched: pv[m] = V[r][m][c], pw[m] = w[r][m] of the lane's own request.
__device__ __forceinline__ float block_sum4(float x, int first, int N) {
    float a[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float v = __shfl_sync(kFullMask, x, (first + k) & 31);
        a[k] = k < N ? v : 0.f;
    }
    return (a[0] + a[2]) + (a[1] + a[3]);
}

template <int MU>
__device__ __forceinline__ void warp_tally_group(const ScoreParams &p, int r0, int G, int n_valid, const float *S, int pitch,
                                                 int lane, const float *pv, const float *pw) {
    const int N = p.N, M = p.M;
    const int g = lane / N, c = lane - g * N; // (N <= 4: a handful of integer ops)
    const bool act = lane < G * N && g < n_valid;
    const int r = r0 + (act ? g : 0);
    const int first = act ? g * N : 0; // lane of the block's choice 0
    const float *Vr = p.V + (size_t)r * M * N;
    const float *wr = p.w + (p.w_per_request ? (size_t)r * M : 0);

    float t = 0.f;
#pragma unroll
    for (int m = 0; m < MU; ++m) {
        if (m < M) {
            const float v0 = __shfl_sync(kFullMask, pv[m], first);
            float wm = pw[m];
            if (isnan(wm)) wm = 0.f;
            if (!isnan(v0) && act) t = fmaf(pv[m], wm, t);
        }
    }
    for (int m = MU; m < M; ++m) { // (few candidates: the voters beyond the prefetch window one at a time)
        const float v0 = act ? __ldg(Vr + (size_t)m * N) : __int_as_float(0x7fc00000);
        float wm = act ? __ldg(wr + m) : 0.f;
        const float vl = act ? __ldg(Vr + (size_t)m * N + c) : 0.f;
        if (isnan(wm)) wm = 0.f;
        if (!isnan(v0)) t = fmaf(vl, wm, t);
    }

    float cwv = t;
    if (p.cosine) {
        const int jl = act ? lane : 0;
        const float d = act ? S[jl * pitch + jl] : 0.f;
        const float rn = d > 0.f ? 1.0f / sqrtf(d) : 0.f;
        const float u = t * rn;
        float acc = 0.f;
#pragma unroll
        for (int k = 0; k < 4; ++k) { // own block only, ascending like the single-request loop
            const float uk = __shfl_sync(kFullMask, u, (first + k) & 31);
            if (k < N) acc = fmaf(uk, S[(first + k) * pitch + jl], acc);
        }
        cwv = acc * rn;
    }
    if (!act) cwv = 0.f;
    const float sum = block_sum4(cwv, first, N);
    const bool positive = sum > 0.f;
    const float cf = (positive && act) ? cwv / sum : 0.f;
    if (act) {
        if (p.cw) p.cw[(size_t)r * N + c] = cwv;
        if (p.conf) p.conf[(size_t)r * N + c] = cf;
    }
    // argmax (ties -> lowest index) and rank inside the block
    float best = -1.f;
    int besti = 0, pos = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float ck = __shfl_sync(kFullMask, cf, (first + k) & 31);
        if (k < N) {
            if (ck > best) best = ck, besti = k;
            pos += (ck > cf || (ck == cf && k < c)) ? 1 : 0;
        }
    }
    if (act) {
        if (p.winner && c == 0) p.winner[r] = positive ? besti : -1;
        if (p.rank) p.rank[(size_t)r * N + pos] = c;
    }
    if (p.agree) {
#pragma unroll
        for (int m = 0; m < MU; ++m) {
            if (m < M) {
                const float a = block_sum4(act ? cf * pv[m] : 0.f, first, N);
                const float v0 = __shfl_sync(kFullMask, pv[m], first);
                if (act && c == 0) p.agree[(size_t)r * M + m] = isnan(v0) ? __int_as_float(0x7fc00000) : a;
            }
        }
        for (int m = MU; m < M; ++m) {
            const float v0 = act ? __ldg(Vr + (size_t)m * N) : 0.f;
            const float a = block_sum4(act ? cf * __ldg(Vr + (size_t)m * N + c) : 0.f, first, N);
            if (act && c == 0) p.agree[(size_t)r * M + m] = isnan(v0) ? __int_as_float(0x7fc00000) : a;
        }
    }
}

template <int NTP>
__device__ __forceinline__ void warp_tally_and_rank(const ScoreParams &p, int r, const float *S, int pitch,
                                                    int lane) {
    warp_tally_and_rank_pf<0, NTP>(p, r, S, pitch, lane, nullptr, nullptr);
}

} // namespace lwc
