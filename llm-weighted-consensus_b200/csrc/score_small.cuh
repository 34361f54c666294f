// score_small.cuh -- fused normalise + similarity + vote kernel for small pools (N <= 32).
//
// One CTA per request; its KW warps split the embedding dimension.  Every byte of the
// request's [N][D] fp32 tile is read from HBM exactly once, straight into registers with
// 128-bit loads laid out so that the loaded registers ARE the tensor-core fragments:
// in mma.m16n8k8 lane (g = lane/4, t = lane%4) holds A[g][t], A[g+8][t], A[g][t+4],
// A[g+8][t+4] and B[t][g], B[t+4][g].  For a Gram matrix B = A^T, so the float4 a lane loads
// from row g at k0+4t feeds both operands; no shared memory, no transposes.  (The k order
// inside a 16-wide chunk is permuted identically for A and B, which a dot product ignores.)
//
// fp32-grade accuracy on tensor cores: x = hi + lo with hi = tf32(x), lo = tf32(x - hi);
//   S ~= hi.hi^T + hi.lo^T + (hi.lo^T)^T        (the lo.lo^T term is < 2^-22 relative)
// so only two MMA streams are needed (HH and HL); HL^T is formed in the epilogue.
//
// Bound: HBM.  Algorithmic bytes/request = 4ND + 4MN + 4M + 12N + 4M + 4.
#pragma once

#include <cuda_fp16.h>

#include "score_common.cuh"

namespace lwc {

__device__ __forceinline__ void mma_m16n8k8_tf32(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                                 uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// hi = x rounded to TF32 (nearest, ties away), lo = the exact remainder x - hi.  The tensor core
// reads only the top 19 bits of an operand, so lo needs no further rounding instruction.
__device__ __forceinline__ void split_tf32(float x, uint32_t &hi, uint32_t &lo) {
    hi = (__float_as_uint(x) + 0x1000u) & 0xffffe000u;
    lo = __float_as_uint(x - __uint_as_float(hi));
}

// NT: padded pool size (16 or 32).  X3: error-compensated 3xTF32 (true) or single TF32 (false).
// KW: warps per CTA (= K-split factor).  U: 16-wide k-chunks in flight per lane.
template <int NT, bool X3, int KW, int U>
struct SmallCfg {
    static constexpr int RPL = NT / 8;  // rows per lane
    static constexpr int MT = NT / 16;  // 16-row MMA tiles
    static constexpr int NTL = NT / 8;  // 8-col MMA tiles
    static constexpr int PITCH = NT + 1;
    // fragment registers per lane that cross warps in the K-split reduction
    static constexpr int NACC = (MT * NTL + (X3 ? MT * NTL : 0)) * 4;
    // HH fragment registers per lane that are actually accumulated (upper blocks of the symmetric tile)
    static constexpr int NHH = (MT == 1 ? NTL : NTL + (NTL - 2)) * 4;
    // floats of the areas below (red, tile, flush), rounded up to 16 bytes: where the optional load pipeline starts
    static constexpr int PIPE_OFF_FLOATS = (((KW - 1) * NACC * 32 + NT * PITCH + (X3 ? KW * NHH * 32 : 0)) + 3) / 4 * 4;
    static constexpr size_t smem_bytes() {
        size_t red = (size_t)(KW - 1) * NACC * 32 * sizeof(float);
        size_t tile = (size_t)NT * PITCH * sizeof(float);
        size_t flush = X3 ? (size_t)KW * NHH * 32 * sizeof(float) : 0;
        return red + tile + flush;
    }
};

template <int NT, bool X3, int KW, int U, int PS>
constexpr size_t small_smem_bytes() {
    return (size_t)SmallCfg<NT, X3, KW, U>::PIPE_OFF_FLOATS * 4 + (size_t)KW * PS * U * (NT / 8) * 32 * 16;
}

// The tensor core's fp32 accumulator truncates when it aligns addends (measured on this part: ~1.1e-8 relative
// drift per accumulated MMA on same-sign sums such as the diagonal |E_i|^2), so an fp32-grade route must not let
// one accumulator run through a whole long row: after this many MMAs the running sums are moved out with
// ordinary round-to-nearest adds and the accumulators restart.
#ifndef LWC_MMA_CHAIN
#define LWC_MMA_CHAIN 64
#endif
constexpr int kMmaChain = LWC_MMA_CHAIN;

__device__ __forceinline__ void mma_m16n8k16_f16(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                                 uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// fp16 pieces of four consecutive values (score_tc16.cuh has the derivation): x = h + 2^-11 l, both fp16, h rounded to
// nearest, l = fp16(2^11 (x - h)) with x - h exact.  h01/h23 and l01/l23 are packed half2 registers.
__device__ __forceinline__ void split_f16x4(const float4 &v, uint32_t &h01, uint32_t &h23, uint32_t &l01, uint32_t &l23) {
    const __half2 a = __floats2half2_rn(v.x, v.y), b = __floats2half2_rn(v.z, v.w);
    const float2 af = __half22float2(a), bf = __half22float2(b);
    const __half2 la = __floats2half2_rn(fmaf(v.x, 2048.0f, -2048.0f * af.x), fmaf(v.y, 2048.0f, -2048.0f * af.y));
    const __half2 lb = __floats2half2_rn(fmaf(v.z, 2048.0f, -2048.0f * bf.x), fmaf(v.w, 2048.0f, -2048.0f * bf.y));
    h01 = *reinterpret_cast<const uint32_t *>(&a);
    h23 = *reinterpret_cast<const uint32_t *>(&b);
    l01 = *reinterpret_cast<const uint32_t *>(&la);
    l23 = *reinterpret_cast<const uint32_t *>(&lb);
}
constexpr float kF16LoScale = 1.0f / 2048.0f;  // the HL accumulators of the fp16-piece route carry 2^11
constexpr float kF16MinNorm2 = 1.0f / 4096.0f; // squared norm below which fp16 pieces may have lost bits

// The Gram pass of one request: K loop over this warp's 16-column chunks, HH flush (kMmaChain), K-split reduction.
// On return warp 0 holds the summed fragments.  Every warp of the CTA must call it (it contains a __syncthreads).
//   F16 = false: tf32 pieces, two m16n8k8 MMAs per accumulator and chunk (hi.hi^T and hi.lo^T).
//   F16 = true (X3 only): fp16 pieces, ONE m16n8k16 MMA per accumulator and chunk -- half the tensor-pipe
//   instructions of the tf32 form, which is what bound the 32-candidate fp32-grade route (0.85 of HBM in round 1).
//   For a Gram matrix the lane's float4 (row g, columns 4t..4t+3) feeds A and B alike: the MMA's k slots {2t, 2t+1}
//   take (x, y) and {2t+8, 2t+9} take (z, w), the same permutation of the chunk's 16 columns on both operands.
template <int NT, bool X3, int KW, int U, bool F16>
__device__ __forceinline__ void small_gram(const ScoreParams &p, int r, float *red, float *fl,
                                           float (&hh)[NT / 16][NT / 8][4], float (&hl)[NT / 16][NT / 8][4]) {
    using C = SmallCfg<NT, X3, KW, U>;
    constexpr int RPL = C::RPL, MT = C::MT, NTL = C::NTL;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int N = p.N, D = p.D;
    const float *Er = p.E + (size_t)r * N * D;
    const float *rowp[RPL];
    bool rv[RPL];
#pragma unroll
    for (int q = 0; q < RPL; ++q) {
        const int row = g + 8 * q;
        rv[q] = row < N;
        rowp[q] = Er + (size_t)(rv[q] ? row : 0) * D + 4 * t;
    }
    // 16-wide chunks, dealt round-robin in groups of U to the KW warps
    const int nchunk = (D + 15) >> 4;
    constexpr int kPerChunk = F16 ? 1 : 2; // MMAs per accumulator and chunk
    constexpr int kFlushIters = kMmaChain / (kPerChunk * U) > 0 ? kMmaChain / (kPerChunk * U) : 1;
    if (X3) {
#pragma unroll
        for (int s = 0; s < C::NHH; ++s) fl[s * 32] = 0.f;
    }
    // segments of kFlushIters iterations: the inner loop is the plain streaming loop, the running HH sums leave the
    // accumulators between segments
    for (int seg0 = warp * U; seg0 < nchunk; seg0 += KW * U * kFlushIters) {
        const int seg1 = min(nchunk, seg0 + KW * U * kFlushIters);
        for (int c0 = seg0; c0 < seg1; c0 += KW * U) {
            float4 buf[U][RPL];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int k = (c0 + u) * 16 + 4 * t;
#pragma unroll
                for (int q = 0; q < RPL; ++q) {
                    if (rv[q] && k < D)
                        buf[u][q] = ld_stream_f4(rowp[q] + (size_t)(c0 + u) * 16);
                    else
                        buf[u][q] = make_float4(0.f, 0.f, 0.f, 0.f);
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (F16) {
                    uint32_t h01[RPL], h23[RPL], l01[RPL], l23[RPL];
#pragma unroll
                    for (int q = 0; q < RPL; ++q) split_f16x4(buf[u][q], h01[q], h23[q], l01[q], l23[q]);
#pragma unroll
                    for (int a = 0; a < MT; ++a) {
#pragma unroll
                        for (int b = 0; b < NTL; ++b) {
                            if (b >= 2 * a) // upper blocks only: S is symmetric
                                mma_m16n8k16_f16(hh[a][b], h01[2 * a], h01[2 * a + 1], h23[2 * a], h23[2 * a + 1], h01[b], h23[b]);
                            mma_m16n8k16_f16(hl[a][b], h01[2 * a], h01[2 * a + 1], h23[2 * a], h23[2 * a + 1], l01[b], l23[b]);
                        }
                    }
                } else {
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        uint32_t xh[RPL], yh[RPL], xl[RPL], yl[RPL];
#pragma unroll
                        for (int q = 0; q < RPL; ++q) {
                            const float x = h ? buf[u][q].z : buf[u][q].x;
                            const float y = h ? buf[u][q].w : buf[u][q].y;
                            if (X3) {
                                split_tf32(x, xh[q], xl[q]);
                                split_tf32(y, yh[q], yl[q]);
                            } else {
                                xh[q] = __float_as_uint(x) + 0x1000u; // RNA; the low 13 bits are ignored
                                yh[q] = __float_as_uint(y) + 0x1000u;
                            }
                        }
#pragma unroll
                        for (int a = 0; a < MT; ++a) {
#pragma unroll
                            for (int b = 0; b < NTL; ++b) {
                                if (b >= 2 * a) // upper blocks only: S is symmetric
                                    mma_m16n8k8_tf32(hh[a][b], xh[2 * a], xh[2 * a + 1], yh[2 * a], yh[2 * a + 1], xh[b],
                                                     yh[b]);
                                if (X3)
                                    mma_m16n8k8_tf32(hl[a][b], xh[2 * a], xh[2 * a + 1], yh[2 * a], yh[2 * a + 1], xl[b],
                                                     yl[b]);
                            }
                        }
                    }
                }
            }
        }
        if (X3 && seg1 < nchunk) { // bound the in-accumulator chain (HL is ~2^-11 of HH: its drift is harmless)
            int s = 0;
#pragma unroll
            for (int a = 0; a < MT; ++a)
#pragma unroll
                for (int b = 0; b < NTL; ++b)
                    if (b >= 2 * a) {
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            fl[s * 32] += hh[a][b][e];
                            hh[a][b][e] = 0.f;
                            ++s;
                        }
                    }
        }
    }
    if (X3) {
        int s = 0;
#pragma unroll
        for (int a = 0; a < MT; ++a)
#pragma unroll
            for (int b = 0; b < NTL; ++b)
                if (b >= 2 * a) {
#pragma unroll
                    for (int e = 0; e < 4; ++e) hh[a][b][e] += fl[(s++) * 32];
                }
    }

    // ---- K-split reduction: warps 1..KW-1 park their fragments, warp 0 adds them in order
    if (KW > 1) {
        if (warp > 0) {
            float *dst = red + (size_t)(warp - 1) * C::NACC * 32 + lane;
            int s = 0;
#pragma unroll
            for (int a = 0; a < MT; ++a)
#pragma unroll
                for (int b = 0; b < NTL; ++b)
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        dst[(s++) * 32] = hh[a][b][e];
                        if (X3) dst[(s++) * 32] = hl[a][b][e];
                    }
        }
        __syncthreads();
        if (warp == 0) {
            for (int wsrc = 0; wsrc < KW - 1; ++wsrc) {
                const float *src = red + (size_t)wsrc * C::NACC * 32 + lane;
                int s = 0;
#pragma unroll
                for (int a = 0; a < MT; ++a)
#pragma unroll
                    for (int b = 0; b < NTL; ++b)
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            hh[a][b][e] += src[(s++) * 32];
                            if (X3) hl[a][b][e] += src[(s++) * 32];
                        }
            }
        }
    }
}

// 16-byte asynchronous global -> shared copy (L2 only); src_bytes = 0 writes zeros without reading
__device__ __forceinline__ void cp_async16(uint32_t dst, const void *src, int src_bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// The fp16-piece Gram pass with the loads software-pipelined through shared memory (fp32-grade route only).
// The direct form above keeps U chunks per warp in flight in registers: at 128 registers 16 warps fit an SM, 64 KB in
// flight, and ncu showed the kernel waiting on its own loads (long-scoreboard stall 6.9, issue slots 36 %).  Here every
// lane owns PS private 16-byte slots per (chunk, row group): it issues cp.async for the chunks PS-1 iterations ahead,
// waits for the oldest group and reads back exactly the bytes it asked for -- no cross-lane synchronisation, no
// bank conflicts (consecutive lanes, consecutive 16 bytes), and the registers hold one iteration's data only.
// NB: 8-row groups (= 8-column MMA tiles) that can hold candidates; NB = 3 serves 17..24-candidate pools without loading,
// splitting or multiplying the fourth group (10 MMAs per chunk instead of 14).
template <int NT, int KW, int U, int PS, int NB = NT / 8>
__device__ __forceinline__ void small_gram_pipe(const ScoreParams &p, int r, float *red, float *fl, float4 *pipe,
                                                float (&hh)[NT / 16][NT / 8][4], float (&hl)[NT / 16][NT / 8][4]) {
    using C = SmallCfg<NT, true, KW, U>;
    constexpr int RPL = C::RPL, MT = C::MT, NTL = C::NTL;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int N = p.N, D = p.D;
    const float *Er = p.E + (size_t)r * N * D;
    const float *rowp[RPL];
    bool rv[RPL];
#pragma unroll
    for (int q = 0; q < RPL; ++q) {
        const int row = g + 8 * q;
        rv[q] = row < N;
        rowp[q] = Er + (size_t)(rv[q] ? row : 0) * D + 4 * t;
    }
    const int nchunk = (D + 15) >> 4;
    const int n_it = (nchunk - warp * U + KW * U - 1) / (KW * U); // iterations of this warp (<= 0: none)
    constexpr int kFlushIters = kMmaChain / U > 0 ? kMmaChain / U : 1; // one MMA per accumulator and chunk
    // this lane's slots: [warp][stage][u][q][lane]
    float4 *mine = pipe + (size_t)warp * PS * U * RPL * 32 + lane;
    const uint32_t mine_u32 = (uint32_t)__cvta_generic_to_shared(mine);
    auto issue = [&](int it) {
        if (it < n_it) {
            const int c0 = warp * U + it * KW * U;
            const uint32_t base = mine_u32 + (uint32_t)((it % PS) * U * RPL * 32 * 16);
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int k = (c0 + u) * 16 + 4 * t;
#pragma unroll
                for (int q = 0; q < NB; ++q) {
                    const bool ok = rv[q] && k < D;
                    cp_async16(base + (uint32_t)((u * RPL + q) * 32 * 16), ok ? (const void *)(rowp[q] + (size_t)(c0 + u) * 16) : (const void *)Er,
                               ok ? 16 : 0);
                }
            }
        }
        cp_async_commit(); // an empty group keeps the group count uniform
    };
#pragma unroll
    for (int s = 0; s < C::NHH; ++s) fl[s * 32] = 0.f;
#pragma unroll
    for (int it = 0; it < PS - 1; ++it) issue(it);
    for (int it = 0; it < n_it; ++it) {
        issue(it + PS - 1);
        cp_async_wait<PS - 1>(); // the group of iteration `it` has landed
        const float4 *slot = mine + (size_t)(it % PS) * U * RPL * 32;
#pragma unroll
        for (int u = 0; u < U; ++u) {
            uint32_t h01[RPL], h23[RPL], l01[RPL], l23[RPL];
#pragma unroll
            for (int q = 0; q < RPL; ++q) {
                if (q < NB)
                    split_f16x4(slot[(u * RPL + q) * 32], h01[q], h23[q], l01[q], l23[q]);
                else
                    h01[q] = h23[q] = l01[q] = l23[q] = 0u; // no candidates in these rows
            }
#pragma unroll
            for (int a = 0; a < MT; ++a) {
#pragma unroll
                for (int b = 0; b < NB; ++b) {
                    if (b >= 2 * a)
                        mma_m16n8k16_f16(hh[a][b], h01[2 * a], h01[2 * a + 1], h23[2 * a], h23[2 * a + 1], h01[b], h23[b]);
                    mma_m16n8k16_f16(hl[a][b], h01[2 * a], h01[2 * a + 1], h23[2 * a], h23[2 * a + 1], l01[b], l23[b]);
                }
            }
        }
        if ((it + 1) % kFlushIters == 0 && it + 1 < n_it) {
            int s = 0;
#pragma unroll
            for (int a = 0; a < MT; ++a)
#pragma unroll
                for (int b = 0; b < NTL; ++b)
                    if (b >= 2 * a) {
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            fl[s * 32] += hh[a][b][e];
                            hh[a][b][e] = 0.f;
                            ++s;
                        }
                    }
        }
    }
    cp_async_wait<0>();
    {
        int s = 0;
#pragma unroll
        for (int a = 0; a < MT; ++a)
#pragma unroll
            for (int b = 0; b < NTL; ++b)
                if (b >= 2 * a) {
#pragma unroll
                    for (int e = 0; e < 4; ++e) hh[a][b][e] += fl[(s++) * 32];
                }
    }
    // ---- K-split reduction, as small_gram
    if (KW > 1) {
        if (warp > 0) {
            float *dst = red + (size_t)(warp - 1) * C::NACC * 32 + lane;
            int s = 0;
#pragma unroll
            for (int a = 0; a < MT; ++a)
#pragma unroll
                for (int b = 0; b < NTL; ++b)
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        dst[(s++) * 32] = hh[a][b][e];
                        dst[(s++) * 32] = hl[a][b][e];
                    }
        }
        __syncthreads();
        if (warp == 0) {
            for (int wsrc = 0; wsrc < KW - 1; ++wsrc) {
                const float *src = red + (size_t)wsrc * C::NACC * 32 + lane;
                int s = 0;
#pragma unroll
                for (int a = 0; a < MT; ++a)
#pragma unroll
                    for (int b = 0; b < NTL; ++b)
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            hh[a][b][e] += src[(s++) * 32];
                            hl[a][b][e] += src[(s++) * 32];
                        }
            }
        }
    }
}

// warp 0: S = HH + scale (HL + HL^T) into the shared tile (fragment (a,b,e) -> (row,col)), both triangles
template <int NT, bool X3>
__device__ __forceinline__ void small_assemble(float *tile, float (&hh)[NT / 16][NT / 8][4], float (&hl)[NT / 16][NT / 8][4],
                                               float lo_scale) {
    constexpr int MT = NT / 16, NTL = NT / 8, PITCH = NT + 1;
    const int lane = threadIdx.x & 31;
    const int g = lane >> 2, t = lane & 3;
    if (X3) {
#pragma unroll
        for (int a = 0; a < MT; ++a)
#pragma unroll
            for (int b = 0; b < NTL; ++b)
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int row = 16 * a + g + 8 * (e >> 1), col = 8 * b + 2 * t + (e & 1);
                    tile[row * PITCH + col] = hl[a][b][e];
                }
        __syncwarp();
#pragma unroll
        for (int a = 0; a < MT; ++a)
#pragma unroll
            for (int b = 0; b < NTL; ++b)
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int row = 16 * a + g + 8 * (e >> 1), col = 8 * b + 2 * t + (e & 1);
                    if (b >= 2 * a) hh[a][b][e] = fmaf(hl[a][b][e] + tile[col * PITCH + row], lo_scale, hh[a][b][e]);
                }
        __syncwarp();
    }
#pragma unroll
    for (int a = 0; a < MT; ++a)
#pragma unroll
        for (int b = 0; b < NTL; ++b)
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int row = 16 * a + g + 8 * (e >> 1), col = 8 * b + 2 * t + (e & 1);
                if (b >= 2 * a) {
                    tile[row * PITCH + col] = hh[a][b][e];
                    if (b >= 2 * a + 2) tile[col * PITCH + row] = hh[a][b][e]; // mirror the off-diagonal block
                }
            }
    __syncwarp();
}

// F16 (fp32-grade only): fp16 pieces first; a request with a candidate whose squared norm is not in
// [2^-12, inf) -- fp16's exponent range was left, or every element underflowed -- is scored again on tf32 pieces by
// the same CTA (rare: unit-norm or O(1)-norm embeddings never are).
// PS > 0 (with F16): the loads of the fp16-piece pass are pipelined through PS shared-memory stages (small_gram_pipe).
template <int NT, bool X3, int KW, int U, bool F16 = false, int PS = 0, int NB = NT / 8>
__global__ void __launch_bounds__(KW * 32, PS > 0 ? 4 : (NT == 16 ? 32 : 16) / KW)
score_small_kernel(const ScoreParams p) {
    static_assert(NB == NT / 8 || PS > 0, "fewer row groups: the pipelined pass only");
    static_assert(!F16 || X3, "fp16 pieces are the fp32-grade route");
    static_assert(PS == 0 || F16, "the pipelined loads belong to the fp16-piece pass");
    using C = SmallCfg<NT, X3, KW, U>;
    constexpr int MT = C::MT, NTL = C::NTL, PITCH = C::PITCH;
    extern __shared__ __align__(16) float smem[];
    float *red = smem;                                    // [KW-1][NACC][32]
    float *tile = smem + (size_t)(KW - 1) * C::NACC * 32; // [NT][PITCH]
    float *fl = tile + NT * PITCH + (size_t)(threadIdx.x >> 5) * C::NHH * 32 + (threadIdx.x & 31); // X3: [KW][NHH][32], lane-private
    float4 *pipe = reinterpret_cast<float4 *>(smem + C::PIPE_OFF_FLOATS); // PS > 0: [KW][PS][U][RPL][32] x 16 B
    __shared__ int redo;

    const int r = blockIdx.x;
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;

    float hh[MT][NTL][4];
    float hl[MT][NTL][4];
#pragma unroll
    for (int a = 0; a < MT; ++a)
#pragma unroll
        for (int b = 0; b < NTL; ++b)
#pragma unroll
            for (int e = 0; e < 4; ++e) hh[a][b][e] = hl[a][b][e] = 0.f;

    if (p.cosine) {
        if (PS > 0)
            small_gram_pipe<NT, KW, U, (PS > 0 ? PS : 2), NB>(p, r, red, fl, pipe, hh, hl);
        else
            small_gram<NT, X3, KW, U, F16>(p, r, red, fl, hh, hl);
        if (warp == 0) small_assemble<NT, X3>(tile, hh, hl, F16 ? kF16LoScale : 1.0f);
        if (F16) {
            if (warp == 0) {
                const float d = lane < p.N ? tile[lane * PITCH + lane] : 1.0f;
                const bool bad = !(d >= kF16MinNorm2 && d < __int_as_float(0x7f800000));
                const bool any = __any_sync(kFullMask, bad);
                if (lane == 0) redo = any ? 1 : 0;
            }
            __syncthreads();
            if (redo) { // CTA-uniform
#pragma unroll
                for (int a = 0; a < MT; ++a)
#pragma unroll
                    for (int b = 0; b < NTL; ++b)
#pragma unroll
                        for (int e = 0; e < 4; ++e) hh[a][b][e] = hl[a][b][e] = 0.f;
                small_gram<NT, X3, KW, U, false>(p, r, red, fl, hh, hl);
                if (warp == 0) small_assemble<NT, X3>(tile, hh, hl, 1.0f);
            }
        }
    }
    if (warp != 0) return;
    warp_tally_and_rank<NT>(p, r, tile, PITCH, lane);
}

} // namespace lwc
