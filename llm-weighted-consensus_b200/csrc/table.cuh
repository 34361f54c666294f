// table.cuh -- training-table weight lookup: the slot the reference ships as
// `unimplemented!()` (src/score/completions/weight.rs:99-117; config `WeightTrainingTable { embeddings, top }`,
// src/score/model/mod.rs:278-306; per-LLM `base_weight / min_weight / max_weight`,
// src/score/llm/mod.rs:663-688).  No arithmetic for it exists in the reference, so everything here is
// BUILD-DEFINED and parity is unpinned by construction; the device part is kept policy-free:
//
//   table_norm_kernel    1 / |row| of the resident table, once at load.
//   table_scores_kernel  cosine of every table row against up to kTableQB prompt embeddings per pass;
//                        HBM-bound stream over the table (one warp per row, 128-bit loads).
//   table_topk_kernel    the `top` best rows per prompt (ties -> lowest row): a reduction tree of per-chunk
//                        top lists held in registers.
//
// The weight policy itself (similarity-weighted local quality of each LLM, scaled around base, clamped
// to [min, max]) is a few flops per LLM and runs in table_weights_kernel.
#pragma once

#include "../../include/lwc_b200.h"
#include "score_common.cuh"
#include "score_stream.cuh" // mbarrier / bulk-copy helpers

namespace lwc {

constexpr int kTableQB = 8;     // prompt embeddings per scoring pass
constexpr int kTableMaxTop = 64;
constexpr int kTableRows = 4;   // table rows per warp step of the score pass

__global__ void __launch_bounds__(256)
table_norm_kernel(const float *__restrict__ T, float *__restrict__ inv_norm, long long Tn, int D) {
    const long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= Tn) return;
    const float *e = T + (size_t)row * D;
    float ss = 0.f;
    for (int k = lane; k < D; k += 32) ss = fmaf(e[k], e[k], ss);
    ss = warp_sum(ss);
    if (lane == 0) inv_norm[row] = ss > 0.f ? 1.0f / sqrtf(ss) : 0.f;
}

// scores[b][t] = <q_b, T_t> / (|q_b| |T_t|) for b < nq <= kTableQB.  q in shared memory.
__global__ void __launch_bounds__(256)
table_scores_kernel(const float *__restrict__ T, const float *__restrict__ inv_norm, const float *__restrict__ q,
                    float *__restrict__ scores, long long Tn, int D, int nq) {
    extern __shared__ float q_s[]; // [nq][D] then [nq] inverse norms
    float *qn = q_s + (size_t)nq * D;
    for (int i = threadIdx.x; i < nq * D; i += blockDim.x) q_s[i] = q[i];
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (warp < nq) {
        float ss = 0.f;
        for (int k = lane; k < D; k += 32) ss = fmaf(q_s[warp * D + k], q_s[warp * D + k], ss);
        ss = warp_sum(ss);
        if (lane == 0) qn[warp] = ss > 0.f ? 1.0f / sqrtf(ss) : 0.f;
    }
    __syncthreads();
    const bool vec = (D % 4 == 0) && ((reinterpret_cast<uintptr_t>(T) & 15u) == 0);
    // kTableRows consecutive rows per warp step: every prompt value read from shared memory is used for
    // all of them, so the pass stays HBM-bound instead of shared-memory-bound.
    const long long groups = (Tn + kTableRows - 1) / kTableRows;
    for (long long g = (long long)blockIdx.x * (blockDim.x >> 5) + warp; g < groups;
         g += (long long)gridDim.x * (blockDim.x >> 5)) {
        const long long row0 = g * kTableRows;
        const float *e[kTableRows];
#pragma unroll
        for (int r = 0; r < kTableRows; ++r) e[r] = T + (size_t)(row0 + r < Tn ? row0 + r : Tn - 1) * D;
        float acc[kTableRows][kTableQB];
#pragma unroll
        for (int r = 0; r < kTableRows; ++r)
#pragma unroll
            for (int b = 0; b < kTableQB; ++b) acc[r][b] = 0.f;
        if (vec) {
#pragma unroll 2
            for (int k = lane * 4; k < D; k += 128) {
                float4 x[kTableRows];
#pragma unroll
                for (int r = 0; r < kTableRows; ++r) x[r] = ld_stream_f4(e[r] + k);
#pragma unroll
                for (int b = 0; b < kTableQB; ++b)
                    if (b < nq) {
                        const float4 y = *reinterpret_cast<const float4 *>(q_s + b * D + k);
#pragma unroll
                        for (int r = 0; r < kTableRows; ++r)
                            acc[r][b] = fmaf(x[r].x, y.x, fmaf(x[r].y, y.y, fmaf(x[r].z, y.z, fmaf(x[r].w, y.w, acc[r][b]))));
                    }
            }
        } else {
            for (int k = lane; k < D; k += 32) {
                float x[kTableRows];
#pragma unroll
                for (int r = 0; r < kTableRows; ++r) x[r] = e[r][k];
#pragma unroll
                for (int b = 0; b < kTableQB; ++b)
                    if (b < nq) {
                        const float y = q_s[b * D + k];
#pragma unroll
                        for (int r = 0; r < kTableRows; ++r) acc[r][b] = fmaf(x[r], y, acc[r][b]);
                    }
            }
        }
#pragma unroll
        for (int r = 0; r < kTableRows; ++r) {
            if (row0 + r >= Tn) break;
            const float rn = inv_norm[row0 + r];
#pragma unroll
            for (int b = 0; b < kTableQB; ++b)
                if (b < nq) {
                    const float sdot = warp_sum(acc[r][b]);
                    if (lane == 0) scores[(size_t)b * Tn + row0 + r] = sdot * rn * qn[b];
                }
        }
    }
}

// The same pass as a persistent producer/consumer pipeline (the route for D % 4 == 0 tables whose ring fits):
// warp 0 streams groups of RS consecutive table rows (one contiguous cp.async.bulk each) into a ring of
// kTableSlots slots; slot s belongs to the consumer warp PAIR (2s+1, 2s+2) outright -- both warps wait on
// full[s] and arrive on empty[s] every lap, so each observes every phase of its barriers.  The pair splits
// the columns (alternate 128-float chunks), every prompt value read from shared memory is used for all RS
// rows, and the RS x kTableQB dot products stay in registers; the odd warp hands its partial sums to the even
// one through a lap-parity double buffer.  kTableSlots * RS * D * 4 bytes are in flight per SM at all times.
constexpr int kTableSlots = 7;
constexpr int kTableRingThreads = (2 * kTableSlots + 1) * 32;

struct TableRingSmem { // byte offsets inside the dynamic shared memory
    size_t qn, bars, part, ring, total;
    __host__ __device__ TableRingSmem(int RS, int D, int nq) {
        qn = (size_t)nq * D * 4;
        bars = qn + kTableQB * 4;                 // full[slots], empty[slots], pair[slots]
        part = bars + 3 * kTableSlots * 8;        // [slots][2][32] floats
        ring = (part + kTableSlots * 2 * 32 * 4 + 127) & ~(size_t)127;
        total = ring + (size_t)kTableSlots * RS * D * 4;
    }
};

template <int RS>
__global__ void __launch_bounds__(kTableRingThreads, 1)
table_scores_ring_kernel(const float *__restrict__ T, const float *__restrict__ inv_norm, const float *__restrict__ q,
                         float *__restrict__ scores, long long Tn, int D, int nq) {
    extern __shared__ __align__(128) unsigned char tsm[];
    const TableRingSmem L(RS, D, nq);
    float *q_s = reinterpret_cast<float *>(tsm); // [nq][D]
    float *qn = reinterpret_cast<float *>(tsm + L.qn);
    uint64_t *bars = reinterpret_cast<uint64_t *>(tsm + L.bars);
    float *part = reinterpret_cast<float *>(tsm + L.part);
    float *ring = reinterpret_cast<float *>(tsm + L.ring); // [slots][RS][D]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < nq * D; i += blockDim.x) q_s[i] = q[i];
    if (threadIdx.x == 0) {
        for (int s = 0; s < kTableSlots; ++s) {
            mbar_init(smem_u32(bars + s), 1);                   // full: the producer's expect_tx arrive
            mbar_init(smem_u32(bars + kTableSlots + s), 2);     // empty: both warps of the pair
            mbar_init(smem_u32(bars + 2 * kTableSlots + s), 1); // pair: the odd warp's partial sums are posted
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (warp < nq) {
        float ss = 0.f;
        for (int k = lane; k < D; k += 32) ss = fmaf(q_s[warp * D + k], q_s[warp * D + k], ss);
        ss = warp_sum(ss);
        if (lane == 0) qn[warp] = ss > 0.f ? 1.0f / sqrtf(ss) : 0.f;
    }
    __syncthreads();
    const long long groups = (Tn + RS - 1) / RS;
    const long long njobs = groups > blockIdx.x ? (groups - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const size_t slot_floats = (size_t)RS * D;
    if (warp == 0) {
        if (lane == 0) {
            for (long long j = 0; j < njobs; ++j) {
                const int slot = (int)(j % kTableSlots);
                const long long lap = j / kTableSlots;
                if (lap > 0) mbar_wait(smem_u32(bars + kTableSlots + slot), (uint32_t)((lap - 1) & 1));
                const long long row0 = (blockIdx.x + j * gridDim.x) * RS;
                const int nrows = (int)(Tn - row0 < RS ? Tn - row0 : RS);
                const uint32_t bytes = (uint32_t)((size_t)nrows * D * 4);
                mbar_arrive_expect_tx(smem_u32(bars + slot), bytes);
                bulk_g2s(smem_u32(ring + slot * slot_floats), T + (size_t)row0 * D, bytes, smem_u32(bars + slot));
            }
        }
        return;
    }
    const int slot = (warp - 1) >> 1, half = (warp - 1) & 1;
    const float *xs = ring + slot * slot_floats;
    constexpr int V = RS * kTableQB;
    for (long long j = slot, lap = 0; j < njobs; j += kTableSlots, ++lap) {
        mbar_wait(smem_u32(bars + slot), (uint32_t)(lap & 1));
        float acc[V];
#pragma unroll
        for (int a = 0; a < V; ++a) acc[a] = 0.f;
#pragma unroll 2
        for (int k = half * 128 + lane * 4; k < D; k += 256) {
            float4 x[RS];
#pragma unroll
            for (int r = 0; r < RS; ++r) x[r] = *reinterpret_cast<const float4 *>(xs + (size_t)r * D + k);
#pragma unroll
            for (int b = 0; b < kTableQB; ++b)
                if (b < nq) {
                    const float4 y = *reinterpret_cast<const float4 *>(q_s + b * D + k);
#pragma unroll
                    for (int r = 0; r < RS; ++r)
                        acc[r * kTableQB + b] =
                            fmaf(x[r].x, y.x, fmaf(x[r].y, y.y, fmaf(x[r].z, y.z, fmaf(x[r].w, y.w, acc[r * kTableQB + b]))));
                }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(bars + kTableSlots + slot)); // this warp no longer reads the slot
        float tot = warp_transpose_sum<V>(acc, lane);
        float *pp = part + ((size_t)slot * 2 + (lap & 1)) * 32;
        if (half == 1) {
            pp[lane] = tot;
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(bars + 2 * kTableSlots + slot));
        } else {
            mbar_wait(smem_u32(bars + 2 * kTableSlots + slot), (uint32_t)(lap & 1));
            tot += pp[lane];
            const int a = lane >> (V == 32 ? 0 : V == 16 ? 1 : 2);
            const int r = a / kTableQB, b = a % kTableQB;
            const long long row = (blockIdx.x + j * gridDim.x) * RS + r;
            const bool first = (lane & ((32 / V) - 1)) == 0;
            if (first && row < Tn && b < nq) scores[(size_t)b * Tn + row] = tot * inv_norm[row] * qn[b];
        }
    }
}

static inline size_t table_ring_smem(int RS, int D, int nq) { return TableRingSmem(RS, D, nq).total; }

// Top-k by repeated arg-max, as a reduction tree.  Grid (chunks, nq), kTopkThreads threads; a CTA holds its
// chunk of kTopkChunk (value, position) entries of one prompt in registers (kTopkPerThread per thread, each
// thread caching its own best), and extracts the chunk's `top` best in `top` rounds: warp arg-max by
// shuffles, one barrier, then every warp reduces the per-warp winners redundantly (round-parity double buffer,
// so one barrier per round).  Ties go to the lowest position, which is the lowest table row at every level:
// level 1 reads raw scores (position = row), later levels read the previous level's lists, which keep equal
// values in row order.  Output [nq][chunks][top] (value, row), -inf / -1 when the chunk runs out; the last
// level (one chunk) writes idx / sim ([nq][top], -1 / 0 past the table's size) instead.
constexpr int kTopkThreads = 256;
constexpr int kTopkPerThread = 32;
constexpr int kTopkChunk = kTopkThreads * kTopkPerThread;

template <bool HAS_IDX>
__global__ void __launch_bounds__(kTopkThreads)
table_topk_kernel(const float *__restrict__ in_val, const int32_t *__restrict__ in_idx, long long n_in, int top,
                  float *__restrict__ out_val, int32_t *__restrict__ out_idx, int32_t *__restrict__ fin_idx,
                  float *__restrict__ fin_sim) {
    constexpr int NW = kTopkThreads / 32;
    __shared__ float s_val[2][NW];
    __shared__ int s_pos[2][NW];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float *iv = in_val + (size_t)blockIdx.y * n_in;
    const long long e0 = (long long)blockIdx.x * kTopkChunk;
    float v[kTopkPerThread];
#pragma unroll
    for (int j = 0; j < kTopkPerThread; ++j) {
        const long long e = e0 + j * kTopkThreads + threadIdx.x;
        v[j] = e < n_in ? iv[e] : -INFINITY;
    }
    float bv = -INFINITY;
    int bj = 0;
#pragma unroll
    for (int j = 0; j < kTopkPerThread; ++j)
        if (v[j] > bv) bv = v[j], bj = j; // positions grow with j: the first maximum is kept
    const size_t out = ((size_t)blockIdx.y * gridDim.x + blockIdx.x) * top;
    for (int k = 0; k < top; ++k) {
        float best = bv;
        int pos = bj * kTopkThreads + threadIdx.x;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(kFullMask, best, o);
            const int op = __shfl_xor_sync(kFullMask, pos, o);
            if (ob > best || (ob == best && op < pos)) best = ob, pos = op;
        }
        if (lane == 0) s_val[k & 1][warp] = best, s_pos[k & 1][warp] = pos;
        __syncthreads();
        best = lane < NW ? s_val[k & 1][lane] : -INFINITY;
        pos = lane < NW ? s_pos[k & 1][lane] : 0x7fffffff;
#pragma unroll
        for (int o = NW / 2; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(kFullMask, best, o);
            const int op = __shfl_xor_sync(kFullMask, pos, o);
            if (ob > best || (ob == best && op < pos)) best = ob, pos = op;
        }
        best = __shfl_sync(kFullMask, best, 0);
        pos = __shfl_sync(kFullMask, pos, 0);
        const bool found = best > -INFINITY;
        if (found && (pos & (kTopkThreads - 1)) == (int)threadIdx.x) { // the winner's owner: emit, mask, rescan
            int32_t row;
            if (HAS_IDX) row = in_idx[(size_t)blockIdx.y * n_in + e0 + pos];
            else row = (int32_t)(e0 + pos);
            if (fin_idx) fin_idx[(size_t)blockIdx.y * top + k] = row, fin_sim[(size_t)blockIdx.y * top + k] = best;
            else out_val[out + k] = best, out_idx[out + k] = row;
            const int wj = pos / kTopkThreads;
            bv = -INFINITY, bj = 0;
#pragma unroll
            for (int j = 0; j < kTopkPerThread; ++j) {
                if (j == wj) v[j] = -INFINITY;
                if (v[j] > bv) bv = v[j], bj = j;
            }
        }
        if (!found && threadIdx.x == 0) {
            if (fin_idx) fin_idx[(size_t)blockIdx.y * top + k] = -1, fin_sim[(size_t)blockIdx.y * top + k] = 0.f;
            else out_val[out + k] = -INFINITY, out_idx[out + k] = -1;
        }
    }
}

// weights[b][m] = clamp(base_m * rho_m / mean_m'(rho_m'), min_m, max_m) with
// rho_m = sum_k max(sim_k, 0) * Q[idx_k][m] / sum_k max(sim_k, 0) over the `top` neighbours;
// base_m when there is no positive neighbour or mean(rho) == 0.  One warp per prompt.
__global__ void __launch_bounds__(128)
table_weights_kernel(const int32_t *__restrict__ idx, const float *__restrict__ sim, const float *__restrict__ Q,
                     const float *__restrict__ base, const float *__restrict__ minw, const float *__restrict__ maxw,
                     float *__restrict__ weights, int nq, int top, int M) {
    const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (b >= nq) return;
    float ssum = 0.f;
    for (int k = 0; k < top; ++k) {
        const int t = idx[(size_t)b * top + k];
        if (t >= 0) ssum += fmaxf(sim[(size_t)b * top + k], 0.f);
    }
    float rho[(LWC_MAX_VOTERS + 31) / 32];
    float part = 0.f;
#pragma unroll
    for (int j = 0; j < (LWC_MAX_VOTERS + 31) / 32; ++j) {
        const int m = j * 32 + lane;
        float r = 0.f;
        if (m < M && ssum > 0.f) {
            for (int k = 0; k < top; ++k) {
                const int t = idx[(size_t)b * top + k];
                if (t >= 0) r = fmaf(fmaxf(sim[(size_t)b * top + k], 0.f), Q[(size_t)t * M + m], r);
            }
            r /= ssum;
        }
        rho[j] = r;
        part += (m < M) ? r : 0.f;
    }
    const float mean = warp_sum(part) / (float)M;
#pragma unroll
    for (int j = 0; j < (LWC_MAX_VOTERS + 31) / 32; ++j) {
        const int m = j * 32 + lane;
        if (m < M) {
            float wv = base[m];
            if (ssum > 0.f && mean > 0.f) wv = base[m] * (rho[j] / mean);
            weights[(size_t)b * M + m] = fminf(fmaxf(wv, minw[m]), maxw[m]);
        }
    }
}

} // namespace lwc
