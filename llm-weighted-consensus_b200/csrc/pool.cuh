// pool.cuh -- one giant pool (BASELINE.json config 5): all-pairs consensus over N candidates, rows
// sharded over the ranks of a node.  Same spec as the batched path applied to one
// request whose pairwise matrix (N x N, up to 2.7e11 entries) is never materialised:
//
//   pool_prep_kernel      own rows: n_i = |E_i|, E^_i = E_i / n_i written as bf16 (scratch), and
//                         t_i = sum_m w_m V_mi (client.rs:410-415).
//   [rows sharded over ranks: the exchange step.  Default: NCCL all-gather of E^ (bf16) and t before the
//    tensor pass.  With peer mappings attached (lwc_pool_p2p_*): FUSED into pool_gemm_kernel -- a gather
//    warp in every CTA pulls the peers' shards over NVLink (P2P loads from cudaIpc-mapped memory) into
//    the local copy while the tensor pass already runs on the own shard's columns, then on each peer
//    shard as soon as its arrival counter is complete.]
//   pool_gemm_kernel      y_i = sum_j t_j <E^_i, E^_j> for own rows i, all columns j: 128 x 256 tiles of
//                         S = E^ E^T on tcgen05 (kind::f16 bf16 operands, fp32 accumulate in TMEM), TMA
//                         SW128 operand stages; the epilogue folds each S tile into y with the staged
//                         t_J and drops it.  Column range split in `nch` chunks -> deterministic
//                         partial sums ypart[chunk][row], no atomics.
//   pool_reduce_kernel /  cw = sum_chunks ypart; block partials of sum(cw) and (max, argmax);
//   pool_finish_kernel    conf = cw / sum (client.rs:422-437), voter agreement partials (438-449).
//
// bf16 operands: stated tolerance 1e-2 of the top confidence (north star: "wider tolerance for
// TF32/bf16 GEMM inputs").  Bound: tensor cores / L2 (intensity N/2 flop per byte).
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>

#include "score_tc.cuh"

namespace lwc {

constexpr int kPoolTM = 128;      // rows of S per tile (UMMA M)
constexpr int kPoolTN = 256;      // columns of S per tile (UMMA N)
constexpr int kPoolKS = 64;       // bf16 columns per pipeline stage = one 128-byte swizzle atom
constexpr int kPoolStages = 4;
constexpr int kPoolABytes = kPoolTM * 128;
constexpr int kPoolBBytes = kPoolTN * 128;
constexpr int kPoolStageBytes = kPoolABytes + kPoolBBytes;
constexpr int kPoolThreads = 256;

// ---------------------------------------------------------------- prep: normalise + tally vector
// one warp per own row
__global__ void __launch_bounds__(256)
pool_prep_kernel(const float *__restrict__ E, const float *__restrict__ V, const float *__restrict__ w,
                 __nv_bfloat16 *__restrict__ Ehat, float *__restrict__ t, long long n_own, int D, int M) {
    const long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= n_own) return;
    const float *e = E + (size_t)row * D;
    float ss = 0.f;
    const bool vec = (D % 4 == 0) && ((reinterpret_cast<uintptr_t>(e) & 15u) == 0);
    if (vec) {
        for (int k = lane * 4; k < D; k += 128) {
            const float4 x = *reinterpret_cast<const float4 *>(e + k);
            ss = fmaf(x.x, x.x, ss);
            ss = fmaf(x.y, x.y, ss);
            ss = fmaf(x.z, x.z, ss);
            ss = fmaf(x.w, x.w, ss);
        }
    } else {
        for (int k = lane; k < D; k += 32) ss = fmaf(e[k], e[k], ss);
    }
    ss = warp_sum(ss);
    const float rn = ss > 0.f ? 1.0f / sqrtf(ss) : 0.f; // zero-norm row: E^_i = 0
    __nv_bfloat16 *o = Ehat + (size_t)row * D;
    for (int k = lane; k < D; k += 32) o[k] = __float2bfloat16_rn(e[k] * rn);
    if (lane == 0) {
        float ti = 0.f;
        for (int m = 0; m < M; ++m) {
            const float v0 = V[(size_t)m * n_own];
            if (isnan(v0)) continue; // voter without a vote (client.rs:410)
            float wm = w[m];
            if (isnan(wm)) wm = 0.f; // weight None -> 0 (client.rs:412-413)
            ti = fmaf(V[(size_t)m * n_own + row], wm, ti);
        }
        t[row] = ti;
    }
}

// ---------------------------------------------------------------- the all-pairs pass
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *tm, int c0, int c1, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
        ::"r"(dst), "l"(tm), "r"(c0), "r"(c1), "r"(bar)
        : "memory");
}

// Instruction descriptor: D = F32, A = B = BF16 (1), both K-major.
__device__ __forceinline__ constexpr uint32_t umma_idesc_bf16(int M, int N) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                          uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}

struct PoolGemmParams {
    const float *t_all;   // [n_total] tally vector of every column
    float *ypart;         // [nch][n_own] partial row sums
    long long n_own;      // rows this rank owns
    long long n_total;    // columns = rows of the whole pool
    long long row0;       // global index of own row 0 (own rows are [row0, row0 + n_own) of E^_all)
    int D;
    int nch;              // column chunks in total (= nshards * cps)
    int jb_per_chunk;     // 256-column blocks per chunk
    // column shards: 1 (single GPU, or gathered beforehand by NCCL) or nranks (fused peer-to-peer gather)
    int nshards, rank;
    int cps;              // chunks per shard
    int jb_per_shard;     // 256-column blocks per shard
    // fused gather only (nshards > 1):
    const __nv_bfloat16 *const *peer_ehat; // [nranks] each rank's E^_all base, mapped into this process
    const float *const *peer_t;            // [nranks] each rank's t_all base
    __nv_bfloat16 *ehat_all;               // local copy being filled
    float *t_all_w;
    unsigned int *shard_done;              // [nranks] CTAs that have finished copying their slice of shard p
};

// unit u -> (row block I, column shard, chunk id for ypart, 256-column block range).  Units are ordered
// shard-major starting with the own shard, then chunk-major, so that all CTAs walk the same columns at
// about the same time (a B stage fetched from HBM by one CTA is an L2 hit for the others) and peer
// shards are needed in the order rank+1, rank+2, ... (each rank starts pulling from a different peer).
struct PoolUnit {
    int I, shard, ychunk, jb0, jb1;
};
__device__ __forceinline__ PoolUnit pool_unit(const PoolGemmParams &p, int u, int nI, int njb) {
    PoolUnit r;
    const int cidx = u / nI;
    r.I = u - cidx * nI;
    const int si = cidx / p.cps, within = cidx - si * p.cps;
    r.shard = (p.rank + si) % p.nshards;
    r.ychunk = r.shard * p.cps + within;
    r.jb0 = r.shard * p.jb_per_shard + within * p.jb_per_chunk;
    r.jb1 = min(min(njb, (r.shard + 1) * p.jb_per_shard), r.jb0 + p.jb_per_chunk);
    return r;
}

__device__ __forceinline__ unsigned int ld_acquire_u32(const unsigned int *p) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// tmA: bf16 [n_total][D] box {64, 128}; tmB: same tensor, box {64, 256}; both SWIZZLE_128B.
__global__ void __launch_bounds__(kPoolThreads, 1)
pool_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const PoolGemmParams p) {
    extern __shared__ __align__(1024) unsigned char smem_pool[];
    unsigned char *ring = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_pool) + 1023) & ~(uintptr_t)1023);
    float *t_s = reinterpret_cast<float *>(ring + (size_t)kPoolStages * kPoolStageBytes); // [2][256]
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(t_s + 512);
    uint64_t *bars = reinterpret_cast<uint64_t *>(tmem_slot + 2);
    const uint32_t full0 = smem_u32(bars), empty0 = full0 + 8 * kPoolStages;
    const uint32_t tfull0 = empty0 + 8 * kPoolStages, tempty0 = tfull0 + 16;

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nI = (int)((p.n_own + kPoolTM - 1) / kPoolTM);
    const int njb = (int)((p.n_total + kPoolTN - 1) / kPoolTN);
    const int nunits = nI * p.nch;
    const int nks = (p.D + kPoolKS - 1) / kPoolKS; // pipeline stages per tile

    if (threadIdx.x == 0) {
        for (int s = 0; s < kPoolStages; ++s) {
            mbar_init(full0 + 8 * s, 1);
            mbar_init(empty0 + 8 * s, 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(tfull0 + 8 * b, 1);
            mbar_init(tempty0 + 8 * b, 128);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) { // all 512 TMEM columns: two 128 x 256 fp32 accumulators
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    // unit u = chunk * nI + I: CTAs walk the same column chunk at about the same time, so a B stage
    // fetched from HBM by one CTA is an L2 hit for the others.
    if (warp == 0) {
        if (lane == 0) {
            const uint32_t ring_u32 = smem_u32(ring);
            int slot = 0;
            uint32_t phase = 0;
            int shard_seen = p.rank; // shards whose arrival this lane has already confirmed
            for (int u = blockIdx.x; u < nunits; u += gridDim.x) {
                const PoolUnit un = pool_unit(p, u, nI, njb);
                const int I = un.I, jb0 = un.jb0, jb1 = un.jb1;
                if (p.nshards > 1 && un.shard != shard_seen) {
                    // fused gather: the shard's rows are being pulled over NVLink by the gather warps of all CTAs
                    const long long t0 = clock64();
                    while (ld_acquire_u32(p.shard_done + un.shard) < gridDim.x)
                        if (clock64() - t0 > 8000000000LL) __trap();
                    asm volatile("fence.proxy.async;" ::: "memory"); // those generic-proxy stores -> visible to TMA
                    shard_seen = un.shard;
                }
                const int arow = (int)(p.row0 + (long long)I * kPoolTM);
                for (int jb = jb0; jb < jb1; ++jb)
                    for (int ks = 0; ks < nks; ++ks) {
                        mbar_wait(empty0 + 8 * slot, phase ^ 1u);
                        mbar_arrive_expect_tx(full0 + 8 * slot, (uint32_t)kPoolStageBytes);
                        const uint32_t dst = ring_u32 + slot * kPoolStageBytes;
                        tma_load_2d(dst, &tmA, ks * kPoolKS, arow, full0 + 8 * slot);
                        tma_load_2d(dst + kPoolABytes, &tmB, ks * kPoolKS, jb * kPoolTN, full0 + 8 * slot);
                        if (++slot == kPoolStages) {
                            slot = 0;
                            phase ^= 1u;
                        }
                    }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t ring_u32 = smem_u32(ring);
            const uint32_t idesc = umma_idesc_bf16(kPoolTM, kPoolTN);
            int slot = 0;
            uint32_t phase = 0, tile = 0;
            for (int u = blockIdx.x; u < nunits; u += gridDim.x) {
                const PoolUnit un = pool_unit(p, u, nI, njb);
                const int jb0 = un.jb0, jb1 = un.jb1;
                for (int jb = jb0; jb < jb1; ++jb, ++tile) {
                    const uint32_t ab = tile & 1u;
                    mbar_wait(tempty0 + 8 * ab, ((tile >> 1) & 1u) ^ 1u);
                    tc_fence_after();
                    const uint32_t d_tmem = tmem_base + ab * (uint32_t)kPoolTN;
                    uint32_t acc = 0;
                    for (int ks = 0; ks < nks; ++ks) {
                        mbar_wait(full0 + 8 * slot, phase);
                        tc_fence_after();
                        const uint32_t a0 = ring_u32 + slot * kPoolStageBytes, b0 = a0 + kPoolABytes;
#pragma unroll
                        for (int kk = 0; kk < 4; ++kk) { // 4 x K=16 inside the 128-byte swizzle atom
                            umma_bf16(d_tmem, umma_desc_sw128(a0 + kk * 32), umma_desc_sw128(b0 + kk * 32), idesc, acc);
                            acc = 1;
                        }
                        umma_commit(empty0 + 8 * slot);
                        if (++slot == kPoolStages) {
                            slot = 0;
                            phase ^= 1u;
                        }
                    }
                    umma_commit(tfull0 + 8 * ab);
                }
            }
        }
        __syncwarp();
    } else if (warp == 3) {
        // ===================== gather (fused collective): pull the peers' shards over NVLink =====================
        if (p.nshards > 1) {
            for (int si = 1; si < p.nshards; ++si) {
                const int src_rank = (p.rank + si) % p.nshards;
                const size_t row_off = (size_t)src_rank * p.n_own;
                // tally vector of the shard (CTA-strided)
                const float *tsrc = p.peer_t[src_rank] + row_off;
                float *tdst = p.t_all_w + row_off;
                for (long long k = (long long)blockIdx.x * 32 + lane; k < p.n_own; k += (long long)gridDim.x * 32)
                    tdst[k] = tsrc[k];
                // normalised rows: this CTA's slice, 16 bytes per lane and load, 8 loads in flight
                const uint4 *src = reinterpret_cast<const uint4 *>(p.peer_ehat[src_rank] + row_off * p.D);
                uint4 *dst = reinterpret_cast<uint4 *>(p.ehat_all + row_off * p.D);
                const long long n16 = (long long)p.n_own * p.D / 8;
                const long long b = n16 * blockIdx.x / gridDim.x, e = n16 * (blockIdx.x + 1) / gridDim.x;
                long long k = b + lane;
                for (; k + 7 * 32 < e; k += 8 * 32) {
                    uint4 v[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j) v[j] = src[k + j * 32];
#pragma unroll
                    for (int j = 0; j < 8; ++j) dst[k + j * 32] = v[j];
                }
                for (; k < e; k += 32) dst[k] = src[k];
                __threadfence(); // this lane's stores are visible device-wide before the counter moves
                __syncwarp();
                if (lane == 0) atomicAdd(p.shard_done + src_rank, 1u);
            }
        }
    } else if (warp >= 4) {
        const int i = threadIdx.x - 128;
        const int ew = warp - 4;
        const uint32_t lane_addr = tmem_base + ((uint32_t)(ew * 32) << 16);
        uint32_t tile = 0;
        for (int u = blockIdx.x; u < nunits; u += gridDim.x) {
            const PoolUnit un = pool_unit(p, u, nI, njb);
            const int chunk = un.ychunk, I = un.I, jb0 = un.jb0, jb1 = un.jb1;
            float y = 0.f;
            for (int jb = jb0; jb < jb1; ++jb, ++tile) {
                const uint32_t ab = tile & 1u;
                float *ts = t_s + ab * 256;
                mbar_wait(tfull0 + 8 * ab, (tile >> 1) & 1u);
                tc_fence_after();
                // stage t_J for this tile (its previous user, tile-2, finished before tempty was released).
                // After the tfull wait: with the fused gather, t of a peer shard exists only once the shard has
                // arrived, which the producer confirmed before loading this tile.  L2 loads: no stale L1 line.
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const long long j = (long long)jb * kPoolTN + h * 128 + i;
                    ts[h * 128 + i] = j < p.n_total ? __ldcg(p.t_all + j) : 0.f;
                }
                epi_sync();
                const uint32_t row_addr = lane_addr + ab * (uint32_t)kPoolTN;
                float v[32];
#pragma unroll 1
                for (int c = 0; c < kPoolTN / 32; ++c) {
                    tmem_ld32(row_addr + (uint32_t)(c * 32), v);
#pragma unroll
                    for (int e = 0; e < 32; ++e) y = fmaf(ts[c * 32 + e], v[e], y);
                }
                tc_fence_before();
                mbar_arrive(tempty0 + 8 * ab);
            }
            const long long row = (long long)I * kPoolTM + i;
            if (row < p.n_own) p.ypart[(size_t)chunk * p.n_own + row] = y;
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
    }
}

inline size_t pool_gemm_smem_bytes() { return 1024 + (size_t)kPoolStages * kPoolStageBytes + 512 * 4 + 8 + (2 * kPoolStages + 4) * 8 + 64; }

// ---------------------------------------------------------------- reductions (fixed order, no atomics)
constexpr int kPoolRedThreads = 256;
constexpr int kPoolRedPerBlock = 512;  // rows per block in the reduce/finish kernels (many small blocks: these passes are latency-bound)

// cw[row] = sum_chunks ypart; per-block partial sum and (max, first argmax) of cw
__global__ void __launch_bounds__(kPoolRedThreads)
pool_reduce_kernel(const float *__restrict__ ypart, int nch, long long n_own, long long row0, float *__restrict__ cw,
                   float *__restrict__ blk_sum, float *__restrict__ blk_max, long long *__restrict__ blk_arg) {
    __shared__ float s_sum[kPoolRedThreads / 32], s_max[kPoolRedThreads / 32];
    __shared__ long long s_arg[kPoolRedThreads / 32];
    const long long base = (long long)blockIdx.x * kPoolRedPerBlock;
    float sum = 0.f, best = -INFINITY;
    long long arg = 0x7fffffffffffffffLL;
    for (int k = threadIdx.x; k < kPoolRedPerBlock; k += kPoolRedThreads) {
        const long long row = base + k;
        if (row >= n_own) break;
        float c = 0.f;
        for (int ch = 0; ch < nch; ++ch) c += ypart[(size_t)ch * n_own + row];
        cw[row] = c;
        sum += c;
        if (c > best) { // rows visited in increasing order per thread: first maximum wins
            best = c;
            arg = row0 + row;
        }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    sum = warp_sum(sum);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ob = __shfl_xor_sync(kFullMask, best, o);
        const long long oa = __shfl_xor_sync(kFullMask, arg, o);
        if (ob > best || (ob == best && oa < arg)) {
            best = ob;
            arg = oa;
        }
    }
    if (lane == 0) {
        s_sum[warp] = sum;
        s_max[warp] = best;
        s_arg[warp] = arg;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        float ts = 0.f, tb = -INFINITY;
        long long ta = 0x7fffffffffffffffLL;
        for (int k = 0; k < kPoolRedThreads / 32; ++k) {
            ts += s_sum[k];
            if (s_max[k] > tb || (s_max[k] == tb && s_arg[k] < ta)) {
                tb = s_max[k];
                ta = s_arg[k];
            }
        }
        blk_sum[blockIdx.x] = ts;
        blk_max[blockIdx.x] = tb;
        blk_arg[blockIdx.x] = ta;
    }
}

// one block: fold the per-block partials -> out3 = {sum, max, argmax(as two floats' worth in arg_out)}
__global__ void __launch_bounds__(32)
pool_fold_kernel(const float *__restrict__ blk_sum, const float *__restrict__ blk_max, const long long *__restrict__ blk_arg,
                 int nblk, float *__restrict__ sum_out, float *__restrict__ max_out, long long *__restrict__ arg_out) {
    if (threadIdx.x != 0) return;
    float ts = 0.f, tb = -INFINITY;
    long long ta = 0x7fffffffffffffffLL;
    for (int k = 0; k < nblk; ++k) {
        ts += blk_sum[k];
        if (blk_max[k] > tb || (blk_max[k] == tb && blk_arg[k] < ta)) {
            tb = blk_max[k];
            ta = blk_arg[k];
        }
    }
    *sum_out = ts;
    *max_out = tb;
    *arg_out = ta;
}

// conf = cw / total (client.rs:428-437); per-block partial voter agreement sum_i conf_i V_mi (438-449)
__global__ void __launch_bounds__(kPoolRedThreads)
pool_finish_kernel(const float *__restrict__ cw, const float *__restrict__ total_ptr, const float *__restrict__ V, int M,
                   long long n_own, float *__restrict__ conf, float *__restrict__ blk_agree) {
    __shared__ float s_part[kPoolRedThreads / 32];
    const float total = *total_ptr;
    const bool positive = total > 0.f;
    const long long base = (long long)blockIdx.x * kPoolRedPerBlock;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int k = threadIdx.x; k < kPoolRedPerBlock; k += kPoolRedThreads) {
        const long long row = base + k;
        if (row < n_own && conf) conf[row] = positive ? cw[row] / total : 0.f;
    }
    for (int m = 0; m < M; ++m) {
        float a = 0.f;
        for (int k = threadIdx.x; k < kPoolRedPerBlock; k += kPoolRedThreads) {
            const long long row = base + k;
            if (row < n_own) {
                const float cf = positive ? cw[row] / total : 0.f;
                a = fmaf(cf, V[(size_t)m * n_own + row], a);
            }
        }
        a = warp_sum(a);
        if (lane == 0) s_part[warp] = a;
        __syncthreads();
        if (threadIdx.x == 0) {
            float s = 0.f;
            for (int k = 0; k < kPoolRedThreads / 32; ++k) s += s_part[k];
            blk_agree[(size_t)blockIdx.x * M + m] = s;
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(128)
pool_fold_agree_kernel(const float *__restrict__ blk_agree, int nblk, int M, const float *__restrict__ V, long long n_own,
                       float *__restrict__ agree) {
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= M) return;
    float s = 0.f;
    for (int k = 0; k < nblk; ++k) s += blk_agree[(size_t)k * M + m];
    agree[m] = isnan(V[(size_t)m * n_own]) ? __int_as_float(0x7fc00000) : s; // partial sum over own rows
}

} // namespace lwc
