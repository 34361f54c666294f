// score_tc16.cuh -- fp32-grade large-pool scoring (32 < N <= 128) on tcgen05 at kind::f16 rate, sm_100a.
//
// score_tc_kernel<true> reaches fp32-grade confidences with two kind::tf32 products per K step and is bound by
// the tf32 tensor rate (an M=128 N=256 K=8 tf32 MMA takes ~141 ns: 0.64 of HBM speed at cfg4).  On this part
// kind::f16 runs ~3x faster per K element (M=128 N=256 K=16: ~187 clk), and an fp16 value carries the same 11
// significant bits as a tf32 one.  So the same compensated scheme is run on fp16 pieces:
//
//     x = h + 2^-11 l,   h = fp16(x) (round to nearest),   l = fp16(2^11 (x - h))   [x - h is exact in fp32]
//     S = X + 2^-11 (Y + Y^T),   X = h h^T,   Y = h l^T            (the l l^T term, < 2^-22, is dropped)
//
// h and l together hold 22-23 bits of x; the 2^11 factor keeps l out of fp16's subnormal range whenever h is in
// its normal range.  fp16 x fp16 products are exact in the fp32 accumulator.  What fp16 lacks is exponent range
// (6e-5 .. 65504 normal), so the kernel CHECKS every request instead of assuming: a candidate whose squared norm
// comes out non-finite (an element beyond fp16's range), or below 2^-12 (elements so small that h is subnormal
// and the pieces hold fewer bits), or exactly zero (all elements may have underflowed) marks its request in
// `fix_flags`, and lwc_score_batch follows this kernel with score_tc_kernel<true> in fix-up mode, which re-scores
// exactly the marked requests with tf32 pieces (fp32's exponent range).  Unit-norm or O(1)-norm embeddings never
// take that pass.
//
// Layout and roles (one persistent CTA per SM, contiguous request ranges, as score_tc.cuh):
//   warp 0 (1 lane)    TMA producer: one box {32, 128, 2, 1} of the rank-4 view of E per stage = a UNIT of 64
//                      columns = two K-major SWIZZLE_128B fp32 slabs [128 rows x 128 B], 32 KB.
//   warps 2, 3, 8, 9   transform, IN PLACE: the stage's 32 KB become [h tile ; l tile], two fp16 K-major
//                      SWIZZLE_128B slabs [128 rows x 64 columns].  A warp owns 8-row atoms: the two fp32 atoms of
//                      rows 8g..8g+7 (2 KB) are read completely into registers, then overwritten by the h atom
//                      and the l atom of the same rows -- the warp's output footprint is its own input footprint,
//                      so no other warp's unread data is ever overwritten.  Lane (r = lane % 8, p = lane / 8)
//                      handles row r, columns 8p..8p+7 of each slab: all 128-bit accesses are conflict-free.
//   warp 1 (1 lane)    MMA issuer: per unit 4 x tcgen05.mma.kind::f16 (M = 128, N = 256, K = 16), A = the h tile,
//                      B = the 256-row pair [h ; l]: accumulator [X | Y], 2 x 256 TMEM columns double-buffered.
//   warps 4..7         epilogue, as score_tc_kernel<true>: X restarts every chunk of kT16ChunkUnits units (the
//                      tensor core's fp32 accumulator truncates; chunks are summed in registers; the restart is the
//                      chunk's first MMA issued with the accumulate flag off), Y runs through the request,
//                      S_ij = X_ij + 2^-11 (Y_ij + Y_ji) via the row / warp-transpose fold.
// Shared-memory traffic per unit: 32 KB TMA in, 32 + 32 KB transform, 16 + 32 KB operand reads = 144 KB against
// the 192 KB per 64 columns of the tf32 scheme; tensor time 4 x 187 clk against ~1,480 clk a unit has at HBM speed.
// Measured (cfg4, 8,192 requests): 2.61 ms against 3.03 ms on tf32 pieces and 1.91 ms for the single-pass TF32 route on
// the same box -- 0.77 of HBM.  Bound: the SHARED-MEMORY pipe (ncu, profiles/r2_tc16_raw.csv: LSU wavefronts 50 % +
// tensor operand reads 24 % + TMA writes; tensor pipe 33 %, DRAM 5.0 TB/s): 144 KB of shared-memory traffic per 32 KB of
// HBM traffic.  Tried and dropped: asking L2 for the units eight ahead of the ring (cp.async.bulk.prefetch.tensor) so
// that HBM latency stops depending on ring depth -- 2.80 ms, i.e. latency was not the limit.  Kept, for ~1 % each on
// cfg4 (2.606 -> 2.585 -> 2.579 ms on one box, scripts/r2_ab_tc16.sh): one "Y read" barrier per accumulator, the one the
// next request starts on released first; and no TMEM clearing stores at all (the first MMA of a chunk runs with the
// accumulate flag off) -- the request-boundary bubble was never the limit either.
#pragma once

#include <cuda_fp16.h>

#include "score_tc.cuh"

namespace lwc {

constexpr int kT16Threads = 320;
constexpr int kT16UnitBytes = 2 * kTcSlabBytes; // 64 columns: two fp32 slabs in, [h ; l] fp16 tiles out
constexpr int kT16ChunkUnits = 4;               // 16 MMAs (256 columns) per in-TMEM accumulation chain of X
constexpr int kT16MaxStages = 6;
// polling back-off (ns) of the producer / MMA issuer / transform warps: every poll is a shared-memory wavefront on the
// pipe that bounds this kernel
#ifndef LWC_T16_BO_P
#define LWC_T16_BO_P 40
#endif
#ifndef LWC_T16_BO_M
#define LWC_T16_BO_M 20
#endif
#ifndef LWC_T16_BO_T
#define LWC_T16_BO_T 40
#endif
constexpr unsigned kT16BackoffP = LWC_T16_BO_P, kT16BackoffM = LWC_T16_BO_M, kT16BackoffT = LWC_T16_BO_T;
constexpr float kT16LoScale = 1.0f / 2048.0f;   // Y carries 2^11
constexpr float kT16MinNorm2 = 1.0f / 4096.0f;  // squared norm below which fp16 pieces may have lost bits

// Instruction descriptor: D = F32 (1 << 4), A = B = F16 (0), both K-major, N >> 3 at bit 17, M >> 4 at bit 24.
__device__ __forceinline__ constexpr uint32_t umma_idesc_f16(int M, int N) {
    return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// acc = 0: D = A B^T (the accumulator's old contents are ignored); acc = 1: D += A B^T
__device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(acc)
        : "memory");
}

// 8 fp32 values -> 8 fp16 h (one 16-byte chunk) and 8 fp16 l
__device__ __forceinline__ void split_f16x8(const float4 &a, const float4 &b, uint4 &h, uint4 &l) {
    const float x[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    uint32_t hp[4], lp[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const __half2 h2 = __floats2half2_rn(x[2 * i], x[2 * i + 1]);
        const float2 hf = __half22float2(h2);
        // (x - h) * 2^11, exact: x - h is exact in fp32 and the scaling is a power of two
        const float l0 = fmaf(x[2 * i], 2048.0f, -2048.0f * hf.x);
        const float l1 = fmaf(x[2 * i + 1], 2048.0f, -2048.0f * hf.y);
        const __half2 l2 = __floats2half2_rn(l0, l1);
        hp[i] = *reinterpret_cast<const uint32_t *>(&h2);
        lp[i] = *reinterpret_cast<const uint32_t *>(&l2);
    }
    h = make_uint4(hp[0], hp[1], hp[2], hp[3]);
    l = make_uint4(lp[0], lp[1], lp[2], lp[3]);
}

// PAIR: pools of at most 64 candidates ride two to a tile.  The tensor map is the rank-4 view {32, N, R, D/32} and one
// box {32, 64, 2, 2} lands requests 2g and 2g+1 as rows 0..63 and 64..127 of the same two slabs; the MMA is unchanged
// (its off-diagonal 64 x 64 blocks are the two requests' cross products and are never read), every epilogue thread
// works on its own request's half of the accumulator columns, and each per-request reduction runs over the two warps
// that own that half.  A 64-candidate pool then costs the shared-memory pipe half of what it costs alone in a tile.
template <bool PAIR>
__global__ void __launch_bounds__(kT16Threads, 1)
score_tc16_kernel(const __grid_constant__ CUtensorMap tmE, const ScoreParams p, const TcGeom geo, int *__restrict__ fix_flags) {
    extern __shared__ __align__(1024) unsigned char smem_t16[];
    const int S = geo.stages;
    constexpr int NTW = 4;
    unsigned char *ring = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_t16) + 1023) & ~(uintptr_t)1023);
    float *u_s = reinterpret_cast<float *>(ring + (size_t)S * kT16UnitBytes); // [128] t_i / n_i
    float *cf_s = u_s + 128;                                                 // [128] confidences
    float *part = cf_s + 128;                                                // [16]
    int *ipart = reinterpret_cast<int *>(part + 16);                         // [8]: [0..3] argmax, [4..7] range check
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(ipart + 8);           // [2]
    uint64_t *bars = reinterpret_cast<uint64_t *>(tmem_slot + 2);            // full, empty, ready [kT16MaxStages]; tfull, tempty, yempty [2]
    const uint32_t full0 = smem_u32(bars), empty0 = full0 + 8 * kT16MaxStages, ready0 = empty0 + 8 * kT16MaxStages;
    const uint32_t tfull0 = ready0 + 8 * kT16MaxStages, tempty0 = tfull0 + 16, yempty0 = tempty0 + 16;
    float *col_s = reinterpret_cast<float *>(bars + 3 * kT16MaxStages + 6);  // [4][128]

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int N = p.N, D = p.D, M = p.M;
    const int ntiles = PAIR ? (p.R + 1) >> 1 : p.R; // a tile holds one request, or a pair of them
    const int r_begin = (int)(((long long)blockIdx.x * ntiles) / gridDim.x);
    const int r_end = (int)(((long long)(blockIdx.x + 1) * ntiles) / gridDim.x);
    const int nreq = r_end - r_begin;
    const int upr = (D + 63) >> 6; // units per request (D % 32 == 0; a last half unit arrives zero-filled)

    if (threadIdx.x == 0) {
        for (int s = 0; s < S; ++s) {
            mbar_init(full0 + 8 * s, 1);
            mbar_init(empty0 + 8 * s, 1);
            mbar_init(ready0 + 8 * s, NTW);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(tfull0 + 8 * b, 1);
            mbar_init(tempty0 + 8 * b, 128);
            mbar_init(yempty0 + 8 * b, 128); // one per accumulator: its Y half drained and zeroed (previous request)
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    constexpr uint32_t kTmemCols = 512; // two accumulators [X | Y] of 128 x 256 fp32
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // Nothing in TMEM is ever cleared by a store: the first MMA of a chunk starts X afresh (accumulate flag off), and Y
    // starts afresh with each accumulator's first chunk of a request (see the MMA issuer).

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            const uint32_t ring_u32 = smem_u32(ring);
            int slot = 0;
            uint32_t phase = 0;
            for (int q = 0; q < nreq; ++q)
                for (int u = 0; u < upr; ++u) {
                    mbar_wait_backoff(empty0 + 8 * slot, phase ^ 1u, kT16BackoffP);
                    mbar_arrive_expect_tx(full0 + 8 * slot, (uint32_t)kT16UnitBytes); // OOB zero-fill counts too
                    if (PAIR)
                        tma_load_4d(ring_u32 + slot * kT16UnitBytes, &tmE, 0, 0, 2 * (r_begin + q), 2 * u, full0 + 8 * slot);
                    else
                        tma_load_4d(ring_u32 + slot * kT16UnitBytes, &tmE, 0, 0, 2 * u, r_begin + q, full0 + 8 * slot);
                    if (++slot == S) {
                        slot = 0;
                        phase ^= 1u;
                    }
                }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            const uint32_t ring_u32 = smem_u32(ring);
            const uint32_t idesc = umma_idesc_f16(128, 256), idesc_half = umma_idesc_f16(128, 128);
            int slot = 0;
            uint32_t phase = 0;
            uint32_t tile = 0; // accumulator hand-offs so far: one per chunk
            for (int q = 0; q < nreq; ++q) {
                uint32_t ab = tile & 1u;
                mbar_wait(tempty0 + 8 * ab, ((tile >> 1) & 1u) ^ 1u);
                tc_fence_after();
                uint32_t d_tmem = tmem_base + ab * 256u;
                // the Y half of THIS accumulator drained and zeroed by the epilogue of the previous request (it drains
                // the half the next request starts on first, so that this wait is as short as it can be)
                mbar_wait(yempty0 + 8 * ab, ((uint32_t)q & 1u) ^ 1u);
                tc_fence_after();
                for (int u = 0; u < upr; ++u) {
                    const bool chunk_start = u % kT16ChunkUnits == 0;
                    if (u > 0 && chunk_start) { // chunk boundary: hand over, switch accumulator
                        umma_commit(tfull0 + 8 * ab);
                        ++tile;
                        ab = tile & 1u;
                        mbar_wait(tempty0 + 8 * ab, ((tile >> 1) & 1u) ^ 1u);
                        if (u == kT16ChunkUnits) mbar_wait(yempty0 + 8 * ab, ((uint32_t)q & 1u) ^ 1u); // first use of the other accumulator
                        tc_fence_after();
                        d_tmem = tmem_base + ab * 256u;
                    }
                    mbar_wait_backoff(ready0 + 8 * slot, phase, kT16BackoffM); // the stage holds [h ; l]
                    tc_fence_after();
                    const uint32_t unit = ring_u32 + slot * kT16UnitBytes;
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk) { // 4 x K = 16 inside the 128-byte swizzle atom
                        const uint64_t d = umma_desc_sw128(unit + kk * 32);
                        if (kk == 0 && chunk_start) {
                            // X restarts with every chunk; Y with this accumulator's first chunk of the request (chunks 0
                            // and 1).  Later chunks: X afresh, Y on top of what the accumulator holds -- two half-width MMAs.
                            if (u < 2 * kT16ChunkUnits) {
                                umma_f16(d_tmem, d, d, idesc, 0u);
                            } else {
                                umma_f16(d_tmem, d, d, idesc_half, 0u);                                                     // X = h h^T
                                umma_f16(d_tmem + 128u, d, umma_desc_sw128(unit + kTcSlabBytes + kk * 32), idesc_half, 1u); // Y += h l^T
                            }
                        } else {
                            umma_f16(d_tmem, d, d, idesc, 1u); // A: 128 rows = h; B: 256 rows = [h ; l]
                        }
                    }
                    umma_commit(empty0 + 8 * slot);
                    if (++slot == S) {
                        slot = 0;
                        phase ^= 1u;
                    }
                }
                umma_commit(tfull0 + 8 * ab);
                ++tile;
            }
        }
        __syncwarp();
    } else if (warp == 2 || warp == 3 || warp >= 8) {
        // ===================== transform, in place: fp32 slabs -> [h ; l] fp16 tiles =====================
        const int tw = warp < 4 ? warp - 2 : warp - 6; // 0..3
        const int r = lane & 7, pq = lane >> 3;        // row within the 8-row atom, 8-column group within a slab
        // byte offsets inside a 1 KB atom (8 rows x 128 B, 16-byte chunks XOR-swizzled by the row)
        const uint32_t in0 = (uint32_t)(r * 128 + (((2 * pq) ^ r) << 4));     // columns 8pq .. 8pq+3
        const uint32_t in1 = (uint32_t)(r * 128 + (((2 * pq + 1) ^ r) << 4)); // columns 8pq+4 .. 8pq+7
        const uint32_t out0 = (uint32_t)(r * 128 + ((pq ^ r) << 4));          // fp16 chunk of slab 0's columns
        const uint32_t out1 = (uint32_t)(r * 128 + (((4 + pq) ^ r) << 4));    // fp16 chunk of slab 1's columns
        // rows past the pool arrive zero-filled, and zero fp32 bits already are zero h and l pieces: their atoms are left alone
        const int natoms = (N + 7) >> 3;
        constexpr int kAtomMask = PAIR ? 7 : 15;
        bool on[16 / NTW];
#pragma unroll
        for (int k = 0; k < 16 / NTW; ++k) on[k] = ((tw + k * NTW) & kAtomMask) < natoms;
        int slot = 0;
        uint32_t phase = 0;
        for (int q = 0; q < nreq; ++q)
            for (int u = 0; u < upr; ++u) {
                mbar_wait_backoff(full0 + 8 * slot, phase, kT16BackoffT);
                const uint32_t base = smem_u32(ring) + (uint32_t)slot * kT16UnitBytes;
                // all four atom pairs of this warp are read first (16 x 128-bit loads in flight per lane), then converted
                // and written back: one atom pair at a time left the warp waiting on its own loads and dependency chains
                // (round 2's ncu: the transform warps were busy ~80 % of the time at an IPC of 0.25)
                float4 x[16 / NTW][4];
#pragma unroll
                for (int k = 0; k < 16 / NTW; ++k) {
                    if (!on[k]) continue;
                    const uint32_t a0 = base + (uint32_t)(tw + k * NTW) * 1024u, a1 = a0 + (uint32_t)kTcSlabBytes;
                    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(x[k][0].x), "=f"(x[k][0].y), "=f"(x[k][0].z), "=f"(x[k][0].w) : "r"(a0 + in0));
                    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(x[k][1].x), "=f"(x[k][1].y), "=f"(x[k][1].z), "=f"(x[k][1].w) : "r"(a0 + in1));
                    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(x[k][2].x), "=f"(x[k][2].y), "=f"(x[k][2].z), "=f"(x[k][2].w) : "r"(a1 + in0));
                    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(x[k][3].x), "=f"(x[k][3].y), "=f"(x[k][3].z), "=f"(x[k][3].w) : "r"(a1 + in1));
                }
                __syncwarp(); // every lane has read its atoms before any lane overwrites them
#pragma unroll
                for (int k = 0; k < 16 / NTW; ++k) {
                    if (!on[k]) continue;
                    const uint32_t a0 = base + (uint32_t)(tw + k * NTW) * 1024u, a1 = a0 + (uint32_t)kTcSlabBytes;
                    uint4 h0, l0, h1, l1;
                    split_f16x8(x[k][0], x[k][1], h0, l0);
                    split_f16x8(x[k][2], x[k][3], h1, l1);
                    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a0 + out0), "r"(h0.x), "r"(h0.y), "r"(h0.z), "r"(h0.w) : "memory");
                    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a0 + out1), "r"(h1.x), "r"(h1.y), "r"(h1.z), "r"(h1.w) : "memory");
                    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a1 + out0), "r"(l0.x), "r"(l0.y), "r"(l0.z), "r"(l0.w) : "memory");
                    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a1 + out1), "r"(l1.x), "r"(l1.y), "r"(l1.z), "r"(l1.w) : "memory");
                }
                // generic-proxy writes must be visible to the tensor core (async proxy) before it is told
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncwarp();
                if (lane == 0) mbar_arrive(ready0 + 8 * slot);
                if (++slot == S) {
                    slot = 0;
                    phase ^= 1u;
                }
            }
    } else if (warp >= 4 && warp < 8) {
        // ===================== epilogue =====================
        const int i = threadIdx.x - 128; // accumulator row
        const int ew = warp - 4;
        constexpr int NCH = PAIR ? 2 : 4;              // 32-column chunks of a request's accumulator columns
        const int hq = PAIR ? (ew >> 1) : 0;           // which request of the tile this row belongs to
        const int ii = PAIR ? (i & 63) : i;            // candidate index
        const int cbase = hq * 64;                     // first accumulator column / row of that request
        const int w0 = PAIR ? 2 * hq : 0;              // first of the epilogue warps that own the request
        const uint32_t lane_addr = tmem_base + ((uint32_t)(ew * 32) << 16);
        uint32_t tile = 0;
        const int nchunk = (upr + kT16ChunkUnits - 1) / kT16ChunkUnits;
        for (int q = 0; q < nreq; ++q) {
            const int r = PAIR ? 2 * (r_begin + q) + hq : r_begin + q;
            const bool exists = r < p.R; // an odd batch ends on a half-empty tile
            const bool act = exists && ii < N;
            const float *Vr = p.V + (size_t)r * M * N;
            const float *wr = p.w + (p.w_per_request ? (size_t)r * M : 0);
            // t_i = sum_m w_m V_mi (client.rs:410-415), eight voters' loads in flight at a time
            float t = 0.f;
            for (int m0 = 0; m0 < M; m0 += 8) {
                float v0[8], wv[8], vi[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const int m = m0 + u;
                    const bool ok = m < M && exists;
                    v0[u] = ok ? __ldg(Vr + (size_t)m * N) : __int_as_float(0x7fc00000);
                    wv[u] = ok ? __ldg(wr + m) : 0.f;
                    vi[u] = (ok && act) ? __ldg(Vr + (size_t)m * N + ii) : 0.f;
                }
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    if (isnan(v0[u])) continue;
                    const float wm = isnan(wv[u]) ? 0.f : wv[u];
                    if (act) t = fmaf(vi[u], wm, t);
                }
            }
            float v[32];
            float row[NCH * 32]; // this candidate's row of S over its own request's columns
#pragma unroll
            for (int e = 0; e < NCH * 32; ++e) row[e] = 0.f;
            const uint32_t first_ab = tile & 1u; // accumulator of this request's even chunks (see score_tc.cuh)
            for (int ch = 0; ch < nchunk; ++ch, ++tile) {
                const uint32_t ab = tile & 1u;
                mbar_wait_backoff(tfull0 + 8 * ab, (tile >> 1) & 1u, 200);
                tc_fence_after();
                const uint32_t row_addr = lane_addr + ab * 256u; // X half
#pragma unroll
                for (int c = 0; c < NCH; ++c) { // the request's own columns
                    tmem_ld32(row_addr + (uint32_t)(cbase + c * 32), v);
#pragma unroll
                    for (int e = 0; e < 32; ++e) row[c * 32 + e] += v[e];
                }
                if (ch + 1 < nchunk) { // X read: the accumulator's next chunk may overwrite it
                    tc_fence_before();
                    mbar_arrive(tempty0 + 8 * ab);
                }
            }
            const uint32_t last_ab = (tile - 1u) & 1u;
            float yd = 0.f;
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
                if (hh == 1 && nchunk == 1) break; // a one-chunk request used one accumulator: the other holds nothing of it
                const uint32_t h = (uint32_t)hh ^ first_ab;
                tmem_ld32(lane_addr + h * 256u + 128u + (uint32_t)(ew * 32), v);
#pragma unroll
                for (int e = 0; e < 32; ++e) yd += (lane == e) ? v[e] : 0.f;
            }
            float d = 0.f;
#pragma unroll
            for (int c = 0; c < NCH; ++c)
#pragma unroll
                for (int e = 0; e < 32; ++e) d = (c == ew - w0 && lane == e) ? row[c * 32 + e] : d;
            d = fmaf(2.f * kT16LoScale, yd, d); // S_ii = X_ii + 2^-11 * 2 Y_ii
            // range check (see the header): this candidate's norm must sit where fp16 pieces are exact enough
            const bool bad = act && !(d >= kT16MinNorm2 && d < __int_as_float(0x7f800000));
            const float rn = (act && d > 0.f) ? 1.0f / sqrtf(d) : 0.f;
            const float ui = t * rn;
            u_s[i] = ui;
            const float uis = ui * kT16LoScale;
            // The accumulator the NEXT request starts on (the one that is not last_ab) is read first and handed back at
            // once; the other one's fold then overlaps that request's first chunk.  With an even chunk count that is the
            // first_ab-then-other order of score_tc.cuh; with an odd count the halves swap -- either way the order is a
            // function of the request's own chunk parities, not of its position in the batch.
            float cs[NCH];
#pragma unroll
            for (int c = 0; c < NCH; ++c) cs[c] = 0.f;
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
                const uint32_t h = (uint32_t)hh ^ last_ab ^ 1u;
                if (hh == 1 || nchunk > 1) { // a one-chunk request never touched the other accumulator
#pragma unroll
                    for (int c = 0; c < NCH; ++c) {
                        tmem_ld32(lane_addr + h * 256u + 128u + (uint32_t)(cbase + c * 32), v);
#pragma unroll
                        for (int e = 0; e < 32; ++e) {
                            row[c * 32 + e] = fmaf(v[e], kT16LoScale, row[c * 32 + e]);
                            v[e] *= uis;
                        }
                        cs[c] += warp_transpose_sum<32>(v, lane);
                    }
                }
                tc_fence_before();
                if (hh == 1) mbar_arrive(tempty0 + 8 * last_ab); // the last chunk's X was read in the chunk loop
                mbar_arrive(yempty0 + 8 * h);
            }
#pragma unroll
            for (int c = 0; c < NCH; ++c) col_s[ew * 128 + c * 32 + lane] = cs[c]; // columns counted from the request's first
            const bool warp_bad = __any_sync(kFullMask, bad);
            if (lane == 0) ipart[4 + ew] = warp_bad ? 1 : 0;
            epi_sync();
            const bool fix = PAIR ? (ipart[4 + w0] | ipart[5 + w0]) != 0 : (ipart[4] | ipart[5] | ipart[6] | ipart[7]) != 0;
            if (ii == 0 && exists && fix_flags) fix_flags[r] = fix ? 1 : 0;
            float acc = 0.f;
#pragma unroll
            for (int e = 0; e < NCH * 32; ++e) acc = fmaf(u_s[cbase + e], row[e], acc);
            if (PAIR)
                acc += col_s[w0 * 128 + ii] + col_s[(w0 + 1) * 128 + ii];
            else
                acc += (col_s[i] + col_s[128 + i]) + (col_s[256 + i] + col_s[384 + i]);
            float c = act ? acc * rn : 0.f;
            // ---- tally (client.rs:422-449) with fixed-order block reductions
            float ws = warp_sum(c);
            if (lane == 0) part[ew] = ws;
            epi_sync();
            const float sum = PAIR ? part[w0] + part[w0 + 1] : (part[0] + part[1]) + (part[2] + part[3]);
            const bool positive = sum > 0.f;
            const float cf = (positive && act) ? c / sum : 0.f;
            cf_s[i] = cf;
            const bool store = !fix && exists; // a marked request is written by the fix-up pass
            if (act && store) {
                if (p.cw) p.cw[(size_t)r * N + ii] = c;
                if (p.conf) p.conf[(size_t)r * N + ii] = cf;
            }
            if (p.winner) {
                float bv = act ? cf : -1.f;
                int bi = act ? ii : 0x7fffffff;
                warp_argmax(bv, bi);
                if (lane == 0) {
                    part[4 + ew] = bv;
                    ipart[ew] = bi;
                }
            }
            epi_sync();
            if (p.winner && ii == 0 && store) {
                float best = part[4 + w0];
                int besti = ipart[w0];
#pragma unroll
                for (int k = 1; k < (PAIR ? 2 : 4); ++k)
                    if (part[4 + w0 + k] > best || (part[4 + w0 + k] == best && ipart[w0 + k] < besti)) {
                        best = part[4 + w0 + k];
                        besti = ipart[w0 + k];
                    }
                p.winner[r] = positive ? besti : -1;
            }
            if (p.rank && act && store) {
                int pos = 0;
                for (int j = 0; j < N; ++j) {
                    const float cj = cf_s[cbase + j];
                    pos += (cj > cf || (cj == cf && j < ii)) ? 1 : 0;
                }
                p.rank[(size_t)r * N + pos] = ii;
            }
            if (p.agree && store) {
                constexpr int NW = PAIR ? 2 : 4; // warps sharing this request's voters
                for (int m0 = ew - w0; m0 < M; m0 += 4 * NW) {
                    float v0[4], s4[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const int m = m0 + NW * u;
                        v0[u] = m < M ? __ldg(Vr + (size_t)m * N) : 0.f;
                        s4[u] = 0.f;
                    }
                    for (int j = lane; j < N; j += 32) {
                        const float cj = cf_s[cbase + j];
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const int m = m0 + NW * u;
                            if (m < M) s4[u] = fmaf(cj, __ldg(Vr + (size_t)m * N + j), s4[u]);
                        }
                    }
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const int m = m0 + NW * u;
                        if (m >= M) break;
                        const float a = isnan(v0[u]) ? __int_as_float(0x7fc00000) : warp_sum(s4[u]);
                        if (lane == 0) p.agree[(size_t)r * M + m] = a;
                    }
                }
            }
            epi_sync(); // u_s / cf_s / part / ipart are reused by the next request
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
    }
}

inline size_t tc16_smem_bytes(int stages) {
    return 1024 + (size_t)stages * kT16UnitBytes + (128 + 128 + 16 + 8 + 2) * 4 + (3 * kT16MaxStages + 6) * 8 + 4 * 128 * 4 + 64;
}

} // namespace lwc
