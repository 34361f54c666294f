// score_tc.cuh -- large-pool scoring kernel (32 < N <= 128): the similarity matrix on the 5th-gen
// tensor cores (tcgen05.mma kind::tf32, fp32 accumulate in TMEM), sm_100a.
//
// One persistent CTA per SM, each owning a contiguous range of requests; per request the whole
// pool is ONE 128 x 128 UMMA tile: S = E E^T with A = B = the request's [N][D] tile.
//
//   warp 0 (1 lane)  TMA producer.  E is described as a rank-4 fp32 tensor {32, N, D/32, R}
//                    (k_in, row, k_out, request); one cp.async.bulk.tensor.4d moves a box
//                    {32, 128, KO, 1} = KO operand slabs of [128 rows x 128 B], 128-byte swizzled,
//                    i.e. exactly the K-major SWIZZLE_128B canonical layout UMMA reads.  Rows >= N
//                    are out of bounds and arrive as zeros.  tools/tma_probe2.cu
//                    (profiles/r1_tma_probe2.txt): 32 KB boxes x 4 stages stream at 6.0 TB/s,
//                    64 KB boxes x 3 stages at 6.58 TB/s.
//   warp 1 (1 lane)  MMA issuer.  Per slab 4 x tcgen05.mma (M=128, N=128, K=8), the SAME shared
//                    tile as A and as B descriptor (X3: N=256, see below); tcgen05.commit releases the stage; the last
//                    commit of a request signals the epilogue.  Two 128-column accumulators in
//                    TMEM alternate, so request q+1 accumulates while q is being finalised.
//   warps 4..7       epilogue: thread i owns TMEM lane i = row i of S (tcgen05.ld 32x32b).
//                    n_i^2 = S_ii; u_i = t_i / n_i goes to shared memory; because S is symmetric,
//                    cw_i = (1/n_i) * sum_j u_j S_ij needs only the thread's own row.  Then the
//                    tally of client.rs:422-449 with block reductions, argmax and rank.
//
// X3 = false: inputs are read by the tensor core as TF32 (the low 13 mantissa bits of each fp32 are
// ignored) -- the `precision = LWC_PREC_TF32` route with the stated 2e-3 confidence tolerance.
// X3 = true: fp32-grade result from two tf32 products per K step in ONE instruction.  With hi = trunc_tf32(x)
// (what the tensor core sees of an fp32 operand -- measured: it truncates) and lo = x - hi,
// S = X + Y + Y^T where X = hi.hi^T and Y = hi.lo^T (lo.lo^T < 2^-20 is dropped).  The third product of the
// classic 3xTF32 scheme is the transpose of the second, so it is never computed: the epilogue folds Y in
// twice, once along its row and once, through a warp transpose-sum, along its column.  Four "transform" warps
// (2, 3, 8, 9) write lo of every landed slab into a twin slab right behind it (same swizzled layout, so the
// transform is purely elementwise), and the issuer runs one M = 128, N = 256 MMA per K step whose B operand is
// the 256-row pair [raw slab ; lo slab]: the accumulator tile is [X | Y] (2 x 256 TMEM columns, double buffered).
// The raw slabs of a stage arrive as one-slab TMA boxes so that each can sit in front of its twin.
// The tensor core's fp32 accumulator truncates when it aligns addends: over the MMAs of a D = 3072 request
// the diagonal drifted by ~1e-5 relative (measured), too much for the 1e-5 bar.  So X accumulates in CHUNKS
// of kTcChunkSlabs slabs: the two accumulators alternate per chunk and the epilogue threads add each finished X
// into fp32 registers (round-to-nearest adds), which bounds every in-TMEM chain to 32 MMAs.  Y is ~2^-11 of X,
// so its drift does not matter and it must NOT restart per chunk (there are no registers left to sum it in):
// the MMAs therefore always accumulate, and the epilogue zeroes what has to restart with tcgen05.st -- the X
// half after each chunk, the two Y halves (even chunks' Y in one accumulator, odd chunks' in the other) after
// reading them at the end of the request.
// Because the epilogue threads are on the accumulator hand-off path every eight slabs, their own global loads
// (votes, weights) are issued in batches: as a load-test-load chain per voter they stalled the MMA issuer.
// This route is TENSOR-bound: an M=128 N=256 K=8 kind::tf32 MMA takes ~141 ns whatever the SM clock (276 clk
// at 1.965 GHz; 550 TFLOP/s per GPU; ncu's tf32 peak is 1,024 FMA/clk/SM, a quarter of bf16), so the four MMAs
// of a 16 KB slab take ~564 ns against the ~345 ns a slab has at HBM speed.
// History (cfg4, 8,192 requests): three N = 128 MMAs, all operands in shared memory 4.02 ms -> two MMAs with
// the symmetric fold 3.6 ms -> A operands from tensor memory (tcgen05.st by row-loader warps, tcgen05.mma with
// a TMEM A) 3.6 ms, which showed that shared-memory bandwidth was not the limit -> one N = 256 MMA 3.36-3.52 ms
// -> batched epilogue loads 2.93-3.11 ms = the tensor roofline of the scheme.  Moving the correction product to
// kind::f16 (bf16 lo as a TMEM A operand x bf16 hi) was as fast but used up to 1.3x the stated tolerance.
// Bound: TF32 route HBM (tensor pipe ~85 % busy at HBM speed); X3 route tensor.
#pragma once

#include <cuda.h>

#include "score_stream.cuh"

namespace lwc {

constexpr int kTcThreads = 256;   // TF32 route: producer, issuer, TMEM allocator, spare, 4 epilogue warps
constexpr int kTcThreadsX3 = 320; // + warps 8, 9; warps 2, 3, 8, 9 transform
constexpr int kTcSlabBytes = 128 * 128;  // one [128 rows x 32 fp32] operand slab, SWIZZLE_128B
constexpr int kTcMaxStages = 8;
constexpr int kTcChunkSlabs = 8;  // X3: operand slabs (32 MMAs into X) per in-TMEM accumulation chunk
#ifndef LWC_TC_ROUND
#define LWC_TC_ROUND 0u
#endif
// what the tensor core does to an fp32 operand of kind::tf32: 0 = truncate (measured: assuming round-to-
// nearest, 0x1000, made the compensated route 20x less accurate)
constexpr uint32_t kTcRound = LWC_TC_ROUND;

struct TcGeom {
    int ko;     // operand slabs (32 columns each) per TMA box / pipeline stage
    int stages; // ring depth
};

__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap *tm, int c0, int c1, int c2, int c3,
                                            uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];"
        ::"r"(dst), "l"(tm), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(bar)
        : "memory");
}

// K-major SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor layout):
// start address >> 4 | LBO (=1, unused for swizzled K-major) << 16 | SBO (1024 B between 8-row
// groups) >> 4 << 32 | version 1 << 46 | layout SWIZZLE_128B (2) << 61.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
    return (uint64_t)((smem_addr & 0x3FFFFu) >> 4) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) |
           (2ull << 61);
}

// Instruction descriptor: D = F32 (1<<4), A = B = TF32 (2<<7, 2<<10), both K-major, N>>3 at bit 17, M>>4 at bit 24.
__device__ __forceinline__ constexpr uint32_t umma_idesc_tf32(int M, int N) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                          uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// 32 consecutive columns of this thread's TMEM lane <- registers.
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
          "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]),
          "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]),
          "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
        : "memory");
}
// All MMAs issued so far by this thread arrive on `bar` when they complete (implies fence::before_thread_sync).
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// 32 consecutive accumulator columns of this thread's TMEM lane.
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

// named barrier over the 128 epilogue threads only
__device__ __forceinline__ void epi_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

// PAIR (TF32 route only): pools of at most 64 candidates ride two to a tile -- the tensor map is the view
// {32, N, R, D/32}, one box {32, 64, 2, KO} lands requests 2g and 2g+1 as rows 0..63 and 64..127 of the same slabs, the
// MMA is unchanged (its off-diagonal 64 x 64 blocks are cross products nobody reads) and every per-request reduction
// of the epilogue runs over the two warps that own the request's rows (see score_tc16.cuh for the fp32-grade twin).
template <bool X3, bool PAIR = false>
__global__ void __launch_bounds__(X3 ? kTcThreadsX3 : kTcThreads, 1)
score_tc_kernel(const __grid_constant__ CUtensorMap tmE, const ScoreParams p, const TcGeom geo, const int *__restrict__ fix_flags) {
    static_assert(!(X3 && PAIR), "the fp32-grade pair route is score_tc16_kernel<true>");
    // fix_flags != nullptr: fix-up pass behind score_tc16_kernel (score_tc16.cuh) -- only the requests it marked are
    // scored here; every role skips the others, so the pipelines stay in step
    extern __shared__ __align__(1024) unsigned char smem_tc[];
    const int KO = geo.ko, S = geo.stages;
    // TF32 route: a stage is one TMA box of KO slabs.  X3: a stage is KO pairs [raw slab][lo slab], each pair
    // the 256-row B operand of one N = 256 MMA; the raw slabs arrive as KO one-slab boxes.
    constexpr int kSlabStride = (X3 ? 2 : 1) * kTcSlabBytes;
    const int stageb = KO * kSlabStride;
    constexpr int NTW = 4;                         // transform warps (X3 only)
    // dynamic shared memory is only guaranteed 16-byte aligned: SWIZZLE_128B tiles need 1024
    unsigned char *ring = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_tc) + 1023) & ~(uintptr_t)1023);
    float *u_s = reinterpret_cast<float *>(ring + (size_t)S * stageb); // [128] t_i / n_i
    float *cf_s = u_s + 128;                                          // [128] confidences
    float *part = cf_s + 128;                                         // [16] block-reduction scratch
    int *ipart = reinterpret_cast<int *>(part + 16);                  // [8]
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(ipart + 8);    // [2]
    uint64_t *bars = reinterpret_cast<uint64_t *>(tmem_slot + 2);     // full[S], empty[S], tfull[2], tempty[2], ready[S], yempty[2]
    const uint32_t full0 = smem_u32(bars), empty0 = full0 + 8 * kTcMaxStages;
    const uint32_t tfull0 = empty0 + 8 * kTcMaxStages, tempty0 = tfull0 + 16;
    const uint32_t ready0 = tempty0 + 16; // X3: the stage's lo slabs are written and visible to the tensor core
    const uint32_t yempty0 = ready0 + 8 * kTcMaxStages; // X3: both Y accumulators have been read and zeroed
    float *col_s = reinterpret_cast<float *>(bars + 3 * kTcMaxStages + 6); // X3: [4][128] column sums of u_i Y_ij per warp

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int N = p.N, D = p.D, M = p.M;
    const int ntiles = PAIR ? (p.R + 1) >> 1 : p.R; // a tile holds one request, or a pair of them
    const int r_begin = (int)(((long long)blockIdx.x * ntiles) / gridDim.x);
    const int r_end = (int)(((long long)(blockIdx.x + 1) * ntiles) / gridDim.x);
    const int nreq = r_end - r_begin;
    const int nslab = D >> 5;                  // 32-column operand slabs per request (D % 32 == 0)
    const int spr = (nslab + KO - 1) / KO;     // stages per request
    const int cst = kTcChunkSlabs / KO > 0 ? kTcChunkSlabs / KO : 1; // stages per accumulation chunk (X3)

    if (threadIdx.x == 0) {
        for (int s = 0; s < S; ++s) {
            mbar_init(full0 + 8 * s, 1);
            mbar_init(empty0 + 8 * s, 1);
            mbar_init(ready0 + 8 * s, NTW);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(tfull0 + 8 * b, 1);
            mbar_init(tempty0 + 8 * b, 128);
            mbar_init(yempty0 + 8 * b, 128); // only [0] is used
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    constexpr uint32_t kTmemCols = X3 ? 512 : 256; // two accumulators: 128 x 128 fp32, or X3: 128 x 256 = [X | Y]
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (X3) {
        // X3 MMAs always accumulate (the N = 256 tile holds X, restarted per chunk, beside Y, which is not),
        // so the epilogue threads zero what must restart; at the start, everything.
        if (warp >= 4 && warp < 8) {
            uint32_t z[32];
#pragma unroll
            for (int e = 0; e < 32; ++e) z[e] = 0u;
            for (int c = 0; c < 16; ++c) tmem_st32(tmem_base + ((uint32_t)((warp - 4) * 32) << 16) + (uint32_t)(c * 32), z);
            asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        }
        tc_fence_before();
        __syncthreads();
        tc_fence_after();
    }

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            const uint32_t ring_u32 = smem_u32(ring);
            int slot = 0;
            uint32_t phase = 0;
            for (int q = 0; q < nreq; ++q) {
                if (fix_flags && !__ldg(fix_flags + r_begin + q)) continue;
                for (int s = 0; s < spr; ++s) {
                    mbar_wait_backoff(empty0 + 8 * slot, phase ^ 1u, 40);
                    if (X3) { // one box per slab, into the first half of its [raw][lo] pair
                        const int nk = min(KO, nslab - s * KO);
                        mbar_arrive_expect_tx(full0 + 8 * slot, (uint32_t)(nk * kTcSlabBytes));
                        for (int k = 0; k < nk; ++k)
                            tma_load_4d(ring_u32 + slot * stageb + k * kSlabStride, &tmE, 0, 0, s * KO + k, r_begin + q,
                                        full0 + 8 * slot);
                    } else {
                        mbar_arrive_expect_tx(full0 + 8 * slot, (uint32_t)(KO * kTcSlabBytes)); // OOB zero-fill counts too
                        if (PAIR)
                            tma_load_4d(ring_u32 + slot * stageb, &tmE, 0, 0, 2 * (r_begin + q), s * KO, full0 + 8 * slot);
                        else
                            tma_load_4d(ring_u32 + slot * stageb, &tmE, 0, 0, s * KO, r_begin + q, full0 + 8 * slot);
                    }
                    if (++slot == S) {
                        slot = 0;
                        phase ^= 1u;
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            const uint32_t ring_u32 = smem_u32(ring);
            const uint32_t idesc = umma_idesc_tf32(128, X3 ? 256 : 128);
            constexpr uint32_t kAccCols = X3 ? 256u : 128u;
            int slot = 0;
            uint32_t phase = 0;
            uint32_t tile = 0; // accumulator hand-offs so far: one per request (TF32) or per chunk (X3)
            uint32_t nproc = 0; // requests scored so far by this CTA
            for (int q = 0; q < nreq; ++q) {
                if (fix_flags && !__ldg(fix_flags + r_begin + q)) continue;
                uint32_t ab = tile & 1u;
                mbar_wait(tempty0 + 8 * ab, ((tile >> 1) & 1u) ^ 1u); // accumulator drained by the epilogue
                tc_fence_after();
                uint32_t d_tmem = tmem_base + ab * kAccCols;
                uint32_t acc = 0;
                if (X3) { // both Y halves drained and zeroed by the epilogue of the previous request
                    mbar_wait(yempty0, (nproc & 1u) ^ 1u);
                    tc_fence_after();
                }
                ++nproc;
                for (int s = 0; s < spr; ++s) {
                    if (X3 && s > 0 && s % cst == 0) { // chunk boundary: hand over, switch accumulator
                        umma_commit(tfull0 + 8 * ab);
                        ++tile;
                        ab = tile & 1u;
                        mbar_wait(tempty0 + 8 * ab, ((tile >> 1) & 1u) ^ 1u);
                        tc_fence_after();
                        d_tmem = tmem_base + ab * kAccCols;
                    }
                    mbar_wait_backoff((X3 ? ready0 : full0) + 8 * slot, phase, 20); // X3: ready implies full
                    tc_fence_after();
                    const int nk = min(KO, nslab - s * KO);
                    for (int k = 0; k < nk; ++k) {
                        const uint32_t slab = ring_u32 + slot * stageb + k * kSlabStride;
#pragma unroll
                        for (int kk = 0; kk < 4; ++kk) { // 4 x K=8 inside the 128-byte swizzle atom
                            // A: the raw slab (hi to the tensor core).  B: the same start address -- 128 rows
                            // (TF32 route: S += hi hi^T) or 256 rows = [raw ; lo] (X3: [X | Y] += hi [hi ; lo]^T)
                            const uint64_t d = umma_desc_sw128(slab + kk * 32);
                            umma_tf32(d_tmem, d, d, idesc, X3 ? 1u : acc);
                            acc = 1;
                        }
                    }
                    umma_commit(empty0 + 8 * slot); // stage reusable once these MMAs have read it
                    if (++slot == S) {
                        slot = 0;
                        phase ^= 1u;
                    }
                }
                umma_commit(tfull0 + 8 * ab); // accumulator (last chunk) complete
                ++tile;
            }
        }
        __syncwarp();
    } else if (X3 && (warp == 2 || warp == 3 || warp >= 8)) {
        // ===================== transform: lo = x - trunc_tf32(x), slab by slab =====================
        // Same swizzled layout in the twin slab, so the transform is purely elementwise (16 B per lane,
        // lanes contiguous: conflict-free).
        const int tw = warp < 4 ? warp - 2 : warp - 6; // 0..3
        int slot = 0;
        uint32_t phase = 0;
        for (int q = 0; q < nreq; ++q) {
            if (fix_flags && !__ldg(fix_flags + r_begin + q)) continue;
            for (int s = 0; s < spr; ++s) {
                mbar_wait_backoff(full0 + 8 * slot, phase, 40);
                const int nk = min(KO, nslab - s * KO);
                for (int k = 0; k < nk; ++k) {
                    const uint32_t src = smem_u32(ring) + (uint32_t)slot * (uint32_t)stageb + (uint32_t)k * kSlabStride;
#pragma unroll 4
                    for (int idx = tw * 32 + lane; idx < kTcSlabBytes / 16; idx += NTW * 32) {
                        float4 x, l;
                        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                                     : "=f"(x.x), "=f"(x.y), "=f"(x.z), "=f"(x.w)
                                     : "r"(src + (uint32_t)idx * 16u));
                        l.x = x.x - __uint_as_float((__float_as_uint(x.x) + kTcRound) & 0xffffe000u);
                        l.y = x.y - __uint_as_float((__float_as_uint(x.y) + kTcRound) & 0xffffe000u);
                        l.z = x.z - __uint_as_float((__float_as_uint(x.z) + kTcRound) & 0xffffe000u);
                        l.w = x.w - __uint_as_float((__float_as_uint(x.w) + kTcRound) & 0xffffe000u);
                        asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(src + (uint32_t)kTcSlabBytes + (uint32_t)idx * 16u),
                                     "f"(l.x), "f"(l.y), "f"(l.z), "f"(l.w)
                                     : "memory");
                    }
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
        }
    } else if (warp >= 4 && warp < 8) {
        // ===================== epilogue =====================
        const int i = threadIdx.x - 128;   // row of S owned by this thread == TMEM lane
        const int ew = warp - 4;           // TMEM lane quarter this warp may access
        constexpr int NCH = PAIR ? 2 : 4;  // 32-column chunks of a request's accumulator columns
        const int hq = PAIR ? (ew >> 1) : 0;   // which request of the tile this row belongs to
        const int ii = PAIR ? (i & 63) : i;    // candidate index
        const int cbase = hq * 64;             // first accumulator column / row of that request
        const int w0 = PAIR ? 2 * hq : 0;      // first of the epilogue warps that own the request
        const uint32_t lane_addr = tmem_base + ((uint32_t)(ew * 32) << 16);
        uint32_t tile = 0;
        const int nchunk = X3 ? (spr + cst - 1) / cst : 1;
        for (int q = 0; q < nreq; ++q) {
            const int r = PAIR ? 2 * (r_begin + q) + hq : r_begin + q;
            if (!PAIR && fix_flags && !__ldg(fix_flags + r)) continue;
            const bool exists = r < p.R; // an odd batch ends on a half-empty tile
            const bool act = exists && ii < N;
            const float *Vr = p.V + (size_t)r * M * N;
            const float *wr = p.w + (p.w_per_request ? (size_t)r * M : 0);
            // t_i = sum_m w_m V_mi (client.rs:410-415): independent of E, overlaps the MMAs.  Eight voters' loads are
            // issued together: a load-test-load chain per voter (~1.5 k clk each) would keep these threads away from
            // the accumulator hand-offs of the X3 route for most of two chunks.
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
            float d = 0.f, acc = 0.f, rn = 0.f;
            if (X3) {
                // add the chunks' accumulators into registers with ordinary (round-to-nearest) fp32 adds
                float row[128];
#pragma unroll
                for (int e = 0; e < 128; ++e) row[e] = 0.f;
                uint32_t z[32];
#pragma unroll
                for (int e = 0; e < 32; ++e) z[e] = 0u;
                // the accumulator this request's even chunks use: with an odd chunk count it alternates from request
                // to request, and the two Y halves are added below in chunk-parity order (even chunks' first) so
                // that a request's bits do not depend on its position in the batch
                const uint32_t first_ab = tile & 1u;
                for (int ch = 0; ch < nchunk; ++ch, ++tile) {
                    const uint32_t ab = tile & 1u;
                    mbar_wait_backoff(tfull0 + 8 * ab, (tile >> 1) & 1u, 200);
                    tc_fence_after();
                    const uint32_t row_addr = lane_addr + ab * 256u; // X half of the [X | Y] accumulator
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        tmem_ld32(row_addr + (uint32_t)(c * 32), v);
                        tmem_st32(row_addr + (uint32_t)(c * 32), z); // X restarts with the accumulator's next chunk
#pragma unroll
                        for (int e = 0; e < 32; ++e) row[c * 32 + e] += v[e];
                    }
                    if (ch + 1 < nchunk) {
                        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                        tc_fence_before();
                        mbar_arrive(tempty0 + 8 * ab);
                    }
                }
                // The last chunk's commit covered every MMA of the request, so Y = hi.lo^T is complete too: it sits
                // in the Y halves of BOTH accumulators (even chunks in one, odd chunks in the other).  The next
                // request's MMAs wait for Y, so it is drained (and zeroed) before any block-wide step: the diagonal
                // gives n_i, hence u_i = t_i / n_i, which is all a thread needs to scale ITS row of Y for the
                // column sums; X + Y stays in row[].
                const uint32_t last_ab = (tile - 1u) & 1u;
                float yd = 0.f;
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                    const uint32_t h = (uint32_t)hh ^ first_ab;
                    tmem_ld32(lane_addr + h * 256u + 128u + (uint32_t)(ew * 32), v);
#pragma unroll
                    for (int e = 0; e < 32; ++e) yd += (lane == e) ? v[e] : 0.f;
                }
#pragma unroll
                for (int c = 0; c < 4; ++c)
#pragma unroll
                    for (int e = 0; e < 32; ++e) d = (c == ew && lane == e) ? row[c * 32 + e] : d;
                d = fmaf(2.f, yd, d); // S_ii = X_ii + 2 Y_ii
                rn = (act && d > 0.f) ? 1.0f / sqrtf(d) : 0.f;
                const float ui = t * rn;
                u_s[i] = ui;
                // sum_j u_j S_ij = sum_j u_j (X_ij + Y_ij)  [own row]  +  sum_j u_j Y_ji  [column i of u_j-scaled Y]
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    float cs = 0.f;
#pragma unroll
                    for (int hh = 0; hh < 2; ++hh) {
                        const uint32_t ya = lane_addr + ((uint32_t)hh ^ first_ab) * 256u + 128u + (uint32_t)(c * 32);
                        tmem_ld32(ya, v);
                        tmem_st32(ya, z);
#pragma unroll
                        for (int e = 0; e < 32; ++e) {
                            row[c * 32 + e] += v[e];
                            v[e] *= ui;
                        }
                        cs += warp_transpose_sum<32>(v, lane); // rows of this warp, column 32c + lane
                    }
                    col_s[ew * 128 + c * 32 + lane] = cs;
                }
                asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                tc_fence_before();
                mbar_arrive(tempty0 + 8 * last_ab); // the last chunk's X half, zeroed in the loop above
                mbar_arrive(yempty0);
                epi_sync();
#pragma unroll
                for (int e = 0; e < 128; ++e) acc = fmaf(u_s[e], row[e], acc);
                acc += (col_s[i] + col_s[128 + i]) + (col_s[256 + i] + col_s[384 + i]);
            } else {
                const uint32_t ab = tile & 1u;
                mbar_wait_backoff(tfull0 + 8 * ab, (tile >> 1) & 1u, 200);
                tc_fence_after();
                const uint32_t row_addr = lane_addr + ab * 128u;
                // diagonal: S_ii sits in column chunk `ew` at register index `lane`
                tmem_ld32(row_addr + (uint32_t)(ew * 32), v);
#pragma unroll
                for (int e = 0; e < 32; ++e) d = (lane == e) ? v[e] : d;
                rn = (act && d > 0.f) ? 1.0f / sqrtf(d) : 0.f;
                u_s[i] = t * rn;
                epi_sync();
#pragma unroll
                for (int c = 0; c < NCH; ++c) { // the request's own columns
                    tmem_ld32(row_addr + (uint32_t)(cbase + c * 32), v);
#pragma unroll
                    for (int e = 0; e < 32; ++e) acc = fmaf(u_s[cbase + c * 32 + e], v[e], acc);
                }
                // accumulator fully read: hand it back to the MMA warp
                tc_fence_before();
                mbar_arrive(tempty0 + 8 * ab);
                ++tile;
            }
            float c = act ? acc * rn : 0.f;
            // ---- tally (client.rs:422-449) with fixed-order block reductions
            float ws = warp_sum(c);
            if (lane == 0) part[ew] = ws;
            epi_sync();
            const float sum = PAIR ? part[w0] + part[w0 + 1] : (part[0] + part[1]) + (part[2] + part[3]);
            const bool positive = sum > 0.f;
            const float cf = (positive && act) ? c / sum : 0.f;
            cf_s[i] = cf;
            if (act) {
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
            if (p.winner && ii == 0 && exists) {
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
            if (p.rank && act) {
                int pos = 0;
                for (int j = 0; j < N; ++j) {
                    const float cj = cf_s[cbase + j];
                    pos += (cj > cf || (cj == cf && j < ii)) ? 1 : 0;
                }
                p.rank[(size_t)r * N + pos] = ii;
            }
            if (p.agree && exists) {
                constexpr int NW = PAIR ? 2 : 4; // warps sharing this request's voters
                for (int m0 = ew - w0; m0 < M; m0 += 4 * NW) { // one warp per voter, four voters' loads in flight
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
            epi_sync(); // u_s / cf_s / part are reused by the next request
        }
    }

    // ---- teardown: everyone done with TMEM before it is freed
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
    }
}

inline size_t tc_smem_bytes(int ko, int stages, bool x3) {
    return 1024 + (size_t)stages * ko * kTcSlabBytes * (x3 ? 2 : 1) + (128 + 128 + 16 + 8 + 2) * 4 + (3 * kTcMaxStages + 6) * 8 +
           (x3 ? 4 * 128 * 4 : 0) + 64;
}

} // namespace lwc
