// score_tc16a.cuh -- score_tc16_kernel with the A operands in tensor memory, sm_100a.
//
// score_tc16_kernel (score_tc16.cuh) is bound by the shared-memory pipe: per 32 KB of HBM traffic it moves 144 KB
// through shared memory -- 32 KB TMA in, 32 + 32 KB transform, 16 + 32 KB operand reads (A = the h tile, B = [h ; l]).
// tcgen05.mma can take A from TMEM instead, and S = X + 2^-11 (Y + Y^T) does not care whether Y is h l^T or its
// transpose l h^T.  So here BOTH pieces go to TMEM as A operands and only h stays in shared memory, as the one B tile:
//     X (+)= h h^T   (A = h from TMEM, B = the h tile)        Y (+)= l h^T   (A = l from TMEM, B = the h tile)
// The transform warps hold every h and l value in registers anyway: they store the row of h into shared memory (16 KB
// per unit instead of 32) and both rows into a ring of two [h | l] tile pairs in TMEM (tcgen05.st: no shared-memory
// traffic).  112 KB per unit instead of 144.  TMEM (512 columns): X0 | X1 (double-buffered per chunk, as before) | Y
// (ONE accumulator for the whole request: the next request's first MMA waits until the epilogue has read it) | 2 x
// (32 + 32) columns of A tiles.  To store a row a thread must own that row's TMEM lane, so the transform is one thread
// per row (warps 2, 3, 8, 9 have warp % 4 = 2, 3, 0, 1: the four lane quarters); it overwrites its own row of slab 0 with
// the row of the h tile, slab 1 is left behind.  Everything else -- pieces, range check and fix-up pass, chunked X, PAIR
// mode, epilogue arithmetic -- is score_tc16.cuh's.
// Measured (same box, LWC_TC16_A=smem against the default): cfg4 2.422 -> 2.349 ms (0.830 -> 0.856 of HBM); 1.6 GB
// launches: 128 candidates 0.802 -> 0.852, two 64-candidate requests per tile 0.875 -> 0.927, two 48-candidate 0.727 ->
// 0.762.  With only h moved to TMEM (l still a second B tile in shared memory, 128 KB per unit) the steps were 2.525 ->
// 2.485 ms, 0.778 -> 0.814, 0.838 -> 0.868.  ncu on the 2,048-request slice (profiles/r2_tc16a_raw.csv): 646 -> 602 us,
// LSU shared-memory wavefronts 49.7 % -> 37.2 % of the pipe, arbitration conflicts on them 23 M -> 9.8 M.  Tried and dropped
// (LWC_T16A_SETS=2): a second set of four transform warps taking the odd units -- no gain with two requests per tile, and
// at 448 threads the 128-register cap spills the one-request epilogue (2.75 ms).  Also tried and dropped: TMA boxes, MMA
// width and ring stages sized to the pool (a 96-candidate tile three quarters as tall, 8 stages instead of 6) -- 65..112
// candidates stayed where they were (80: 0.666, 96: 0.706 of HBM): a tile costs ~0.7-0.9 us per 64-column unit almost
// regardless of its height, so the limit there is a per-unit latency chain, not bytes through any pipe.
// Timing probes on cfg4 (results deliberately wrong, same box, 2.48 ms for the real kernel): without the Y product -- half
// the MMAs and half the B reads -- 2.21 ms (0.91 of HBM); with ONE [h | l] tile pair in TMEM instead of two
// (LWC_T16A_TILES=1: the transform may not run ahead of the MMAs at all) 2.49 ms against 2.44.  So what is left is the
// tensor core's B reads contending with the transform for the shared-memory pipe, not the hand-offs: the next step has
// to take bytes off that pipe again (the fp32 rows reaching registers without passing through shared memory).
#pragma once

#include "score_tc16.cuh"

namespace lwc {

constexpr uint32_t kT16aCol = 384;  // first TMEM column of the A tiles
#ifndef LWC_T16A_TILES
#define LWC_T16A_TILES 2
#endif
constexpr uint32_t kT16aTiles = LWC_T16A_TILES;  // A tile pairs in flight: [h | l] of one unit, 32 + 32 columns
#ifndef LWC_T16A_SETS
#define LWC_T16A_SETS 1
#endif
constexpr int kT16aSets = LWC_T16A_SETS;               // transform warp sets (4 warps each) taking the units in turn
constexpr int kT16aThreads = 320 + (kT16aSets - 1) * 128; // set 0: warps 2, 3, 8, 9; set 1: warps 10..13

// D (+)= A B^T with A in tensor memory (row i = lane i, K along the columns, two fp16 per column), B in shared memory
__device__ __forceinline__ void umma_f16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(acc)
        : "memory");
}

template <bool PAIR>
__global__ void __launch_bounds__(kT16aThreads, 1)
score_tc16a_kernel(const __grid_constant__ CUtensorMap tmE, const ScoreParams p, const TcGeom geo, int *__restrict__ fix_flags) {
    extern __shared__ __align__(1024) unsigned char smem_t16a[];
    const int S = geo.stages;
    constexpr int NTW = 4;
    unsigned char *ring = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_t16a) + 1023) & ~(uintptr_t)1023);
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
            mbar_init(yempty0 + 8 * b, 128); // [0]: Y read by the epilogue (previous request); [1] unused
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    constexpr uint32_t kTmemCols = 512; // X0 | X1 | Y (128 columns each) | four A tiles of 32 columns
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // Nothing in TMEM is ever cleared by a store: the first MMA of a chunk starts X afresh (accumulate flag off), the
    // first MMA of a request starts Y afresh.

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
            const uint32_t idesc = umma_idesc_f16(128, 128);
            int slot = 0;
            uint32_t phase = 0;
            uint32_t tile = 0; // X accumulator hand-offs so far: one per chunk
            uint32_t unit_no = 0; // units issued so far: unit u's A tile is ring entry u % kT16aTiles
            for (int q = 0; q < nreq; ++q) {
                uint32_t ab = tile & 1u;
                mbar_wait(tempty0 + 8 * ab, ((tile >> 1) & 1u) ^ 1u);
                mbar_wait(yempty0, ((uint32_t)q & 1u) ^ 1u); // Y read by the epilogue of the previous request
                tc_fence_after();
                uint32_t x_tmem = tmem_base + ab * 128u;
                for (int u = 0; u < upr; ++u, ++unit_no) {
                    const bool chunk_start = u % kT16ChunkUnits == 0;
                    if (u > 0 && chunk_start) { // chunk boundary: hand X over, switch X accumulator
                        umma_commit(tfull0 + 8 * ab);
                        ++tile;
                        ab = tile & 1u;
                        mbar_wait(tempty0 + 8 * ab, ((tile >> 1) & 1u) ^ 1u);
                        tc_fence_after();
                        x_tmem = tmem_base + ab * 128u;
                    }
                    mbar_wait_backoff(ready0 + 8 * slot, phase, kT16BackoffM); // the stage holds [h ; l], TMEM holds h
                    tc_fence_after();
                    const uint32_t unit = ring_u32 + slot * kT16UnitBytes;
                    const uint32_t a_tmem = tmem_base + kT16aCol + (unit_no % kT16aTiles) * 64u;
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk) { // 4 x K = 16: 8 TMEM columns of A, 32 bytes inside the swizzle atom of B
                        const uint64_t bh = umma_desc_sw128(unit + kk * 32); // B is the h tile for both products
                        umma_f16_ts(x_tmem, a_tmem + kk * 8u, bh, idesc, (kk == 0 && chunk_start) ? 0u : 1u);          // X (+)= h h^T
                        umma_f16_ts(tmem_base + 256u, a_tmem + 32u + kk * 8u, bh, idesc, (kk == 0 && u == 0) ? 0u : 1u); // Y (+)= l h^T
                    }
                    umma_commit(empty0 + 8 * slot); // the stage's smem AND its A tile are free once these MMAs are done
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
        // ===================== transform, in place, one thread per row =====================
        // fp32 slabs -> the h tile in shared memory (the B operand) and the rows of h and l into this unit's A tiles in TMEM.
        // Thread t of the warp with warp % 4 == g owns row 32g + t: tcgen05.st lets a warp reach exactly that quarter of
        // the lanes, and a thread its own lane.  The row's 2 x 128 bytes are read completely (16 x 128-bit loads: the
        // 16-byte chunks of 8 consecutive rows are XOR-swizzled onto 8 different bank groups, so a quarter-warp is
        // conflict-free), converted, and the row of slab 0 is overwritten with the row of the h tile.
        const int tset = warp >= 10 ? 1 : 0; // which units this warp takes: unit_no % kT16aSets == tset
        const int row = 32 * (warp & 3) + lane;
        const bool on = (PAIR ? (row & 63) : row) < N; // rows past the pool arrive zero-filled: zero pieces, nothing to do
        const uint32_t sw = (uint32_t)(row & 7);
        const uint32_t lane_taddr = ((uint32_t)(32 * (warp & 3)) << 16) + kT16aCol;
        int slot = 0;
        uint32_t phase = 0;
        uint32_t unit_no = 0;
        // the A tile of unit u is free once the MMAs of unit u - kT16aTiles are done: they commit to that unit's `empty`
        int slot_a = 0;
        uint32_t phase_a = 0;
        for (int q = 0; q < nreq; ++q)
            for (int u = 0; u < upr; ++u, ++unit_no) {
                if (kT16aSets > 1 && (int)(unit_no % kT16aSets) != tset) { // the other set's unit: keep the ring positions in step
                    if (unit_no >= kT16aTiles && ++slot_a == S) slot_a = 0, phase_a ^= 1u;
                    if (++slot == S) slot = 0, phase ^= 1u;
                    continue;
                }
                mbar_wait_backoff(full0 + 8 * slot, phase, kT16BackoffT);
                const uint32_t r0 = smem_u32(ring) + (uint32_t)slot * kT16UnitBytes + (uint32_t)row * 128u, r1 = r0 + (uint32_t)kTcSlabBytes;
                uint32_t hreg[32], lreg[32];
#pragma unroll
                for (int e = 0; e < 32; ++e) hreg[e] = lreg[e] = 0u;
                if (on) {
                    float4 x[16];
#pragma unroll
                    for (int c = 0; c < 8; ++c) {
                        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(x[c].x), "=f"(x[c].y), "=f"(x[c].z), "=f"(x[c].w) : "r"(r0 + (((uint32_t)c ^ sw) << 4)));
                        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(x[8 + c].x), "=f"(x[8 + c].y), "=f"(x[8 + c].z), "=f"(x[8 + c].w) : "r"(r1 + (((uint32_t)c ^ sw) << 4)));
                    }
#pragma unroll
                    for (int j = 0; j < 4; ++j) { // 8 columns of slab 0 -> chunk j, 8 columns of slab 1 -> chunk 4 + j
                        uint4 h0, l0, h1, l1;
                        split_f16x8(x[2 * j], x[2 * j + 1], h0, l0);
                        split_f16x8(x[8 + 2 * j], x[8 + 2 * j + 1], h1, l1);
                        const uint32_t o0 = (((uint32_t)j ^ sw) << 4), o1 = (((uint32_t)(4 + j) ^ sw) << 4);
                        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(r0 + o0), "r"(h0.x), "r"(h0.y), "r"(h0.z), "r"(h0.w) : "memory");
                        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(r0 + o1), "r"(h1.x), "r"(h1.y), "r"(h1.z), "r"(h1.w) : "memory");
                        lreg[4 * j] = l0.x, lreg[4 * j + 1] = l0.y, lreg[4 * j + 2] = l0.z, lreg[4 * j + 3] = l0.w;
                        lreg[16 + 4 * j] = l1.x, lreg[16 + 4 * j + 1] = l1.y, lreg[16 + 4 * j + 2] = l1.z, lreg[16 + 4 * j + 3] = l1.w;
                        hreg[4 * j] = h0.x, hreg[4 * j + 1] = h0.y, hreg[4 * j + 2] = h0.z, hreg[4 * j + 3] = h0.w;
                        hreg[16 + 4 * j] = h1.x, hreg[16 + 4 * j + 1] = h1.y, hreg[16 + 4 * j + 2] = h1.z, hreg[16 + 4 * j + 3] = h1.w;
                    }
                }
                if (unit_no >= kT16aTiles) { // warp-uniform
                    mbar_wait_backoff(empty0 + 8 * slot_a, phase_a, kT16BackoffT);
                    tc_fence_after();
                    if (++slot_a == S) {
                        slot_a = 0;
                        phase_a ^= 1u;
                    }
                }
                tmem_st32(tmem_base + lane_taddr + (unit_no % kT16aTiles) * 64u, hreg);       // row of h: 64 fp16 = 32 columns
                tmem_st32(tmem_base + lane_taddr + (unit_no % kT16aTiles) * 64u + 32u, lreg); // row of l beside it
                asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                // generic-proxy writes must be visible to the tensor core (async proxy) before it is told
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                tc_fence_before();
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
            for (int ch = 0; ch < nchunk; ++ch, ++tile) {
                const uint32_t ab = tile & 1u;
                mbar_wait_backoff(tfull0 + 8 * ab, (tile >> 1) & 1u, 200);
                tc_fence_after();
                const uint32_t row_addr = lane_addr + ab * 128u; // this chunk's X accumulator
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
            tmem_ld32(lane_addr + 256u + (uint32_t)(ew * 32), v); // Y is one accumulator for the whole request
#pragma unroll
            for (int e = 0; e < 32; ++e) yd += (lane == e) ? v[e] : 0.f;
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
            float cs[NCH];
#pragma unroll
            for (int c = 0; c < NCH; ++c) {
                tmem_ld32(lane_addr + 256u + (uint32_t)(cbase + c * 32), v);
#pragma unroll
                for (int e = 0; e < 32; ++e) {
                    row[c * 32 + e] = fmaf(v[e], kT16LoScale, row[c * 32 + e]);
                    v[e] *= uis;
                }
                cs[c] = warp_transpose_sum<32>(v, lane);
            }
            tc_fence_before();
            mbar_arrive(yempty0);                // Y read: the next request may overwrite it
            mbar_arrive(tempty0 + 8 * last_ab);  // the last chunk's X was read in the chunk loop
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

} // namespace lwc
