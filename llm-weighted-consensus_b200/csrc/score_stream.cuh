// score_stream.cuh -- persistent, warp-specialised small-pool kernel (N <= 32), sm_100a.
//
// One CTA per SM, resident for the whole launch; each CTA owns a contiguous range of requests.
//
//   warp 0              producer.  A stage is NT candidate rows x KS columns of one request.  Each
//                       lane < N issues ONE bulk async copy (cp.async.bulk, the 1-D TMA path) for
//                       its row -- KS*4 = 3 KB per copy -- and the stage's mbarrier collects the
//                       bytes.  tools/tma_probe.cu (profiles/r1_tma_probe.txt) measured what the
//                       copy engine can do on this tensor: ~50 ns per bulk copy and ~18 ns per
//                       tensor-box row *whatever their size*, and ~2.5 us of loaded HBM latency.
//                       So copies must be >= 3 KB to stream at HBM speed (512 B copies ran at
//                       1.5 TB/s) and the ring must keep >= 150 KB in flight.  Rows are placed at
//                       pitch KS*4+64 bytes = 64 (mod 128), which makes the fragment-shaped
//                       128-bit shared loads below bank-conflict free.
//   warps 1..NCW        consumers.  Every consumer waits on every stage (plain in-order ring) and
//                       takes a fixed contiguous run of its 16-column chunks.  A lane (g,t) loads
//                       row g, bytes 16t..16t+15 of a 64-byte chunk: those registers are at once
//                       the A and B operands of mma.m16n8k8.tf32 (Gram matrix, B = A^T), see
//                       score_small.cuh.  At the end of a request each consumer parks its partial
//                       fragments in shared memory and streams on.
//   warps NCW+1..+NF    finalisers.  One park buffer; finaliser f takes requests q = f (mod NF).
//                       A token (mbarrier) travels round the finalisers so that each one waits on
//                       the shared park barrier only when it is its turn -- a warp that skipped
//                       phases could not trust a parity wait.  The owner adds the parked partials
//                       in consumer order (fixed chunk ownership => bit-reproducible), frees the
//                       buffer, passes the token on, then forms S = HH + (HL + HL^T) and runs the
//                       tally (client.rs:410-449) -- a ~3.5 us dependent chain per request, which
//                       is why several finalisers overlap.  Votes/weights are prefetched ahead.
//
// Bound: HBM.  Every byte of E crosses HBM->SMEM once; nothing is written but the results.
#pragma once

#include <cuda.h>

#include "score_small.cuh"

namespace lwc {

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}
// Bounded wait: a pipeline bug must surface as a launch failure, never as a hung GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    if (mbar_try_wait(bar, parity)) return;
    const long long t0 = clock64();
    while (!mbar_try_wait(bar, parity)) {
        if (clock64() - t0 > 4000000000LL) __trap(); // ~2 s at 2 GHz
    }
}
// The same wait for warps that expect to wait long: every try_wait of a spinning warp is a shared-memory
// wavefront (ncu: ten spinning warps took 16 % of the SM's shared-memory pipe in score_tc_kernel<x3>), so
// back off between polls.
__device__ __forceinline__ void mbar_wait_backoff(uint32_t bar, uint32_t parity, unsigned ns) {
    if (mbar_try_wait(bar, parity)) return;
    const long long t0 = clock64();
    do {
        __nanosleep(ns);
        if (clock64() - t0 > 4000000000LL) __trap();
    } while (!mbar_try_wait(bar, parity));
}
// TMA: one [1][rows][cols] box of a rank-3 tensor map -> shared memory, completion on `bar`.
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap *tm, int c0, int c1, int c2, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
        ::"r"(dst), "l"(tm), "r"(c0), "r"(c1), "r"(c2), "r"(bar)
        : "memory");
}

// global -> shared bulk async copy; bytes % 16 == 0, both addresses 16-byte aligned.
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}

// Launch geometry chosen on the host (lwc_api.cu) and passed by value.
struct StreamGeom {
    int ks;     // columns per stage (multiple of 16)
    int stages; // ring depth
    int parks;  // park buffers (1 or 2): 2 lets the consumers park request q while a finaliser still drains q - 1
};

template <int NT, bool X3, int NCW, int NFIN>
struct StreamCfg {
    static constexpr int RPL = NT / 8, MT = NT / 16, NTL = NT / 8;
    static constexpr int TPITCH = NT + 1; // pitch of the S tile (floats)
    static constexpr int NACC = (MT * NTL + (X3 ? MT * NTL : 0)) * 4;
    static constexpr int NF = NFIN; // finaliser warps (they share one park buffer)
    static constexpr int MAXS = 16;               // ring depth (upper bound)
    static constexpr int THREADS = (1 + NCW + NF) * 32;
    static constexpr size_t RED_BYTES = (size_t)NCW * NACC * 32 * sizeof(float);
    static constexpr size_t TILE_BYTES = (size_t)NF * NT * TPITCH * sizeof(float);
    static constexpr size_t BAR_BYTES = (size_t)(2 * MAXS + 4 + NF) * sizeof(uint64_t);
    static constexpr size_t fixed_bytes(int parks = 1) { return (size_t)parks * RED_BYTES + TILE_BYTES + BAR_BYTES + 128; }
    // rows: candidate rows a stage has room for (NT, or 8 in the half-tile form: twice the stages, twice the bytes in flight)
    static constexpr size_t stage_bytes(int ks, int rows = NT) { return (size_t)rows * (ks * sizeof(float) + 64); }
    static constexpr size_t smem_bytes(int ks, int stages, int rows = NT, int parks = 1) { return fixed_bytes(parks) + (size_t)stages * stage_bytes(ks, rows); }
};

constexpr bool kStreamFlush = kMmaChain < (1 << 20); // A/B builds (-DLWC_MMA_CHAIN=1000000) compile the flush out
constexpr int kVotePrefetch = 4; // voters whose rows a finaliser keeps in registers

// A list of same-shaped batches scored by ONE launch (lwc_score_batches): request g of the launch is request
// g % R_each of batch g / R_each.  Passed by value as a kernel parameter (constant bank), so no device allocation or
// copy is needed for it.
constexpr int kMaxList = 32;
struct ScoreList {
    int n, R_each;
    const float *E[kMaxList], *V[kMaxList], *w[kMaxList];
    float *conf[kMaxList], *cw[kMaxList], *agree[kMaxList];
    int32_t *winner[kMaxList], *rank[kMaxList];
};

// `p` with the pointers of the batch that holds launch-wide request g; *r = g's index inside that batch
__device__ __forceinline__ ScoreParams list_params(const ScoreParams &p, const ScoreList *list, int g, int *r) {
    ScoreParams q = p;
    const int b = g / list->R_each;
    *r = g - b * list->R_each;
    q.E = list->E[b], q.V = list->V[b], q.w = list->w[b];
    q.conf = list->conf[b], q.cw = list->cw[b], q.agree = list->agree[b];
    q.winner = list->winner[b], q.rank = list->rank[b];
    return q;
}

// NB: 8-column MMA tiles (= 8-row groups) that can hold candidates.  NB = NT / 8 is the full tile; NB = 1 with NT = 16 is
// the form for pools of at most 8 candidates: rows 8..15 are neither loaded nor multiplied (their A registers are
// constant zeros), the second column tile is never computed -- half the shared-memory loads, piece splits and MMAs of
// a request, which is what bounds this kernel when a request brings half the bytes.
// GROUP (half-tile form only, N <= 4): G = 8 / N consecutive requests share one 8-row tile.  Their rows are contiguous in
// E, so producer and consumers treat the group as one request of G * N rows; only the finaliser knows better
// (warp_tally_group: every sum over the lane's own N x N block of the 8 x 8 Gram matrix).  A 4-candidate pool then costs
// the consumers half, a 2-candidate pool a quarter, of what it costs alone in a tile.
template <int NT, bool X3, int NCW, int NFIN, bool LIST, int NB = NT / 8, bool GROUP = false>
__device__ __forceinline__ void score_stream_body(const ScoreParams &p, const StreamGeom &geo, const ScoreList *list) {
    using C = StreamCfg<NT, X3, NCW, NFIN>;
    static_assert(!GROUP || NB == 1, "request groups: the half-tile form only");
    constexpr int RPL = C::RPL, MT = C::MT, NF = C::NF;
    constexpr int NTL = NB; // column tiles in use (C::NTL of them have room in the park buffer and the S tile)
    static_assert(NB >= 1 && NB <= C::NTL && (NB == C::NTL || NB == C::NTL - 1), "NB: the full tile or one 8-row group less");
    constexpr int TPITCH = C::TPITCH, NACC = C::NACC, MAXS = C::MAXS;

    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int KS = geo.ks, S = geo.stages;
    const int pitchb = KS * 4 + 64;  // bytes between candidate rows of a stage: 64 (mod 128)
    constexpr int SROWS = NB * 8;        // candidate rows a stage holds
    const int stageb = SROWS * pitchb;   // bytes per stage
    unsigned char *ring = smem_raw;                                    // [S][SROWS][pitchb]
    const int P = geo.parks;                                           // park buffers: request q uses buffer q % P
    float *red = reinterpret_cast<float *>(ring + (size_t)S * stageb); // [P][NCW][NACC][32]
    float *tiles = red + (size_t)P * NCW * NACC * 32;                  // [NF][NT][TPITCH]
    uint64_t *bars = reinterpret_cast<uint64_t *>(tiles + (size_t)NF * NT * TPITCH);
    const uint32_t full0 = smem_u32(bars), empty0 = full0 + 8 * MAXS;
    const uint32_t redfull0 = empty0 + 8 * MAXS, redempty0 = redfull0 + 16, turn0 = redempty0 + 16; // [2] each

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int D = p.D, M = p.M;
    const int G = GROUP ? 8 / p.N : 1;    // requests per tile
    const int N = GROUP ? G * p.N : p.N;  // rows per tile: what the producer copies and the consumers multiply
    const int ntiles = GROUP ? (p.R + G - 1) / G : p.R;
    const int r_begin = (int)(((long long)blockIdx.x * ntiles) / gridDim.x); // in tiles
    const int r_end = (int)(((long long)(blockIdx.x + 1) * ntiles) / gridDim.x);
    const int nreq = r_end - r_begin;
    const int spr = (D + KS - 1) / KS; // stages per tile

    // rows >= N are never written by a copy and must read as zeros
    if (N < SROWS) {
        float4 *z = reinterpret_cast<float4 *>(ring);
        const int n16 = S * stageb / 16;
        for (int i = threadIdx.x; i < n16; i += C::THREADS) z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (threadIdx.x == 0) {
        for (int s = 0; s < S; ++s) {
            mbar_init(full0 + 8 * s, 1);
            mbar_init(empty0 + 8 * s, NCW);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(redfull0 + 8 * b, NCW);
            mbar_init(redempty0 + 8 * b, 1);
        }
        for (int b = 0; b < NF; ++b) mbar_init(turn0 + 8 * b, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    // Back-to-back launches: the next grid may take this SM as soon as this CTA retires.  Unless the
    // caller vouched for resident inputs, nothing is read before the previous grid has finished.
    pdl_launch_dependents();
    if (!p.inputs_resident) pdl_wait_prior_grid();

    if (warp == 0) {
        // ===================== producer =====================
        const uint32_t ring_u32 = smem_u32(ring);
        int slot = 0;
        uint32_t phase = 0;
        for (int q = 0; q < nreq; ++q) {
            const float *Er;
            const int g0 = (r_begin + q) * G; // first request of the tile
            if (LIST) {
                int rl;
                const ScoreParams pb = list_params(p, list, g0, &rl); // (a group never straddles two batches: lwc_api.cu)
                Er = pb.E + (size_t)rl * p.N * D;
            } else {
                Er = p.E + (size_t)g0 * p.N * D;
            }
            const int rows = GROUP ? min(G, p.R - g0) * p.N : N; // the last group of a batch may be short
            for (int s = 0; s < spr; ++s) {
                mbar_wait(empty0 + 8 * slot, phase ^ 1u);
                const int kv = min(KS, D - s * KS);
                if (lane == 0) mbar_arrive_expect_tx(full0 + 8 * slot, (uint32_t)(kv * 4 * rows));
                __syncwarp();
                if (lane < rows)
                    bulk_g2s(ring_u32 + slot * stageb + lane * pitchb, Er + (size_t)lane * D + (size_t)s * KS,
                             (uint32_t)(kv * 4), full0 + 8 * slot);
                if (++slot == S) {
                    slot = 0;
                    phase ^= 1u;
                }
            }
        }
        return;
    }

    const int g8 = lane >> 2, t4 = lane & 3;

    if (warp <= NCW) {
        // ===================== consumers =====================
        const int cw = warp - 1;
        float hh[MT][NTL][4], hl[MT][NTL][4];
#pragma unroll
        for (int a = 0; a < MT; ++a)
#pragma unroll
            for (int b = 0; b < NTL; ++b)
#pragma unroll
                for (int e = 0; e < 4; ++e) hh[a][b][e] = hl[a][b][e] = 0.f;

        const int cpw = ((KS >> 4) + NCW - 1) / NCW; // chunks of a full stage owned by one consumer
        const int c_lo = cw * cpw;
        // X3: running sums of HH moved out of the tensor-core accumulators every few stages (kMmaChain, score_small.cuh)
        float fh[MT][NTL][4];
#pragma unroll
        for (int a = 0; a < MT; ++a)
#pragma unroll
            for (int b = 0; b < NTL; ++b)
#pragma unroll
                for (int e = 0; e < 4; ++e) fh[a][b][e] = 0.f;
        const int flush_stages = max(1, kMmaChain / (2 * cpw)); // 2 MMAs per accumulator and chunk
        int slot = 0;
        uint32_t phase = 0;
        for (int q = 0; q < nreq; ++q) {
            int since = 0;
            for (int s = 0; s < spr; ++s) {
                mbar_wait(full0 + 8 * slot, phase);
                const int kv = min(KS, D - s * KS);        // valid columns of this stage
                const int c_hi = min(c_lo + cpw, (kv + 15) >> 4);
                const unsigned char *base = ring + (size_t)slot * stageb + g8 * pitchb + t4 * 16;
#pragma unroll 2
                for (int c = c_lo; c < c_hi; ++c) {
                    float4 v[RPL];
#pragma unroll
                    for (int qq = 0; qq < RPL; ++qq)
                        v[qq] = qq < NB ? *reinterpret_cast<const float4 *>(base + qq * 8 * pitchb + c * 64) : make_float4(0.f, 0.f, 0.f, 0.f);
                    if (c * 16 + 16 > kv && c * 16 + 4 * t4 >= kv) { // ragged tail: stale bytes beyond D
#pragma unroll
                        for (int qq = 0; qq < RPL; ++qq) v[qq] = make_float4(0.f, 0.f, 0.f, 0.f);
                    }
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        uint32_t xh[RPL], yh[RPL], xl[RPL], yl[RPL];
#pragma unroll
                        for (int qq = 0; qq < RPL; ++qq) {
                            if (qq >= NB) { // no candidates in these rows: constant zero operands
                                xh[qq] = yh[qq] = xl[qq] = yl[qq] = 0u;
                                continue;
                            }
                            const float x = h ? v[qq].z : v[qq].x;
                            const float y = h ? v[qq].w : v[qq].y;
                            if (X3) {
                                split_tf32(x, xh[qq], xl[qq]);
                                split_tf32(y, yh[qq], yl[qq]);
                            } else {
                                xh[qq] = __float_as_uint(x) + 0x1000u;
                                yh[qq] = __float_as_uint(y) + 0x1000u;
                            }
                        }
#pragma unroll
                        for (int a = 0; a < MT; ++a)
#pragma unroll
                            for (int bb = 0; bb < NTL; ++bb) {
                                if (bb >= 2 * a)
                                    mma_m16n8k8_tf32(hh[a][bb], xh[2 * a], xh[2 * a + 1], yh[2 * a], yh[2 * a + 1],
                                                     xh[bb], yh[bb]);
                                if (X3)
                                    mma_m16n8k8_tf32(hl[a][bb], xh[2 * a], xh[2 * a + 1], yh[2 * a], yh[2 * a + 1],
                                                     xl[bb], yl[bb]);
                            }
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(empty0 + 8 * slot);
                if (++slot == S) {
                    slot = 0;
                    phase ^= 1u;
                }
                if (X3 && kStreamFlush && ++since == flush_stages && s + 1 < spr) {
                    since = 0;
#pragma unroll
                    for (int a = 0; a < MT; ++a)
#pragma unroll
                        for (int b = 0; b < NTL; ++b)
                            if (b >= 2 * a) {
#pragma unroll
                                for (int e = 0; e < 4; ++e) {
                                    fh[a][b][e] += hh[a][b][e];
                                    hh[a][b][e] = 0.f;
                                }
                            }
                }
            }
            if (X3 && kStreamFlush) {
#pragma unroll
                for (int a = 0; a < MT; ++a)
#pragma unroll
                    for (int b = 0; b < NTL; ++b)
                        if (b >= 2 * a) {
#pragma unroll
                            for (int e = 0; e < 4; ++e) {
                                hh[a][b][e] += fh[a][b][e];
                                fh[a][b][e] = 0.f;
                            }
                        }
            }
            // ---- request end: park this warp's partial fragments (buffer free once q-1 was drained)
            const int pb = P == 2 ? (q & 1) : 0;       // park buffer of this request
            const uint32_t pk = (uint32_t)(P == 2 ? q >> 1 : q); // its use count
            mbar_wait(redempty0 + 8 * pb, (pk & 1u) ^ 1u);
            {
                float *dst = red + ((size_t)pb * NCW + cw) * NACC * 32 + lane;
                int k = 0;
#pragma unroll
                for (int a = 0; a < MT; ++a)
#pragma unroll
                    for (int b = 0; b < NTL; ++b)
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            dst[(k++) * 32] = hh[a][b][e];
                            if (X3) dst[(k++) * 32] = hl[a][b][e];
                            hh[a][b][e] = 0.f;
                            hl[a][b][e] = 0.f;
                        }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(redfull0 + 8 * pb);
        }
        return;
    }

    // ===================== finalisers =====================
    const int f = warp - 1 - NCW;
    float *tile = tiles + (size_t)f * NT * TPITCH;
    float pv[kVotePrefetch], pw[kVotePrefetch];
    auto prefetch_votes = [&](int r) { // r: tile index (= request index unless GROUP)
        ScoreParams pb = p;
        const int pn = p.N;
        const int gl = GROUP ? lane / pn : 0, cl = GROUP ? lane - gl * pn : lane; // this lane's request of the group, choice
        int rq = r * G;
        const bool ok = GROUP ? (lane < G * pn && rq + gl < p.R) : lane < pn;
        if (LIST) {
            const int g = rq;
            pb = list_params(p, list, g, &rq);
        }
        rq += ok ? gl : 0;
        const float *Vr = pb.V + (size_t)rq * M * pn;
        const float *wr = pb.w + (pb.w_per_request ? (size_t)rq * M : 0);
#pragma unroll
        for (int m = 0; m < kVotePrefetch; ++m) {
            pv[m] = (m < M && ok) ? __ldg(Vr + (size_t)m * pn + cl) : 0.f;
            pw[m] = (m < M && (ok || !GROUP)) ? __ldg(wr + m) : 0.f;
        }
    };
    if (f < nreq) prefetch_votes(r_begin + f);
    // results may only be stored once the previous grid (which may write the same buffers) is complete
    if (p.inputs_resident) pdl_wait_prior_grid();
    for (int q = f, k = 0; q < nreq; q += NF, ++k) {
        // wait for the token: the owner of q-1 has drained the park buffer, so the park barrier is
        // exactly in phase q (this warp did not watch the phases in between)
        if (q > 0) mbar_wait(turn0 + 8 * f, (uint32_t)(f == 0 ? k - 1 : k) & 1u);
        // (two park buffers: buffer q % 2 was last drained by the owner of q - 2, before the token got here, and cannot
        // fill for q + 2 until this warp has drained it -- its barrier is exactly in use q / 2)
        const int pb = P == 2 ? (q & 1) : 0;
        mbar_wait(redfull0 + 8 * pb, (uint32_t)(P == 2 ? q >> 1 : q) & 1u);
        const int r = r_begin + q;
        float sh[MT][NTL][4], sl[MT][NTL][4];
#pragma unroll
        for (int a = 0; a < MT; ++a)
#pragma unroll
            for (int b = 0; b < NTL; ++b)
#pragma unroll
                for (int e = 0; e < 4; ++e) sh[a][b][e] = sl[a][b][e] = 0.f;
        for (int ws = 0; ws < NCW; ++ws) { // consumer order: timing-independent sums
            const float *src = red + ((size_t)pb * NCW + ws) * NACC * 32 + lane;
            int k = 0;
#pragma unroll
            for (int a = 0; a < MT; ++a)
#pragma unroll
                for (int b = 0; b < NTL; ++b)
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        sh[a][b][e] += src[(k++) * 32];
                        if (X3) sl[a][b][e] += src[(k++) * 32];
                    }
        }
        __syncwarp();
        if (lane == 0) {
            mbar_arrive(redempty0 + 8 * pb); // partials are in registers: consumers may park again
            if (q + 1 < nreq) mbar_arrive(turn0 + 8 * ((f + 1) % NF)); // pass the token
        }

        if (X3) {
#pragma unroll
            for (int a = 0; a < MT; ++a)
#pragma unroll
                for (int b = 0; b < NTL; ++b)
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const int row = 16 * a + g8 + 8 * (e >> 1), col = 8 * b + 2 * t4 + (e & 1);
                        tile[row * TPITCH + col] = sl[a][b][e];
                    }
            __syncwarp();
#pragma unroll
            for (int a = 0; a < MT; ++a)
#pragma unroll
                for (int b = 0; b < NTL; ++b)
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const int row = 16 * a + g8 + 8 * (e >> 1), col = 8 * b + 2 * t4 + (e & 1);
                        if (b >= 2 * a) sh[a][b][e] += sl[a][b][e] + tile[col * TPITCH + row];
                    }
            __syncwarp();
        }
#pragma unroll
        for (int a = 0; a < MT; ++a)
#pragma unroll
            for (int b = 0; b < NTL; ++b)
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int row = 16 * a + g8 + 8 * (e >> 1), col = 8 * b + 2 * t4 + (e & 1);
                    if (b >= 2 * a) {
                        tile[row * TPITCH + col] = sh[a][b][e];
                        if (b >= 2 * a + 2) tile[col * TPITCH + row] = sh[a][b][e];
                    }
                }
        __syncwarp();
        if (GROUP) {
            const int g0 = r * G, nv = min(G, p.R - g0);
            if (LIST) {
                int rl;
                const ScoreParams pb = list_params(p, list, g0, &rl);
                warp_tally_group<kVotePrefetch>(pb, rl, G, nv, tile, TPITCH, lane, pv, pw);
            } else {
                warp_tally_group<kVotePrefetch>(p, g0, G, nv, tile, TPITCH, lane, pv, pw);
            }
        } else if (LIST) {
            int rl;
            const ScoreParams pb = list_params(p, list, r, &rl);
            warp_tally_and_rank_pf<kVotePrefetch, NB * 8>(pb, rl, tile, TPITCH, lane, pv, pw);
        } else {
            warp_tally_and_rank_pf<kVotePrefetch, NB * 8>(p, r, tile, TPITCH, lane, pv, pw);
        }
        __syncwarp();
        if (q + NF < nreq) prefetch_votes(r + NF);
    }
}

template <int NT, bool X3, int NCW, int NFIN, int NB = NT / 8, bool GROUP = false>
__global__ void __launch_bounds__((1 + NCW + NFIN) * 32, 1)
score_stream_kernel(const ScoreParams p, const StreamGeom geo) {
    score_stream_body<NT, X3, NCW, NFIN, false, NB, GROUP>(p, geo, nullptr);
}

// p.R = list.n * list.R_each; p's own pointers are unused
template <int NT, bool X3, int NCW, int NFIN, int NB = NT / 8, bool GROUP = false>
__global__ void __launch_bounds__((1 + NCW + NFIN) * 32, 1)
score_stream_list_kernel(const ScoreParams p, const StreamGeom geo, const __grid_constant__ ScoreList list) {
    score_stream_body<NT, X3, NCW, NFIN, true, NB, GROUP>(p, geo, &list);
}

} // namespace lwc
