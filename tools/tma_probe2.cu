// tma_probe2.cu -- microbenchmark for the large-pool (N = 128) staging scheme: HBM -> SMEM rate of
// E[R][128][3072] fp32 by TMA flavour, one persistent CTA per SM, no math.
//   mode 0: rank-3 map, box {32, 128, 1}, SWIZZLE_128B (the UMMA K-major operand tile, 16 KB)
//   mode 1: rank-4 map {32, N, D/32, R}, box {32, 128, KO, 1}, SWIZZLE_128B (KO operand tiles per copy)
//   mode 2: rank-3 map, box {256, ROWS, 1}, no swizzle (wide rows)
//   mode 3: 1-D bulk, one copy per candidate row, KC columns each, 32 lanes issuing
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count)); }
__device__ __forceinline__ void mbar_arrive(uint32_t bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory"); }
__device__ __forceinline__ void mbar_expect(uint32_t bar, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory"); }
__device__ __forceinline__ bool mbar_try(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    long long t0 = clock64();
    while (!mbar_try(bar, parity)) if (clock64() - t0 > 2000000000LL) __trap();
}

__global__ void __launch_bounds__(64, 1)
probe(const __grid_constant__ CUtensorMap tm, const float *E, int R, int N, int D, int mode, int p0, int p1, int slots, int slotb) {
    extern __shared__ __align__(1024) unsigned char smem[];
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + (size_t)slots * slotb);
    const uint32_t full0 = smem_u32(bars), empty0 = full0 + 8 * slots;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int r0 = (int)((long long)blockIdx.x * R / gridDim.x), r1 = (int)((long long)(blockIdx.x + 1) * R / gridDim.x);
    if (threadIdx.x == 0) {
        for (int s = 0; s < slots; ++s) { mbar_init(full0 + 8 * s, 1); mbar_init(empty0 + 8 * s, 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const uint32_t ring = smem_u32(smem);
    // stages per request
    int spr;
    if (mode == 0) spr = D / 32;
    else if (mode == 1) spr = D / (32 * p0);
    else if (mode == 2) spr = (D / 256) * (N / p0);
    else spr = D / p0;
    if (warp == 0) {
        uint32_t g = 0;
        for (int r = r0; r < r1; ++r)
            for (int s = 0; s < spr; ++s, ++g) {
                const uint32_t slot = g % slots, round = g / slots;
                if (mode != 3) {
                    if (lane == 0) {
                        mbar_wait(empty0 + 8 * slot, (round & 1) ^ 1);
                        mbar_expect(full0 + 8 * slot, slotb);
                        if (mode == 0)
                            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                                         ::"r"(ring + slot * slotb), "l"(&tm), "r"(s * 32), "r"(0), "r"(r), "r"(full0 + 8 * slot) : "memory");
                        else if (mode == 1)
                            asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];"
                                         ::"r"(ring + slot * slotb), "l"(&tm), "r"(0), "r"(0), "r"(s * p0), "r"(r), "r"(full0 + 8 * slot) : "memory");
                        else {
                            const int rb = s % (N / p0), cb = s / (N / p0);
                            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                                         ::"r"(ring + slot * slotb), "l"(&tm), "r"(cb * 256), "r"(rb * p0), "r"(r), "r"(full0 + 8 * slot) : "memory");
                        }
                    }
                } else {
                    mbar_wait(empty0 + 8 * slot, (round & 1) ^ 1);
                    if (lane == 0) mbar_expect(full0 + 8 * slot, N * p0 * 4);
                    __syncwarp();
                    for (int row = lane; row < N; row += 32)
                        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                     ::"r"(ring + slot * slotb + row * p0 * 4), "l"(E + ((size_t)r * N + row) * D + s * p0),
                                       "r"(p0 * 4), "r"(full0 + 8 * slot) : "memory");
                }
            }
    } else if (lane == 0) {
        uint32_t g = 0;
        unsigned acc = 0;
        for (int r = r0; r < r1; ++r)
            for (int s = 0; s < spr; ++s, ++g) {
                const uint32_t slot = g % slots, round = g / slots;
                mbar_wait(full0 + 8 * slot, round & 1);
                acc += *reinterpret_cast<volatile unsigned int *>(smem + (size_t)slot * slotb);
                mbar_arrive(empty0 + 8 * slot);
            }
        if (acc == 0x12345678u) bars[0] = acc;
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
    const int R = 592, N = 128, D = 3072; // 4 requests per SM, 931 MB
    const size_t bytes = (size_t)R * N * D * 4;
    float *E[2];
    for (int i = 0; i < 2; ++i) { CK(cudaMalloc(&E[i], bytes)); CK(cudaMemset(E[i], i + 1, bytes)); }
    int sms = 0, maxsmem = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    CK(cudaDeviceGetAttribute(&maxsmem, cudaDevAttrMaxSharedMemoryPerBlockOptin, 0));
    CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, maxsmem));
    void *sym = nullptr; cudaDriverEntryPointQueryResult qr;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qr));
    EncodeTiledFn enc = (EncodeTiledFn)sym;
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    printf("%-44s %6s %8s %10s %8s\n", "mode", "slots", "slotKB", "us/pass", "GB/s");
    struct Cfg { int mode, p0, p1; const char *name; };
    Cfg cfgs[] = {
        {0, 0, 0, "3d box{32,128,1} SW128 (16 KB)"},
        {1, 2, 0, "4d box{32,128,2,1} SW128 (32 KB)"},
        {1, 4, 0, "4d box{32,128,4,1} SW128 (64 KB)"},
        {2, 32, 0, "3d box{256,32,1} noswz (32 KB)"},
        {2, 64, 0, "3d box{256,64,1} noswz (64 KB)"},
        {3, 96, 0, "bulk1d 128 rows x 384 B"},
        {3, 192, 0, "bulk1d 128 rows x 768 B"},
        {3, 384, 0, "bulk1d 128 rows x 1536 B"},
    };
    for (auto &c : cfgs) {
        size_t slotb = c.mode == 0 ? 16384 : c.mode == 1 ? (size_t)16384 * c.p0 : c.mode == 2 ? (size_t)256 * 4 * c.p0 : (size_t)N * c.p0 * 4;
        for (int slots : {2, 3, 4, 6, 8, 12}) {
            if (slots * slotb + 1024 > (size_t)maxsmem - 2048) continue;
            CUtensorMap tms[2];
            for (int i = 0; i < 2; ++i) {
                CUresult rc = CUDA_SUCCESS;
                if (c.mode == 0) {
                    cuuint64_t dims[3] = {(cuuint64_t)D, (cuuint64_t)N, (cuuint64_t)R};
                    cuuint64_t strides[2] = {(cuuint64_t)D * 4, (cuuint64_t)N * D * 4};
                    cuuint32_t box[3] = {32, 128, 1}; cuuint32_t es[3] = {1, 1, 1};
                    rc = enc(&tms[i], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, E[i], dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
                } else if (c.mode == 1) {
                    cuuint64_t dims[4] = {32, (cuuint64_t)N, (cuuint64_t)D / 32, (cuuint64_t)R};
                    cuuint64_t strides[3] = {(cuuint64_t)D * 4, 128, (cuuint64_t)N * D * 4};
                    cuuint32_t box[4] = {32, 128, (cuuint32_t)c.p0, 1}; cuuint32_t es[4] = {1, 1, 1, 1};
                    rc = enc(&tms[i], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, E[i], dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
                } else {
                    cuuint64_t dims[3] = {(cuuint64_t)D, (cuuint64_t)N, (cuuint64_t)R};
                    cuuint64_t strides[2] = {(cuuint64_t)D * 4, (cuuint64_t)N * D * 4};
                    cuuint32_t box[3] = {256, (cuuint32_t)(c.mode == 2 ? c.p0 : 16), 1}; cuuint32_t es[3] = {1, 1, 1};
                    rc = enc(&tms[i], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, E[i], dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
                }
                if (rc != CUDA_SUCCESS) { printf("encode failed rc=%d for %s\n", (int)rc, c.name); }
            }
            const size_t smem = slots * slotb + 16 * slots + 64;
            for (int w = 0; w < 2; ++w) probe<<<sms, 64, smem>>>(tms[w % 2], E[w % 2], R, N, D, c.mode, c.p0, c.p1, slots, (int)slotb);
            CK(cudaDeviceSynchronize());
            const int reps = 6;
            CK(cudaEventRecord(e0));
            for (int w = 0; w < reps; ++w) probe<<<sms, 64, smem>>>(tms[w % 2], E[w % 2], R, N, D, c.mode, c.p0, c.p1, slots, (int)slotb);
            CK(cudaEventRecord(e1));
            CK(cudaDeviceSynchronize());
            float ms = 0; CK(cudaEventElapsedTime(&ms, e0, e1));
            const double us = ms * 1e3 / reps;
            printf("%-44s %6d %8.1f %10.1f %8.0f\n", c.name, slots, slotb / 1024.0, us, bytes / us / 1e3);
            fflush(stdout);
        }
    }
    return 0;
}
