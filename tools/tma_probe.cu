// tma_probe.cu -- microbenchmark: how fast can one persistent CTA per SM stream the cfg2 tensor
// E[R][16][1536] fp32 from HBM into shared memory, by copy flavour?  No math: a single consumer
// thread per CTA releases each stage as soon as it lands.  Used to pick the staging scheme of
// score_stream.cuh; results are recorded in profiles/.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tools/tma_probe tools/tma_probe.cu
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x)                                                                                  \
    do {                                                                                       \
        cudaError_t e_ = (x);                                                                  \
        if (e_ != cudaSuccess) {                                                               \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__);    \
            exit(1);                                                                           \
        }                                                                                      \
    } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    long long t0 = clock64();
    while (!mbar_try(bar, parity))
        if (clock64() - t0 > 2000000000LL) __trap();
}

// mode 0: 3-D tensor box {kc, 16, 1}; mode 1: 1-D bulk, 16 row copies of kc*4 bytes per stage (one lane);
// mode 2: 1-D bulk, the 16 row copies issued by 16 lanes; mode 3: 2-D tensor box {kc,16} over [R*16][D]
__global__ void __launch_bounds__(64, 1)
probe(const __grid_constant__ CUtensorMap tm, const float *E, int R, int N, int D, int kc, int slots, int mode,
      unsigned long long *bytes_out) {
    extern __shared__ __align__(1024) unsigned char smem[];
    const int slotb = 16 * kc * 4;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + (size_t)slots * slotb);
    const uint32_t full0 = smem_u32(bars), empty0 = full0 + 8 * slots;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int r0 = (int)((long long)blockIdx.x * R / gridDim.x), r1 = (int)((long long)(blockIdx.x + 1) * R / gridDim.x);
    const int spr = (D + kc - 1) / kc;
    if (threadIdx.x == 0) {
        for (int s = 0; s < slots; ++s) {
            mbar_init(full0 + 8 * s, 1);
            mbar_init(empty0 + 8 * s, 1);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const uint32_t ring = smem_u32(smem);
    if (warp == 0) {
        uint32_t g = 0;
        for (int r = r0; r < r1; ++r)
            for (int s = 0; s < spr; ++s, ++g) {
                const uint32_t slot = g % slots, round = g / slots;
                const int kv = min(kc, D - s * kc);
                if (mode == 0 || mode == 3) {
                    if (lane == 0) {
                        mbar_wait(empty0 + 8 * slot, (round & 1) ^ 1);
                        mbar_expect(full0 + 8 * slot, slotb);
                        if (mode == 0)
                            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                                         ::"r"(ring + slot * slotb), "l"(&tm), "r"(s * kc), "r"(0), "r"(r), "r"(full0 + 8 * slot) : "memory");
                        else
                            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                                         ::"r"(ring + slot * slotb), "l"(&tm), "r"(s * kc), "r"(r * 16), "r"(full0 + 8 * slot) : "memory");
                    }
                } else if (mode == 1) {
                    if (lane == 0) {
                        mbar_wait(empty0 + 8 * slot, (round & 1) ^ 1);
                        mbar_expect(full0 + 8 * slot, 16 * kv * 4);
                        for (int row = 0; row < 16; ++row)
                            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                         ::"r"(ring + slot * slotb + row * kc * 4), "l"(E + ((size_t)r * 16 + row) * D + s * kc),
                                           "r"(kv * 4), "r"(full0 + 8 * slot) : "memory");
                    }
                } else {
                    mbar_wait(empty0 + 8 * slot, (round & 1) ^ 1);
                    if (lane == 0) mbar_expect(full0 + 8 * slot, 16 * kv * 4);
                    __syncwarp();
                    if (lane < 16)
                        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                     ::"r"(ring + slot * slotb + lane * kc * 4), "l"(E + ((size_t)r * 16 + lane) * D + s * kc),
                                       "r"(kv * 4), "r"(full0 + 8 * slot) : "memory");
                }
            }
    } else if (lane == 0) {
        uint32_t g = 0;
        unsigned long long acc = 0;
        for (int r = r0; r < r1; ++r)
            for (int s = 0; s < spr; ++s, ++g) {
                const uint32_t slot = g % slots, round = g / slots;
                mbar_wait(full0 + 8 * slot, round & 1);
                acc += *reinterpret_cast<volatile unsigned int *>(smem + (size_t)slot * slotb); // touch
                mbar_arrive(empty0 + 8 * slot);
            }
        if (acc == 0x12345678ULL) bytes_out[0] = acc;
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
    const int R = 4096, N = 16, D = 1536;
    const size_t bytes = (size_t)R * N * D * 4;
    float *E[3];
    for (int i = 0; i < 3; ++i) {
        CK(cudaMalloc(&E[i], bytes));
        CK(cudaMemset(E[i], i + 1, bytes));
    }
    unsigned long long *out;
    CK(cudaMalloc(&out, 8));
    int sms = 0, maxsmem = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    CK(cudaDeviceGetAttribute(&maxsmem, cudaDevAttrMaxSharedMemoryPerBlockOptin, 0));
    CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, maxsmem));
    void *sym = nullptr;
    cudaDriverEntryPointQueryResult qr;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qr));
    EncodeTiledFn enc = (EncodeTiledFn)sym;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    printf("SMs=%d maxsmem=%d  tensor bytes=%zu\n", sms, maxsmem, bytes);
    printf("%-28s %6s %6s %10s %10s\n", "mode", "kc", "slots", "us/pass", "GB/s");
    const char *names[4] = {"tensor3d {kc,16,1}", "bulk1d one lane x16 rows", "bulk1d 16 lanes", "tensor2d {kc,16}"};
    int kcs[] = {64, 128, 192, 256, 384, 768, 1536};
    for (int mode = 0; mode < 4; ++mode)
        for (int kc : kcs) {
            if ((mode == 0 || mode == 3) && kc > 256) continue;
            for (int want : {2, 4, 8, 16}) {
                const size_t slotb = (size_t)16 * kc * 4;
                int slots = want;
                if (slots * slotb + 1024 > (size_t)maxsmem - 2048) continue;
                CUtensorMap tms[3];
                for (int i = 0; i < 3; ++i) {
                    if (mode == 3) {
                        cuuint64_t dims[2] = {(cuuint64_t)D, (cuuint64_t)R * N};
                        cuuint64_t strides[1] = {(cuuint64_t)D * 4};
                        cuuint32_t box[2] = {(cuuint32_t)kc, 16};
                        cuuint32_t es[2] = {1, 1};
                        enc(&tms[i], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, E[i], dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                            CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
                    } else {
                        cuuint64_t dims[3] = {(cuuint64_t)D, (cuuint64_t)N, (cuuint64_t)R};
                        cuuint64_t strides[2] = {(cuuint64_t)D * 4, (cuuint64_t)N * D * 4};
                        cuuint32_t box[3] = {(cuuint32_t)(kc > 256 ? 256 : kc), 16, 1};
                        cuuint32_t es[3] = {1, 1, 1};
                        enc(&tms[i], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, E[i], dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                            CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
                    }
                }
                const size_t smem = slots * slotb + 16 * slots + 64;
                for (int w = 0; w < 3; ++w) probe<<<sms, 64, smem>>>(tms[w % 3], E[w % 3], R, N, D, kc, slots, mode, out);
                CK(cudaDeviceSynchronize());
                const int reps = 12;
                CK(cudaEventRecord(e0));
                for (int w = 0; w < reps; ++w) probe<<<sms, 64, smem>>>(tms[w % 3], E[w % 3], R, N, D, kc, slots, mode, out);
                CK(cudaEventRecord(e1));
                CK(cudaDeviceSynchronize());
                float ms = 0;
                CK(cudaEventElapsedTime(&ms, e0, e1));
                const double us = ms * 1e3 / reps;
                printf("%-28s %6d %6d %10.1f %10.0f\n", names[mode], kc, slots, us, bytes / us / 1e3);
                fflush(stdout);
            }
        }
    return 0;
}
