// convert.cuh -- element-type conversion of embeddings on the device (wire/staging formats).
// Embedding.embedding is Vec<f64> in the reference (src/embeddings/response.rs:21); the scoring kernels read
// fp32.  These two streaming kernels turn what crossed PCIe -- f64 (8 B per element) or bf16 (2 B) -- into the
// fp32 [R][N][D] tile the kernels stream.  Both are HBM-bound copies: f64 -> fp32 moves 12 B per element,
// bf16 -> fp32 6 B; 128-bit accesses, grid = a multiple of the SM count, grid-stride.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace lwc {

// round to nearest even, like the host's static_cast<float>(double): device and host conversion give the same bits
__global__ void __launch_bounds__(256) convert_f64_f32_kernel(const double *__restrict__ src, float *__restrict__ dst, size_t n) {
    const size_t n4 = n / 4;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        const double2 a = reinterpret_cast<const double2 *>(src)[2 * i];
        const double2 b = reinterpret_cast<const double2 *>(src)[2 * i + 1];
        reinterpret_cast<float4 *>(dst)[i] = make_float4(__double2float_rn(a.x), __double2float_rn(a.y), __double2float_rn(b.x),
                                                         __double2float_rn(b.y));
    }
    if (blockIdx.x == 0 && threadIdx.x < (unsigned)(n - 4 * n4)) dst[4 * n4 + threadIdx.x] = __double2float_rn(src[4 * n4 + threadIdx.x]);
}

// exact: a bf16 is the upper half of an fp32
__global__ void __launch_bounds__(256) expand_bf16_f32_kernel(const uint16_t *__restrict__ src, float *__restrict__ dst, size_t n) {
    const size_t n8 = n / 8;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += stride) {
        const uint4 v = reinterpret_cast<const uint4 *>(src)[i];
        float4 lo, hi;
        lo.x = __uint_as_float(v.x << 16), lo.y = __uint_as_float(v.x & 0xffff0000u);
        lo.z = __uint_as_float(v.y << 16), lo.w = __uint_as_float(v.y & 0xffff0000u);
        hi.x = __uint_as_float(v.z << 16), hi.y = __uint_as_float(v.z & 0xffff0000u);
        hi.z = __uint_as_float(v.w << 16), hi.w = __uint_as_float(v.w & 0xffff0000u);
        reinterpret_cast<float4 *>(dst)[2 * i] = lo;
        reinterpret_cast<float4 *>(dst)[2 * i + 1] = hi;
    }
    if (blockIdx.x == 0 && threadIdx.x < (unsigned)(n - 8 * n8))
        dst[8 * n8 + threadIdx.x] = __uint_as_float((uint32_t)src[8 * n8 + threadIdx.x] << 16);
}

} // namespace lwc
