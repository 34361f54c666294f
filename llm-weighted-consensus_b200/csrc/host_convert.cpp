// host_convert.cpp -- host-side element conversion for the staging entry points (lwc_stage_f64,
// lwc_stage_f64_bf16): Embedding.embedding is Vec<f64> (src/embeddings/response.rs:21), the device wants fp32
// (or, opt-in, bf16).  Called per embedding row by many submitting threads, so the loops are single-threaded,
// branch-free and vectorised (AVX2 when the CPU has it, chosen once at run time; plain loops otherwise).
#include <cstddef>
#include <cstdint>
#include <cstring>

#if defined(__x86_64__)
#include <immintrin.h>
#endif

namespace lwc_host {

static void f64_to_f32_plain(const double *src, size_t n, float *dst) {
    for (size_t i = 0; i < n; ++i) dst[i] = static_cast<float>(src[i]);
}

static inline uint16_t bf16_bits(float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    // nearest even on the upper 16 bits; NaN keeps a quiet payload
    return (u & 0x7fffffffu) > 0x7f800000u ? (uint16_t)((u >> 16) | 0x40u) : (uint16_t)((u + 0x7fffu + ((u >> 16) & 1u)) >> 16);
}

static void f64_to_bf16_plain(const double *src, size_t n, uint16_t *dst) {
    for (size_t i = 0; i < n; ++i) dst[i] = bf16_bits(static_cast<float>(src[i]));
}

#if defined(__x86_64__)
__attribute__((target("avx2"))) static void f64_to_f32_avx2(const double *src, size_t n, float *dst) {
    size_t i = 0;
    // A whole embedding row going into a pinned staging arena is written once and next read by the DMA engine:
    // streaming stores skip the read-for-ownership of every destination line (a third of the pass's memory traffic).
    if (n >= 256 && (reinterpret_cast<uintptr_t>(dst) & 31u) == 0) {
        for (; i + 8 <= n; i += 8) {
            const __m128 a = _mm256_cvtpd_ps(_mm256_loadu_pd(src + i));
            const __m128 b = _mm256_cvtpd_ps(_mm256_loadu_pd(src + i + 4));
            _mm256_stream_ps(dst + i, _mm256_set_m128(b, a));
        }
        _mm_sfence();
    }
    for (; i + 8 <= n; i += 8) {
        const __m128 a = _mm256_cvtpd_ps(_mm256_loadu_pd(src + i));
        const __m128 b = _mm256_cvtpd_ps(_mm256_loadu_pd(src + i + 4));
        _mm256_storeu_ps(dst + i, _mm256_set_m128(b, a));
    }
    for (; i < n; ++i) dst[i] = static_cast<float>(src[i]);
}

__attribute__((target("avx2"))) static void f64_to_bf16_avx2(const double *src, size_t n, uint16_t *dst) {
    size_t i = 0;
    const __m256i bias = _mm256_set1_epi32(0x7fff), one = _mm256_set1_epi32(1), absmask = _mm256_set1_epi32(0x7fffffff),
                  inf = _mm256_set1_epi32(0x7f800000), quiet = _mm256_set1_epi32(0x40);
    for (; i + 8 <= n; i += 8) {
        const __m128 a = _mm256_cvtpd_ps(_mm256_loadu_pd(src + i));
        const __m128 b = _mm256_cvtpd_ps(_mm256_loadu_pd(src + i + 4));
        const __m256i u = _mm256_castps_si256(_mm256_set_m128(b, a));
        const __m256i hi = _mm256_srli_epi32(u, 16);
        const __m256i rounded = _mm256_srli_epi32(_mm256_add_epi32(_mm256_add_epi32(u, bias), _mm256_and_si256(hi, one)), 16);
        const __m256i isnan = _mm256_cmpgt_epi32(_mm256_and_si256(u, absmask), inf);
        const __m256i r = _mm256_blendv_epi8(rounded, _mm256_or_si256(hi, quiet), isnan);
        // 8 x u32 (values < 65536) -> 8 x u16, in order
        const __m256i packed = _mm256_permute4x64_epi64(_mm256_packus_epi32(r, r), 0xD8);
        _mm_storeu_si128(reinterpret_cast<__m128i *>(dst + i), _mm256_castsi256_si128(packed));
    }
    for (; i < n; ++i) dst[i] = bf16_bits(static_cast<float>(src[i]));
}

static bool has_avx2() {
    static const bool v = __builtin_cpu_supports("avx2");
    return v;
}
#endif

void convert_f64_f32(const double *src, size_t n, float *dst) {
#if defined(__x86_64__)
    if (has_avx2()) return f64_to_f32_avx2(src, n, dst);
#endif
    f64_to_f32_plain(src, n, dst);
}

void convert_f64_bf16(const double *src, size_t n, uint16_t *dst) {
#if defined(__x86_64__)
    if (has_avx2()) return f64_to_bf16_avx2(src, n, dst);
#endif
    f64_to_bf16_plain(src, n, dst);
}

} // namespace lwc_host
