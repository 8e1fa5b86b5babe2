// streamcopy.hpp -- the memcpy of the staging ring (hostcopy.hpp), on its own so that the host unit-test shim can exercise it.
#pragma once
#include <cstdint>
#include <cstring>
#if defined(__x86_64__) && !defined(__CUDA_ARCH__)
#include <immintrin.h>
#define H2B_HAVE_AVX2_STREAM 1
#endif

namespace h2b {

// memcpy for staging pieces: both sides are touched once and are far larger than the caches, so the destination is written with
// non-temporal stores (no read-for-ownership of the destination lines: 2 bytes of DRAM traffic per byte copied instead of 3).
#if defined(H2B_HAVE_AVX2_STREAM)
__attribute__((target("avx2"))) inline void stream_copy_avx2(char* dst, const char* src, size_t len) {
    size_t head = (32 - (reinterpret_cast<uintptr_t>(dst) & 31)) & 31;
    if (head > len) head = len;
    if (head) { std::memcpy(dst, src, head); dst += head; src += head; len -= head; }
    size_t i = 0;
    for (; i + 128 <= len; i += 128) {
        __m256i a = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(src + i));
        __m256i b = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(src + i + 32));
        __m256i c = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(src + i + 64));
        __m256i d = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(src + i + 96));
        _mm256_stream_si256(reinterpret_cast<__m256i*>(dst + i), a);
        _mm256_stream_si256(reinterpret_cast<__m256i*>(dst + i + 32), b);
        _mm256_stream_si256(reinterpret_cast<__m256i*>(dst + i + 64), c);
        _mm256_stream_si256(reinterpret_cast<__m256i*>(dst + i + 96), d);
    }
    _mm_sfence();
    if (i < len) std::memcpy(dst + i, src + i, len - i);
}
#endif
inline void stream_copy(void* dst, const void* src, size_t len) {
#if defined(H2B_HAVE_AVX2_STREAM)
    static const bool avx2 = __builtin_cpu_supports("avx2");
    if (avx2 && len >= 4096) { stream_copy_avx2(static_cast<char*>(dst), static_cast<const char*>(src), len); return; }
#endif
    std::memcpy(dst, src, len);
}

}  // namespace h2b
