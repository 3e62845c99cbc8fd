#include "hostcopy.h"

#include <cstdint>
#include <cstring>
#if defined(__x86_64__)
#include <immintrin.h>
#endif

namespace phmm {
#if defined(__x86_64__)
__attribute__((target("avx2"))) static void copy_stream_avx2(uint8_t *d, const uint8_t *s, std::size_t n) {
    std::size_t head = (32 - ((uintptr_t)d & 31)) & 31;       // streaming stores need an aligned destination
    std::memcpy(d, s, head);
    d += head; s += head; n -= head;
    std::size_t i = 0;
    for (; i + 128 <= n; i += 128) {
        const __m256i a = _mm256_loadu_si256((const __m256i *)(s + i)), b = _mm256_loadu_si256((const __m256i *)(s + i + 32)),
                      c = _mm256_loadu_si256((const __m256i *)(s + i + 64)), e = _mm256_loadu_si256((const __m256i *)(s + i + 96));
        _mm256_stream_si256((__m256i *)(d + i), a); _mm256_stream_si256((__m256i *)(d + i + 32), b);
        _mm256_stream_si256((__m256i *)(d + i + 64), c); _mm256_stream_si256((__m256i *)(d + i + 96), e);
    }
    std::memcpy(d + i, s + i, n - i);
}
static const bool g_avx2 = __builtin_cpu_supports("avx2");
void copy_stream(void *dst, const void *src, std::size_t n) {
    if (n < 512 || !g_avx2) { std::memcpy(dst, src, n); return; }
    copy_stream_avx2(static_cast<uint8_t *>(dst), static_cast<const uint8_t *>(src), n);
}
void copy_stream_fence() { _mm_sfence(); }
#else
void copy_stream(void *dst, const void *src, std::size_t n) { std::memcpy(dst, src, n); }
void copy_stream_fence() {}
#endif
}  // namespace phmm
