// See hostcopy.h.  Plain C++ (compiled by the host compiler, no device code).
#include "hostcopy.h"

#include <string.h>

#if defined(__x86_64__)
#include <immintrin.h>
#endif

namespace cas {

#if defined(__x86_64__)
__attribute__((target("avx2"))) static void copy_nt_avx2(uint8_t* dst, const uint8_t* src, size_t bytes) {
    // head: up to the first 32-byte boundary of the destination
    const size_t head = (32 - ((uintptr_t)dst & 31)) & 31;
    if (head) {
        memcpy(dst, src, head);
        dst += head; src += head; bytes -= head;
    }
    size_t i = 0;
    for (; i + 128 <= bytes; i += 128) {
        const __m256i a = _mm256_loadu_si256((const __m256i*)(src + i));
        const __m256i b = _mm256_loadu_si256((const __m256i*)(src + i + 32));
        const __m256i c = _mm256_loadu_si256((const __m256i*)(src + i + 64));
        const __m256i d = _mm256_loadu_si256((const __m256i*)(src + i + 96));
        _mm256_stream_si256((__m256i*)(dst + i), a);
        _mm256_stream_si256((__m256i*)(dst + i + 32), b);
        _mm256_stream_si256((__m256i*)(dst + i + 64), c);
        _mm256_stream_si256((__m256i*)(dst + i + 96), d);
    }
    for (; i + 32 <= bytes; i += 32)
        _mm256_stream_si256((__m256i*)(dst + i), _mm256_loadu_si256((const __m256i*)(src + i)));
    _mm_sfence();  // the stores must be globally visible before the DMA that follows the pack
    if (i < bytes) memcpy(dst + i, src + i, bytes - i);
}

static bool have_avx2() {
    static const bool yes = __builtin_cpu_supports("avx2");
    return yes;
}
#endif

void copy_to_staging(uint8_t* dst, const uint8_t* src, size_t bytes) {
#if defined(__x86_64__)
    if (bytes >= 4096 && have_avx2()) {
        copy_nt_avx2(dst, src, bytes);
        return;
    }
#endif
    memcpy(dst, src, bytes);
}

}  // namespace cas
