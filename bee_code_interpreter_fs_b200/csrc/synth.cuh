// Device generator for the synthetic benchmark objects.
// Emits exactly the bytes of oracle/sha256_ref.c:oracle_gen_synthetic and
// oracle/sha256_oracle.py:synth_object, so GPU and CPU hash the same input
// without a PCIe copy:
//   word(seed, obj, j) = splitmix64_finalise(seed + GOLDEN * ((obj << 32) | j))
// stored little-endian, 8 bytes per word, truncated to the object length.
#pragma once
#include <stdint.h>

namespace cas {

__device__ __forceinline__ uint64_t synth_word(uint64_t seed, uint64_t obj, uint64_t j) {
    uint64_t z = seed + 0x9E3779B97F4A7C15ull * ((obj << 32) | (j & 0xFFFFFFFFull));
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// One CTA writes one 4 KiB segment of one object: 256 threads x 16 bytes,
// consecutive lanes on consecutive 16-byte pieces (fully coalesced stores).
constexpr int SYNTH_THREADS = 256;
constexpr int SYNTH_SEG = SYNTH_THREADS * 16;

__global__ void __launch_bounds__(SYNTH_THREADS) kernel_synth(uint8_t* dst, uint64_t seed, uint64_t first_obj,
                                                              uint64_t n, uint64_t len, uint64_t stride,
                                                              uint64_t segs_per_obj) {
    const uint64_t blk = blockIdx.x;
    const uint64_t i = blk / segs_per_obj;
    const uint64_t seg = blk - i * segs_per_obj;
    if (i >= n) return;
    const uint64_t b = seg * SYNTH_SEG + (uint64_t)threadIdx.x * 16u;
    if (b >= len) return;
    const uint64_t obj = first_obj + i;
    const uint64_t w0 = synth_word(seed, obj, b >> 3);
    const uint64_t w1 = synth_word(seed, obj, (b >> 3) + 1);
    uint8_t* p = dst + i * stride + b;
    if (b + 16 <= len && ((uintptr_t)p & 15u) == 0) {
        uint4 v;
        v.x = (uint32_t)w0; v.y = (uint32_t)(w0 >> 32); v.z = (uint32_t)w1; v.w = (uint32_t)(w1 >> 32);
        *reinterpret_cast<uint4*>(p) = v;
    } else {
        const uint64_t k = len - b < 16 ? len - b : 16;
        for (uint64_t q = 0; q < k; ++q) p[q] = (uint8_t)((q < 8 ? w0 : w1) >> (8 * (q & 7)));
    }
}

}  // namespace cas
