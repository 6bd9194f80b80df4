// Host-side copy into pinned staging (csrc/hostcopy.cpp).
#pragma once
#include <stddef.h>
#include <stdint.h>

namespace cas {

// memcpy for the pack threads.  The destination is pinned staging that the CPU never reads
// back (the DMA engine does), so large copies use non-temporal stores: no read-for-ownership of
// the destination lines and no pollution of the caches the other pack threads work in.
void copy_to_staging(uint8_t* dst, const uint8_t* src, size_t bytes);

}  // namespace cas
