// Timing harness for the warp-per-message kernel (csrc/sha256_long.cuh) against the
// lane-per-message kernel on the same few long objects.  One JSON line per run.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include <vector>

#include "../bee_code_interpreter_fs_b200/csrc/sha256_long.cuh"
#include "../bee_code_interpreter_fs_b200/csrc/synth.cuh"

#ifndef KB_ADDS
#define KB_ADDS 0
#endif
#ifndef KB_FMA
#define KB_FMA 3
#endif
#ifndef KB_M
#define KB_M 1
#endif
#ifndef KB_ACT
#define KB_ACT 32
#endif
#ifndef KB_ROTF
#define KB_ROTF 0
#endif
#ifndef KB_NAME
#define KB_NAME "long"
#endif

#define CK(x)                                                               \
    do {                                                                    \
        cudaError_t e_ = (x);                                               \
        if (e_ != cudaSuccess) {                                            \
            fprintf(stderr, "%s failed: %s\n", #x, cudaGetErrorString(e_)); \
            return 1;                                                       \
        }                                                                   \
    } while (0)

static uint64_t checksum(const uint8_t* d_out, uint64_t n) {
    std::vector<uint32_t> h(n * 8);
    cudaMemcpy(h.data(), d_out, n * 32, cudaMemcpyDeviceToHost);
    uint32_t x = 0, s = 0;
    for (uint64_t i = 0; i < n * 8; ++i) { x ^= h[i] * (uint32_t)(2 * i + 1); s += h[i]; }
    return ((uint64_t)x << 32) | s;
}

int main(int argc, char** argv) {
    const uint64_t L = argc > 1 ? strtoull(argv[1], 0, 10) : (16ull << 20) + 5;
    uint8_t *buf, *out;
    const uint64_t nmax = argc > 2 ? strtoull(argv[2], 0, 10) : 2368;
    const uint64_t stride = (L + 15) & ~15ull;
    CK(cudaMalloc(&buf, nmax * stride));
    CK(cudaMalloc(&out, nmax * 32));
    const uint64_t segs = (L + cas::SYNTH_SEG - 1) / cas::SYNTH_SEG;
    cas::kernel_synth<<<(unsigned)(nmax * segs), cas::SYNTH_THREADS>>>(buf, 0xB200, 0, nmax, L, stride, segs);
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    std::vector<uint64_t> counts;
    if (argc > 3) {
        for (int i = 3; i < argc; ++i) counts.push_back(strtoull(argv[i], 0, 10));
    } else {
        counts = {1, 64, 148, 592, 1184, 2368};
    }
    for (uint64_t n : counts) {
        if (n > nmax) continue;
        cas::BatchArgs a{};
        a.base = buf; a.stride = stride; a.ulen = L; a.n = n; a.out = out; a.one = 1; a.zero = 0;
        float ms_long = 0, ms_lane = 0;
        CK(cudaMemset(out, 0, n * 32));
        CK(cudaEventRecord(e0));
        cas::kernel_long<KB_FMA, KB_ADDS, KB_M, KB_ACT, KB_ROTF><<<(unsigned)((n + KB_M - 1) / KB_M), 32, cas::LONG_SMEM_PER_WARP>>>(a);
        CK(cudaGetLastError());
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        CK(cudaEventElapsedTime(&ms_long, e0, e1));
        const uint64_t c_long = checksum(out, n);
        CK(cudaMemset(out, 0, n * 32));
        CK(cudaEventRecord(e0));
        cas::kernel_ldg<3, 32, 16><<<(unsigned)((n + 31) / 32), 32>>>(a);  // one warp per CTA, as the engine launches small batches
        CK(cudaGetLastError());
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        CK(cudaEventElapsedTime(&ms_lane, e0, e1));
        const uint64_t c_lane = checksum(out, n);
        printf("{\"variant\": \"%s\", \"objects\": %llu, \"object_bytes\": %llu, \"warp_per_msg_ms\": %.2f, "
               "\"warp_per_msg_gb_s\": %.2f, \"per_chain_mb_s\": %.1f, \"cycles_per_block_at_1965\": %.0f, "
               "\"lane_per_msg_ms\": %.2f, \"lane_per_msg_gb_s\": %.2f, \"same_digests\": %s}\n",
               KB_NAME, (unsigned long long)n, (unsigned long long)L, ms_long, n * L / ms_long / 1e6, L / ms_long / 1e3,
               ms_long * 1e-3 * 1.965e9 / ((L + 72) / 64), ms_lane, n * L / ms_lane / 1e6, c_long == c_lane ? "true" : "false");
        fflush(stdout);
    }
    return 0;
}
