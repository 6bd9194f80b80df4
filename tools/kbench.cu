// Standalone timing harness for variants of kernel_tma (csrc/sha256_batch.cuh).
// Built many times with different -DKB_* values by tools/kbench_build.py; each binary
// prints one JSON line: variant, GB/s (best of reps), and a checksum of all digests.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "../bee_code_interpreter_fs_b200/csrc/sha256_batch.cuh"
#include "../bee_code_interpreter_fs_b200/csrc/synth.cuh"

#ifndef KB_FMA
#define KB_FMA 3
#endif
#ifndef KB_NSTAGE
#define KB_NSTAGE 2
#endif
#ifndef KB_THREADS
#define KB_THREADS 256
#endif
#ifndef KB_MINCTAS
#define KB_MINCTAS 2
#endif
#ifndef KB_MSGS
#define KB_MSGS 1
#endif
#ifndef KB_SCHED
#define KB_SCHED 0
#endif
#ifndef KB_PADV
#define KB_PADV 1
#endif
#ifndef KB_REGCAP
#define KB_REGCAP 0
#endif
#ifndef KB_NAME
#define KB_NAME "variant"
#endif

#define CK(x)                                                                                  \
    do {                                                                                       \
        cudaError_t e_ = (x);                                                                  \
        if (e_ != cudaSuccess) {                                                               \
            fprintf(stderr, "%s failed: %s\n", #x, cudaGetErrorString(e_));                    \
            return 1;                                                                          \
        }                                                                                      \
    } while (0)

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main(int argc, char** argv) {
    const uint64_t n = argc > 1 ? strtoull(argv[1], 0, 10) : 303104;
    const uint64_t L = argc > 2 ? strtoull(argv[2], 0, 10) : 65536;
    const int reps = argc > 3 ? atoi(argv[3]) : 5;
    uint8_t *buf, *out;
    CK(cudaMalloc(&buf, n * L));
    CK(cudaMalloc(&out, n * 32));
    const uint64_t segs = (L + cas::SYNTH_SEG - 1) / cas::SYNTH_SEG;
    cas::kernel_synth<<<(unsigned)(n * segs), cas::SYNTH_THREADS>>>(buf, 0xB200, 0, n, L, L, segs);
    CK(cudaDeviceSynchronize());
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
    constexpr int MSGS = KB_MSGS;
    CUtensorMap tm;
    const cuuint64_t gdim[2] = {L, n};
    const cuuint64_t gstride[1] = {L};
    const cuuint32_t box[2] = {64u, 32u * MSGS};
    const cuuint32_t estr[2] = {1u, 1u};
    if (((EncodeTiledFn)fn)(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, buf, gdim, gstride, box, estr,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
        fprintf(stderr, "tensor map encode failed\n");
        return 1;
    }
    cas::TensorMapBlob blob;
    memcpy(&blob, &tm, sizeof tm);
    cas::BatchArgs a{};
    a.base = buf; a.stride = L; a.ulen = L; a.n = n; a.out = out; a.one = 1; a.zero = 0;
#if KB_REGCAP
    auto k = cas::kernel_tma_regcap<KB_FMA, KB_NSTAGE, KB_THREADS, KB_REGCAP, KB_MSGS, KB_SCHED, KB_PADV>;
#else
    auto k = cas::kernel_tma<KB_FMA, KB_NSTAGE, KB_THREADS, KB_MINCTAS, KB_MSGS, KB_SCHED, KB_PADV>;
#endif
    constexpr int smem = cas::tma_smem_bytes<KB_NSTAGE, KB_THREADS, KB_MSGS>();
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    CK(cudaFuncSetAttribute(k, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    const uint64_t per_cta = (uint64_t)KB_THREADS * MSGS;
    const unsigned grid = (unsigned)((n + per_cta - 1) / per_cta);
    cudaFuncAttributes fa;
    CK(cudaFuncGetAttributes(&fa, k));
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k, KB_THREADS, smem));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int r = 0; r < reps + 1; ++r) {
        CK(cudaEventRecord(e0));
        k<<<grid, KB_THREADS, smem>>>(blob, a);
        CK(cudaGetLastError());
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (r > 0 && ms < best) best = ms;
    }
    std::vector<uint32_t> h(n * 8);
    CK(cudaMemcpy(h.data(), out, n * 32, cudaMemcpyDeviceToHost));
    uint32_t x = 0, s = 0;
    for (uint64_t i = 0; i < n * 8; ++i) { x ^= h[i] * (uint32_t)(2 * i + 1); s += h[i]; }
    printf("{\"variant\": \"%s\", \"gb_s\": %.1f, \"ms\": %.3f, \"regs\": %d, \"ctas_per_sm\": %d, \"smem\": %d, "
           "\"checksum\": \"%08x%08x\", \"objects\": %llu, \"object_bytes\": %llu}\n",
           KB_NAME, n * L / best / 1e6, best, fa.numRegs, occ, smem, x, s, (unsigned long long)n, (unsigned long long)L);
    return 0;
}
