// Integer-pipe throughput probes: the measured denominator of the INT32-ALU
// roofline (MEASURED_PEAKS.json carries no integer entry).  Eight independent
// chains per thread so that neither the 4-cycle ALU latency nor a single
// register dependency limits issue.  Modes:
//   0  ALU only       : SHF + LOP3 + IADD3 per chain per step   (24 instr/step)
//   1  ALU + FMA pipe : the same plus one IMAD per chain        (32 instr/step)
//   2  FMA pipe only  : IMAD per chain                          ( 8 instr/step)
//   3  IMAD.HI only   : mul.hi-add per chain                    ( 8 instr/step)
//   4  ALU + 2 IMAD   : SHF + LOP3 + IADD3 + 2 IMAD per chain   (40 instr/step)
//   5  IMAD.WIDE only : 32x32+64 -> 64 per chain                ( 8 instr/step)
//   6  ALU + IMAD 1:1 : SHF + LOP3 + 2 IMAD per chain           (32 instr/step)
// The instruction mix per step is verified against the SASS in
// tests/test_sass.py; `ops_per_thread_step` below must match it.
#pragma once
#include <stdint.h>

namespace cas {

constexpr int PEAK_CHAINS = 8;
constexpr int PEAK_THREADS = 256;
constexpr int PEAK_UNROLL = 8;

__host__ __device__ constexpr int peak_ops_per_step(int mode) {
    return mode == 0 ? 24 : (mode == 1 || mode == 6) ? 32 : mode == 4 ? 40 : 8;
}

template <int MODE>
__global__ void __launch_bounds__(PEAK_THREADS) kernel_peak(uint32_t* sink, int steps, uint32_t one, uint32_t seed) {
    uint32_t x[PEAK_CHAINS], y[PEAK_CHAINS];
#pragma unroll
    for (int c = 0; c < PEAK_CHAINS; ++c) {
        x[c] = seed + threadIdx.x * 0x9E3779B9u + c * 0x85EBCA6Bu;
        y[c] = x[c] ^ (blockIdx.x + 0x27D4EB2Fu * c);
    }
#pragma unroll 1
    for (int s = 0; s < steps; s += PEAK_UNROLL) {
#pragma unroll
        for (int u = 0; u < PEAK_UNROLL; ++u) {
            if (MODE == 0 || MODE == 1 || MODE == 4) {
#pragma unroll
                for (int c = 0; c < PEAK_CHAINS; ++c)
                    asm volatile("shf.r.wrap.b32 %0, %0, %0, 7;" : "+r"(x[c]));
#pragma unroll
                for (int c = 0; c < PEAK_CHAINS; ++c)
                    asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(y[c]) : "r"(x[c]), "r"(one));
#pragma unroll
                for (int c = 0; c < PEAK_CHAINS; ++c)
                    asm volatile("{.reg .u32 t; add.u32 t, %0, %1; add.u32 %0, t, %2;}" : "+r"(x[c]) : "r"(y[c]), "r"(one));
            }
            if (MODE == 1 || MODE == 2 || MODE == 4) {
#pragma unroll
                for (int c = 0; c < PEAK_CHAINS; ++c)
                    asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(y[c]) : "r"(one), "r"(x[c]));
            }
            if (MODE == 4) {
#pragma unroll
                for (int c = 0; c < PEAK_CHAINS; ++c)
                    asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[c]) : "r"(one), "r"(y[c]));
            }
            if (MODE == 5) {
#pragma unroll
                for (int c = 0; c < PEAK_CHAINS; c += 2)
                    asm volatile("{.reg .u64 t; mov.b64 t, {%0, %1}; mad.wide.u32 t, %0, %2, t; mov.b64 {%0, %1}, t;"
                                 " mov.b64 t, {%3, %4}; mad.wide.u32 t, %3, %2, t; mov.b64 {%3, %4}, t;}"
                                 : "+r"(x[c]), "+r"(y[c]), "+r"(seed), "+r"(x[c + 1]), "+r"(y[c + 1]));
            }
            if (MODE == 6) {
#pragma unroll
                for (int c = 0; c < PEAK_CHAINS; ++c)
                    asm volatile("shf.r.wrap.b32 %0, %0, %0, 7;" : "+r"(x[c]));
#pragma unroll
                for (int c = 0; c < PEAK_CHAINS; ++c)
                    asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(y[c]) : "r"(one), "r"(x[c]));
#pragma unroll
                for (int c = 0; c < PEAK_CHAINS; ++c)
                    asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[c]) : "r"(y[c]), "r"(one));
#pragma unroll
                for (int c = 0; c < PEAK_CHAINS; ++c)
                    asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(y[c]) : "r"(one), "r"(x[c]));
            }
            if (MODE == 3) {
#pragma unroll
                for (int c = 0; c < PEAK_CHAINS; ++c)
                    asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(y[c]) : "r"(seed), "r"(x[c]));
            }
        }
    }
    uint32_t acc = 0;
#pragma unroll
    for (int c = 0; c < PEAK_CHAINS; ++c) acc ^= x[c] ^ y[c];
    if (acc == 0x12345678u) sink[0] = acc;  // keeps the chains alive; practically never taken
}

}  // namespace cas
