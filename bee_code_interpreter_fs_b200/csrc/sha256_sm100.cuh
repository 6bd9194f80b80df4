// SHA-256 round/schedule primitives shaped for the sm_100a integer pipes.
//
// The function is FIPS 180-4 SHA-256, i.e. what CPython's hashlib.sha256
// computes and what the reference's Storage docstring promises for object ids
// (/root/reference/src/code_interpreter/services/storage.py:36; the reference
// itself assigns random ids at storage.py:52).
//
// Instruction shapes this file aims for (one per line of the FIPS formulae):
//   rotate          -> SHF.R.W.U32        (funnel shift, ALU pipe)
//   3-input boolean -> LOP3.LUT           (ALU pipe)
//   3-input add     -> IADD3              (ALU pipe)
//   2-input add     -> IMAD  x*1+y        (FMA pipe, when FMA_LEVEL >= 1)
// B200's SM sub-partition issues one instruction per clock but the ALU pipe and
// the FMA pipe each accept one warp instruction every two clocks, so moving the
// adds that do not need IADD3's third operand to the FMA pipe takes them off
// the pipe that bounds this kernel.  `one` is a run-time 1 so that ptxas cannot
// fold the multiply back into an ALU add.
#pragma once
#include <stdint.h>

namespace cas {

__device__ __forceinline__ uint32_t rotr32(uint32_t x, int n) { return __funnelshift_r(x, x, n); }

__device__ __forceinline__ uint32_t lop_xor3(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
// Ch(e,f,g) = (e & f) ^ (~e & g): LUT 0xCA
__device__ __forceinline__ uint32_t lop_ch(uint32_t e, uint32_t f, uint32_t g) {
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0xCA;" : "=r"(d) : "r"(e), "r"(f), "r"(g));
    return d;
}
// Maj(a,b,c) = (a & b) ^ (a & c) ^ (b & c): LUT 0xE8
__device__ __forceinline__ uint32_t lop_maj(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
// a + b on the FMA pipe: a * one + b with one == 1 at run time.  (With a literal 1 ptxas folds the
// multiply away and merges the adds back into IADD3 on the ALU pipe; with the 1 in constant memory or
// without inline asm it still keeps the multiplier in a vector register -- both tried, same SASS.)
__device__ __forceinline__ uint32_t fma_add(uint32_t a, uint32_t b, uint32_t one) {
    uint32_t d;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(one), "r"(b));
    return d;
}

__device__ __forceinline__ uint32_t big_sigma0(uint32_t x) { return lop_xor3(rotr32(x, 2), rotr32(x, 13), rotr32(x, 22)); }
__device__ __forceinline__ uint32_t big_sigma1(uint32_t x) { return lop_xor3(rotr32(x, 6), rotr32(x, 11), rotr32(x, 25)); }
__device__ __forceinline__ uint32_t small_sigma0(uint32_t x) { return lop_xor3(rotr32(x, 7), rotr32(x, 18), x >> 3); }
__device__ __forceinline__ uint32_t small_sigma1(uint32_t x) { return lop_xor3(rotr32(x, 17), rotr32(x, 19), x >> 10); }
// x >> n as the high half of x * 2^(32-n): IMAD.HI on the FMA pipe instead of SHF on the ALU pipe.
template <int N>
__device__ __forceinline__ uint32_t shr_fma(uint32_t x) {
    uint32_t d;
    asm("mul.hi.u32 %0, %1, %2;" : "=r"(d) : "r"(x), "n"(1u << (32 - N)));
    return d;
}
__device__ __forceinline__ uint32_t small_sigma0_fma(uint32_t x) { return lop_xor3(rotr32(x, 7), rotr32(x, 18), shr_fma<3>(x)); }
__device__ __forceinline__ uint32_t small_sigma1_fma(uint32_t x) { return lop_xor3(rotr32(x, 17), rotr32(x, 19), shr_fma<10>(x)); }

// FMA_LEVEL: 0 = every add where ptxas puts it; 1 = the 2-input adds on the FMA
// pipe; 2 = additionally h+W+K as two FMA-pipe adds; 3 = every add on the FMA pipe;
// 4 = level 3 plus the schedule's plain right shifts as IMAD.HI.
template <int FMA_LEVEL>
__device__ __forceinline__ void sha_round(uint32_t a, uint32_t b, uint32_t c, uint32_t& d, uint32_t e,
                                          uint32_t f, uint32_t g, uint32_t& h, uint32_t k, uint32_t w,
                                          uint32_t one) {
    const uint32_t s1 = big_sigma1(e);
    const uint32_t ch = lop_ch(e, f, g);
    const uint32_t s0 = big_sigma0(a);
    const uint32_t mj = lop_maj(a, b, c);
    if (FMA_LEVEL == 0) {
        const uint32_t t1 = h + s1 + ch + k + w;
        d += t1;
        h = t1 + s0 + mj;
    } else if (FMA_LEVEL == 1) {
        const uint32_t t1 = (h + k + w) + s1 + ch;
        d = fma_add(d, t1, one);
        h = t1 + s0 + mj;
    } else if (FMA_LEVEL == 2) {
        const uint32_t hw = fma_add(fma_add(h, w, one), k, one);
        const uint32_t t1 = hw + s1 + ch;
        d = fma_add(d, t1, one);
        h = t1 + s0 + mj;
    } else {
        const uint32_t hw = fma_add(fma_add(h, w, one), k, one);
        const uint32_t t1 = fma_add(fma_add(hw, ch, one), s1, one);
        d = fma_add(d, t1, one);
        h = fma_add(fma_add(t1, mj, one), s0, one);
    }
}

template <int FMA_LEVEL>
__device__ __forceinline__ uint32_t sha_sched(uint32_t w16, uint32_t w15, uint32_t w7, uint32_t w2, uint32_t one) {
    const uint32_t s0 = FMA_LEVEL >= 4 ? small_sigma0_fma(w15) : small_sigma0(w15);
    const uint32_t s1 = FMA_LEVEL >= 4 ? small_sigma1_fma(w2) : small_sigma1(w2);
    if (FMA_LEVEL == 0) return w16 + s0 + w7 + s1;
    if (FMA_LEVEL <= 2) return fma_add(w16 + w7 + s0, s1, one);
    return fma_add(fma_add(fma_add(w16, w7, one), s0, one), s1, one);
}

#define CAS_RND(a, b, c, d, e, f, g, h, i, kc) sha_round<FMA_LEVEL>(a, b, c, d, e, f, g, h, kc, w[(i)&15], one)
#define CAS_SCH(i) w[(i)&15] = sha_sched<FMA_LEVEL>(w[(i)&15], w[((i) + 1) & 15], w[((i) + 9) & 15], w[((i) + 14) & 15], one)
#define CAS_RND8(i, k0, k1, k2, k3, k4, k5, k6, k7) \
    CAS_RND(a, b, c, d, e, f, g, h, (i) + 0, k0);    \
    CAS_RND(h, a, b, c, d, e, f, g, (i) + 1, k1);    \
    CAS_RND(g, h, a, b, c, d, e, f, (i) + 2, k2);    \
    CAS_RND(f, g, h, a, b, c, d, e, (i) + 3, k3);    \
    CAS_RND(e, f, g, h, a, b, c, d, (i) + 4, k4);    \
    CAS_RND(d, e, f, g, h, a, b, c, (i) + 5, k5);    \
    CAS_RND(c, d, e, f, g, h, a, b, (i) + 6, k6);    \
    CAS_RND(b, c, d, e, f, g, h, a, (i) + 7, k7)
#define CAS_SCH8(i) \
    CAS_SCH((i) + 0); CAS_SCH((i) + 1); CAS_SCH((i) + 2); CAS_SCH((i) + 3); \
    CAS_SCH((i) + 4); CAS_SCH((i) + 5); CAS_SCH((i) + 6); CAS_SCH((i) + 7)

// One compression of the 16 big-endian words in w[] into st[8].  w[] is
// clobbered (it is the rolling 16-word window of the message schedule).
template <int FMA_LEVEL>
__device__ __forceinline__ void sha256_compress(uint32_t (&st)[8], uint32_t (&w)[16], uint32_t one) {
    uint32_t a = st[0], b = st[1], c = st[2], d = st[3], e = st[4], f = st[5], g = st[6], h = st[7];
    CAS_RND8(0, 0x428a2f98u, 0x71374491u, 0xb5c0fbcfu, 0xe9b5dba5u, 0x3956c25bu, 0x59f111f1u, 0x923f82a4u, 0xab1c5ed5u);
    CAS_RND8(8, 0xd807aa98u, 0x12835b01u, 0x243185beu, 0x550c7dc3u, 0x72be5d74u, 0x80deb1feu, 0x9bdc06a7u, 0xc19bf174u);
    CAS_SCH8(16);
    CAS_RND8(16, 0xe49b69c1u, 0xefbe4786u, 0x0fc19dc6u, 0x240ca1ccu, 0x2de92c6fu, 0x4a7484aau, 0x5cb0a9dcu, 0x76f988dau);
    CAS_SCH8(24);
    CAS_RND8(24, 0x983e5152u, 0xa831c66du, 0xb00327c8u, 0xbf597fc7u, 0xc6e00bf3u, 0xd5a79147u, 0x06ca6351u, 0x14292967u);
    CAS_SCH8(32);
    CAS_RND8(32, 0x27b70a85u, 0x2e1b2138u, 0x4d2c6dfcu, 0x53380d13u, 0x650a7354u, 0x766a0abbu, 0x81c2c92eu, 0x92722c85u);
    CAS_SCH8(40);
    CAS_RND8(40, 0xa2bfe8a1u, 0xa81a664bu, 0xc24b8b70u, 0xc76c51a3u, 0xd192e819u, 0xd6990624u, 0xf40e3585u, 0x106aa070u);
    CAS_SCH8(48);
    CAS_RND8(48, 0x19a4c116u, 0x1e376c08u, 0x2748774cu, 0x34b0bcb5u, 0x391c0cb3u, 0x4ed8aa4au, 0x5b9cca4fu, 0x682e6ff3u);
    CAS_SCH8(56);
    CAS_RND8(56, 0x748f82eeu, 0x78a5636fu, 0x84c87814u, 0x8cc70208u, 0x90befffau, 0xa4506cebu, 0xbef9a3f7u, 0xc67178f2u);
    if (FMA_LEVEL == 0) {
        st[0] += a; st[1] += b; st[2] += c; st[3] += d; st[4] += e; st[5] += f; st[6] += g; st[7] += h;
    } else {
        st[0] = fma_add(st[0], a, one); st[1] = fma_add(st[1], b, one);
        st[2] = fma_add(st[2], c, one); st[3] = fma_add(st[3], d, one);
        st[4] = fma_add(st[4], e, one); st[5] = fma_add(st[5], f, one);
        st[6] = fma_add(st[6], g, one); st[7] = fma_add(st[7], h, one);
    }
}

__host__ __device__ constexpr uint32_t k256(int i) {
    constexpr uint32_t K[64] = {
        0x428a2f98u, 0x71374491u, 0xb5c0fbcfu, 0xe9b5dba5u, 0x3956c25bu, 0x59f111f1u, 0x923f82a4u, 0xab1c5ed5u,
        0xd807aa98u, 0x12835b01u, 0x243185beu, 0x550c7dc3u, 0x72be5d74u, 0x80deb1feu, 0x9bdc06a7u, 0xc19bf174u,
        0xe49b69c1u, 0xefbe4786u, 0x0fc19dc6u, 0x240ca1ccu, 0x2de92c6fu, 0x4a7484aau, 0x5cb0a9dcu, 0x76f988dau,
        0x983e5152u, 0xa831c66du, 0xb00327c8u, 0xbf597fc7u, 0xc6e00bf3u, 0xd5a79147u, 0x06ca6351u, 0x14292967u,
        0x27b70a85u, 0x2e1b2138u, 0x4d2c6dfcu, 0x53380d13u, 0x650a7354u, 0x766a0abbu, 0x81c2c92eu, 0x92722c85u,
        0xa2bfe8a1u, 0xa81a664bu, 0xc24b8b70u, 0xc76c51a3u, 0xd192e819u, 0xd6990624u, 0xf40e3585u, 0x106aa070u,
        0x19a4c116u, 0x1e376c08u, 0x2748774cu, 0x34b0bcb5u, 0x391c0cb3u, 0x4ed8aa4au, 0x5b9cca4fu, 0x682e6ff3u,
        0x748f82eeu, 0x78a5636fu, 0x84c87814u, 0x8cc70208u, 0x90befffau, 0xa4506cebu, 0xbef9a3f7u, 0xc67178f2u};
    return K[i];
}

// M interleaved compressions: round i of every message before round i+1 of any
// (M = 2 measured slower than M = 1 with twice the warps on B200; kept for the record).
// SCHED 0: schedule words 8 at a time ahead of 8 rounds; 1: one word right before its round.
template <int FMA_LEVEL, int M, int SCHED>
__device__ __forceinline__ void sha256_compress_multi(uint32_t (&st)[M][8], uint32_t (&w)[M][16], uint32_t one) {
    uint32_t v[M][8];
#pragma unroll
    for (int m = 0; m < M; ++m)
#pragma unroll
        for (int j = 0; j < 8; ++j) v[m][j] = st[m][j];
#pragma unroll
    for (int i = 0; i < 64; ++i) {
        if (SCHED == 0) {
            if (i >= 16 && (i & 7) == 0) {
#pragma unroll
                for (int j = i; j < i + 8; ++j)
#pragma unroll
                    for (int m = 0; m < M; ++m)
                        w[m][j & 15] = sha_sched<FMA_LEVEL>(w[m][j & 15], w[m][(j + 1) & 15], w[m][(j + 9) & 15],
                                                            w[m][(j + 14) & 15], one);
            }
        } else if (i >= 16) {
#pragma unroll
            for (int m = 0; m < M; ++m)
                w[m][i & 15] = sha_sched<FMA_LEVEL>(w[m][i & 15], w[m][(i + 1) & 15], w[m][(i + 9) & 15],
                                                    w[m][(i + 14) & 15], one);
        }
#pragma unroll
        for (int m = 0; m < M; ++m)
            sha_round<FMA_LEVEL>(v[m][(64 - i) & 7], v[m][(65 - i) & 7], v[m][(66 - i) & 7], v[m][(67 - i) & 7],
                                 v[m][(68 - i) & 7], v[m][(69 - i) & 7], v[m][(70 - i) & 7], v[m][(71 - i) & 7],
                                 k256(i), w[m][i & 15], one);
    }
#pragma unroll
    for (int m = 0; m < M; ++m)
#pragma unroll
        for (int j = 0; j < 8; ++j) st[m][j] = FMA_LEVEL == 0 ? st[m][j] + v[m][j] : fma_add(st[m][j], v[m][j], one);
}

__device__ __forceinline__ void sha256_init(uint32_t (&st)[8]) {
    st[0] = 0x6a09e667u; st[1] = 0xbb67ae85u; st[2] = 0x3c6ef372u; st[3] = 0xa54ff53au;
    st[4] = 0x510e527fu; st[5] = 0x9b05688cu; st[6] = 0x1f83d9abu; st[7] = 0x5be0cd19u;
}

__device__ __forceinline__ uint32_t bswap32(uint32_t x) { return __byte_perm(x, 0, 0x0123); }

}  // namespace cas
