// Long messages, few of them: one WARP per M messages (M = 1, 2, 4, 8, 16, 32).
//
// With lane-per-message a batch of a few large objects (BASELINE config 5, the big
// files of config 4, a single upload) leaves the GPU almost empty and every chain
// runs at the speed of one lane.  SHA-256 cannot be split inside a message, but the
// message schedule can: W[16..63] of a block depends only on that block's bytes.
// A warp therefore serves M messages with G = 32 / M lanes each and alternates between
//   phase A  lane l of a message's group expands the schedule of that message's block
//            (G g + l) and stores W[t] + K[t] (64 words) to shared memory: 32 blocks in
//            parallel, coalesced loads;
//   phase B  the 64 rounds of each message's G blocks, one block after the other, the
//            G lanes of a group all running their message's chain from broadcast
//            LDS.128 reads: only 10 logic instructions + the adds per round stay on
//            the serial path instead of ~27.
// One warp instruction of phase B advances M chains at once, so M trades per-chain
// speed (phase A is amortised over 32 / M blocks per chain) against chains per
// sub-partition.  Measured on a B200 (profiles/r02_kbench_long.jsonl, 1 MiB objects, one
// warp per sub-partition): M = 1 70-74 MB/s per chain, M = 2 69, M = 4 66, M = 8 61,
// M = 16 52, M = 32 41, against 31 MB/s for one object per lane (kernel_ldg) -- so the warp
// form wins up to ~38 000 long objects and the engine picks M from the object count.
// The digest is unchanged.
//
// The serial path is bounded by one warp's ALU-pipe occupancy: ptxas' schedule of the 64
// rounds is 11 ALU-pipe instructions per round (6 SHF + 4 LOP3 + 1 IADD3; the other adds
// run as IMAD on the FMA pipe), issued every second clock.  A warp instruction takes the
// pipe for two clocks whatever its active mask (phase B on 16, 8 or 1 lanes: same time),
// and rotations moved to the FMA pipe (IMAD + IMAD.HI) are slower than the SHF they save.
#pragma once
#include "sha256_batch.cuh"

namespace cas {

constexpr int LONG_GROUP = 32;             // blocks expanded per phase A (one per lane)
constexpr int LONG_PITCH = 64 * 4 + 16;    // bytes per expanded block; +16 keeps STS.128 / LDS.128 conflict-free
constexpr int LONG_SMEM_PER_WARP = LONG_GROUP * LONG_PITCH;

__device__ __forceinline__ void sts_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// rotr(x, N) on the FMA pipe: hi32(x * 2^(32-N)) + (x << (32-N)); the two halves have no
// bit in common, so the sum is the rotation.  IMAD + IMAD.HI (the latter quarter rate).
template <int N>
__device__ __forceinline__ uint32_t rotr_fma(uint32_t x) {
    uint32_t lo, d;
    asm("mul.lo.u32 %0, %1, %2;" : "=r"(lo) : "r"(x), "n"(1u << (32 - N)));
    asm("mad.hi.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(x), "n"(1u << (32 - N)), "r"(lo));
    return d;
}

// One round given wk = W[t] + K[t].  ADDS 0: plain adds (ptxas picks IADD3 / IMAD.IADD);
// 1: shortest dependency chain -- both new values are one 3-input add after Sigma;
// 2: every add a 2-input add on the FMA pipe; 3: the same pipe, but with the data flow
// rearranged so that both new values are one or two FMA-pipe adds after their Sigma:
// a' = e' - d + Sigma0 + Maj (minus_one * d on the FMA pipe); 4: as 3 with the last add as IADD3.
// ROTF: how many of Sigma0's rotations are computed on the FMA pipe (0..2).
template <int ADDS, int ROTF = 0>
__device__ __forceinline__ void long_round(uint32_t a, uint32_t b, uint32_t c, uint32_t& d, uint32_t e, uint32_t f,
                                           uint32_t g, uint32_t& h, uint32_t wk, uint32_t one) {
    const uint32_t s1 = big_sigma1(e);
    const uint32_t ch = lop_ch(e, f, g);
    uint32_t s0;
    if (ROTF == 0) s0 = big_sigma0(a);
    else if (ROTF == 1) s0 = lop_xor3(rotr32(a, 2), rotr32(a, 13), rotr_fma<22>(a));
    else s0 = lop_xor3(rotr32(a, 2), rotr_fma<13>(a), rotr_fma<22>(a));
    const uint32_t mj = lop_maj(a, b, c);
    if (ADDS == 0) {
        const uint32_t t1 = h + wk + s1 + ch;
        d += t1;
        h = t1 + s0 + mj;
    } else if (ADDS == 1) {
        const uint32_t x = fma_add(h, wk, one);   // off the critical path: h and d are three rounds old
        const uint32_t x2 = fma_add(x, d, one);
        const uint32_t p = x + ch + mj;
        d = x2 + ch + s1;
        h = p + s1 + s0;
    } else if (ADDS == 2) {
        const uint32_t x = fma_add(h, wk, one);
        const uint32_t t1 = fma_add(fma_add(x, ch, one), s1, one);
        d = fma_add(d, t1, one);
        h = fma_add(fma_add(t1, mj, one), s0, one);
    } else {
        const uint32_t x2 = fma_add(fma_add(h, wk, one), d, one);  // h + wk + d: all three rounds old
        const uint32_t m2 = fma_add(d, mj, 0u - one);               // Maj - d   (d * -1 + Maj)
        const uint32_t dn = fma_add(fma_add(x2, ch, one), s1, one);
        if (ADDS == 3) {
            h = fma_add(fma_add(m2, s0, one), dn, one);
        } else if (ADDS == 4) {
            h = dn + m2 + s0;
        } else {  // 5: e' by one IADD3 (shortest e recurrence), a' by one IADD3 off Maj - d
            const uint32_t dn2 = x2 + ch + s1;
            h = dn2 + m2 + s0;
            d = dn2;
            return;
        }
        d = dn;
    }
}

// M messages per warp, G = 32 / M lanes per message.  ACT < 32 (experiment, M = 1 only): only
// the first ACT lanes run phase B.
template <int FMA_LEVEL, int ADDS, int M = 1, int ACT = 32, int ROTF = 0>
__global__ void __launch_bounds__(32) kernel_long(const BatchArgs A) {
    static_assert(M == 1 || M == 2 || M == 4 || M == 8 || M == 16 || M == 32, "M must be a power of two up to 32");
    static_assert(ACT == 32 || M == 1, "ACT is an experiment knob of the one-message form");
    constexpr int G = 32 / M;
    extern __shared__ __align__(16) uint8_t smem_long[];
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t grp = lane / G, sub = lane % G;            // message of this lane, lane's place in its group
    const uint64_t t = (uint64_t)blockIdx.x * M + grp;        // M consecutive (schedule-order) messages per warp
    uint64_t idx = 0, L = 0;
    const uint8_t* p = nullptr;
    const bool live = locate(A, t, idx, p, L);
    const uint32_t one = A.one;
    const uint64_t nfull = L >> 6;
    const uint32_t rem = (uint32_t)(L & 63u);
    const uint32_t ntail = rem + 9u <= 64u ? 1u : 2u;
    const uint64_t nb = live ? nfull + ntail : 0;
    const bool al = ((uintptr_t)p & 15u) == 0;
    const uint32_t rows = (uint32_t)__cvta_generic_to_shared(smem_long);
    const uint32_t my_row = rows + lane * LONG_PITCH;          // lane (grp, sub) expands into row grp*G + sub = lane
    const uint32_t grp_rows = rows + grp * G * LONG_PITCH;

    // Warp-uniform trip count: the longest message of the warp decides.
    uint64_t nb_max = nb;
    if (M > 1) {
        uint32_t hi32 = (uint32_t)(nb >> 32), lo32 = (uint32_t)nb;
        const uint32_t mhi = __reduce_max_sync(0xffffffffu, hi32);
        lo32 = (hi32 == mhi) ? lo32 : 0u;
        nb_max = ((uint64_t)mhi << 32) | __reduce_max_sync(0xffffffffu, lo32);
    }

    uint32_t st[8];
    sha256_init(st);

    uint4 q0, q1, q2, q3;
    q0 = q1 = q2 = q3 = make_uint4(0, 0, 0, 0);
    if (al && sub < nfull) {
        const uint8_t* s = p + ((uint64_t)sub << 6);
        q0 = ldg_nc_v4(s); q1 = ldg_nc_v4(s + 16); q2 = ldg_nc_v4(s + 32); q3 = ldg_nc_v4(s + 48);
    }
#pragma unroll 1
    for (uint64_t g0 = 0; g0 < nb_max; g0 += G) {
        // ---- phase A: this lane expands block j of its message ----------------------------
        const uint64_t j = g0 + sub;
        if (j < nb) {
            uint32_t w[16];
            if (j < nfull) {
                if (al) {
                    w[0] = bswap32(q0.x); w[1] = bswap32(q0.y); w[2] = bswap32(q0.z); w[3] = bswap32(q0.w);
                    w[4] = bswap32(q1.x); w[5] = bswap32(q1.y); w[6] = bswap32(q1.z); w[7] = bswap32(q1.w);
                    w[8] = bswap32(q2.x); w[9] = bswap32(q2.y); w[10] = bswap32(q2.z); w[11] = bswap32(q2.w);
                    w[12] = bswap32(q3.x); w[13] = bswap32(q3.y); w[14] = bswap32(q3.z); w[15] = bswap32(q3.w);
                } else {
                    load_block_unaligned(p + (j << 6), w);
                }
            } else {
                uint32_t tw[16];
                build_tail_block(p + (nfull << 6), rem, L, (uint32_t)(j - nfull), ntail, tw);
#pragma unroll
                for (int i = 0; i < 16; ++i) w[i] = tw[i];
            }
#pragma unroll
            for (int i = 0; i < 64; i += 4) {
                uint32_t o[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int r = i + u;
                    if (r >= 16)
                        w[r & 15] = sha_sched<FMA_LEVEL>(w[r & 15], w[(r + 1) & 15], w[(r + 9) & 15], w[(r + 14) & 15], one);
                    o[u] = w[r & 15] + k256(r);
                }
                sts_v4(my_row + i * 4, o[0], o[1], o[2], o[3]);
            }
        }
        // The next group's bytes travel while this group's rounds run.
        {
            const uint64_t jn = j + G;
            if (al && jn < nfull) {
                const uint8_t* s = p + (jn << 6);
                q0 = ldg_nc_v4(s); q1 = ldg_nc_v4(s + 16); q2 = ldg_nc_v4(s + 32); q3 = ldg_nc_v4(s + 48);
            }
        }
        __syncwarp();
        // ---- phase B: the serial chains over the expanded blocks ---------------------------
        const uint64_t left_max = nb_max - g0;
        const uint32_t cnt_max = (uint32_t)(left_max < (uint64_t)G ? left_max : (uint64_t)G);
        const uint32_t cnt = nb > g0 ? (uint32_t)(nb - g0 < (uint64_t)G ? nb - g0 : (uint64_t)G) : 0u;
        if (ACT == 32 || lane < ACT) {
#pragma unroll 1
            for (uint32_t b = 0; b < cnt_max; ++b) {
                if (M > 1 && b >= cnt) continue;  // this message is done; others in the warp are not
                const uint32_t row = grp_rows + b * LONG_PITCH;
                uint32_t v[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) v[i] = st[i];
#pragma unroll
                for (int i = 0; i < 64; i += 4) {
                    const uint4 k4 = lds_v4(row + i * 4);
                    const uint32_t wk[4] = {k4.x, k4.y, k4.z, k4.w};
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const int r = i + u;
                        long_round<ADDS, ROTF>(v[(64 - r) & 7], v[(65 - r) & 7], v[(66 - r) & 7], v[(67 - r) & 7],
                                               v[(68 - r) & 7], v[(69 - r) & 7], v[(70 - r) & 7], v[(71 - r) & 7], wk[u], one);
                    }
                }
#pragma unroll
                for (int i = 0; i < 8; ++i) st[i] += v[i];
            }
        }
        __syncwarp();
    }
    if (ACT != 32) {
#pragma unroll
        for (int i = 0; i < 8; ++i) st[i] = __shfl_sync(0xffffffffu, st[i], 0);
    }
    if (live && sub == 0) store_digest(A.out + idx * 32u, st);
}

}  // namespace cas
