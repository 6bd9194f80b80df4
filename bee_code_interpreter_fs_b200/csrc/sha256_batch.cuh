// Batched SHA-256: one message per lane, many messages per launch.
//
// SHA-256 of one object is a serial Merkle-Damgard chain, so all parallelism
// comes from hashing many objects at once (BASELINE.json north_star).  Two load
// paths feed the same compression code:
//
//   kernel_ldg   each lane reads its own next 64-byte block with four 16-byte
//                read-only loads, one block ahead of the compression.
//   kernel_bulk  each lane asks the TMA unit (cp.async.bulk) to drop its next
//                block into a private, bank-conflict-free slot of shared memory
//                and waits on a per-warp mbarrier; no load instruction and no
//                staging register is spent in the ALU-bound loop.
//
// Objects whose first byte is not 16-byte aligned take a byte-exact unaligned
// path inside the same kernels (aligned word loads merged with PRMT).
#pragma once
#include <stdint.h>

#include "sha256_sm100.cuh"

namespace cas {

struct BatchArgs {
    const uint8_t* base;     // device base address of the object bytes
    const uint64_t* off;     // per-object byte offset from base, or nullptr => uniform
    const uint64_t* len;     // per-object length in bytes (used when off != nullptr)
    const uint32_t* order;   // optional: lane t hashes object order[t] (longest-first schedule)
    uint64_t stride;         // uniform layout: object i starts at base + i*stride
    uint64_t ulen;           // uniform layout: every object is ulen bytes
    uint64_t n;              // number of objects
    uint8_t* out;            // n * 32 bytes of digests, object-major
    uint32_t one;            // run-time 1 for the FMA-pipe adds
    uint32_t zero;           // run-time 0 (makes an opaque data dependency, see kernel_bulk)
};

// Big-endian 32-bit word made of the 4 bytes at p (any alignment); only the
// first `nvalid` (1..4) bytes are read as message bytes, the rest come back 0.
// Touches only aligned words that contain at least one valid byte.
__device__ __forceinline__ uint32_t load_be32_partial(const uint8_t* p, uint32_t nvalid) {
    const uint32_t sh = (uint32_t)((uintptr_t)p & 3u);
    const uint32_t* q = reinterpret_cast<const uint32_t*>(p - sh);
    const uint32_t lo = __ldg(q);
    const uint32_t hi = (sh + nvalid > 4u) ? __ldg(q + 1) : 0u;
    const uint32_t v = __byte_perm(lo, hi, 0x0123u + 0x1111u * sh);
    return v & (0xFFFFFFFFu << (8u * (4u - nvalid)));
}

// 16 big-endian words of the full 64-byte block at p, p not 16-byte aligned.
__device__ __forceinline__ void load_block_unaligned(const uint8_t* p, uint32_t (&w)[16]) {
    const uint32_t sh = (uint32_t)((uintptr_t)p & 3u);
    const uint32_t* q = reinterpret_cast<const uint32_t*>(p - sh);
    const uint32_t sel = 0x0123u + 0x1111u * sh;
    uint32_t x[17];
#pragma unroll
    for (int j = 0; j < 16; ++j) x[j] = __ldg(q + j);
    x[16] = sh ? __ldg(q + 16) : 0u;
#pragma unroll
    for (int j = 0; j < 16; ++j) w[j] = __byte_perm(x[j], x[j + 1], sel);
}

// Words of tail block `t` (0 or 1) of a message of L bytes whose last partial
// block starts at `tailp` and holds `rem` = L % 64 bytes: data, 0x80, zeros and
// the 64-bit big-endian bit length in the last 8 bytes of the last tail block.
__device__ __noinline__ void build_tail_block(const uint8_t* tailp, uint32_t rem, uint64_t L, uint32_t t,
                                              uint32_t ntail, uint32_t* w) {
#pragma unroll 1
    for (uint32_t j = 0; j < 16; ++j) {
        uint32_t v = 0;
        if (t == 0) {
            const uint32_t b0 = 4u * j;
            if (b0 < rem) {
                const uint32_t nv = rem - b0 < 4u ? rem - b0 : 4u;
                v = load_be32_partial(tailp + b0, nv);
            }
            if (rem >= b0 && rem < b0 + 4u) v |= 0x80000000u >> (8u * (rem - b0));
        }
        if (t + 1 == ntail) {
            const uint64_t bits = L << 3;
            if (j == 14) v = (uint32_t)(bits >> 32);
            if (j == 15) v = (uint32_t)bits;
        }
        w[j] = v;
    }
}

__device__ __forceinline__ void store_digest(uint8_t* out, const uint32_t (&st)[8]) {
    uint4 lo, hi;
    lo.x = bswap32(st[0]); lo.y = bswap32(st[1]); lo.z = bswap32(st[2]); lo.w = bswap32(st[3]);
    hi.x = bswap32(st[4]); hi.y = bswap32(st[5]); hi.z = bswap32(st[6]); hi.w = bswap32(st[7]);
    if (((uintptr_t)out & 15u) == 0) {
        reinterpret_cast<uint4*>(out)[0] = lo;
        reinterpret_cast<uint4*>(out)[1] = hi;
    } else {
        const uint32_t v[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
#pragma unroll
        for (int i = 0; i < 32; ++i) out[i] = (uint8_t)(v[i >> 2] >> (8 * (i & 3)));
    }
}

__device__ __forceinline__ bool locate(const BatchArgs& A, uint64_t t, uint64_t& idx, const uint8_t*& p, uint64_t& L) {
    if (t >= A.n) return false;
    idx = A.order ? (uint64_t)A.order[t] : t;
    if (A.off) {
        p = A.base + A.off[idx];
        L = A.len[idx];
    } else {
        p = A.base + idx * A.stride;
        L = A.ulen;
    }
    return true;
}

__device__ __forceinline__ uint4 ldg_nc_v4(const uint8_t* p) {
    uint4 v;
    asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p));
    return v;
}

// ---------------------------------------------------------------- kernel_ldg --
template <int FMA_LEVEL, int THREADS, int MIN_CTAS>
__global__ void __launch_bounds__(THREADS, MIN_CTAS) kernel_ldg(const BatchArgs A) {
    const uint64_t t = (uint64_t)blockIdx.x * THREADS + threadIdx.x;
    uint64_t idx, L;
    const uint8_t* p;
    if (!locate(A, t, idx, p, L)) return;
    const uint32_t one = A.one;
    const uint64_t nfull = L >> 6;
    const uint32_t rem = (uint32_t)(L & 63u);
    const uint32_t ntail = rem + 9u <= 64u ? 1u : 2u;
    const uint64_t nb = nfull + ntail;
    const bool al = ((uintptr_t)p & 15u) == 0;

    uint32_t st[8];
    sha256_init(st);
    uint4 q0, q1, q2, q3;
    q0 = q1 = q2 = q3 = make_uint4(0, 0, 0, 0);
    if (al && nfull) {
        q0 = ldg_nc_v4(p); q1 = ldg_nc_v4(p + 16); q2 = ldg_nc_v4(p + 32); q3 = ldg_nc_v4(p + 48);
    }
#pragma unroll 1
    for (uint64_t i = 0; i < nb; ++i) {
        uint32_t w[16];
        if (i < nfull) {
            if (al) {
                w[0] = bswap32(q0.x); w[1] = bswap32(q0.y); w[2] = bswap32(q0.z); w[3] = bswap32(q0.w);
                w[4] = bswap32(q1.x); w[5] = bswap32(q1.y); w[6] = bswap32(q1.z); w[7] = bswap32(q1.w);
                w[8] = bswap32(q2.x); w[9] = bswap32(q2.y); w[10] = bswap32(q2.z); w[11] = bswap32(q2.w);
                w[12] = bswap32(q3.x); w[13] = bswap32(q3.y); w[14] = bswap32(q3.z); w[15] = bswap32(q3.w);
                if (i + 1 < nfull) {
                    const uint8_t* pn = p + ((i + 1) << 6);
                    q0 = ldg_nc_v4(pn); q1 = ldg_nc_v4(pn + 16); q2 = ldg_nc_v4(pn + 32); q3 = ldg_nc_v4(pn + 48);
                }
            } else {
                load_block_unaligned(p + (i << 6), w);
            }
        } else {
            // Cold path: the builder fills a stack array so that w[] itself never
            // has its address taken and stays in registers in the hot loop.
            uint32_t tw[16];
            build_tail_block(p + (nfull << 6), rem, L, (uint32_t)(i - nfull), ntail, tw);
#pragma unroll
            for (int j = 0; j < 16; ++j) w[j] = tw[j];
        }
        sha256_compress<FMA_LEVEL>(st, w, one);
    }
    store_digest(A.out + idx * 32u, st);
}

// -------------------------------------------------------------- kernel_resume --
// Streaming form: lane t continues the chain of stream t from the 8-word state it is given
// over `nblk[t]` whole 64-byte blocks at base + off[t] (16-byte aligned) and writes the state
// back.  No padding here: the caller appends the FIPS 180-4 tail as ordinary blocks when the
// stream ends (cas_stream_finish).  States are host-order 32-bit words, 8 per stream.
template <int FMA_LEVEL, int THREADS>
__global__ void __launch_bounds__(THREADS) kernel_resume(const uint8_t* __restrict__ base, const uint64_t* __restrict__ off,
                                                         const uint64_t* __restrict__ nblk, uint32_t* __restrict__ state,
                                                         uint64_t n, uint32_t one) {
    const uint64_t t = (uint64_t)blockIdx.x * THREADS + threadIdx.x;
    if (t >= n) return;
    const uint64_t nb = nblk[t];
    if (nb == 0) return;
    const uint8_t* p = base + off[t];
    uint32_t st[8];
    const uint4 s0 = reinterpret_cast<const uint4*>(state + 8 * t)[0], s1 = reinterpret_cast<const uint4*>(state + 8 * t)[1];
    st[0] = s0.x; st[1] = s0.y; st[2] = s0.z; st[3] = s0.w; st[4] = s1.x; st[5] = s1.y; st[6] = s1.z; st[7] = s1.w;
    uint4 q0 = ldg_nc_v4(p), q1 = ldg_nc_v4(p + 16), q2 = ldg_nc_v4(p + 32), q3 = ldg_nc_v4(p + 48);
#pragma unroll 1
    for (uint64_t i = 0; i < nb; ++i) {
        uint32_t w[16];
        w[0] = bswap32(q0.x); w[1] = bswap32(q0.y); w[2] = bswap32(q0.z); w[3] = bswap32(q0.w);
        w[4] = bswap32(q1.x); w[5] = bswap32(q1.y); w[6] = bswap32(q1.z); w[7] = bswap32(q1.w);
        w[8] = bswap32(q2.x); w[9] = bswap32(q2.y); w[10] = bswap32(q2.z); w[11] = bswap32(q2.w);
        w[12] = bswap32(q3.x); w[13] = bswap32(q3.y); w[14] = bswap32(q3.z); w[15] = bswap32(q3.w);
        if (i + 1 < nb) {
            const uint8_t* pn = p + ((i + 1) << 6);
            q0 = ldg_nc_v4(pn); q1 = ldg_nc_v4(pn + 16); q2 = ldg_nc_v4(pn + 32); q3 = ldg_nc_v4(pn + 48);
        }
        sha256_compress<FMA_LEVEL>(st, w, one);
    }
    reinterpret_cast<uint4*>(state + 8 * t)[0] = make_uint4(st[0], st[1], st[2], st[3]);
    reinterpret_cast<uint4*>(state + 8 * t)[1] = make_uint4(st[4], st[5], st[6], st[7]);
}

// --------------------------------------------------------------- kernel_bulk --
// Shared-memory plan per warp: NSTAGE stages, each 32 slots of SLOT_PITCH bytes
// (64 data bytes + 16 pad so that the 8 lanes of an LDS.128 phase hit 8
// different 16-byte bank groups), plus one mbarrier per stage (32 arrivals).
constexpr int SLOT_PITCH = 80;

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(bar),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const uint8_t* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ uint4 lds_v4(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}

template <int NSTAGE, int THREADS>
constexpr int bulk_smem_bytes() {
    return (THREADS / 32) * NSTAGE * (32 * SLOT_PITCH + 8);
}

template <int FMA_LEVEL, int NSTAGE, int THREADS, int MIN_CTAS>
__global__ void __launch_bounds__(THREADS, MIN_CTAS) kernel_bulk(const BatchArgs A) {
    extern __shared__ __align__(128) uint8_t smem[];
    constexpr int WARPS = THREADS / 32;
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t smem_base = (uint32_t)__cvta_generic_to_shared(smem);
    const uint32_t slots = smem_base + (warp * NSTAGE) * (32 * SLOT_PITCH) + lane * SLOT_PITCH;
    const uint32_t bars = smem_base + WARPS * NSTAGE * (32 * SLOT_PITCH) + warp * NSTAGE * 8;

    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < NSTAGE; ++s) mbar_init(bars + 8 * s, 32);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncwarp();

    const uint64_t t = (uint64_t)blockIdx.x * THREADS + threadIdx.x;
    uint64_t idx = 0, L = 0;
    const uint8_t* p = nullptr;
    const bool live = locate(A, t, idx, p, L);
    const uint32_t one = A.one;
    const uint64_t nfull = live ? (L >> 6) : 0;
    const uint32_t rem = (uint32_t)(L & 63u);
    const uint32_t ntail = rem + 9u <= 64u ? 1u : 2u;
    const uint64_t nb = live ? nfull + ntail : 0;
    const bool al = ((uintptr_t)p & 15u) == 0;
    const uint64_t nbulk = al ? nfull : 0;  // blocks this lane fetches through the TMA unit

    // Warp-uniform trip count: every lane keeps the mbarrier protocol going
    // until the longest message of the warp is done.
    uint32_t hi32 = (uint32_t)(nb >> 32), lo32 = (uint32_t)nb;
    const uint32_t mhi = __reduce_max_sync(0xffffffffu, hi32);
    lo32 = (hi32 == mhi) ? lo32 : 0u;
    const uint64_t trips = ((uint64_t)mhi << 32) | __reduce_max_sync(0xffffffffu, lo32);

    uint32_t st[8];
    sha256_init(st);

#pragma unroll
    for (int s = 0; s < NSTAGE; ++s) {
        if ((uint64_t)s < nbulk) {
            mbar_arrive_tx(bars + 8 * s, 64);
            bulk_g2s(slots + s * (32 * SLOT_PITCH), p + ((uint64_t)s << 6), 64, bars + 8 * s);
        } else {
            mbar_arrive(bars + 8 * s);
        }
    }

    uint32_t stage = 0, parity = 0;
#pragma unroll 1
    for (uint64_t i = 0; i < trips; ++i) {
        const uint32_t bar = bars + 8 * stage;
        const uint32_t slot = slots + stage * (32 * SLOT_PITCH);
        mbar_wait(bar, parity);
        if (i < nb) {
            uint32_t w[16];
            if (i < nfull) {
                if (al) {
                    const uint4 q0 = lds_v4(slot), q1 = lds_v4(slot + 16), q2 = lds_v4(slot + 32), q3 = lds_v4(slot + 48);
                    w[0] = bswap32(q0.x); w[1] = bswap32(q0.y); w[2] = bswap32(q0.z); w[3] = bswap32(q0.w);
                    w[4] = bswap32(q1.x); w[5] = bswap32(q1.y); w[6] = bswap32(q1.z); w[7] = bswap32(q1.w);
                    w[8] = bswap32(q2.x); w[9] = bswap32(q2.y); w[10] = bswap32(q2.z); w[11] = bswap32(q2.w);
                    w[12] = bswap32(q3.x); w[13] = bswap32(q3.y); w[14] = bswap32(q3.z); w[15] = bswap32(q3.w);
                } else {
                    load_block_unaligned(p + (i << 6), w);
                }
            } else {
                uint32_t tw[16];
                build_tail_block(p + (nfull << 6), rem, L, (uint32_t)(i - nfull), ntail, tw);
#pragma unroll
                for (int j = 0; j < 16; ++j) w[j] = tw[j];
            }
            sha256_compress<FMA_LEVEL>(st, w, one);
            // Hand the slot back to the TMA unit.  The byte count is made to depend on
            // the compression result (st[0] & 0 | 64, one LOP3), so neither the compiler
            // nor the in-order issue logic can start the copy before this lane's
            // shared-memory reads of the slot have returned.
            if (i + NSTAGE < nbulk) {
                uint32_t bytes;
                asm volatile("lop3.b32 %0, %1, %2, 64, 0xEA;" : "=r"(bytes) : "r"(st[0]), "r"(A.zero));
                mbar_arrive_tx(bar, bytes);
                bulk_g2s(slot, p + ((i + NSTAGE) << 6), bytes, bar);
            } else {
                mbar_arrive(bar);
            }
        } else {
            mbar_arrive(bar);
        }
        if (++stage == NSTAGE) { stage = 0; parity ^= 1u; }
    }
    if (live) store_digest(A.out + idx * 32u, st);
}

// ---------------------------------------------------------------- kernel_tma --
// Uniform layouts (object i at base + i*stride, all `ulen` bytes long) are a 2-D
// byte tensor [n objects][ulen bytes].  One TMA tile load per warp per block
// index brings the next 64-byte block of 32 consecutive objects (a 64 B x 32 row
// box, 2 KiB) into shared memory: one instruction from one elected lane instead
// of 128 per-lane loads.  The tensor map uses the 64-byte swizzle, so the four
// 16-byte pieces a lane reads with LDS.128 are spread over all bank groups.
// Bytes beyond `ulen` (the partial last block) and rows beyond `n` are
// zero-filled by the TMA unit, which is exactly the zero padding SHA-256 wants:
// the lane only ORs in the 0x80 marker and the bit length.
constexpr int TMA_STAGE_BYTES = 32 * 64;

__device__ __forceinline__ void tma_tile_2d(uint32_t dst, const void* tmap, int32_t c0, int32_t c1, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst),
        "l"(tmap), "r"(c0), "r"(c1), "r"(bar)
        : "memory");
}

struct alignas(64) TensorMapBlob {
    uint64_t opaque[16];  // a CUtensorMap (128 bytes), passed by value as __grid_constant__
};

template <int NSTAGE, int THREADS, int MSGS>
constexpr int tma_smem_bytes() {
    return 1024 + (THREADS / 32) * NSTAGE * (TMA_STAGE_BYTES * MSGS) + (THREADS / 32) * NSTAGE * 8;
}

template <int FMA_LEVEL, int NSTAGE, int THREADS, int MSGS, int SCHED, int PADV>
__device__ __forceinline__ void tma_body(const TensorMapBlob& tmap, const BatchArgs& A) {
    extern __shared__ uint8_t smem_raw[];
    constexpr int WARPS = THREADS / 32;
    constexpr uint32_t STAGE = TMA_STAGE_BYTES * MSGS;  // the tile is 32*MSGS rows of 64 bytes
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t smem_base = ((uint32_t)__cvta_generic_to_shared(smem_raw) + 1023u) & ~1023u;
    const uint32_t tiles = smem_base + warp * NSTAGE * STAGE;
    const uint32_t bars = smem_base + WARPS * NSTAGE * STAGE + warp * NSTAGE * 8;
    const uint32_t row = lane * 64u + (((lane >> 1) & 3u) << 4);  // message m of this lane: row + m*2048

    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < NSTAGE; ++s) mbar_init(bars + 8 * s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncwarp();

    const uint64_t first = ((uint64_t)blockIdx.x * WARPS + warp) * (32u * MSGS);
    const uint32_t one = A.one;
    const uint64_t L = A.ulen;
    const uint32_t nfull = (uint32_t)(L >> 6);
    const uint32_t rem = (uint32_t)(L & 63u);
    const uint32_t ntail = rem + 9u <= 64u ? 1u : 2u;
    const uint32_t nb = nfull + ntail;
    const uint32_t bits_hi = (uint32_t)(L >> 29), bits_lo = (uint32_t)(L << 3);

    uint32_t st[MSGS][8];
#pragma unroll
    for (int m = 0; m < MSGS; ++m) sha256_init(st[m]);

    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < NSTAGE; ++s)
            if ((uint32_t)s < nb) {
                mbar_arrive_tx(bars + 8 * s, STAGE);
                tma_tile_2d(tiles + s * STAGE, &tmap, (int32_t)(s * 64), (int32_t)first, bars + 8 * s);
            }
    }

    uint32_t stage = 0, parity = 0;
#pragma unroll 1
    for (uint32_t i = 0; i < nb; ++i) {
        uint32_t w[MSGS][16];
        const uint32_t tile = tiles + stage * STAGE + row;
        mbar_wait(bars + 8 * stage, parity);
#pragma unroll
        for (int m = 0; m < MSGS; ++m) {
            const uint32_t t = tile + m * TMA_STAGE_BYTES;
            const uint4 q0 = lds_v4(t), q1 = lds_v4(t ^ 16u), q2 = lds_v4(t ^ 32u), q3 = lds_v4(t ^ 48u);
            w[m][0] = bswap32(q0.x); w[m][1] = bswap32(q0.y); w[m][2] = bswap32(q0.z); w[m][3] = bswap32(q0.w);
            w[m][4] = bswap32(q1.x); w[m][5] = bswap32(q1.y); w[m][6] = bswap32(q1.z); w[m][7] = bswap32(q1.w);
            w[m][8] = bswap32(q2.x); w[m][9] = bswap32(q2.y); w[m][10] = bswap32(q2.z); w[m][11] = bswap32(q2.w);
            w[m][12] = bswap32(q3.x); w[m][13] = bswap32(q3.y); w[m][14] = bswap32(q3.z); w[m][15] = bswap32(q3.w);
        }
        if (i >= nfull) {  // tail blocks (kernel-uniform branch, once or twice per message)
            if (i == nfull) {
                uint32_t r = rem;
                if (PADV == 1) asm volatile("" : "+r"(r));  // keeps the 16 masks out of loop-invariant registers
                const uint32_t pad_word = r >> 2, pad_mask = 0x80000000u >> (8u * (r & 3u));
#pragma unroll
                for (int m = 0; m < MSGS; ++m)
#pragma unroll
                    for (int j = 0; j < 16; ++j) w[m][j] |= ((uint32_t)j == pad_word) ? pad_mask : 0u;
            }
            if (i + 1u == nb) {
#pragma unroll
                for (int m = 0; m < MSGS; ++m) { w[m][14] = bits_hi; w[m][15] = bits_lo; }
            }
        }
        sha256_compress_multi<FMA_LEVEL, MSGS, SCHED>(st, w, one);
        // Every lane's LDS reads of this stage must have completed before the async proxy may
        // overwrite it: the warp reconverges here explicitly (independent thread scheduling
        // gives no such guarantee after the divergent try_wait spin above).
        __syncwarp();
        if (lane == 0 && i + NSTAGE < nb) {
            uint32_t bytes;
            asm volatile("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(bytes) : "r"(st[0][0] | st[MSGS - 1][0]), "r"(A.zero), "r"(STAGE));
            mbar_arrive_tx(bars + 8 * stage, bytes);
            tma_tile_2d(tiles + stage * STAGE, &tmap, (int32_t)((i + NSTAGE) * 64u), (int32_t)first, bars + 8 * stage);
        }
        if (++stage == NSTAGE) { stage = 0; parity ^= 1u; }
    }
#pragma unroll
    for (int m = 0; m < MSGS; ++m) {
        const uint64_t t = first + lane + 32u * m;
        if (t < A.n) store_digest(A.out + t * 32u, st[m]);
    }
}


// Register budget either from __launch_bounds__(THREADS, MIN_CTAS) or, finer, from
// __maxnreg__(REGCAP).  ptxas' instruction schedule (and with it 5-8 % of the
// throughput) changes with the budget; tools/kbench_build.py sweeps both.
template <int FMA_LEVEL, int NSTAGE, int THREADS, int MIN_CTAS, int MSGS = 1, int SCHED = 0, int PADV = 1>
__global__ void __launch_bounds__(THREADS, MIN_CTAS)
    kernel_tma(const __grid_constant__ TensorMapBlob tmap, const BatchArgs A) {
    tma_body<FMA_LEVEL, NSTAGE, THREADS, MSGS, SCHED, PADV>(tmap, A);
}
template <int FMA_LEVEL, int NSTAGE, int THREADS, int REGCAP, int MSGS = 1, int SCHED = 0, int PADV = 1>
__global__ void __launch_bounds__(THREADS) __maxnreg__(REGCAP)
    kernel_tma_regcap(const __grid_constant__ TensorMapBlob tmap, const BatchArgs A) {
    tma_body<FMA_LEVEL, NSTAGE, THREADS, MSGS, SCHED, PADV>(tmap, A);
}

}  // namespace cas
