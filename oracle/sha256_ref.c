/*
 * oracle/sha256_ref.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain scalar C restatement of SHA-256 (FIPS 180-4, sections 4.1.2, 4.2.2,
 * 5.1.1, 5.3.3, 6.2) and of the synthetic object generator this repository
 * benchmarks with.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load it, and only as the checker.
 *
 * What it restates.  The reference's object store documents its ids as
 * "identified by their SHA-256 hash"
 * (/root/reference/src/code_interpreter/services/storage.py:36) but at the
 * surveyed commit assigns them with secrets.token_hex(32) (storage.py:52); no
 * reference source file computes a digest.  The arithmetic therefore lives in
 * a dependency that is not under /root/reference: CPython's hashlib.sha256
 * (the stdlib `_hashlib` module over OpenSSL 3.0.13 in this image), which
 * BASELINE.json names as the oracle.  The Rust executor pins `sha2 = 0.10.8`
 * (executor/Cargo.lock:1158-1161) but never calls it.
 *
 * Parity pin.  "parity unpinned BY THE REFERENCE": no reference test holds a
 * digest constant (test/e2e/test_http.py:47-85 only round-trips an opaque id).
 * This restatement is pinned instead to the FIPS 180-4 / NIST known answers
 * and to hashlib itself in tests/test_oracle.py and tests/golden/.
 */
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* FIPS 180-4 section 4.2.2: first 32 bits of the fractional parts of the cube
 * roots of the first 64 primes. */
static const uint32_t K256[64] = {
    0x428a2f98u, 0x71374491u, 0xb5c0fbcfu, 0xe9b5dba5u, 0x3956c25bu, 0x59f111f1u,
    0x923f82a4u, 0xab1c5ed5u, 0xd807aa98u, 0x12835b01u, 0x243185beu, 0x550c7dc3u,
    0x72be5d74u, 0x80deb1feu, 0x9bdc06a7u, 0xc19bf174u, 0xe49b69c1u, 0xefbe4786u,
    0x0fc19dc6u, 0x240ca1ccu, 0x2de92c6fu, 0x4a7484aau, 0x5cb0a9dcu, 0x76f988dau,
    0x983e5152u, 0xa831c66du, 0xb00327c8u, 0xbf597fc7u, 0xc6e00bf3u, 0xd5a79147u,
    0x06ca6351u, 0x14292967u, 0x27b70a85u, 0x2e1b2138u, 0x4d2c6dfcu, 0x53380d13u,
    0x650a7354u, 0x766a0abbu, 0x81c2c92eu, 0x92722c85u, 0xa2bfe8a1u, 0xa81a664bu,
    0xc24b8b70u, 0xc76c51a3u, 0xd192e819u, 0xd6990624u, 0xf40e3585u, 0x106aa070u,
    0x19a4c116u, 0x1e376c08u, 0x2748774cu, 0x34b0bcb5u, 0x391c0cb3u, 0x4ed8aa4au,
    0x5b9cca4fu, 0x682e6ff3u, 0x748f82eeu, 0x78a5636fu, 0x84c87814u, 0x8cc70208u,
    0x90befffau, 0xa4506cebu, 0xbef9a3f7u, 0xc67178f2u};

/* FIPS 180-4 section 5.3.3: initial hash value. */
static const uint32_t H256[8] = {0x6a09e667u, 0xbb67ae85u, 0x3c6ef372u, 0xa54ff53au,
                                 0x510e527fu, 0x9b05688cu, 0x1f83d9abu, 0x5be0cd19u};

static inline uint32_t ror(uint32_t x, unsigned n) { return (x >> n) | (x << (32u - n)); }

/* One compression (FIPS 180-4 section 6.2.2) over a 64-byte block. */
static void compress(uint32_t st[8], const uint8_t blk[64]) {
    uint32_t w[64];
    for (int t = 0; t < 16; ++t)
        w[t] = ((uint32_t)blk[4 * t] << 24) | ((uint32_t)blk[4 * t + 1] << 16) |
               ((uint32_t)blk[4 * t + 2] << 8) | (uint32_t)blk[4 * t + 3];
    for (int t = 16; t < 64; ++t) {
        uint32_t s0 = ror(w[t - 15], 7) ^ ror(w[t - 15], 18) ^ (w[t - 15] >> 3);
        uint32_t s1 = ror(w[t - 2], 17) ^ ror(w[t - 2], 19) ^ (w[t - 2] >> 10);
        w[t] = w[t - 16] + s0 + w[t - 7] + s1;
    }
    uint32_t a = st[0], b = st[1], c = st[2], d = st[3];
    uint32_t e = st[4], f = st[5], g = st[6], h = st[7];
    for (int t = 0; t < 64; ++t) {
        uint32_t S1 = ror(e, 6) ^ ror(e, 11) ^ ror(e, 25);
        uint32_t ch = (e & f) ^ (~e & g);
        uint32_t t1 = h + S1 + ch + K256[t] + w[t];
        uint32_t S0 = ror(a, 2) ^ ror(a, 13) ^ ror(a, 22);
        uint32_t mj = (a & b) ^ (a & c) ^ (b & c);
        uint32_t t2 = S0 + mj;
        h = g; g = f; f = e; e = d + t1;
        d = c; c = b; b = a; a = t1 + t2;
    }
    st[0] += a; st[1] += b; st[2] += c; st[3] += d;
    st[4] += e; st[5] += f; st[6] += g; st[7] += h;
}

/* Whole-message digest: padding per FIPS 180-4 section 5.1.1 (0x80, zeros,
 * 64-bit big-endian bit length). */
void oracle_sha256(const uint8_t* data, uint64_t len, uint8_t out[32]) {
    uint32_t st[8];
    memcpy(st, H256, sizeof st);
    uint64_t full = len / 64;
    for (uint64_t i = 0; i < full; ++i) compress(st, data + 64 * i);
    uint8_t tail[128];
    uint64_t rem = len - 64 * full;
    memset(tail, 0, sizeof tail);
    if (rem) memcpy(tail, data + 64 * full, rem);
    tail[rem] = 0x80;
    unsigned tl = (rem + 9 <= 64) ? 64 : 128;
    uint64_t bits = len * 8u;
    for (int i = 0; i < 8; ++i) tail[tl - 1 - i] = (uint8_t)(bits >> (8 * i));
    compress(st, tail);
    if (tl == 128) compress(st, tail + 64);
    for (int i = 0; i < 8; ++i) {
        out[4 * i] = (uint8_t)(st[i] >> 24);
        out[4 * i + 1] = (uint8_t)(st[i] >> 16);
        out[4 * i + 2] = (uint8_t)(st[i] >> 8);
        out[4 * i + 3] = (uint8_t)st[i];
    }
}

/* Number of compressions a message of `len` bytes costs: ceil((len+9)/64). */
uint64_t oracle_sha256_blocks(uint64_t len) { return (len + 9 + 63) / 64; }

/* Batch form used by the tests: object i is base[off[i] .. off[i]+len[i]). */
void oracle_sha256_batch(const uint8_t* base, const uint64_t* off, const uint64_t* len,
                         uint64_t n, uint8_t* out32n) {
    for (uint64_t i = 0; i < n; ++i) oracle_sha256(base + off[i], len[i], out32n + 32 * i);
}

struct span {
    const uint8_t* base;
    const uint64_t* off;
    const uint64_t* len;
    uint8_t* out;
    uint64_t lo, hi;
};

static void* span_main(void* p) {
    struct span* s = (struct span*)p;
    for (uint64_t i = s->lo; i < s->hi; ++i)
        oracle_sha256(s->base + s->off[i], s->len[i], s->out + 32 * i);
    return NULL;
}

/* Same, striped over `threads` host threads (the "port" CPU baseline). */
int oracle_sha256_batch_mt(const uint8_t* base, const uint64_t* off, const uint64_t* len,
                           uint64_t n, uint8_t* out32n, int threads) {
    if (threads < 1) threads = 1;
    if ((uint64_t)threads > n) threads = n ? (int)n : 1;
    pthread_t* tid = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)threads);
    struct span* sp = (struct span*)malloc(sizeof(struct span) * (size_t)threads);
    if (!tid || !sp) { free(tid); free(sp); return -1; }
    for (int t = 0; t < threads; ++t) {
        sp[t].base = base; sp[t].off = off; sp[t].len = len; sp[t].out = out32n;
        sp[t].lo = n * (uint64_t)t / (uint64_t)threads;
        sp[t].hi = n * (uint64_t)(t + 1) / (uint64_t)threads;
        if (pthread_create(&tid[t], NULL, span_main, &sp[t]) != 0) {
            for (int j = 0; j < t; ++j) pthread_join(tid[j], NULL);
            free(tid); free(sp);
            return -2;
        }
    }
    for (int t = 0; t < threads; ++t) pthread_join(tid[t], NULL);
    free(tid); free(sp);
    return 0;
}

/*
 * Synthetic object bytes: 8-byte little-endian words
 *   word(seed, obj, j) = splitmix64_finalise(seed + GOLDEN * ((obj << 32) | j))
 * so that numpy, this file and the device generator kernel
 * (bee_code_interpreter_fs_b200/csrc/synth.cuh) emit identical bytes.
 */
static inline uint64_t synth_word(uint64_t seed, uint64_t obj, uint64_t j) {
    uint64_t z = seed + 0x9E3779B97F4A7C15ull * ((obj << 32) | (j & 0xFFFFFFFFull));
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

/* Object i (global index first_obj+i) occupies dst[i*stride .. i*stride+len). */
void oracle_gen_synthetic(uint8_t* dst, uint64_t seed, uint64_t first_obj, uint64_t n,
                          uint64_t len, uint64_t stride) {
    for (uint64_t i = 0; i < n; ++i) {
        uint8_t* p = dst + i * stride;
        for (uint64_t b = 0; b < len; b += 8) {
            uint64_t w = synth_word(seed, first_obj + i, b / 8);
            uint64_t k = len - b < 8 ? len - b : 8;
            for (uint64_t q = 0; q < k; ++q) p[b + q] = (uint8_t)(w >> (8 * q));
        }
    }
}
