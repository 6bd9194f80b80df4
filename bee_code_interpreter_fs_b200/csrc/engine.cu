// C ABI of the B200 SHA-256 content-addressing engine (see include/b200cas.h).
//
// Host side of the path that the reference leaves to Storage.writer/Storage.write
// (/root/reference/src/code_interpreter/services/storage.py:44-63): take a batch
// of objects, return their SHA-256 object ids.  Device side: csrc/sha256_batch.cuh.
//
// Host-buffer pipeline (cas_hash_host): objects are ordered longest-first and laid
// out 16-byte aligned in a device arena; the arena's byte range is cut into chunks
// (independent of object boundaries) that are packed into pinned staging by a few
// host threads -- or DMA'd straight from the caller's memory when that is pinned
// and contiguous -- and copied with cudaMemcpyAsync on a copy stream.  As soon as
// the last byte of a group of objects has landed, a kernel for that group starts
// on one of several compute streams, so long serial chains from early chunks
// overlap both later copies and later kernels.  Digests come back once at the end.
#include <cuda.h>
#include <cuda_runtime.h>
#include <errno.h>
#include <fcntl.h>
#include <semaphore.h>
#include <sys/stat.h>
#include <sys/syscall.h>
#include <sys/time.h>
#include <unistd.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <chrono>
#include <atomic>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <numeric>
#include <string>
#include <thread>
#include <vector>

#include "../../include/b200cas.h"
#include "hostcopy.h"
#include "peak.cuh"
#include "sha256_batch.cuh"
#include "sha256_long.cuh"
#include "synth.cuh"

namespace {

constexpr uint64_t MiB = 1ull << 20;

thread_local std::string g_create_error;  // errors that have no context (cas_create, cas_hash_host_static), per calling thread

// Small fork-join pool for packing objects into pinned staging.
class PackPool {
public:
    explicit PackPool(int n) : stop_(false), gen_(0), pending_(0) {
        for (int i = 0; i < n; ++i) th_.emplace_back([this, i] { loop(i); });
    }
    ~PackPool() {
        {
            std::lock_guard<std::mutex> lk(m_);
            stop_ = true;
            ++gen_;
        }
        cv_.notify_all();
        for (auto& t : th_) t.join();
    }
    int size() const { return (int)th_.size(); }
    // Runs fn(worker_index) on every worker and waits for all of them.
    void run(const std::function<void(int)>& fn) {
        start(fn);
        wait();
    }
    // The same in two halves, so that the caller can work meanwhile.  `fn` must stay alive until wait() returns.
    void start(const std::function<void(int)>& fn) {
        std::lock_guard<std::mutex> lk(m_);
        fn_ = &fn;
        pending_ = (int)th_.size();
        ++gen_;
        cv_.notify_all();
    }
    void wait() {
        std::unique_lock<std::mutex> lk(m_);
        done_.wait(lk, [this] { return pending_ == 0; });
        fn_ = nullptr;
    }

private:
    void loop(int idx) {
        uint64_t seen = 0;
        for (;;) {
            const std::function<void(int)>* fn;
            {
                std::unique_lock<std::mutex> lk(m_);
                cv_.wait(lk, [&] { return gen_ != seen; });
                seen = gen_;
                if (stop_) return;
                fn = fn_;
            }
            (*fn)(idx);
            {
                std::lock_guard<std::mutex> lk(m_);
                if (--pending_ == 0) done_.notify_all();
            }
        }
    }
    std::vector<std::thread> th_;
    std::mutex m_;
    std::condition_variable cv_, done_;
    const std::function<void(int)>* fn_ = nullptr;
    bool stop_;
    uint64_t gen_;
    int pending_;
};

constexpr int NSTG = 3;   // pinned staging buffers in flight
constexpr int NCOMP = 128;  // compute streams: launch groups overlap, so long serial chains do not queue behind each other

struct Stage {
    uint8_t* h = nullptr;  // pinned staging
    uint64_t cap = 0;
    cudaEvent_t ev = nullptr;  // recorded after the chunk's last H2D copy
    bool used = false;
};

}  // namespace

struct cas_ctx {
    int device = 0;
    int sm_count = 0;
    int clock_khz = 0;
    uint64_t mem_bytes = 0;
    std::string err;
    int kernel = 2;
    int fma_level = 3;
    int ctas_per_sm = 0;  // 0 = per load path: 2 for TMA tiles, 4 for per-lane loads (measured best)
    void* encode_tiled = nullptr;  // cuTensorMapEncodeTiled, resolved through the runtime
    uint64_t wave_bytes = 256 * MiB;
    int pack_threads = 0;
    uint64_t launches = 0;
    uint64_t arena_max = 32ull << 30;  // device bytes the host path may hold (two halves)
    uint64_t long_max_objects = 592;   // batches up to this many objects may go one warp per object (set to 4 x SMs)
    uint64_t long_min_bytes = 32768;   // ... if the objects are at least this long
    int long_max_m = 32;               // groups of up to long_max_objects x this many long objects run M objects per warp
    uint64_t first_chunk = 2 * MiB;    // host path: size of the first copy chunk (chunks double up to the cap)
    uint64_t chunk_div = 16;           // host path: a batch is cut into at least about this many chunks
    cudaStream_t s_copy = nullptr;
    cudaStream_t s_comp[NCOMP] = {};
    cudaEvent_t ev_epoch[2][NCOMP] = {};  // kernels of the epoch that last used an arena half
    cudaEvent_t ev_join[NCOMP] = {};
    Stage stage[NSTG];
    uint8_t* d_arena = nullptr;
    uint64_t arena_cap = 0;
    uint64_t* h_meta = nullptr;  // pinned: offsets[n] then lengths[n], schedule order
    uint64_t* d_meta = nullptr;
    uint8_t* h_dig = nullptr;    // pinned: 32 n
    uint8_t* d_dig = nullptr;
    uint64_t obj_cap = 0;
    PackPool* pool = nullptr;
    PackPool* io_pool = nullptr;  // file writers of cas_store_host (they run while the pack pool feeds the GPU)
    int io_threads = 0;
    bool trace = false;           // B200CAS_TRACE=1: one line of host-pipeline timings per call on stderr
    sem_t* gate = SEM_FAILED;     // optional cross-process limit on concurrent host->device copy chunks (cas_set_h2d_gate)
    std::string gate_name;
    uint64_t gate_timeouts = 0;
    uint32_t* d_sink = nullptr;
    bool host_path_dirty = false;  // set while a host-path call is in progress; stays set if it failed
};

namespace {

int fail(cas_ctx* c, int code, const char* what, cudaError_t e = cudaSuccess) {
    char msg[512];
    if (e != cudaSuccess)
        snprintf(msg, sizeof msg, "%s: %s (%s)", what, cudaGetErrorName(e), cudaGetErrorString(e));
    else
        snprintf(msg, sizeof msg, "%s", what);
    if (c) c->err = msg; else g_create_error = msg;
    return code;
}

#define CU(call)                                                   \
    do {                                                           \
        cudaError_t e_ = (call);                                   \
        if (e_ != cudaSuccess) return fail(ctx, CAS_E_CUDA, #call, e_); \
    } while (0)

constexpr int THREADS = 256;

// cuTensorMapEncodeTiled is a driver-API entry point; it is looked up through the
// runtime so that the library has no link-time dependency on libcuda.
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

bool tma_eligible(const cas::BatchArgs& a) {
    return a.off == nullptr && a.order == nullptr && a.ulen >= 64 && a.ulen < (1ull << 31) && a.n < (1ull << 31) &&
           (a.stride % 16) == 0 && a.stride < (1ull << 40) && (((uintptr_t)a.base) % 16) == 0;
}

template <typename K>
cudaError_t configure_smem(K k, int smem, bool (&done)[16]) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 16 && done[dev]) return cudaSuccess;
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(k, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if (e != cudaSuccess) return e;
    if (dev < 16) done[dev] = true;
    return cudaSuccess;
}

template <int FMA, int MIN_CTAS, int THREADS = 256>
cudaError_t launch_one(cas_ctx* ctx, int kernel, const cas::BatchArgs& a, cudaStream_t st) {
    constexpr int NSTAGE = 2;
    const unsigned grid = (unsigned)((a.n + THREADS - 1) / THREADS);
    if (kernel == 2 && tma_eligible(a)) {
        CUtensorMap tm;
        const cuuint64_t gdim[2] = {(cuuint64_t)a.ulen, (cuuint64_t)a.n};
        const cuuint64_t gstride[1] = {(cuuint64_t)a.stride};
        const cuuint32_t box[2] = {64u, 32u};
        const cuuint32_t estr[2] = {1u, 1u};
        CUresult r = reinterpret_cast<EncodeTiledFn>(ctx->encode_tiled)(
            &tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<uint8_t*>(a.base), gdim, gstride, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return cudaErrorInvalidValue;
        cas::TensorMapBlob blob;
        static_assert(sizeof(blob) == sizeof(tm), "CUtensorMap is 128 bytes");
        memcpy(&blob, &tm, sizeof tm);
        constexpr int smem = cas::tma_smem_bytes<NSTAGE, THREADS, 1>();
        auto k = cas::kernel_tma<FMA, NSTAGE, THREADS, MIN_CTAS>;
        static bool done[16] = {};
        cudaError_t e = configure_smem(k, smem, done);
        if (e != cudaSuccess) return e;
        k<<<grid, THREADS, smem, st>>>(blob, a);
    } else if (kernel == 1) {
        constexpr int smem = cas::bulk_smem_bytes<NSTAGE, THREADS>();
        auto k = cas::kernel_bulk<FMA, NSTAGE, THREADS, MIN_CTAS>;
        static bool done[16] = {};
        cudaError_t e = configure_smem(k, smem, done);
        if (e != cudaSuccess) return e;
        k<<<grid, THREADS, smem, st>>>(a);
    } else {
        cas::kernel_ldg<FMA, THREADS, MIN_CTAS><<<grid, THREADS, 0, st>>>(a);
    }
    return cudaGetLastError();
}

template <int FMA>
cudaError_t launch_occ(cas_ctx* ctx, int kernel, const cas::BatchArgs& a, cudaStream_t st) {
    int ctas = ctx->ctas_per_sm;
    if (ctas == 0) ctas = (kernel == 2 && tma_eligible(a)) ? 2 : 4;
    switch (ctas) {
        case 2: return launch_one<FMA, 2>(ctx, kernel, a, st);
        case 3: return launch_one<FMA, 3>(ctx, kernel, a, st);
        default: return launch_one<FMA, 4>(ctx, kernel, a, st);
    }
}

int launch_batch(cas_ctx* ctx, cas::BatchArgs a, cudaStream_t st) {
    if (a.n == 0) return CAS_OK;
    if (a.n > (1ull << 36)) return fail(ctx, CAS_E_ARG, "too many objects in one launch");
    a.one = 1u;
    a.zero = 0u;
    cudaError_t e;
    // Small batches: one warp per CTA, so that the block scheduler spreads the warps over
    // all SM sub-partitions instead of packing eight of them onto a few SMs (a lone warp on
    // a sub-partition runs its chains ~1.5x faster than two sharing its ALU pipe).
    if (ctx->fma_level == 3 && ctx->ctas_per_sm == 0 && a.n <= 2ull * 4ull * (uint64_t)ctx->sm_count * 32ull) {
        e = launch_one<3, 16, 32>(ctx, ctx->kernel, a, st);
        if (e != cudaSuccess) return fail(ctx, CAS_E_CUDA, "sha256 batch kernel launch (small batch)", e);
        ++ctx->launches;
        return CAS_OK;
    }
    // Uniform batches of just over one wave (3 x 256-thread CTAs = 24 warps per SM): the part that
    // does not fit would run as a nearly empty second wave (512 CTAs on 444 slots cost 1.33 waves,
    // profiles/r02_tail_wave.jsonl).  Seven 128-thread CTAs per SM hold up to 28 warps at the same
    // register budget, so everything up to 1.17 waves is resident at once.  The shard size of
    // BASELINE config 3 split over eight GPUs (131 072 objects) falls in this window.
    if (ctx->fma_level == 3 && ctx->ctas_per_sm == 0 && ctx->kernel == 2 && tma_eligible(a)) {
        const uint64_t warps = (a.n + 31) / 32, sms = (uint64_t)ctx->sm_count;
        if (warps > 24 * sms && warps <= 28 * sms) {
            e = launch_one<3, 7, 128>(ctx, ctx->kernel, a, st);
            if (e != cudaSuccess) return fail(ctx, CAS_E_CUDA, "sha256 batch kernel launch (one-wave shape)", e);
            ++ctx->launches;
            return CAS_OK;
        }
    }
    switch (ctx->fma_level) {
        case 0: e = launch_occ<0>(ctx, ctx->kernel, a, st); break;
        case 1: e = launch_occ<1>(ctx, ctx->kernel, a, st); break;
        case 2: e = launch_occ<2>(ctx, ctx->kernel, a, st); break;
        case 3: e = launch_occ<3>(ctx, ctx->kernel, a, st); break;
        default: e = launch_occ<4>(ctx, ctx->kernel, a, st); break;
    }
    if (e != cudaSuccess) return fail(ctx, CAS_E_CUDA, "sha256 batch kernel launch", e);
    ++ctx->launches;
    return CAS_OK;
}

// One warp per M objects (csrc/sha256_long.cuh): for few, long objects.
template <int M>
cudaError_t launch_long_m(const cas::BatchArgs& a, cudaStream_t st) {
    cas::kernel_long<3, 4, M><<<(unsigned)((a.n + M - 1) / M), 32, cas::LONG_SMEM_PER_WARP, st>>>(a);
    return cudaGetLastError();
}

int launch_long(cas_ctx* ctx, cas::BatchArgs a, cudaStream_t st, int m = 1) {
    if (a.n == 0) return CAS_OK;
    if (a.n > 0x7FFFFFFFull) return fail(ctx, CAS_E_ARG, "too many objects for the warp-per-object kernel");
    a.one = 1u;
    a.zero = 0u;
    cudaError_t e;
    switch (m) {
        case 1: e = launch_long_m<1>(a, st); break;
        case 2: e = launch_long_m<2>(a, st); break;
        case 4: e = launch_long_m<4>(a, st); break;
        case 8: e = launch_long_m<8>(a, st); break;
        case 16: e = launch_long_m<16>(a, st); break;
        case 32: e = launch_long_m<32>(a, st); break;
        default: return fail(ctx, CAS_E_ARG, "objects per warp must be 1, 2, 4, 8, 16 or 32");
    }
    if (e != cudaSuccess) return fail(ctx, CAS_E_CUDA, "sha256 warp-per-object kernel launch", e);
    ++ctx->launches;
    return CAS_OK;
}

// One group of objects whose bytes are on the device: pick the kernel.  `min_len` is the
// shortest object of the group.  Groups small enough to leave SM sub-partitions idle and
// made of objects long enough to amortise the cooperative schedule expansion run M objects
// per warp (kernel_long<M>, csrc/sha256_long.cuh); everything else one object per lane.
int launch_group(cas_ctx* ctx, const cas::BatchArgs& a, uint64_t min_len, cudaStream_t cs) {
    if (ctx->long_max_objects && min_len >= ctx->long_min_bytes && a.n <= ctx->long_max_objects * (uint64_t)ctx->long_max_m) {
        int m = 1;
        while ((uint64_t)m * ctx->long_max_objects < a.n) m *= 2;
        return launch_long(ctx, a, cs, m);
    }
    return launch_batch(ctx, a, cs);
}

inline uint64_t align16(uint64_t x) { return (x + 15u) & ~15ull; }
inline uint64_t nblocks(uint64_t len) { return (len + 9 + 63) / 64; }

void free_host_path(cas_ctx* c) {
    for (auto& st : c->stage) {
        if (st.h) cudaFreeHost(st.h);
        if (st.ev) cudaEventDestroy(st.ev);
        st = Stage();
    }
    if (c->d_arena) cudaFree(c->d_arena);
    if (c->h_meta) cudaFreeHost(c->h_meta);
    if (c->d_meta) cudaFree(c->d_meta);
    if (c->h_dig) cudaFreeHost(c->h_dig);
    if (c->d_dig) cudaFree(c->d_dig);
    c->d_arena = nullptr; c->arena_cap = 0;
    c->h_meta = nullptr; c->d_meta = nullptr; c->h_dig = nullptr; c->d_dig = nullptr; c->obj_cap = 0;
    for (auto& half : c->ev_epoch)
        for (auto& ev : half)
            if (ev) { cudaEventDestroy(ev); ev = nullptr; }
    for (auto& ev : c->ev_join)
        if (ev) { cudaEventDestroy(ev); ev = nullptr; }
}

// The pool serves two kinds of work: packing host memory into pinned staging (memory-bound)
// and reading files into it (syscall-bound).  Both did best with three quarters of a 16-vCPU
// host's threads (profiles/r02_host_sweep.jsonl: 12 pack threads 16.6 ms vs 18.0 with 8 and
// 17.3 with 16 for 10 000 x 64 KiB; file reads 23 ms with 12, 27 with 8, 25 with 16).
int pack_width(const cas_ctx* ctx) {
    if (ctx->pack_threads > 0) return ctx->pack_threads;
    return std::max(1, std::min((int)std::thread::hardware_concurrency() * 3 / 4, 12));
}

int file_width(const cas_ctx* ctx) {
    if (ctx->io_threads > 0) return ctx->io_threads;
    return std::max(1, std::min((int)std::thread::hardware_concurrency() * 3 / 4, 16));
}

void ensure_pack_pool(cas_ctx* ctx) {
    const int want = std::max(pack_width(ctx), file_width(ctx));
    if (ctx->pool && ctx->pool->size() != want) {
        delete ctx->pool;
        ctx->pool = nullptr;
    }
    if (!ctx->pool) ctx->pool = new PackPool(want);
}

// Grow-only resources of the host path: nothing is allocated per call once the
// largest batch shape has been seen.
int ensure_host_path(cas_ctx* ctx, uint64_t arena_bytes, uint64_t stage_bytes, uint64_t objs) {
    if (!ctx->s_copy) CU(cudaStreamCreateWithFlags(&ctx->s_copy, cudaStreamNonBlocking));
    for (auto& st : ctx->s_comp)
        if (!st) CU(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    for (auto& half : ctx->ev_epoch)
        for (auto& ev : half)
            if (!ev) CU(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    for (auto& ev : ctx->ev_join)
        if (!ev) CU(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    ensure_pack_pool(ctx);
    if (arena_bytes > ctx->arena_cap) {
        if (ctx->d_arena) cudaFree(ctx->d_arena);
        ctx->d_arena = nullptr; ctx->arena_cap = 0;
        const uint64_t cap = (arena_bytes + 2 * MiB - 1) & ~(2 * MiB - 1);
        if (cudaMalloc((void**)&ctx->d_arena, cap) != cudaSuccess) {
            cudaGetLastError();
            return fail(ctx, CAS_E_NOMEM, "device arena allocation failed");
        }
        ctx->arena_cap = cap;
    }
    for (auto& st : ctx->stage) {
        if (!st.ev) CU(cudaEventCreateWithFlags(&st.ev, cudaEventDisableTiming));
        if (stage_bytes > st.cap) {
            if (st.h) cudaFreeHost(st.h);
            st.h = nullptr; st.cap = 0;
            const uint64_t cap = (stage_bytes + MiB - 1) & ~(MiB - 1);
            if (cudaHostAlloc((void**)&st.h, cap, cudaHostAllocDefault) != cudaSuccess) {
                cudaGetLastError();
                return fail(ctx, CAS_E_NOMEM, "pinned staging allocation failed");
            }
            st.cap = cap;
        }
    }
    if (objs > ctx->obj_cap) {
        if (ctx->h_meta) cudaFreeHost(ctx->h_meta);
        if (ctx->d_meta) cudaFree(ctx->d_meta);
        if (ctx->h_dig) cudaFreeHost(ctx->h_dig);
        if (ctx->d_dig) cudaFree(ctx->d_dig);
        ctx->h_meta = nullptr; ctx->d_meta = nullptr; ctx->h_dig = nullptr; ctx->d_dig = nullptr; ctx->obj_cap = 0;
        const uint64_t cap = std::max<uint64_t>(1024, objs + objs / 4);
        if (cudaHostAlloc((void**)&ctx->h_meta, cap * 16, cudaHostAllocDefault) != cudaSuccess ||
            cudaHostAlloc((void**)&ctx->h_dig, cap * 32, cudaHostAllocDefault) != cudaSuccess) {
            cudaGetLastError();
            return fail(ctx, CAS_E_NOMEM, "pinned metadata allocation failed");
        }
        if (cudaMalloc((void**)&ctx->d_meta, cap * 16) != cudaSuccess ||
            cudaMalloc((void**)&ctx->d_dig, cap * 32) != cudaSuccess) {
            cudaGetLastError();
            return fail(ctx, CAS_E_NOMEM, "device metadata allocation failed");
        }
        ctx->obj_cap = cap;
    }
    return CAS_OK;
}

bool host_ptr_is_pinned(const void* p) {
    if (!p) return false;
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return at.type == cudaMemoryTypeHost;
}

}  // namespace

extern "C" {

const char* cas_version(void) { return "b200cas 0.1 sm_100a"; }

const char* cas_last_error(const cas_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int cas_create(int device, cas_ctx** out) {
    if (!out) return fail(nullptr, CAS_E_ARG, "cas_create: out is NULL");
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(nullptr, CAS_E_NODEVICE, "cas_create: no CUDA device is visible (this engine has no CPU fallback)", e);
    }
    if (device < 0 || device >= ndev) return fail(nullptr, CAS_E_ARG, "cas_create: device index out of range");
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess)
        return fail(nullptr, CAS_E_CUDA, "cudaGetDeviceProperties", e);
    if (prop.major != 10) {
        char msg[160];
        snprintf(msg, sizeof msg, "cas_create: device %d is sm_%d%d; this library carries sm_100a code only", device,
                 prop.major, prop.minor);
        return fail(nullptr, CAS_E_NODEVICE, msg);
    }
    if ((e = cudaSetDevice(device)) != cudaSuccess) return fail(nullptr, CAS_E_CUDA, "cudaSetDevice", e);
    cas_ctx* ctx = new (std::nothrow) cas_ctx();
    if (!ctx) return fail(nullptr, CAS_E_NOMEM, "cas_create: out of host memory");
    ctx->device = device;
    ctx->sm_count = prop.multiProcessorCount;
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, device);
    ctx->clock_khz = khz;
    ctx->mem_bytes = prop.totalGlobalMem;
    // One warp per SM sub-partition is where the warp-per-object kernel keeps its full per-chain speed;
    // beyond that, lanes (spread one warp per CTA) are faster again.
    ctx->long_max_objects = 4ull * (uint64_t)prop.multiProcessorCount;
    cudaDriverEntryPointQueryResult qres;
    e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ctx->encode_tiled, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !ctx->encode_tiled) {
        delete ctx;
        return fail(nullptr, CAS_E_CUDA, "cuTensorMapEncodeTiled not available from this driver", e);
    }
    if (const char* v = getenv("B200CAS_KERNEL")) ctx->kernel = std::max(0, std::min(2, atoi(v)));
    if (const char* v = getenv("B200CAS_FMA")) ctx->fma_level = std::max(0, std::min(4, atoi(v)));
    if (const char* v = getenv("B200CAS_CTAS")) ctx->ctas_per_sm = atoi(v) == 0 ? 0 : std::max(2, std::min(4, atoi(v)));
    if (const char* v = getenv("B200CAS_TRACE")) ctx->trace = atoi(v) != 0;
    *out = ctx;
    return CAS_OK;
}

void cas_destroy(cas_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    free_host_path(ctx);
    for (auto& st : ctx->s_comp)
        if (st) cudaStreamDestroy(st);
    if (ctx->s_copy) cudaStreamDestroy(ctx->s_copy);
    if (ctx->d_sink) cudaFree(ctx->d_sink);
    if (ctx->gate != SEM_FAILED) sem_close(ctx->gate);
    delete ctx->pool;
    delete ctx->io_pool;
    delete ctx;
}

int cas_set_option(cas_ctx* ctx, int key, int64_t v) {
    if (!ctx) return CAS_E_ARG;
    switch (key) {
        case CAS_OPT_KERNEL:
            if (v < 0 || v > 2) return fail(ctx, CAS_E_ARG, "CAS_OPT_KERNEL must be 0, 1 or 2");
            ctx->kernel = (int)v;
            return CAS_OK;
        case CAS_OPT_FMA_LEVEL:
            if (v < 0 || v > 4) return fail(ctx, CAS_E_ARG, "CAS_OPT_FMA_LEVEL must be 0..4");
            ctx->fma_level = (int)v;
            return CAS_OK;
        case CAS_OPT_WAVE_BYTES:
            if (v < (int64_t)MiB) return fail(ctx, CAS_E_ARG, "CAS_OPT_WAVE_BYTES must be >= 1 MiB");
            ctx->wave_bytes = ((uint64_t)v + MiB - 1) & ~(MiB - 1);
            return CAS_OK;
        case CAS_OPT_PACK_THREADS:
            if (v < 1 || v > 256) return fail(ctx, CAS_E_ARG, "CAS_OPT_PACK_THREADS must be 1..256");
            ctx->pack_threads = (int)v;
            return CAS_OK;
        case CAS_OPT_ARENA_BYTES:
            if (v < (int64_t)(64 * MiB)) return fail(ctx, CAS_E_ARG, "CAS_OPT_ARENA_BYTES must be >= 64 MiB");
            ctx->arena_max = (uint64_t)v;
            return CAS_OK;
        case CAS_OPT_LONG_MAX_OBJECTS:
            if (v < 0 || v > (1 << 20)) return fail(ctx, CAS_E_ARG, "CAS_OPT_LONG_MAX_OBJECTS must be 0..2^20");
            ctx->long_max_objects = (uint64_t)v;
            return CAS_OK;
        case CAS_OPT_LONG_MIN_BYTES:
            if (v < 64) return fail(ctx, CAS_E_ARG, "CAS_OPT_LONG_MIN_BYTES must be >= 64");
            ctx->long_min_bytes = (uint64_t)v;
            return CAS_OK;
        case CAS_OPT_CTAS_PER_SM:
            if (v != 0 && (v < 2 || v > 4)) return fail(ctx, CAS_E_ARG, "CAS_OPT_CTAS_PER_SM must be 0 (auto), 2, 3 or 4");
            ctx->ctas_per_sm = (int)v;
            return CAS_OK;
        case CAS_OPT_LONG_MAX_PER_WARP:
            if (v != 1 && v != 2 && v != 4 && v != 8 && v != 16 && v != 32)
                return fail(ctx, CAS_E_ARG, "CAS_OPT_LONG_MAX_PER_WARP must be 1, 2, 4, 8, 16 or 32");
            ctx->long_max_m = (int)v;
            return CAS_OK;
        case CAS_OPT_FIRST_CHUNK_BYTES:
            if (v < 65536) return fail(ctx, CAS_E_ARG, "CAS_OPT_FIRST_CHUNK_BYTES must be >= 64 KiB");
            ctx->first_chunk = ((uint64_t)v + 65535) & ~65535ull;
            return CAS_OK;
        case CAS_OPT_CHUNK_DIVISOR:
            if (v < 1 || v > 1024) return fail(ctx, CAS_E_ARG, "CAS_OPT_CHUNK_DIVISOR must be 1..1024");
            ctx->chunk_div = (uint64_t)v;
            return CAS_OK;
        case CAS_OPT_IO_THREADS:
            if (v < 1 || v > 256) return fail(ctx, CAS_E_ARG, "CAS_OPT_IO_THREADS must be 1..256");
            ctx->io_threads = (int)v;
            return CAS_OK;
    }
    return fail(ctx, CAS_E_ARG, "unknown option key");
}

int cas_get_option(const cas_ctx* ctx, int key, int64_t* v) {
    if (!ctx || !v) return CAS_E_ARG;
    switch (key) {
        case CAS_OPT_KERNEL: *v = ctx->kernel; return CAS_OK;
        case CAS_OPT_FMA_LEVEL: *v = ctx->fma_level; return CAS_OK;
        case CAS_OPT_WAVE_BYTES: *v = (int64_t)ctx->wave_bytes; return CAS_OK;
        case CAS_OPT_PACK_THREADS: *v = pack_width(ctx); return CAS_OK;
        case CAS_OPT_CTAS_PER_SM: *v = ctx->ctas_per_sm; return CAS_OK;
        case CAS_OPT_ARENA_BYTES: *v = (int64_t)ctx->arena_max; return CAS_OK;
        case CAS_OPT_LONG_MAX_OBJECTS: *v = (int64_t)ctx->long_max_objects; return CAS_OK;
        case CAS_OPT_LONG_MIN_BYTES: *v = (int64_t)ctx->long_min_bytes; return CAS_OK;
        case CAS_OPT_LONG_MAX_PER_WARP: *v = ctx->long_max_m; return CAS_OK;
        case CAS_OPT_FIRST_CHUNK_BYTES: *v = (int64_t)ctx->first_chunk; return CAS_OK;
        case CAS_OPT_IO_THREADS: *v = file_width(ctx); return CAS_OK;
        case CAS_OPT_CHUNK_DIVISOR: *v = (int64_t)ctx->chunk_div; return CAS_OK;
    }
    return CAS_E_ARG;
}

int cas_device_info(const cas_ctx* ctx, int* sm_count, int* sm_clock_khz, uint64_t* mem_bytes) {
    if (!ctx) return CAS_E_ARG;
    if (sm_count) *sm_count = ctx->sm_count;
    if (sm_clock_khz) *sm_clock_khz = ctx->clock_khz;
    if (mem_bytes) *mem_bytes = ctx->mem_bytes;
    return CAS_OK;
}

uint64_t cas_launch_count(const cas_ctx* ctx) { return ctx ? ctx->launches : 0; }

int cas_hash_device(cas_ctx* ctx, const void* d_base, const uint64_t* d_off, const uint64_t* d_len,
                    const uint32_t* d_order, uint64_t n, void* d_out, void* stream) {
    if (!ctx) return CAS_E_ARG;
    if (n == 0) return CAS_OK;
    if (!d_off || !d_len || !d_out) return fail(ctx, CAS_E_ARG, "cas_hash_device: NULL offset/length/output array");
    CU(cudaSetDevice(ctx->device));
    cas::BatchArgs a{};
    a.base = static_cast<const uint8_t*>(d_base);
    a.off = d_off;
    a.len = d_len;
    a.order = d_order;
    a.n = n;
    a.out = static_cast<uint8_t*>(d_out);
    return launch_batch(ctx, a, static_cast<cudaStream_t>(stream));
}

int cas_hash_device_long(cas_ctx* ctx, const void* d_base, const uint64_t* d_off, const uint64_t* d_len,
                         const uint32_t* d_order, uint64_t n, void* d_out, void* stream) {
    if (!ctx) return CAS_E_ARG;
    if (n == 0) return CAS_OK;
    if (!d_off || !d_len || !d_out) return fail(ctx, CAS_E_ARG, "cas_hash_device_long: NULL offset/length/output array");
    CU(cudaSetDevice(ctx->device));
    cas::BatchArgs a{};
    a.base = static_cast<const uint8_t*>(d_base);
    a.off = d_off;
    a.len = d_len;
    a.order = d_order;
    a.n = n;
    a.out = static_cast<uint8_t*>(d_out);
    // As many objects per warp as it takes to keep to about one warp per SM sub-partition.
    int m = 1;
    while (m < ctx->long_max_m && (uint64_t)m * std::max<uint64_t>(ctx->long_max_objects, 1) < n) m *= 2;
    return launch_long(ctx, a, static_cast<cudaStream_t>(stream), m);
}

int cas_hash_device_uniform(cas_ctx* ctx, const void* d_base, uint64_t stride, uint64_t len, uint64_t n,
                            void* d_out, void* stream) {
    if (!ctx) return CAS_E_ARG;
    if (n == 0) return CAS_OK;
    if (!d_out || (!d_base && len)) return fail(ctx, CAS_E_ARG, "cas_hash_device_uniform: NULL buffer");
    if (stride < len && n > 1) return fail(ctx, CAS_E_ARG, "cas_hash_device_uniform: stride smaller than length");
    CU(cudaSetDevice(ctx->device));
    cas::BatchArgs a{};
    a.base = static_cast<const uint8_t*>(d_base);
    a.stride = stride;
    a.ulen = len;
    a.n = n;
    a.out = static_cast<uint8_t*>(d_out);
    // Few long objects: lane-per-object would leave the GPU empty and every chain slow.
    return launch_group(ctx, a, len, static_cast<cudaStream_t>(stream));
}

int cas_gen_synthetic(cas_ctx* ctx, void* d_dst, uint64_t seed, uint64_t first_obj, uint64_t n, uint64_t len,
                      uint64_t stride, void* stream) {
    if (!ctx) return CAS_E_ARG;
    if (n == 0 || len == 0) return CAS_OK;
    if (!d_dst) return fail(ctx, CAS_E_ARG, "cas_gen_synthetic: NULL destination");
    if (stride < len && n > 1) return fail(ctx, CAS_E_ARG, "cas_gen_synthetic: stride smaller than length");
    CU(cudaSetDevice(ctx->device));
    const uint64_t segs = (len + cas::SYNTH_SEG - 1) / cas::SYNTH_SEG;
    // Launch in slices so that the 1-D grid stays below 2^31 CTAs.
    const uint64_t max_objs = std::max<uint64_t>(1, ((1ull << 31) - 1) / segs);
    for (uint64_t done = 0; done < n; done += max_objs) {
        const uint64_t m = std::min(max_objs, n - done);
        cas::kernel_synth<<<(unsigned)(m * segs), cas::SYNTH_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(
            static_cast<uint8_t*>(d_dst) + done * stride, seed, first_obj + done, m, len, stride, segs);
        CU(cudaGetLastError());
        ++ctx->launches;
    }
    return CAS_OK;
}

int cas_int32_peak(cas_ctx* ctx, int mode, int steps, double* lane_ops_per_s, double* ms_out) {
    if (!ctx) return CAS_E_ARG;
    if (mode < 0 || mode > 6 || steps < cas::PEAK_UNROLL) return fail(ctx, CAS_E_ARG, "cas_int32_peak: bad mode/steps");
    CU(cudaSetDevice(ctx->device));
    if (!ctx->d_sink) CU(cudaMalloc((void**)&ctx->d_sink, 256));
    steps -= steps % cas::PEAK_UNROLL;
    const unsigned grid = (unsigned)ctx->sm_count * 8u;
    cudaEvent_t e0, e1;
    CU(cudaEventCreate(&e0));
    CU(cudaEventCreate(&e1));
    float best = 0.f;
    for (int rep = 0; rep < 3; ++rep) {  // first repetition is the warm-up
        CU(cudaEventRecord(e0, 0));
        switch (mode) {
            case 0: cas::kernel_peak<0><<<grid, cas::PEAK_THREADS>>>(ctx->d_sink, steps, 1u, 12345u + rep); break;
            case 1: cas::kernel_peak<1><<<grid, cas::PEAK_THREADS>>>(ctx->d_sink, steps, 1u, 12345u + rep); break;
            case 2: cas::kernel_peak<2><<<grid, cas::PEAK_THREADS>>>(ctx->d_sink, steps, 1u, 12345u + rep); break;
            case 3: cas::kernel_peak<3><<<grid, cas::PEAK_THREADS>>>(ctx->d_sink, steps, 1u, 12345u + rep); break;
            case 4: cas::kernel_peak<4><<<grid, cas::PEAK_THREADS>>>(ctx->d_sink, steps, 1u, 12345u + rep); break;
            case 5: cas::kernel_peak<5><<<grid, cas::PEAK_THREADS>>>(ctx->d_sink, steps, 1u, 12345u + rep); break;
            default: cas::kernel_peak<6><<<grid, cas::PEAK_THREADS>>>(ctx->d_sink, steps, 1u, 12345u + rep); break;
        }
        CU(cudaGetLastError());
        CU(cudaEventRecord(e1, 0));
        CU(cudaEventSynchronize(e1));
        float ms = 0.f;
        CU(cudaEventElapsedTime(&ms, e0, e1));
        ++ctx->launches;
        if (rep > 0 && (best == 0.f || ms < best)) best = ms;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    const double ops = (double)grid * cas::PEAK_THREADS * (double)steps * cas::peak_ops_per_step(mode);
    if (lane_ops_per_s) *lane_ops_per_s = ops / (best * 1e-3);
    if (ms_out) *ms_out = best;
    return CAS_OK;
}

}  // extern "C"

namespace {

// Where the bytes of object i come from: host memory (hbase + off[i]) or a file (paths[i]).
struct ByteSource {
    const uint8_t* hbase = nullptr;
    const uint64_t* off = nullptr;
    const char* const* paths = nullptr;
};

// Reads [offset, offset+bytes) of a file into dst; false if the file is shorter or unreadable.
bool read_file_range(const char* path, uint64_t offset, uint8_t* dst, uint64_t bytes) {
    const int fd = open(path, O_RDONLY | O_CLOEXEC);
    if (fd < 0) return false;
    uint64_t done = 0;
    while (done < bytes) {
        const ssize_t r = pread(fd, dst + done, bytes - done, (off_t)(offset + done));
        if (r <= 0) {
            if (r < 0 && errno == EINTR) continue;
            break;
        }
        done += (uint64_t)r;
    }
    close(fd);
    return done == bytes;
}

// ------------------------------------------------------------------ store-side file helpers --
void ensure_io_pool(cas_ctx* ctx) {
    const int want = file_width(ctx);
    if (ctx->io_pool && ctx->io_pool->size() != want) {
        delete ctx->io_pool;
        ctx->io_pool = nullptr;
    }
    if (!ctx->io_pool) ctx->io_pool = new PackPool(want);
}

// mkdir -p
bool make_dirs(const char* dir) {
    struct stat sb;
    if (stat(dir, &sb) == 0) return S_ISDIR(sb.st_mode);
    std::string p(dir);
    for (size_t i = 1; i <= p.size(); ++i) {
        if (i != p.size() && p[i] != '/') continue;
        const std::string part = p.substr(0, i);
        if (mkdir(part.c_str(), 0777) != 0 && errno != EEXIST) return false;
    }
    return stat(dir, &sb) == 0 && S_ISDIR(sb.st_mode);
}

uint64_t fresh_token() {
    static std::atomic<uint64_t> counter{0};
    uint64_t t = 0;
    const int fd = open("/dev/urandom", O_RDONLY | O_CLOEXEC);
    if (fd >= 0) {
        if (read(fd, &t, sizeof t) != (ssize_t)sizeof t) t = 0;
        close(fd);
    }
    struct timeval tv;
    gettimeofday(&tv, nullptr);
    return t ^ ((uint64_t)getpid() << 40) ^ ((uint64_t)tv.tv_usec << 20) ^ (uint64_t)tv.tv_sec ^ (counter.fetch_add(1) << 56);
}

// Temporary files live in per-writer-thread directories "<dir>/.tmp-<token>.<t>/": creating a
// file takes its parent directory's lock, and with one directory per thread the writers do
// not queue on it (nor on the store directory, which then only sees the final links).
// The names start with '.', which no object id can (the reference's Hash pattern,
// utils/validation.py:19), so a reader can never mistake anything in there for an object.
void tmp_dir_name(char* dst, size_t cap, uint64_t token, unsigned writer) {
    snprintf(dst, cap, ".tmp-%016llx.%u", (unsigned long long)token, writer);
}

std::string tmp_path(const std::string& root, uint64_t token, unsigned writer, uint64_t i) {
    char name[96];
    snprintf(name, sizeof name, "/.tmp-%016llx.%u/%llu", (unsigned long long)token, writer, (unsigned long long)i);
    return root + name;
}

bool write_whole_file(const char* path, const uint8_t* src, uint64_t bytes) {
    const int fd = open(path, O_WRONLY | O_CREAT | O_EXCL | O_CLOEXEC, 0666);
    if (fd < 0) return false;
    uint64_t done = 0;
    while (done < bytes) {
        const ssize_t w = write(fd, src + done, bytes - done);
        if (w < 0) {
            if (errno == EINTR) continue;
            break;
        }
        done += (uint64_t)w;
    }
    const bool ok = close(fd) == 0 && done == bytes;
    if (!ok) unlink(path);
    return ok;
}

void hex_name(char* dst, const uint8_t* digest) {
    static const char* H = "0123456789abcdef";
    for (int i = 0; i < 32; ++i) { dst[2 * i] = H[digest[i] >> 4]; dst[2 * i + 1] = H[digest[i] & 15]; }
    dst[64] = 0;
}

// Gives `tmp` its final name unless that name is taken: 0 = done, 1 = an object with this id
// was there already (tmp removed), < 0 = error.  Never replaces a file, so a reader holding
// the existing object open is not disturbed, and the final name only ever appears complete.
// A hard link plus unlink is tried first: unlike a rename between two directories it does not
// take the file system's rename lock, so many threads can publish at once; rename (without
// replacement) is the fallback for file systems without links.  On an error the temporary file
// is removed if it is one of ours; a caller's spool file is left for the caller to deal with.
int publish_one(const char* tmp, const char* target, bool ours) {
    if (link(tmp, target) == 0) { unlink(tmp); return CAS_STORED_NEW; }
    if (errno == EEXIST) { unlink(tmp); return CAS_STORED_EXISTED; }
#if defined(SYS_renameat2)
    if (syscall(SYS_renameat2, AT_FDCWD, tmp, AT_FDCWD, target, 1u /* RENAME_NOREPLACE */) == 0) return CAS_STORED_NEW;
    if (errno == EEXIST) { unlink(tmp); return CAS_STORED_EXISTED; }
#endif
    if (ours) unlink(tmp);
    return CAS_E_IO;
}

// Final names for the caller's spool files (cas_publish_files).  Files whose status is already
// negative are skipped.
void publish_paths(cas_ctx* ctx, const std::string& root, const char* const* paths, uint64_t n, const uint8_t* digests,
                   int32_t* status) {
    std::atomic<uint64_t> next{0};
    const std::function<void(int)> job = [&](int) {
        char hex[65];
        for (;;) {
            const uint64_t i0 = next.fetch_add(16, std::memory_order_relaxed);
            if (i0 >= n) return;
            for (uint64_t i = i0; i < std::min(n, i0 + 16); ++i) {
                if (status[i] < 0) continue;
                hex_name(hex, digests + 32 * i);
                status[i] = publish_one(paths[i], (root + "/" + hex).c_str(), false);
            }
        }
    };
    ctx->io_pool->run(job);
}

// Sizes of n files, looked up by the pack pool's threads.  With status == NULL a file that is
// missing or not regular fails the call; otherwise it gets status CAS_E_IO and length 0.
int stat_files(cas_ctx* ctx, const char* const* paths, uint64_t n, uint64_t* len, int32_t* status) {
    ensure_pack_pool(ctx);
    std::atomic<uint64_t> next{0};
    std::atomic<int64_t> first_bad{-1};
    const std::function<void(int)> job = [&](int) {
        for (;;) {
            const uint64_t i0 = next.fetch_add(64, std::memory_order_relaxed);
            if (i0 >= n) return;
            for (uint64_t i = i0; i < std::min(n, i0 + 64); ++i) {
                struct stat sb;
                if (paths[i] && stat(paths[i], &sb) == 0 && S_ISREG(sb.st_mode)) {
                    len[i] = (uint64_t)sb.st_size;
                } else {
                    len[i] = 0;
                    if (status) status[i] = CAS_E_IO;
                    int64_t expect = -1;
                    first_bad.compare_exchange_strong(expect, (int64_t)i);
                }
            }
        }
    };
    ctx->pool->run(job);
    if (!status && first_bad.load() >= 0) {
        const char* bad = paths[first_bad.load()];
        char msg[400];
        snprintf(msg, sizeof msg, "cas_hash_files: cannot stat regular file '%.300s'", bad ? bad : "(null)");
        return fail(ctx, CAS_E_IO, msg);
    }
    return CAS_OK;
}

// Host->device gate: a named POSIX semaphore shared by the processes (or contexts) that copy
// through the same host uplink.  A chunk's copies are enqueued only while a slot is held; the
// slot is given back by a host function that runs in the copy stream right after them.
bool gate_acquire(cas_ctx* ctx) {
    if (ctx->gate == SEM_FAILED) return false;
    timespec ts;
    clock_gettime(CLOCK_REALTIME, &ts);
    ts.tv_sec += 1;  // never wait for ever on a slot somebody may have died with
    for (;;) {
        if (sem_timedwait(ctx->gate, &ts) == 0) return true;
        if (errno == EINTR) continue;
        ++ctx->gate_timeouts;
        return false;
    }
}

void CUDART_CB gate_release(void* sem) { sem_post(static_cast<sem_t*>(sem)); }

double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

int hash_host_impl(cas_ctx* ctx, const ByteSource& srcs, const uint64_t* len, uint64_t n, uint8_t* out,
                   int32_t* status = nullptr) {
    const uint8_t* hbase = srcs.hbase;
    const uint64_t* off = srcs.off;
    const bool from_files = srcs.paths != nullptr;
    const double t_begin = ctx->trace ? now_s() : 0.0;
    double t_pack = 0.0, t_stage_wait = 0.0;
    // A call that failed half-way may have left copies and kernels in flight on buffers
    // this call is about to reuse: wait them out before touching anything.
    if (ctx->host_path_dirty) CU(cudaDeviceSynchronize());
    ctx->host_path_dirty = true;

    // Schedule: longest first when lengths differ, so that the 32 lanes of a warp
    // carry messages of (nearly) equal block counts and the long chains start early.
    uint64_t total = 0, maxlen = 0;
    bool uniform = true;
    for (uint64_t i = 0; i < n; ++i) {
        total += align16(len[i]);
        maxlen = std::max(maxlen, len[i]);
        uniform = uniform && len[i] == len[0];
    }
    std::vector<uint32_t> order(n);
    std::iota(order.begin(), order.end(), 0u);
    if (!uniform)
        std::stable_sort(order.begin(), order.end(),
                         [&](uint32_t a, uint32_t b) { return nblocks(len[a]) > nblocks(len[b]); });

    // The longest objects go one warp per object when their serial chains, not the
    // batch's total work, would set the time (a few big files among many small ones).
    uint64_t n_long = 0;
    if (!uniform || n <= ctx->long_max_objects) {
        const double t_serial = (double)maxlen / 25e6, t_bulk = (double)total / 900e9;
        if (t_serial > 2.0 * t_bulk) {
            const uint64_t floor_len = std::max<uint64_t>(ctx->long_min_bytes, maxlen / 4);  // measured better than /3 on config 4
            while (n_long < n && n_long < ctx->long_max_objects && len[order[n_long]] >= floor_len) ++n_long;
        }
    }

    // Chunk (copy granule) and arena sizes.  The first chunk is small so that the first
    // copy and the first kernel start almost at once; chunks then double up to a cap that
    // keeps at least ~8 of them in a batch (copy and hash overlap) and at most wave_bytes.
    // The arena has two halves so that the copies of one epoch run while the kernels of
    // the previous one finish.
    const uint64_t chunk_cap = std::min<uint64_t>(ctx->wave_bytes, std::max<uint64_t>(8 * MiB, (total / ctx->chunk_div + MiB - 1) & ~(MiB - 1)));
    uint64_t chunk_now = std::min<uint64_t>(chunk_cap, ctx->first_chunk);
    uint64_t half = std::max<uint64_t>(std::min<uint64_t>(ctx->arena_max / 2, (total + MiB - 1) & ~(MiB - 1)),
                                       std::max<uint64_t>(align16(maxlen), 16));
    half = (half + 255) & ~255ull;
    const bool one_epoch = total <= half;
    int rc = ensure_host_path(ctx, one_epoch ? half : 2 * half, chunk_cap, n);
    if (rc) return rc;
    // Page-locked source memory is DMA'd straight from where it lies.  With absolute addresses
    // (hbase == NULL) the first, middle and last object stand for the batch; should other
    // objects be pageable after all, their copies still work, only slower.
    bool pinned_src = false;
    if (!from_files) {
        if (hbase) {
            pinned_src = host_ptr_is_pinned(hbase);
        } else if (total >= 8 * MiB) {
            pinned_src = true;
            for (uint64_t k : {uint64_t(0), n / 2, n - 1})
                pinned_src = pinned_src && (len[k] == 0 || host_ptr_is_pinned(reinterpret_cast<const void*>(off[k])));
        }
    }

    // Device layout in schedule order; epochs are runs of objects that fit one arena half.
    uint64_t* h_off = ctx->h_meta;
    uint64_t* h_len = ctx->h_meta + n;
    std::vector<uint64_t> epoch_first;  // schedule index where each epoch starts
    {
        uint64_t cursor = 0;
        epoch_first.push_back(0);
        for (uint64_t k = 0; k < n; ++k) {
            const uint64_t need = align16(len[order[k]]);
            if (cursor + need > half) {
                epoch_first.push_back(k);
                cursor = 0;
            }
            h_off[k] = cursor;
            h_len[k] = len[order[k]];
            cursor += need;
        }
        epoch_first.push_back(n);
    }
    CU(cudaMemcpyAsync(ctx->d_meta, h_off, n * 8, cudaMemcpyHostToDevice, ctx->s_copy));
    CU(cudaMemcpyAsync(ctx->d_meta + n, h_len, n * 8, cudaMemcpyHostToDevice, ctx->s_copy));
    for (auto& st : ctx->stage) st.used = false;
    const double t_prep = ctx->trace ? now_s() : 0.0;

    // A run of bytes of one object: from memory (src) or from file `obj` at `fpos`.
    struct Piece { const uint8_t* src; uint64_t dst; uint64_t bytes; uint64_t obj; uint64_t fpos; };
    std::atomic<bool> io_failed{false};
    std::vector<Piece> pieces;
    uint64_t chunk_no = 0, launch_no = 0;
    for (size_t e = 0; e + 1 < epoch_first.size(); ++e) {
        const uint64_t k0 = epoch_first[e], k1 = epoch_first[e + 1];
        uint8_t* d_base = ctx->d_arena + (one_epoch ? 0 : (e & 1) * half);
        if (e >= 2)  // this half was last read by the kernels of epoch e-2
            for (int i = 0; i < NCOMP; ++i) CU(cudaStreamWaitEvent(ctx->s_copy, ctx->ev_epoch[e & 1][i], 0));
        const uint64_t bytes = h_off[k1 - 1] + align16(h_len[k1 - 1]);
        uint64_t kp = k0;       // first object that still has bytes to copy
        uint64_t klaunch = k0;  // first object not yet handed to a kernel
        for (uint64_t c0 = 0; c0 < bytes || klaunch < k1;) {
            const uint64_t c1 = std::min(bytes, c0 + chunk_now);
            chunk_now = std::min(chunk_cap, chunk_now * 2);
            Stage& sg = ctx->stage[chunk_no % NSTG];
            // Pieces of objects that fall into [c0, c1), merged when contiguous on both sides.
            pieces.clear();
            while (kp < k1 && h_off[kp] + h_len[kp] <= c0) ++kp;
            for (uint64_t k = kp; k < k1 && h_off[k] < c1; ++k) {
                const uint64_t lo = std::max(h_off[k], c0), hi = std::min(h_off[k] + h_len[k], c1);
                if (lo >= hi) continue;
                if (from_files) {
                    pieces.push_back({nullptr, lo, hi - lo, order[k], lo - h_off[k]});
                    continue;
                }
                const uint8_t* src = hbase + off[order[k]] + (lo - h_off[k]);
                if (!pieces.empty() && pieces.back().src + pieces.back().bytes == src &&
                    pieces.back().dst + pieces.back().bytes == lo)
                    pieces.back().bytes += hi - lo;
                else
                    pieces.push_back({src, lo, hi - lo, order[k], 0});
            }
            if (c1 > c0) {
                if (pinned_src && pieces.size() <= 64) {
                    const bool held = gate_acquire(ctx);
                    cudaError_t ce = cudaSuccess;
                    for (auto& pc : pieces)
                        if (ce == cudaSuccess)
                            ce = cudaMemcpyAsync(d_base + pc.dst, pc.src, pc.bytes, cudaMemcpyHostToDevice, ctx->s_copy);
                    if (held && (ce != cudaSuccess || cudaLaunchHostFunc(ctx->s_copy, gate_release, ctx->gate) != cudaSuccess))
                        sem_post(ctx->gate);  // the slot must not be lost with the failed call
                    if (ce != cudaSuccess) return fail(ctx, CAS_E_CUDA, "host->device copy of a chunk", ce);
                } else if (!pieces.empty()) {
                    const double tw0 = ctx->trace ? now_s() : 0.0;
                    if (sg.used) CU(cudaEventSynchronize(sg.ev));  // its previous H2D has left the staging
                    const double tw1 = ctx->trace ? now_s() : 0.0;
                    const int T = std::min(ctx->pool->size(), from_files ? file_width(ctx) : pack_width(ctx));
                    const uint64_t span = c1 - c0, share = ((span + T - 1) / (uint64_t)T + 63) & ~63ull;
                    uint8_t* stage = sg.h;
                    const Piece* pv = pieces.data();
                    const size_t np = pieces.size();
                    ctx->pool->run([&](int t) {
                        if (t >= T) return;
                        const uint64_t lo = c0 + share * (uint64_t)t, hi = std::min(c1, lo + share);
                        if (lo >= hi) return;
                        // first piece that ends after lo
                        size_t i = std::partition_point(pv, pv + np, [&](const Piece& q) { return q.dst + q.bytes <= lo; }) - pv;
                        for (; i < np && pv[i].dst < hi; ++i) {
                            const uint64_t a = std::max(pv[i].dst, lo), b = std::min(pv[i].dst + pv[i].bytes, hi);
                            if (pv[i].src) {
                                cas::copy_to_staging(stage + (a - c0), pv[i].src + (a - pv[i].dst), b - a);
                            } else if (status && status[pv[i].obj] < 0) {
                                memset(stage + (a - c0), 0, b - a);  // already failed: keep the staging defined
                            } else if (!read_file_range(srcs.paths[pv[i].obj], pv[i].fpos + (a - pv[i].dst), stage + (a - c0), b - a)) {
                                // shorter than stat() said, or unreadable
                                if (status) {
                                    status[pv[i].obj] = CAS_E_IO;  // this object only; its digest is not reported
                                    memset(stage + (a - c0), 0, b - a);
                                } else {
                                    io_failed.store(true);
                                }
                            }
                        }
                    });
                    if (ctx->trace) { t_stage_wait += tw1 - tw0; t_pack += now_s() - tw1; }
                    if (io_failed.load()) {
                        cudaDeviceSynchronize();
                        return fail(ctx, CAS_E_IO, "a file changed or could not be read while it was being hashed");
                    }
                    CU(cudaMemcpyAsync(d_base + c0, stage, span, cudaMemcpyHostToDevice, ctx->s_copy));
                    sg.used = true;
                }
            }
            CU(cudaEventRecord(sg.ev, ctx->s_copy));
            ++chunk_no;
            // Objects whose last byte is now on the device.
            uint64_t kend = klaunch;
            while (kend < k1 && h_off[kend] + h_len[kend] <= c1) ++kend;
            if (c1 >= bytes) kend = k1;
            while (kend > klaunch) {
                // [klaunch, split) goes one warp per object, [split, kend) by group size.
                const bool is_long = klaunch < n_long;
                const uint64_t split = is_long ? std::min(kend, n_long) : kend;
                cudaStream_t cs = ctx->s_comp[launch_no % NCOMP];
                CU(cudaStreamWaitEvent(cs, sg.ev, 0));
                cas::BatchArgs a{};
                a.base = d_base;
                a.off = ctx->d_meta + klaunch;
                a.len = ctx->d_meta + n + klaunch;
                a.n = split - klaunch;
                a.out = ctx->d_dig + 32ull * klaunch;
                // objects are in descending order of compressions: the group's last one is (within 63 bytes) its shortest
                const uint64_t min_len = h_len[split - 1] & ~63ull;
                if ((rc = is_long ? launch_long(ctx, a, cs, 1) : launch_group(ctx, a, min_len, cs))) return rc;
                ++launch_no;
                klaunch = split;
            }
            if (c1 >= bytes) break;
            c0 = c1;
        }
        if (e + 3 < epoch_first.size())  // a later epoch will reuse this half
            for (int i = 0; i < NCOMP; ++i) CU(cudaEventRecord(ctx->ev_epoch[e & 1][i], ctx->s_comp[i]));
    }
    // Join the compute streams that were used, bring the digests back, restore the caller's order.
    const int used_streams = (int)std::min<uint64_t>(launch_no, NCOMP);
    for (int i = 1; i < used_streams; ++i) {
        CU(cudaEventRecord(ctx->ev_join[i], ctx->s_comp[i]));
        CU(cudaStreamWaitEvent(ctx->s_comp[0], ctx->ev_join[i], 0));
    }
    const double t_issued = ctx->trace ? now_s() : 0.0;
    CU(cudaMemcpyAsync(ctx->h_dig, ctx->d_dig, n * 32, cudaMemcpyDeviceToHost, ctx->s_comp[0]));
    CU(cudaStreamSynchronize(ctx->s_comp[0]));
    CU(cudaStreamSynchronize(ctx->s_copy));
    if (uniform) {
        memcpy(out, ctx->h_dig, n * 32);
    } else {
        for (uint64_t k = 0; k < n; ++k) memcpy(out + 32ull * order[k], ctx->h_dig + 32ull * k, 32);
    }
    ctx->host_path_dirty = false;
    if (ctx->trace) {
        const double t_end = now_s();
        fprintf(stderr,
                "{\"b200cas_trace\": \"host_path\", \"objects\": %llu, \"bytes\": %llu, \"chunks\": %llu, \"launches\": %llu, "
                "\"pinned_src\": %s, \"files\": %s, \"ms_total\": %.3f, \"ms_prepare\": %.3f, \"ms_pack\": %.3f, "
                "\"ms_wait_for_staging\": %.3f, \"ms_issue_other\": %.3f, \"ms_final_wait\": %.3f}\n",
                (unsigned long long)n, (unsigned long long)total, (unsigned long long)chunk_no, (unsigned long long)launch_no,
                pinned_src ? "true" : "false", from_files ? "true" : "false", (t_end - t_begin) * 1e3, (t_prep - t_begin) * 1e3,
                t_pack * 1e3, t_stage_wait * 1e3, (t_issued - t_prep - t_pack - t_stage_wait) * 1e3, (t_end - t_issued) * 1e3);
    }
    return CAS_OK;
}

}  // namespace

extern "C" {

int cas_hash_host(cas_ctx* ctx, const void* base, const uint64_t* off, const uint64_t* len, uint64_t n,
                  void* out32n) {
    if (!ctx) return CAS_E_ARG;
    if (n == 0) return CAS_OK;
    if (!off || !len || !out32n) return fail(ctx, CAS_E_ARG, "cas_hash_host: NULL offset/length/output array");
    if (n > 0xFFFFFFFFull) return fail(ctx, CAS_E_ARG, "cas_hash_host: more than 2^32-1 objects in one call");
    CU(cudaSetDevice(ctx->device));
    ByteSource src;
    src.hbase = static_cast<const uint8_t*>(base);
    src.off = off;
    return hash_host_impl(ctx, src, len, n, static_cast<uint8_t*>(out32n));
}

int cas_hash_files(cas_ctx* ctx, const char* const* paths, uint64_t n, void* out32n, uint64_t* sizes_out) {
    if (!ctx) return CAS_E_ARG;
    if (n == 0) return CAS_OK;
    if (!paths || !out32n) return fail(ctx, CAS_E_ARG, "cas_hash_files: NULL path or output array");
    if (n > 0xFFFFFFFFull) return fail(ctx, CAS_E_ARG, "cas_hash_files: more than 2^32-1 files in one call");
    CU(cudaSetDevice(ctx->device));
    std::vector<uint64_t> len(n);
    int rc = stat_files(ctx, paths, n, len.data(), nullptr);
    if (rc) return rc;
    if (sizes_out) memcpy(sizes_out, len.data(), n * sizeof(uint64_t));
    ByteSource src;
    src.paths = paths;
    return hash_host_impl(ctx, src, len.data(), n, static_cast<uint8_t*>(out32n));
}

int cas_set_h2d_gate(cas_ctx* ctx, const char* name, int slots) {
    if (!ctx) return CAS_E_ARG;
    if (ctx->gate != SEM_FAILED) {
        if (ctx->s_copy) cudaStreamSynchronize(ctx->s_copy);  // a failed call may have left a release queued
        sem_close(ctx->gate);
        ctx->gate = SEM_FAILED;
        ctx->gate_name.clear();
    }
    if (!name) return CAS_OK;
    if (name[0] != '/' || slots < 1) return fail(ctx, CAS_E_ARG, "cas_set_h2d_gate: name must start with '/' and slots be >= 1");
    sem_t* s = sem_open(name, O_CREAT, 0600, (unsigned)slots);
    if (s == SEM_FAILED) return fail(ctx, CAS_E_IO, "cas_set_h2d_gate: sem_open failed");
    ctx->gate = s;
    ctx->gate_name = name;
    return CAS_OK;
}

int cas_reserve(cas_ctx* ctx, uint64_t batch_bytes, uint64_t max_objects) {
    if (!ctx) return CAS_E_ARG;
    CU(cudaSetDevice(ctx->device));
    if (ctx->host_path_dirty) CU(cudaDeviceSynchronize());
    const uint64_t total = std::max<uint64_t>(batch_bytes, MiB);
    const uint64_t chunk_cap = std::min<uint64_t>(ctx->wave_bytes, std::max<uint64_t>(8 * MiB, (total / ctx->chunk_div + MiB - 1) & ~(MiB - 1)));
    uint64_t half = std::min<uint64_t>(ctx->arena_max / 2, (total + MiB - 1) & ~(MiB - 1));
    half = (half + 255) & ~255ull;
    return ensure_host_path(ctx, total <= half ? half : 2 * half, chunk_cap, std::max<uint64_t>(max_objects, 1));
}

int cas_store_host(cas_ctx* ctx, const char* dir, const void* base, const uint64_t* off, const uint64_t* len,
                   uint64_t n, void* out32n, int32_t* status_out) {
    if (!ctx) return CAS_E_ARG;
    if (n == 0) return CAS_OK;
    if (!dir || !off || !len || !out32n) return fail(ctx, CAS_E_ARG, "cas_store_host: NULL directory/offset/length/output array");
    if (n > 0xFFFFFFFFull) return fail(ctx, CAS_E_ARG, "cas_store_host: more than 2^32-1 objects in one call");
    CU(cudaSetDevice(ctx->device));
    if (!make_dirs(dir)) return fail(ctx, CAS_E_IO, "cas_store_host: cannot create the store directory");
    ensure_io_pool(ctx);
    std::vector<int32_t> local_status;
    int32_t* status = status_out;
    if (!status) { local_status.resize(n); status = local_status.data(); }
    for (uint64_t i = 0; i < n; ++i) status[i] = CAS_STORED_NEW;
    const uint8_t* hbase = static_cast<const uint8_t*>(base);
    const std::string root(dir);
    const uint64_t token = fresh_token();

    // File writers run while this thread drives the GPU pipeline over the same bytes.  A writer
    // publishes its own files: at once for files written after the digests have arrived, in
    // one sweep for those written before (so there is no separate rename phase at the end).
    std::atomic<uint64_t> next{0};
    std::atomic<int> outcome{0};  // 0 = still hashing, 1 = digests are in out32n, 2 = hashing failed
    const uint8_t* digests = static_cast<const uint8_t*>(out32n);
    const std::function<void(int)> write_all = [&](int t) {
        bool have_dir = false;
        char name[64], hex[65];
        std::vector<uint64_t> waiting;  // written, not yet published
        auto publish = [&](uint64_t i) {
            hex_name(hex, digests + 32 * i);
            status[i] = publish_one(tmp_path(root, token, (unsigned)t, i).c_str(), (root + "/" + hex).c_str(), true);
        };
        for (;;) {
            const uint64_t i = next.fetch_add(1, std::memory_order_relaxed);
            if (i >= n) break;
            if (!have_dir) {
                tmp_dir_name(name, sizeof name, token, (unsigned)t);
                have_dir = mkdir((root + "/" + name).c_str(), 0777) == 0 || errno == EEXIST;
            }
            if (!have_dir || !write_whole_file(tmp_path(root, token, (unsigned)t, i).c_str(), hbase + off[i], len[i])) {
                status[i] = CAS_E_IO;
                continue;
            }
            if (outcome.load(std::memory_order_acquire) == 1) publish(i);
            else waiting.push_back(i);
        }
        int how;
        while ((how = outcome.load(std::memory_order_acquire)) == 0) std::this_thread::sleep_for(std::chrono::microseconds(50));
        for (uint64_t i : waiting) {
            if (how == 1) publish(i);
            else unlink(tmp_path(root, token, (unsigned)t, i).c_str());  // no ids: nothing may appear under a final name
        }
        if (have_dir) rmdir((root + "/" + name).c_str());
    };
    ctx->io_pool->start(write_all);
    ByteSource src;
    src.hbase = hbase;
    src.off = off;
    const int rc = hash_host_impl(ctx, src, len, n, static_cast<uint8_t*>(out32n));
    outcome.store(rc ? 2 : 1, std::memory_order_release);
    ctx->io_pool->wait();
    return rc;
}

int cas_publish_files(cas_ctx* ctx, const char* dir, const char* const* paths, uint64_t n, void* out32n,
                      uint64_t* sizes_out, int32_t* status_out) {
    if (!ctx) return CAS_E_ARG;
    if (n == 0) return CAS_OK;
    if (!paths || !out32n || !status_out) return fail(ctx, CAS_E_ARG, "cas_publish_files: NULL path, output or status array");
    if (n > 0xFFFFFFFFull) return fail(ctx, CAS_E_ARG, "cas_publish_files: more than 2^32-1 files in one call");
    CU(cudaSetDevice(ctx->device));
    if (dir && !make_dirs(dir)) return fail(ctx, CAS_E_IO, "cas_publish_files: cannot create the store directory");
    ensure_io_pool(ctx);
    for (uint64_t i = 0; i < n; ++i) status_out[i] = CAS_STORED_NEW;
    std::vector<uint64_t> len(n);
    int rc = stat_files(ctx, paths, n, len.data(), status_out);  // a missing file fails that file only
    if (rc) return rc;
    if (sizes_out) memcpy(sizes_out, len.data(), n * sizeof(uint64_t));
    ByteSource src;
    src.paths = paths;
    rc = hash_host_impl(ctx, src, len.data(), n, static_cast<uint8_t*>(out32n), status_out);
    if (rc) return rc;
    uint8_t* out = static_cast<uint8_t*>(out32n);
    for (uint64_t i = 0; i < n; ++i)
        if (status_out[i] < 0) memset(out + 32 * i, 0, 32);  // never a digest of bytes that were not all read
    if (dir) publish_paths(ctx, std::string(dir), paths, n, out, status_out);
    return CAS_OK;
}

int cas_verify_names(cas_ctx* ctx, const char* dir, const char* names, uint64_t n, uint8_t* result_out) {
    if (!ctx) return CAS_E_ARG;
    if (n == 0) return CAS_OK;
    if (!dir || !names || !result_out) return fail(ctx, CAS_E_ARG, "cas_verify_names: NULL directory, names or result array");
    if (n > 0xFFFFFFFFull) return fail(ctx, CAS_E_ARG, "cas_verify_names: more than 2^32-1 names in one call");
    CU(cudaSetDevice(ctx->device));
    const std::string root(dir);
    std::vector<const char*> name(n);
    std::vector<std::string> path(n);
    std::vector<const char*> cpath(n);
    const char* p = names;
    for (uint64_t i = 0; i < n; ++i) {
        name[i] = p;
        path[i] = root + "/" + p;
        cpath[i] = path[i].c_str();
        p += strlen(p) + 1;
    }
    std::vector<uint64_t> len(n);
    std::vector<int32_t> status(n, 0);
    int rc = stat_files(ctx, cpath.data(), n, len.data(), status.data());
    if (rc) return rc;
    // Only what can be an object is read: a regular file with a 64-character name.
    std::vector<uint32_t> pick;
    for (uint64_t i = 0; i < n; ++i) {
        if (status[i] < 0) {
            struct stat sb;
            result_out[i] = (stat(cpath[i], &sb) == 0 && !S_ISREG(sb.st_mode)) ? CAS_VERIFY_NOT_A_FILE : CAS_VERIFY_BAD;
        } else if (strlen(name[i]) != 64) {
            result_out[i] = CAS_VERIFY_BAD;  // cannot be the hex form of a SHA-256 digest
        } else {
            pick.push_back((uint32_t)i);
        }
    }
    const uint64_t m = pick.size();
    if (m == 0) return CAS_OK;
    std::vector<const char*> ppath(m);
    std::vector<uint64_t> plen(m);
    std::vector<int32_t> pstatus(m, 0);
    std::vector<uint8_t> dig(m * 32);
    for (uint64_t k = 0; k < m; ++k) { ppath[k] = cpath[pick[k]]; plen[k] = len[pick[k]]; }
    ByteSource src;
    src.paths = ppath.data();
    rc = hash_host_impl(ctx, src, plen.data(), m, dig.data(), pstatus.data());
    if (rc) return rc;
    char hex[65];
    for (uint64_t k = 0; k < m; ++k) {
        hex_name(hex, dig.data() + 32 * k);
        result_out[pick[k]] = (pstatus[k] >= 0 && memcmp(hex, name[pick[k]], 64) == 0) ? CAS_VERIFY_OK : CAS_VERIFY_BAD;
    }
    return CAS_OK;
}

}  // extern "C"

namespace {

// Copies `bytes` of the virtual input "a followed by b" starting at `pos` to dst.
void gather2(uint8_t* dst, const uint8_t* a, uint64_t alen, const uint8_t* b, uint64_t pos, uint64_t bytes) {
    if (pos < alen) {
        const uint64_t k = std::min(bytes, alen - pos);
        memcpy(dst, a + pos, k);
        dst += k; pos += k; bytes -= k;
    }
    if (bytes) cas::copy_to_staging(dst, b + (pos - alen), bytes);
}

// Advances n chains over whole 64-byte blocks.  The input of stream i is a[i] (alen[i] bytes)
// followed by b[i] (blen[i] bytes), alen + blen a multiple of 64.  Goes in rounds of at most one
// staging buffer; a stream's blocks may be split between rounds (that is what the state is for).
int stream_blocks(cas_ctx* ctx, cas_stream_state* const* s, const uint8_t* const* a, const uint64_t* alen,
                  const uint8_t* const* b, const uint64_t* blen, uint64_t n) {
    if (ctx->host_path_dirty) CU(cudaDeviceSynchronize());
    uint64_t total = 0;
    for (uint64_t i = 0; i < n; ++i) total += alen[i] + blen[i];
    if (total == 0) return CAS_OK;
    // One staging buffer's worth per round, and no more staging than this call needs.
    const uint64_t round_cap =
        std::max<uint64_t>(MiB, std::min<uint64_t>({ctx->wave_bytes, 256 * MiB, (total + MiB - 1) & ~(MiB - 1)})) & ~63ull;
    int rc = ensure_host_path(ctx, round_cap, round_cap, n);
    if (rc) return rc;
    ctx->host_path_dirty = true;
    std::vector<uint64_t> done(n, 0);  // bytes of each stream's input already hashed
    uint64_t first = 0;                // first stream that still has blocks
    struct Entry { uint64_t stream, stage_off, bytes; };
    std::vector<Entry> entries;
    cudaStream_t st = ctx->s_comp[0];
    for (;;) {
        while (first < n && done[first] == alen[first] + blen[first]) ++first;
        if (first >= n) break;
        entries.clear();
        uint64_t used = 0;
        for (uint64_t i = first; i < n && used < round_cap; ++i) {
            const uint64_t left = alen[i] + blen[i] - done[i];
            if (!left) continue;
            const uint64_t take = std::min(left, round_cap - used);
            entries.push_back({i, used, take});
            used += take;
        }
        const uint64_t m = entries.size();
        uint8_t* stage = ctx->stage[0].h;
        uint64_t* h_off = ctx->h_meta;
        uint64_t* h_nblk = ctx->h_meta + m;
        uint32_t* h_state = reinterpret_cast<uint32_t*>(ctx->h_dig);
        const Entry* ev = entries.data();
        ctx->pool->run([&](int t) {
            const int T = ctx->pool->size();
            for (uint64_t k = (uint64_t)t; k < m; k += (uint64_t)T) {
                const Entry& e = ev[k];
                gather2(stage + e.stage_off, a[e.stream], alen[e.stream], b[e.stream], done[e.stream], e.bytes);
            }
        });
        for (uint64_t k = 0; k < m; ++k) {
            h_off[k] = entries[k].stage_off;
            h_nblk[k] = entries[k].bytes / 64;
            memcpy(h_state + 8 * k, s[entries[k].stream]->h, 32);
        }
        CU(cudaMemcpyAsync(ctx->d_arena, stage, used, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(ctx->d_meta, h_off, m * 16, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(ctx->d_dig, h_state, m * 32, cudaMemcpyHostToDevice, st));
        cas::kernel_resume<3, 32><<<(unsigned)((m + 31) / 32), 32, 0, st>>>(
            ctx->d_arena, ctx->d_meta, ctx->d_meta + m, reinterpret_cast<uint32_t*>(ctx->d_dig), m, 1u);
        CU(cudaGetLastError());
        ++ctx->launches;
        CU(cudaMemcpyAsync(h_state, ctx->d_dig, m * 32, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        for (uint64_t k = 0; k < m; ++k) {
            memcpy(s[entries[k].stream]->h, h_state + 8 * k, 32);
            done[entries[k].stream] += entries[k].bytes;
        }
    }
    ctx->host_path_dirty = false;
    return CAS_OK;
}

}  // namespace

extern "C" {

void cas_stream_begin(cas_stream_state* s) {
    if (!s) return;
    static const uint32_t iv[8] = {0x6a09e667u, 0xbb67ae85u, 0x3c6ef372u, 0xa54ff53au,
                                   0x510e527fu, 0x9b05688cu, 0x1f83d9abu, 0x5be0cd19u};  // FIPS 180-4 section 5.3.3
    memcpy(s->h, iv, sizeof iv);
    s->length = 0;
    memset(s->tail, 0, sizeof s->tail);
}

int cas_stream_update(cas_ctx* ctx, cas_stream_state* const* streams, const void* const* data, const uint64_t* len,
                      uint64_t n) {
    if (!ctx) return CAS_E_ARG;
    if (n == 0) return CAS_OK;
    if (!streams || !data || !len) return fail(ctx, CAS_E_ARG, "cas_stream_update: NULL stream, data or length array");
    CU(cudaSetDevice(ctx->device));
    std::vector<const uint8_t*> a(n), b(n);
    std::vector<uint64_t> alen(n), blen(n);
    for (uint64_t i = 0; i < n; ++i) {
        if (!streams[i] || (len[i] && !data[i])) return fail(ctx, CAS_E_ARG, "cas_stream_update: NULL stream or data pointer");
        const uint64_t have = streams[i]->length & 63u;
        const uint64_t whole = (have + len[i]) & ~63ull;  // bytes that make whole blocks
        a[i] = streams[i]->tail;
        b[i] = static_cast<const uint8_t*>(data[i]);
        alen[i] = whole ? have : 0;
        blen[i] = whole ? whole - have : 0;
    }
    const int rc = stream_blocks(ctx, streams, a.data(), alen.data(), b.data(), blen.data(), n);
    if (rc) return rc;  // states may be part-way through this update: the streams are no longer usable
    for (uint64_t i = 0; i < n; ++i) {
        const uint64_t have = streams[i]->length & 63u;
        const uint8_t* src = static_cast<const uint8_t*>(data[i]);
        if (alen[i] + blen[i]) {  // the old tail was consumed; what is left of data becomes the new tail
            const uint64_t rest = len[i] - blen[i];
            memcpy(streams[i]->tail, src + blen[i], rest);
        } else {
            memcpy(streams[i]->tail + have, src, len[i]);
        }
        streams[i]->length += len[i];
    }
    return CAS_OK;
}

int cas_stream_finish(cas_ctx* ctx, cas_stream_state* const* streams, uint64_t n, void* out32n) {
    if (!ctx) return CAS_E_ARG;
    if (n == 0) return CAS_OK;
    if (!streams || !out32n) return fail(ctx, CAS_E_ARG, "cas_stream_finish: NULL stream or output array");
    CU(cudaSetDevice(ctx->device));
    // FIPS 180-4 section 5.1.1: 0x80, zeros, the bit length as a 64-bit big-endian number.
    std::vector<uint8_t> pad(n * 128, 0);
    std::vector<const uint8_t*> a(n), b(n);
    std::vector<uint64_t> alen(n), blen(n, 0);
    for (uint64_t i = 0; i < n; ++i) {
        if (!streams[i]) return fail(ctx, CAS_E_ARG, "cas_stream_finish: NULL stream");
        const uint64_t have = streams[i]->length & 63u;
        uint8_t* p = pad.data() + 128 * i;
        memcpy(p, streams[i]->tail, have);
        p[have] = 0x80;
        const uint64_t total = have + 9 <= 64 ? 64 : 128;
        const uint64_t bits = streams[i]->length << 3;
        for (int k = 0; k < 8; ++k) p[total - 1 - k] = (uint8_t)(bits >> (8 * k));
        a[i] = p;
        alen[i] = total;
        b[i] = nullptr;
    }
    const int rc = stream_blocks(ctx, streams, a.data(), alen.data(), b.data(), blen.data(), n);
    if (rc) return rc;
    uint8_t* out = static_cast<uint8_t*>(out32n);
    for (uint64_t i = 0; i < n; ++i)
        for (int w = 0; w < 8; ++w) {
            const uint32_t v = streams[i]->h[w];
            out[32 * i + 4 * w + 0] = (uint8_t)(v >> 24);
            out[32 * i + 4 * w + 1] = (uint8_t)(v >> 16);
            out[32 * i + 4 * w + 2] = (uint8_t)(v >> 8);
            out[32 * i + 4 * w + 3] = (uint8_t)v;
        }
    return CAS_OK;
}

// ---------------------------------------------------------------------------
// cas_hash_host_static: the host-buffer path for callers that must not allocate.
//
// cas_hash_host stages each wave in device buffers it gets from cudaMalloc and in
// pinned memory from cudaHostAlloc.  Some callers cannot afford that: a request
// handler on a latency budget (cudaMalloc synchronises the whole device and can
// stall every other stream), or a process whose device memory is already handed out
// to another allocator or capped by its operator.  For them the device staging lives in the
// file-scope array below: fixed size, part of the loaded module, held by every GPU
// that loads this library until the process exits.
//
// Layout of the array: [ packed object bytes | offsets | lengths | digests ].
// In host code `g_static_scratch` is only a stand-in symbol, not a device
// address; the address is looked up once per device with cudaGetSymbolAddress and
// that pointer is what the copies and the kernel get.
static __device__ __align__(256) uint8_t
    g_static_scratch[CAS_STATIC_MAX_BYTES + CAS_STATIC_MAX_OBJECTS * (8 + 8 + 32)];

namespace {
constexpr int STATIC_MAX_DEVICES = 64;
std::mutex g_static_mutex;                          // one call at a time through the shared array
uint8_t* g_static_ptr[STATIC_MAX_DEVICES] = {};     // device address of the array, per device

int static_fail(int code, const char* what, cudaError_t e = cudaSuccess) { return fail(nullptr, code, what, e); }
}  // namespace

int cas_hash_host_static(int device, const void* base, const uint64_t* off, const uint64_t* len, uint64_t n,
                         void* out32n, void* stream) {
    if (n == 0) return CAS_OK;
    if (!off || !len || !out32n) return static_fail(CAS_E_ARG, "cas_hash_host_static: NULL offset/length/output array");
    if (device < 0 || device >= STATIC_MAX_DEVICES) return static_fail(CAS_E_ARG, "cas_hash_host_static: bad device index");
    if (n > CAS_STATIC_MAX_OBJECTS) {
        char msg[160];
        snprintf(msg, sizeof msg, "cas_hash_host_static: %llu objects exceed the fixed scratch (at most %d per call)",
                 (unsigned long long)n, CAS_STATIC_MAX_OBJECTS);
        return static_fail(CAS_E_ARG, msg);
    }
    // Slot of every object inside the data region, 16-byte aligned (the kernel's fast path).
    uint64_t slot[CAS_STATIC_MAX_OBJECTS], length[CAS_STATIC_MAX_OBJECTS];
    uint64_t cursor = 0;
    for (uint64_t i = 0; i < n; ++i) {
        if (len[i] && !base && !off[i])
            return static_fail(CAS_E_ARG, "cas_hash_host_static: a non-empty object has a NULL address");
        if (len[i] > CAS_STATIC_MAX_BYTES || cursor + align16(len[i]) > CAS_STATIC_MAX_BYTES) {
            char msg[200];
            snprintf(msg, sizeof msg,
                     "cas_hash_host_static: batch needs more than the fixed %u-byte scratch "
                     "(objects are padded to 16 bytes); split it or use cas_hash_host",
                     (unsigned)CAS_STATIC_MAX_BYTES);
            return static_fail(CAS_E_ARG, msg);
        }
        slot[i] = cursor;
        length[i] = len[i];
        cursor += align16(len[i]);
    }
    std::lock_guard<std::mutex> lock(g_static_mutex);
    // The calling thread gets its current device back on every return path.
    struct DeviceGuard {
        int prev = -1;
        ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
    } guard;
    if (cudaGetDevice(&guard.prev) != cudaSuccess) { cudaGetLastError(); guard.prev = -1; }
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return static_fail(CAS_E_CUDA, "cas_hash_host_static: cudaSetDevice", e);
    if (guard.prev == device) guard.prev = -1;  // nothing to restore
    if (!g_static_ptr[device]) {
        // Once per device, with that device current.
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || !p) return static_fail(CAS_E_CUDA, "cas_hash_host_static: cudaGetSymbolAddress", e);
        g_static_ptr[device] = static_cast<uint8_t*>(p);
    }
    uint8_t* d_data = g_static_ptr[device];
    uint64_t* d_off = reinterpret_cast<uint64_t*>(d_data + CAS_STATIC_MAX_BYTES);
    uint64_t* d_len = d_off + CAS_STATIC_MAX_OBJECTS;
    uint8_t* d_dig = reinterpret_cast<uint8_t*>(d_len + CAS_STATIC_MAX_OBJECTS);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const uint8_t* hbase = static_cast<const uint8_t*>(base);

    // Once something is queued on the stream, a failure must not hand the shared array to
    // the next caller while this call's copies or kernel may still be running on it: every
    // error return below waits for the stream first (the result of that wait is not the
    // error being reported, so it is ignored).
    auto fail_drained = [&](const char* what, cudaError_t err) {
        cudaStreamSynchronize(st);
        cudaGetLastError();
        return static_fail(CAS_E_CUDA, what, err);
    };

    // The array keeps whatever the previous call left.  Object bytes, offsets and
    // lengths are all rewritten below before the kernel reads them; the digest
    // slots are the one thing that must not carry over, so they start at zero.
    e = cudaMemsetAsync(d_dig, 0, n * 32, st);
    if (e != cudaSuccess) return fail_drained("cas_hash_host_static: cudaMemsetAsync", e);
    for (uint64_t i = 0; i < n; ++i) {
        if (!length[i]) continue;
        e = cudaMemcpyAsync(d_data + slot[i], hbase + off[i], length[i], cudaMemcpyHostToDevice, st);
        if (e != cudaSuccess) return fail_drained("cas_hash_host_static: copy of an object to the device", e);
    }
    e = cudaMemcpyAsync(d_off, slot, n * 8, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_len, length, n * 8, cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return fail_drained("cas_hash_host_static: copy of offsets/lengths", e);
    cas::BatchArgs a{};
    a.base = d_data;
    a.off = d_off;
    a.len = d_len;
    a.n = n;
    a.out = d_dig;
    a.one = 1u;
    a.zero = 0u;
    cas::kernel_ldg<3, THREADS, 4><<<(unsigned)((n + THREADS - 1) / THREADS), THREADS, 0, st>>>(a);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail_drained("cas_hash_host_static: kernel launch", e);
    e = cudaMemcpyAsync(out32n, d_dig, n * 32, cudaMemcpyDeviceToHost, st);
    if (e != cudaSuccess) return fail_drained("cas_hash_host_static: digest copy", e);
    e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return static_fail(CAS_E_CUDA, "cas_hash_host_static: synchronize", e);
    return CAS_OK;
}

}  // extern "C"
