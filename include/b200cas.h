/*
 * b200cas.h -- C ABI of the B200 batched SHA-256 content-addressing engine.
 *
 * Drop-in boundary.  The reference (bee-code-interpreter-fs) has no FFI: its
 * object store is the Python class `Storage`
 * (/root/reference/src/code_interpreter/services/storage.py:34-101), built once
 * in ApplicationContext.file_storage (application_context.py:67-70).  The id of
 * an object is fixed in `Storage.writer` (storage.py:44-55, storage.py:52) and
 * returned by `Storage.write` (storage.py:57-63); the Execute call maps
 * path -> id for every changed file (services/kubernetes_code_executor.py:138-154).
 * This library is what a binding for that id-assignment step links against:
 * plain pointers and sizes, no torch or Python types.  Our own host code binds
 * it with ctypes (bee_code_interpreter_fs_b200/_ffi.py); INTEGRATION.md shows
 * the stub a maintainer of the reference would add.
 *
 * Conventions: every function returns 0 on success or a negative CAS_E_* code;
 * the message is available from cas_last_error().  A context belongs to one
 * device and must not be used from two host threads at once.  Digests are the
 * 32 raw bytes of SHA-256 (FIPS 180-4), bit-exact with hashlib.sha256; hex
 * encoding (the 64-character id of storage.py:52's shape) is done by the caller.
 * There is no CPU fallback: without a usable CUDA device every call fails.
 * Calls that take a context make that context's device the calling thread's current
 * CUDA device and leave it so (cudaSetDevice); cas_hash_host_static restores the caller's.
 */
#ifndef B200CAS_H
#define B200CAS_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CAS_OK 0
#define CAS_E_CUDA (-1)     /* a CUDA runtime call failed */
#define CAS_E_ARG (-2)      /* invalid argument */
#define CAS_E_NOMEM (-3)    /* host or device allocation failed */
#define CAS_E_NODEVICE (-4) /* no CUDA device / wrong architecture */
#define CAS_E_IO (-5)       /* a file could not be read (cas_hash_files) */

typedef struct cas_ctx cas_ctx;

/* Option keys for cas_set_option / cas_get_option. */
#define CAS_OPT_KERNEL 1       /* load path: 0 = per-lane read-only loads, 1 = per-lane TMA bulk copies,
                                  2 = TMA 2-D tiles for uniform layouts (per-lane loads otherwise) */
#define CAS_OPT_FMA_LEVEL 2    /* 0..4: how much of the arithmetic runs on the FMA pipe (sha256_sm100.cuh) */
#define CAS_OPT_WAVE_BYTES 3   /* host path: bytes per host->device copy chunk (rounded to 1 MiB) */
#define CAS_OPT_PACK_THREADS 4 /* host path: threads that pack objects into pinned staging */
#define CAS_OPT_CTAS_PER_SM 5  /* resident 256-thread CTAs per SM the kernels are compiled for: 4, 3 or 2
                                  (register budget 64, 80 or 128 per thread); 0 = best measured per load path */

#define CAS_OPT_ARENA_BYTES 6   /* host path: most device memory the staging arena may take (default 32 GiB) */

#define CAS_OPT_LONG_MAX_OBJECTS 7 /* batches of at most this many objects may run one warp per object (default 4 x SM count = 592; 0 = never) */
#define CAS_OPT_LONG_MIN_BYTES 8   /* ... when the objects are at least this long (default 32 KiB) */
#define CAS_OPT_LONG_MAX_PER_WARP 9 /* ... and up to this many objects share a warp (1, 2, 4 ... 32; default 32) when the
                                       batch or group has up to CAS_OPT_LONG_MAX_OBJECTS x this many objects */
#define CAS_OPT_FIRST_CHUNK_BYTES 10 /* host path: size of the first host->device copy chunk; chunks double from there
                                        up to CAS_OPT_WAVE_BYTES (default 2 MiB) */
#define CAS_OPT_IO_THREADS 11      /* threads that read, write and rename files (cas_hash_files, cas_store_host,
                                      cas_publish_files, cas_verify_names); default: every core, at most 32 */
#define CAS_OPT_CHUNK_DIVISOR 12   /* host path: a batch is cut into at least about this many copy chunks (default 8) */

/* Create / destroy the per-device context (streams, staging buffers are made
 * lazily).  Replaces nothing in the reference; it is the state a
 * `Storage(storage_path)` instance would own (storage.py:41-42). */
int cas_create(int device, cas_ctx** out);
void cas_destroy(cas_ctx* ctx);
/* Last error text of `ctx`; with ctx == NULL, of the calling thread's last failed call that had no
 * context (cas_create, cas_hash_host_static). */
const char* cas_last_error(const cas_ctx* ctx);
int cas_set_option(cas_ctx* ctx, int key, int64_t value);
int cas_get_option(const cas_ctx* ctx, int key, int64_t* value);
/* sm_count, max SM clock (kHz), total device memory (bytes). Any pointer may be NULL. */
int cas_device_info(const cas_ctx* ctx, int* sm_count, int* sm_clock_khz, uint64_t* mem_bytes);

/*
 * Id assignment for a batch of objects held in HOST memory: the arithmetic the
 * reference's docstring assigns to Storage.writer/Storage.write
 * (storage.py:36, :44-63), for the n objects base[off[i] .. off[i]+len[i]).
 * `base` may be NULL, in which case off[i] are absolute host addresses (how a
 * list of Python `bytes` is passed without copying).  Synchronous: host->device
 * copies (pinned staging, overlapped with the kernels), the kernels and the
 * device->host copy of the digests all happen inside the call.  Objects are
 * borrowed for the call only.  out32n receives n*32 bytes, object-major.
 */
int cas_hash_host(cas_ctx* ctx, const void* base, const uint64_t* off, const uint64_t* len, uint64_t n,
                  void* out32n);

/*
 * Ids of objects that live in FILES: the store directory itself (one file per object,
 * storage.py:53, :69-75) or a workspace about to be stored.  Same pipeline as
 * cas_hash_host, except that the pinned staging is filled by pread() from a few host
 * threads, so file bytes go disk/page cache -> pinned memory -> GPU without an
 * intermediate copy.  sizes_out (may be NULL) receives the file sizes.  A file that is
 * missing, not regular, or changes size during the call fails the call with CAS_E_IO.
 * Serves Storage.verify (an integrity scrub of the store) and bulk imports.
 */
int cas_hash_files(cas_ctx* ctx, const char* const* paths, uint64_t n, void* out32n, uint64_t* sizes_out);

/*
 * Pre-sizes the host path's grow-only buffers (device arena, pinned staging, metadata) for
 * batches of up to `batch_bytes` bytes and `max_objects` objects, so that the first real
 * call does not pay for cudaMalloc / cudaHostAlloc.  Optional; replaces nothing in the
 * reference (the state a Storage instance would set up in storage.py:41-42).
 */
int cas_reserve(cas_ctx* ctx, uint64_t batch_bytes, uint64_t max_objects);

/*
 * Optional limit on how many host->device copy chunks run at once across the contexts and
 * processes that name the same gate (a POSIX named semaphore, created with `slots` if it does
 * not exist; name must start with '/').  On a host whose aggregate host->device bandwidth
 * drops when every GPU copies at the same time (profiles/r02_h2d_matrix.jsonl: 220 GB/s with
 * four B200s copying, 186 GB/s with eight) the ranks that share an uplink take turns instead.
 * Applies to cas_hash_host's direct copies from page-locked memory; a slot is never waited for
 * longer than a second.  name == NULL removes the gate.  Whoever created the name removes it
 * with sem_unlink when the job is over.  No counterpart in the reference (it has no GPU path).
 */
int cas_set_h2d_gate(cas_ctx* ctx, const char* name, int slots);

/* Per-object results of cas_store_host / cas_publish_files (status_out[i]); negative = CAS_E_* for that object. */
#define CAS_STORED_NEW 0     /* the object now exists under its id */
#define CAS_STORED_EXISTED 1 /* an object with this id was in the store already; it was left untouched */

/*
 * Storage.write for a batch (storage.py:44-63: open "<dir>/<id>", write, close), with
 * content-derived ids: hashes the n host objects exactly as cas_hash_host does and, while the
 * GPU pipeline runs, writes each object to a temporary file under "<dir>/.tmp-<token>.<thread>/"
 * from several host threads; as soon as the digests are back every temporary file is given its
 * final name "<dir>/<64 hex digits>" (hard link, then the temporary name is dropped), or just
 * removed if an object of that id exists already.  A final name only ever appears complete, an
 * existing object is never replaced, and a failed call leaves no temporary file behind.  `dir`
 * is created if missing.  status_out (may be NULL) receives CAS_STORED_NEW, CAS_STORED_EXISTED
 * or CAS_E_IO per object; an object that could not be written does not fail the call, its id is
 * still returned.  No fsync, as in the reference.
 */
int cas_store_host(cas_ctx* ctx, const char* dir, const void* base, const uint64_t* off, const uint64_t* len,
                   uint64_t n, void* out32n, int32_t* status_out);

/*
 * The close of Storage.writer for a batch (storage.py:44-55 plus the executor's download leg,
 * services/kubernetes_code_executor.py:138-146): hashes n spooled files as cas_hash_files does
 * and gives each its final name "<dir>/<id>" (or removes it if that id exists).  Unlike cas_hash_files,
 * a file that is missing, not regular or shrinks while it is read fails only itself:
 * status_out[i] = CAS_E_IO, its digest slot is zeroed, and the other files are still hashed
 * and published.  With dir == NULL nothing is renamed (hash only, per-file status).
 */
int cas_publish_files(cas_ctx* ctx, const char* dir, const char* const* paths, uint64_t n, void* out32n,
                      uint64_t* sizes_out, int32_t* status_out);

/*
 * Integrity scrub of stored objects: the reference names a file by its id (storage.py:53), so
 * an object is intact when the SHA-256 of "<dir>/<name>" spells <name>.  `names` holds n
 * NUL-terminated names back to back; result_out[i] receives CAS_VERIFY_OK, CAS_VERIFY_BAD
 * (content does not match the name, the file is missing, unreadable or shrank while it was
 * read, or the name is not 64 characters) or CAS_VERIFY_NOT_A_FILE (a directory or other
 * non-regular entry, which a scrub of a whole directory should skip).  Files are read by the
 * host threads straight into pinned staging, as in cas_hash_files.
 */
#define CAS_VERIFY_BAD 0
#define CAS_VERIFY_OK 1
#define CAS_VERIFY_NOT_A_FILE 2
int cas_verify_names(cas_ctx* ctx, const char* dir, const char* names, uint64_t n, uint8_t* result_out);

/*
 * Streaming form: objects that arrive in pieces -- what the reference's writers do, chunk by
 * chunk (storage.py:44-55; the executor's download leg feeds `stored_file.write(chunk)`,
 * services/kubernetes_code_executor.py:139-145) -- or that are larger than device memory.
 * The running state of a stream is plain data owned by the caller: the 8 chaining words, the
 * byte count and the up to 63 bytes that do not fill a block yet.  cas_stream_update takes n
 * streams and one piece of host memory for each, runs all whole 64-byte blocks through the GPU
 * from each stream's own state (one launch for the n chains; pieces larger than the staging
 * buffer go in rounds, which is the same thing as more updates) and keeps the remainder in the
 * state.  cas_stream_finish appends the FIPS 180-4 padding as ordinary blocks, runs them on
 * the GPU too, and writes the n digests.  The same stream must not be in two calls at once.
 * The throughput path for whole objects is cas_hash_host; this one pays a device round trip per
 * update and exists for correctness with unbounded inputs, not for speed.
 */
typedef struct cas_stream_state {
    uint32_t h[8];     /* chaining value, host byte order */
    uint64_t length;   /* bytes absorbed so far */
    uint8_t tail[64];  /* the last length % 64 bytes, not yet compressed */
} cas_stream_state;
void cas_stream_begin(cas_stream_state* s); /* the FIPS 180-4 initial hash value, length 0 */
int cas_stream_update(cas_ctx* ctx, cas_stream_state* const* streams, const void* const* data, const uint64_t* len,
                      uint64_t n);
int cas_stream_finish(cas_ctx* ctx, cas_stream_state* const* streams, uint64_t n, void* out32n);

/*
 * Same arithmetic for objects already resident in DEVICE memory (kernel only,
 * asynchronous on `stream`, a cudaStream_t or NULL for the default stream).
 * d_off/d_len are device arrays; d_order is NULL or a device permutation: lane t
 * hashes object d_order[t] (a longest-first order keeps lanes of a warp the same
 * length).  d_out receives n*32 bytes.  Objects starting on a 16-byte boundary
 * take the fast path; any alignment is correct.
 */
int cas_hash_device(cas_ctx* ctx, const void* d_base, const uint64_t* d_off, const uint64_t* d_len,
                    const uint32_t* d_order, uint64_t n, void* d_out, void* stream);

/*
 * Same contract as cas_hash_device, but a WARP serves M = 1, 2, 4 ... 32 objects instead of 32:
 * the message schedules of 32 blocks are expanded in parallel by the warp's lanes and only the
 * 64 rounds of each object stay on its serial chain (csrc/sha256_long.cuh).  M = 1 is about
 * 1.7x faster per object than one lane per object; larger M trade some of that for more
 * objects per warp.  M is the smallest power of two that keeps the launch to about
 * CAS_OPT_LONG_MAX_OBJECTS warps (default: one per SM sub-partition, 592), capped by
 * CAS_OPT_LONG_MAX_PER_WARP.  For batches with so few objects that lanes would leave the GPU
 * empty -- BASELINE.json config 5 (64 x 1 GiB), the big files of config 4, a single large
 * upload, a few thousand files of a megabyte.  Digests are identical.
 */
int cas_hash_device_long(cas_ctx* ctx, const void* d_base, const uint64_t* d_off, const uint64_t* d_len,
                         const uint32_t* d_order, uint64_t n, void* d_out, void* stream);

/* Uniform layout: object i is d_base[i*stride .. i*stride+len). The shape of
 * BASELINE.json configs 2, 3 and 5 (1M x 4 KiB / 64 KiB, 64 x 1 GiB).  Small batches of
 * long objects (CAS_OPT_LONG_*) run on the warp kernel, 1 to 32 objects per warp. */
int cas_hash_device_uniform(cas_ctx* ctx, const void* d_base, uint64_t stride, uint64_t len, uint64_t n,
                            void* d_out, void* stream);

/* Synthetic benchmark objects generated on the device:
 * object first_obj+i at d_dst[i*stride .. i*stride+len), byte-identical to
 * oracle/sha256_oracle.py:synth_object(seed, first_obj+i, len). */
int cas_gen_synthetic(cas_ctx* ctx, void* d_dst, uint64_t seed, uint64_t first_obj, uint64_t n, uint64_t len,
                      uint64_t stride, void* stream);

/* Integer-pipe throughput probe (roofline denominator). mode: see csrc/peak.cuh.
 * Returns thread-level instructions per second and the kernel time. */
int cas_int32_peak(cas_ctx* ctx, int mode, int steps, double* lane_ops_per_s, double* ms);

/*
 * No counterpart in the reference: a variant of cas_hash_host that allocates nothing.
 *
 * cas_hash_host gets its device and pinned staging from cudaMalloc/cudaHostAlloc.  A caller
 * that cannot allocate -- a request handler that must not take cudaMalloc's device-wide
 * synchronisation, or a process whose GPU memory is fully owned by another allocator or
 * capped by its operator -- uses this entry point instead: the device staging is a fixed array inside the library
 * (CAS_STATIC_MAX_BYTES + 48 bytes per object slot, about 8.05 MiB), resident on every GPU
 * that loads the library until the process exits.  No context is needed.
 *
 * Same inputs and outputs as cas_hash_host (host memory; base may be NULL for absolute
 * addresses), plus the device index and a cudaStream_t (NULL = default stream).  The call
 * is synchronous.  Limits: at most CAS_STATIC_MAX_OBJECTS objects and CAS_STATIC_MAX_BYTES
 * bytes per call, counting each object rounded up to 16 bytes; larger batches are rejected
 * with CAS_E_ARG and a message saying so (cas_last_error(NULL)) -- split them, or use
 * cas_hash_host.  It calls no cudaMalloc, cudaFree, cudaHostAlloc or any other allocating
 * routine, creates no stream or event, and copies straight from the caller's memory.
 *
 * The first call on a device looks up the array's device address (cudaGetSymbolAddress);
 * make that first call before starting any stream capture.  All calls on a device share
 * the one array, which still holds the previous call's bytes: everything the kernel reads
 * is rewritten first and the digest slots are zeroed first.  Calls are serialised by a
 * lock inside the library, because two calls in flight on the same GPU would overwrite
 * each other's staging and race.
 *
 * `stream` must be a stream of `device` (or NULL).  The calling thread's current device is
 * the same after the call as before it.  If a copy or the launch fails part-way, the call
 * waits for whatever it had already queued on `stream` before it returns the error, so a
 * failed call never leaves work running on the array the next caller is about to reuse.
 * A non-empty object with a NULL address (base == NULL and off[i] == 0) is rejected.
 */
#define CAS_STATIC_MAX_OBJECTS 1024
#define CAS_STATIC_MAX_BYTES (8u << 20)
int cas_hash_host_static(int device, const void* base, const uint64_t* off, const uint64_t* len, uint64_t n,
                         void* out32n, void* stream);

/* Number of kernels this context has launched since creation. */
uint64_t cas_launch_count(const cas_ctx* ctx);

/* Library build tag, e.g. "b200cas 0.1 sm_100a". */
const char* cas_version(void);

#ifdef __cplusplus
}
#endif
#endif /* B200CAS_H */
