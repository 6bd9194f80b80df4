import sys, time, numpy as np, torch, hashlib
sys.path.insert(0, ".")
from bee_code_interpreter_fs_b200 import GpuHasher, _ffi
from oracle import sha256_oracle as oracle
h = GpuHasher(0)
n, L = 131072, 65536   # 8 GiB
dev = torch.empty(n * L, dtype=torch.uint8, device="cuda:0"); h.gen_synthetic(dev, 0xB200, 0, n, L)
pinned = torch.empty(n * L, dtype=torch.uint8, pin_memory=True); pinned.copy_(dev); torch.cuda.synchronize()
pageable = np.empty(n * L, dtype=np.uint8); pageable[:] = pinned.numpy()
off = np.arange(n, dtype=np.uint64) * L; lens = np.full(n, L, dtype=np.uint64)
want0 = hashlib.sha256(oracle.synth_object(0xB200, 0, L)).digest()
for name, src in (("pinned", pinned), ("pageable", pageable)):
    for threads in ((16,) if name == "pinned" else (4, 8, 16)):
        h.set_option(_ffi.OPT_PACK_THREADS, threads)
        d = h.digest_packed(src, off, lens)
        assert d[0].tobytes() == want0
        best = 1e9
        for _ in range(3):
            t0 = time.perf_counter(); h.digest_packed(src, off, lens); best = min(best, time.perf_counter() - t0)
        print(f"{name} pack_threads={threads}: {n*L/best/1e9:.1f} GB/s", flush=True)
# raw memcpy bandwidth of the host, 16 threads
import threading
dst = np.empty(n * L, dtype=np.uint8)
def cp(i, T): 
    lo, hi = n*L*i//T, n*L*(i+1)//T
    dst[lo:hi] = pageable[lo:hi]
for T in (1, 4, 8, 16):
    ts = [threading.Thread(target=cp, args=(i, T)) for i in range(T)]
    t0 = time.perf_counter(); [t.start() for t in ts]; [t.join() for t in ts]
    print(f"numpy memcpy {T} threads: {n*L/(time.perf_counter()-t0)/1e9:.1f} GB/s", flush=True)
