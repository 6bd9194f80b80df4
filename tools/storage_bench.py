#!/usr/bin/env python
"""The Storage-shaped path on one B200 (BASELINE config 1 shape: 10 000 x 64 KiB objects):
write_many, gathered write(), and an integrity scrub of the store through cas_hash_files,
next to the same steps with hashlib on the host.  Writes gpurun_out/storage_bench.json."""
import asyncio
import hashlib
import json
import os
import shutil
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from bee_code_interpreter_fs_b200 import GpuHasher, Storage  # noqa: E402
from oracle import sha256_oracle as oracle  # noqa: E402


async def main():
    n, length = 10_000, 65536
    blob = oracle.synth_batch(0xB200, 0, n, length)
    objs = [blob[i * length:(i + 1) * length].tobytes() for i in range(n)]
    total = n * length
    h = GpuHasher(0)
    h.hash_many(objs)  # buffers reach their steady-state size
    base = tempfile.mkdtemp(prefix="b200cas-store-", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    res = {"objects": n, "object_bytes": length, "store_dir": base, "host_cores": os.cpu_count()}
    try:
        st = Storage(os.path.join(base, "a"), hasher=h)
        t0 = time.perf_counter()
        ids = await st.write_many(objs)
        dt = time.perf_counter() - t0
        res["write_many_gb_s"] = total / dt / 1e9
        res["write_many_objects_s"] = n / dt
        assert ids[7] == hashlib.sha256(objs[7]).hexdigest()

        st2 = Storage(os.path.join(base, "b"), hasher=h, batch_window=0.01, max_batch=4096)
        t0 = time.perf_counter()
        ids2 = await asyncio.gather(*(st2.write(o) for o in objs[:2000]))
        dt = time.perf_counter() - t0
        res["gathered_write_2000_gb_s"] = 2000 * length / dt / 1e9
        res["gathered_write_2000_objects_s"] = 2000 / dt
        res["gathered_write_gpu_batches"] = st2.gpu_batches
        assert list(ids2) == ids[:2000]

        t0 = time.perf_counter()
        scrub = await st.verify_all()
        dt = time.perf_counter() - t0
        assert len(scrub) == n and all(scrub.values())
        res["scrub_cas_hash_files_gb_s"] = total / dt / 1e9

        paths = [os.path.join(base, "a", i) for i in ids]
        t0 = time.perf_counter()
        for p, want in zip(paths, ids):
            with open(p, "rb") as f:
                assert hashlib.sha256(f.read()).hexdigest() == want
        dt = time.perf_counter() - t0
        res["scrub_hashlib_1_thread_gb_s"] = total / dt / 1e9

        def chunk(lo, hi):
            for p, want in zip(paths[lo:hi], ids[lo:hi]):
                with open(p, "rb") as f:
                    assert hashlib.sha256(f.read()).hexdigest() == want

        cores = os.cpu_count() or 1
        from concurrent.futures import ThreadPoolExecutor

        t0 = time.perf_counter()
        with ThreadPoolExecutor(cores) as ex:
            list(ex.map(lambda t: chunk(n * t // cores, n * (t + 1) // cores), range(cores)))
        dt = time.perf_counter() - t0
        res["scrub_hashlib_all_cores_gb_s"] = total / dt / 1e9
    finally:
        shutil.rmtree(base, ignore_errors=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "storage_bench.json"), "w") as f:
        json.dump(res, f, indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    asyncio.run(main())
