#!/usr/bin/env python
"""Sweeps the host pipeline's knobs on mid-size batches of Python bytes (BASELINE config 1's shape
and the 4 096 x 64 KiB point of the latency table), and the store / scrub paths over thread counts.
One JSON line per setting on stdout."""
import asyncio
import json
import os
import shutil
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from bee_code_interpreter_fs_b200 import GpuHasher, Storage, _ffi  # noqa: E402
from oracle import sha256_oracle as oracle  # noqa: E402


def best(fn, reps=5):
    fn()
    t = 1e30
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        t = min(t, time.perf_counter() - t0)
    return t


def main():
    h = GpuHasher(0)
    length = 65536
    blob = oracle.synth_batch(0xB200, 0, 10000, length)
    objs = [blob[i * length:(i + 1) * length].tobytes() for i in range(10000)]
    for n in (10000, 4096):
        sub = objs[:n]
        for pack in (8, 12, 16):
            for div in (8, 16):
                for first in (2,):
                    h.set_option(_ffi.OPT_PACK_THREADS, pack)
                    h.set_option(_ffi.OPT_CHUNK_DIVISOR, div)
                    h.set_option(_ffi.OPT_FIRST_CHUNK_BYTES, first << 20)
                    t = best(lambda: h.hash_many(sub), reps=4)
                    print(json.dumps({"what": "hash_many", "objects": n, "object_bytes": length, "pack_threads": pack,
                                      "chunk_divisor": div, "first_chunk_mib": first, "ms": t * 1e3,
                                      "gb_s": n * length / t / 1e9}), flush=True)
    h.set_option(_ffi.OPT_PACK_THREADS, 8)
    h.set_option(_ffi.OPT_CHUNK_DIVISOR, 8)
    h.set_option(_ffi.OPT_FIRST_CHUNK_BYTES, 2 << 20)
    base = tempfile.mkdtemp(prefix="b200cas-sweep-", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    try:
        for io in (4, 8, 12, 16):
            h.set_option(_ffi.OPT_IO_THREADS, io)
            root = os.path.join(base, f"s{io}")
            t0 = time.perf_counter()
            ids, status = h.store_many(root, objs)
            t_store = time.perf_counter() - t0
            assert status.tolist() == [0] * len(objs)
            t_scrub = best(lambda: h.verify_names(root, ids), reps=3)
            st = Storage(root, hasher=h)
            t_all = best(lambda: asyncio.run(st.verify_all()), reps=2)
            print(json.dumps({"what": "store_and_scrub", "objects": len(objs), "object_bytes": length, "io_threads": io,
                              "store_many_ms": t_store * 1e3, "store_many_gb_s": len(objs) * length / t_store / 1e9,
                              "verify_names_ms": t_scrub * 1e3, "verify_names_gb_s": len(objs) * length / t_scrub / 1e9,
                              "storage_verify_all_gb_s": len(objs) * length / t_all / 1e9}), flush=True)
            shutil.rmtree(root, ignore_errors=True)
    finally:
        shutil.rmtree(base, ignore_errors=True)
    h.close()


if __name__ == "__main__":
    main()
