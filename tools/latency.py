#!/usr/bin/env python
"""Latency of small and medium batches: where the GPU path starts to pay off.

For each (objects, size): kernel-only time (device-resident), end-to-end time of hash_many on
Python bytes, and hashlib on one thread / all cores.  Writes gpurun_out/latency.json.
"""
import hashlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from bee_code_interpreter_fs_b200 import GpuHasher  # noqa: E402
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
    cores = os.cpu_count() or 1
    rows = []
    for length in (4096, 65536, 1 << 20):
        for n in (1, 8, 64, 592, 4096, 18944, 37888):
            if n * length > (8 << 30):
                continue
            blob = oracle.synth_batch(1, 0, n, length)
            objs = [blob[i * length:(i + 1) * length].tobytes() for i in range(n)]
            dev = torch.from_numpy(blob).cuda()
            out = torch.empty((n, 32), dtype=torch.uint8, device="cuda:0")

            def kernel():
                h.digest_device_uniform(dev, length, length, n, out=out)
                torch.cuda.synchronize()

            t_kernel = best(kernel)
            ids = h.hash_many(objs)
            assert ids[0] == hashlib.sha256(objs[0]).hexdigest() and ids[-1] == hashlib.sha256(objs[-1]).hexdigest()
            t_e2e = best(lambda: h.hash_many(objs), reps=3)
            t_cpu1 = best(lambda: [hashlib.sha256(o).digest() for o in objs], reps=2)
            t_cpun = best(lambda: oracle.hashlib_digests_mt(objs, cores), reps=2)
            row = {"objects": n, "object_bytes": length, "kernel_ms": t_kernel * 1e3, "hash_many_ms": t_e2e * 1e3,
                   "hashlib_1_thread_ms": t_cpu1 * 1e3, f"hashlib_{cores}_threads_ms": t_cpun * 1e3,
                   "hash_many_gb_s": n * length / t_e2e / 1e9, "kernel_gb_s": n * length / t_kernel / 1e9}
            rows.append(row)
            print(json.dumps(row), flush=True)
            del dev, out
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "latency.json"), "w") as f:
        json.dump({"host_cores": cores, "rows": rows}, f, indent=1)


if __name__ == "__main__":
    main()
