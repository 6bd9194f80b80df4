#!/usr/bin/env python
"""Runs the five BASELINE.json configs on one B200 and writes a JSON report.

    python tools/configs.py [--out gpurun_out/configs.json] [--only c1,c2,...]

c1  10k x 64 KiB through the Storage-shaped host API (hash_many on Python bytes) vs hashlib
c2  2^20 x 4 KiB on one GPU: kernel-only and end-to-end
c3  2^20 x 64 KiB: kernel-only (bench.py carries the full line for this one)
c4  4096 mixed-size objects, log-uniform 1 KiB..64 MiB: end-to-end from pageable and pinned memory
c5  64 x 1 GiB: kernel-only, where batching parallelism runs out
Every GPU result is checked against hashlib (full or sampled) before a number is recorded.
"""
import argparse
import hashlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

from bee_code_interpreter_fs_b200 import GpuHasher, _ffi  # noqa: E402
from oracle import sha256_oracle as oracle  # noqa: E402  (checker and CPU baseline only)

SEED = 0xB200


def timed_kernel(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


def check_sample(dig, seed, first, length, idx):
    for i in idx:
        want = hashlib.sha256(oracle.synth_object(seed, first + i, length)).digest()
        assert bytes(dig[i]) == want, f"digest mismatch at object {i}"


def uniform(h, n, length, e2e=True):
    buf = torch.empty(n * length, dtype=torch.uint8, device="cuda:0")
    h.gen_synthetic(buf, SEED, 0, n, length)
    out = torch.empty((n, 32), dtype=torch.uint8, device="cuda:0")
    ms = timed_kernel(lambda: h.digest_device_uniform(buf, length, length, n, out=out))
    dig = out.cpu().numpy()
    rng = np.random.default_rng(0)
    check_sample(dig, SEED, 0, length, [0, 1, n - 1] + rng.integers(0, n, size=256).tolist())
    nb = (length + 72) // 64
    res = {"objects": n, "object_bytes": length, "kernel_ms": ms, "kernel_gb_s": n * length / ms / 1e6,
           "kernel_mobj_s": n / ms / 1e3, "int32_tinstr_s": n * nb * 1400 / ms / 1e9}
    if e2e:
        host = torch.empty(n * length, dtype=torch.uint8, pin_memory=True)
        host.copy_(buf)
        torch.cuda.synchronize()
        off = np.arange(n, dtype=np.uint64) * length
        lens = np.full(n, length, dtype=np.uint64)
        h.digest_packed(host, off, lens)
        t0 = time.perf_counter()
        d2 = h.digest_packed(host, off, lens)
        dt = time.perf_counter() - t0
        assert (d2 == dig).all()
        res.update({"e2e_pinned_s": dt, "e2e_pinned_gb_s": n * length / dt / 1e9})
        cores = os.cpu_count() or 1
        k = min(n, max(1024, int(2e9 // length)))
        mv = memoryview(host.numpy())
        views = [mv[i * length:(i + 1) * length] for i in range(k)]
        t0 = time.perf_counter()
        oracle.hashlib_digests_mt(views, cores)
        dt = time.perf_counter() - t0
        res.update({"hashlib_all_cores_gb_s": k * length / dt / 1e9, "hashlib_cores": cores, "hashlib_sample_objects": k})
    return res


def c1(h):
    n, length = 10_000, 65536
    blob = oracle.synth_batch(SEED, 0, n, length)
    objs = [blob[i * length:(i + 1) * length].tobytes() for i in range(n)]
    t0 = time.perf_counter()
    h.reserve(n * length, n)  # what Storage(reserve_bytes=...) does at start-up: arena and pinned staging
    t_reserve = time.perf_counter() - t0
    t0 = time.perf_counter()
    h.hash_many(objs)  # first call of this shape, buffers already there
    t_cold = time.perf_counter() - t0
    t0 = time.perf_counter()
    ids = h.hash_many(objs)
    t_gpu = time.perf_counter() - t0
    t0 = time.perf_counter()
    want = [hashlib.sha256(o).hexdigest() for o in objs]
    t_cpu1 = time.perf_counter() - t0
    assert ids == want
    cores = os.cpu_count() or 1
    t0 = time.perf_counter()
    oracle.hashlib_digests_mt(objs, cores)
    t_cpun = time.perf_counter() - t0
    tot = n * length
    return {"objects": n, "object_bytes": length, "engine_hash_many_gb_s": tot / t_gpu / 1e9,
            "engine_hash_many_first_call_gb_s": tot / t_cold / 1e9, "reserve_s": t_reserve,
            "hashlib_1_thread_gb_s": tot / t_cpu1 / 1e9, "hashlib_all_cores_gb_s": tot / t_cpun / 1e9, "cores": cores,
            "note": "host API on a list of Python bytes (pageable memory): packed by host threads into pinned staging"}


def c4(h):
    rng = np.random.default_rng(SEED)
    lengths = np.floor(2.0 ** rng.uniform(10, 26, size=4096)).astype(np.uint64)
    total = int(lengths.sum())
    off = np.concatenate([[0], np.cumsum((lengths + 15) // 16 * 16)[:-1]]).astype(np.uint64)
    span = int(off[-1] + lengths[-1])
    dev = torch.empty(span, dtype=torch.uint8, device="cuda:0")
    h.gen_synthetic(dev, SEED, 0, 1, span, span)  # one long pseudo-random stream, cut into objects
    pinned = torch.empty(span, dtype=torch.uint8, pin_memory=True)
    pinned.copy_(dev)
    torch.cuda.synchronize()
    res = {"objects": 4096, "total_bytes": total, "max_object": int(lengths.max()), "min_object": int(lengths.min())}
    # kernel-only, longest-first schedule
    d_off = torch.from_numpy(off.astype(np.int64)).cuda()
    d_len = torch.from_numpy(lengths.astype(np.int64)).cuda()
    order = torch.from_numpy(np.argsort(-lengths.astype(np.int64), kind="stable").astype(np.int32)).cuda()
    out = torch.empty((4096, 32), dtype=torch.uint8, device="cuda:0")
    ms = timed_kernel(lambda: h.digest_device(dev, d_off, d_len, order=order, out=out), reps=2)
    res.update({"kernel_ms_lpt": ms, "kernel_gb_s_lpt": total / ms / 1e6})
    ms2 = timed_kernel(lambda: h.digest_device(dev, d_off, d_len, out=out), reps=1)
    res.update({"kernel_ms_unsorted": ms2, "kernel_gb_s_unsorted": total / ms2 / 1e6})
    # the hybrid the host path uses: the longest objects one warp each, the rest one lane each
    srt = np.argsort(-lengths.astype(np.int64), kind="stable")
    n_long = int((lengths >= max(32768, int(lengths.max()) // 4)).sum())
    o_long = torch.from_numpy(srt[:n_long].astype(np.int32)).cuda()
    o_rest = torch.from_numpy(srt[n_long:].astype(np.int32)).cuda()
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    out2 = torch.empty((4096, 32), dtype=torch.uint8, device="cuda:0")

    def hybrid():
        # digest_device writes object-major via `order`, so both launches can share `out2`
        s1.wait_stream(torch.cuda.current_stream())
        s2.wait_stream(torch.cuda.current_stream())
        h._check(h._lib.cas_hash_device_long(h._ctx, dev.data_ptr(), d_off.data_ptr(), d_len.data_ptr(), o_long.data_ptr(),
                                             n_long, out2.data_ptr(), s1.cuda_stream))
        h._check(h._lib.cas_hash_device(h._ctx, dev.data_ptr(), d_off.data_ptr(), d_len.data_ptr(), o_rest.data_ptr(),
                                        4096 - n_long, out2.data_ptr(), s2.cuda_stream))
        torch.cuda.current_stream().wait_stream(s1)
        torch.cuda.current_stream().wait_stream(s2)

    ms3 = timed_kernel(hybrid, reps=2)
    res.update({"kernel_ms_hybrid": ms3, "kernel_gb_s_hybrid": total / ms3 / 1e6, "hybrid_warp_objects": n_long})
    assert (out2.cpu().numpy() == out.cpu().numpy()).all()
    dig = out.cpu().numpy()
    hp = pinned.numpy()
    for i in rng.integers(0, 4096, size=64).tolist() + [int(np.argmax(lengths))]:
        assert bytes(dig[i]) == hashlib.sha256(hp[int(off[i]):int(off[i]) + int(lengths[i])].tobytes()).digest()
    # end-to-end from pinned memory
    h.digest_packed(pinned, off, lengths)
    t0 = time.perf_counter()
    d2 = h.digest_packed(pinned, off, lengths)
    dt = time.perf_counter() - t0
    assert (d2 == dig).all()
    res.update({"e2e_pinned_s": dt, "e2e_pinned_gb_s": total / dt / 1e9})
    # end-to-end from pageable memory
    pageable = np.empty(span, dtype=np.uint8)
    pageable[:] = hp
    h.digest_packed(pageable, off, lengths)
    t0 = time.perf_counter()
    d3 = h.digest_packed(pageable, off, lengths)
    dt = time.perf_counter() - t0
    assert (d3 == dig).all()
    res.update({"e2e_pageable_s": dt, "e2e_pageable_gb_s": total / dt / 1e9})
    cores = os.cpu_count() or 1
    sel = np.argsort(-lengths.astype(np.int64))[::8][:256]
    views = [memoryview(hp)[int(off[i]):int(off[i]) + int(lengths[i])] for i in sel]
    t0 = time.perf_counter()
    oracle.hashlib_digests_mt(views, cores)
    dt = time.perf_counter() - t0
    res.update({"hashlib_all_cores_gb_s": sum(len(v) for v in views) / dt / 1e9, "hashlib_cores": cores})
    return res


def c5(h):
    n, length = 64, 1 << 30
    buf = torch.empty(n * length, dtype=torch.uint8, device="cuda:0")
    h.gen_synthetic(buf, SEED, 0, n, length)
    out = torch.empty((n, 32), dtype=torch.uint8, device="cuda:0")
    res = {"objects": n, "object_bytes": length}
    first = buf[:length].cpu().numpy()
    t0 = time.perf_counter()
    want_first = hashlib.sha256(first).digest()
    res["hashlib_1_thread_gb_s"] = length / (time.perf_counter() - t0) / 1e9
    want_last = hashlib.sha256(buf[(n - 1) * length:].cpu().numpy()).digest()
    del first
    sums = []
    for kernel, name in ((2, "warp_per_object"), (0, "lane_per_object")):
        h.set_option(_ffi.OPT_LONG_MAX_OBJECTS, 592 if kernel == 2 else 0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        h.digest_device_uniform(buf, length, length, n, out=out)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        res[f"kernel_ms_{name}"] = ms
        res[f"kernel_gb_s_{name}"] = n * length / ms / 1e6
        res[f"per_chain_gb_s_{name}"] = length / ms / 1e6
        dig = out.cpu().numpy()  # this kernel's own output, checked before the next one overwrites it
        assert bytes(dig[0]) == want_first and bytes(dig[n - 1]) == want_last, name
        sums.append(dig.tobytes())
        out.zero_()
    assert sums[0] == sums[1], "warp-per-object and lane-per-object digests differ"
    h.set_option(_ffi.OPT_LONG_MAX_OBJECTS, 592)
    res["active_chains"] = n
    res["note"] = ("64 serial chains: one warp each (schedule expanded in parallel, rounds serial) on 64 of 592 SM "
                   "sub-partitions; SHA-256 cannot be split inside a message without changing the digest, so a host "
                   "with many SHA-NI cores wins this config")
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "configs.json"))
    ap.add_argument("--only", default="c1,c2,c3,c4,c5")
    a = ap.parse_args()
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    h = GpuHasher(0)
    report = {"gpu": torch.cuda.get_device_name(0), "host_cores": os.cpu_count(), "seed": SEED}
    for name in a.only.split(","):
        t0 = time.perf_counter()
        if name == "c1":
            report[name] = c1(h)
        elif name == "c2":
            report[name] = uniform(h, 1 << 20, 4096)
        elif name == "c3":
            report[name] = uniform(h, 1 << 20, 65536, e2e=False)
        elif name == "c4":
            report[name] = c4(h)
        elif name == "c5":
            report[name] = c5(h)
        torch.cuda.empty_cache()
        print(name, json.dumps(report[name]), f"({time.perf_counter() - t0:.1f}s)", flush=True)
        with open(a.out, "w") as f:
            json.dump(report, f, indent=1)
    h.close()


if __name__ == "__main__":
    main()
