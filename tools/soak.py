#!/usr/bin/env python
"""Differential soak: random batches through every entry point, every digest compared with
hashlib.  Shapes, alignments, chunk and arena sizes are randomised so that the host
pipeline's chunk/epoch boundaries and the kernels' tail paths are hit from all sides.

    python tools/soak.py --seconds 180 [--seed 1]      ->  gpurun_out/soak.json
"""
import argparse
import hashlib
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

from bee_code_interpreter_fs_b200 import GpuHasher, _ffi  # noqa: E402
from bee_code_interpreter_fs_b200.static_path import STATIC_MAX_BYTES, STATIC_MAX_OBJECTS, hash_many_static  # noqa: E402


def random_lengths(rng, n):
    kind = rng.integers(0, 5)
    if kind == 0:
        return rng.integers(0, 200, size=n)
    if kind == 1:
        return np.floor(2.0 ** rng.uniform(0, 22, size=n)).astype(np.int64)
    if kind == 2:
        base = int(rng.integers(1, 1 << 16))
        return np.full(n, base) + rng.integers(-1, 2, size=n).clip(min=-base)
    if kind == 3:
        return rng.choice([0, 1, 55, 56, 63, 64, 65, 119, 120, 127, 128, 4095, 4096, 4097, 65535, 65536, 65537], size=n)
    return rng.integers(0, 1 << 14, size=n)


def want_hex(blobs):
    return [hashlib.sha256(b).hexdigest() for b in blobs]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=120)
    ap.add_argument("--seed", type=int, default=1)
    a = ap.parse_args()
    rng = np.random.default_rng(a.seed)
    h = GpuHasher(0)
    counts = {}
    objects = 0
    nbytes = 0
    t_end = time.time() + a.seconds
    it = 0
    with tempfile.TemporaryDirectory(dir="/dev/shm" if os.path.isdir("/dev/shm") else None) as tmp:
        while time.time() < t_end:
            it += 1
            h.set_option(_ffi.OPT_WAVE_BYTES, int(rng.choice([1 << 20, 4 << 20, 64 << 20, 256 << 20])))
            h.set_option(_ffi.OPT_ARENA_BYTES, int(rng.choice([64 << 20, 256 << 20, 32 << 30])))
            h.set_option(_ffi.OPT_KERNEL, int(rng.integers(0, 3)))
            h.set_option(_ffi.OPT_FMA_LEVEL, int(rng.choice([0, 1, 2, 3, 3, 3, 4])))
            h.set_option(_ffi.OPT_CTAS_PER_SM, int(rng.choice([0, 0, 2, 3, 4])))
            h.set_option(_ffi.OPT_LONG_MAX_OBJECTS, int(rng.choice([0, 592, 592, 100000])))
            n = int(rng.choice([1, 2, 7, 33, 200, 1000, 5000]))
            lengths = random_lengths(rng, n).astype(np.int64)
            total = int(lengths.sum())
            if total > (1 << 30):
                lengths = lengths // 8
                total = int(lengths.sum())
            data = rng.integers(0, 256, size=max(total, 1) + 64, dtype=np.uint8)
            off = np.concatenate([[0], np.cumsum(lengths)[:-1]]).astype(np.int64)
            blobs = [data[int(o):int(o) + int(k)].tobytes() for o, k in zip(off, lengths)]
            want = want_hex(blobs)
            scenario = ["bytes", "packed_pageable", "packed_pinned", "device", "device_order", "device_warp",
                        "uniform", "files", "static"][int(rng.integers(0, 9))]
            if scenario == "bytes":
                got = h.hash_many(blobs)
            elif scenario == "packed_pageable":
                got = [bytes(r).hex() for r in h.digest_packed(data, off.astype(np.uint64), lengths.astype(np.uint64))]
            elif scenario == "packed_pinned":
                pinned = torch.from_numpy(data).pin_memory()
                got = [bytes(r).hex() for r in h.digest_packed(pinned, off.astype(np.uint64), lengths.astype(np.uint64))]
            elif scenario in ("device", "device_order", "device_warp"):
                lead = int(rng.integers(0, 17))
                d = torch.from_numpy(np.concatenate([np.zeros(lead, np.uint8), data])).cuda()
                d_off = torch.from_numpy(off + lead).cuda()
                d_len = torch.from_numpy(lengths).cuda()
                order = None
                if scenario == "device_order":
                    order = torch.from_numpy(np.argsort(-lengths, kind="stable").astype(np.int32)).cuda()
                got = [bytes(r).hex() for r in h.digest_device(d, d_off, d_len, order=order,
                                                               warp_per_object=(scenario == "device_warp" and n <= 1000)).cpu().numpy()]
            elif scenario == "uniform":
                length = int(lengths[0])
                stride = length + int(rng.choice([0, 0, 1, 15, 16, 48])) if n > 1 else max(length, 1)
                stride = max(stride, 1)
                m = min(n, max(1, (256 << 20) // max(stride, 1)))
                buf = rng.integers(0, 256, size=m * stride + 32, dtype=np.uint8)
                d = torch.from_numpy(buf).cuda()
                got = [bytes(r).hex() for r in h.digest_device_uniform(d, stride, length, m).cpu().numpy()]
                want = want_hex([buf[i * stride:i * stride + length].tobytes() for i in range(m)])
            elif scenario == "files":
                m = min(n, 300)
                paths = []
                for i in range(m):
                    p = os.path.join(tmp, f"f{i}")
                    with open(p, "wb") as f:
                        f.write(blobs[i])
                    paths.append(p)
                got, want = h.hash_files(paths), want[:m]
            else:  # static
                keep, used = [], 0
                for b in blobs[:STATIC_MAX_OBJECTS]:
                    need = (len(b) + 15) // 16 * 16
                    if used + need > STATIC_MAX_BYTES:
                        break
                    keep.append(b)
                    used += need
                got, want = hash_many_static(keep), want_hex(keep)
            if got != want:
                bad = next(i for i, (g, w) in enumerate(zip(got, want)) if g != w)
                print(json.dumps({"FAILED": scenario, "iteration": it, "seed": a.seed, "object": bad,
                                  "length": int(lengths[bad]) if bad < len(lengths) else None}), flush=True)
                sys.exit(1)
            counts[scenario] = counts.get(scenario, 0) + 1
            objects += len(want)
            nbytes += total
    res = {"seconds": a.seconds, "seed": a.seed, "iterations": it, "objects": objects, "bytes": nbytes,
           "scenarios": counts, "mismatches": 0}
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", f"soak_seed{a.seed}.json"), "w") as f:
        json.dump(res, f, indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
