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
import shutil
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
            h.set_option(_ffi.OPT_LONG_MAX_OBJECTS, int(rng.choice([0, 1, 8, 592, 592, 100000])))
            h.set_option(_ffi.OPT_LONG_MAX_PER_WARP, int(rng.choice([1, 2, 4, 8, 16, 32, 32])))
            h.set_option(_ffi.OPT_LONG_MIN_BYTES, int(rng.choice([64, 4096, 32768])))
            h.set_option(_ffi.OPT_FIRST_CHUNK_BYTES, int(rng.choice([65536, 1 << 20, 2 << 20, 16 << 20])))
            h.set_option(_ffi.OPT_CHUNK_DIVISOR, int(rng.choice([1, 8, 16, 64])))
            h.set_option(_ffi.OPT_PACK_THREADS, int(rng.choice([1, 3, 8, 12])))
            h.set_option(_ffi.OPT_IO_THREADS, int(rng.choice([1, 5, 12, 16])))
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
                        "uniform", "files", "static", "store", "publish", "stream"][int(rng.integers(0, 12))]
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
            elif scenario == "store":
                m = min(n, 400)
                root = os.path.join(tmp, f"store{it}")
                got, status = h.store_many(root, blobs[:m])
                want = want[:m]
                assert all(s in (0, 1) for s in status.tolist()) and sorted(os.listdir(root)) == sorted(set(want))
                for name in list(set(want))[:20]:
                    with open(os.path.join(root, name), "rb") as f:
                        assert hashlib.sha256(f.read()).hexdigest() == name
                assert h.verify_names(root, got).tolist() == [1] * m
                shutil.rmtree(root)
            elif scenario == "publish":
                m = min(n, 300)
                root, spool = os.path.join(tmp, f"pub{it}"), os.path.join(tmp, f"spool{it}")
                os.makedirs(spool)
                paths = []
                for i in range(m):
                    paths.append(os.path.join(spool, f".tmp-{i}"))
                    with open(paths[-1], "wb") as f:
                        f.write(blobs[i])
                missing = int(rng.integers(0, m)) if m > 1 and rng.integers(0, 2) else -1
                if missing >= 0:
                    os.unlink(paths[missing])
                got, _, status = h.publish_files(root, paths)
                want = want[:m]
                if missing >= 0:
                    assert status[missing] < 0
                    got[missing] = want[missing]
                live = {w for i, w in enumerate(want) if i != missing}
                assert sorted(os.listdir(root)) == sorted(live) and os.listdir(spool) == []
                shutil.rmtree(root)
                shutil.rmtree(spool)
            elif scenario == "stream":
                m = min(n, 200)
                states = [h.stream_begin() for _ in range(m)]
                pos = [0] * m
                while any(pos[i] < len(blobs[i]) for i in range(m)):
                    step = int(rng.choice([1, 63, 64, 65, 1000, 100000, 1 << 22]))
                    pieces = [blobs[i][pos[i]:pos[i] + step] for i in range(m)]
                    for i in range(m):
                        pos[i] += len(pieces[i])
                    h.stream_update(states, pieces)
                got, want = h.stream_finish(states), want[:m]
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
