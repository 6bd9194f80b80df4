#!/usr/bin/env python
"""Kernel-variant sweep on one B200: load path x FMA level x register budget, plus the
integer-pipe probes.  Writes JSON lines to gpurun_out/sweep.jsonl (scratch) and prints a table.

    python tools/sweep.py [--objects N] [--object-bytes L] [--reps R]
"""
import argparse
import hashlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from bee_code_interpreter_fs_b200 import GpuHasher, _ffi  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--objects", type=int, default=148 * 1024 * 2)
    ap.add_argument("--object-bytes", type=int, default=65536)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--kernels", default="0,1,2")
    ap.add_argument("--fmas", default="0,1,2,3,4")
    ap.add_argument("--ctas", default="4,3,2")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "sweep.jsonl"))
    a = ap.parse_args()
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    h = GpuHasher(0)
    info = h.device_info()
    n, L = a.objects, a.object_bytes
    buf = torch.empty(n * L, dtype=torch.uint8, device="cuda:0")
    h.gen_synthetic(buf, 0xB200, 0, n, L)
    out = torch.empty((n, 32), dtype=torch.uint8, device="cuda:0")
    torch.cuda.synchronize()
    rows = []
    with open(a.out, "a") as f:
        for mode in range(7):
            ops, ms = h.int32_peak(mode, 8192)
            row = {"probe": mode, "lane_ops_per_s": ops, "ms": ms,
                   "lanes_per_clk_per_sm_at_max_clock": ops / (info["sm_count"] * info["sm_clock_khz"] * 1e3)}
            print(json.dumps(row))
            f.write(json.dumps(row) + "\n")
        ref = None
        nb = (L + 9 + 63) // 64
        for kernel in map(int, a.kernels.split(",")):
            for fma in map(int, a.fmas.split(",")):
                for ctas in map(int, a.ctas.split(",")):
                    h.set_option(_ffi.OPT_KERNEL, kernel)
                    h.set_option(_ffi.OPT_FMA_LEVEL, fma)
                    h.set_option(_ffi.OPT_CTAS_PER_SM, ctas)
                    h.digest_device_uniform(buf, L, L, n, out=out)
                    torch.cuda.synchronize()
                    best = 1e30
                    for _ in range(a.reps):
                        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        e0.record()
                        h.digest_device_uniform(buf, L, L, n, out=out)
                        e1.record()
                        torch.cuda.synchronize()
                        best = min(best, e0.elapsed_time(e1))
                    chk = hashlib.sha256(out.cpu().numpy().tobytes()).hexdigest()
                    ref = ref or chk
                    row = {"kernel": kernel, "fma": fma, "ctas": ctas, "objects": n, "object_bytes": L, "ms": best,
                           "gb_s": n * L / best / 1e6, "tinstr_s": n * nb * 1400 / best / 1e9, "same_digests": chk == ref}
                    rows.append(row)
                    print(json.dumps(row), flush=True)
                    f.write(json.dumps(row) + "\n")
    rows.sort(key=lambda r: -r["gb_s"])
    print("best:", rows[0])
    assert all(r["same_digests"] for r in rows), "variants disagree"


if __name__ == "__main__":
    main()
