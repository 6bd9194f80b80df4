#!/usr/bin/env python
"""Strong-scaling shard sizes of BASELINE config 3 (2^20 / N objects of 64 KiB per GPU) against the
register budget / resident CTAs per SM: does a launch that fits one wave beat the default when
the default needs 1.15 waves (N = 8: 512 CTAs on 444 slots)?  One JSON line per setting."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from bee_code_interpreter_fs_b200 import GpuHasher, _ffi  # noqa: E402


def main():
    h = GpuHasher(0)
    length = 65536
    nmax = 1 << 19
    buf = torch.empty(nmax * length, dtype=torch.uint8, device="cuda:0")
    h.gen_synthetic(buf, 0xB200, 0, nmax, length)
    out = torch.empty((nmax, 32), dtype=torch.uint8, device="cuda:0")
    ref = {}
    for n in (1 << 17, 148 * 3 * 256, 1 << 18, 1 << 19):
        for ctas in (0, 2, 3, 4):
            h.set_option(_ffi.OPT_CTAS_PER_SM, ctas)
            for _ in range(2):
                h.digest_device_uniform(buf, length, length, n, out=out)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5):
                h.digest_device_uniform(buf, length, length, n, out=out)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 5
            chk = int(out[:n].to(torch.int64).sum().item())
            ref.setdefault(n, chk)
            assert ref[n] == chk, "digests differ between register budgets"
            print(json.dumps({"objects": n, "object_bytes": length, "ctas_per_sm_option": ctas, "ms": ms,
                              "gb_s": n * length / ms / 1e6, "ctas_256": -(-n // 256)}), flush=True)
    h.close()


if __name__ == "__main__":
    main()
