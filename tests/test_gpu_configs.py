"""BASELINE.json configs 4 and 5 at full size, and a 2^30-byte object on every kernel, against
hashlib.sha256 -- every digest, not a sample.  These are the shapes the host pipeline and the
warp-per-object kernel exist for (one request's files = one batch:
/root/reference/src/code_interpreter/services/kubernetes_code_executor.py:149-154); round 1 only
measured them in tools/configs.py.
"""
import hashlib
import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest
import torch

from bee_code_interpreter_fs_b200 import _ffi

pytestmark = pytest.mark.gpu

SEED = 0xB200


def hashlib_mt(views):
    with ThreadPoolExecutor(os.cpu_count() or 1) as ex:
        return list(ex.map(lambda v: hashlib.sha256(v).digest(), views))


def test_config4_mixed_sizes_end_to_end_pinned_and_pageable_every_digest(hasher):
    """4 096 objects, sizes floor(2^u), u ~ U[10, 26) with numpy's default_rng(0xB200): exactly the
    batch tools/configs.py times.  Through cas_hash_host from pinned and from pageable memory."""
    rng = np.random.default_rng(SEED)
    lengths = np.floor(2.0 ** rng.uniform(10, 26, size=4096)).astype(np.uint64)
    off = np.concatenate([[0], np.cumsum((lengths + 15) // 16 * 16)[:-1]]).astype(np.uint64)
    span = int(off[-1] + lengths[-1])
    free, _ = torch.cuda.mem_get_info(0)
    if free < 2 * span + (8 << 30):
        pytest.skip("not enough free device memory for config 4")
    dev = torch.empty(span, dtype=torch.uint8, device="cuda:0")
    hasher.gen_synthetic(dev, SEED, 0, 1, span, span)
    pinned = torch.empty(span, dtype=torch.uint8, pin_memory=True)
    pinned.copy_(dev)
    torch.cuda.synchronize()
    del dev
    torch.cuda.empty_cache()
    hp = pinned.numpy()
    mv = memoryview(hp)
    want = np.frombuffer(b"".join(hashlib_mt([mv[int(o):int(o) + int(n)] for o, n in zip(off, lengths)])),
                         dtype=np.uint8).reshape(-1, 32)
    got = hasher.digest_packed(pinned, off, lengths)
    assert (got == want).all(), f"{int((got != want).any(axis=1).sum())} digests differ (pinned)"
    pageable = np.empty(span, dtype=np.uint8)
    pageable[:] = hp
    got = hasher.digest_packed(pageable, off, lengths)
    assert (got == want).all(), f"{int((got != want).any(axis=1).sum())} digests differ (pageable)"


def test_one_gib_object_on_every_kernel(hasher):
    """Bit length 2^33: the high word of the length field is exercised by the warp-per-object
    kernel, by the per-lane-load kernel and by the TMA-tile kernel (the latter two with the
    warp-per-object switch turned off, so they really run).  The three run side by side."""
    n = 1 << 30
    buf = torch.empty(n, dtype=torch.uint8, device="cuda:0")
    hasher.gen_synthetic(buf, 77, 3, 1, n, n)
    torch.cuda.synchronize()
    want = hashlib.sha256(buf.cpu().numpy()).hexdigest()
    saved = hasher.get_option(_ffi.OPT_LONG_MAX_OBJECTS), hasher.get_option(_ffi.OPT_KERNEL)
    streams = [torch.cuda.Stream() for _ in range(3)]
    outs = []
    try:
        outs.append(hasher.digest_device_uniform(buf, n, n, 1, stream=streams[0]))  # 1 object >= 32 KiB: kernel_long
        hasher.set_option(_ffi.OPT_LONG_MAX_OBJECTS, 0)
        launches = hasher.launch_count
        hasher.set_option(_ffi.OPT_KERNEL, 0)
        outs.append(hasher.digest_device_uniform(buf, n, n, 1, stream=streams[1]))  # kernel_ldg
        hasher.set_option(_ffi.OPT_KERNEL, 2)
        outs.append(hasher.digest_device_uniform(buf, n, n, 1, stream=streams[2]))  # kernel_tma
        assert hasher.launch_count == launches + 2
        torch.cuda.synchronize()
    finally:
        hasher.set_option(_ffi.OPT_LONG_MAX_OBJECTS, saved[0])
        hasher.set_option(_ffi.OPT_KERNEL, saved[1])
    for name, out in zip(("kernel_long", "kernel_ldg", "kernel_tma"), outs):
        assert bytes(out.cpu().numpy()[0]).hex() == want, name


def test_config5_64_objects_of_1_gib_every_digest(hasher):
    """64 x 1 GiB, one warp per object; all 64 digests against hashlib."""
    n, length = 64, 1 << 30
    free, _ = torch.cuda.mem_get_info(0)
    if free < n * length + (4 << 30):
        pytest.skip("not enough free device memory for 64 x 1 GiB")
    buf = torch.empty(n * length, dtype=torch.uint8, device="cuda:0")
    hasher.gen_synthetic(buf, SEED, 0, n, length)
    out = hasher.digest_device_uniform(buf, length, length, n)  # asynchronous: the CPU hashes meanwhile
    host = torch.empty(n * length, dtype=torch.uint8)
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        host.copy_(buf)
    side.synchronize()
    mv = memoryview(host.numpy())
    want = hashlib_mt([mv[i * length:(i + 1) * length] for i in range(n)])
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    for i in range(n):
        assert bytes(got[i]) == want[i], f"object {i}"
