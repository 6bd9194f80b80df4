"""cas_hash_host_static: the entry point for callers that cannot allocate GPU memory.

Some users of this library must not let it call cudaMalloc/cudaHostAlloc for them -- a
request handler that cannot take cudaMalloc's device-wide synchronisation, a process whose
GPU memory is already owned by another allocator or capped by its operator.  This test
calls the entry point through the C ABI and checks it, bit for bit, against the regular
host path (cas_hash_host) and against hashlib.sha256 on the same objects, twice in a row:
the second call finds the first call's bytes still in the shared scratch array and must not
be affected by them.
Collected last on purpose (file name), and self-contained so that it also passes alone in
a fresh process.
"""
import ctypes
import hashlib

import numpy as np
import pytest

from bee_code_interpreter_fs_b200 import GpuHasher, _ffi
from bee_code_interpreter_fs_b200.static_path import STATIC_MAX_BYTES, STATIC_MAX_OBJECTS, hash_many_static

pytestmark = pytest.mark.gpu


def _batch(seed: int, n: int, max_len: int) -> list[bytes]:
    rng = np.random.default_rng(seed)
    lengths = rng.integers(0, max_len, size=n).tolist()
    lengths[:6] = [0, 1, 55, 56, 64, 119][: min(6, n)]
    return [rng.integers(0, 256, size=k, dtype=np.uint8).tobytes() for k in lengths]


def _static_digests(lib, objs, device=0):
    """Straight through the C ABI: absolute host addresses, no context, default stream."""
    n = len(objs)
    ptrs = (ctypes.c_char_p * n)(*objs)
    off = np.frombuffer(ptrs, dtype=np.uint64, count=n).copy()
    length = np.fromiter(map(len, objs), dtype=np.uint64, count=n)
    out = np.full((n, 32), 0xAB, dtype=np.uint8)
    rc = lib.cas_hash_host_static(device, None, off.ctypes.data, length.ctypes.data, n, out.ctypes.data, None)
    return rc, out


def _hashlib_digests(objs):
    return np.frombuffer(b"".join(hashlib.sha256(o).digest() for o in objs), dtype=np.uint8).reshape(-1, 32)


def test_static_entry_point_matches_regular_host_path_twice():
    lib = _ffi.load()
    with GpuHasher(0) as h:
        # first call: a full-width batch that leaves the scratch array well used
        first = _batch(1, 700, 9000)
        rc, got = _static_digests(lib, first)
        assert rc == _ffi.CAS_OK, lib.cas_last_error(None)
        assert (got == h.digest_many(first)).all()
        assert (got == _hashlib_digests(first)).all()
        # second call, straight after: fewer, shorter, different objects over the stale bytes
        second = _batch(2, 300, 3000)
        rc, got = _static_digests(lib, second)
        assert rc == _ffi.CAS_OK, lib.cas_last_error(None)
        assert (got == h.digest_many(second)).all()
        assert (got == _hashlib_digests(second)).all()
        # and the Python helper gives the same ids as the regular path
        assert hash_many_static(second) == h.hash_many(second) == [hashlib.sha256(o).hexdigest() for o in second]


def test_static_entry_point_rejects_what_does_not_fit():
    lib = _ffi.load()
    too_many = [b"x"] * (STATIC_MAX_OBJECTS + 1)
    rc, _ = _static_digests(lib, too_many)
    assert rc == _ffi.CAS_E_ARG and b"at most" in lib.cas_last_error(None)
    too_big = [bytes(STATIC_MAX_BYTES // 2), bytes(STATIC_MAX_BYTES // 2), b"one more byte"]
    rc, _ = _static_digests(lib, too_big)
    assert rc == _ffi.CAS_E_ARG and b"scratch" in lib.cas_last_error(None)
    # exactly at the limit still works, and still agrees with the regular path
    edge = [bytes(STATIC_MAX_BYTES // 2), np.arange(STATIC_MAX_BYTES // 2, dtype=np.uint8).tobytes()]
    rc, got = _static_digests(lib, edge)
    assert rc == _ffi.CAS_OK, lib.cas_last_error(None)
    with GpuHasher(0) as h:
        assert (got == h.digest_many(edge)).all()
    assert (got == _hashlib_digests(edge)).all()


def test_static_entry_point_rejections_leave_it_usable_and_the_device_alone():
    """Rejected calls (NULL arrays, a NULL object address) must not disturb the next call, and the
    calling thread's current device must be what it was.  (The drain-before-return on a CUDA error
    after work was queued cannot be provoked from outside on a healthy one-GPU box -- a device
    pointer as output is accepted by the runtime under unified addressing -- so it is covered by
    reading csrc/engine.cu, not by this test.)"""
    import torch

    lib = _ffi.load()
    objs = _batch(3, 500, 6000)
    n = len(objs)
    length = np.fromiter(map(len, objs), dtype=np.uint64, count=n)
    out = np.zeros((n, 32), dtype=np.uint8)
    before = torch.cuda.current_device()
    rc = lib.cas_hash_host_static(0, None, None, length.ctypes.data, n, out.ctypes.data, None)
    assert rc == _ffi.CAS_E_ARG
    null_src = np.zeros(1, dtype=np.uint64)
    one_len = np.ones(1, dtype=np.uint64)
    rc = lib.cas_hash_host_static(0, None, null_src.ctypes.data, one_len.ctypes.data, 1, out.ctypes.data, None)
    assert rc == _ffi.CAS_E_ARG and b"NULL address" in lib.cas_last_error(None)
    rc = lib.cas_hash_host_static(4096, None, null_src.ctypes.data, one_len.ctypes.data, 1, out.ctypes.data, None)
    assert rc == _ffi.CAS_E_ARG
    if torch.cuda.device_count() > 1:  # the call runs on device 1 and hands device 0 back
        rc, got = _static_digests(lib, objs, device=1)
        assert rc == _ffi.CAS_OK, lib.cas_last_error(None)
        assert (got == _hashlib_digests(objs)).all()
    assert torch.cuda.current_device() == before
    # straight after the rejections: a different batch over the stale bytes, exact
    again = _batch(4, 400, 5000)
    rc, got = _static_digests(lib, again)
    assert rc == _ffi.CAS_OK, lib.cas_last_error(None)
    assert (got == _hashlib_digests(again)).all()
