"""Python face of ``cas_hash_host_static`` (include/b200cas.h): hashing without allocating.

For callers that cannot let the library call cudaMalloc/cudaHostAlloc on their behalf --
a request handler that must not take cudaMalloc's device-wide synchronisation, a process
whose GPU memory is already owned by another allocator or capped by its operator.  The
device staging is a fixed ~8 MiB array inside libb200cas.so; batches are limited to
``STATIC_MAX_OBJECTS`` objects and ``STATIC_MAX_BYTES`` bytes (objects rounded up to 16 bytes)
and larger ones raise ``CasError``.  Only numpy host arrays are used here, no torch tensors.
Make the first call on a device before any CUDA stream capture; calls are serialised inside
the library because they all share the one array.
"""
from __future__ import annotations

import numpy as np

from . import _ffi
from ._ffi import STATIC_MAX_BYTES, STATIC_MAX_OBJECTS, CasError  # noqa: F401  (re-exported limits)
from .hasher import _addresses, hex_ids


def digest_many_static(objs, device: int = 0, stream: int = 0) -> np.ndarray:
    """(n, 32) uint8 SHA-256 digests of a small batch of bytes-like objects; allocates no GPU memory."""
    lib = _ffi.load()
    n = len(objs)
    out = np.empty((n, 32), dtype=np.uint8)
    if n == 0:
        return out
    addr, length, keep = _addresses(objs)
    rc = lib.cas_hash_host_static(int(device), None, addr.ctypes.data, length.ctypes.data, n, out.ctypes.data, int(stream))
    del keep
    if rc != _ffi.CAS_OK:
        raise CasError(rc, (lib.cas_last_error(None) or b"").decode())
    return out


def hash_many_static(objs, device: int = 0, stream: int = 0) -> list[str]:
    """Object ids (64 hex characters) of a small batch; see ``digest_many_static``."""
    return hex_ids(digest_many_static(objs, device, stream))
