"""GPU SHA-256 hashers over the C ABI (include/b200cas.h).

``GpuHasher`` is one device; ``MultiGpuHasher`` shards a batch by object over
several devices from one process (one host thread per device, no collective).
PyTorch is used only for device buffers and stream handles.
"""
from __future__ import annotations

import ctypes
import os
import threading
from typing import Sequence

import numpy as np

from . import _ffi, shard
from ._ffi import CasError

BytesLike = bytes | bytearray | memoryview | np.ndarray


def hex_ids(digests: np.ndarray) -> list[str]:
    """64-character lowercase hex ids of an (n, 32) digest array (one hex() call, then slices)."""
    flat = np.ascontiguousarray(digests, dtype=np.uint8).tobytes().hex()
    return [flat[i:i + 64] for i in range(0, len(flat), 64)]


def _bytes_data_offset() -> int:
    """Offset of a ``bytes`` object's characters from ``id(obj)`` on this interpreter, or 0 if it
    cannot be confirmed (then addresses are taken one by one through ctypes).  Saves about
    0.3 us per object on the way into the engine, which matters for batches of small objects."""
    try:
        guess = bytes.__basicsize__ - 1
        for probe in (b"probe-one", b"x" * 5000, b""):
            want = ctypes.cast(ctypes.c_char_p(probe), ctypes.c_void_p).value
            if id(probe) + guess != want:
                return 0
        return guess
    except Exception:  # noqa: BLE001 - not CPython, or a layout this check does not know
        return 0


_BYTES_DATA_OFFSET = _bytes_data_offset()


def _addresses(objs: Sequence[BytesLike]) -> tuple[np.ndarray, np.ndarray, list]:
    """Host addresses and lengths of the objects, without copying them."""
    n = len(objs)
    keep: list = []
    if set(map(type, objs)) <= {bytes}:
        length = np.fromiter(map(len, objs), dtype=np.uint64, count=n)
        if _BYTES_DATA_OFFSET:  # CPython: the characters sit at a fixed offset inside the object
            addr = np.fromiter(map(id, objs), dtype=np.uint64, count=n) + np.uint64(_BYTES_DATA_OFFSET)
            keep.append(objs)
        else:
            arr = (ctypes.c_char_p * n)(*objs)  # pointers into the bytes objects themselves
            addr = np.frombuffer(arr, dtype=np.uint64, count=n).copy() if n else np.zeros(0, np.uint64)
            keep.append(arr)
    else:
        length = np.fromiter((len(o) if not isinstance(o, np.ndarray) else o.nbytes for o in objs), dtype=np.uint64,
                             count=n)
        addr = np.zeros(n, dtype=np.uint64)
        for i, o in enumerate(objs):
            if isinstance(o, np.ndarray):
                a = np.ascontiguousarray(o).view(np.uint8).reshape(-1)
            else:
                a = np.frombuffer(o, dtype=np.uint8)
            keep.append(a)
            addr[i] = a.ctypes.data if a.size else 0
    return addr, length, keep


class GpuHasher:
    """SHA-256 of batches of objects on one B200."""

    def __init__(self, device: int = 0):
        self._lib = _ffi.load()
        self._ctx = ctypes.c_void_p()
        rc = self._lib.cas_create(int(device), ctypes.byref(self._ctx))
        if rc != _ffi.CAS_OK:
            msg = (self._lib.cas_last_error(None) or b"").decode()
            self._ctx = ctypes.c_void_p()
            raise CasError(rc, msg)
        self.device = int(device)
        self._lock = threading.Lock()

    # -- plumbing ---------------------------------------------------------
    def _check(self, rc: int) -> None:
        if rc != _ffi.CAS_OK:
            raise CasError(rc, (self._lib.cas_last_error(self._ctx) or b"").decode())

    def close(self) -> None:
        if getattr(self, "_ctx", None) and self._ctx.value:
            self._lib.cas_destroy(self._ctx)
            self._ctx = ctypes.c_void_p()

    def __del__(self):  # pragma: no cover - best effort
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def set_option(self, key: int, value: int) -> None:
        self._check(self._lib.cas_set_option(self._ctx, key, int(value)))

    def get_option(self, key: int) -> int:
        v = ctypes.c_int64()
        self._check(self._lib.cas_get_option(self._ctx, key, ctypes.byref(v)))
        return v.value

    def device_info(self) -> dict:
        sm, khz, mem = ctypes.c_int(), ctypes.c_int(), ctypes.c_uint64()
        self._check(self._lib.cas_device_info(self._ctx, ctypes.byref(sm), ctypes.byref(khz), ctypes.byref(mem)))
        return {"sm_count": sm.value, "sm_clock_khz": khz.value, "mem_bytes": mem.value}

    @property
    def launch_count(self) -> int:
        return int(self._lib.cas_launch_count(self._ctx))

    @staticmethod
    def _stream_handle(stream) -> int:
        if stream is None:
            return 0
        return int(getattr(stream, "cuda_stream", stream))

    # -- host buffers (end-to-end path) -------------------------------------
    def digest_packed(self, base, off, length) -> np.ndarray:
        """Digests of ``base[off[i]:off[i]+length[i]]``; ``base`` is host memory
        (numpy array, pinned torch tensor, or None for absolute addresses)."""
        off = np.ascontiguousarray(off, dtype=np.uint64)
        length = np.ascontiguousarray(length, dtype=np.uint64)
        n = int(off.shape[0])
        if length.shape[0] != n:
            raise ValueError("off and length differ in size")
        out = np.empty((n, 32), dtype=np.uint8)
        if base is None:
            bptr = 0
        elif isinstance(base, np.ndarray):
            bptr = base.ctypes.data
        elif hasattr(base, "data_ptr"):
            if base.is_cuda:
                raise ValueError("digest_packed takes host memory; use digest_device for device tensors")
            bptr = base.data_ptr()
        else:
            bptr = int(base)
        with self._lock:
            self._check(self._lib.cas_hash_host(self._ctx, bptr, off.ctypes.data, length.ctypes.data, n, out.ctypes.data))
        return out

    def digest_many(self, objs: Sequence[BytesLike]) -> np.ndarray:
        """(n, 32) uint8 digests of a list of bytes-like objects (no Python-side copy)."""
        if len(objs) == 0:
            return np.empty((0, 32), dtype=np.uint8)
        addr, length, keep = _addresses(objs)
        out = self.digest_packed(None, addr, length)
        del keep
        return out

    def hash_many(self, objs: Sequence[BytesLike]) -> list[str]:
        """Object ids (64 lowercase hex characters) of a batch of objects."""
        return hex_ids(self.digest_many(objs))

    def digest_files(self, paths: Sequence[str | os.PathLike]) -> tuple[np.ndarray, np.ndarray]:
        """((n, 32) digests, sizes) of files, read straight into pinned staging by the engine."""
        n = len(paths)
        out = np.empty((n, 32), dtype=np.uint8)
        sizes = np.zeros(n, dtype=np.uint64)
        if n == 0:
            return out, sizes
        raw = [os.fsencode(p) for p in paths]
        arr = (ctypes.c_char_p * n)(*raw)
        with self._lock:
            self._check(self._lib.cas_hash_files(self._ctx, ctypes.cast(arr, ctypes.c_void_p), n, out.ctypes.data,
                                                 sizes.ctypes.data))
        return out, sizes

    def hash_files(self, paths: Sequence[str | os.PathLike]) -> list[str]:
        """Object ids of files on disk (the store directory, a workspace)."""
        return hex_ids(self.digest_files(paths)[0])

    def hash_one(self, data: BytesLike) -> str:
        return self.hash_many([data])[0]

    def reserve(self, batch_bytes: int, max_objects: int) -> None:
        """Allocate the host path's buffers now for batches up to this size (cas_reserve), so the
        first request does not pay for cudaMalloc / cudaHostAlloc."""
        with self._lock:
            self._check(self._lib.cas_reserve(self._ctx, int(batch_bytes), int(max_objects)))

    def set_h2d_gate(self, name: str | None, slots: int = 2) -> None:
        """Take turns with the other hashers (any process) that name the same gate when copying
        from pinned host memory: at most ``slots`` copy chunks run at once (cas_set_h2d_gate)."""
        with self._lock:
            self._check(self._lib.cas_set_h2d_gate(self._ctx, None if name is None else name.encode(), int(slots)))

    # -- hashing and storing in one call (the Storage-shaped path) ----------------------------
    def store_many(self, root: str | os.PathLike, objs: Sequence[BytesLike]) -> tuple[list[str], np.ndarray]:
        """Hash a batch and store every object as ``<root>/<id>`` (cas_store_host: file writers run
        beside the GPU pipeline, final names by atomic rename).  Returns (ids, status) with status
        0 = stored now, 1 = an object of that id was there already, < 0 = that object failed."""
        n = len(objs)
        status = np.zeros(n, dtype=np.int32)
        if n == 0:
            return [], status
        addr, length, keep = _addresses(objs)
        out = np.empty((n, 32), dtype=np.uint8)
        with self._lock:
            self._check(self._lib.cas_store_host(self._ctx, os.fsencode(root), None, addr.ctypes.data, length.ctypes.data, n,
                                                 out.ctypes.data, status.ctypes.data))
        del keep
        return hex_ids(out), status

    def publish_files(self, root: str | os.PathLike | None, paths: Sequence[str | os.PathLike]):
        """Hash files and rename each to ``<root>/<id>`` (cas_publish_files); with ``root`` None
        they are only hashed.  A file that cannot be read fails alone.  Returns (ids, sizes, status)."""
        n = len(paths)
        out = np.empty((n, 32), dtype=np.uint8)
        sizes = np.zeros(n, dtype=np.uint64)
        status = np.zeros(n, dtype=np.int32)
        if n == 0:
            return [], sizes, status
        raw = [os.fsencode(p) for p in paths]
        arr = (ctypes.c_char_p * n)(*raw)
        with self._lock:
            self._check(self._lib.cas_publish_files(self._ctx, None if root is None else os.fsencode(root),
                                                    ctypes.cast(arr, ctypes.c_void_p), n, out.ctypes.data, sizes.ctypes.data,
                                                    status.ctypes.data))
        return hex_ids(out), sizes, status

    # -- device buffers (kernel-only path) ------------------------------------
    def digest_device(self, base, off, length, order=None, out=None, stream=None, warp_per_object=False):
        """Hash objects resident in device memory. ``base`` is a CUDA uint8 tensor (or
        pointer), ``off``/``length`` CUDA int64/uint64 tensors; returns a CUDA (n, 32)
        uint8 tensor. Asynchronous on ``stream`` (a torch stream) or the current stream.
        ``warp_per_object`` selects the kernel for few, long objects (cas_hash_device_long)."""
        import torch

        n = int(off.numel())
        dev = off.device
        if out is None:
            out = torch.empty((n, 32), dtype=torch.uint8, device=dev)
        if stream is None:
            stream = torch.cuda.current_stream(dev)
        bptr = base.data_ptr() if hasattr(base, "data_ptr") else int(base)
        optr = order.data_ptr() if order is not None else 0
        fn = self._lib.cas_hash_device_long if warp_per_object else self._lib.cas_hash_device
        with self._lock:
            self._check(fn(self._ctx, bptr, off.data_ptr(), length.data_ptr(), optr, n, out.data_ptr(),
                           self._stream_handle(stream)))
        return out

    def digest_device_uniform(self, base, stride: int, length: int, n: int, out=None, stream=None):
        """Uniform layout: object i at ``base + i*stride``, ``length`` bytes each."""
        import torch

        dev = base.device if hasattr(base, "device") else torch.device("cuda", self.device)
        if out is None:
            out = torch.empty((n, 32), dtype=torch.uint8, device=dev)
        if stream is None:
            stream = torch.cuda.current_stream(dev)
        bptr = base.data_ptr() if hasattr(base, "data_ptr") else int(base)
        with self._lock:
            self._check(self._lib.cas_hash_device_uniform(self._ctx, bptr, int(stride), int(length), int(n),
                                                          out.data_ptr(), self._stream_handle(stream)))
        return out

    def gen_synthetic(self, dst, seed: int, first_obj: int, n: int, length: int, stride: int | None = None, stream=None):
        """Fill ``dst`` (CUDA uint8 tensor) with synthetic objects first_obj..first_obj+n."""
        import torch

        stride = length if stride is None else stride
        need = (n - 1) * stride + length if n else 0
        if dst.numel() * dst.element_size() < need:
            raise ValueError("destination tensor too small")
        if stream is None:
            stream = torch.cuda.current_stream(dst.device)
        with self._lock:
            self._check(self._lib.cas_gen_synthetic(self._ctx, dst.data_ptr(), int(seed), int(first_obj), int(n),
                                                    int(length), int(stride), self._stream_handle(stream)))
        return dst

    def int32_peak(self, mode: int = 0, steps: int = 4096) -> tuple[float, float]:
        """(thread-level integer instructions per second, kernel ms) of the pipe probe."""
        ops, ms = ctypes.c_double(), ctypes.c_double()
        with self._lock:
            self._check(self._lib.cas_int32_peak(self._ctx, int(mode), int(steps), ctypes.byref(ops), ctypes.byref(ms)))
        return ops.value, ms.value


    # -- objects that arrive in pieces ---------------------------------------------------------
    def stream_begin(self) -> "_ffi.StreamState":
        """A fresh streaming state (plain data: it can be kept, copied or dropped freely)."""
        st = _ffi.StreamState()
        self._lib.cas_stream_begin(ctypes.byref(st))
        return st

    def stream_update(self, states: Sequence["_ffi.StreamState"], pieces: Sequence[BytesLike]) -> None:
        """Feeds one piece to each stream; all streams advance in one GPU launch (cas_stream_update)."""
        n = len(states)
        if n != len(pieces):
            raise ValueError("one piece per stream")
        if n == 0:
            return
        addr, length, keep = _addresses(pieces)
        ptrs = (ctypes.c_void_p * n)(*[ctypes.addressof(s) for s in states])
        with self._lock:
            self._check(self._lib.cas_stream_update(self._ctx, ptrs, addr.ctypes.data, length.ctypes.data, n))
        del keep

    def stream_finish(self, states: Sequence["_ffi.StreamState"]) -> list[str]:
        """Pads and closes the streams; returns their ids.  The states must not be used again."""
        n = len(states)
        if n == 0:
            return []
        out = np.empty((n, 32), dtype=np.uint8)
        ptrs = (ctypes.c_void_p * n)(*[ctypes.addressof(s) for s in states])
        with self._lock:
            self._check(self._lib.cas_stream_finish(self._ctx, ptrs, n, out.ctypes.data))
        return hex_ids(out)

    def verify_names(self, root: str | os.PathLike, names: Sequence[str]) -> np.ndarray:
        """Scrub: per name 1 if sha256(<root>/<name>) spells the name, 0 if not (or unreadable),
        2 if the entry is not a regular file (cas_verify_names)."""
        n = len(names)
        out = np.zeros(n, dtype=np.uint8)
        if n == 0:
            return out
        blob = "\0".join(names).encode() + b"\0"
        with self._lock:
            self._check(self._lib.cas_verify_names(self._ctx, os.fsencode(root), blob, n, out.ctypes.data))
        return out


class MultiGpuHasher:
    """Shards a batch by object across several devices of one node (no collective)."""

    def __init__(self, devices: Sequence[int] | None = None):
        if devices is None:  # every visible GPU, in the measured uplink order (dist.gpu_order)
            import torch

            from .dist import gpu_order

            devices = gpu_order(torch.cuda.device_count())
        if not devices:
            raise CasError(_ffi.CAS_E_NODEVICE, "no CUDA device is visible (this engine has no CPU fallback)")
        self.hashers = [GpuHasher(d) for d in devices]

    def close(self) -> None:
        for h in self.hashers:
            h.close()

    def digest_packed(self, base, off, length) -> np.ndarray:
        """Digests of ``base[off[i]:off[i]+length[i]]`` (host memory), sharded by object."""
        off = np.ascontiguousarray(off, dtype=np.uint64)
        length = np.ascontiguousarray(length, dtype=np.uint64)
        if base is None:
            bptr = 0
        elif isinstance(base, np.ndarray):
            bptr = base.ctypes.data
        elif hasattr(base, "data_ptr"):
            bptr = base.data_ptr()
        else:
            bptr = int(base)
        return self._digest_addresses(off + np.uint64(bptr), length)

    def digest_many(self, objs: Sequence[BytesLike]) -> np.ndarray:
        n = len(objs)
        if n == 0:
            return np.empty((0, 32), dtype=np.uint8)
        addr, length, keep = _addresses(objs)
        out = self._digest_addresses(addr, length)
        del keep
        return out

    def _digest_addresses(self, addr: np.ndarray, length: np.ndarray) -> np.ndarray:
        n = int(addr.shape[0])
        if n == 0:
            return np.empty((0, 32), dtype=np.uint8)
        parts = shard.assign(length, len(self.hashers))
        results: list = [None] * len(self.hashers)
        errors: list = []

        def run(r: int) -> None:
            try:
                idx = parts[r]
                results[r] = self.hashers[r].digest_packed(None, addr[idx], length[idx])
            except Exception as e:  # noqa: BLE001 - re-raised below
                errors.append(e)

        threads = [threading.Thread(target=run, args=(r,)) for r in range(len(self.hashers))]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        if errors:
            raise errors[0]
        return shard.scatter_back(parts, results, n)

    def hash_many(self, objs: Sequence[BytesLike]) -> list[str]:
        return hex_ids(self.digest_many(objs))
