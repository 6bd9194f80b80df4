"""oracle/sha256_oracle.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Python face of the CPU oracle.  Two independent statements of the same
function are exposed so each can check the other:

* ``hashlib_digests`` -- CPython ``hashlib.sha256`` (OpenSSL EVP), the oracle
  BASELINE.json names.  The reference itself computes no digest at the
  surveyed commit (``/root/reference/src/code_interpreter/services/storage.py:52``
  assigns ``secrets.token_hex(32)``); its docstring (storage.py:36) promises
  SHA-256 ids, which is what this path delivers.
* ``c_digests`` -- the plain-C restatement in ``sha256_ref.c`` (FIPS 180-4),
  loaded with ctypes from ``oracle/liboracle_sha256.so`` (``make -C oracle``).

Parity is unpinned by the reference (no digest constant in its tests); it is
pinned by the FIPS 180-4 known answers in tests/golden/sha256_kat.json.

Also here: the numpy statement of the synthetic-object generator that the
device kernel ``cas_gen_synthetic`` must reproduce byte for byte.
"""
from __future__ import annotations

import ctypes
import hashlib
import os
import subprocess
from concurrent.futures import ThreadPoolExecutor
from typing import Iterable, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle_sha256.so")
_lib = None

SEED_DEFAULT = 0xB200
_GOLDEN = np.uint64(0x9E3779B97F4A7C15)
_M1 = np.uint64(0xBF58476D1CE4E5B9)
_M2 = np.uint64(0x94D049BB133111EB)


def build(force: bool = False) -> str:
    """Compile sha256_ref.c with gcc (seconds).  Returns the .so path."""
    src = os.path.join(_HERE, "sha256_ref.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B", "liboracle_sha256.so"])
    return _LIB_PATH


def _load():
    global _lib
    if _lib is None:
        build()
        lib = ctypes.CDLL(_LIB_PATH)
        u8p = ctypes.c_void_p
        lib.oracle_sha256.argtypes = [u8p, ctypes.c_uint64, u8p]
        lib.oracle_sha256.restype = None
        lib.oracle_sha256_blocks.argtypes = [ctypes.c_uint64]
        lib.oracle_sha256_blocks.restype = ctypes.c_uint64
        lib.oracle_sha256_batch.argtypes = [u8p, u8p, u8p, ctypes.c_uint64, u8p]
        lib.oracle_sha256_batch.restype = None
        lib.oracle_sha256_batch_mt.argtypes = [u8p, u8p, u8p, ctypes.c_uint64, u8p, ctypes.c_int]
        lib.oracle_sha256_batch_mt.restype = ctypes.c_int
        lib.oracle_gen_synthetic.argtypes = [u8p] + [ctypes.c_uint64] * 5
        lib.oracle_gen_synthetic.restype = None
        _lib = lib
    return _lib


def nblocks(length: int) -> int:
    """Compressions a message of ``length`` bytes costs: ceil((L+9)/64)."""
    return (length + 9 + 63) // 64


def hashlib_digests(objs: Iterable[bytes | memoryview]) -> list[bytes]:
    return [hashlib.sha256(o).digest() for o in objs]


def hashlib_digests_mt(objs: Sequence[bytes | memoryview], threads: int) -> list[bytes]:
    """hashlib over a thread pool (hashlib drops the GIL for buffers >= 2 KiB)."""
    if threads <= 1:
        return hashlib_digests(objs)
    n = len(objs)
    spans = [(n * t // threads, n * (t + 1) // threads) for t in range(threads)]

    def run(span):
        lo, hi = span
        return [hashlib.sha256(objs[i]).digest() for i in range(lo, hi)]

    with ThreadPoolExecutor(threads) as ex:
        parts = list(ex.map(run, spans))
    return [d for p in parts for d in p]


def c_digest(data: bytes) -> bytes:
    lib = _load()
    out = ctypes.create_string_buffer(32)
    buf = np.frombuffer(data, dtype=np.uint8) if len(data) else np.zeros(1, np.uint8)
    lib.oracle_sha256(buf.ctypes.data, len(data), ctypes.addressof(out))
    return out.raw


def c_digests_packed(base: np.ndarray, off: np.ndarray, length: np.ndarray, threads: int = 1) -> np.ndarray:
    """Digests of objects ``base[off[i]:off[i]+length[i]]``; returns (n, 32) uint8."""
    lib = _load()
    base = np.ascontiguousarray(base, dtype=np.uint8)
    off = np.ascontiguousarray(off, dtype=np.uint64)
    length = np.ascontiguousarray(length, dtype=np.uint64)
    n = int(off.shape[0])
    out = np.empty((n, 32), dtype=np.uint8)
    bptr = base.ctypes.data if base.size else 0
    if threads > 1:
        rc = lib.oracle_sha256_batch_mt(bptr, off.ctypes.data, length.ctypes.data, n, out.ctypes.data, threads)
        if rc != 0:
            raise RuntimeError(f"oracle_sha256_batch_mt failed: {rc}")
    else:
        lib.oracle_sha256_batch(bptr, off.ctypes.data, length.ctypes.data, n, out.ctypes.data)
    return out


def c_digests(objs: Sequence[bytes]) -> list[bytes]:
    return [c_digest(bytes(o)) for o in objs]


# ---------------------------------------------------------------- synthetic --
def synth_words(seed: int, obj: int, nwords: int, first_word: int = 0) -> np.ndarray:
    """uint64 words ``first_word .. first_word+nwords`` of synthetic object ``obj``."""
    with np.errstate(over="ignore"):
        j = np.arange(first_word, first_word + nwords, dtype=np.uint64)
        ctr = (np.uint64(obj) << np.uint64(32)) | (j & np.uint64(0xFFFFFFFF))
        z = np.uint64(seed) + _GOLDEN * ctr
        z = (z ^ (z >> np.uint64(30))) * _M1
        z = (z ^ (z >> np.uint64(27))) * _M2
        return z ^ (z >> np.uint64(31))


def synth_object(seed: int, obj: int, length: int) -> bytes:
    """Bytes of synthetic object ``obj`` (little-endian words, truncated to length)."""
    w = synth_words(seed, obj, (length + 7) // 8)
    return w.astype("<u8").tobytes()[:length]


def synth_batch(seed: int, first_obj: int, n: int, length: int, stride: int | None = None) -> np.ndarray:
    """``n`` objects of ``length`` bytes laid out at ``stride`` (C generator)."""
    lib = _load()
    stride = length if stride is None else stride
    buf = np.zeros(max(1, n * stride), dtype=np.uint8)
    lib.oracle_gen_synthetic(buf.ctypes.data, seed, first_obj, n, length, stride)
    return buf[: n * stride]
