"""ctypes binding of include/b200cas.h.  No fallback: a missing library is an error."""
from __future__ import annotations

import ctypes
import os

from ._build import LIB_PATH

CAS_OK = 0
CAS_E_CUDA, CAS_E_ARG, CAS_E_NOMEM, CAS_E_NODEVICE, CAS_E_IO = -1, -2, -3, -4, -5
OPT_KERNEL, OPT_FMA_LEVEL, OPT_WAVE_BYTES, OPT_PACK_THREADS, OPT_CTAS_PER_SM, OPT_ARENA_BYTES = 1, 2, 3, 4, 5, 6
OPT_LONG_MAX_OBJECTS, OPT_LONG_MIN_BYTES = 7, 8
OPT_LONG_MAX_PER_WARP, OPT_FIRST_CHUNK_BYTES, OPT_IO_THREADS, OPT_CHUNK_DIVISOR = 9, 10, 11, 12
STORED_NEW, STORED_EXISTED = 0, 1

# Every symbol include/b200cas.h declares (tests check the library exports them all).
SYMBOLS = (
    "cas_create", "cas_destroy", "cas_last_error", "cas_set_option", "cas_get_option", "cas_device_info",
    "cas_hash_host", "cas_hash_files", "cas_hash_device", "cas_hash_device_long", "cas_hash_device_uniform", "cas_gen_synthetic", "cas_int32_peak",
    "cas_launch_count", "cas_version", "cas_hash_host_static", "cas_reserve", "cas_store_host", "cas_publish_files", "cas_verify_names", "cas_stream_begin", "cas_stream_update", "cas_stream_finish", "cas_set_h2d_gate",
)
STATIC_MAX_OBJECTS = 1024
STATIC_MAX_BYTES = 8 << 20


class StreamState(ctypes.Structure):
    """cas_stream_state of include/b200cas.h: 8 chaining words, byte count, unhashed tail."""

    _fields_ = [("h", ctypes.c_uint32 * 8), ("length", ctypes.c_uint64), ("tail", ctypes.c_uint8 * 64)]


class CasError(RuntimeError):
    """The engine reported a failure; never swallowed into a placeholder digest."""

    def __init__(self, code: int, message: str):
        super().__init__(f"b200cas error {code}: {message}")
        self.code = code


_lib = None


def load() -> ctypes.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). This engine has no CPU fallback."
        )
    lib = ctypes.CDLL(LIB_PATH)
    vp, u64, i64, ci = ctypes.c_void_p, ctypes.c_uint64, ctypes.c_int64, ctypes.c_int
    lib.cas_create.argtypes = [ci, ctypes.POINTER(vp)]
    lib.cas_create.restype = ci
    lib.cas_destroy.argtypes = [vp]
    lib.cas_destroy.restype = None
    lib.cas_last_error.argtypes = [vp]
    lib.cas_last_error.restype = ctypes.c_char_p
    lib.cas_set_option.argtypes = [vp, ci, i64]
    lib.cas_set_option.restype = ci
    lib.cas_get_option.argtypes = [vp, ci, ctypes.POINTER(i64)]
    lib.cas_get_option.restype = ci
    lib.cas_device_info.argtypes = [vp, ctypes.POINTER(ci), ctypes.POINTER(ci), ctypes.POINTER(u64)]
    lib.cas_device_info.restype = ci
    lib.cas_hash_host.argtypes = [vp, vp, vp, vp, u64, vp]
    lib.cas_hash_host.restype = ci
    lib.cas_hash_files.argtypes = [vp, vp, u64, vp, vp]
    lib.cas_hash_files.restype = ci
    lib.cas_hash_device.argtypes = [vp, vp, vp, vp, vp, u64, vp, vp]
    lib.cas_hash_device.restype = ci
    lib.cas_hash_device_long.argtypes = [vp, vp, vp, vp, vp, u64, vp, vp]
    lib.cas_hash_device_long.restype = ci
    lib.cas_hash_device_uniform.argtypes = [vp, vp, u64, u64, u64, vp, vp]
    lib.cas_hash_device_uniform.restype = ci
    lib.cas_gen_synthetic.argtypes = [vp, vp, u64, u64, u64, u64, u64, vp]
    lib.cas_gen_synthetic.restype = ci
    lib.cas_int32_peak.argtypes = [vp, ci, ci, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double)]
    lib.cas_int32_peak.restype = ci
    lib.cas_hash_host_static.argtypes = [ci, vp, vp, vp, u64, vp, vp]
    lib.cas_hash_host_static.restype = ci
    lib.cas_reserve.argtypes = [vp, u64, u64]
    lib.cas_reserve.restype = ci
    lib.cas_store_host.argtypes = [vp, ctypes.c_char_p, vp, vp, vp, u64, vp, vp]
    lib.cas_store_host.restype = ci
    lib.cas_publish_files.argtypes = [vp, ctypes.c_char_p, vp, u64, vp, vp, vp]
    lib.cas_publish_files.restype = ci
    lib.cas_verify_names.argtypes = [vp, ctypes.c_char_p, ctypes.c_char_p, u64, vp]
    lib.cas_verify_names.restype = ci
    lib.cas_stream_begin.argtypes = [vp]
    lib.cas_stream_begin.restype = None
    lib.cas_stream_update.argtypes = [vp, vp, vp, vp, u64]
    lib.cas_stream_update.restype = ci
    lib.cas_stream_finish.argtypes = [vp, vp, u64, vp]
    lib.cas_stream_finish.restype = ci
    lib.cas_set_h2d_gate.argtypes = [vp, ctypes.c_char_p, ci]
    lib.cas_set_h2d_gate.restype = ci
    lib.cas_launch_count.argtypes = [vp]
    lib.cas_launch_count.restype = u64
    lib.cas_version.argtypes = []
    lib.cas_version.restype = ctypes.c_char_p
    _lib = lib
    return lib
