"""B200-native batched SHA-256 content addressing behind the bee-code-interpreter Storage API.

Public surface:
  Storage            drop-in for code_interpreter.services.storage.Storage (ids = SHA-256 hex)
  GpuHasher          one device: hash_many / digest_device / digest_device_uniform / gen_synthetic
  MultiGpuHasher     shards a batch by object over several devices (no collective)
  Hash, AbsolutePath the reference's id / path types (utils/validation.py:19-22)

The CUDA library (libb200cas.so, built in-tree by ``__graft_entry__.build()``) is
required: there is no CPU fallback, constructing a hasher without it raises.
"""
from ._ffi import CasError
from .validation import AbsolutePath, Hash

__all__ = ["Storage", "GpuHasher", "MultiGpuHasher", "Hash", "AbsolutePath", "CasError"]


def __getattr__(name):  # lazy: keeps `import bee_code_interpreter_fs_b200` free of anyio/torch imports
    if name in ("GpuHasher", "MultiGpuHasher"):
        from . import hasher

        return getattr(hasher, name)
    if name == "Storage":
        from .storage import Storage

        return Storage
    raise AttributeError(name)
