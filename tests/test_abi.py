"""The C-ABI library: loads here (no GPU), exports what include/b200cas.h declares, fails loudly."""
import ctypes
import os
import re

import pytest

import __graft_entry__ as entry
from bee_code_interpreter_fs_b200 import CasError, _ffi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    entry.build()
    return _ffi.load()


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "b200cas.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(cas_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_all_exported(lib):
    names = declared_symbols()
    assert len(names) >= 13
    for name in names:
        assert hasattr(lib, name), f"{name} declared in include/b200cas.h but not exported"
    assert sorted(_ffi.SYMBOLS) == names


def test_version(lib):
    assert lib.cas_version().decode().startswith("b200cas")


def test_no_cpu_fallback_without_device(lib):
    import torch

    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    ctx = ctypes.c_void_p()
    rc = lib.cas_create(0, ctypes.byref(ctx))
    assert rc == _ffi.CAS_E_NODEVICE and not ctx.value
    assert b"no CPU fallback" in lib.cas_last_error(None)
    from bee_code_interpreter_fs_b200 import GpuHasher, Storage

    with pytest.raises(CasError):
        GpuHasher(0)
    with pytest.raises(CasError):
        Storage("/tmp/never-created")


def test_product_code_never_imports_oracle():
    pkg = os.path.join(ROOT, "bee_code_interpreter_fs_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "from oracle" not in text, f
                assert "import hashlib" not in text, f"{f} computes digests on the CPU"
