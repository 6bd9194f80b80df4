import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with `pytest -m gpu` on the GPU box)")


def _cuda_available() -> bool:
    try:
        import torch

        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _cuda_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def kat():
    with open(os.path.join(GOLDEN, "sha256_kat.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def storage_behaviour():
    with open(os.path.join(GOLDEN, "storage_behaviour.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def hasher():
    """One GpuHasher for the GPU tests. No fallback: a missing library or device fails the test."""
    from bee_code_interpreter_fs_b200 import GpuHasher

    h = GpuHasher(0)
    yield h
    h.close()


class OracleHasher:
    """Stand-in for GpuHasher in CPU-only tests of the host logic (tests may use the oracle)."""

    def __init__(self):
        self.calls = 0

    def hash_many(self, objs):
        from oracle import sha256_oracle

        self.calls += 1
        return [d.hex() for d in sha256_oracle.hashlib_digests(objs)]

    def digest_packed(self, base, off, length):
        import ctypes

        import numpy as np
        from oracle import sha256_oracle

        assert base is None
        objs = [ctypes.string_at(int(a), int(n)) for a, n in zip(off, length)]
        return np.frombuffer(b"".join(sha256_oracle.hashlib_digests(objs)), dtype=np.uint8).reshape(-1, 32)


class OracleFileHasher(OracleHasher):
    """Same, but also able to hash files by path, like GpuHasher.hash_files."""

    def __init__(self):
        super().__init__()
        self.file_calls = 0

    def hash_files(self, paths):
        from oracle import sha256_oracle

        self.file_calls += 1
        return [d.hex() for d in sha256_oracle.hashlib_digests([open(p, "rb").read() for p in paths])]


@pytest.fixture
def oracle_hasher():
    return OracleHasher()


@pytest.fixture
def oracle_file_hasher():
    return OracleFileHasher()
