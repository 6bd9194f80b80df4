"""Streaming form of the C ABI (cas_stream_begin / update / finish): objects that arrive in
pieces, as the reference's writers receive them (storage.py:44-55,
services/kubernetes_code_executor.py:139-145).  Every digest against hashlib.sha256."""
import ctypes
import hashlib

import numpy as np
import pytest

from bee_code_interpreter_fs_b200 import _ffi

pytestmark = pytest.mark.gpu


def test_streams_fed_in_arbitrary_pieces_match_hashlib(hasher):
    rng = np.random.default_rng(17)
    sizes = [0, 1, 55, 56, 63, 64, 65, 119, 120, 127, 128, 1000, 4096, 65536, 100001, (1 << 20) + 13]
    blobs = [rng.integers(0, 256, size=n, dtype=np.uint8).tobytes() for n in sizes]
    states = [hasher.stream_begin() for _ in blobs]
    pos = [0] * len(blobs)
    before = hasher.launch_count
    calls = 0
    while any(p < len(b) for p, b in zip(pos, blobs)):
        pieces = []
        for i, b in enumerate(blobs):
            k = int(rng.choice([0, 1, 7, 63, 64, 65, 128, 1000, 50000, 300000]))
            pieces.append(b[pos[i]:pos[i] + k])
            pos[i] += len(pieces[-1])
        hasher.stream_update(states, pieces)  # all streams advance in one call
        calls += 1
    assert [s.length for s in states] == sizes
    assert hasher.stream_finish(states) == [hashlib.sha256(b).hexdigest() for b in blobs]
    assert hasher.launch_count - before <= 2 * calls + 2  # one launch per update unless nothing filled a block


def test_a_stream_larger_than_the_staging_buffer_goes_in_rounds(hasher):
    saved = hasher.get_option(_ffi.OPT_WAVE_BYTES)
    hasher.set_option(_ffi.OPT_WAVE_BYTES, 1 << 20)
    try:
        rng = np.random.default_rng(18)
        big = rng.integers(0, 256, size=(9 << 20) + 77, dtype=np.uint8).tobytes()
        small = [rng.integers(0, 256, size=int(n), dtype=np.uint8).tobytes() for n in rng.integers(0, 3 << 20, size=5)]
        states = [hasher.stream_begin() for _ in range(6)]
        hasher.stream_update(states, [big[:5 << 20]] + small)
        hasher.stream_update(states[:1], [big[5 << 20:]])
        assert hasher.stream_finish(states) == [hashlib.sha256(b).hexdigest() for b in [big] + small]
    finally:
        hasher.set_option(_ffi.OPT_WAVE_BYTES, saved)


def test_stream_state_is_plain_data_through_the_c_abi(hasher):
    """A state can be copied mid-stream and both copies continued: nothing is hidden in the context."""
    lib = _ffi.load()
    st = _ffi.StreamState()
    lib.cas_stream_begin(ctypes.byref(st))
    assert list(st.h) == [0x6a09e667, 0xbb67ae85, 0x3c6ef372, 0xa54ff53a, 0x510e527f, 0x9b05688c, 0x1f83d9ab, 0x5be0cd19]
    hasher.stream_update([st], [b"common prefix " * 100])
    fork = _ffi.StreamState.from_buffer_copy(bytes(st))
    hasher.stream_update([st, fork], [b"left", b"right"])
    assert hasher.stream_finish([st, fork]) == [hashlib.sha256(b"common prefix " * 100 + t).hexdigest()
                                               for t in (b"left", b"right")]
    assert lib.cas_stream_update(hasher._ctx, None, None, None, 3) == _ffi.CAS_E_ARG
    assert hasher.stream_finish([hasher.stream_begin()]) == [hashlib.sha256(b"").hexdigest()]
