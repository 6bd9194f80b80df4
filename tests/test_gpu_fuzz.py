"""Property-based parity (hypothesis): arbitrary batches of arbitrary bytes through the host
path, the ragged device path and the warp-per-object kernel must equal hashlib, bit for bit."""
import hashlib

import numpy as np
import pytest
import torch
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

pytestmark = pytest.mark.gpu

blob = st.binary(min_size=0, max_size=700)
batch = st.lists(blob, min_size=1, max_size=40)


@settings(max_examples=60, deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture])
@given(objs=batch)
def test_host_path_equals_hashlib(hasher, objs):
    assert hasher.hash_many(objs) == [hashlib.sha256(o).hexdigest() for o in objs]


@settings(max_examples=40, deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture])
@given(objs=batch, lead=st.integers(min_value=0, max_value=19), gaps=st.lists(st.integers(0, 5), min_size=40, max_size=40),
       warp=st.booleans())
def test_device_paths_equal_hashlib_at_any_offset(hasher, objs, lead, gaps, warp):
    off, cursor = [], lead
    for o, g in zip(objs, gaps):
        off.append(cursor)
        cursor += len(o) + g
    data = np.zeros(cursor + 8, dtype=np.uint8)
    for o, a in zip(objs, off):
        data[a:a + len(o)] = np.frombuffer(o, dtype=np.uint8)
    d = torch.from_numpy(data).cuda()
    d_off = torch.tensor(off, dtype=torch.int64, device="cuda:0")
    d_len = torch.tensor([len(o) for o in objs], dtype=torch.int64, device="cuda:0")
    got = hasher.digest_device(d, d_off, d_len, warp_per_object=warp).cpu().numpy()
    assert [bytes(r).hex() for r in got] == [hashlib.sha256(o).hexdigest() for o in objs]


@settings(max_examples=25, deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture])
@given(objs=batch, per_warp_budget=st.sampled_from([1, 2, 5, 40]))
def test_every_objects_per_warp_form_equals_hashlib(hasher, objs, per_warp_budget):
    """The warp kernel with 1 ... 32 objects per warp (picked from the object count and the warp
    budget), on arbitrary mixes of lengths inside one warp."""
    from bee_code_interpreter_fs_b200 import _ffi

    saved = hasher.get_option(_ffi.OPT_LONG_MAX_OBJECTS)
    hasher.set_option(_ffi.OPT_LONG_MAX_OBJECTS, per_warp_budget)
    try:
        data = np.frombuffer(b"".join(o + b"\0" * (-len(o) % 16) for o in objs) + b"\0" * 16, dtype=np.uint8)
        off = np.cumsum([0] + [len(o) + (-len(o) % 16) for o in objs[:-1]])
        d = torch.from_numpy(data.copy()).cuda()
        d_off = torch.tensor(off, dtype=torch.int64, device="cuda:0")
        d_len = torch.tensor([len(o) for o in objs], dtype=torch.int64, device="cuda:0")
        got = hasher.digest_device(d, d_off, d_len, warp_per_object=True).cpu().numpy()
    finally:
        hasher.set_option(_ffi.OPT_LONG_MAX_OBJECTS, saved)
    assert [bytes(r).hex() for r in got] == [hashlib.sha256(o).hexdigest() for o in objs]


@settings(max_examples=25, deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture])
@given(objs=st.lists(st.binary(min_size=0, max_size=400), min_size=1, max_size=12),
       cuts=st.lists(st.integers(min_value=0, max_value=200), min_size=1, max_size=6))
def test_streams_cut_anywhere_equal_hashlib(hasher, objs, cuts):
    states = [hasher.stream_begin() for _ in objs]
    pos = [0] * len(objs)
    for c in cuts:
        hasher.stream_update(states, [o[p:p + c] for o, p in zip(objs, pos)])
        pos = [min(len(o), p + c) for o, p in zip(objs, pos)]
    hasher.stream_update(states, [o[p:] for o, p in zip(objs, pos)])
    assert hasher.stream_finish(states) == [hashlib.sha256(o).hexdigest() for o in objs]


@settings(max_examples=15, deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture])
@given(objs=batch)
def test_store_many_equals_hashlib_and_the_files_hold_the_bytes(hasher, objs, tmp_path_factory):
    root = tmp_path_factory.mktemp("store")
    ids, status = hasher.store_many(root, objs)
    assert ids == [hashlib.sha256(o).hexdigest() for o in objs]
    assert all(s in (0, 1) for s in status.tolist())
    for object_id, o in zip(ids, objs):
        assert (root / object_id).read_bytes() == o
    assert sorted(p.name for p in root.iterdir()) == sorted(set(ids))
    assert hasher.verify_names(root, ids).tolist() == [1] * len(ids)
