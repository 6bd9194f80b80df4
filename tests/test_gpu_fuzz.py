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
