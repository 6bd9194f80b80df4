"""MultiGpuHasher on more than one device: the one-process, one-batch, shard-by-object call
(the reference's gather over one request's files,
/root/reference/src/code_interpreter/services/kubernetes_code_executor.py:149-154).
Skipped on a one-GPU box; run with `gpurun --gpus N` (log kept under profiles/)."""
import hashlib

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

needs_two = pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs at least two GPUs")


@pytest.fixture(scope="module")
def multi():
    from bee_code_interpreter_fs_b200 import MultiGpuHasher

    m = MultiGpuHasher()
    yield m
    m.close()


@needs_two
def test_mixed_size_batch_is_lpt_sharded_over_all_devices_and_exact(multi):
    from bee_code_interpreter_fs_b200 import shard

    rng = np.random.default_rng(0xB200)
    lengths = np.floor(2.0 ** rng.uniform(10, 23, size=4096)).astype(np.uint64)  # config 4's shape, 1 KiB-8 MiB
    off = np.concatenate([[0], np.cumsum(lengths)[:-1]]).astype(np.uint64)
    data = rng.integers(0, 256, size=int(lengths.sum()), dtype=np.uint8)
    want = [hashlib.sha256(data[int(o):int(o) + int(n)]).hexdigest() for o, n in zip(off, lengths)]
    world = len(multi.hashers)
    parts = shard.assign(lengths, world)
    assert all(len(p) for p in parts)
    loads = [int(shard.nblocks(lengths[p]).sum()) for p in parts]
    assert max(loads) - min(loads) <= int(shard.nblocks(lengths.max()))  # LPT balance, by compressions
    before = [h.launch_count for h in multi.hashers]
    got = multi.digest_packed(data, off, lengths)
    assert [bytes(d).hex() for d in got] == want
    assert all(h.launch_count > b for h, b in zip(multi.hashers, before)), "every device must have launched kernels"
    mv = memoryview(data)
    assert multi.hash_many([mv[int(o):int(o) + int(n)] for o, n in zip(off[:300], lengths[:300])]) == want[:300]


@needs_two
def test_uniform_batch_is_cut_into_contiguous_ranges_and_exact(multi):
    n, length = 20000, 65536
    rng = np.random.default_rng(5)
    host = torch.empty(n * length, dtype=torch.uint8).pin_memory()
    host.numpy()[:] = rng.integers(0, 256, size=n * length, dtype=np.uint8)
    off = np.arange(n, dtype=np.uint64) * length
    got = multi.digest_packed(host, off, np.full(n, length, dtype=np.uint64))
    hp = host.numpy()
    for i in list(range(0, n, 97)) + [n - 1] + [n * r // len(multi.hashers) for r in range(1, len(multi.hashers))]:
        assert bytes(got[i]) == hashlib.sha256(hp[i * length:(i + 1) * length]).digest(), i
