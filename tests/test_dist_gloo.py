"""The N>1 host path on CPU: two gloo ranks shard a batch by object, hash their shards
(the oracle stands in for the device), and the gathered digests equal the unsharded answer."""
import os
import socket

import numpy as np
import pytest
import torch.multiprocessing as mp


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, mixed: bool, q):
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    import torch.distributed as dist

    from bee_code_interpreter_fs_b200 import dist as cdist
    from bee_code_interpreter_fs_b200 import shard
    from oracle import sha256_oracle as oracle

    cdist.init("gloo")
    try:
        n = 101
        lengths = [(i * 977) % 5000 for i in range(n)] if mixed else [4096] * n
        objs = [oracle.synth_object(7, i, L) for i, L in enumerate(lengths)]
        parts = shard.assign(lengths, world)
        mine = parts[rank]
        local = np.frombuffer(b"".join(oracle.hashlib_digests([objs[i] for i in mine])), dtype=np.uint8).reshape(-1, 32)
        cdist.barrier()
        gathered = cdist.gather_digests(local, [len(p) for p in parts])
        full = shard.scatter_back(parts, np.split(gathered, np.cumsum([len(p) for p in parts])[:-1]), n)
        want = np.frombuffer(b"".join(oracle.hashlib_digests(objs)), dtype=np.uint8).reshape(-1, 32)
        slowest = cdist.max_over_ranks(1.0 + rank)
        total = cdist.sum_over_ranks(float(len(mine)))
        q.put((rank, bool((full == want).all()), slowest, total))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("mixed", [False, True])
def test_two_rank_shard_and_gather(mixed):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, mixed, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, ok, slowest, total in results:
        assert ok, f"rank {rank}: gathered digests differ from the unsharded oracle"
        assert slowest == 2.0 and total == 101.0
