"""Host-side logic that needs no GPU: sharding, id/path types, the Storage drop-in's
file semantics (with the oracle standing in for the device, as tests may)."""
import asyncio
import hashlib
import os
import re

import numpy as np
import pytest
from pydantic import TypeAdapter, ValidationError

from bee_code_interpreter_fs_b200 import shard
from bee_code_interpreter_fs_b200.storage import Storage
from bee_code_interpreter_fs_b200.validation import AbsolutePath, Hash


def run(coro):
    return asyncio.run(coro)


# ----------------------------------------------------------------- id types --
def test_hash_and_path_types_match_reference(storage_behaviour):
    ha, pa = TypeAdapter(Hash), TypeAdapter(AbsolutePath)

    def ok(adapter, v):
        try:
            adapter.validate_python(v)
            return True
        except ValidationError:
            return False

    for v, want in storage_behaviour["hash_accepts"].items():
        assert ok(ha, v) == want, v
    for v, want in storage_behaviour["path_accepts"].items():
        assert ok(pa, v) == want, v


# ------------------------------------------------------------------ sharding --
def test_nblocks():
    assert shard.nblocks([0, 55, 56, 64, 4096, 65536, 1 << 30]).tolist() == [1, 1, 2, 2, 65, 1025, 16777217]


def test_contiguous_ranges_cover():
    for n in (0, 1, 7, 1 << 20):
        for w in (1, 2, 4, 8):
            r = shard.contiguous_ranges(n, w)
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(r, r[1:]))
            assert max(hi - lo for lo, hi in r) - min(hi - lo for lo, hi in r) <= 1


def test_lpt_balances_by_compressions():
    rng = np.random.default_rng(0xB200)
    lengths = np.floor(2 ** rng.uniform(10, 26, size=4096)).astype(np.uint64)
    for w in (2, 4, 8):
        parts = shard.assign(lengths, w)
        assert sorted(np.concatenate(parts).tolist()) == list(range(4096))
        loads = [int(shard.nblocks(lengths[p]).sum()) for p in parts]
        assert max(loads) <= 1.02 * (sum(loads) / w) + int(shard.nblocks(lengths.max()))


def test_scatter_back_roundtrip():
    lengths = np.array([5, 100000, 3, 70000, 64, 1], dtype=np.uint64)
    parts = shard.assign(lengths, 3)
    digs = [np.tile((p % 251).astype(np.uint8)[:, None], (1, 32)) for p in parts]
    out = shard.scatter_back(parts, digs, 6)
    assert out[:, 0].tolist() == [0, 1, 2, 3, 4, 5]
    with pytest.raises(ValueError):
        shard.scatter_back(parts[:-1], digs[:-1], 6)


# ------------------------------------------------------------------- Storage --
def test_storage_matches_reference_behaviour(tmp_path, oracle_hasher, storage_behaviour):
    ref = storage_behaviour["storage"]

    async def go():
        st = Storage(str(tmp_path / "nested" / "store"), hasher=oracle_hasher)
        rec = {}
        i1 = await st.write(b"abc")
        i2 = await st.write(b"abc")
        rec["write_returns_str"] = isinstance(i1, str)
        rec["id_matches_regex"] = bool(re.fullmatch(ref["id_regex"], i1))
        rec["file_named_by_id"] = os.path.isfile(tmp_path / "nested" / "store" / i1)
        rec["read_roundtrip"] = (await st.read(i1)) == b"abc"
        rec["exists_true"] = await st.exists(i1)
        rec["exists_missing"] = await st.exists("0" * 64)
        async with st.writer() as w:
            await w.write(b"he")
            await w.write(b"llo")
        rec["writer_hash_after_close_is_str"] = isinstance(w.hash, str)
        rec["writer_chunks_concatenate"] = (await st.read(w.hash)) == b"hello"
        async with st.reader(i1) as r:
            rec["reader_read_size"] = (await r.read(2)).decode()
            rec["reader_read_rest"] = (await r.read()).decode()
        for name, fn in (("read", st.read), ("delete", st.delete)):
            try:
                await fn("0" * 64)
                rec[f"{name}_missing"] = None
            except Exception as e:  # noqa: BLE001
                rec[f"{name}_missing"] = {"type": type(e).__name__, "msg": str(e)}
        try:
            async with st.reader("0" * 64):
                pass
        except Exception as e:  # noqa: BLE001
            rec["reader_missing"] = {"type": type(e).__name__, "msg": str(e)}
        for name, fn in (("read", st.read), ("exists", st.exists), ("delete", st.delete)):
            try:
                await fn("a/b")
                rec[f"{name}_bad_id"] = None
            except Exception as e:  # noqa: BLE001
                rec[f"{name}_bad_id"] = {"type": type(e).__name__}
        await st.delete(i1)
        # "abc" was written twice: in the reference that is two objects, here two holders of one
        assert await st.read(i2) == b"abc"
        await st.delete(i2)
        rec["exists_after_delete"] = await st.exists(i1)
        rec["empty_object_roundtrip"] = (await st.read(await st.write(b""))) == b""
        # the deliberate difference: ids are the content's SHA-256, hence repeatable
        assert i1 == i2 == hashlib.sha256(b"abc").hexdigest()
        assert w.hash == hashlib.sha256(b"hello").hexdigest()
        return rec

    rec = run(go())
    for key, got in rec.items():
        assert got == ref[key], key
    assert ref["reference_ids_are_random"] and not ref["reference_id_is_sha256"]


def test_storage_leaves_no_temp_files_and_dedups(tmp_path, oracle_hasher):
    async def go():
        st = Storage(str(tmp_path), hasher=oracle_hasher)
        ids = await asyncio.gather(*(st.write(b"same") for _ in range(5)))
        assert len(set(ids)) == 1
        with pytest.raises(RuntimeError):
            async with st.writer() as w:
                await w.write(b"partial")
                raise RuntimeError("caller failed mid-stream")
        return ids[0]

    object_id = run(go())
    # one object, plus the count of its five holders; no temporary file survives
    assert sorted(os.listdir(tmp_path)) == [f".refs-{object_id}", object_id]
    assert (tmp_path / f".refs-{object_id}").read_text() == "5"


def test_concurrent_finalisations_coalesce_into_one_batch(tmp_path, oracle_hasher):
    """The executor's download leg gathers over changed files
    (kubernetes_code_executor.py:149-154): those writers must share a GPU batch."""

    async def go():
        st = Storage(str(tmp_path), hasher=oracle_hasher, batch_window=0.05)

        async def download(i):
            async with st.writer() as f:
                for chunk in (b"chunk-%d-" % i, b"x" * i):
                    await f.write(chunk)
            return f.hash

        ids = await asyncio.gather(*(download(i) for i in range(16)))
        return st, ids

    st, ids = run(go())
    assert st.gpu_batches == 1 and oracle_hasher.calls == 1
    for i, object_id in enumerate(ids):
        assert object_id == hashlib.sha256(b"chunk-%d-" % i + b"x" * i).hexdigest()


def test_streaming_writers_are_hashed_from_their_spool_files(tmp_path, oracle_file_hasher):
    """With a hasher that reads files (GpuHasher.hash_files) a writer keeps no chunk in memory."""

    async def go():
        st = Storage(str(tmp_path), hasher=oracle_file_hasher, batch_window=0.05)

        async def download(i):
            async with st.writer() as f:
                assert f._chunks is None
                for k in range(3):
                    await f.write(b"%d-%d;" % (i, k) * 1000)
            return f.hash

        ids = await asyncio.gather(*(download(i) for i in range(8)), st.hash_many([b"in memory"]))
        return st, ids

    st, ids = run(go())
    assert oracle_file_hasher.file_calls == 1 and st.gpu_batches == 1  # eight spool files, one engine call
    for i in range(8):
        want = hashlib.sha256(b"".join(b"%d-%d;" % (i, k) * 1000 for k in range(3))).hexdigest()
        assert ids[i] == want and (tmp_path / want).read_bytes().startswith(b"%d-0;" % i)
    assert ids[8] == [hashlib.sha256(b"in memory").hexdigest()]
    assert sorted(os.listdir(tmp_path)) == sorted(ids[:8])


def test_group_commit_batches_and_per_request_errors(tmp_path, oracle_hasher):
    """A lone writer is not delayed; writers that queue behind a running batch go out together;
    max_batch caps a batch; a failure while storing one object reaches that caller only."""
    import time

    class SlowHasher:
        def __init__(self):
            self.sizes = []

        def hash_many(self, objs):
            self.sizes.append(len(objs))
            time.sleep(0.05)
            return [hashlib.sha256(o).hexdigest() for o in objs]

    async def go():
        slow = SlowHasher()
        st = Storage(str(tmp_path / "a"), hasher=slow, max_batch=8)
        t0 = time.perf_counter()
        await st.write(b"alone")
        assert time.perf_counter() - t0 < 0.5 and slow.sizes == [1]
        ids = await asyncio.gather(*(st.write(b"obj-%d" % i) for i in range(20)))
        assert ids == [hashlib.sha256(b"obj-%d" % i).hexdigest() for i in range(20)]
        assert sum(slow.sizes) == 21 and max(slow.sizes) <= 8 and len(slow.sizes) <= 6

        st2 = Storage(str(tmp_path / "b"), hasher=oracle_hasher)
        real = st2._store_sync

        def flaky(data, object_id):
            if data == b"bad":
                raise OSError("disk full")
            real(data, object_id)

        st2._store_sync = flaky
        results = await asyncio.gather(st2.write(b"ok"), st2.write(b"bad"), return_exceptions=True)
        assert results[0] == hashlib.sha256(b"ok").hexdigest() and isinstance(results[1], OSError)
        assert await st2.exists(results[0]) and st2.gpu_batches == 1

    run(go())


def test_write_many_and_verify(tmp_path, oracle_hasher):
    async def go():
        st = Storage(str(tmp_path), hasher=oracle_hasher)
        objs = [os.urandom(n) for n in (0, 1, 100, 5000)]
        ids = await st.write_many(objs)
        assert ids == [hashlib.sha256(o).hexdigest() for o in objs]
        assert [await st.read(i) for i in ids] == objs
        (tmp_path / ids[2]).write_bytes(b"corrupted")
        ok = await st.verify(ids + ["f" * 64])
        assert ok == {ids[0]: True, ids[1]: True, ids[2]: False, ids[3]: True, "f" * 64: False}

    run(go())


def test_refcounted_delete_keeps_shared_content(tmp_path, oracle_hasher):
    """Equal content shares one file; by default only the last holder's delete unlinks it, so one
    client's DELETE /v1/files/{hash} cannot take the object away from another uploader."""

    async def go():
        st = Storage(str(tmp_path), hasher=oracle_hasher)
        a = await st.write(b"shared")
        b = await st.write(b"shared")
        await st.delete(a)
        assert a == b and await st.exists(a) and await st.read(a) == b"shared"  # written twice, deleted once
        (c,) = await st.write_many([b"shared"])
        async with st.writer() as w:
            await w.write(b"sha")
            await w.write(b"red")
        assert a == c == w.hash
        await st.delete(a)
        await st.delete(a)
        assert await st.exists(a)
        await st.delete(a)
        assert not await st.exists(a)
        with pytest.raises(FileNotFoundError):
            await st.delete(a)
        assert await st.verify_all() == {}
        once = await st.write(b"written once")  # the common case keeps no count file
        assert sorted(os.listdir(tmp_path)) == [once]
        await st.delete(once)
        # refcount=False: the reference's literal semantics, the first delete removes the object
        plain = Storage(str(tmp_path / "plain"), hasher=oracle_hasher, refcount=False)
        x = await plain.write(b"shared")
        await plain.write(b"shared")
        await plain.delete(x)
        assert not await plain.exists(x)

    run(go())
    assert [n for n in os.listdir(tmp_path) if n != "plain"] == []


def test_cancelled_and_broken_requests_do_not_fail_their_batch(tmp_path, oracle_file_hasher):
    """One client going away (its coroutine cancelled, its spool file removed) must not turn other
    clients' Execute calls into errors: the cancelled request is dropped before the batch is built,
    and a spool file that vanished fails its own writer only."""

    async def go():
        st = Storage(str(tmp_path), hasher=oracle_file_hasher, batch_window=0.1)

        async def download(data):
            async with st.writer() as f:
                await f.write(data)
            return f.hash

        tasks = [asyncio.ensure_future(download(b"file-%d" % i)) for i in range(6)]
        doomed = asyncio.ensure_future(st.write(b"never stored"))
        await asyncio.sleep(0.03)  # everything is queued, the batch window is still open
        tasks[2].cancel()
        doomed.cancel()
        victim = next(n for n in os.listdir(tmp_path) if n.startswith(".tmp-"))
        os.unlink(tmp_path / victim)  # some other writer's spool file disappears under it
        results = await asyncio.gather(*tasks, doomed, return_exceptions=True)
        return st, results

    st, results = run(go())
    assert isinstance(results[2], asyncio.CancelledError) and isinstance(results[6], asyncio.CancelledError)
    good = [r for r in results if isinstance(r, str)]
    failed = [r for r in results if isinstance(r, Exception) and not isinstance(r, asyncio.CancelledError)]
    assert len(good) + len(failed) == 5 and len(failed) <= 1 and all(isinstance(f, OSError) for f in failed)
    assert set(good) <= {hashlib.sha256(b"file-%d" % i).hexdigest() for i in range(6)}
    assert hashlib.sha256(b"never stored").hexdigest() not in os.listdir(tmp_path)
    assert sorted(os.listdir(tmp_path)) == sorted(good)


def test_engine_failure_reaches_the_caller(tmp_path):
    class Broken:
        def hash_many(self, objs):
            raise RuntimeError("device lost")

    async def go():
        st = Storage(str(tmp_path), hasher=Broken())
        with pytest.raises(RuntimeError, match="device lost"):
            await st.write(b"abc")

    run(go())
    assert os.listdir(tmp_path) == []


# ------------------------------------------------- rank -> GPU order, gate names, addresses --
def test_gpu_order_uses_the_measured_map_only_when_it_fits_the_node(monkeypatch):
    from bee_code_interpreter_fs_b200 import dist as cdist

    monkeypatch.delenv("B200CAS_GPU_ORDER", raising=False)
    order = cdist.gpu_order(8)  # the committed map was measured on an 8-GPU node
    assert sorted(order) == list(range(8))
    assert {g // 4 for g in order[:2]} == {0, 1} and {g // 4 for g in order[:4]} == {0, 1}  # both uplink halves early
    assert cdist.gpu_order(4) == [0, 1, 2, 3] and cdist.gpu_order(1) == [0]  # another node shape: identity
    monkeypatch.setenv("B200CAS_GPU_ORDER", "3,2,1,0")
    assert cdist.gpu_order(4) == [3, 2, 1, 0] and cdist.gpu_for_rank(1, 4) == 2
    assert cdist.gpu_order(8) == list(range(8))  # an override that does not name every GPU is ignored
    monkeypatch.setenv("B200CAS_GPU_ORDER", "nonsense")
    assert cdist.gpu_order(2) == [0, 1]


def test_gate_names_group_ranks_by_uplink_and_by_launch():
    from bee_code_interpreter_fs_b200 import dist as cdist

    a, b, c = cdist.h2d_gate_for(0, 8), cdist.h2d_gate_for(3, 8), cdist.h2d_gate_for(5, 8)
    assert a == b and a != c and a[1] == c[1] == 2
    assert a[0].startswith("/b200cas-h2d-") and str(os.getppid()) in a[0]
    assert cdist.h2d_gate_for(0, 4) is None  # the map describes another node shape
    cdist.remove_h2d_gate("/b200cas-h2d-does-not-exist")  # never raises


def test_addresses_of_bytes_objects_match_ctypes():
    import ctypes

    import numpy as np

    from bee_code_interpreter_fs_b200 import hasher

    objs = [b"", b"a", os.urandom(100), os.urandom(70000), b"x" * 4096]
    addr, length, keep = hasher._addresses(objs)
    assert length.tolist() == [len(o) for o in objs]
    for a, o in zip(addr.tolist(), objs):
        assert ctypes.string_at(a, len(o)) == o
    if hasher._BYTES_DATA_OFFSET:
        want = np.frombuffer((ctypes.c_char_p * len(objs))(*objs), dtype=np.uint64, count=len(objs))
        assert (addr == want).all()
    mixed, mlen, _ = hasher._addresses([bytearray(b"abc"), memoryview(b"defg"), np.arange(5, dtype=np.uint8)])
    assert mlen.tolist() == [3, 4, 5] and ctypes.string_at(int(mixed[1]), 4) == b"defg"
