"""The Storage drop-in on a B200: the reference's call patterns with real GPU hashing."""
import asyncio
import hashlib
import os

import pytest

from bee_code_interpreter_fs_b200 import MultiGpuHasher, Storage

pytestmark = pytest.mark.gpu


def test_storage_roundtrip_like_the_reference_e2e_test(tmp_path, hasher):
    """Shape of test_create_file_in_interpreter (reference test/e2e/test_http.py:47-85): an id
    returned for a written file is fed back and the content round-trips."""

    async def go():
        st = Storage(str(tmp_path), hasher=hasher)
        object_id = await st.write(b"Hello, World!")
        assert object_id == hashlib.sha256(b"Hello, World!").hexdigest()
        assert await st.exists(object_id)
        assert await st.read(object_id) == b"Hello, World!"
        async with st.reader(object_id) as r:
            assert await r.read() == b"Hello, World!"
        await st.delete(object_id)
        assert not await st.exists(object_id)
        with pytest.raises(FileNotFoundError, match=f"File not found: {object_id}"):
            await st.read(object_id)

    asyncio.run(go())


def test_gathered_writers_share_a_gpu_batch(tmp_path, hasher):
    async def go():
        st = Storage(str(tmp_path), hasher=hasher, batch_window=0.05)
        blobs = [os.urandom(1000 + 977 * i) for i in range(24)]

        async def download(blob):
            async with st.writer() as f:
                for k in range(0, len(blob), 4096):
                    await f.write(blob[k:k + 4096])
            return f.hash

        before = hasher.launch_count
        ids = await asyncio.gather(*(download(b) for b in blobs))
        assert ids == [hashlib.sha256(b).hexdigest() for b in blobs]
        assert st.gpu_batches == 1 and hasher.launch_count == before + 1
        many = await st.write_many(blobs)
        assert many == ids
        assert all((await st.verify(ids)).values())

    asyncio.run(go())


def test_hash_files_and_store_scrub(tmp_path, hasher):
    """cas_hash_files: the engine reads the files itself (pread into pinned staging)."""
    sizes = [0, 1, 55, 64, 4096, 100001, (5 << 20) + 3, 3 << 20]
    blobs = [os.urandom(n) for n in sizes]
    paths = []
    for i, b in enumerate(blobs):
        p = tmp_path / f"f{i}.bin"
        p.write_bytes(b)
        paths.append(str(p))
    dig, got_sizes = hasher.digest_files(paths)
    assert got_sizes.tolist() == sizes
    assert [bytes(d).hex() for d in dig] == [hashlib.sha256(b).hexdigest() for b in blobs]
    from bee_code_interpreter_fs_b200 import CasError

    with pytest.raises(CasError, match="cannot stat"):
        hasher.hash_files(paths + [str(tmp_path / "missing.bin")])

    async def go():
        st = Storage(str(tmp_path / "store"), hasher=hasher)
        ids = await st.write_many(blobs)
        assert all((await st.verify_all()).values())
        (tmp_path / "store" / ids[5]).write_bytes(b"bit rot")
        scrub = await st.verify_all()
        assert scrub[ids[5]] is False and sum(scrub.values()) == len(set(ids)) - 1

    asyncio.run(go())


def test_http_round_trip_on_the_gpu_store(tmp_path, hasher):
    """The reference's test_create_file_in_interpreter flow over the GPU-backed store."""
    from starlette.testclient import TestClient

    from bee_code_interpreter_fs_b200.service_harness import create_app

    app = create_app(Storage(str(tmp_path / "store"), hasher=hasher, batch_window=0.05))
    with TestClient(app) as client:
        script = b"for i in range(4):\n    open(f'o{i}.txt', 'w').write('x' * (i + 1) * 1000)\n"
        sid = client.put("/v1/files", files={"file": ("main.py", script)}).json()["hash"]
        assert sid == hashlib.sha256(script).hexdigest()
        before = hasher.launch_count
        r = client.post("/v1/execute", json={"source_file": "/workspace/main.py", "files": {"/workspace/main.py": sid}})
        assert r.status_code == 200 and r.json()["exit_code"] == 0
        assert r.json()["files"] == {f"/workspace/o{i}.txt": hashlib.sha256(b"x" * (i + 1) * 1000).hexdigest()
                                     for i in range(4)}
        assert hasher.launch_count == before + 1  # four changed files, one kernel launch
        got = client.get("/v1/files/" + r.json()["files"]["/workspace/o2.txt"])
        assert got.content == b"x" * 3000


def test_two_contexts_on_one_device_from_two_threads():
    """Contexts are independent (own streams, arena, staging): two host threads may hash at once."""
    import threading

    from bee_code_interpreter_fs_b200 import GpuHasher

    blobs = [[os.urandom(1000 + 37 * i) for i in range(400)] + [os.urandom(3 << 20)] for _ in range(2)]
    want = [[hashlib.sha256(b).hexdigest() for b in bl] for bl in blobs]
    got = [None, None]
    errors = []

    def run(k):
        try:
            with GpuHasher(0) as h:
                for _ in range(5):
                    got[k] = h.hash_many(blobs[k])
        except Exception as e:  # noqa: BLE001
            errors.append(e)

    threads = [threading.Thread(target=run, args=(k,)) for k in range(2)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    assert got == want


def test_multi_gpu_hasher_on_visible_devices():
    mh = MultiGpuHasher()
    blobs = [os.urandom(n) for n in (0, 5, 64, 100000, 3, 70000, 1 << 20)] * 3
    assert mh.hash_many(blobs) == [hashlib.sha256(b).hexdigest() for b in blobs]
    uniform = [os.urandom(4096) for _ in range(100)]
    assert mh.hash_many(uniform) == [hashlib.sha256(b).hexdigest() for b in uniform]
    mh.close()


# ------------------------------------------------- batched store / publish in the engine --
def test_store_many_writes_hashes_and_renames_in_one_call(tmp_path, hasher):
    """cas_store_host (the batch form of storage.py:44-63): ids are hashlib's, every object is a
    complete file under its id, nothing temporary is left, content that is already stored (or
    occurs twice in the batch) is reported and left alone."""
    import numpy as np

    root = tmp_path / "deep" / "store"  # created on demand, like writer()'s mkdir -p
    rng = np.random.default_rng(1)
    blobs = [rng.integers(0, 256, size=int(n), dtype=np.uint8).tobytes()
             for n in [0, 1, 55, 64, 65536, 100001, 3 << 20] + rng.integers(0, 200000, size=300).tolist()]
    blobs.append(blobs[4])  # the same content twice in one batch
    ids, status = hasher.store_many(root, blobs)
    assert ids == [hashlib.sha256(b).hexdigest() for b in blobs]
    assert sorted(os.listdir(root)) == sorted(set(ids))
    for object_id, blob in zip(ids, blobs):
        assert (root / object_id).read_bytes() == blob
    assert sorted(status.tolist()) == [0] * (len(blobs) - 1) + [1]
    assert {status[4], status[-1]} == {0, 1}  # one of the twins created the file, the other found it
    (root / ids[5]).write_bytes(b"kept as it is")  # an existing object is never replaced
    ids2, status2 = hasher.store_many(root, blobs[:10])
    assert ids2 == ids[:10] and status2.tolist() == [1] * 10
    assert (root / ids[5]).read_bytes() == b"kept as it is"
    from bee_code_interpreter_fs_b200 import CasError

    (tmp_path / "plainfile").write_bytes(b"x")
    with pytest.raises(CasError, match="cannot create the store directory"):
        hasher.store_many(tmp_path / "plainfile" / "store", blobs[:3])


def test_a_final_name_never_shows_partial_content(tmp_path, hasher):
    """Crash safety of the store: while a batch is being stored, whatever is visible under an id
    is the whole object (rename after the last byte), so a crash at any point leaves no truncated
    object behind -- only '.tmp-' files, which no reader mistakes for objects."""
    import threading

    import numpy as np

    root = tmp_path / "store"
    rng = np.random.default_rng(2)
    blobs = [rng.integers(0, 256, size=1 << 20, dtype=np.uint8).tobytes() for _ in range(200)]
    want = {hashlib.sha256(b).hexdigest() for b in blobs}
    seen, bad, done = set(), [], threading.Event()

    def watch():
        while not done.is_set():
            try:
                names = os.listdir(root)
            except FileNotFoundError:
                continue
            for name in names:
                if name.startswith(".") or name in seen:
                    continue
                with open(root / name, "rb") as f:
                    if hashlib.sha256(f.read()).hexdigest() != name:
                        bad.append(name)
                seen.add(name)

    t = threading.Thread(target=watch)
    t.start()
    try:
        ids, status = hasher.store_many(root, blobs)
    finally:
        done.set()
        t.join()
    assert not bad and set(ids) == want and status.tolist() == [0] * len(blobs)
    assert sorted(os.listdir(root)) == sorted(want)


def test_publish_files_fails_only_the_file_that_is_missing(tmp_path, hasher):
    """cas_publish_files (the close of Storage.writer for a batch): spool files get their final
    names; one that vanished is reported for itself and the rest are still published."""
    root = tmp_path / "store"
    spool = tmp_path / "spool"
    spool.mkdir()
    blobs = [os.urandom(n) for n in (10, 0, 70000, 4096, 2 << 20)]
    paths = []
    for i, b in enumerate(blobs):
        (spool / f".tmp-{i}").write_bytes(b)
        paths.append(str(spool / f".tmp-{i}"))
    paths.insert(2, str(spool / ".tmp-gone"))
    ids, sizes, status = hasher.publish_files(root, paths)
    assert status.tolist() == [0, 0, -5, 0, 0, 0]
    want = [hashlib.sha256(b).hexdigest() for b in blobs]
    assert ids[:2] + ids[3:] == want and ids[2] == "00" * 32
    assert sizes.tolist() == [10, 0, 0, 70000, 4096, 2 << 20]
    assert sorted(os.listdir(root)) == sorted(want) and os.listdir(spool) == []
    # hash only (no directory): nothing moves, statuses are per file
    ids3, _, status3 = hasher.publish_files(None, [str(root / want[0]), str(root / "nope")])
    assert ids3[0] == want[0] and status3.tolist() == [0, -5] and (root / want[0]).exists()


def test_storage_on_the_engine_store_path_counts_holders(tmp_path, hasher):
    """Storage over the C batch calls: dedup, holder counts, delete, scrub, and a reserve that
    makes the first batch allocation-free."""

    async def go():
        st = Storage(str(tmp_path / "store"), hasher=hasher, reserve_bytes=64 << 20)
        blobs = [os.urandom(1000 * (i + 1)) for i in range(50)]
        ids = await st.write_many(blobs)
        assert ids == [hashlib.sha256(b).hexdigest() for b in blobs]
        again = await asyncio.gather(*(st.write(b) for b in blobs[:5]))
        assert list(again) == ids[:5]
        async with st.writer() as w:
            await w.write(blobs[0][:500])
            await w.write(blobs[0][500:])
        assert w.hash == ids[0]
        await st.delete(ids[0])
        await st.delete(ids[0])
        assert await st.read(ids[0]) == blobs[0]  # three holders, two deletes
        await st.delete(ids[0])
        assert not await st.exists(ids[0])
        await st.delete(ids[7])
        assert not await st.exists(ids[7])  # written once: the first delete removes it
        scrub = await st.verify_all()
        assert len(scrub) == 48 and all(scrub.values())
        assert not [n for n in os.listdir(tmp_path / "store") if n.startswith(".tmp-")]

    asyncio.run(go())


def test_scrub_in_the_engine_tells_objects_from_everything_else(tmp_path, hasher):
    """cas_verify_names: per name ok / bad / not a file; Storage.verify_all skips what cannot be
    an object (sub-directories, dot files) and flags what is wrong under an id."""
    root = tmp_path / "store"
    blobs = [os.urandom(n) for n in (0, 1, 4096, 65536, 300001)]
    ids, _ = hasher.store_many(root, blobs)
    (root / "subdir").mkdir()
    (root / ".refs-something").write_text("2")
    (root / ("a" * 64)).write_bytes(b"not what the name says")
    (root / "short-name").write_bytes(b"x")
    got = hasher.verify_names(root, ids + ["subdir", "a" * 64, "short-name", "f" * 64])
    assert got.tolist() == [1] * 5 + [2, 0, 0, 0]

    async def go():
        st = Storage(str(root), hasher=hasher)
        scrub = await st.verify_all()
        assert scrub == {**{i: True for i in ids}, "a" * 64: False, "short-name": False}
        assert await st.verify(ids[:2] + ["0" * 64]) == {ids[0]: True, ids[1]: True, "0" * 64: False}

    asyncio.run(go())


def test_h2d_gate_lets_contexts_take_turns_and_changes_no_digest(hasher):
    """cas_set_h2d_gate: two contexts that name the same one-slot gate copy their chunks in turn;
    the digests are the same, the slot always comes back (the second round would otherwise crawl
    along on one-second time-outs), and removing the gate restores the ungated path."""
    import threading
    import time

    import numpy as np
    import torch

    from bee_code_interpreter_fs_b200 import GpuHasher, dist as cdist

    name = f"/b200cas-test-gate-{os.getpid()}"
    n, length = 3000, 65536
    host = torch.empty(n * length, dtype=torch.uint8).pin_memory()
    host.numpy()[:] = np.random.default_rng(8).integers(0, 256, size=n * length, dtype=np.uint8)
    off = np.arange(n, dtype=np.uint64) * length
    lens = np.full(n, length, dtype=np.uint64)
    want = hasher.digest_packed(host, off, lens)
    assert bytes(want[7]) == hashlib.sha256(host.numpy()[7 * length:8 * length]).digest()
    got, errors = [None, None], []

    def run(k):
        try:
            with GpuHasher(0) as h:
                h.set_option(3, 16 << 20)  # CAS_OPT_WAVE_BYTES: a dozen gated chunks per call
                h.set_h2d_gate(name, 1)
                for _ in range(3):
                    got[k] = h.digest_packed(host, off, lens)
                h.set_h2d_gate(None)
                got[k] = h.digest_packed(host, off, lens)
        except Exception as e:  # noqa: BLE001
            errors.append(e)

    try:
        t0 = time.perf_counter()
        threads = [threading.Thread(target=run, args=(k,)) for k in range(2)]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        assert not errors, errors
        assert (got[0] == want).all() and (got[1] == want).all()
        assert time.perf_counter() - t0 < 10.0  # no slot was lost: nothing waited out a time-out
    finally:
        cdist.remove_h2d_gate(name)
