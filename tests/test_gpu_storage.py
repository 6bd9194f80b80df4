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
