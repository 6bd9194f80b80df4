"""Content-addressed object store: a drop-in for the reference's ``Storage``.

Mirrors ``/root/reference/src/code_interpreter/services/storage.py:34-101`` --
same constructor, same six methods, same error behaviour -- with one change the
reference's own docstring (storage.py:36) asks for: the id of an object is the
SHA-256 of its bytes (64 lowercase hex characters, the shape
``secrets.token_hex(32)`` has at storage.py:52), computed on the GPU in batches.

Consequences of content-derived ids:
* the id is known only when the last byte has arrived, so ``writer()`` spools
  to a temporary file and renames it to ``<dir>/<sha256hex>`` on close;
  ``.hash`` is set at close, which is when the reference's caller reads it
  (services/kubernetes_code_executor.py:146);
* equal content gives equal ids: a second write of the same bytes is a no-op.

* ``delete`` removes the file, as in the reference; because equal content now shares
  one file, ``Storage(..., refcount=True)`` keeps a per-object holder count instead
  (one ``write`` = one holder) and only the last ``delete`` unlinks.

With the GPU hasher a streaming writer keeps nothing in memory: chunks go to the spool file and
the engine reads that file itself at close (``cas_hash_files``: pread into pinned staging).

Concurrent finalisations (the ``asyncio.gather`` over changed files at
kubernetes_code_executor.py:149-154) are coalesced into one GPU batch by
``_Batcher``; the blocking CUDA call runs in a worker thread, off the event loop.
Additive API: ``hash_many`` and ``write_many`` take whole batches.
"""
from __future__ import annotations

import asyncio
import os
import re
import secrets
from contextlib import asynccontextmanager
from typing import AsyncIterator, Protocol, Sequence

from anyio import Path
from pydantic import validate_call

from .hasher import GpuHasher
from .validation import HASH_PATTERN, Hash

_ID_RE = re.compile(HASH_PATTERN)


class ObjectReader(Protocol):
    async def read(self, size: int = -1) -> bytes: ...


class ObjectWriter(Protocol):
    hash: str

    async def write(self, data: bytes) -> None: ...


class _Batcher:
    """Collects id requests that arrive within ``window`` seconds of each other (or until
    ``max_batch`` are waiting) and hashes them in one engine call.  A request is either
    bytes in memory (``hash``) or a file the engine reads itself (``hash_path``, only with
    a hasher that offers ``hash_files``); a mixed batch costs one call per kind."""

    def __init__(self, hasher, window: float, max_batch: int):
        self._hasher = hasher
        self._window = window
        self._max_batch = max_batch
        self._pending: list[tuple[bytes | str, bool, asyncio.Future]] = []  # (payload, is_path, future)
        self._timer: asyncio.TimerHandle | None = None
        self._lock: asyncio.Lock | None = None
        self.batches = 0  # number of GPU batches issued (observable in tests)

    @property
    def reads_files(self) -> bool:
        return hasattr(self._hasher, "hash_files")

    async def hash(self, data: bytes) -> str:
        return await self._submit(data, False)

    async def hash_path(self, path: str) -> str:
        return await self._submit(path, True)

    async def _submit(self, payload, is_path: bool) -> str:
        loop = asyncio.get_running_loop()
        fut: asyncio.Future = loop.create_future()
        self._pending.append((payload, is_path, fut))
        if len(self._pending) >= self._max_batch:
            self._fire(loop)
        elif self._timer is None:
            self._timer = loop.call_later(self._window, self._fire, loop)
        return await fut

    def _fire(self, loop: asyncio.AbstractEventLoop) -> None:
        if self._timer is not None:
            self._timer.cancel()
            self._timer = None
        batch, self._pending = self._pending, []
        if batch:
            loop.create_task(self._run(batch))

    def _hash_batch(self, batch) -> list[str]:
        """Worker thread: one engine call per kind of payload, ids back in request order."""
        ids: list = [None] * len(batch)
        for want_path, fn in ((False, getattr(self._hasher, "hash_many")), (True, getattr(self._hasher, "hash_files", None))):
            idx = [i for i, (_, is_path, _f) in enumerate(batch) if is_path == want_path]
            if idx:
                for i, object_id in zip(idx, fn([batch[i][0] for i in idx])):
                    ids[i] = object_id
        return ids

    async def _run(self, batch) -> None:
        if self._lock is None:
            self._lock = asyncio.Lock()
        try:
            async with self._lock:  # one batch on the device context at a time
                self.batches += 1
                ids = await asyncio.get_running_loop().run_in_executor(None, self._hash_batch, batch)
            for (_, _p, fut), object_id in zip(batch, ids):
                if not fut.done():
                    fut.set_result(object_id)
        except BaseException as e:  # noqa: BLE001 - an engine failure must reach every waiter
            for _, _p, fut in batch:
                if not fut.done():
                    fut.set_exception(e if isinstance(e, Exception) else RuntimeError(str(e)))
            if not isinstance(e, Exception):
                raise


class _SpoolWriter:
    """What ``Storage.writer()`` yields: ``write`` appends a chunk, ``hash`` is set on close.
    Chunks are kept in memory only when the hasher cannot read the spooled file itself."""

    def __init__(self, file, keep_chunks: bool):
        self._file = file
        self._chunks: list[bytes] | None = [] if keep_chunks else None
        self.hash: str = ""

    async def write(self, data: bytes) -> None:
        data = bytes(data)
        if self._chunks is not None:
            self._chunks.append(data)
        await self._file.write(data)

    def _content(self) -> bytes:
        return self._chunks[0] if len(self._chunks) == 1 else b"".join(self._chunks)


class Storage:
    """
    Storage is a collection of objects. Objects consist of binary data and are identified by their SHA-256 hash.

    Backed by the filesystem: each object is a file named by its id under ``storage_path``.
    """

    def __init__(self, storage_path: str, hasher=None, device: int = 0, batch_window: float = 0.002,
                 max_batch: int = 4096, refcount: bool = False):
        self.storage_path = Path(storage_path)
        # No fallback: if the CUDA library or a B200 is missing this raises.
        self._hasher = hasher if hasher is not None else GpuHasher(device)
        self._batcher = _Batcher(self._hasher, batch_window, max_batch)
        self._refcount = refcount
        self._ref_lock: asyncio.Lock | None = None
        self._dir_ready = False

    # -- holder counts (only with refcount=True) -------------------------------------------
    def _refs_path(self, object_id: str) -> Path:
        return self.storage_path / f".refs-{object_id}"  # '.' cannot occur in an id (validation.Hash)

    async def _read_refs(self, object_id: str) -> int:
        try:
            return int((await self._refs_path(object_id).read_text()).strip() or 0)
        except (FileNotFoundError, ValueError):
            return 1 if await (self.storage_path / object_id).exists() else 0

    def _lock_refs(self) -> asyncio.Lock:
        if self._ref_lock is None:
            self._ref_lock = asyncio.Lock()
        return self._ref_lock

    @property
    def hasher(self):
        return self._hasher

    @property
    def gpu_batches(self) -> int:
        return self._batcher.batches

    def _publish_sync(self, tmp: str, object_id: str) -> None:
        """Worker thread: give the spooled file its content-derived name (or drop it if that
        content is already stored) and, with refcount=True, count one more holder."""
        target = os.path.join(str(self.storage_path), object_id)
        present = os.path.exists(target)
        if present:
            if os.path.exists(tmp):
                os.unlink(tmp)  # same content already stored
        else:
            os.rename(tmp, target)  # atomic within the store directory
        if self._refcount:
            refs = os.path.join(str(self.storage_path), f".refs-{object_id}")
            try:
                holders = int(open(refs).read().strip() or 0)
            except (FileNotFoundError, ValueError):
                holders = 1 if present else 0
            with open(refs, "w") as f:
                f.write(str(holders + 1))

    async def _publish(self, tmp: Path, object_id: str) -> None:
        loop = asyncio.get_running_loop()
        if self._refcount:
            async with self._lock_refs():
                await loop.run_in_executor(None, self._publish_sync, str(tmp), object_id)
        else:
            await loop.run_in_executor(None, self._publish_sync, str(tmp), object_id)

    @asynccontextmanager
    async def writer(self) -> AsyncIterator[ObjectWriter]:
        """
        Async context manager for writing a new object to the storage.

        The final hash can be retrieved by using the `.hash` attribute once the
        context has been left (the id depends on every byte written).
        """
        if not self._dir_ready:  # one hop, the first time only
            await self.storage_path.mkdir(parents=True, exist_ok=True)
            self._dir_ready = True
        tmp = self.storage_path / f".tmp-{secrets.token_hex(16)}"
        try:
            file = await tmp.open("wb")
        except FileNotFoundError:  # the directory went away since the first writer: make it again
            await self.storage_path.mkdir(parents=True, exist_ok=True)
            file = await tmp.open("wb")
        from_file = self._batcher.reads_files  # the engine preads the spooled file: O(1) memory per writer
        spool = _SpoolWriter(file, keep_chunks=not from_file)
        try:
            try:
                yield spool
            finally:
                await file.aclose()
            if from_file:
                spool.hash = await self._batcher.hash_path(str(tmp))
            else:
                spool.hash = await self._batcher.hash(spool._content())
            await self._publish(tmp, spool.hash)
        except BaseException:
            await tmp.unlink(missing_ok=True)
            raise

    async def write(self, data: bytes) -> str:
        """
        Writes the data to the storage and returns the hash of the object.
        """
        async with self.writer() as f:
            await f.write(data)
        return f.hash

    async def hash_many(self, objects: Sequence[bytes]) -> list[str]:
        """Ids of a batch of objects, one GPU batch, nothing stored."""
        if not objects:
            return []
        return await asyncio.get_running_loop().run_in_executor(None, self._hasher.hash_many, list(objects))

    def _store_batch_sync(self, objects: Sequence[bytes], ids: Sequence[str]) -> None:
        """Worker thread: write every object of a batch that is not stored yet (temp name, then
        rename), with plain os calls -- one thread hop for the batch instead of four per object."""
        root = str(self.storage_path)
        os.makedirs(root, exist_ok=True)
        for data, object_id in zip(objects, ids):
            target = os.path.join(root, object_id)
            present = os.path.exists(target)
            if not present:
                tmp = os.path.join(root, f".tmp-{secrets.token_hex(16)}")
                try:
                    with open(tmp, "wb") as f:
                        f.write(data)
                    os.rename(tmp, target)
                except BaseException:
                    if os.path.exists(tmp):
                        os.unlink(tmp)
                    raise
            if self._refcount:
                refs = os.path.join(root, f".refs-{object_id}")
                try:
                    holders = int(open(refs).read().strip() or 0)
                except (FileNotFoundError, ValueError):
                    holders = 1 if present else 0
                with open(refs, "w") as f:
                    f.write(str(holders + 1))

    async def write_many(self, objects: Sequence[bytes]) -> list[str]:
        """Stores a batch of objects and returns their ids: one GPU batch, one worker-thread
        hop for all the file writes."""
        objects = list(objects)
        ids = await self.hash_many(objects)
        loop = asyncio.get_running_loop()
        if self._refcount:
            async with self._lock_refs():  # holder counts are read-modify-write
                await loop.run_in_executor(None, self._store_batch_sync, objects, ids)
        else:
            await loop.run_in_executor(None, self._store_batch_sync, objects, ids)
        return ids

    @asynccontextmanager
    @validate_call
    async def reader(self, object_hash: Hash) -> AsyncIterator[ObjectReader]:
        """
        Async context manager that opens an object for reading.
        """
        target_file = self.storage_path / object_hash
        if not object_hash or not await target_file.exists():
            raise FileNotFoundError(f"File not found: {object_hash}")
        async with await target_file.open("rb") as f:
            yield f

    @validate_call
    async def read(self, object_hash: Hash) -> bytes:
        """
        Reads the object with the given hash and returns it.
        """
        async with self.reader(object_hash) as f:
            return await f.read()

    @validate_call
    async def exists(self, object_hash: Hash) -> bool:
        """
        Check if an object with the given hash exists in the storage.
        """
        return await (self.storage_path / object_hash).exists()

    @validate_call
    async def delete(self, object_hash: Hash) -> None:
        """
        Delete an object with the given hash from the storage.
        """
        file_path = self.storage_path / object_hash
        if not await file_path.exists():
            raise FileNotFoundError(f"File not found: {object_hash}")
        if self._refcount:
            async with self._lock_refs():
                holders = await self._read_refs(object_hash) - 1
                if holders > 0:  # another writer of the same content still holds it
                    await self._refs_path(object_hash).write_text(str(holders))
                    return
                await self._refs_path(object_hash).unlink(missing_ok=True)
                await file_path.unlink()
        else:
            await file_path.unlink()

    def _verify_sync(self, object_hashes: Sequence[str]) -> dict[str, bool]:
        root = str(self.storage_path)
        present = [h for h in object_hashes if os.path.isfile(os.path.join(root, h))]
        if hasattr(self._hasher, "hash_files"):
            ids = self._hasher.hash_files([os.path.join(root, h) for h in present]) if present else []
        else:
            blobs = []
            for h in present:
                with open(os.path.join(root, h), "rb") as f:
                    blobs.append(f.read())
            ids = self._hasher.hash_many(blobs) if blobs else []
        ok = {h: (h == got) for h, got in zip(present, ids)}
        return {h: ok.get(h, False) for h in object_hashes}

    async def verify(self, object_hashes: Sequence[str]) -> dict[str, bool]:
        """Re-hashes stored objects in one GPU batch: id -> (content still matches id).
        With a GPU hasher the engine reads the files itself (cas_hash_files: pread into
        pinned staging), so the bytes never pass through Python; the whole check is one
        worker-thread hop."""
        for h in object_hashes:
            if not _ID_RE.fullmatch(h):
                raise ValueError(f"not an object id: {h!r}")
        return await asyncio.get_running_loop().run_in_executor(None, self._verify_sync, list(object_hashes))

    async def verify_all(self) -> dict[str, bool]:
        """Integrity scrub of the whole store directory (temporary and holder-count files skipped)."""

        def scan() -> dict[str, bool]:
            root = str(self.storage_path)
            if not os.path.isdir(root):
                return {}
            names = sorted(e.name for e in os.scandir(root) if not e.name.startswith(".") and e.is_file())
            return self._verify_sync(names)

        return await asyncio.get_running_loop().run_in_executor(None, scan)
