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
* equal content gives equal ids: a second write of the same bytes is a no-op;
* ``delete`` removes the file, as in the reference; because equal content now shares
  one file, ``Storage(..., refcount=True)`` keeps a per-object holder count instead
  (one ``write`` = one holder) and only the last ``delete`` unlinks.

With the GPU hasher a streaming writer keeps nothing in memory: chunks go to the spool file and
the engine reads that file itself at close (``cas_hash_files``: pread into pinned staging).

Concurrent finalisations (the ``asyncio.gather`` over changed files at
kubernetes_code_executor.py:149-154) are coalesced by ``_Batcher`` (group commit: no added
latency for a lone writer, one engine call for everything that queued up behind a running
one; ``batch_window`` optionally waits to collect more).  The blocking CUDA call runs in a
worker thread, off the event loop, and the same worker gives each object its final name, so
a batch of N writes costs one thread hop, not N.
Additive API: ``hash_many``, ``write_many``, ``verify``, ``verify_all``.
"""
from __future__ import annotations

import asyncio
import os
import re
import secrets
import threading
from contextlib import asynccontextmanager
from functools import partial
from typing import AsyncIterator, Callable, Protocol, Sequence

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


class _Request:
    __slots__ = ("payload", "is_path", "future", "on_done")

    def __init__(self, payload, is_path: bool, future: asyncio.Future, on_done: Callable[[str], None] | None):
        self.payload, self.is_path, self.future, self.on_done = payload, is_path, future, on_done


class _Batcher:
    """Group commit for id requests.  A request that arrives while the engine is idle starts a
    batch at once (after ``window`` seconds if a window is set); requests that arrive while a
    batch is running wait for it and then go out together.  A lone sequential caller pays no
    batching delay, concurrent finalisations share engine calls.  A request is either bytes in
    memory (``hash``) or a file the engine reads itself (``hash_path``, only with a hasher that
    offers ``hash_files``); a mixed batch costs one call per kind.  ``on_done(id)`` runs in the
    worker thread right after hashing (it stores or renames the object there)."""

    def __init__(self, hasher, window: float, max_batch: int):
        self._hasher = hasher
        self._window = window
        self._max_batch = max_batch
        self._pending: list[_Request] = []
        self._timer: asyncio.TimerHandle | None = None
        self._running = False
        self.batches = 0  # number of engine batches issued (observable in tests)

    @property
    def reads_files(self) -> bool:
        return hasattr(self._hasher, "hash_files")

    async def hash(self, data: bytes, on_done: Callable[[str], None] | None = None) -> str:
        return await self._submit(data, False, on_done)

    async def hash_path(self, path: str, on_done: Callable[[str], None] | None = None) -> str:
        return await self._submit(path, True, on_done)

    async def _submit(self, payload, is_path: bool, on_done) -> str:
        loop = asyncio.get_running_loop()
        request = _Request(payload, is_path, loop.create_future(), on_done)
        self._pending.append(request)
        if not self._running:
            if len(self._pending) >= self._max_batch:
                self._fire(loop)
            elif self._timer is None:
                # window == 0: end of this loop iteration, so same-tick arrivals still coalesce
                self._timer = loop.call_later(self._window, self._fire, loop)
        return await request.future

    def _fire(self, loop: asyncio.AbstractEventLoop) -> None:
        if self._timer is not None:
            self._timer.cancel()
            self._timer = None
        if self._running or not self._pending:
            return
        batch, self._pending = self._pending[: self._max_batch], self._pending[self._max_batch:]
        self._running = True
        loop.create_task(self._run(batch))

    def _hash_batch(self, batch: list[_Request]) -> list:
        """Worker thread: one engine call per kind of payload, then each request's ``on_done``.
        Returns, per request, the id or the exception its ``on_done`` raised."""
        ids: list = [None] * len(batch)
        for want_path, fn in ((False, getattr(self._hasher, "hash_many")), (True, getattr(self._hasher, "hash_files", None))):
            idx = [i for i, r in enumerate(batch) if r.is_path == want_path]
            if idx:
                for i, object_id in zip(idx, fn([batch[i].payload for i in idx])):
                    ids[i] = object_id
        for i, r in enumerate(batch):
            if r.on_done is not None:
                try:
                    r.on_done(ids[i])
                except Exception as e:  # noqa: BLE001 - reported to that request only
                    ids[i] = e
        return ids

    async def _run(self, batch: list[_Request]) -> None:
        loop = asyncio.get_running_loop()
        try:
            self.batches += 1
            results = await loop.run_in_executor(None, self._hash_batch, batch)
            for r, result in zip(batch, results):
                if r.future.done():
                    continue
                if isinstance(result, Exception):
                    r.future.set_exception(result)
                else:
                    r.future.set_result(result)
        except BaseException as e:  # noqa: BLE001 - an engine failure must reach every waiter
            for r in batch:
                if not r.future.done():
                    r.future.set_exception(e if isinstance(e, Exception) else RuntimeError(str(e)))
            if not isinstance(e, Exception):
                raise
        finally:
            self._running = False
            if self._pending:  # whatever queued up behind this batch goes out together, now
                self._fire(loop)


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

    def __init__(self, storage_path: str, hasher=None, device: int = 0, batch_window: float = 0.0,
                 max_batch: int = 4096, refcount: bool = False):
        self.storage_path = Path(storage_path)
        # No fallback: if the CUDA library or a B200 is missing this raises.
        self._hasher = hasher if hasher is not None else GpuHasher(device)
        self._batcher = _Batcher(self._hasher, batch_window, max_batch)
        self._refcount = refcount
        self._refs_mutex = threading.Lock()  # holder counts are read-modify-write, done in worker threads
        self._dir_ready = False

    @property
    def hasher(self):
        return self._hasher

    @property
    def gpu_batches(self) -> int:
        return self._batcher.batches

    # -- worker-thread halves (plain os calls: one thread hop covers a whole batch) -----------
    def _root(self) -> str:
        return str(self.storage_path)

    def _count_holder_sync(self, object_id: str, present_before: bool, delta: int) -> int:
        """With refcount=True: adjust the holder count of an object; returns the new count."""
        refs = os.path.join(self._root(), f".refs-{object_id}")  # '.' cannot occur in an id (validation.Hash)
        try:
            with open(refs) as f:
                holders = int(f.read().strip() or 0)
        except (FileNotFoundError, ValueError):
            holders = 1 if present_before else 0
        holders += delta
        if holders > 0:
            with open(refs, "w") as f:
                f.write(str(holders))
        elif os.path.exists(refs):
            os.unlink(refs)
        return holders

    def _publish_sync(self, tmp: str, object_id: str) -> None:
        """Give a spooled file its content-derived name, or drop it if that content is stored already."""
        target = os.path.join(self._root(), object_id)
        with self._refs_mutex:
            present = os.path.exists(target)
            if present:
                if os.path.exists(tmp):
                    os.unlink(tmp)
            else:
                os.rename(tmp, target)  # atomic within the store directory
            if self._refcount:
                self._count_holder_sync(object_id, present, +1)

    def _store_sync(self, data: bytes, object_id: str) -> None:
        """Store one in-memory object under its id (temp name, then rename)."""
        root = self._root()
        target = os.path.join(root, object_id)
        if not os.path.exists(target):
            tmp = os.path.join(root, f".tmp-{secrets.token_hex(16)}")
            try:
                try:
                    f = open(tmp, "wb")
                except FileNotFoundError:
                    os.makedirs(root, exist_ok=True)
                    f = open(tmp, "wb")
                with f:
                    f.write(data)
            except BaseException:
                if os.path.exists(tmp):
                    os.unlink(tmp)
                raise
            self._publish_sync(tmp, object_id)
        elif self._refcount:
            with self._refs_mutex:
                self._count_holder_sync(object_id, True, +1)

    def _delete_sync(self, object_id: str) -> None:
        target = os.path.join(self._root(), object_id)
        with self._refs_mutex:
            if not os.path.exists(target):
                raise FileNotFoundError(f"File not found: {object_id}")
            if self._refcount and self._count_holder_sync(object_id, True, -1) > 0:
                return  # another writer of the same content still holds it
            os.unlink(target)

    def _verify_sync(self, object_hashes: Sequence[str]) -> dict[str, bool]:
        root = self._root()
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

    # -- the reference's interface ------------------------------------------------------------
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
            publish = partial(self._publish_sync, str(tmp))  # runs in the batch's worker thread
            if from_file:
                spool.hash = await self._batcher.hash_path(str(tmp), publish)
            else:
                spool.hash = await self._batcher.hash(spool._content(), publish)
        except BaseException:
            await tmp.unlink(missing_ok=True)
            raise

    async def write(self, data: bytes) -> str:
        """
        Writes the data to the storage and returns the hash of the object.
        """
        data = bytes(data)
        return await self._batcher.hash(data, partial(self._store_sync, data))

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
        await asyncio.get_running_loop().run_in_executor(None, self._delete_sync, object_hash)

    # -- additive batch API ---------------------------------------------------------------------
    async def hash_many(self, objects: Sequence[bytes]) -> list[str]:
        """Ids of a batch of objects, one GPU batch, nothing stored."""
        if not objects:
            return []
        return await asyncio.get_running_loop().run_in_executor(None, self._hasher.hash_many, list(objects))

    async def write_many(self, objects: Sequence[bytes]) -> list[str]:
        """Stores a batch of objects and returns their ids: one GPU batch, one worker-thread
        hop for all the file writes."""
        objects = [bytes(o) for o in objects]
        ids = await self.hash_many(objects)

        def store_all() -> None:
            os.makedirs(self._root(), exist_ok=True)
            for data, object_id in zip(objects, ids):
                self._store_sync(data, object_id)

        await asyncio.get_running_loop().run_in_executor(None, store_all)
        return ids

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
            root = self._root()
            if not os.path.isdir(root):
                return {}
            names = sorted(e.name for e in os.scandir(root) if not e.name.startswith(".") and e.is_file())
            return self._verify_sync(names)

        return await asyncio.get_running_loop().run_in_executor(None, scan)
