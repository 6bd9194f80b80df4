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
* equal content gives equal ids: a second write of the same bytes stores nothing new;
* because equal content shares one file, every write counts as one *holder* of the object
  and ``delete`` (storage.py:92-101; reached over HTTP by ``DELETE /v1/files/{hash}`` and
  ``GET ...?delete=true``, http_server.py:88-109) removes the file only when the last
  holder lets go -- otherwise one client's delete would take the object away from every
  other client that uploaded the same bytes, which the reference's random ids cannot do.
  ``refcount=False`` gives the reference's literal behaviour (first delete unlinks).

The work of a batch -- hashing, writing the bytes out, giving files their final names --
happens in one C call per batch (``cas_store_host`` for in-memory objects,
``cas_publish_files`` for spooled writers: include/b200cas.h) on a worker thread, off the
event loop.  Concurrent calls (the ``asyncio.gather`` over changed files at
kubernetes_code_executor.py:149-154) are coalesced by ``_Batcher`` (group commit: no added
latency for a lone writer, one engine call for everything that queued up behind a running
one; ``batch_window`` optionally waits to collect more).  A request that fails (its spool
file vanished, the disk is full) fails alone; a request whose caller was cancelled before
its batch started is dropped from the batch.
Additive API: ``hash_many``, ``write_many``, ``verify``, ``verify_all``.
"""
from __future__ import annotations

import asyncio
import os
import re
import secrets
import threading
from contextlib import asynccontextmanager
from typing import AsyncIterator, Callable, Protocol, Sequence

from anyio import Path
from pydantic import validate_call

from .hasher import GpuHasher
from .validation import HASH_PATTERN, Hash

_ID_RE = re.compile(HASH_PATTERN)

_HASH, _STORE, _PUBLISH = "hash", "store", "publish"  # what a queued request wants done


class ObjectReader(Protocol):
    async def read(self, size: int = -1) -> bytes: ...


class ObjectWriter(Protocol):
    hash: str

    async def write(self, data: bytes) -> None: ...


class _Request:
    __slots__ = ("payload", "kind", "future")

    def __init__(self, payload, kind: str, future: asyncio.Future):
        self.payload, self.kind, self.future = payload, kind, future


class _Batcher:
    """Group commit.  A request that arrives while the engine is idle starts a batch at once
    (after ``window`` seconds if a window is set); requests that arrive while a batch is running
    wait for it and then go out together.  A lone sequential caller pays no batching delay,
    concurrent finalisations share engine calls.  ``process(batch)`` runs in a worker thread and
    returns one result per request: the object id, or the exception for that request alone."""

    def __init__(self, process: Callable[[list[_Request]], list], window: float, max_batch: int):
        self._process = process
        self._window = window
        self._max_batch = max_batch
        self._pending: list[_Request] = []
        self._timer: asyncio.TimerHandle | None = None
        self._running = False
        self.batches = 0  # number of engine batches issued (observable in tests)

    async def submit(self, payload, kind: str) -> str:
        loop = asyncio.get_running_loop()
        request = _Request(payload, kind, loop.create_future())
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
        if self._running:
            return
        # A caller that was cancelled while it waited has nothing to store any more (a cancelled
        # writer removes its own spool file): it must not reach the engine, let alone fail others.
        self._pending = [r for r in self._pending if not r.future.done()]
        if not self._pending:
            return
        batch, self._pending = self._pending[: self._max_batch], self._pending[self._max_batch:]
        self._running = True
        loop.create_task(self._run(batch))

    async def _run(self, batch: list[_Request]) -> None:
        loop = asyncio.get_running_loop()
        try:
            self.batches += 1
            results = await loop.run_in_executor(None, self._process, batch)
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
    """Flat store of immutable binary objects, each named by the SHA-256 of its content.

    One file per object under ``storage_path`` (the layout of the reference's store,
    storage.py:53), so an unmodified reader of that directory keeps working.
    """

    def __init__(self, storage_path: str, hasher=None, device: int = 0, batch_window: float = 0.0,
                 max_batch: int = 4096, refcount: bool = True, reserve_bytes: int = 0):
        self.storage_path = Path(storage_path)
        # No fallback: if the CUDA library or a B200 is missing this raises.
        self._hasher = hasher if hasher is not None else GpuHasher(device)
        if reserve_bytes and hasattr(self._hasher, "reserve"):
            self._hasher.reserve(reserve_bytes, max_batch)  # the first request then allocates nothing
        self._batcher = _Batcher(self._process_batch, batch_window, max_batch)
        self._refcount = refcount
        # Guards every step that gives an object its final name, counts a holder or removes an
        # object (all done in worker threads), so that a delete cannot slip between "this id is
        # stored already" and the new holder being counted.
        self._refs_mutex = threading.Lock()
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

    def _refs_path(self, object_id: str) -> str:
        return os.path.join(self._root(), f".refs-{object_id}")  # '.' cannot occur in an id (validation.Hash)

    def _holders_sync(self, object_id: str) -> int:
        """Holder count of a stored object.  An object without a count file has one holder, so
        the common case -- content written once -- costs no extra file."""
        try:
            with open(self._refs_path(object_id)) as f:
                return max(1, int(f.read().strip() or 1))
        except (FileNotFoundError, ValueError):
            return 1

    def _set_holders_sync(self, object_id: str, holders: int) -> None:
        refs = self._refs_path(object_id)
        if holders > 1:
            with open(refs, "w") as f:
                f.write(str(holders))
        elif os.path.exists(refs):
            os.unlink(refs)

    def _note_existing_sync(self, object_id: str) -> None:
        """A write found its content stored already: that is one more holder (mutex held)."""
        if self._refcount:
            self._set_holders_sync(object_id, self._holders_sync(object_id) + 1)

    def _publish_sync(self, tmp: str, object_id: str) -> None:
        """Python form of the publish step (mutex held): give a spooled file its content-derived
        name, or drop it if that content is stored already."""
        target = os.path.join(self._root(), object_id)
        if os.path.exists(target):
            if os.path.exists(tmp):
                os.unlink(tmp)
            self._note_existing_sync(object_id)
        else:
            os.rename(tmp, target)  # atomic within the store directory

    def _store_sync(self, data: bytes, object_id: str) -> None:
        """Python form of the store step (mutex held): temp name, then rename."""
        root = self._root()
        if os.path.exists(os.path.join(root, object_id)):
            self._note_existing_sync(object_id)
            return
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

    def _finish_sync(self, ids: Sequence[str], status: Sequence[int], what: Sequence) -> list:
        """Per-object outcome of a C batch call (mutex held): count holders of content that was
        already there, turn a per-object failure into that request's exception."""
        out: list = []
        for object_id, st, item in zip(ids, status, what):
            if st < 0:
                out.append(OSError(f"object could not be stored (engine status {int(st)}): {item!r:.80}"))
                continue
            if st == 1:
                self._note_existing_sync(object_id)
            out.append(object_id)
        return out

    def _store_batch_sync(self, datas: list[bytes]) -> list:
        """Ids of the objects, all of them stored; per object the id or its exception."""
        with self._refs_mutex:
            if hasattr(self._hasher, "store_many"):  # one C call: hash, write, rename
                ids, status = self._hasher.store_many(self._root(), datas)
                return self._finish_sync(ids, status, ["<bytes>"] * len(ids))
            # A hasher without the batched store (the CPU tests' stand-in): the same steps here.
            out: list = []
            for data, object_id in zip(datas, self._hasher.hash_many(datas)):
                try:
                    self._store_sync(data, object_id)
                    out.append(object_id)
                except Exception as e:  # noqa: BLE001 - reported to that request only
                    out.append(e)
            return out

    def _publish_batch_sync(self, payloads: list) -> list:
        """Spooled writers: payload = (spool path, content or None)."""
        paths = [p for p, _ in payloads]
        with self._refs_mutex:
            if hasattr(self._hasher, "publish_files"):  # one C call: pread into pinned staging, hash, rename
                ids, _, status = self._hasher.publish_files(self._root(), paths)
                return self._finish_sync(ids, status, paths)
            if hasattr(self._hasher, "hash_files"):
                present = [p for p in paths if os.path.isfile(p)]
                got = dict(zip(present, self._hasher.hash_files(present))) if present else {}
            else:
                got = dict(zip(paths, self._hasher.hash_many([c for _, c in payloads])))
            out: list = []
            for p in paths:
                try:
                    if p not in got:
                        raise FileNotFoundError(f"spooled object vanished before it was stored: {p}")
                    self._publish_sync(p, got[p])
                    out.append(got[p])
                except Exception as e:  # noqa: BLE001 - reported to that request only
                    out.append(e)
            return out

    def _process_batch(self, batch: list[_Request]) -> list:
        """Worker thread: one engine call per kind of request in the batch."""
        results: list = [None] * len(batch)
        for kind, fn in ((_HASH, self._hasher.hash_many), (_STORE, self._store_batch_sync),
                         (_PUBLISH, self._publish_batch_sync)):
            idx = [i for i, r in enumerate(batch) if r.kind == kind]
            if idx:
                for i, res in zip(idx, fn([batch[i].payload for i in idx])):
                    results[i] = res
        return results

    def _delete_sync(self, object_id: str) -> None:
        target = os.path.join(self._root(), object_id)
        with self._refs_mutex:
            if not os.path.exists(target):
                raise FileNotFoundError(f"File not found: {object_id}")
            if self._refcount:
                holders = self._holders_sync(object_id)
                if holders > 1:  # another writer of the same content still holds it
                    self._set_holders_sync(object_id, holders - 1)
                    return
                self._set_holders_sync(object_id, 0)
            os.unlink(target)

    def _verify_sync(self, object_hashes: Sequence[str]) -> dict[str, bool]:
        root = self._root()
        if hasattr(self._hasher, "verify_names"):  # read, hashed and compared inside the engine call
            ok = self._hasher.verify_names(root, object_hashes)
            return dict(zip(object_hashes, (ok == 1).tolist()))
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

    @property
    def _reads_files(self) -> bool:
        return hasattr(self._hasher, "publish_files") or hasattr(self._hasher, "hash_files")

    # -- the reference's interface ------------------------------------------------------------
    @asynccontextmanager
    async def writer(self) -> AsyncIterator[ObjectWriter]:
        """Streams a new object into the store (same shape as the reference's ``writer``,
        storage.py:44-55): ``await w.write(chunk)`` inside the block, ``w.hash`` after it.

        The id depends on every byte, so it exists once the block has been left -- which is
        where the executor's download leg reads it (kubernetes_code_executor.py:146).
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
        from_file = self._reads_files  # the engine preads the spooled file: O(1) memory per writer
        spool = _SpoolWriter(file, keep_chunks=not from_file)
        try:
            try:
                yield spool
            finally:
                await file.aclose()
            spool.hash = await self._batcher.submit((str(tmp), None if from_file else spool._content()), _PUBLISH)
        except BaseException:
            await tmp.unlink(missing_ok=True)
            raise

    async def write(self, data: bytes) -> str:
        """Stores ``data`` and returns its id (storage.py:57-63).  Calls made at the same time
        share one engine batch."""
        return await self._batcher.submit(bytes(data), _STORE)

    @asynccontextmanager
    @validate_call
    async def reader(self, object_hash: Hash) -> AsyncIterator[ObjectReader]:
        """Opens a stored object for chunked reads (storage.py:65-75); an unknown or empty id
        raises ``FileNotFoundError("File not found: <id>")`` as in the reference."""
        target_file = self.storage_path / object_hash
        if not object_hash or not await target_file.exists():
            raise FileNotFoundError(f"File not found: {object_hash}")
        async with await target_file.open("rb") as f:
            yield f

    @validate_call
    async def read(self, object_hash: Hash) -> bytes:
        """The whole content of one object (storage.py:77-83)."""
        async with self.reader(object_hash) as f:
            return await f.read()

    @validate_call
    async def exists(self, object_hash: Hash) -> bool:
        """Whether an object of this id is stored (storage.py:85-90)."""
        return await (self.storage_path / object_hash).exists()

    @validate_call
    async def delete(self, object_hash: Hash) -> None:
        """Lets go of one holder of the object; the file goes with the last holder
        (storage.py:92-101 removes it at once: there ids are unique per write)."""
        await asyncio.get_running_loop().run_in_executor(None, self._delete_sync, object_hash)

    # -- additive batch API ---------------------------------------------------------------------
    async def hash_many(self, objects: Sequence[bytes]) -> list[str]:
        """Ids of a batch of objects, one GPU batch, nothing stored."""
        if not objects:
            return []
        return await asyncio.get_running_loop().run_in_executor(None, self._hasher.hash_many, list(objects))

    async def write_many(self, objects: Sequence[bytes]) -> list[str]:
        """Stores a batch of objects and returns their ids: one GPU batch and one worker-thread
        hop; hashing, the file writes and the renames overlap inside the engine call."""
        objects = [bytes(o) for o in objects]
        if not objects:
            return []
        results = await asyncio.get_running_loop().run_in_executor(None, self._store_batch_sync, objects)
        for r in results:
            if isinstance(r, Exception):
                raise r
        return results

    async def verify(self, object_hashes: Sequence[str]) -> dict[str, bool]:
        """Re-hashes stored objects in one GPU batch: id -> (content still matches id).
        With a GPU hasher the engine reads the files itself (pread into pinned staging), so
        the bytes never pass through Python; the whole check is one worker-thread hop."""
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
            if hasattr(self._hasher, "verify_names"):  # the engine tells files from other entries itself
                names = [n for n in os.listdir(root) if not n.startswith(".")]
                ok = self._hasher.verify_names(root, names).tolist()
                if 2 in ok:  # sub-directories and the like are not objects
                    return {n: v == 1 for n, v in zip(names, ok) if v != 2}
                return dict(zip(names, map(bool, ok)))
            names = sorted(e.name for e in os.scandir(root) if not e.name.startswith(".") and e.is_file())
            return self._verify_sync(names)

        return await asyncio.get_running_loop().run_in_executor(None, scan)
