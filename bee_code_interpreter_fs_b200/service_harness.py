"""The callers on either side of the object store, without Kubernetes.

SURVEY.md section 8(f) rank 3: the same HTTP file routes and ``files`` maps the
reference exposes (``/root/reference/src/code_interpreter/services/http_server.py:36-45,
88-160``) over this repo's ``Storage``, plus an in-process stand-in for the two file legs
of ``KubernetesCodeExecutor.execute`` (``services/kubernetes_code_executor.py:109-154``)
and for the executor pod's file endpoints (``executor/server.rs:76-137``: write the
uploaded files into a workspace, run, report files whose ctime is newer than the start).

It exists so that ``test_create_file_in_interpreter``-style round trips
(``test/e2e/test_http.py:47-85``) run against the GPU-backed store on one machine.  It is
a test harness, not a sandbox: the default runner executes the given source file with
the current Python interpreter as a plain subprocess.
"""
from __future__ import annotations

import asyncio
import os
import sys
import tempfile
import time
from dataclasses import dataclass
from typing import Awaitable, Callable, Dict, Mapping

from fastapi import FastAPI, HTTPException, Response, UploadFile
from pydantic import BaseModel, validate_call

from .storage import Storage
from .validation import AbsolutePath, Hash

CHUNK = 64 * 1024

Runner = Callable[[str, str, int], Awaitable[tuple[str, str, int]]]


async def python_subprocess_runner(workspace: str, source_file: str, timeout: int) -> tuple[str, str, int]:
    """Runs ``source_file`` (workspace-relative) with this interpreter; (stdout, stderr, exit_code)."""
    proc = await asyncio.create_subprocess_exec(
        sys.executable, os.path.join(workspace, source_file), cwd=workspace,
        stdout=asyncio.subprocess.PIPE, stderr=asyncio.subprocess.PIPE)
    try:
        out, err = await asyncio.wait_for(proc.communicate(), timeout=timeout)
    except asyncio.TimeoutError:
        proc.kill()
        await proc.wait()
        return "", "Execution timed out", -1
    return out.decode(), err.decode(), proc.returncode if proc.returncode is not None else -1


class InProcessExecutor:
    """Upload leg -> run -> download leg, against a throw-away workspace directory."""

    @dataclass
    class Result:
        stdout: str
        stderr: str
        exit_code: int
        files: Mapping[str, str]

    def __init__(self, file_storage: Storage, runner: Runner = python_subprocess_runner):
        self.file_storage = file_storage
        self.runner = runner

    @staticmethod
    def _relative(path: str) -> str:
        return path.removeprefix("/workspace/").lstrip("/")

    @validate_call
    async def execute(self, source_file: AbsolutePath, files: Mapping[AbsolutePath, Hash] = {}, timeout: int = 60):
        with tempfile.TemporaryDirectory(prefix="b200cas-workspace-") as workspace:

            async def upload(path: str, object_id: str) -> None:
                target = os.path.join(workspace, self._relative(path))
                os.makedirs(os.path.dirname(target), exist_ok=True)
                async with self.file_storage.reader(object_id) as reader:
                    with open(target, "wb") as f:
                        while chunk := await reader.read(CHUNK):
                            f.write(chunk)

            await asyncio.gather(*(upload(p, h) for p, h in files.items()))
            await asyncio.sleep(0.01)  # ctime granularity: files written above must not count as changed
            started = time.time_ns()
            stdout, stderr, exit_code = await self.runner(workspace, self._relative(source_file), timeout)

            changed = []
            for entry in os.scandir(workspace):  # top level only, as the pod's server does
                if entry.is_file() and entry.stat().st_ctime_ns > started:
                    changed.append(entry.name)

            async def download(name: str) -> tuple[str, str]:
                # Same shape as the reference's download leg: chunks stream into a writer and
                # the id is read once the writer has closed.  Concurrent writers share a GPU batch.
                async with self.file_storage.writer() as stored:
                    with open(os.path.join(workspace, name), "rb") as f:
                        while chunk := f.read(CHUNK):
                            await stored.write(chunk)
                return f"/workspace/{name}", stored.hash

            stored_files = dict(await asyncio.gather(*(download(n) for n in sorted(changed))))
            return InProcessExecutor.Result(stdout=stdout, stderr=stderr, exit_code=exit_code, files=stored_files)


class ExecuteRequest(BaseModel):
    source_file: AbsolutePath
    files: Dict[AbsolutePath, Hash] = {}
    timeout: int = 60


class ExecuteResponse(BaseModel):
    stdout: str
    stderr: str
    exit_code: int
    files: Dict[AbsolutePath, Hash]


def create_app(file_storage: Storage, code_executor: InProcessExecutor | None = None) -> FastAPI:
    """FastAPI app with the reference's file routes and Execute route over ``file_storage``."""
    app = FastAPI()
    executor = code_executor or InProcessExecutor(file_storage)

    @app.delete("/v1/files/{file_hash}")
    async def delete_file(file_hash: str):
        await file_storage.delete(file_hash)  # a missing id is not caught here either: 500
        return {"message": "File deleted"}

    @app.get("/v1/files/{file_hash}")
    async def get_file(file_hash: str, delete: bool = False):
        try:
            async with file_storage.reader(file_hash) as reader:
                content = await reader.read()
            if delete:
                await file_storage.delete(file_hash)
            return Response(content=content, media_type="application/octet-stream",
                            headers={"Content-Disposition": f"attachment; filename={file_hash}"})
        except FileNotFoundError:
            raise HTTPException(status_code=404, detail=f"File with hash {file_hash} not found")
        except Exception as e:  # noqa: BLE001 - the route answers 500 with the text
            raise HTTPException(status_code=500, detail=str(e))

    @app.put("/v1/files")
    async def write_file(file: UploadFile):
        try:
            return {"hash": await file_storage.write(await file.read())}
        except Exception as e:  # noqa: BLE001
            raise HTTPException(status_code=500, detail=str(e))

    @app.post("/v1/execute", response_model=ExecuteResponse)
    async def execute(request: ExecuteRequest):
        try:
            return await executor.execute(source_file=request.source_file, files=request.files, timeout=request.timeout)
        except Exception as e:  # noqa: BLE001
            raise HTTPException(status_code=500, detail=str(e))

    return app
