"""The reference's HTTP file routes and Execute file map over this repo's Storage
(SURVEY.md 8f rank 3), driven the way test/e2e/test_http.py:47-85 drives the real service.
CPU-only here (the oracle stands in for the device); tests/test_gpu_storage.py repeats the
round trip on a B200."""
import hashlib

import pytest
from starlette.testclient import TestClient

from bee_code_interpreter_fs_b200.service_harness import InProcessExecutor, create_app
from bee_code_interpreter_fs_b200.storage import Storage

WRITE_SCRIPT = b"""
with open('file.txt', 'w') as f:
    f.write("Hello, World!")
"""
READ_SCRIPT = b"""
with open('file.txt', 'r') as f:
    print(f.read())
"""
MANY_SCRIPT = b"""
for i in range(6):
    with open(f'out-{i}.bin', 'wb') as f:
        f.write(bytes([i]) * (1000 * (i + 1)))
"""


def upload(client, content: bytes) -> str:
    r = client.put("/v1/files", files={"file": ("blob", content)})
    assert r.status_code == 200
    return r.json()["hash"]


@pytest.fixture
def client(tmp_path, oracle_hasher):
    storage = Storage(str(tmp_path / "store"), hasher=oracle_hasher, batch_window=0.05)
    app = create_app(storage)
    with TestClient(app, raise_server_exceptions=False) as c:
        c.storage = storage
        yield c


def test_file_routes_status_codes_and_ids(client):
    object_id = upload(client, b"some bytes")
    assert object_id == hashlib.sha256(b"some bytes").hexdigest()
    r = client.get(f"/v1/files/{object_id}")
    assert r.status_code == 200 and r.content == b"some bytes"
    assert r.headers["content-type"] == "application/octet-stream"
    assert client.get("/v1/files/" + "0" * 64).status_code == 404
    assert client.get(f"/v1/files/{object_id}?delete=true").content == b"some bytes"
    assert client.get(f"/v1/files/{object_id}").status_code == 404
    again = upload(client, b"some bytes")
    assert again == object_id  # content-addressed: the same id comes back
    assert client.delete(f"/v1/files/{object_id}").json() == {"message": "File deleted"}
    assert client.delete(f"/v1/files/{object_id}").status_code == 500  # as in the reference: not caught


def test_create_file_in_interpreter_round_trip(client):
    """Shape of the reference's only test on this path (test/e2e/test_http.py:47-85)."""
    write_id = upload(client, WRITE_SCRIPT)
    r = client.post("/v1/execute", json={"source_file": "/workspace/main.py", "files": {"/workspace/main.py": write_id}})
    assert r.status_code == 200, r.text
    body = r.json()
    assert body["exit_code"] == 0
    assert body["files"].keys() == {"/workspace/file.txt"}
    file_id = body["files"]["/workspace/file.txt"]
    assert file_id == hashlib.sha256(b"Hello, World!").hexdigest()

    read_id = upload(client, READ_SCRIPT)
    r = client.post("/v1/execute", json={
        "source_file": "/workspace/main.py",
        "files": {"/workspace/main.py": read_id, "/workspace/file.txt": file_id},
    })
    assert r.status_code == 200, r.text
    body = r.json()
    assert body["exit_code"] == 0
    assert body["stdout"] == "Hello, World!\n"
    assert not body["files"]


def test_changed_files_of_one_execute_share_one_batch(client):
    script_id = upload(client, MANY_SCRIPT)
    before = client.storage.gpu_batches
    r = client.post("/v1/execute", json={"source_file": "/workspace/main.py", "files": {"/workspace/main.py": script_id}})
    body = r.json()
    assert r.status_code == 200 and body["exit_code"] == 0
    assert client.storage.gpu_batches == before + 1
    want = {f"/workspace/out-{i}.bin": hashlib.sha256(bytes([i]) * (1000 * (i + 1))).hexdigest() for i in range(6)}
    assert body["files"] == want


def test_bad_requests(client):
    r = client.post("/v1/execute", json={"source_file": "relative.py", "files": {}})
    assert r.status_code == 422
    r = client.post("/v1/execute", json={"source_file": "/workspace/main.py", "files": {"/workspace/main.py": "bad/id"}})
    assert r.status_code == 422
    r = client.post("/v1/execute", json={"source_file": "/workspace/main.py", "files": {"/workspace/main.py": "0" * 64}})
    assert r.status_code == 500  # unknown id: the upload leg raises FileNotFoundError
