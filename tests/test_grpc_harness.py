"""The gRPC Execute leg over the content-addressed store.

The message layout and the two map patterns are checked against facts recorded from the
reference's generated descriptor (tests/golden/grpc_wire.json, made by gen_golden.py) and,
where /root/reference is present, against the reference's own pb2 classes and generated client
stub.  The round trip is the reference's test_create_file_in_interpreter
(test/e2e/test_grpc.py:78-106) without Kubernetes."""
import asyncio
import hashlib
import json
import os
import sys

import grpc
import pytest

from bee_code_interpreter_fs_b200 import Storage, grpc_harness as gh, validation
from conftest import GOLDEN

REF_GEN = "/root/reference/submodules/bee-proto/gen/python"


@pytest.fixture(scope="module")
def wire():
    with open(os.path.join(GOLDEN, "grpc_wire.json")) as f:
        return json.load(f)


def test_messages_match_the_reference_descriptor(wire):
    assert gh.SERVICE == wire["service"] and wire["execute_method"]["name"] == "Execute"
    for cls, key in ((gh.ExecuteRequest, "ExecuteRequest"), (gh.ExecuteResponse, "ExecuteResponse")):
        assert cls.DESCRIPTOR.full_name == wire["execute_method"]["input" if key == "ExecuteRequest" else "output"]
        got = [{"name": f.name, "number": f.number, "type": f.type, "label": 3 if f.is_repeated else 1,
                "map_entry": bool(f.message_type and f.message_type.GetOptions().map_entry)} for f in cls.DESCRIPTOR.fields]
        assert got == wire[key]["fields"]
        # the rules the servicer applies by hand are the ones in the proto's buf.validate options
        assert wire[key]["files_patterns"] == {"keys": validation.ABSOLUTE_PATH_PATTERN, "values": validation.HASH_PATTERN}
    s = wire["sample"]
    req = gh.ExecuteRequest.FromString(bytes.fromhex(s["request_hex"]))
    assert req.source_code == s["request_source_code"] and dict(req.files) == s["files"]
    assert gh.ExecuteRequest(source_code=s["request_source_code"], files=s["files"]).SerializeToString(
        deterministic=True).hex() == s["request_hex"]
    resp = gh.ExecuteResponse.FromString(bytes.fromhex(s["response_hex"]))
    assert (resp.stdout, resp.stderr, resp.exit_code, dict(resp.files)) == ("hi\n", "", 3, s["files"])
    assert gh.ExecuteResponse(files=s["files"], **s["response"]).SerializeToString(deterministic=True).hex() == s["response_hex"]


def test_map_rules():
    good = {"/workspace/a.txt": "0" * 64, "/x": "A-b_9"}
    assert gh.files_map_violations(good) == []
    bad = gh.files_map_violations({"relative": "0" * 64, "/ok": "has space", "//double": "x", "/p": ""})
    assert len(bad) == 4


async def _serve(storage):
    server = grpc.aio.server()
    gh.add_servicer_to_server(gh.CodeInterpreterServicer(storage), server)
    port = server.add_insecure_port("127.0.0.1:0")
    await server.start()
    return server, port


def _round_trip(storage, make_stub):
    """Create a file in one Execute, read it back in the next through its id."""

    async def go():
        server, port = await _serve(storage)
        try:
            async with grpc.aio.insecure_channel(f"127.0.0.1:{port}") as channel:
                execute, Request = make_stub(channel)
                content = "Hello, World!"
                r1 = await execute(Request(source_code=f"with open('file.txt', 'w') as f:\n    f.write(\"{content}\")\n"))
                assert r1.exit_code == 0, r1.stderr
                assert set(r1.files.keys()) == {"/workspace/file.txt"}
                object_id = r1.files["/workspace/file.txt"]
                assert object_id == hashlib.sha256(content.encode()).hexdigest()  # content-derived id
                r2 = await execute(Request(source_code="with open('file.txt', 'r') as f:\n    print(f.read())\n",
                                           files={"/workspace/file.txt": object_id}))
                assert r2.exit_code == 0 and r2.stdout == content + "\n" and not r2.files
                # a script that is itself one of the mapped files (this fork's source_file form)
                script = await storage.write(b"print(open('file.txt').read().upper())\n")
                r3 = await execute(Request(source_code="/workspace/main.py",
                                           files={"/workspace/main.py": script, "/workspace/file.txt": object_id}))
                assert r3.exit_code == 0 and r3.stdout == content.upper() + "\n"
                with pytest.raises(grpc.aio.AioRpcError) as err:
                    await execute(Request(source_code="print(1)", files={"not-absolute": object_id}))
                assert err.value.code() == grpc.StatusCode.INVALID_ARGUMENT
                with pytest.raises(grpc.aio.AioRpcError) as err:
                    await execute(Request(source_code="print(1)", files={"/workspace/x": "bad id!"}))
                assert err.value.code() == grpc.StatusCode.INVALID_ARGUMENT
        finally:
            await server.stop(None)

    asyncio.run(go())


def test_create_file_round_trip_over_grpc(tmp_path, oracle_hasher):
    storage = Storage(str(tmp_path / "store"), hasher=oracle_hasher, batch_window=0.02)
    _round_trip(storage, lambda ch: (gh.execute_stub(ch), gh.ExecuteRequest))


@pytest.mark.skipif(not os.path.isdir(REF_GEN), reason="the reference's generated stubs are not on this machine")
def test_a_client_built_from_the_reference_stubs_talks_to_this_server(tmp_path, oracle_hasher):
    sys.path.insert(0, REF_GEN)
    try:
        import proto.code_interpreter.v1.code_interpreter_service_pb2 as pb
        import proto.code_interpreter.v1.code_interpreter_service_pb2_grpc as pb_grpc
    finally:
        sys.path.remove(REF_GEN)
    storage = Storage(str(tmp_path / "store"), hasher=oracle_hasher, batch_window=0.02)
    _round_trip(storage, lambda ch: (pb_grpc.CodeInterpreterServiceStub(ch).Execute, pb.ExecuteRequest))


@pytest.mark.gpu
def test_create_file_round_trip_over_grpc_on_the_gpu_store(tmp_path, hasher):
    storage = Storage(str(tmp_path / "store"), hasher=hasher, batch_window=0.02)
    before = hasher.launch_count
    _round_trip(storage, lambda ch: (gh.execute_stub(ch), gh.ExecuteRequest))
    assert hasher.launch_count > before
