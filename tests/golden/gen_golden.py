"""Regenerates the fixtures in tests/golden/ (run here, where /root/reference exists).

    python tests/golden/gen_golden.py

* sha256_kat.json      -- FIPS 180-4 known answers (typed in from the standard,
                          re-derived with hashlib and the two must agree), the
                          padding-boundary all-zero messages of SURVEY.md
                          appendix A, and digests of synthetic objects.
* storage_behaviour.json -- what the UNMODIFIED reference classes do at the
                          drop-in boundary, recorded by importing
                          /root/reference/src/code_interpreter/{services/storage,
                          utils/validation}.py: which ids/paths validate, the
                          id shape, the error types and messages.

* grpc_wire.json       -- the Execute messages as the reference's generated protobuf
                          descriptor declares them (numbers, types, map patterns from the
                          buf.validate options) and sample wire bytes from its classes.

/root/reference does not exist on the GPU box, hence fixtures.
"""
from __future__ import annotations

import asyncio
import hashlib
import json
import os
import re
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import sha256_oracle as oracle  # noqa: E402

# FIPS 180-4 / NIST CAVP "SHA-256 examples" known answers, typed from the standard.
FIPS = [
    ("", 1, "e3b0c44298fc1c149afbf4c8996fb92427ae41e4649b934ca495991b7852b855"),
    ("abc", 1, "ba7816bf8f01cfea414140de5dae2223b00361a396177a9cb410ff61f20015ad"),
    ("abcdbcdecdefdefgefghfghighijhijkijkljklmklmnlmnomnopnopq", 1,
     "248d6a61d20638b8e5c026930c3e6039a33ce45964ff2167f6ecedd419db06c1"),
    ("abcdefghbcdefghicdefghijdefghijkefghijklfghijklmghijklmnhijklmnoijklmnopjklmnopqklmnopqrlmnopqrsmnopqrstnopqrstu",
     1, "cf5b16a778af8380036ce59e7b0492370b249b11e8f07a51afac45037afee9d1"),
    ("a", 1_000_000, "cdc76e5c9914fb9281a1c7e284d73e67f1809a48a497200e046d39ccc7112cd0"),
]
ZERO_LENGTHS = [0, 1, 3, 4, 15, 16, 17, 31, 32, 55, 56, 57, 63, 64, 65, 119, 120, 127, 128,
                129, 183, 184, 191, 192, 255, 256, 1023, 1024, 4095, 4096, 4097, 65535, 65536, 65537]
SYNTH = [(0xB200, 0, 4096), (0xB200, 1, 4096), (0xB200, 1048575, 4096), (0xB200, 0, 65536),
         (0xB200, 12345, 65536), (0xB200, 7, 1), (0xB200, 7, 55), (0xB200, 7, 56), (0xB200, 7, 1000),
         (1, 2, 3), (0xDEADBEEF, 4000000000, 777)]


def kat() -> dict:
    fips = []
    for unit, reps, want in FIPS:
        msg = unit.encode() * reps
        assert hashlib.sha256(msg).hexdigest() == want, unit
        assert oracle.c_digest(msg).hex() == want, unit
        fips.append({"unit": unit, "repeat": reps, "sha256": want})
    zeros = [{"len": n, "nblocks": oracle.nblocks(n), "sha256": hashlib.sha256(bytes(n)).hexdigest()}
             for n in ZERO_LENGTHS]
    synth = []
    for seed, obj, n in SYNTH:
        data = oracle.synth_object(seed, obj, n)
        synth.append({"seed": seed, "object": obj, "len": n, "first16": data[:16].hex(),
                      "sha256": hashlib.sha256(data).hexdigest()})
    return {"source": "FIPS 180-4 examples + hashlib.sha256 (OpenSSL %s)" % __import__("ssl").OPENSSL_VERSION,
            "fips": fips, "zeros": zeros, "synthetic": synth}


def storage_behaviour() -> dict:
    sys.path.insert(0, "/root/reference/src")
    from pydantic import TypeAdapter, ValidationError
    from code_interpreter.services.storage import Storage
    from code_interpreter.utils.validation import AbsolutePath, Hash

    ids = ["a", "A-b_9", "0" * 64, "f" * 255, "f" * 256, "", "a/b", "a.b", "a b", "../x", "é", "abc\n",
           hashlib.sha256(b"abc").hexdigest()]
    paths = ["/workspace/file.txt", "/a", "/", "//a", "a", "", "/ a", "relative/path", "/workspace/x/../y"]
    ha, pa = TypeAdapter(Hash), TypeAdapter(AbsolutePath)

    def ok(adapter, v):
        try:
            adapter.validate_python(v)
            return True
        except ValidationError:
            return False

    out = {"hash_pattern": r"^[0-9a-zA-Z_-]{1,255}$",  # utils/validation.py:19
           "path_pattern": r"^/[^/].*$",  # utils/validation.py:20-22
           "hash_accepts": {v: ok(ha, v) for v in ids},
           "path_accepts": {v: ok(pa, v) for v in paths}}

    async def probe():
        with tempfile.TemporaryDirectory() as d:
            st = Storage(os.path.join(d, "nested", "store"))
            rec = {}
            i1 = await st.write(b"abc")
            i2 = await st.write(b"abc")
            rec["write_returns_str"] = isinstance(i1, str)
            rec["id_regex"] = "^[0-9a-f]{64}$"
            rec["id_matches_regex"] = bool(re.fullmatch(rec["id_regex"], i1))
            rec["reference_ids_are_random"] = i1 != i2
            rec["reference_id_is_sha256"] = i1 == hashlib.sha256(b"abc").hexdigest()
            rec["file_named_by_id"] = os.path.isfile(os.path.join(d, "nested", "store", i1))
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
            for name, coro in (("read", st.read), ("delete", st.delete)):
                try:
                    await coro("0" * 64)
                    rec[f"{name}_missing"] = None
                except Exception as e:  # noqa: BLE001
                    rec[f"{name}_missing"] = {"type": type(e).__name__, "msg": str(e)}
            try:
                async with st.reader("0" * 64):
                    pass
            except Exception as e:  # noqa: BLE001
                rec["reader_missing"] = {"type": type(e).__name__, "msg": str(e)}
            for name, coro in (("read", st.read), ("exists", st.exists), ("delete", st.delete)):
                try:
                    await coro("a/b")
                    rec[f"{name}_bad_id"] = None
                except Exception as e:  # noqa: BLE001
                    rec[f"{name}_bad_id"] = {"type": type(e).__name__}
            await st.delete(i1)
            rec["exists_after_delete"] = await st.exists(i1)
            rec["empty_object_roundtrip"] = (await st.read(await st.write(b""))) == b""
            return rec

    out["storage"] = asyncio.run(probe())
    return out


def grpc_wire() -> dict:
    """Facts about the Execute messages, read from the reference's generated descriptor
    (submodules/bee-proto/gen/python): field numbers and types, the two map patterns taken from
    the buf.validate options, the service path, and the bytes the reference's classes produce
    for one request and one response (so that a wire-compatible restatement can be checked
    where the reference is absent)."""
    sys.path.insert(0, "/root/reference/submodules/bee-proto/gen/python")
    import proto.code_interpreter.v1.code_interpreter_service_pb2 as pb
    from buf.validate import validate_pb2

    def fields(msg):
        return [{"name": f.name, "number": f.number, "type": f.type, "label": f.label,
                 "map_entry": bool(f.message_type and f.message_type.GetOptions().map_entry)} for f in msg.DESCRIPTOR.fields]

    def patterns(msg):
        rules = msg.DESCRIPTOR.fields_by_name["files"].GetOptions().Extensions[validate_pb2.field].map
        return {"keys": rules.keys.string.pattern, "values": rules.values.string.pattern}

    files = {"/workspace/file.txt": hashlib.sha256(b"Hello, World!").hexdigest(), "/workspace/b": "A-b_9"}
    req = pb.ExecuteRequest(source_code="print('hi')", files=files)
    resp = pb.ExecuteResponse(stdout="hi\n", stderr="", exit_code=3, files=files)
    svc = pb.DESCRIPTOR.services_by_name["CodeInterpreterService"]
    return {"package": pb.DESCRIPTOR.package, "service": svc.full_name,
            "execute_method": {"name": "Execute", "input": svc.methods_by_name["Execute"].input_type.full_name,
                               "output": svc.methods_by_name["Execute"].output_type.full_name},
            "ExecuteRequest": {"fields": fields(pb.ExecuteRequest), "files_patterns": patterns(pb.ExecuteRequest)},
            "ExecuteResponse": {"fields": fields(pb.ExecuteResponse), "files_patterns": patterns(pb.ExecuteResponse)},
            "sample": {"files": files, "request_source_code": "print('hi')",
                       "request_hex": req.SerializeToString(deterministic=True).hex(),
                       "response": {"stdout": "hi\n", "stderr": "", "exit_code": 3},
                       "response_hex": resp.SerializeToString(deterministic=True).hex()}}


if __name__ == "__main__":
    with open(os.path.join(HERE, "grpc_wire.json"), "w") as f:
        json.dump(grpc_wire(), f, indent=1)
    with open(os.path.join(HERE, "sha256_kat.json"), "w") as f:
        json.dump(kat(), f, indent=1)
    with open(os.path.join(HERE, "storage_behaviour.json"), "w") as f:
        json.dump(storage_behaviour(), f, indent=1, ensure_ascii=True)
    print("wrote fixtures")
