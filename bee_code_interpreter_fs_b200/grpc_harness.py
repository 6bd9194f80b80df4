"""The gRPC side of the callers: ``CodeInterpreterService/Execute`` over this repo's store.

SURVEY.md section 8 rows a13 / f3.  The reference's servicer
(``/root/reference/src/code_interpreter/services/grpc_servicers/code_interpreter_servicer.py:55-81``)
validates the request, hands ``files`` (path -> object id) to the executor and returns the
executor's ``files`` map; the wire form is two ``map<string, string>`` fields whose keys and
values carry regex rules
(``/root/reference/submodules/bee-proto/proto/code_interpreter/v1/code_interpreter_service.proto:27-44``).

Nothing of the reference is imported here: the two messages are declared at run time from a
descriptor written out below (same package, names, field numbers and types, so a client built
from the reference's generated stubs talks to this server unchanged -- tests/test_grpc_harness.py
checks that against the reference's ``pb2`` where it is present), and the handler is registered
as a generic unary-unary method, so no generated ``*_pb2_grpc`` module is needed either.

Validation: the reference calls ``protovalidate.validate`` (not installed in this image); the
only rules on these messages are the two map patterns, which are the same two patterns
``validation.py`` holds for the Python layer (``utils/validation.py:19-22``).  They are applied
by hand and a violation aborts with INVALID_ARGUMENT, as ``_validate_request`` does (:45-53).

One difference of this fork is kept visible rather than papered over: the servicer passes
``source_code=...`` (:67-70) while this fork's executor takes ``source_file`` (an absolute
path inside the workspace, ``kubernetes_code_executor.py:82-88``), so the reference's gRPC
Execute cannot run as written.  Here ``source_code`` is taken as a workspace path when it is a
key of ``files``; otherwise it is inline code, stored as one more object (through the same
batched hasher) and run as ``/workspace/.main.py``.
"""
from __future__ import annotations

import re

import grpc
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

from .service_harness import InProcessExecutor
from .storage import Storage
from .validation import ABSOLUTE_PATH_PATTERN, HASH_PATTERN

SERVICE = "code_interpreter.v1.CodeInterpreterService"
INLINE_MAIN = "/workspace/.main.py"

_KEY_RE = re.compile(ABSOLUTE_PATH_PATTERN)
_VALUE_RE = re.compile(HASH_PATTERN)


def _file_descriptor() -> descriptor_pb2.FileDescriptorProto:
    """ExecuteRequest / ExecuteResponse as the proto declares them (:27-44)."""
    F = descriptor_pb2.FieldDescriptorProto
    fd = descriptor_pb2.FileDescriptorProto(name="b200cas/code_interpreter_execute.proto", package="code_interpreter.v1",
                                            syntax="proto3")

    def message(name: str, scalars: list[tuple[str, int, int]], map_number: int):
        m = fd.message_type.add(name=name)
        for fname, number, ftype in scalars:
            m.field.add(name=fname, number=number, type=ftype, label=F.LABEL_OPTIONAL, json_name=fname)
        entry = m.nested_type.add(name="FilesEntry")
        entry.options.map_entry = True
        entry.field.add(name="key", number=1, type=F.TYPE_STRING, label=F.LABEL_OPTIONAL, json_name="key")
        entry.field.add(name="value", number=2, type=F.TYPE_STRING, label=F.LABEL_OPTIONAL, json_name="value")
        m.field.add(name="files", number=map_number, type=F.TYPE_MESSAGE, label=F.LABEL_REPEATED, json_name="files",
                    type_name=f".code_interpreter.v1.{name}.FilesEntry")
        return m

    req = message("ExecuteRequest", [("source_code", 2, F.TYPE_STRING)], 3)
    req.reserved_range.add(start=1, end=2)  # executor_id, no longer used
    message("ExecuteResponse", [("stdout", 1, F.TYPE_STRING), ("stderr", 2, F.TYPE_STRING), ("exit_code", 3, F.TYPE_INT32)], 4)
    return fd


_pool = descriptor_pool.DescriptorPool()  # a private pool: the reference's pb2 may be loaded beside it in tests
_file = _pool.Add(_file_descriptor())
ExecuteRequest = message_factory.GetMessageClass(_pool.FindMessageTypeByName("code_interpreter.v1.ExecuteRequest"))
ExecuteResponse = message_factory.GetMessageClass(_pool.FindMessageTypeByName("code_interpreter.v1.ExecuteResponse"))


def files_map_violations(files) -> list[str]:
    """The proto's map rules, applied by hand: keys ``^/[^/].*$``, values ``^[0-9a-zA-Z_-]{1,255}$``."""
    bad = []
    for key, value in files.items():
        if not _KEY_RE.fullmatch(key):
            bad.append(f"files key {key!r} does not match {ABSOLUTE_PATH_PATTERN}")
        if not _VALUE_RE.fullmatch(value):
            bad.append(f"files[{key!r}] value {value!r} does not match {HASH_PATTERN}")
    return bad


class CodeInterpreterServicer:
    """``Execute`` only (the custom-tool calls do not touch the object store: out of scope)."""

    def __init__(self, file_storage: Storage, code_executor: InProcessExecutor | None = None):
        self.file_storage = file_storage
        self.code_executor = code_executor or InProcessExecutor(file_storage)

    async def Execute(self, request, context: grpc.aio.ServicerContext):
        bad = files_map_violations(request.files)
        if bad:
            await context.abort(grpc.StatusCode.INVALID_ARGUMENT, str(bad))
        files = dict(request.files)
        if request.source_code in files:
            source_file = request.source_code
        else:
            files[INLINE_MAIN] = await self.file_storage.write(request.source_code.encode())
            source_file = INLINE_MAIN
        result = await self.code_executor.execute(source_file=source_file, files=files)
        out = {k: v for k, v in result.files.items() if k != INLINE_MAIN}
        bad = files_map_violations(out)
        if bad:  # the response map carries the same rules (:40-43)
            await context.abort(grpc.StatusCode.INTERNAL, str(bad))
        return ExecuteResponse(stdout=result.stdout, stderr=result.stderr, exit_code=result.exit_code, files=out)


def add_servicer_to_server(servicer: CodeInterpreterServicer, server: grpc.aio.Server) -> None:
    handlers = {"Execute": grpc.unary_unary_rpc_method_handler(
        servicer.Execute, request_deserializer=ExecuteRequest.FromString,
        response_serializer=ExecuteResponse.SerializeToString)}
    server.add_generic_rpc_handlers((grpc.method_handlers_generic_handler(SERVICE, handlers),))


def execute_stub(channel: grpc.aio.Channel):
    """Client side of Execute for code that has no generated stub."""
    return channel.unary_unary(f"/{SERVICE}/Execute", request_serializer=ExecuteRequest.SerializeToString,
                               response_deserializer=ExecuteResponse.FromString)
