#!/usr/bin/env python
"""File-I/O cost of the drop-in Storage next to the UNMODIFIED reference Storage, in the
CPU-only build container (the reference is not available on the GPU box).  The reference
assigns random ids and hashes nothing; ours is given a hashlib-backed stand-in for the device
(tests/conftest.py::OracleHasher) so that only the host-side path is compared.

    python tools/storage_vs_reference.py   ->  profiles/r01_storage_vs_reference_cpu_container.json
"""
import asyncio
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, "/root/reference/src")

from code_interpreter.services.storage import Storage as RefStorage  # noqa: E402

from bee_code_interpreter_fs_b200.storage import Storage  # noqa: E402
from conftest import OracleHasher  # noqa: E402

N, L = 2000, 65536


async def timed(label, coro_fn, res):
    t0 = time.perf_counter()
    await coro_fn()
    dt = time.perf_counter() - t0
    res[label + "_objects_s"] = N / dt
    res[label + "_gb_s"] = N * L / dt / 1e9


async def main():
    objs = [os.urandom(L) for _ in range(N)]
    res = {"objects": N, "object_bytes": L, "where": "build container, 8 vCPU, no GPU, store on tmpfs; ours hashes with hashlib via the oracle stand-in"}
    with tempfile.TemporaryDirectory(dir="/dev/shm" if os.path.isdir("/dev/shm") else None) as d:
        ref = RefStorage(os.path.join(d, "ref"))
        ours = Storage(os.path.join(d, "ours"), hasher=OracleHasher())

        async def seq(st):
            for o in objs:
                await st.write(o)

        async def gathered(st):
            await asyncio.gather(*(st.write(o) for o in objs))

        await timed("reference_write_sequential", lambda: seq(ref), res)
        await timed("ours_write_sequential", lambda: seq(ours), res)
        await timed("reference_write_gathered", lambda: gathered(RefStorage(os.path.join(d, "ref2"))), res)
        await timed("ours_write_gathered", lambda: gathered(Storage(os.path.join(d, "ours2"), hasher=OracleHasher())), res)
        await timed("ours_write_many", lambda: Storage(os.path.join(d, "ours3"), hasher=OracleHasher()).write_many(objs), res)
    out = os.path.join(ROOT, "profiles", "r01_storage_vs_reference_cpu_container.json")
    with open(out, "w") as f:
        json.dump(res, f, indent=1)
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    asyncio.run(main())
