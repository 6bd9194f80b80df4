#!/usr/bin/env python
"""Builds tools/kbench_long.cu once per variant of the warp-per-object kernel into tools/kbench_bin/.

    python tools/kbench_long_build.py
    bash tools/kbench_bin/run_long.sh > gpurun_out/kbench_long.jsonl     # on the GPU box
"""
import os
import subprocess
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tools", "kbench_bin")
MIB = 1 << 20

# (name, defines, argv): argv = object bytes, buffer capacity in objects, object counts
VARIANTS = []
for adds in (1, 4, 5):  # add placement on the serial path
    VARIANTS.append((f"long_adds{adds}", dict(KB_ADDS=adds), [4 * MIB + 5, 592, 1, 592]))
for act in (16, 8, 1):  # fewer active lanes in phase B: does a half-empty warp free ALU-pipe cycles?
    VARIANTS.append((f"long_act{act}", dict(KB_ADDS=1, KB_ACT=act), [4 * MIB + 5, 592, 1, 592]))
for rotf in (1, 2):  # Sigma0 rotations on the FMA pipe
    for adds in (1, 4, 5):
        VARIANTS.append((f"long_adds{adds}_rotf{rotf}", dict(KB_ADDS=adds, KB_ROTF=rotf), [4 * MIB + 5, 592, 1, 592]))
for m in (2, 4, 8, 16, 32):  # M messages per warp, n = 592 M and 1184 M objects of 1 MiB
    VARIANTS.append((f"long_m{m}", dict(KB_ADDS=1, KB_M=m), [MIB + 5, 1184 * m, 592 * m, 1184 * m]))
VARIANTS.append(("long_m1_1mib", dict(KB_ADDS=1, KB_M=1), [MIB + 5, 2368, 592, 1184, 2368]))


def build(item):
    name, defs, _ = item
    exe = os.path.join(OUT, name)
    cmd = ["nvcc", "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
           f'-DKB_NAME="{name}"'] + [f"-D{k}={v}" for k, v in defs.items()] + ["-o", exe, os.path.join(ROOT, "tools", "kbench_long.cu")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    return name, r.returncode, r.stderr[-2000:]


def main():
    os.makedirs(OUT, exist_ok=True)
    with ThreadPoolExecutor(8) as ex:
        results = list(ex.map(build, VARIANTS))
    ok = {n for n, rc, _ in results if rc == 0}
    for n, rc, err in results:
        if rc:
            print("FAILED", n, err)
    with open(os.path.join(OUT, "run_long.sh"), "w") as f:
        f.write("#!/bin/bash\ncd $(dirname $0)\n")
        for n, _, argv in VARIANTS:
            if n in ok:
                f.write(f"timeout 300 ./{n} {' '.join(map(str, argv))} || echo '{{\"variant\": \"{n}\", \"failed\": true}}'\n")
    os.chmod(os.path.join(OUT, "run_long.sh"), 0o755)
    print(f"built {len(ok)}/{len(VARIANTS)} variants into {OUT}")


if __name__ == "__main__":
    main()
