#!/usr/bin/env python
"""Builds tools/kbench.cu once per kernel variant into tools/kbench_bin/ (git-ignored; travels with gpurun).

    python tools/kbench_build.py            # build all variants (parallel)
    bash tools/kbench_bin/run_all.sh        # on the GPU box: one JSON line per variant
"""
import itertools
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tools", "kbench_bin")


def variants():
    v = []
    for c in (2, 3, 4):  # register budget through __launch_bounds__(256, c)
        v.append((f"lb_c{c}", dict(KB_MINCTAS=c)))
    for cap in (64, 72, 80, 96, 128):  # ... and through __maxnreg__
        v.append((f"cap{cap}", dict(KB_REGCAP=cap)))
        v.append((f"cap{cap}_t128", dict(KB_REGCAP=cap, KB_THREADS=128)))
    for cap in (80, 96):
        v.append((f"cap{cap}_s1", dict(KB_REGCAP=cap, KB_SCHED=1)))
        v.append((f"cap{cap}_fma2", dict(KB_REGCAP=cap, KB_FMA=2)))
        v.append((f"cap{cap}_n3", dict(KB_REGCAP=cap, KB_NSTAGE=3)))
    for threads, cap in ((160, 80), (192, 80), (224, 72), (384, 80), (512, 64)):  # other warp counts per SM
        v.append((f"t{threads}_cap{cap}", dict(KB_REGCAP=cap, KB_THREADS=threads)))
    v.append(("m2_c2", dict(KB_MSGS=2, KB_MINCTAS=2)))  # two messages per lane
    return v


def build(item):
    name, defs = item
    exe = os.path.join(OUT, name)
    cmd = ["nvcc", "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
           f'-DKB_NAME="{name}"'] + [f"-D{k}={v}" for k, v in defs.items()] + ["-o", exe, os.path.join(ROOT, "tools", "kbench.cu")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    return name, r.returncode, r.stderr[-2000:]


def main():
    os.makedirs(OUT, exist_ok=True)
    vs = variants()
    if len(sys.argv) > 1:
        vs = [x for x in vs if any(s in x[0] for s in sys.argv[1:])]
    with ThreadPoolExecutor(8) as ex:
        results = list(ex.map(build, vs))
    ok = [n for n, rc, _ in results if rc == 0]
    for n, rc, err in results:
        if rc:
            print("FAILED", n, err)
    with open(os.path.join(OUT, "run_all.sh"), "w") as f:
        f.write("#!/bin/bash\ncd $(dirname $0)\n")
        for n in ok:
            f.write(f"timeout 120 ./{n} \"$@\" || echo '{{\"variant\": \"{n}\", \"failed\": true}}'\n")
    os.chmod(os.path.join(OUT, "run_all.sh"), 0o755)
    print(f"built {len(ok)}/{len(vs)} variants into {OUT}")


if __name__ == "__main__":
    main()
