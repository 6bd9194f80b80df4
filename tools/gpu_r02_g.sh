#!/bin/bash
# Round 2: the reference arm exactly as the driver launches it (defaults), on the final commit.
mkdir -p gpurun_out
SECONDS=0
timeout 500 python bench.py --impl reference > gpurun_out/r02g_bench_reference.json 2> gpurun_out/r02g_bench_reference.err; echo "reference exit $? after ${SECONDS}s"
