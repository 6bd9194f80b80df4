#!/bin/bash
# Round 2: the two bench arms exactly as the driver launches them (defaults), on the final commit.
mkdir -p gpurun_out
/usr/bin/time -v timeout 400 python bench.py --impl reference > gpurun_out/r02g_bench_reference.json 2> gpurun_out/r02g_bench_reference.err; echo "reference exit $?"
timeout 400 python bench.py > gpurun_out/r02g_bench.json 2> gpurun_out/r02g_bench.err; echo "bench exit $?"
