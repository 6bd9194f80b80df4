#!/bin/bash
# Round 2, second GPU call: suite on the new store path + M-per-warp kernel, host-pipeline traces,
# storage bench, latency table, ncu source-level capture of the warp kernel's serial chain.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/r02b_pytest_gpu.log 2>&1; echo "pytest exit $?"
tail -3 gpurun_out/r02b_pytest_gpu.log
B200CAS_TRACE=1 timeout 300 python tools/configs.py --only c1,c4 --out gpurun_out/r02b_configs.json > gpurun_out/r02b_configs.log 2>&1; echo "configs exit $?"
B200CAS_TRACE=1 timeout 300 python tools/storage_bench.py > gpurun_out/r02b_storage_bench.log 2>&1; echo "storage exit $?"
cp gpurun_out/storage_bench.json gpurun_out/r02b_storage_bench.json
timeout 300 python tools/latency.py > gpurun_out/r02b_latency.log 2>&1; echo "latency exit $?"
cp gpurun_out/latency.json gpurun_out/r02b_latency.json 2>/dev/null
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 1 > gpurun_out/r02b_bench.json 2> gpurun_out/r02b_bench.err; echo "bench exit $?"
timeout 200 ncu --set full --clock-control none --import-source on -k regex:kernel_long -c 1 -f -o gpurun_out/prof_long_r02b tools/kbench_bin/long_adds4 4194309 592 592 > gpurun_out/r02b_ncu_long.log 2>&1; echo "ncu exit $?"
