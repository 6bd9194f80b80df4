#!/bin/bash
# Round 2, first GPU call: the suite with the new full-size config tests, the warp-kernel variants,
# and this round's baseline numbers for the configs the round works on.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > gpurun_out/r02_gpu_info.txt
timeout 900 python -m pytest tests -m gpu -x -q --durations=15 > gpurun_out/r02a_pytest_gpu.log 2>&1; echo "pytest exit $?"
tail -3 gpurun_out/r02a_pytest_gpu.log
timeout 600 bash tools/kbench_bin/run_long.sh > gpurun_out/r02a_kbench_long.jsonl 2> gpurun_out/r02a_kbench_long.err; echo "kbench exit $?"
timeout 300 python tools/configs.py --only c1,c4 --out gpurun_out/r02a_configs.json > gpurun_out/r02a_configs.log 2>&1; echo "configs exit $?"
timeout 300 python tools/storage_bench.py > gpurun_out/r02a_storage_bench.log 2>&1; echo "storage exit $?"
