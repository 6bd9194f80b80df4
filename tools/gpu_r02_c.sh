#!/bin/bash
# Round 2, third single-GPU call: full suite, host-pipeline sweeps, store/scrub, tail-wave shapes.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --durations=5 > gpurun_out/r02c_pytest_gpu.log 2>&1; echo "pytest exit $?"
tail -3 gpurun_out/r02c_pytest_gpu.log
timeout 600 python tools/host_sweep.py > gpurun_out/r02c_host_sweep.jsonl 2> gpurun_out/r02c_host_sweep.err; echo "host_sweep exit $?"
timeout 300 python tools/tail_wave.py > gpurun_out/r02c_tail_wave.jsonl 2> gpurun_out/r02c_tail_wave.err; echo "tail_wave exit $?"
B200CAS_TRACE=1 timeout 300 python tools/configs.py --only c1 --out gpurun_out/r02c_configs.json > gpurun_out/r02c_configs.log 2>&1; echo "configs exit $?"
B200CAS_TRACE=1 timeout 300 python tools/storage_bench.py > gpurun_out/r02c_storage_bench.log 2>&1; echo "storage exit $?"
cp gpurun_out/storage_bench.json gpurun_out/r02c_storage_bench.json
