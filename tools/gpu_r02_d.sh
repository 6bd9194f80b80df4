#!/bin/bash
# Round 2 evidence run on one B200: suite, all five configs, storage path, latency table, soak,
# the bench line with its reference arm, and the ncu launch list + full capture of the bench.
mkdir -p gpurun_out
bash tools/gds_probe.sh > gpurun_out/r02d_gds_probe.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -x -q --durations=5 > gpurun_out/r02d_pytest_gpu.log 2>&1; echo "pytest exit $?"
tail -3 gpurun_out/r02d_pytest_gpu.log
timeout 600 python tools/configs.py --out gpurun_out/r02d_configs.json > gpurun_out/r02d_configs.log 2>&1; echo "configs exit $?"
timeout 300 python tools/storage_bench.py > gpurun_out/r02d_storage_bench.log 2>&1; echo "storage exit $?"
cp gpurun_out/storage_bench.json gpurun_out/r02d_storage_bench.json
timeout 300 python tools/latency.py > gpurun_out/r02d_latency.log 2>&1; echo "latency exit $?"
cp gpurun_out/latency.json gpurun_out/r02d_latency.json
timeout 200 python tools/soak.py --seconds 45 --seed 11 > gpurun_out/r02d_soak11.log 2>&1; echo "soak11 exit $?"
timeout 200 python tools/soak.py --seconds 45 --seed 12 > gpurun_out/r02d_soak12.log 2>&1; echo "soak12 exit $?"
timeout 400 python bench.py > gpurun_out/r02d_bench.json 2> gpurun_out/r02d_bench.err; echo "bench exit $?"
timeout 300 python bench.py --impl reference --ref-seconds 30 > gpurun_out/r02d_bench_reference.json 2> gpurun_out/r02d_bench_reference.err; echo "reference exit $?"
bash tools/gpu_ncu.sh r02 > gpurun_out/r02d_ncu_script.log 2>&1; echo "ncu exit $?"
