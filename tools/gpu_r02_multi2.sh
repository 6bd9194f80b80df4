#!/bin/bash
# Round 2, second multi-GPU call (gpurun --gpus 8): does the host->device gate raise the
# eight-GPU end-to-end number, and the one-process product call after the pinned-source fix.
mkdir -p gpurun_out
run_bench() {  # $1 = ranks, $2 = tag, $3 = gate on/off
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29542 \
    bench.py --gpus $1 --steps 3 --warmup 3 --e2e-steps 2 --verify-objects 256 --h2d-gate $3 > gpurun_out/r02_bench_n$1_$2.json 2> gpurun_out/r02_bench_n$1_$2.err
  echo "bench n=$1 $2 exit $?"
}
run_bench 8 gate_on on
run_bench 8 gate_off off
timeout 240 python tools/multi_gpu.py --out gpurun_out/r02_multi_gpu_b.json > gpurun_out/r02_multi_gpu_b.log 2>&1; echo "multi_gpu exit $?"
