#!/bin/bash
# Round 2, multi-GPU call (gpurun --gpus 8): host->device matrix, measured GPU order, the
# one-process multi-GPU product call, multi-device tests, weak + strong scaling bench lines.
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r02_topo.txt 2>&1
nproc >> gpurun_out/r02_topo.txt; free -g >> gpurun_out/r02_topo.txt
timeout 300 tools/kbench_bin/h2d_matrix > gpurun_out/r02_h2d_matrix.jsonl 2> gpurun_out/r02_h2d_matrix.err; echo "h2d exit $?"
ORDER=$(python - <<'PY'
import json
order = None
for line in open("gpurun_out/r02_h2d_matrix.jsonl"):
    d = json.loads(line)
    if d.get("section") == "greedy":
        order = d["order"]
        json.dump({"order": order, "total_gb_s_with_first_k": d["total_gb_s_with_first_k"],
                   "source": "tools/h2d_matrix.cu greedy section, gpurun --gpus 8, round 2"}, open("gpurun_out/h2d_gpu_order.json", "w"))
print(",".join(map(str, order)) if order else "")
PY
)
echo "greedy order: $ORDER"
timeout 600 python -m pytest tests/test_gpu_multi.py tests/test_gpu_storage.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r02_pytest_multi_gpu.log 2>&1; echo "pytest exit $?"
tail -3 gpurun_out/r02_pytest_multi_gpu.log
B200CAS_GPU_ORDER=$ORDER timeout 600 python tools/multi_gpu.py --out gpurun_out/r02_multi_gpu.json > gpurun_out/r02_multi_gpu.log 2>&1; echo "multi_gpu exit $?"
run_bench() {  # $1 = ranks, $2 = tag, $3 = order ("" = identity)
  B200CAS_GPU_ORDER=$3 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29541 \
    bench.py --gpus $1 --steps 5 --warmup 3 --e2e-steps 2 > gpurun_out/r02_bench_n$1_$2.json 2> gpurun_out/r02_bench_n$1_$2.err
  echo "bench n=$1 $2 exit $?"
}
run_bench 8 ordered "$ORDER"
run_bench 4 ordered "$ORDER"
run_bench 4 identity "0,1,2,3,4,5,6,7"
