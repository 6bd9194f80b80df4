#!/bin/bash
# ncu evidence for the bench command: launch list (shares of the step) and one full capture
# of the dominant kernel.  Run under gpurun, one GPU.  $1 = tag for output names.
TAG=${1:-r01}
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 1"
$CMD > gpurun_out/ncu_plain_${TAG}.json 2> gpurun_out/ncu_plain_${TAG}.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_${TAG}.csv $CMD > gpurun_out/ncu_launches_${TAG}.log 2>&1
echo "launch list exit $?"
SMALL="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --e2e-steps 1 --objects 151552 --verify-objects 64"
$SMALL > gpurun_out/ncu_small_plain_${TAG}.json 2> gpurun_out/ncu_small_plain_${TAG}.err &&
ncu --set full --clock-control none --import-source on -k regex:kernel_tma -s 3 -c 2 -f -o gpurun_out/prof_${TAG} $SMALL > gpurun_out/ncu_full_${TAG}.log 2>&1
echo "full capture exit $?"
ls -la gpurun_out/
