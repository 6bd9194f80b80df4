#!/bin/bash
# Round 2, last check on the final commit: the GPU suite and smoke().
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02f_pytest_gpu.log 2>&1; echo "pytest exit $?"
tail -2 gpurun_out/r02f_pytest_gpu.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02f_smoke.log 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/r02f_smoke.log
timeout 120 python tools/tail_wave.py > gpurun_out/r02f_tail_wave.jsonl 2> gpurun_out/r02f_tail_wave.err; echo "tail_wave exit $?"
