#!/bin/bash
# Round 2: host-pipeline sweeps and tail-wave shapes (single GPU), kept under profiles/.
mkdir -p gpurun_out
timeout 400 python tools/host_sweep.py > gpurun_out/r02_host_sweep.jsonl 2> gpurun_out/r02_host_sweep.err; echo "host_sweep exit $?"
timeout 200 python tools/tail_wave.py > gpurun_out/r02_tail_wave.jsonl 2> gpurun_out/r02_tail_wave.err; echo "tail_wave exit $?"
