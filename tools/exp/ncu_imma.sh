#!/bin/bash
# gpurun payload: ncu --set full of the integer-MMA matcher experiment (MCS_MATCH_IMMA=1) in the 64-frame sequence
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --frames 64 --distinct 2 --no-cpu --no-latency"
MCS_MATCH_IMMA=1 timeout 400 ncu --set full --clock-control none --import-source on -k regex:hamming_match_imma -s 2 -c 1 -o gpurun_out/ncu_imma -f $CMD > gpurun_out/ncu_run_imma.log 2>&1
tail -1 gpurun_out/ncu_run_imma.log
