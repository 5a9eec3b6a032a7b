#!/bin/bash
set -u
mkdir -p gpurun_out
K=${1:-fast_strips_kernel}
timeout 300 python bench.py --steps 2 --warmup 1 --frames 64 --distinct 2 --no-cpu > gpurun_out/plain.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K -s 1 -c 1 -o gpurun_out/ncu_$K -f python bench.py --steps 2 --warmup 1 --frames 64 --distinct 2 --no-cpu > gpurun_out/ncu_run.log 2>&1
tail -3 gpurun_out/ncu_run.log
ls -la gpurun_out/*.ncu-rep
