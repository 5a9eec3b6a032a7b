#!/bin/bash
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --frames 64 --distinct 2 --no-cpu"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 || { tail -5 gpurun_out/plain.log; exit 1; }
for K in fast_strips_kernel octree_kernel describe_fused_kernel; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$K -s 2 -c 1 -o gpurun_out/ncu_$K -f $CMD > gpurun_out/ncu_run_$K.log 2>&1
  tail -1 gpurun_out/ncu_run_$K.log
done
