#!/bin/bash
# gpurun payload: launch list (durations + instruction counts) of one-rig-frame calls, then ncu --set full of kernel $1 in such a call
set -u
mkdir -p gpurun_out
K=${1:-merge_kernel}
CMD="python bench.py --steps 3 --warmup 3 --frames 1 --distinct 1 --no-cpu --no-latency"
timeout 300 $CMD > gpurun_out/plain_lat.log 2>&1 || { tail -5 gpurun_out/plain_lat.log; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none --csv --log-file gpurun_out/launches_lat.csv $CMD > gpurun_out/ncu_lat_list.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:$K -s 4 -c 1 -o gpurun_out/ncu_lat_$K -f $CMD > gpurun_out/ncu_lat_run.log 2>&1
tail -1 gpurun_out/ncu_lat_run.log
