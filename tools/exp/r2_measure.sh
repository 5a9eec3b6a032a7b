#!/bin/bash
# gpurun payload (round 2): GPU suite, the bench lines of every configuration, the reference arm, the ncu launch list with instruction
# counts, --set full captures of the main kernels
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 400 python bench.py > gpurun_out/bench.json 2>gpurun_out/bench.err; echo "bench rc=$?"; tail -c 600 gpurun_out/bench.json
timeout 200 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference_arm.json 2>gpurun_out/bench_ref.err; echo "ref rc=$?"
for c in 3 4 noise; do
  timeout 400 python bench.py --config $c --steps 10 --cpu-seconds 8 > gpurun_out/bench_config$c.json 2>gpurun_out/bench_config$c.err; echo "config $c rc=$?"
done
CMD="python bench.py --steps 2 --warmup 1 --frames 64 --distinct 2 --no-cpu --no-latency"
timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_l.log 2>&1
for K in fast_strips_kernel hamming_match_kernel octree_kernel describe_fused_kernel tri_solve_kernel; do
  timeout 400 ncu --set full --clock-control none --import-source on -k regex:$K -s 2 -c 1 -o gpurun_out/ncu_$K -f $CMD > gpurun_out/ncu_run_$K.log 2>&1
  tail -1 gpurun_out/ncu_run_$K.log
done
ls -la gpurun_out/*.ncu-rep | awk '{print $5, $9}'
