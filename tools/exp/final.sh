#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench_final.json 2>gpurun_out/bench_final.err; tail -c 1800 gpurun_out/bench_final.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2>>gpurun_out/bench_final.err; tail -c 600 gpurun_out/bench_ref.json
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
CMD="python bench.py --steps 2 --warmup 1 --frames 64 --distinct 2 --no-cpu"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_l.log 2>&1
