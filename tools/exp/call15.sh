#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu 2>>gpurun_out/err.log | tail -1 > gpurun_out/b_cur.json
python -c "import json; d=json.load(open('gpurun_out/b_cur.json')); s=d['stage_ms_per_launch_sequence']; print(round(d['value']), round(d['e2e']['value']), s)"
bash tools/exp/ncu_one.sh octree_kernel 1
