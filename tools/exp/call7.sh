#!/bin/bash
set -u
mkdir -p gpurun_out
for v in rs4 rs2; do
  MCS_LIB=$PWD/mc-slam_b200/libmcs_$v.so timeout 600 python -m pytest tests/test_gpu_extract.py -m gpu -x -q 2>&1 | tail -1
  MCS_LIB=$PWD/mc-slam_b200/libmcs_$v.so timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu 2>>gpurun_out/err.log | tail -1 > gpurun_out/b_$v.json
  python -c "import json; d=json.load(open('gpurun_out/b_$v.json')); s=d['stage_ms_per_launch_sequence']; print('$v', round(d['value']), round(d['e2e']['value']), s['pyramid'], s['fast'])"
done
