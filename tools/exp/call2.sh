#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
for v in 0 1; do
  MCS_FAST_CELLS=$v timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu 2>>gpurun_out/err.log | tail -1 > gpurun_out/bench_cells$v.json
  python -c "import json; d=json.load(open('gpurun_out/bench_cells$v.json')); print($v, d['value'], d['e2e']['value'], d['stage_ms_per_launch_sequence'], d.get('parity'))"
done
timeout 600 compute-sanitizer --tool memcheck python -m pytest tests/test_gpu_extract.py -m gpu -x -q -k "fallback or vs_golden" > gpurun_out/sanitizer.log 2>&1; tail -5 gpurun_out/sanitizer.log
