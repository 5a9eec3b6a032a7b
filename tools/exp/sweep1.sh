#!/bin/bash
# chunk size x frames-per-step sweep of the device-resident and host-buffer paths (no CPU baseline)
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
: > gpurun_out/sweep1.jsonl
for frames in 256 512 1024; do
  for chunk in 32 64 128; do
    MCS_CHUNK_FRAMES=$chunk timeout 300 python bench.py --steps 10 --warmup 3 --frames $frames --no-cpu 2>gpurun_out/sweep_err.log | tail -1 | \
      python -c "import sys,json; d=json.loads(sys.stdin.read()); print(json.dumps({'frames':$frames,'chunk':$chunk,'value':d['value'],'e2e':d['e2e']['value']}))" | tee -a gpurun_out/sweep1.jsonl
  done
done
