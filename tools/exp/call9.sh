#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_extract.py -m gpu -x -q 2>&1 | tail -1
for nit in 0 1 2 4 8; do
  MCS_RS_NIT=$nit timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu 2>>gpurun_out/err.log | tail -1 > gpurun_out/b_nit$nit.json
  python -c "import json; d=json.load(open('gpurun_out/b_nit$nit.json')); s=d['stage_ms_per_launch_sequence']; print('nit$nit', round(d['value']), round(d['e2e']['value']), s['pyramid'], s['fast'])"
done
