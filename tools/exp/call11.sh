#!/bin/bash
set -u
mkdir -p gpurun_out
for nt in 128 256 512; do
  MCS_OCT_THREADS=$nt timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu 2>>gpurun_out/err.log | tail -1 > gpurun_out/b_oct$nt.json
  python -c "import json; d=json.load(open('gpurun_out/b_oct$nt.json')); s=d['stage_ms_per_launch_sequence']; print('cfg2 oct$nt', round(d['value']), round(d['e2e']['value']), s['octree'])"
  MCS_OCT_THREADS=$nt timeout 300 python bench_configs.py --config 3 --steps 5 2>>gpurun_out/err.log | tail -1 > gpurun_out/c3_oct$nt.json
  python -c "import json; d=json.load(open('gpurun_out/c3_oct$nt.json')); s=d['stage_ms_per_launch_sequence']; print('cfg3 oct$nt', round(d['value']), round(d['e2e']['value']), s['octree'])"
  MCS_OCT_THREADS=$nt timeout 400 python bench_configs.py --config 4 --steps 5 2>>gpurun_out/err.log | tail -1 > gpurun_out/c4_oct$nt.json
  python -c "import json; d=json.load(open('gpurun_out/c4_oct$nt.json')); s=d['stage_ms_per_launch_sequence']; print('cfg4 oct$nt', round(d['value'],1), round(d['e2e']['value'],1), s['octree'])"
done
