#!/bin/bash
set -u
mkdir -p gpurun_out
run() { name=$1; shift
  env "$@" timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu 2>>gpurun_out/err.log | tail -1 > gpurun_out/b_$name.json
  python -c "import json; d=json.load(open('gpurun_out/b_$name.json')); s=d['stage_ms_per_launch_sequence']; print('$name', round(d['value']), round(d['e2e']['value']), s['triangulate'])"
}
run base X=1
run tri1 MCS_TRI_PER_SM=1
run tri2 MCS_TRI_PER_SM=2
run tri3 MCS_TRI_PER_SM=3
run tri1hi MCS_TRI_PER_SM=1 MCS_HIPRIO=1
run tri2hi MCS_TRI_PER_SM=2 MCS_HIPRIO=1
