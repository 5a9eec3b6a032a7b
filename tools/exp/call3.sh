#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
run() { # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu 2>>gpurun_out/err.log | tail -1 > gpurun_out/b_$name.json
  python -c "import json; d=json.load(open('gpurun_out/b_$name.json')); s=d['stage_ms_per_launch_sequence']; print('$name', round(d['value']), round(d['e2e']['value']), s['fast'], s['octree'])"
}
run base X=1
run hi MCS_HIPRIO=1
run fsb5 MCS_FS_BLOCKS=5
run fsb4 MCS_FS_BLOCKS=4
run hi_fsb5 MCS_HIPRIO=1 MCS_FS_BLOCKS=5
run hi_fsb4 MCS_HIPRIO=1 MCS_FS_BLOCKS=4
run hi_fsb3 MCS_HIPRIO=1 MCS_FS_BLOCKS=3
