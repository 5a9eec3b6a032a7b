#!/bin/bash
set -u
mkdir -p gpurun_out
run() { name=$1; shift
  env "$@" timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu 2>>gpurun_out/err.log | tail -1 > gpurun_out/b_$name.json
  python -c "import json; d=json.load(open('gpurun_out/b_$name.json')); print('$name', round(d['value']), round(d['e2e']['value']))"
}
run base X=1
run hi MCS_HIPRIO=1
run base2 X=1
run hi2 MCS_HIPRIO=1
