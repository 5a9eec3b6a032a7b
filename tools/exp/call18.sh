#!/bin/bash
set -u
mkdir -p gpurun_out
run() { name=$1; shift
  env "$@" timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu 2>>gpurun_out/err.log | tail -1 > gpurun_out/b_$name.json
  python -c "import json; d=json.load(open('gpurun_out/b_$name.json')); print('$name', round(d['value']), round(d['e2e']['value']))"
}
run ns4_div2 X=1
run ns4_div4 MCS_HOST_CHUNK_DIV=4
run ns6_div4 MCS_LIB=$PWD/mc-slam_b200/libmcs_ns6.so MCS_HOST_CHUNK_DIV=4
run ns6_div3 MCS_LIB=$PWD/mc-slam_b200/libmcs_ns6.so MCS_HOST_CHUNK_DIV=3
run ns8_div4 MCS_LIB=$PWD/mc-slam_b200/libmcs_ns8.so MCS_HOST_CHUNK_DIV=4
run ns8_div8 MCS_LIB=$PWD/mc-slam_b200/libmcs_ns8.so MCS_HOST_CHUNK_DIV=8
