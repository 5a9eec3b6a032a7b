#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python bench_configs.py --config 3 > gpurun_out/bench_config3.json 2>gpurun_out/c3.err; tail -c 600 gpurun_out/bench_config3.json; tail -3 gpurun_out/c3.err
timeout 900 python bench_configs.py --config 4 > gpurun_out/bench_config4.json 2>gpurun_out/c4.err; tail -c 600 gpurun_out/bench_config4.json; tail -3 gpurun_out/c4.err
