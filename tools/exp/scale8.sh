#!/bin/bash
set -u
mkdir -p gpurun_out
free -g | head -2
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/scale_n8.json 2>gpurun_out/scale_n8.err
tail -c 900 gpurun_out/scale_n8.json; tail -3 gpurun_out/scale_n8.err
free -g | head -2
