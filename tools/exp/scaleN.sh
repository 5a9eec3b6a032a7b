#!/bin/bash
# gpurun --gpus N payload: the bench line at N GPUs (torchrun, one rank per GPU), as the driver launches it
set -u
N=$1
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 2>gpurun_out/scale_n$N.err | grep '^{' | tail -1 > gpurun_out/scale_n$N.json
python -c "import json; d=json.load(open('gpurun_out/scale_n$N.json')); print(d['n_gpus'], round(d['value']), round(d['e2e']['value']))"
