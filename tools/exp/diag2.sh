#!/bin/bash
# gpurun --gpus 2 payload: a short two-rank bench run with progress lines, bounded tightly
set -u
mkdir -p gpurun_out
NCCL_DEBUG=WARN timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29577 bench.py --gpus 2 --steps 3 --warmup 3 --frames 64 --distinct 1 > gpurun_out/diag2.json 2>gpurun_out/diag2.err
echo "rc=$?"; tail -c 700 gpurun_out/diag2.json; echo; grep -E "bench rank|Error|error|Traceback|NCCL WARN" gpurun_out/diag2.err | tail -30
