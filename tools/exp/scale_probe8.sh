#!/bin/bash
# gpurun --gpus 8 payload: the copy-only probe at 1 / 2 / 4 / 8 ranks (args: rank counts, default "1 2 4 8"), then the bench line at 2, 4 and 8 GPUs
set -u
mkdir -p gpurun_out
: > gpurun_out/pcie_probe.jsonl
for N in ${1:-1 2 4 8}; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29540 + N)) tools/pcie_probe.py 2>gpurun_out/probe_n$N.err | grep '^{' | tee -a gpurun_out/pcie_probe.jsonl | cut -c1-300
done
for N in ${2:-2 4 8}; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29510 + N)) bench.py --gpus $N --steps 10 --warmup 3 2>gpurun_out/scale_n$N.err | grep '^{' | tail -1 > gpurun_out/scale_n$N.json
  python -c "import json; d=json.load(open('gpurun_out/scale_n$N.json')); print(d['n_gpus'], round(d['value']), round(d['e2e']['value']), d['e2e']['d2h_bytes_per_step'], d.get('gathered_counts'))"
done
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1; lscpu | head -20 >> gpurun_out/topo.txt
