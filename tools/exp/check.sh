#!/bin/bash
# gpurun payload: GPU parity suite, the default bench line, smoke()
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench.json 2>gpurun_out/bench.err; tail -c 2500 gpurun_out/bench.json
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
