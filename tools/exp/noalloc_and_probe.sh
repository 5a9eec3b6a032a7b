#!/bin/bash
# gpurun payload: the no-allocation entry point's test alone (fresh process), the TMA hand-over probe, then the whole GPU suite
set -u
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/pytest_noalloc.log 2>&1; echo "noalloc alone rc=$?" | tee -a gpurun_out/pytest_noalloc.log
tail -3 gpurun_out/pytest_noalloc.log
for v in A B C; do timeout 60 ./tools/tma_probe $v 2>&1 | tail -2; done | tee gpurun_out/tma_probe.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench.json 2>gpurun_out/bench.err; tail -c 1500 gpurun_out/bench.json
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
