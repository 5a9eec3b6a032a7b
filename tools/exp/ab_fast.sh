#!/bin/bash
# gpurun payload: GPU parity suite, then the bench line with the TMA strip kernel (default) and with the previous strip kernel
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 120 python -m pytest tests -m gpu -q -s -k "orientation_flag or edge_cases_and_standalone" 2>&1 | grep -E "orientation mode|2-view|passed|failed" | tee gpurun_out/tols.log
MCS_FAST_TMA=0 timeout 240 python bench.py --no-cpu > gpurun_out/bench_strips.json 2>gpurun_out/bench_strips.err; echo "old rc=$?"
timeout 240 python bench.py --no-cpu > gpurun_out/bench_tma.json 2>gpurun_out/bench_tma.err; echo "tma rc=$?"
for f in strips tma; do python - <<PY
import json
try:
    d = json.load(open("gpurun_out/bench_$f.json"))
    print("$f", round(d["value"]), round(d["e2e"]["value"]), d["stage_ms_per_launch_sequence"])
except Exception as e:
    print("$f", "no line", e)
PY
done
