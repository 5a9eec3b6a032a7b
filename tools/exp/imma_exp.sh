#!/bin/bash
# gpurun payload: the integer-MMA matcher experiment (MCS_MATCH_IMMA=1): rig parity tests with it, then bench lines with and without
set -u
mkdir -p gpurun_out
MCS_MATCH_IMMA=1 timeout 900 python -m pytest tests/test_gpu_rig.py tests/test_gpu_vs_ref.py -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/pytest_imma.log
for cfg in 2 3 4; do
  for v in 0 1; do
    MCS_MATCH_IMMA=$v timeout 300 python bench.py --config $cfg --no-cpu --no-latency --steps 6 > gpurun_out/imma_c${cfg}_$v.json 2>gpurun_out/imma_c${cfg}_$v.err
    python - <<PY
import json
try:
    d = json.load(open("gpurun_out/imma_c${cfg}_$v.json"))
    s = d["stage_ms_per_launch_sequence"]
    print("config $cfg imma=$v value", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "match ms", s["match"], "frames", s["frames"])
except Exception as e:
    print("config $cfg imma=$v no line", e)
PY
  done
done
