#!/bin/bash
# gpurun payload: parity of the extract / rig tests in the three FAST staging modes, A/B bench lines
set -u
mkdir -p gpurun_out
for m in 1 2 0; do
MCS_FAST_TMA=$m timeout 600 python -m pytest tests/test_gpu_extract.py tests/test_gpu_rig.py -m gpu -x -q 2>&1 | tail -2
done | tee gpurun_out/pytest_modes.log
run() {  # name, env...
  local name=$1; shift
  env "$@" timeout 200 python bench.py --no-cpu --steps 10 > gpurun_out/bench_$name.json 2>gpurun_out/bench_$name.err; local rc=$?
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/bench_$name.json"))
    s = d["stage_ms_per_launch_sequence"]
    print("$name", "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "fast", s["fast"], "sum", round(sum(v for k, v in s.items() if k != "frames"), 3))
except Exception as e:
    print("$name", "no line rc=$rc", e)
PY
}
run tma1 MCS_FAST_TMA=1
run tma2 MCS_FAST_TMA=2
run tma0 MCS_FAST_TMA=0
run tma1b MCS_FAST_TMA=1
