#!/bin/bash
# gpurun payload: extract / rig parity with the 8-pixel pre-test, then bench A/B against the 4-pixel one
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_extract.py tests/test_gpu_rig.py -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/pytest_ab5.log
run() {  # name, env...
  local name=$1; shift
  env "$@" timeout 200 python bench.py --no-cpu --no-latency --steps 10 > gpurun_out/bench_$name.json 2>gpurun_out/bench_$name.err; local rc=$?
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/bench_$name.json"))
    s = d["stage_ms_per_launch_sequence"]
    print("$name", "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "fast", s["fast"], d["overlap"])
except Exception as e:
    print("$name", "no line rc=$rc", e)
PY
}
run px8 A=1
run px4 MCS_FAST_PX8=0
run px8b A=1
CMD="python bench.py --steps 2 --warmup 1 --frames 64 --distinct 2 --no-cpu --no-latency"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:fast_strips_kernel -s 2 -c 1 -o gpurun_out/ncu_fast_px8 -f $CMD > gpurun_out/ncu_run_px8.log 2>&1
tail -1 gpurun_out/ncu_run_px8.log
