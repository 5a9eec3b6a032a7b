#!/bin/bash
# gpurun payload: parity of the extract / rig tests, A/B of the TMA strip kernel's launch shapes, ncu capture of the TMA kernel
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_extract.py tests/test_gpu_rig.py -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
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
run tma_dyn A=1
run tma_b5 MCS_FT_BLOCKS=5
run tma_b4 MCS_FT_BLOCKS=4
run tma_oneitem MCS_FT_GRID_MULT=100000
run tma_x2 MCS_FT_GRID_MULT=2
run strips MCS_FAST_TMA=0
CMD="python bench.py --steps 2 --warmup 1 --frames 64 --distinct 2 --no-cpu"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fast_strips_tma -s 1 -c 1 -o gpurun_out/ncu_fast_tma -f $CMD > gpurun_out/ncu_run.log 2>&1
tail -2 gpurun_out/ncu_run.log
