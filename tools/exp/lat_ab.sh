#!/bin/bash
# gpurun payload: batch-1 / 8 latency under different octree block sizes (and any other knob given as name:ENV=V,...)
set -u
mkdir -p gpurun_out
for spec in "$@"; do
  name=${spec%%:*}; envs=${spec#*:}
  env $(echo "$envs" | tr ',' ' ') timeout 240 python bench.py --no-cpu --steps 3 --frames 64 --distinct 2 --latency-iters 100 > gpurun_out/lat_$name.json 2>gpurun_out/lat_$name.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/lat_$name.json"))
    L = d["latency"]
    print("$name", {b: {m: L[b][m].get("p50_ms") for m in L[b] if m != "stage_ms"} for b in L if b.startswith("batch")}, d["stage_ms_per_launch_sequence"]["octree"])
    print("   batch-1 stages", L["batch_1"].get("stage_ms"))
    print("   batch-8 stages", L["batch_8"].get("stage_ms"))
except Exception as e:
    print("$name", "no line", e)
PY
done
