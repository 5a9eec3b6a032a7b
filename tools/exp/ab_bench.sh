#!/bin/bash
# gpurun payload: A/B of tuning knobs.  usage: ab_bench.sh [-t "pytest args"] name1:ENV=V,ENV2=V  name2:  ...
# Runs the named GPU tests first (parity before speed), then one short bench line per variant with the variant's environment.
set -u
mkdir -p gpurun_out
if [ "${1:-}" = "-t" ]; then
  timeout 1200 python -m pytest $2 -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/pytest_ab.log
  shift 2
fi
for spec in "$@"; do
  name=${spec%%:*}; envs=${spec#*:}
  env $(echo "$envs" | tr ',' ' ') timeout 240 python bench.py --no-cpu --no-latency --steps 10 > gpurun_out/bench_$name.json 2>gpurun_out/bench_$name.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/bench_$name.json"))
    s = d["stage_ms_per_launch_sequence"]
    print("$name", "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), {k: s[k] for k in ("pyramid", "fast", "octree", "describe", "match", "triangulate")}, d["overlap"]["amortised_ms"])
except Exception as e:
    print("$name", "no line", e)
PY
done
