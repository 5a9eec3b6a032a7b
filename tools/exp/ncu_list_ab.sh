#!/bin/bash
# gpurun payload: ncu launch list (duration + warp instructions) of kernel regex $1 for calls of $2 rig frames, one list per library
# variant given as name:path.so (name: alone = the product library)
set -u
mkdir -p gpurun_out
K=$1; NF=$2; shift 2
for spec in "$@"; do
  name=${spec%%:*}; lib=${spec#*:}
  CMD="python bench.py --steps 2 --warmup 2 --frames $NF --distinct 2 --no-cpu --no-latency"
  if [ -n "$lib" ]; then export MCS_LIB=$lib; else unset MCS_LIB; fi
  timeout 300 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -k regex:$K --csv --log-file gpurun_out/list_${name}.csv $CMD > gpurun_out/list_${name}.log 2>&1
  python - <<PY
import csv
rows = list(csv.reader(open("gpurun_out/list_${name}.csv")))
hdr = None; d = {}
for r in rows:
    if "Kernel Name" in r: hdr = r; continue
    if hdr is None or len(r) != len(hdr): continue
    x = dict(zip(hdr, r)); d.setdefault(int(x["ID"]), {"k": x["Kernel Name"][:30], "g": x["Grid Size"]})[x["Metric Name"][:9]] = x["Metric Value"]
for i in sorted(d)[-3:]: print("$name", d[i])
PY
done
