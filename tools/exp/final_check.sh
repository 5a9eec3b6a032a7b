#!/bin/bash
# gpurun payload: the whole GPU suite, smoke() and the default bench line (value, e2e, fractions, latency sweep, parity block)
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 400 python bench.py > gpurun_out/bench.json 2>gpurun_out/bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench.json"))
L = d["latency"]
print("value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "issue frac", round(d["roofline"]["issue"]["path"]["frac"], 4), "hbm frac", round(d["roofline"]["frac"], 4))
print({b: {m: L[b][m]["p50_ms"] for m in ("device", "device_graph", "host")} for b in ("batch_1", "batch_8", "batch_64")})
print(L["batch_1"]["stage_ms"]); print(d["parity"])
PY
