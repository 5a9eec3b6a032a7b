"""GPU: the bench line the driver reads — one short run of bench.py's own arm on a small workload, checked for the contract's keys."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_bench_line_contract_small_run():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--frames", "24", "--distinct", "1", "--steps", "2", "--warmup", "3",
                          "--latency-iters", "8", "--cpu-seconds", "1"], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-3000:]
    lines = [x for x in out.stdout.splitlines() if x.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["unit"] == "rig-frames/s" and d["higher_is_better"] is True and d["scaling"] == "weak" and d["dtype"] == "u8" and d["n_gpus"] == 1
    assert d["value"] > 100 and d["e2e"]["value"] > 100 and d["e2e"]["h2d_bytes_per_step"] == 24 * 5 * 720 * 540 and d["e2e"]["d2h_bytes_per_step"] > 0
    assert d["steps"] == 2 and d["warmup"] == 3 and d["gpu_launches"] >= 2 * 15
    assert d["config"]["workload"].startswith("config2") and d["vs_baseline"] is None
    r = d["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-12 and r["path"]["bytes_per_rig_frame"] == 13185090
    assert set(d["clocks"]) >= {"sm_mhz", "sm_max_mhz", "reasons"}
    p = d["parity"]
    assert p["mismatched_keypoints"] == p["mismatched_descriptor_bits"] == p["mismatched_intramatch_rows"] == p["mismatched_depth_gate"] == 0
    assert p["xyz_max_rel_err"] <= p["xyz_tolerance"]
    c = d["cpu_baseline"]
    assert c["kind"] in ("reference", "port") and c["value"] > 0 and c["cores"] >= 1 and set(c["stage_ms_one_frame"]) == {"extract", "intra_match", "triangulate"}
    lat = d["latency"]["batch_1"]
    assert lat["device"]["p50_ms"] > 0 and lat["device_graph"]["p50_ms"] > 0 and lat["host"]["p50_ms"] > 0 and "error" not in lat["device_graph"]
    g = d["gathered_counts"]
    assert g["frames"] == 24 and g["keypoints"] > 24 * 5 * 500
