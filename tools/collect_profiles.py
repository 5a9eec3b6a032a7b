#!/usr/bin/env python
"""Copies the outputs of tools/exp/r2_measure.sh from gpurun_out/ into profiles/ (tracked): bench lines, launch list -> issue.json,
ncu raw pages and per-line summaries of the five main kernels, traffic.json (DRAM bytes per launch, unit-aware).

    python tools/collect_profiles.py [prefix, default r2]
"""
import csv
import json
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
KERNELS = {"fast": "fast_strips_kernel", "match": "hamming_match_kernel", "describe": "describe_fused_kernel", "octree": "octree_kernel",
           "triangulate": "tri_solve_kernel"}
UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main():
    pre = sys.argv[1] if len(sys.argv) > 1 else "r2"
    for n in ("bench", "bench_config3", "bench_config4", "bench_confignoise", "bench_reference_arm"):
        shutil.copy(os.path.join(G, n + ".json"), os.path.join(P, "%s_%s.json" % (pre, n)))
    shutil.copy(os.path.join(G, "launches.csv"), os.path.join(P, pre + "_launches.csv"))
    subprocess.check_call([sys.executable, os.path.join(ROOT, "tools", "issue_table.py"), os.path.join(P, pre + "_launches.csv"), os.path.join(P, "issue.json")])
    traffic = json.load(open(os.path.join(P, "traffic.json")))
    for stage, k in KERNELS.items():
        rep = os.path.join(G, "ncu_%s.ncu-rep" % k)
        raw = os.path.join(P, "%s_%s_raw.csv" % (pre, k))
        with open(raw, "w") as f:
            subprocess.check_call(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=f, stderr=subprocess.DEVNULL)
        src = "/tmp/src_%s.csv" % k
        with open(src, "w") as f:
            subprocess.check_call(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], stdout=f, stderr=subprocess.DEVNULL)
        with open(os.path.join(P, "%s_%s_lines.txt" % (pre, k)), "w") as f:
            subprocess.check_call([sys.executable, os.path.join(ROOT, "tools", "ncu_lines.py"), src, "0.8"], stdout=f)
        r = list(csv.reader(open(raw)))
        h, un, v = r[0], r[1], r[2]
        b = sum(float(v[h.index(m)].replace(",", "")) * UNIT[un[h.index(m)]] for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
        traffic[stage]["dram_bytes_per_launch"] = int(b)
        traffic[stage]["source"] = "profiles/%s_%s_raw.csv" % (pre, k)
        print(stage, int(b), "B DRAM per launch,", v[h.index("gpu__time_duration.sum")], un[h.index("gpu__time_duration.sum")])
    json.dump(traffic, open(os.path.join(P, "traffic.json"), "w"), indent=2)


if __name__ == "__main__":
    main()
