#!/usr/bin/env python
"""profiles/issue.json — warp instructions per launch sequence, per stage, for bench.py's instruction-issue roofline.

Input: the ncu launch list of the 64-frame bench command,
    ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -c 600 --csv --log-file X.csv \
        python bench.py --steps 2 --warmup 1 --frames 64 --distinct 2 --no-cpu --no-latency
(long CSV: one row per launch and metric).  Every launch sequence of that command is one unchunked pass over 64 rig frames: 7 resize
launches, FAST, octree, describe, match, merge and 3 triangulation kernels.  Per stage the table holds the mean over the sequences.

    python tools/issue_table.py profiles/r2_launches.csv [profiles/issue.json]
"""
import csv
import json
import os
import sys

STAGE_OF = [("resize_level", "pyramid"), ("fast_strips", "fast"), ("fast_cells", "fast"), ("octree_kernel", "octree"), ("describe", "describe"),
            ("blur_kernel", "describe"), ("hamming_match", "match"), ("merge_kernel", "merge"), ("tri_", "triangulate"),
            ("lapping_permute", "describe"), ("undistort", "pyramid"), ("stage_f32", "pyramid")]


def stage_of(kernel):
    for key, st in STAGE_OF:
        if key in kernel:
            return st
    return None


def build(launch_csv, frames_per_sequence=64, sms=148):
    rows = [r for r in csv.reader(open(launch_csv)) if len(r) >= 15 and r[0].isdigit()]
    per_launch = {}
    for r in rows:
        d = per_launch.setdefault(int(r[0]), {"kernel": r[4]})
        d[r[12]] = float(r[14].replace(",", ""))
    n_seq = sum(1 for d in per_launch.values() if stage_of(d["kernel"]) == "fast")
    if n_seq < 1:
        raise SystemExit("no FAST launch in %s" % launch_csv)
    acc = {}
    for d in per_launch.values():
        st = stage_of(d["kernel"])
        if st is None or "smsp__inst_executed.sum" not in d:
            continue
        a = acc.setdefault(st, {"warp_insts": 0.0, "ns": 0.0, "launches": 0})
        a["warp_insts"] += d["smsp__inst_executed.sum"]
        a["ns"] += d.get("gpu__time_duration.sum", 0.0)
        a["launches"] += 1
    kernels = {st: {"warp_insts": int(round(a["warp_insts"] / n_seq)), "launches_per_sequence": round(a["launches"] / n_seq, 2),
                    "ncu_ms": round(a["ns"] / n_seq / 1e6, 4)} for st, a in sorted(acc.items())}
    return {"_comment": "warp instructions (smsp__inst_executed.sum) per launch sequence of 64 rig frames (320 images), mean over the "
                        "sequences of one ncu launch list; written by tools/issue_table.py, read by bench.py (roofline.issue)",
            "source": os.path.relpath(launch_csv, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")),
            "frames_per_sequence": frames_per_sequence, "sequences_in_capture": n_seq, "sms": sms, "sm_mhz_fallback": 1965.0, "kernels": kernels}


if __name__ == "__main__":
    src = sys.argv[1]
    dst = sys.argv[2] if len(sys.argv) > 2 else os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "profiles", "issue.json")
    t = build(src)
    json.dump(t, open(dst, "w"), indent=1)
    tot = sum(k["warp_insts"] for k in t["kernels"].values())
    print("%d sequences; %.3f G warp instructions per sequence = %.3f ms at %d SMs x 4 x 1965 MHz" % (t["sequences_in_capture"], tot / 1e9, 1e3 * tot / (t["sms"] * 4 * 1965e6), t["sms"]))
    for k, v in t["kernels"].items():
        print("  %-12s %12d  %5.2f launches  %.4f ms (ncu)" % (k, v["warp_insts"], v["launches_per_sequence"], v["ncu_ms"]))
