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


def build(launch_csv, frames_per_sequence=64, images_per_frame=5, sms=148):
    rows = [r for r in csv.reader(open(launch_csv)) if len(r) >= 15 and r[0].isdigit()]
    per_launch = {}
    for r in rows:
        d = per_launch.setdefault(int(r[0]), {"kernel": r[4], "grid": tuple(int(x) for x in r[8].strip("()").split(","))})
        d[r[12]] = float(r[14].replace(",", ""))
    # launch sequences in launch order: a new one starts where a pyramid launch follows a launch of another stage.  The bench command
    # also issues short sequences (the host-buffer leg ramps its chunks up), so only the sequences whose FAST launch covers
    # frames_per_sequence x images_per_frame images (grid y of the strip kernel) are kept.
    seqs, cur, prev = [], [], None
    for i in sorted(per_launch):
        d = per_launch[i]
        st = stage_of(d["kernel"])
        if st is None:
            continue
        if st == "pyramid" and prev not in (None, "pyramid") and cur:
            seqs.append(cur)
            cur = []
        cur.append((st, d))
        prev = st
    if cur:
        seqs.append(cur)
    want = frames_per_sequence * images_per_frame
    full = [q for q in seqs if any(st == "fast" and d["grid"][1] == want for st, d in q) and {st for st, _ in q} >= {"pyramid", "fast", "octree", "describe", "match", "merge", "triangulate"}]
    if not full:
        raise SystemExit("no complete %d-frame launch sequence in %s" % (frames_per_sequence, launch_csv))
    acc = {}
    for q in full:
        for st, d in q:
            if "smsp__inst_executed.sum" not in d:
                continue
            a = acc.setdefault(st, {"warp_insts": 0.0, "ns": 0.0, "launches": 0})
            a["warp_insts"] += d["smsp__inst_executed.sum"]
            a["ns"] += d.get("gpu__time_duration.sum", 0.0)
            a["launches"] += 1
    n_seq = len(full)
    kernels = {st: {"warp_insts": int(round(a["warp_insts"] / n_seq)), "launches_per_sequence": round(a["launches"] / n_seq, 2),
                    "ncu_ms": round(a["ns"] / n_seq / 1e6, 4)} for st, a in sorted(acc.items())}
    return {"_comment": "warp instructions (smsp__inst_executed.sum) per launch sequence of 64 rig frames (320 images), mean over the "
                        "complete 64-frame sequences of one ncu launch list; written by tools/issue_table.py, read by bench.py (roofline.issue)",
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
