#!/usr/bin/env python
"""Randomised soak of the CUDA path against the CPU oracle (test infrastructure: this tool imports oracle/, the product never does).

    python tools/soak.py [--seconds 240] [--seed 0] [--out gpurun_out/soak.json]

Two loops, each for half of the time budget, every case drawn from a seeded generator and printed on failure so that it can be replayed:
  extractor : random image size, nfeatures, scaleFactor, nlevels, iniThFAST / minThFAST and image statistics (ray-cast view blended
              with noise at a random ratio, random contrast) — mono flag, keypoints (order included) and descriptor bits against
              the oracle's ORBextractor::operator() restatement;
  rig       : random camera count (2..6), size, quota, `old` flag, masks on / off, 1..3 frames per call — features, IntraMatch rows,
              mask / depth-gate decisions, representative cameras exactly, points to 1e-6 relative.
Sizes are kept where the reference itself is defined (every pyramid level keeps at least one 35-pixel cell row and column).
Exit code 1 and a JSON report with the failing cases when anything differs.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def blend(M, calib, w, h, rng):
    base = M.synth.make_rig_frame(calib, w, h, frame_index=int(rng.integers(0, 10000)), n_rects=int(rng.integers(3, 10)))[0].astype(np.int32)
    noise = M.synth.noise_image(w, h, int(rng.integers(0, 1 << 30))).astype(np.int32)
    a = int(rng.integers(0, 5))                      # 0: pure noise ... 4: pure view
    img = (a * base + (4 - a) * noise) // 4
    gain = float(rng.choice([0.25, 0.5, 1.0, 1.0, 1.5]))
    return np.clip((img - 128) * gain + 128, 0, 255).astype(np.uint8)


def draw_extractor(rng):
    while True:
        w, h = int(rng.integers(96, 900)), int(rng.integers(96, 700))
        sf = float(rng.choice([1.1, 1.2, 1.2, 1.25, 1.3, 1.5, 2.0]))
        nl = int(rng.integers(1, 11))
        if min(w, h) / sf ** (nl - 1) >= 80:
            break
    nf = int(rng.integers(50, 3000))
    ini = int(rng.integers(5, 60))
    mn = int(rng.integers(2, ini + 1))
    return dict(w=w, h=h, nf=nf, sf=sf, nl=nl, ini=ini, mn=mn)


def run_extractor(M, O, case, rng):
    w, h = case["w"], case["h"]
    calib = M.synth.make_rig(1, w, h)
    img = blend(M, calib, w, h, rng)
    ex = M.Extractor(case["nf"], scale_factor=case["sf"], nlevels=case["nl"], ini_th=case["ini"], min_th=case["mn"], max_width=w, max_height=h)
    orc = O.Extractor(case["nf"], case["sf"], case["nl"], case["ini"], case["mn"])
    mono, kps, desc = ex(img)
    m2, k2, d2 = orc(img)
    if mono != m2 or len(kps) != len(k2):
        return "counts: gpu %d/%d oracle %d/%d" % (mono, len(kps), m2, len(k2))
    if not (kps == k2).all():
        return "keypoints differ at %s" % np.nonzero(kps != k2)[0][:5].tolist()
    if not np.array_equal(desc, d2):
        return "descriptor bits differ: %d" % int(np.unpackbits(desc ^ d2).sum())
    case["keypoints"] = len(kps)
    return None


def draw_rig(rng):
    ncam = int(rng.integers(2, 7))
    w, h = int(rng.choice([200, 256, 320, 417, 512, 720])), int(rng.choice([160, 192, 240, 312, 384, 540]))
    return dict(ncam=ncam, w=w, h=h, nf=int(rng.integers(100, 1200)), old=bool(rng.integers(0, 2)), masks=bool(rng.integers(0, 2)),
                frames=int(rng.integers(1, 4)), spacing=float(rng.choice([0.08, 0.12, 0.2])))


def run_rig(M, O, case, rng):
    ncam, w, h, nf = case["ncam"], case["w"], case["h"], case["nf"]
    calib = M.synth.make_rig(ncam, w, h, spacing=case["spacing"])
    n = case["frames"]
    fi = int(rng.integers(0, 10000))
    frames = np.stack([M.synth.make_rig_frame(calib, w, h, frame_index=fi + i, n_rects=int(rng.integers(3, 9))) for i in range(n)])
    masks = np.stack([M.synth.make_masks(ncam, w, h, frame_index=fi + 17 + i) for i in range(n)]) if case["masks"] else None
    rig = M.RigFrames(calib, w, h, nfeatures=nf, max_frames=n)
    rig.process(frames, masks=masks, old=case["old"])
    total = 0
    for f in range(n):
        ex = [O.Extractor(nf) for _ in range(ncam)]
        K, D = [], []
        for c in range(ncam):
            _, k, d = ex[c](frames[f][c])
            K.append(k)
            D.append(d)
        rows, _ = O.intra_matches(D, K, calib, ex[0].sigma2, use_epipolar=case["old"])
        tri = O.triangulate_rows(calib, K, D, rows, masks=None if masks is None else masks[f])
        for c in range(ncam):
            k, d = rig.features(f, c)
            if len(k) != len(K[c]) or not (k == K[c]).all() or not np.array_equal(d, D[c]):
                return "features frame %d cam %d" % (f, c)
        r = rig.intra_matches(f)
        if r.shape != rows.shape or not np.array_equal(r, rows):
            return "IntraMatch rows frame %d: gpu %s oracle %s" % (f, r.shape, rows.shape)
        lm = rig.landmarks(f)
        for key in ("rows", "nviews", "status", "rep_cam"):
            if not np.array_equal(lm[key], tri[key]):
                return "landmark %s frame %d" % (key, f)
        ok = tri["nviews"] >= 2
        if not np.allclose(lm["xyz"][ok], tri["xyz"][ok], rtol=1e-6, atol=1e-9):
            return "xyz frame %d: %g" % (f, float(np.abs(lm["xyz"][ok] - tri["xyz"][ok]).max()))
        total += len(rows)
    case["rows"] = total
    return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=240.0)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    import mcslam_b200 as M
    import oracle_lib as O
    report = {"seed": args.seed, "extractor_cases": 0, "rig_cases": 0, "failures": []}
    for name, draw, run in (("extractor", draw_extractor, run_extractor), ("rig", draw_rig, run_rig)):
        rng = np.random.default_rng([args.seed, len(name)])
        t_end = time.time() + args.seconds / 2
        i = 0
        while time.time() < t_end:
            case = draw(rng)
            sub = np.random.default_rng([args.seed, len(name), i])
            try:
                err = run(M, O, case, sub)
            except Exception as e:  # a library error on a legal configuration is a finding too
                err = "exception: %r" % (e,)
            if err:
                report["failures"].append({"loop": name, "index": i, "case": case, "what": err})
                print("FAIL", name, i, case, err, flush=True)
            i += 1
        report[name + "_cases"] = i
    print(json.dumps({k: (v if k != "failures" else len(v)) for k, v in report.items()}), flush=True)
    if args.out:
        with open(args.out, "w") as f:
            json.dump(report, f, indent=1)
    return 1 if report["failures"] else 0


if __name__ == "__main__":
    sys.exit(main())
