"""Generates tests/golden/*.npz with the cv2-based tier-1 oracle (oracle/cv2_tier1.py, OpenCV 4.13.0 here).
Run from the repo root in a container that has cv2:  python oracle/gen_golden.py
The fixtures pin the portable C++ oracle (tests/test_oracle_golden.py) and, through it, the CUDA path."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(HERE, ".."))
import cv2  # noqa: E402
import cv2_tier1 as T  # noqa: E402
import mcslam_b200  # noqa: E402
from mcslam_b200 import synth  # noqa: E402

OUT = os.path.join(HERE, "..", "tests", "golden")
os.makedirs(OUT, exist_ok=True)


def single(name, img, nfeatures, **kw):
    t = T.ORBextractorCv2(nfeatures, **kw)
    mono, kps, desc = t(img)
    d = dict(image=img, nfeatures=nfeatures, mono=mono, kps=kps, desc=desc, quota=t.quota, scale=t.scale, sigma2=t.sigma2,
             inv_scale=t.inv_scale, inv_sigma2=t.inv_sigma2, level_counts=np.array([len(k) for k in t.level_kps], np.int32))
    for l in range(t.nlevels):
        d["cand%d" % l] = t.cands[l]
        d["pyr%d" % l] = t.pyr[l]
        if t.blurred[l] is not None:
            d["blur%d" % l] = t.blurred[l]
    d["pyr_padded3"] = t.pyr_padded[3]
    np.savez_compressed(os.path.join(OUT, name), **d)
    print(name, img.shape, "kps", len(kps), "cands", [len(c) for c in t.cands])


def main():
    print("cv2", cv2.__version__)
    calib = synth.make_rig(3, 480, 360)
    frame = synth.make_rig_frame(calib, 480, 360, frame_index=0, n_rects=10)
    # (1) one small image with every intermediate
    single("single_256x192.npz", np.ascontiguousarray(frame[0][60:252, 100:356]), 300)
    # (2) pure noise (worst-case candidate counts, many score ties), odd size (non multiple-of-4 width)
    single("noise_203x157.npz", synth.noise_image(203, 157, 5), 400)
    # (3) low-contrast image: exercises the minThFAST fallback cells
    low = (frame[1][40:280, 60:380].astype(np.float32) * 0.18 + 100).astype(np.uint8)
    single("lowcontrast_320x240.npz", low, 500)
    # (4) a 3-camera rig frame: features, pair matches, IntraMatch rows, triangulated points
    t = [T.ORBextractorCv2(600) for _ in range(3)]
    K, D = [], []
    for c in range(3):
        _, k, d = t[c](frame[c])
        K.append(k)
        D.append(d)
    rows, pairs = T.compute_intra_matches(D)
    xyz = T.triangulate_numpy(calib, K, rows)
    d = dict(images=frame, calib=calib, rows=rows, xyz=xyz, sigma2=t[0].sigma2)
    for c in range(3):
        d["kps%d" % c] = K[c]
        d["desc%d" % c] = D[c]
    for p, (i1, i2) in enumerate(pairs):
        d["pair%d_i1" % p] = i1
        d["pair%d_i2" % p] = i2
    np.savez_compressed(os.path.join(OUT, "rig3_480x360.npz"), **d)
    print("rig3: kps", [len(k) for k in K], "rows", len(rows), "pairs", [len(p[0]) for p in pairs])
    # (5) orientation primitive: cv2.fastAtan2 samples for the IC_Angle flag mode
    rng = np.random.default_rng(3)
    yx = rng.integers(-40000, 40000, (4000, 2)).astype(np.float32)
    ang = np.array([cv2.fastAtan2(float(a), float(b)) for a, b in yx], np.float32)
    np.savez_compressed(os.path.join(OUT, "fast_atan2.npz"), yx=yx, ang=ang)


if __name__ == "__main__":
    main()
