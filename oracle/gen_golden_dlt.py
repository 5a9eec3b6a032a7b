"""Golden vectors for the N-view DLT (cv::sfm::triangulatePoints as MultiCameraFrame::triangulateIntraMatches calls it,
/root/reference/MCSlam/src/MultiCameraFrame.cpp:411-441): the null vector of the (3n) x (4+n) design from OpenCV's own SVD.

TEST INFRASTRUCTURE ONLY.  Run in this container (cv2 4.13.0):  python oracle/gen_golden_dlt.py  ->  tests/golden/dlt_cv2.npz
200 points seen by 2..5 of the 5 cameras of the synthetic config-2 rig, pixel noise sigma 0.3; per trial the view list, the pixel
coordinates and X = homogeneousToEuclidean(last row of vt) from cv2.SVDecomp(A, SVD_FULL_UV).  The oracle's one-sided Jacobi and the
CUDA kernels must reproduce X to 1e-10 relative (measured 1.6e-13); the product's own bar stays 1e-6.  Plus 300 two-view f32 solves
(cameras 0 and 4) from cv2.SVDecomp on the f32 4 x 4 system: 1e-5 relative (measured 9.6e-7; the product's bar is 1e-3).
"""
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, ".."))
import mcslam_b200 as M  # noqa: E402

calib = M.synth.make_rig(5, 720, 540)
rng = np.random.default_rng(3)
views_all, xs_all, ys_all, xyz = [], [], [], []
for trial in range(200):
    X = np.array([rng.uniform(-3, 3), rng.uniform(-2, 2), rng.uniform(1.5, 30)])
    nv = int(rng.integers(2, 6))
    views = sorted(rng.choice(5, nv, replace=False).tolist())
    xs, ys = [], []
    for c in views:
        K, R, t = calib[c]["K"].reshape(3, 3), calib[c]["R"].reshape(3, 3), calib[c]["t"]
        p = K @ (R @ X + t)
        xs.append(p[0] / p[2] + rng.normal(0, 0.3))
        ys.append(p[1] / p[2] + rng.normal(0, 0.3))
    A = np.zeros((3 * nv, 4 + nv))
    for i, c in enumerate(views):
        K, R, t = calib[c]["K"].reshape(3, 3), calib[c]["R"].reshape(3, 3), calib[c]["t"]
        A[3 * i:3 * i + 3, :4] = -(K @ np.hstack([R, t.reshape(3, 1)]))
        A[3 * i, 4 + i], A[3 * i + 1, 4 + i], A[3 * i + 2, 4 + i] = xs[i], ys[i], 1.0
    _, _, vt = cv2.SVDecomp(A, flags=cv2.SVD_FULL_UV)
    z = vt[-1]
    v = np.full(5, -1, np.int32); v[:nv] = views
    a = np.zeros(5); a[:nv] = xs
    b = np.zeros(5); b[:nv] = ys
    views_all.append(v); xs_all.append(a); ys_all.append(b); xyz.append(z[:3] / z[3])
# 2-view f32 DLT (MultiCameraFrame::triangulate, :481-507): 4 x 4 f32 A from x * P.row(2) - P.row(0) ..., SVD::compute(MODIFY_A | FULL_UV),
# last row of vt divided by w — cameras 0 and 4, 300 exact projections narrowed to f32
rng = np.random.default_rng(5)
P3 = np.stack([rng.uniform(-3, 3, 300), rng.uniform(-2, 2, 300), rng.uniform(2, 30, 300)], 1)
pair_pts, Ps = [], []
for c in (0, 4):
    K, R, t = calib[c]["K"].reshape(3, 3), calib[c]["R"].reshape(3, 3), calib[c]["t"]
    uv = (P3 @ R.T + t) @ K.T
    pair_pts.append((uv[:, :2] / uv[:, 2:]).astype(np.float32))
    Ps.append(K @ np.hstack([R, t.reshape(3, 1)]))
pair_xyz = np.zeros((300, 3), np.float32)
for i in range(300):
    A = np.zeros((4, 4), np.float32)
    for c in range(4):
        A[0, c] = np.float32(np.float64(pair_pts[0][i, 0]) * Ps[0][2, c] - Ps[0][0, c])
        A[1, c] = np.float32(np.float64(pair_pts[0][i, 1]) * Ps[0][2, c] - Ps[0][1, c])
        A[2, c] = np.float32(np.float64(pair_pts[1][i, 0]) * Ps[1][2, c] - Ps[1][0, c])
        A[3, c] = np.float32(np.float64(pair_pts[1][i, 1]) * Ps[1][2, c] - Ps[1][1, c])
    _, _, vt = cv2.SVDecomp(A, flags=cv2.SVD_MODIFY_A | cv2.SVD_FULL_UV)
    pair_xyz[i] = vt[3][:3] / vt[3][3]
out = os.path.join(HERE, "..", "tests", "golden", "dlt_cv2.npz")
np.savez_compressed(out, calib=calib, views=np.array(views_all), xs=np.array(xs_all), ys=np.array(ys_all), xyz=np.array(xyz),
                    pair_p1=pair_pts[0], pair_p2=pair_pts[1], pair_xyz=pair_xyz, cv2_version=cv2.__version__)
print("wrote", out, len(xyz), "points")
