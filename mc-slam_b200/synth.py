"""Deterministic synthetic rig frames (numpy only) — the workload BASELINE.json's configs name.

Scene: a far textured plane plus textured fronto-parallel rectangles at 1.5-15 m, ray-cast into every camera of a
linear rig through its K[R|t], so that cross-camera matches exist and triangulate to known depths.  Rig: cameras on
a line, 0.165 m spacing (the baseline in the reference's testTriangulateIntraMatches, MultiCameraFrame.cpp:384),
identity rotations perturbed by <= 1 degree, fx = fy = 0.8 W, principal point at the centre, zero distortion (so
UndistortKeyPoints is a copy, MultiCameraFrame.cpp:302-305).  seed = 1234 + frame_index.
"""
import numpy as np

CALIB_DTYPE = np.dtype([("K", "<f8", (9,)), ("R", "<f8", (9,)), ("t", "<f8", (3,)), ("dist", "<f8", (5,))])


def _rot(rx, ry, rz):
    cx, sx, cy, sy, cz, sz = np.cos(rx), np.sin(rx), np.cos(ry), np.sin(ry), np.cos(rz), np.sin(rz)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def make_rig(num_cams=5, w=720, h=540, spacing=0.165, seed=7, perturb_deg=1.0):
    """Returns a CALIB_DTYPE array [num_cams]; camera 0 is the reference (R = I, t = 0)."""
    rng = np.random.default_rng(seed)
    calib = np.zeros(num_cams, CALIB_DTYPE)
    f = 0.8 * w
    for c in range(num_cams):
        K = np.array([[f, 0, w / 2.0], [0, f, h / 2.0], [0, 0, 1.0]])
        ang = np.deg2rad(rng.uniform(-perturb_deg, perturb_deg, 3)) if c > 0 else np.zeros(3)
        R = _rot(*ang)
        pos = np.array([c * spacing, 0.0, 0.0])
        calib[c]["K"] = K.ravel()
        calib[c]["R"] = R.ravel()
        calib[c]["t"] = -R @ pos
    return calib


def _smooth(a):
    k = np.array([1, 4, 6, 4, 1], np.float32) / 16
    for ax in (0, 1):
        p = np.pad(a, [(2, 2) if i == ax else (0, 0) for i in (0, 1)], mode="wrap")
        a = sum(k[i] * np.take(p, np.arange(i, i + a.shape[ax]), axis=ax) for i in range(5))
    return a


def _texture(rng, n=256):
    t = _smooth(rng.uniform(0, 1, (n, n)).astype(np.float32))
    t = (t - t.mean()) / (t.std() + 1e-6) * 38.0 + float(rng.uniform(90, 170))
    for _ in range(40):  # blocky patches give strong, repeatable corners
        x, y = rng.integers(0, n - 8, 2)
        ww, hh = rng.integers(3, 24, 2)
        t[y:y + hh, x:x + ww] += float(rng.uniform(-70, 70))
    return t.astype(np.float32)


def _sample(tex, u, v):
    n = tex.shape[0]
    u0 = np.floor(u)
    v0 = np.floor(v)
    fu = (u - u0).astype(np.float32)
    fv = (v - v0).astype(np.float32)
    iu = u0.astype(np.int64) % n
    iv = v0.astype(np.int64) % n
    iu1 = (iu + 1) % n
    iv1 = (iv + 1) % n
    return (tex[iv, iu] * (1 - fu) * (1 - fv) + tex[iv, iu1] * fu * (1 - fv) + tex[iv1, iu] * (1 - fu) * fv + tex[iv1, iu1] * fu * fv)


def make_rig_frame(calib, w=720, h=540, frame_index=0, n_rects=14, noise_sigma=2.0):
    """Returns uint8 images [num_cams, h, w] for one rig frame."""
    seed = 1234 + frame_index
    rng = np.random.default_rng(seed)
    nc = len(calib)
    f = calib[0]["K"][0]
    planes = [(float(rng.uniform(18, 30)), None, _texture(rng))]  # (depth, rect or None, texture)
    for _ in range(n_rects):
        z = float(rng.uniform(1.5, 15.0))
        half_w = z * w / f / 2
        half_h = z * h / f / 2
        cx = float(rng.uniform(-half_w, half_w + 0.7))
        cy = float(rng.uniform(-half_h, half_h))
        sx = float(rng.uniform(0.15, 0.6)) * half_w
        sy = float(rng.uniform(0.15, 0.6)) * half_h
        planes.append((z, (cx - sx, cx + sx, cy - sy, cy + sy), _texture(rng)))
    planes.sort(key=lambda p: -p[0])  # far to near: nearer overwrite
    uu, vv = np.meshgrid(np.arange(w, dtype=np.float64), np.arange(h, dtype=np.float64))
    out = np.zeros((nc, h, w), np.uint8)
    for c in range(nc):
        K = calib[c]["K"].reshape(3, 3)
        R = calib[c]["R"].reshape(3, 3)
        pos = -R.T @ calib[c]["t"]
        d_cam = np.stack([(uu - K[0, 2]) / K[0, 0], (vv - K[1, 2]) / K[1, 1], np.ones_like(uu)], -1)
        d = d_cam @ R  # rows: R^T d_cam
        img = np.zeros((h, w), np.float32)
        for z, rect, tex in planes:
            lam = (z - pos[2]) / d[..., 2]
            X = pos[0] + lam * d[..., 0]
            Y = pos[1] + lam * d[..., 1]
            texel = z / f * 2.5
            val = _sample(tex, X / texel, Y / texel)
            if rect is None:
                img = val
            else:
                m = (X >= rect[0]) & (X <= rect[1]) & (Y >= rect[2]) & (Y <= rect[3])
                img = np.where(m, val, img)
        nrng = np.random.default_rng(seed * 1000 + c)
        img = img + nrng.normal(0, noise_sigma, (h, w)).astype(np.float32)
        out[c] = np.clip(np.rint(img), 0, 255).astype(np.uint8)
    return out


def make_masks(num_cams, w, h, frame_index=0):
    """Config-3 semantic masks: float32 [num_cams, h, w], 1.0 inside 3-6 random rectangles/ellipses, else 0.0
    (CV_32FC1, keep-rule mask < 0.7: FrontEnd.cpp:266)."""
    rng = np.random.default_rng(4321 + frame_index)
    yy, xx = np.mgrid[0:h, 0:w]
    masks = np.zeros((num_cams, h, w), np.float32)
    for c in range(num_cams):
        for _ in range(int(rng.integers(3, 7))):
            cx, cy = rng.uniform(0, w), rng.uniform(0, h)
            sx, sy = rng.uniform(0.06, 0.16) * w, rng.uniform(0.06, 0.16) * h
            if rng.integers(0, 2):
                m = (np.abs(xx - cx) < sx) & (np.abs(yy - cy) < sy)
            else:
                m = ((xx - cx) / sx) ** 2 + ((yy - cy) / sy) ** 2 < 1.0
            masks[c][m] = 1.0
    return masks


def noise_image(w, h, seed):
    """Pure-noise worst case for FAST candidate counts."""
    return np.random.default_rng(seed).integers(0, 256, (h, w), dtype=np.uint8)
