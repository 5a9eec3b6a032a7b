"""Tier-1 oracle: a Python transliteration of the reference's ORBextractor / BruteForceMatch that calls the SAME
OpenCV primitives the reference calls, through cv2 (4.13.0 in this container; the reference pins 4.5.5).

TEST INFRASTRUCTURE ONLY.  Runs only where cv2 is importable; its outputs are committed as tests/golden/*.npz by
oracle/gen_golden.py, and the portable C++ oracle (oracle/mcs_oracle.cpp) is pinned against those.

Follows /root/reference/MCSlam/src/ORBextractor.cpp :408-468, :479-535, :537-552, :554-778, :780-895, :106-145,
:1085-1198 and MCSlam/src/MultiCameraFrame.cpp :1024-1086.
"""
import math

import cv2
import numpy as np

import oracle_lib as O

f32 = np.float32
EDGE_THRESHOLD, PATCH_SIZE = 19, 31


def _pattern():
    import re, os
    txt = open(os.path.join(os.path.dirname(__file__), "..", "include", "mcs_orb_pattern.inc")).read()
    body = txt[txt.index("{") + 1: txt.rindex("}")]
    v = np.array([int(x) for x in re.findall(r"-?\d+", body)], np.int32)
    assert v.size == 1024
    return v.reshape(512, 2)


PATTERN = _pattern()


class Node:
    __slots__ = ("keys", "ULx", "ULy", "BRx", "BRy", "noMore", "alive")

    def __init__(self):
        self.keys, self.noMore, self.alive = [], False, True


def divide_node(n, cx, cy):
    halfX = int(math.ceil(float(f32(n.BRx - n.ULx) / f32(2))))
    halfY = int(math.ceil(float(f32(n.BRy - n.ULy) / f32(2))))
    ch = [Node() for _ in range(4)]
    n1, n2, n3, n4 = ch
    n1.ULx, n1.ULy, n1.BRx, n1.BRy = n.ULx, n.ULy, n.ULx + halfX, n.ULy + halfY
    n2.ULx, n2.ULy, n2.BRx, n2.BRy = n1.BRx, n.ULy, n.BRx, n.ULy + halfY
    n3.ULx, n3.ULy, n3.BRx, n3.BRy = n.ULx, n1.BRy, n1.BRx, n.BRy
    n4.ULx, n4.ULy, n4.BRx, n4.BRy = n1.BRx, n1.BRy, n.BRx, n.BRy
    for k in n.keys:
        if cx[k] < n1.BRx:
            (n1 if cy[k] < n1.BRy else n3).keys.append(k)
        elif cy[k] < n1.BRy:
            n2.keys.append(k)
        else:
            n4.keys.append(k)
    for c in ch:
        if len(c.keys) == 1:
            c.noMore = True
    return ch


def distribute_octtree(cx, cy, resp, minX, maxX, minY, maxY, N):
    """std::list emulated by a Python list (front = index 0)."""
    nIni = int(round(float(f32(maxX - minX) / f32(maxY - minY))))  # C round(): half away from zero; values here are not .5
    nIni = max(nIni, 1)
    hX = f32(maxX - minX) / f32(nIni)
    nodes = []
    for i in range(nIni):
        n = Node()
        n.ULx, n.BRx, n.ULy, n.BRy = int(hX * f32(i)), int(hX * f32(i + 1)), 0, maxY - minY
        nodes.append(n)
    for k in range(len(cx)):
        nodes[min(int(f32(cx[k]) / hX), nIni - 1)].keys.append(k)
    lst = []
    for n in nodes:
        if len(n.keys) == 1:
            n.noMore = True
        if n.keys:
            lst.append(n)
    finish = False
    while not finish:
        prevSize = len(lst)
        nToExpand = 0
        sizeAndNode = []
        new_front = []  # children pushed to the front, in push order
        rest = []
        for n in lst:
            if n.noMore:
                rest.append(n)
                continue
            for c in divide_node(n, cx, cy):
                if c.keys:
                    new_front.append(c)
                    if len(c.keys) > 1:
                        nToExpand += 1
                        sizeAndNode.append(c)
        lst = new_front[::-1] + rest
        if len(lst) >= N or len(lst) == prevSize:
            finish = True
        elif len(lst) + nToExpand * 3 > N:
            while not finish:
                prevSize = len(lst)
                prev = sizeAndNode
                sizeAndNode = []
                perm = O.std_sort_nodes([len(p.keys) for p in prev], [p.ULx for p in prev])
                prev = [prev[i] for i in perm]
                for j in range(len(prev) - 1, -1, -1):
                    for c in divide_node(prev[j], cx, cy):
                        if c.keys:
                            lst.insert(0, c)
                            if len(c.keys) > 1:
                                sizeAndNode.append(c)
                    lst.remove(prev[j])  # identity-based: Node has no __eq__
                    if len(lst) >= N:
                        break
                if len(lst) >= N or len(lst) == prevSize:
                    finish = True
    keep = []
    for n in lst:
        best = n.keys[0]
        for k in n.keys[1:]:
            if resp[k] > resp[best]:
                best = k
        keep.append(best)
    return keep


class ORBextractorCv2:
    def __init__(self, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7):
        self.nfeatures, self.nlevels, self.iniTh, self.minTh = nfeatures, nlevels, iniThFAST, minThFAST
        sf = float(f32(scaleFactor))  # double member initialised from a float argument
        self.scale = np.ones(nlevels, f32)
        for i in range(1, nlevels):
            self.scale[i] = f32(float(self.scale[i - 1]) * sf)
        self.sigma2 = self.scale * self.scale
        self.sigma2[0] = 1.0
        self.inv_scale = f32(1.0) / self.scale
        self.inv_sigma2 = f32(1.0) / self.sigma2
        factor = f32(1.0 / sf)
        nd = f32(f32(nfeatures) * (f32(1) - factor)) / (f32(1) - f32(math.pow(float(factor), float(nlevels))))
        self.quota = np.zeros(nlevels, np.int32)
        s = 0
        for l in range(nlevels - 1):
            self.quota[l] = int(np.rint(nd))
            s += self.quota[l]
            nd = f32(nd * factor)
        self.quota[nlevels - 1] = max(nfeatures - s, 0)
        self.fast_ini = cv2.FastFeatureDetector_create(iniThFAST, True)
        self.fast_min = cv2.FastFeatureDetector_create(minThFAST, True)

    def compute_pyramid(self, image):
        self.pyr, self.pyr_padded = [], []
        h, w = image.shape
        for l in range(self.nlevels):
            s = self.inv_scale[l]
            sz = (int(np.rint(f32(w) * s)), int(np.rint(f32(h) * s)))
            if l == 0:
                lvl = image.copy()
            else:
                lvl = cv2.resize(self.pyr[l - 1], sz, interpolation=cv2.INTER_LINEAR)
            self.pyr.append(lvl)
            self.pyr_padded.append(cv2.copyMakeBorder(lvl, EDGE_THRESHOLD, EDGE_THRESHOLD, EDGE_THRESHOLD, EDGE_THRESHOLD,
                                                      cv2.BORDER_REFLECT_101))

    def compute_keypoints(self):
        self.cands, self.level_kps = [], []
        W = f32(35)
        for level in range(self.nlevels):
            img = self.pyr[level]
            rows, cols = img.shape
            minBX = minBY = EDGE_THRESHOLD - 3
            maxBX, maxBY = cols - EDGE_THRESHOLD + 3, rows - EDGE_THRESHOLD + 3
            width, height = f32(maxBX - minBX), f32(maxBY - minBY)
            nCols, nRows = int(width / W), int(height / W)
            cx, cy, cr = [], [], []
            if nCols >= 1 and nRows >= 1:
                wCell, hCell = int(math.ceil(float(width / f32(nCols)))), int(math.ceil(float(height / f32(nRows))))
                for i in range(nRows):
                    iniY = minBY + i * hCell
                    maxY = iniY + hCell + 6
                    if iniY >= maxBY - 3:
                        continue
                    maxY = min(maxY, maxBY)
                    for j in range(nCols):
                        iniX = minBX + j * wCell
                        maxX = iniX + wCell + 6
                        if iniX >= maxBX - 6:
                            continue
                        maxX = min(maxX, maxBX)
                        cell = img[iniY:maxY, iniX:maxX]
                        kps = self.fast_ini.detect(cell)
                        if len(kps) == 0:
                            kps = self.fast_min.detect(cell)
                        for k in kps:
                            cx.append(f32(k.pt[0]) + f32(j * wCell))
                            cy.append(f32(k.pt[1]) + f32(i * hCell))
                            cr.append(f32(k.response))
            self.cands.append(np.array([cx, cy, cr], f32).T.reshape(-1, 3))
            keep = distribute_octtree(cx, cy, cr, minBX, maxBX, minBY, maxBY, int(self.quota[level])) if cx else []
            sps = int(f32(PATCH_SIZE) * self.scale[level])
            kp = np.zeros(len(keep), O.KP_DTYPE)
            for n, k in enumerate(keep):
                kp[n] = (cx[k] + f32(minBX), cy[k] + f32(minBY), sps, 0.0, cr[k], level, -1)
            self.level_kps.append(kp)

    def __call__(self, image, lapping=(0, 0)):
        self.compute_pyramid(image)
        self.compute_keypoints()
        n = sum(len(k) for k in self.level_kps)
        kps = np.zeros(n, O.KP_DTYPE)
        desc = np.zeros((n, 32), np.uint8)
        mono, stereo = 0, n - 1
        self.blurred = [None] * self.nlevels
        for level in range(self.nlevels):
            lk = self.level_kps[level]
            if len(lk) == 0:
                continue
            work = cv2.GaussianBlur(self.pyr[level].copy(), (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
            self.blurred[level] = work
            s = self.scale[level]
            for kp in lk:
                cxi, cyi = int(np.rint(kp["x"])), int(np.rint(kp["y"]))
                vals = work[cyi + PATTERN[:, 1], cxi + PATTERN[:, 0]].astype(np.int32)  # angle 0: un-steered
                bits = (vals[0::2] < vals[1::2]).astype(np.uint8)
                d = np.packbits(bits.reshape(32, 8), axis=1, bitorder="little").ravel()
                kp = kp.copy()
                if level != 0:
                    kp["x"] = f32(kp["x"] * s)
                    kp["y"] = f32(kp["y"] * s)
                if lapping[0] <= kp["x"] <= lapping[1]:
                    slot = stereo
                    stereo -= 1
                else:
                    slot = mono
                    mono += 1
                kps[slot] = kp
                desc[slot] = d
        return mono, kps, desc


def brute_force_match(d1, d2, dist_thresh=75.0, ratio=0.85):
    """MultiCameraFrame.cpp:1024-1086 with cv2.BFMatcher(NORM_HAMMING).knnMatch(k=2)."""
    m = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(d1, d2, k=2)
    i1, i2 = [], []
    for pair in m:
        if len(pair) < 2:
            continue
        if f32(pair[0].distance) < f32(ratio) * f32(pair[1].distance):
            if pair[0].distance > dist_thresh:
                continue
            i1.append(pair[0].queryIdx)
            i2.append(pair[0].trainIdx)
    return np.array(i1, np.uint32), np.array(i2, np.uint32)


def compute_intra_matches(descs, dist_thresh=75.0, ratio=0.85):
    """computeIntraMatches(matches, old=false) merge (MultiCameraFrame.cpp:1100-1288) over cv2 BFMatcher pair lists."""
    nc = len(descs)
    inv = [np.full(len(d), -1, np.int64) for d in descs]
    rows = []
    pair_lists = []
    for i in range(nc - 1):
        for j in range(i + 1, nc):
            i1, i2 = brute_force_match(descs[i], descs[j], dist_thresh, ratio) if len(descs[i]) and len(descs[j]) >= 2 else ([], [])
            pair_lists.append((np.array(i1, np.uint32), np.array(i2, np.uint32)))
            for fi, fj in zip(i1, i2):
                fi, fj = int(fi), int(fj)
                mi, mj = inv[i][fi], inv[j][fj]
                if mi == -1 and mj == -1:
                    r = [-1] * nc
                    r[i], r[j] = fi, fj
                    inv[i][fi] = inv[j][fj] = len(rows)
                    rows.append(r)
                else:
                    if mi == -1 and mj != -1:
                        if rows[mj][i] == -1:
                            rows[mj][i] = fi
                            inv[i][fi] = mj
                    if mi != -1 and mj == -1:
                        rows[mi][j] = fj
                        inv[j][fj] = mi
    return np.array(rows, np.int32).reshape(-1, nc), pair_lists


def triangulate_numpy(calib, kps, rows):
    """FrontEnd.cpp:249-370 triangulation with numpy.linalg.svd as the solveZ (independent of the oracle's Jacobi)."""
    nc = len(kps)
    out = np.zeros((len(rows), 3))
    for r, row in enumerate(rows):
        views = [c for c in range(nc) if row[c] != -1]
        if len(views) < 2:
            continue
        Ps, xs = [], []
        for c in views:
            K = calib[c]["K"].reshape(3, 3)
            Rt = np.hstack([calib[c]["R"].reshape(3, 3), calib[c]["t"].reshape(3, 1)])
            p = kps[c][row[c]]
            xs.append(((float(p["x"]) - K[0, 2]) / K[0, 0], (float(p["y"]) - K[1, 2]) / K[1, 1]))
            Ps.append(Rt)
        n = len(views)
        if n == 2:
            A = np.zeros((4, 4))
            for v in range(2):
                A[2 * v] = xs[v][0] * Ps[v][2] - Ps[v][0]
                A[2 * v + 1] = xs[v][1] * Ps[v][2] - Ps[v][1]
        else:
            A = np.zeros((3 * n, 4 + n))
            for v in range(n):
                A[3 * v:3 * v + 3, :4] = -Ps[v]
                A[3 * v, 4 + v], A[3 * v + 1, 4 + v], A[3 * v + 2, 4 + v] = xs[v][0], xs[v][1], 1.0
        z = np.linalg.svd(A)[2][-1]
        out[r] = z[:3] / z[3]
    return out
