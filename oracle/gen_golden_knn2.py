"""Generates tests/golden/knn2_cv2.npz: cv2.BFMatcher(NORM_HAMMING).knnMatch(q, t, k=2) (OpenCV 4.13.0 here) on descriptor sets with
many distance ties (duplicated and near-duplicated rows), the primitive behind FrontEnd::findMatchesMono / findInterMatches
(FrontEnd.cpp:3134, :3200).  Run from the repo root:  python oracle/gen_golden_knn2.py"""
import os

import cv2
import numpy as np

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")
rng = np.random.default_rng(11)
t = rng.integers(0, 256, (700, 32), dtype=np.uint8)
t[100:140] = t[0:40]                      # exact duplicates: ties on (distance 0)
t[300:340, :4] = t[200:240, :4]
q = t[rng.integers(0, 700, 500)].copy()
flip = rng.integers(0, 256, (500, 32), dtype=np.uint8) & rng.integers(0, 256, (500, 32), dtype=np.uint8) & rng.integers(0, 256, (500, 32), dtype=np.uint8)
q[100:] ^= flip[100:]                     # first 100 queries are exact copies of train rows
knn = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(q, t, k=2)
out = np.array([[m[0].trainIdx, int(m[0].distance), m[1].trainIdx, int(m[1].distance)] for m in knn], np.int32)
np.savez_compressed(os.path.join(OUT, "knn2_cv2.npz"), q=q, t=t, knn=out)
print("cv2", cv2.__version__, out.shape, "ties:", int((out[:, 1] == out[:, 3]).sum()))


# ---- inter-frame host logic on top of cv2's knnMatch: a second, independent restatement (python, cv2.DMatch / cv2.norm) of
# FrontEnd::findMatchesMono (:3114-3154) and both FrontEnd::findInterMatches overloads (:3157-3343, :3344-3470).  The reference
# itself cannot run here (needs GTSAM / DBoW2 / ROS), so these fixtures pin oracle/mcs_oracle.cpp against cv2 + this restatement.
def py_find_matches_mono(d1, d2, xy1, xy2, m1, m2):
    out = []
    for m in cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(d1, d2, k=2):
        if float(m[0].distance) < 0.7 * float(m[1].distance):
            if m[0].distance > 50:
                continue
            p, c = xy1[m[0].queryIdx], xy2[m[0].trainIdx]
            if float(m1[int(p[1]), int(p[0])]) < 0.7 and float(m2[int(c[1]), int(c[0])]) < 0.7:  # float -> double like C++ (numpy would compare in f32)
                out.append((m[0].queryIdx, m[0].trainIdx, int(m[0].distance)))
    return np.array(out, np.int32).reshape(-1, 3)


def py_find_inter_matches(d1, d2, mono1, mono2, p1, p2, lids1, isz1, isz2, split):
    inds1, inds2, lf, mi1, mi2, mo = [], [], [], [], [], []
    for m in cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(d1, d2, k=2):
        q, t = m[0].queryIdx, m[0].trainIdx
        if lids1[q] == -1 and float(m[0].distance) > 0.7 * float(m[1].distance):
            continue
        if not mono1[q] and not mono2[t]:
            if not (float(np.float32(cv2.norm(p1[q] - p2[t]))) <= 2.0):
                continue
        in_lf = (t in inds2) if (not split or t < isz2) else False
        in_mo = (t in mi2) if split else False
        rec = (q, t, int(m[0].distance))
        if not in_lf and not in_mo:
            if not split or (q < isz1 and t < isz2):
                inds2.append(t); inds1.append(q); lf.append(rec)
            else:
                mi2.append(t); mi1.append(q); mo.append(rec)
        elif not in_mo:
            k = inds2.index(t)
            if rec[2] < lf[k][2]:
                inds1[k] = q; lf[k] = rec
        else:
            k = mi2.index(t)
            if rec[2] < mo[k][2]:
                mi1[k] = q; mo[k] = rec
    return np.array(lf, np.int32).reshape(-1, 3), np.array(mo, np.int32).reshape(-1, 3)


rng = np.random.default_rng(12)
W, H = 720, 540
n1, n2 = 600, 640
d2 = rng.integers(0, 256, (n2, 32), dtype=np.uint8)
src = rng.integers(0, n2, n1)
noise = np.zeros((n1, 32), np.uint8)
for k in range(n1):                        # 0..70 flipped bits: both sides of the distance-50 gate
    bits = rng.choice(256, int(rng.integers(0, 70)), replace=False)
    np.bitwise_or.at(noise[k], bits // 8, (1 << (bits % 8)).astype(np.uint8))
d1 = d2[src] ^ noise
d1[500:] = rng.integers(0, 256, (100, 32), dtype=np.uint8)
xy1 = (rng.random((n1, 2)) * [W - 1, H - 1]).astype(np.float32)
xy2 = (rng.random((n2, 2)) * [W - 1, H - 1]).astype(np.float32)
blk = lambda: np.kron(np.round(rng.random((H // 12, W // 12)) * 0.9, 1), np.ones((12, 12))).astype(np.float32)  # 12x12-pixel blocks: the fixture stays small
m1, m2 = blk(), blk()
mono_out = py_find_matches_mono(d1, d2, xy1, xy2, m1, m2)

isz1, isz2 = 350, 380                      # rows >= isz are mono features appended after the triangulated intra-matches
mono1 = (np.arange(n1) >= isz1).astype(np.uint8)
mono2 = (np.arange(n2) >= isz2).astype(np.uint8)
p2 = rng.normal(0, 5, (n2, 3)) + [0, 0, 12]
p1 = p2[src] + rng.normal(0, 0.9, (n1, 3))  # ~half inside the 2 m gate
lids1 = np.where(rng.random(n1) < 0.3, rng.integers(0, 1000, n1), -1).astype(np.int32)
lf_a, mo_a = py_find_inter_matches(d1, d2, mono1, mono2, p1, p2, lids1, isz1, isz2, True)
lf_b, _ = py_find_inter_matches(d1, d2, mono1, mono2, p1, p2, lids1, isz1, isz2, False)
np.savez_compressed(os.path.join(OUT, "inter_frame_cv2.npz"), d1=d1, d2=d2, xy1=xy1, xy2=xy2, m1=m1, m2=m2, mono_out=mono_out,
                    mono1=mono1, mono2=mono2, p1=p1, p2=p2, lids1=lids1, isz=np.array([isz1, isz2]), lf_split=lf_a, mono_split=mo_a,
                    lf_single=lf_b)
print("findMatchesMono", len(mono_out), "findInterMatches split", len(lf_a), len(mo_a), "single", len(lf_b))


# ---- BoW-guided subset matcher: ORBextractor::getMatches_distRatio (ORBextractor.cpp:1228-1290) + the node walk of BowMatching /
# InterMatchingBow (MultiCameraFrame.cpp:946-1022, FrontEnd.cpp:3791-3965), restated in python over numpy popcounts.  The DBoW2
# vocabulary is absent, so node ids are synthetic: matching features mostly share a node, with some scattered to other nodes.
def py_dist(a, b):
    return int(np.unpackbits(a ^ b).sum())


def py_matches_dist_ratio(A, iA, B, iB, th_low=75, ratio=0.85):
    ma, mb, book = [], [], 0
    for ia in iA:
        best_j, b1, b2 = -1, 1e9, 1e9
        for j, ib in enumerate(iB):
            d = float(py_dist(A[ia], B[ib])); book += 1
            if d < b1:
                best_j, b2, b1 = j, b1, d
            elif d < b2:
                b2 = d
        if b1 <= th_low and b1 / b2 <= ratio:
            idx_b = iB[best_j]
            if idx_b not in mb:
                mb.append(idx_b); ma.append(ia)
            else:
                k = mb.index(idx_b)
                d = float(py_dist(A[ma[k]], B[idx_b])); book += 1
                if b1 < d:
                    ma[k] = ia
    return ma, mb, book


rng = np.random.default_rng(13)
nB = 900
B = rng.integers(0, 256, (nB, 32), dtype=np.uint8)
srcA = rng.permutation(nB)[:800]
A = B[srcA].copy()
for k in range(len(A)):
    bits = rng.choice(256, int(rng.integers(0, 90)), replace=False)
    np.bitwise_xor.at(A[k], bits // 8, (1 << (bits % 8)).astype(np.uint8))
A[700:] = A[rng.integers(0, 700, 100)]    # repeated A features: several queries claim the same B feature
node_of_B = rng.integers(0, 60, nB) * 7 + 3
node_of_A = node_of_B[srcA].copy()
node_of_A[700:] = node_of_B[srcA[:700]][rng.integers(0, 700, 100)]
scatter = rng.random(800) < 0.15
node_of_A[scatter] = rng.integers(0, 70, int(scatter.sum())) * 7 + 3   # includes node ids that do not exist in B
fvA, fvB = {}, {}
for i, nd in enumerate(node_of_A):
    fvA.setdefault(int(nd), []).append(i)
for i, nd in enumerate(node_of_B):
    fvB.setdefault(int(nd), []).append(i)
i1, i2, words, book = [], [], [], 0
for nd in sorted(set(fvA) & set(fvB)):     # the lower_bound walk visits exactly the common node ids in ascending order
    a, b, k = py_matches_dist_ratio(A, fvA[nd], B, fvB[nd])
    book += k
    i1 += a; i2 += b
    if a:
        words.append(nd)
np.savez_compressed(os.path.join(OUT, "bow_subsets.npz"), A=A, B=B, node_of_A=node_of_A.astype(np.int32), node_of_B=node_of_B.astype(np.int32),
                    indices_1=np.array(i1, np.uint32), indices_2=np.array(i2, np.uint32), words=np.array(words, np.uint32),
                    book_k=np.array([book], np.int64))
print("bow subsets: nodes", len(fvA), len(fvB), "matches", len(i1), "words", len(words), "BookK", book)


# ---- setData input conversion through cv2 (MultiCameraFrame.cpp:107-116): multiply(img, 255); convertTo(CV_8U); cvtColor(BGR2GRAY)
rng = np.random.default_rng(14)
gray = rng.random((48, 64), dtype=np.float32)
gray.ravel()[:16] = [0, 1, 0.5, 0.0019607844, 0.005882353, 1.2, -0.3, 0.998, 0.5019608, 0.49803922, 2.5 / 255, 3.5 / 255, 254.5 / 255, 255.5 / 255,
                     1e-8, 0.9999999]
bgr = rng.random((40, 52, 3), dtype=np.float32)
bgr.reshape(-1)[:6] = [0.5, 1.5 / 255, 2.5 / 255, 1.0, 1.0, 1.0]


def cv_set_data(img):
    m = cv2.multiply(img, 255.0)
    u8 = cv2.add(m, np.zeros_like(m), dtype=cv2.CV_8U)       # saturate_cast<uchar>(cvRound(.)) like Mat::convertTo
    return cv2.cvtColor(u8, cv2.COLOR_BGR2GRAY) if u8.ndim == 3 else u8


np.savez_compressed(os.path.join(OUT, "set_data_f32.npz"), gray=gray, gray_u8=cv_set_data(gray), bgr=bgr, bgr_u8=cv_set_data(bgr))
print("setData fixture", cv_set_data(gray).ravel()[:16])


# ---- cv2.undistort (RECTIFY branch of setData, MultiCameraFrame.cpp:123-136) with two of the shipped calibrations' coefficient
# sets scaled to a 320 x 240 image: strong barrel (nuance cam0) and a pincushion-like set that samples outside the source image.
rng = np.random.default_rng(15)
uw, uh = 320, 240
uimg = cv2.GaussianBlur(rng.integers(0, 256, (uh, uw), dtype=np.uint8), (0, 0), 1.2)
umask = np.kron(np.round(rng.random((uh // 8, uw // 8)), 1), np.ones((8, 8))).astype(np.float32)
K1 = np.array([[203.8, 0, 163.2], [0, 203.2, 110.4], [0, 0, 1]], np.float64)
d1 = np.array([-0.376399267634092, 0.2684942610034098, 0.001628309011323219, 0.0014062772667959773, 0.0])
K2 = np.array([[311.2, 0, 157.8], [0, 311.0, 120.4], [0, 0, 1]], np.float64)
d2 = np.array([0.12, 0.05, -0.0023147852033182366, 0.001098765060765587, -0.01])
np.savez_compressed(os.path.join(OUT, "undistort_cv2.npz"), img=uimg, mask=umask, K1=K1, d1=d1, K2=K2, d2=d2,
                    img_u1=cv2.undistort(uimg, K1, d1), img_u2=cv2.undistort(uimg, K2, d2),
                    mask_u1=cv2.undistort(umask, K1, d1), mask_u2=cv2.undistort(umask, K2, d2))
print("undistort fixture written")


# ---- cv2.undistortPoints as UndistortKeyPoints calls it (MultiCameraFrame.cpp:318-334): K and dist narrowed to f32, P = K
rng = np.random.default_rng(16)
upts = (rng.random((1500, 2)) * [uw, uh]).astype(np.float32)
und_pts = {}
for name, Kx, dx in (("1", K1, d1), ("2", K2, d2)):
    K32, d32 = Kx.astype(np.float32), dx.astype(np.float32).reshape(1, -1)
    und_pts[name] = cv2.undistortPoints(upts.reshape(-1, 1, 2).copy(), K32, d32, None, K32).reshape(-1, 2)
np.savez_compressed(os.path.join(OUT, "undistort_points_cv2.npz"), pts=upts, K1=K1, d1=d1, K2=K2, d2=d2, out1=und_pts["1"], out2=und_pts["2"])
print("undistortPoints fixture written")
