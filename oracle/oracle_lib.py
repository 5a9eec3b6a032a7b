"""ctypes binding of oracle/liboracle.so (the CPU oracle — TEST INFRASTRUCTURE ONLY).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
CALIB_DTYPE = np.dtype([("K", "<f8", (9,)), ("R", "<f8", (9,)), ("t", "<f8", (3,)), ("dist", "<f8", (5,))])
assert KP_DTYPE.itemsize == 28 and CALIB_DTYPE.itemsize == 26 * 8


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    src = os.path.join(_HERE, "mcs_oracle.cpp")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        L = _LIB
        L.orc_extractor_create.restype = C.c_void_p
        L.orc_extractor_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.orc_extractor_destroy.argtypes = [C.c_void_p]
        L.orc_extractor_set_orientation.argtypes = [C.c_void_p, C.c_int]
        L.orc_extractor_tables.argtypes = [C.c_void_p] + [C.c_void_p] * 6
        L.orc_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
        L.orc_extractor_last_total.argtypes = [C.c_void_p]
        L.orc_level_size.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_get_level.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.orc_get_level_padded.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.orc_num_candidates.argtypes = [C.c_void_p, C.c_int]
        L.orc_get_candidates.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.orc_num_level_kps.argtypes = [C.c_void_p, C.c_int]
        L.orc_resize_linear_u8.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.orc_gauss7_u8.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.orc_gauss_kernel_fixed.argtypes = [C.c_int, C.c_double, C.c_void_p]
        L.orc_fast9.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.orc_distribute_octtree.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.orc_std_sort_nodes.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.orc_fast_atan2.restype = C.c_float
        L.orc_fast_atan2.argtypes = [C.c_float, C.c_float]
        L.orc_hamming256.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_bf_match.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_float, C.c_float] + [C.c_void_p] * 5
        L.orc_knn2.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int] + [C.c_void_p] * 4
        L.orc_find_matches_mono.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int] + [C.c_void_p] * 4 + [C.c_int, C.c_void_p]
        L.orc_find_inter_matches.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int] + [C.c_void_p] * 5 + [C.c_int] * 3 + [C.c_void_p] * 3
        L.orc_matches_dist_ratio.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_double] + [C.c_void_p] * 3
        L.orc_bow_matching.argtypes = [C.c_void_p] * 5 + [C.c_int] + [C.c_void_p] * 3 + [C.c_int, C.c_int, C.c_double] + [C.c_void_p] * 5
        L.orc_argsorte_desc.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.orc_undistort_u8.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_undistort_f32.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_undistort_keypoints.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_vocab_transform.argtypes = [C.c_int, C.c_int] + [C.c_void_p] * 5 + [C.c_int, C.c_int] + [C.c_void_p] * 3
        L.orc_intra_matches_bow.argtypes = [C.c_int] + [C.c_void_p] * 6 + [C.c_int, C.c_double, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_set_data_f32.argtypes = [C.c_void_p, C.c_int, C.c_size_t, C.c_void_p]
        L.orc_intra_matches.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_float,
                                        C.c_float, C.c_void_p, C.c_int, C.c_void_p]
        L.orc_triangulate_pixels.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.orc_triangulate_pair_f32.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.orc_triangulate_rows.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p,
                                           C.c_int] + [C.c_void_p] * 6
        L.orc_process_rig.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                      C.c_void_p, C.c_void_p]
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def resize_linear(src, dw, dh):
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.empty((dh, dw), np.uint8)
    lib().orc_resize_linear_u8(_p(src), src.shape[1], src.shape[0], src.shape[1], _p(dst), dw, dh, dw)
    return dst


def gauss7(src):
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.empty_like(src)
    lib().orc_gauss7_u8(_p(src), src.shape[1], src.shape[0], src.shape[1], _p(dst), src.shape[1])
    return dst


def gauss_kernel_fixed(n=7, sigma=2.0):
    out = np.zeros(n, np.int32)
    lib().orc_gauss_kernel_fixed(n, sigma, _p(out))
    return out


def fast9(img, threshold):
    """cv::FAST(img, th, nms=True) on a (possibly strided) 2-D uint8 view; returns int32 [n,3] (x, y, score)."""
    assert img.dtype == np.uint8 and img.strides[1] == 1
    h, w = img.shape
    cap = (w * h) // 4 + 16
    out = np.empty((cap, 3), np.int32)
    n = lib().orc_fast9(C.c_void_p(img.ctypes.data), w, h, img.strides[0], threshold, _p(out), cap)
    return out[:n].copy()


class Extractor:
    """Mirror of ORBextractor (MCSlam/include/MCSlam/ORBextractor.h:44-112)."""

    def __init__(self, nfeatures=1000, scale=1.2, nlevels=8, ini_th=20, min_th=7):
        self.nfeatures, self.nlevels = nfeatures, nlevels
        self.h = lib().orc_extractor_create(nfeatures, scale, nlevels, ini_th, min_th)
        self.scale = np.zeros(nlevels, np.float32)
        self.inv_scale = np.zeros(nlevels, np.float32)
        self.sigma2 = np.zeros(nlevels, np.float32)
        self.inv_sigma2 = np.zeros(nlevels, np.float32)
        self.quota = np.zeros(nlevels, np.int32)
        self.umax = np.zeros(16, np.int32)
        lib().orc_extractor_tables(self.h, _p(self.scale), _p(self.inv_scale), _p(self.sigma2), _p(self.inv_sigma2),
                                   _p(self.quota), _p(self.umax))

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_extractor_destroy(self.h)
            self.h = None

    def set_orientation(self, on):
        lib().orc_extractor_set_orientation(self.h, int(on))

    def __call__(self, image, lapping=(0, 0)):
        """Returns (monoIndex, keypoints[KP_DTYPE], descriptors[n,32])."""
        image = np.ascontiguousarray(image, np.uint8)
        h, w = image.shape
        cap = self.nfeatures + 64
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        mono = lib().orc_extract(self.h, _p(image), w, h, w, lapping[0], lapping[1], _p(kps), _p(desc), cap)
        if mono < 0:
            return mono, kps[:0], desc[:0]
        n = lib().orc_extractor_last_total(self.h)
        return mono, kps[:n].copy(), desc[:n].copy()

    def level(self, l, blurred=False):
        w, h = C.c_int(), C.c_int()
        lib().orc_level_size(self.h, l, C.byref(w), C.byref(h))
        out = np.zeros((h.value, w.value), np.uint8)
        n = lib().orc_get_level(self.h, l, int(blurred), _p(out))
        return out if n else None

    def level_padded(self, l):
        w, h = C.c_int(), C.c_int()
        lib().orc_level_size(self.h, l, C.byref(w), C.byref(h))
        out = np.zeros((h.value + 38, w.value + 38), np.uint8)
        lib().orc_get_level_padded(self.h, l, _p(out))
        return out

    def candidates(self, l):
        n = lib().orc_num_candidates(self.h, l)
        out = np.zeros((max(n, 1), 3), np.float32)
        lib().orc_get_candidates(self.h, l, _p(out))
        return out[:n]

    def level_counts(self):
        return np.array([lib().orc_num_level_kps(self.h, l) for l in range(self.nlevels)], np.int32)


def distribute_octtree(xyr, minX, maxX, minY, maxY, N):
    xyr = np.ascontiguousarray(xyr, np.float32)
    keep = np.zeros(len(xyr) + 8, np.int32)
    n = lib().orc_distribute_octtree(_p(xyr), len(xyr), minX, maxX, minY, maxY, N, _p(keep))
    return keep[:n].copy()


def std_sort_nodes(sizes, ulx):
    """libstdc++ std::sort with compareNodes (ORBextractor.cpp:537-552); returns the permutation."""
    sizes = np.ascontiguousarray(sizes, np.int32)
    ulx = np.ascontiguousarray(ulx, np.int32)
    perm = np.zeros(len(sizes), np.int32)
    lib().orc_std_sort_nodes(_p(sizes), _p(ulx), len(sizes), _p(perm))
    return perm


def bf_match(q, t, dist_thresh=75.0, ratio=0.85):
    """BruteForceMatch (MultiCameraFrame.cpp:1024-1086). Returns idx1, idx2, best_d, second_d, best_i."""
    q = np.ascontiguousarray(q, np.uint8)
    t = np.ascontiguousarray(t, np.uint8)
    nq, nt = len(q), len(t)
    i1 = np.zeros(max(nq, 1), np.uint32)
    i2 = np.zeros(max(nq, 1), np.uint32)
    bd = np.zeros(max(nq, 1), np.int32)
    sd = np.zeros(max(nq, 1), np.int32)
    bi = np.zeros(max(nq, 1), np.int32)
    n = lib().orc_bf_match(_p(q), nq, _p(t), nt, dist_thresh, ratio, _p(i1), _p(i2), _p(bd), _p(sd), _p(bi))
    return i1[:n].copy(), i2[:n].copy(), bd[:nq], sd[:nq], bi[:nq]


def knn2(q, t):
    """knnMatch(q, t, k=2) (FrontEnd.cpp:3134, :3200): int32 [nq, 4] = (trainIdx0, dist0, trainIdx1, dist1), -1 where missing."""
    q = np.ascontiguousarray(q, np.uint8).reshape(-1, 32)
    t = np.ascontiguousarray(t, np.uint8).reshape(-1, 32)
    out = [np.zeros(max(len(q), 1), np.int32) for _ in range(4)]
    lib().orc_knn2(_p(q), len(q), _p(t), len(t), *[_p(o) for o in out])
    return np.stack([o[:len(q)] for o in out], 1)


def find_matches_mono(d1, d2, xy1, xy2, mask1=None, mask2=None, width=0):
    """FrontEnd::findMatchesMono (FrontEnd.cpp:3114-3154): int32 [n, 3] = (queryIdx, trainIdx, distance)."""
    d1 = np.ascontiguousarray(d1, np.uint8).reshape(-1, 32)
    d2 = np.ascontiguousarray(d2, np.uint8).reshape(-1, 32)
    xy1 = np.ascontiguousarray(xy1, np.float32)
    xy2 = np.ascontiguousarray(xy2, np.float32)
    m1 = None if mask1 is None else np.ascontiguousarray(mask1, np.float32)
    m2 = None if mask2 is None else np.ascontiguousarray(mask2, np.float32)
    out = np.zeros((max(len(d1), 1), 3), np.int32)
    n = lib().orc_find_matches_mono(_p(d1), len(d1), _p(d2), len(d2), _p(xy1), _p(xy2), _p(m1), _p(m2), width, _p(out))
    return out[:n]


def find_inter_matches(d1, d2, mono1, mono2, p1, p2, lids1, isz1, isz2, split_mono=True):
    """FrontEnd::findInterMatches (FrontEnd.cpp:3157-3343 / :3344-3470): (matches_z_filtered, matches_mono_z_filtered) as int32 [n, 3]."""
    d1 = np.ascontiguousarray(d1, np.uint8).reshape(-1, 32)
    d2 = np.ascontiguousarray(d2, np.uint8).reshape(-1, 32)
    mono1 = np.ascontiguousarray(mono1, np.uint8)
    mono2 = np.ascontiguousarray(mono2, np.uint8)
    p1 = np.ascontiguousarray(p1, np.float64)
    p2 = np.ascontiguousarray(p2, np.float64)
    lids1 = np.ascontiguousarray(lids1, np.int32)
    lf = np.zeros((max(len(d1), 1), 3), np.int32)
    mo = np.zeros((max(len(d1), 1), 3), np.int32)
    nm = C.c_int(0)
    n = lib().orc_find_inter_matches(_p(d1), len(d1), _p(d2), len(d2), _p(mono1), _p(mono2), _p(p1), _p(p2), _p(lids1), isz1, isz2,
                                     int(split_mono), _p(lf), _p(mo), C.byref(nm))
    return lf[:n], mo[:nm.value]


def feature_vector_csr(fv):
    """{node id: [feature indices]} (a DBoW2::FeatureVector) -> (ascending node ids u32, CSR offsets i32, indices u32)."""
    nodes = sorted(fv)
    ofs = np.zeros(len(nodes) + 1, np.int32)
    for k, nd in enumerate(nodes):
        ofs[k + 1] = ofs[k] + len(fv[nd])
    idx = np.array([i for nd in nodes for i in fv[nd]], np.uint32)
    return np.array(nodes, np.uint32), ofs, idx


def matches_dist_ratio(A, iA, B, iB, th_low=75, max_ratio=0.85):
    """ORBextractor::getMatches_distRatio (ORBextractor.cpp:1228-1290): (i_match_A, i_match_B, BookK)."""
    A = np.ascontiguousarray(A, np.uint8).reshape(-1, 32)
    B = np.ascontiguousarray(B, np.uint8).reshape(-1, 32)
    iA = np.ascontiguousarray(iA, np.uint32)
    iB = np.ascontiguousarray(iB, np.uint32)
    ma = np.zeros(max(len(iA), 1), np.uint32)
    mb = np.zeros(max(len(iA), 1), np.uint32)
    bk = C.c_longlong(0)
    n = lib().orc_matches_dist_ratio(_p(A), _p(iA), len(iA), _p(B), _p(iB), len(iB), th_low, max_ratio, _p(ma), _p(mb), C.byref(bk))
    return ma[:n], mb[:n], bk.value


def bow_matching(A, B, fv_a, fv_b, th_low=75, max_ratio=0.85):
    """BowMatching / InterMatchingBow node walk (MultiCameraFrame.cpp:946-1022, FrontEnd.cpp:3791-3965) over two feature vectors
    given as {node id: [indices]}: (indices_1, indices_2, words, BookK)."""
    A = np.ascontiguousarray(A, np.uint8).reshape(-1, 32)
    B = np.ascontiguousarray(B, np.uint8).reshape(-1, 32)
    na, oa, ia = feature_vector_csr(fv_a)
    nb, ob, ib = feature_vector_csr(fv_b)
    ma = np.zeros(max(len(ia), 1), np.uint32)
    mb = np.zeros(max(len(ia), 1), np.uint32)
    words = np.zeros(max(len(na), 1), np.uint32)
    nw, bk = C.c_int(0), C.c_longlong(0)
    n = lib().orc_bow_matching(_p(A), _p(B), _p(na), _p(oa), _p(ia), len(na), _p(nb), _p(ob), _p(ib), len(nb), th_low, max_ratio, _p(ma),
                               _p(mb), _p(words), C.byref(nw), C.byref(bk))
    return ma[:n], mb[:n], words[:nw.value], bk.value


def set_data_f32(images):
    """setData conversion (MultiCameraFrame.cpp:107-116): float32 [..., h, w] or [..., h, w, 3] (BGR) in [0, 1] -> uint8 gray."""
    images = np.ascontiguousarray(images, np.float32)
    ch = 3 if images.shape[-1] == 3 and images.ndim >= 3 else 1
    shape = images.shape[:-1] if ch == 3 else images.shape
    out = np.zeros(shape, np.uint8)
    lib().orc_set_data_f32(_p(images), ch, out.size, _p(out))
    return out


def argsorte_desc(values):
    values = np.ascontiguousarray(values, np.float32)
    idx = np.zeros(max(len(values), 1), np.int32)
    lib().orc_argsorte_desc(_p(values), len(values), _p(idx))
    return idx[:len(values)]


def obtain_lf_features(kps, tri, total_feats=3000):
    """FrontEnd::obtainLfFeatures bookkeeping around the triangulation block (FrontEnd.cpp:213-533, the shipped branch: tiling and
    refview off): accepted rows go to intraMatches in row order and consume their keypoints (:355-365); single-view rows (:371-384)
    and then every keypoint still unused, camera-major (:487-505), become mono candidates; the candidates are ordered by descending
    response with argsorte (:513) and the first total_feats - intramatch_size are appended (:515-521).
    Returns (row indices of the accepted rows, [(cam, keypoint index)] of the appended mono features)."""
    used = [np.zeros(len(k), bool) for k in kps]
    accepted, mono, resp = [], [], []
    for r in range(len(tri["rows"])):
        row = tri["rows"][r]
        if tri["status"][r] == 1:
            accepted.append(r)
            for c, j in enumerate(row):
                if j >= 0:
                    used[c][j] = True
        elif tri["nviews"][r] == 1:
            c = int(np.flatnonzero(row >= 0)[0])
            mono.append((c, int(row[c])))
            resp.append(kps[c][row[c]]["response"])
            used[c][row[c]] = True
    for c in range(len(kps)):
        for j in range(len(kps[c])):
            if not used[c][j]:
                mono.append((c, j))
                resp.append(kps[c][j]["response"])
    order = argsorte_desc(resp)
    take = max(0, total_feats - len(accepted))
    return accepted, [mono[i] for i in order[:take]]


def undistort(img, K, dist):
    """cv::undistort(img, K, dist) as setData applies it for RECTIFY rigs (MultiCameraFrame.cpp:123-136); uint8 or float32 [h, w]."""
    K = np.ascontiguousarray(K, np.float64).reshape(9)
    d5 = np.zeros(5, np.float64)
    d5[:len(np.ravel(dist))] = np.ravel(dist)[:5]
    img = np.ascontiguousarray(img)
    out = np.zeros_like(img)
    h, w = img.shape
    if img.dtype == np.uint8:
        lib().orc_undistort_u8(_p(img), w, h, _p(K), _p(d5), _p(out))
    else:
        assert img.dtype == np.float32
        lib().orc_undistort_f32(_p(img), w, h, _p(K), _p(d5), _p(out))
    return out


def undistort_keypoints(kps, K, dist):
    """UndistortKeyPoints (MultiCameraFrame.cpp:300-347): KP_DTYPE array with pt replaced by cv::undistortPoints(pt, K32, dist32, P=K32)."""
    kps = np.ascontiguousarray(kps, KP_DTYPE)
    K = np.ascontiguousarray(K, np.float64).reshape(9)
    d5 = np.zeros(5, np.float64)
    d5[:len(np.ravel(dist))] = np.ravel(dist)[:5]
    out = np.zeros_like(kps)
    lib().orc_undistort_keypoints(_p(kps), len(kps), _p(K), _p(d5), _p(out))
    return out


def vocab_transform(L, parent, is_leaf, node_desc, weight, feats, levelsup=4):
    """DBoW2 transform per feature on a vocabulary in loadFromTextFile layout: (word_id, node_id, weight) arrays."""
    parent = np.ascontiguousarray(parent, np.int32)
    is_leaf = np.ascontiguousarray(is_leaf, np.int32)
    node_desc = np.ascontiguousarray(node_desc, np.uint8).reshape(-1, 32)
    weight = np.ascontiguousarray(weight, np.float64)
    feats = np.ascontiguousarray(feats, np.uint8).reshape(-1, 32)
    n = len(feats)
    wid, nid, w = np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.float64)
    lib().orc_vocab_transform(L, len(parent), _p(parent), _p(is_leaf), _p(node_desc), _p(weight), _p(feats), n, levelsup, _p(wid), _p(nid), _p(w))
    return wid[:n], nid[:n], w[:n]


def feature_vector(node_id, weight):
    """FeatureVector of DBoW2's transform(features, bowVec, featVec, levelsup): {node id: [feature indices]}, features of weight 0
    (stopped words) left out."""
    fv = {}
    for i, (n, w) in enumerate(zip(node_id, weight)):
        if w > 0:
            fv.setdefault(int(n), []).append(i)
    return fv


def intra_matches_bow(descs, kps_undist, fvs, th_low=75, ratio=0.85):
    """computeIntraMatches(matches, words_) (MultiCameraFrame.cpp:586-943) over per-camera feature vectors {node id: [indices]}.
    Returns (rows [M, num_cams] int32, n_rays [M], words [n])."""
    nc = len(descs)
    descs = [np.ascontiguousarray(d, np.uint8).reshape(-1, 32) for d in descs]
    kps = [np.ascontiguousarray(k, KP_DTYPE) for k in kps_undist]
    csr = [feature_vector_csr(fv) for fv in fvs]
    nn = np.array([len(c[0]) for c in csr], np.int32)
    cap = int(sum(len(d) for d in descs)) + 1
    rows = np.full((cap, nc), -1, np.int32)
    nr = np.zeros(cap, np.int32)
    words = np.zeros(cap, np.uint32)
    nw = C.c_int(0)
    M = lib().orc_intra_matches_bow(nc, _ptr_array(descs), _ptr_array(kps), _p(nn), _ptr_array([c[0] for c in csr]), _ptr_array([c[1] for c in csr]),
                                    _ptr_array([c[2] for c in csr]), th_low, ratio, _p(rows), _p(nr), cap, _p(words), C.byref(nw))
    return rows[:M].copy(), nr[:M].copy(), words[:nw.value].copy()


def intra_matches_chained(descs, kps_undist, fvs):
    """computeIntraMatches_chained (MultiCameraFrame.cpp:1290-1355): BowMatching between cameras (j-1, j), |dy| < 50 on the undistorted
    keypoints, rows chained through the shared camera.  int32 [M, num_cams]."""
    nc = len(descs)
    inv = [np.full(len(d), -1, np.int64) for d in descs]
    rows = []
    for j in range(1, nc):
        i = j - 1
        i1, i2, _, _ = bow_matching(descs[i], descs[j], fvs[i], fvs[j])
        for fi, fj in zip(i1.tolist(), i2.tolist()):
            ri, rj = inv[i][fi], inv[j][fj]
            if abs(np.float32(kps_undist[i]["y"][fi]) - np.float32(kps_undist[j]["y"][fj])) >= 50:
                continue
            if ri == -1 and rj == -1:
                r = [-1] * nc
                r[i], r[j] = fi, fj
                inv[i][fi] = inv[j][fj] = len(rows)
                rows.append(r)
            elif rj == -1:
                rows[ri][j] = fj
                inv[j][fj] = ri
    return np.array(rows, np.int32).reshape(-1, nc)


def intra_matches_dmatch(descs, kps, fvs):
    """computeIntraMatches(matches, kps_1, kps_2) (MultiCameraFrame.cpp:1365-1466): {(i, j): (queryIdx, trainIdx, pts_i, pts_j)} per
    camera pair i < j, points taken from the distorted image_kps."""
    out = {}
    for i in range(len(descs) - 1):
        for j in range(i + 1, len(descs)):
            i1, i2, _, _ = bow_matching(descs[i], descs[j], fvs[i], fvs[j])
            out[(i, j)] = (i1, i2, np.stack([kps[i]["x"][i1], kps[i]["y"][i1]], 1), np.stack([kps[j]["x"][i2], kps[j]["y"][i2]], 1))
    return out


def _ptr_array(arrs):
    return (C.c_void_p * len(arrs))(*[a.ctypes.data if a is not None else None for a in arrs])


def intra_matches(descs, kps, calib, sigma2, use_epipolar=False, dist_thresh=75.0, ratio=0.85):
    """computeIntraMatches(matches, old) (MultiCameraFrame.cpp:1100-1288). Returns rows [M, num_cams] int32, pair_counts."""
    nc = len(descs)
    descs = [np.ascontiguousarray(d, np.uint8) for d in descs]
    kps = [np.ascontiguousarray(k, KP_DTYPE) for k in kps]
    n = np.array([len(d) for d in descs], np.int32)
    cap = int(n.sum()) + 1
    rows = np.full((cap, nc), -1, np.int32)
    pc = np.zeros(nc * (nc - 1) // 2, np.int32)
    calib = np.ascontiguousarray(calib, CALIB_DTYPE)
    sigma2 = np.ascontiguousarray(sigma2, np.float32)
    M = lib().orc_intra_matches(nc, _ptr_array(descs), _ptr_array(kps), _p(n), _p(calib), _p(sigma2), int(use_epipolar), dist_thresh,
                                ratio, _p(rows), cap, _p(pc))
    assert M >= 0
    return rows[:M].copy(), pc


def triangulate_pixels(calib, view_inds, xs, ys):
    calib = np.ascontiguousarray(calib, CALIB_DTYPE)
    vi = np.ascontiguousarray(view_inds, np.int32)
    xs = np.ascontiguousarray(xs, np.float64)
    ys = np.ascontiguousarray(ys, np.float64)
    X = np.zeros(3, np.float64)
    lib().orc_triangulate_pixels(_p(calib), _p(vi), _p(xs), _p(ys), len(vi), _p(X))
    return X


def triangulate_pair_f32(calib, cam1, cam2, p1, p2):
    calib = np.ascontiguousarray(calib, CALIB_DTYPE)
    p1 = np.ascontiguousarray(p1, np.float32)
    p2 = np.ascontiguousarray(p2, np.float32)
    out = np.zeros((len(p1), 3), np.float32)
    lib().orc_triangulate_pair_f32(_p(calib), cam1, cam2, _p(p1), _p(p2), len(p1), _p(out))
    return out


def triangulate_rows(calib, kps, descs, rows, masks=None):
    """obtainLfFeatures triangulation block (FrontEnd.cpp:249-370). masks: list of float32 [h,w] or None."""
    nc = len(kps)
    calib = np.ascontiguousarray(calib, CALIB_DTYPE)
    kps = [np.ascontiguousarray(k, KP_DTYPE) for k in kps]
    descs = [np.ascontiguousarray(d, np.uint8) for d in descs]
    rows = np.ascontiguousarray(rows, np.int32).reshape(-1, nc)
    M = len(rows)
    mw = mh = 0
    mp = None
    if masks is not None:
        masks = [np.ascontiguousarray(m, np.float32) for m in masks]
        mh, mw = masks[0].shape
        mp = _ptr_array(masks)
    rows_out = np.zeros((max(M, 1), nc), np.int32)
    status = np.zeros(max(M, 1), np.uint8)
    nviews = np.zeros(max(M, 1), np.int32)
    xyz = np.zeros((max(M, 1), 3), np.float64)
    uv = np.zeros((max(M, 1), 2), np.float32)
    rep = np.zeros(max(M, 1), np.int32)
    acc = lib().orc_triangulate_rows(nc, _p(calib), _ptr_array(kps), _ptr_array(descs), mp, mw, mh, _p(rows), M, _p(rows_out),
                                     _p(status), _p(nviews), _p(xyz), _p(uv), _p(rep))
    return dict(accepted=acc, rows=rows_out[:M], status=status[:M], nviews=nviews[:M], xyz=xyz[:M], uv_ref=uv[:M], rep_cam=rep[:M])


def process_rig(extractors, images, calib, threads=None, do_match=True, do_triangulate=True):
    """Whole CPU front-end for one rig frame (timed cpu_baseline). images: [C,h,w] uint8."""
    images = np.ascontiguousarray(images, np.uint8)
    nc, h, w = images.shape
    calib = np.ascontiguousarray(calib, CALIB_DTYPE)
    hs = (C.c_void_p * nc)(*[e.h for e in extractors])
    counts = np.zeros(nc, np.int32)
    acc = C.c_int(0)
    M = lib().orc_process_rig(hs, nc, _p(images), w, h, _p(calib), nc if threads is None else threads, int(do_match),
                              int(do_triangulate), _p(counts), C.byref(acc))
    return M, counts, acc.value
