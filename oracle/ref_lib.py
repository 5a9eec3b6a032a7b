"""ctypes binding of oracle/_ref/libmcsref.so — the REFERENCE's own front-end code compiled here (oracle/ref_build.py).

TEST INFRASTRUCTURE ONLY: tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
The classes mirror oracle_lib's (same call shapes) so a test can run against either checker.
"""
import ctypes as C
import os

import numpy as np

import oracle_lib as O
import ref_build

KP_DTYPE = O.KP_DTYPE
CALIB_DTYPE = O.CALIB_DTYPE
_LIB = None
_p = O._p
_ptr_array = O._ptr_array


def available():
    """True when libmcsref.so exists or can be built (needs /root/reference; on the GPU box only the prebuilt file)."""
    try:
        return ref_build.build() is not None
    except Exception:
        return os.path.exists(ref_build.SO)


def lib():
    global _LIB
    if _LIB is None:
        so = ref_build.build()
        if so is None:
            raise RuntimeError("oracle/_ref/libmcsref.so is missing and /root/reference is not present to build it")
        O.lib()  # liboracle.so first (libmcsref.so links against it)
        _LIB = C.CDLL(so)
        L = _LIB
        L.ref_last_error.restype = C.c_char_p
        L.ref_extractor_create.restype = C.c_void_p
        L.ref_extractor_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.ref_extractor_destroy.argtypes = [C.c_void_p]
        L.ref_extractor_tables.argtypes = [C.c_void_p] * 7
        L.ref_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.ref_level_size.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.ref_get_level.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.ref_distribute_octtree.argtypes = [C.c_void_p, C.c_void_p] + [C.c_int] * 7 + [C.c_void_p]
        L.ref_descriptor_distance.argtypes = [C.c_void_p] * 3
        L.ref_matches_dist_ratio.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int,
                                             C.c_double, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_rig_create.restype = C.c_void_p
        L.ref_rig_create.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int]
        L.ref_rig_destroy.argtypes = [C.c_void_p]
        L.ref_rig_extract.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_rig_set_features.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p]
        L.ref_rig_num_features.argtypes = [C.c_void_p, C.c_int]
        L.ref_rig_get_features.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_rig_bf_match.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float] + [C.c_void_p] * 4
        L.ref_rig_intra_matches.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        L.ref_rig_set_rows.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        L.ref_rig_obtain_lf_features.argtypes = [C.c_void_p] * 8 + [C.c_int, C.c_void_p, C.c_void_p]
        L.ref_rig_triangulate_nview.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.ref_rig_triangulate_pair.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.ref_rig_representative_desc.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.ref_rig_process.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    return _LIB


def _check(rc):
    if rc <= -1000 or rc in (-2, -3, -4):
        raise RuntimeError("libmcsref: " + lib().ref_last_error().decode())
    return rc


class Extractor:
    """The reference's ORBextractor object (MCSlam/src/ORBextractor.cpp compiled whole)."""

    def __init__(self, nfeatures=1000, scale=1.2, nlevels=8, ini_th=20, min_th=7):
        self.nfeatures, self.nlevels = nfeatures, nlevels
        self.h = lib().ref_extractor_create(nfeatures, scale, nlevels, ini_th, min_th)
        self.scale = np.zeros(nlevels, np.float32)
        self.inv_scale = np.zeros(nlevels, np.float32)
        self.sigma2 = np.zeros(nlevels, np.float32)
        self.inv_sigma2 = np.zeros(nlevels, np.float32)
        self.quota = np.zeros(nlevels, np.int32)
        self.umax = np.zeros(16, np.int32)
        lib().ref_extractor_tables(self.h, _p(self.scale), _p(self.inv_scale), _p(self.sigma2), _p(self.inv_sigma2), _p(self.quota), _p(self.umax))

    def __del__(self):
        if getattr(self, "h", None):
            lib().ref_extractor_destroy(self.h)
            self.h = None

    def __call__(self, image, lapping=(0, 0)):
        """operator()(image, mask, keypoints, descriptors, vLappingArea): (monoIndex, keypoints[KP_DTYPE], descriptors[n, 32])."""
        if image is None or image.size == 0:
            tot = C.c_int(0)
            return lib().ref_extract(self.h, None, 0, 0, 0, lapping[0], lapping[1], None, None, 0, C.byref(tot)), np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8)
        assert image.dtype == np.uint8 and image.strides[1] == 1
        h, w = image.shape
        cap = self.nfeatures + 64
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        tot = C.c_int(0)
        mono = _check(lib().ref_extract(self.h, C.c_void_p(image.ctypes.data), w, h, image.strides[0], lapping[0], lapping[1], _p(kps), _p(desc), cap,
                                        C.byref(tot)))
        return mono, kps[:tot.value].copy(), desc[:tot.value].copy()

    def level(self, l, padded=False):
        w, h = C.c_int(), C.c_int()
        if lib().ref_level_size(self.h, l, C.byref(w), C.byref(h)):
            return None
        pad = 38 if padded else 0
        out = np.zeros((h.value + pad, w.value + pad), np.uint8)
        lib().ref_get_level(self.h, l, int(padded), _p(out))
        return out

    def level_padded(self, l):
        return self.level(l, True)

    def distribute_octtree(self, xyr, minX, maxX, minY, maxY, N, level=0):
        """DistributeOctTree on (x, y, response) candidates: the retained keys (x, y, response) in the reference's list order."""
        xyr = np.ascontiguousarray(xyr, np.float32).reshape(-1, 3)
        out = np.zeros((len(xyr) + 8, 3), np.float32)
        n = _check(lib().ref_distribute_octtree(self.h, _p(xyr), len(xyr), minX, maxX, minY, maxY, N, level, _p(out)))
        return out[:n].copy()

    def descriptor_distance(self, a, b):
        a = np.ascontiguousarray(a, np.uint8)
        b = np.ascontiguousarray(b, np.uint8)
        return lib().ref_descriptor_distance(self.h, _p(a), _p(b))

    def matches_dist_ratio(self, A, iA, B, iB, max_ratio=0.85):
        A = np.ascontiguousarray(A, np.uint8).reshape(-1, 32)
        B = np.ascontiguousarray(B, np.uint8).reshape(-1, 32)
        iA = np.ascontiguousarray(iA, np.uint32)
        iB = np.ascontiguousarray(iB, np.uint32)
        ma = np.zeros(max(len(iA), 1), np.uint32)
        mb = np.zeros(max(len(iA), 1), np.uint32)
        bk = C.c_longlong(0)
        n = _check(lib().ref_matches_dist_ratio(self.h, _p(A), len(A), _p(iA), len(iA), _p(B), len(B), _p(iB), len(iB), max_ratio, _p(ma), _p(mb),
                                                C.byref(bk)))
        return ma[:n], mb[:n], bk.value


class Rig:
    """The reference's MultiCameraFrame + FrontEnd::obtainLfFeatures over one rig (method bodies extracted by line range)."""

    def __init__(self, calib, w, h, nfeatures=1000, scale=1.2, nlevels=8, ini_th=20, min_th=7, rectify=False):
        self.calib = np.ascontiguousarray(calib, CALIB_DTYPE)
        self.nc, self.w, self.hh, self.nfeatures = len(self.calib), w, h, nfeatures
        self.h = lib().ref_rig_create(self.nc, _p(self.calib), w, h, nfeatures, scale, nlevels, ini_th, min_th, int(rectify))

    def __del__(self):
        if getattr(self, "h", None):
            lib().ref_rig_destroy(self.h)
            self.h = None

    def extract(self, images):
        """extractFeaturesParallel (one thread per camera) on [C, h, w] uint8."""
        images = np.ascontiguousarray(images, np.uint8)
        assert images.shape == (self.nc, self.hh, self.w)
        _check(lib().ref_rig_extract(self.h, _p(images)))

    def set_features(self, cam, kps, desc):
        kps = np.ascontiguousarray(kps, KP_DTYPE)
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        _check(lib().ref_rig_set_features(self.h, cam, _p(kps), len(kps), _p(desc)))

    def features(self, cam):
        """(image_kps, image_kps_undist, image_descriptors) of one camera."""
        n = lib().ref_rig_num_features(self.h, cam)
        kps, und, desc = np.zeros(max(n, 1), KP_DTYPE), np.zeros(max(n, 1), KP_DTYPE), np.zeros((max(n, 1), 32), np.uint8)
        lib().ref_rig_get_features(self.h, cam, _p(kps), _p(und), _p(desc))
        return kps[:n], und[:n], desc[:n]

    def bf_match(self, i, j, dist_thresh=75.0, ratio=0.85):
        """BruteForceMatch: (indices_1, indices_2, kps1, kps2)."""
        n = max(lib().ref_rig_num_features(self.h, i), 1)
        i1, i2 = np.zeros(n, np.uint32), np.zeros(n, np.uint32)
        k1, k2 = np.zeros(n, KP_DTYPE), np.zeros(n, KP_DTYPE)
        m = _check(lib().ref_rig_bf_match(self.h, i, j, dist_thresh, ratio, _p(i1), _p(i2), _p(k1), _p(k2)))
        return i1[:m].copy(), i2[:m].copy(), k1[:m].copy(), k2[:m].copy()

    def intra_matches(self, old=False):
        """computeIntraMatches(matches, old): rows int32 [M, num_cams] (the reference's matchIndex is 5 wide)."""
        cap = sum(lib().ref_rig_num_features(self.h, c) for c in range(self.nc)) + 1
        rows = np.full((cap, 5), -1, np.int32)
        M = _check(lib().ref_rig_intra_matches(self.h, int(old), _p(rows), cap))
        return rows[:M, :self.nc].copy()

    def set_rows(self, rows):
        rows = np.ascontiguousarray(rows, np.int32).reshape(-1, self.nc)
        r5 = np.full((len(rows), 5), -1, np.int32)
        r5[:, :self.nc] = rows
        lib().ref_rig_set_rows(self.h, _p(r5), len(r5))

    def obtain_lf_features(self, masks=None):
        """FrontEnd::obtainLfFeatures on the stored rows: dict of arrays describing currentFrame->intraMatches in order."""
        cap = sum(lib().ref_rig_num_features(self.h, c) for c in range(self.nc)) + 16
        mi = np.full((cap, 5), -1, np.int32)
        uv = np.zeros((cap, 2), np.float32)
        mono = np.zeros(cap, np.uint8)
        nr = np.zeros(cap, np.int32)
        xyz = np.zeros((cap, 3), np.float64)
        desc = np.zeros((cap, 32), np.uint8)
        isz, msz = C.c_int(0), C.c_int(0)
        mp = None
        if masks is not None:
            masks = [np.ascontiguousarray(m, np.float32) for m in masks]
            assert all(m.shape == (self.hh, self.w) for m in masks)
            mp = _ptr_array(masks)
        n = _check(lib().ref_rig_obtain_lf_features(self.h, mp, _p(mi), _p(uv), _p(mono), _p(nr), _p(xyz), _p(desc), cap, C.byref(isz), C.byref(msz)))
        return dict(match_index=mi[:n, :self.nc].copy(), uv_ref=uv[:n].copy(), mono=mono[:n].astype(bool), n_rays=nr[:n].copy(), xyz=xyz[:n].copy(),
                    desc=desc[:n].copy(), intramatch_size=isz.value, mono_size=msz.value)

    def triangulate_nview(self, view_inds, xs, ys):
        vi = np.ascontiguousarray(view_inds, np.int32)
        xs = np.ascontiguousarray(xs, np.float64)
        ys = np.ascontiguousarray(ys, np.float64)
        X = np.zeros(3, np.float64)
        _check(lib().ref_rig_triangulate_nview(self.h, _p(vi), _p(xs), _p(ys), len(vi), _p(X)))
        return X

    def triangulate_pair(self, cam1, cam2, p1, p2):
        p1 = np.ascontiguousarray(p1, np.float32).reshape(-1, 2)
        p2 = np.ascontiguousarray(p2, np.float32).reshape(-1, 2)
        out = np.zeros((len(p1), 3), np.float32)
        _check(lib().ref_rig_triangulate_pair(self.h, cam1, cam2, _p(p1), _p(p2), len(p1), _p(out)))
        return out

    def representative_desc(self, descs):
        descs = np.ascontiguousarray(descs, np.uint8).reshape(-1, 32)
        out = np.zeros(32, np.uint8)
        _check(lib().ref_rig_representative_desc(self.h, _p(descs), len(descs), _p(out)))
        return out

    def process(self, images, do_match=True, do_triangulate=True):
        """Whole CPU front-end for one rig frame through the reference's calls (timed cpu_baseline): (rows, kp counts, accepted)."""
        images = np.ascontiguousarray(images, np.uint8)
        counts = np.zeros(self.nc, np.int32)
        acc = C.c_int(0)
        M = _check(lib().ref_rig_process(self.h, _p(images), int(do_match), int(do_triangulate), _p(counts), C.byref(acc)))
        return M, counts, acc.value
