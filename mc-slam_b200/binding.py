"""ctypes binding of libmcslam_b200.so — the host-side mirror of the reference's ORBextractor / MultiCameraFrame call shapes
for Python harnesses (tests, bench).  Fails loudly when the CUDA library is missing or no GPU is present: there is no CPU path.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
CALIB_DTYPE = np.dtype([("K", "<f8", (9,)), ("R", "<f8", (9,)), ("t", "<f8", (3,)), ("dist", "<f8", (5,))])


class McsError(RuntimeError):
    pass


class _RigConfig(C.Structure):
    _fields_ = [("num_cams", C.c_int), ("width", C.c_int), ("height", C.c_int), ("nfeatures", C.c_int), ("scale_factor", C.c_float),
                ("nlevels", C.c_int), ("ini_th_fast", C.c_int), ("min_th_fast", C.c_int), ("max_frames", C.c_int), ("device", C.c_int)]


class _Landmarks(C.Structure):
    _fields_ = [("rows", C.c_void_p), ("accepted", C.c_void_p), ("n_views", C.c_void_p), ("xyz", C.c_void_p), ("uv_ref", C.c_void_p),
                ("rep_cam", C.c_void_p)]


EXPORTS = [
    "mcs_last_error", "mcs_device_count", "mcs_bind_host_thread", "mcs_version", "mcs_extractor_create", "mcs_extractor_destroy", "mcs_extractor_set_orientation",
    "mcs_extractor_capacity", "mcs_extract", "mcs_extract_batch", "mcs_extractor_get_tables", "mcs_extractor_level_size",
    "mcs_extractor_download_level", "mcs_extractor_download_candidates", "mcs_rig_create", "mcs_rig_destroy", "mcs_rig_feature_capacity",
    "mcs_rig_row_capacity", "mcs_rig_set_orientation", "mcs_rig_extract", "mcs_rig_extract_device", "mcs_rig_get_features",
    "mcs_rig_brute_force_match", "mcs_rig_compute_intra_matches", "mcs_rig_get_intra_matches", "mcs_rig_triangulate",
    "mcs_rig_triangulate_device_masks", "mcs_rig_get_landmarks", "mcs_rig_triangulate_nview", "mcs_rig_triangulate_pair", "mcs_rig_process",
    "mcs_rig_process_device", "mcs_rig_get_counts", "mcs_rig_stream", "mcs_rig_synchronize", "mcs_rig_kernel_launches", "mcs_rig_stage_times",
    "mcs_rig_set_profiling", "mcs_rig_timer_start", "mcs_rig_timer_stop", "mcs_rig_set_chunk_frames", "mcs_device_malloc", "mcs_device_free", "mcs_device_upload", "mcs_device_download", "mcs_host_alloc_pinned",
    "mcs_host_free_pinned", "mcs_rig_download_level", "mcs_rig_level_size", "mcs_matcher_create", "mcs_matcher_destroy", "mcs_knn2",
    "mcs_knn2_device", "mcs_knn2_device_static", "mcs_matcher_stream", "mcs_matcher_synchronize", "mcs_matcher_kernel_launches", "mcs_rig_knn2_images", "mcs_match_subsets", "mcs_rig_process_f32", "mcs_rig_process_device_f32", "mcs_rig_set_features", "mcs_rig_set_undistort", "mcs_rig_set_undistort_keypoints", "mcs_rig_get_keypoints_undist", "mcs_vocab_load_text", "mcs_vocab_create", "mcs_vocab_destroy", "mcs_vocab_info",
    "mcs_vocab_transform", "mcs_rig_vocab_transform", "mcs_rig_bow_best2", "mcs_plan_fast_tables", "mcs_rig_process_masks_u8",
    "mcs_rig_process_device_masks_u8", "mcs_rig_set_graph", "mcs_rig_graph_replays", "mcs_rig_last_d2h_bytes",
]


def library_path():
    # MCS_LIB: another build of the same library (kernel tuning experiments, tools/exp/)
    return os.environ.get("MCS_LIB") or os.path.join(_HERE, "libmcslam_b200.so")


def lib():
    """Loads the CUDA library.  Raises if it has not been built (python mc-slam_b200/build.py / __graft_entry__.build())."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = library_path()
    if not os.path.exists(path):
        raise McsError("libmcslam_b200.so is missing: build it with __graft_entry__.build() — there is no CPU fallback")
    L = C.CDLL(path)
    vp, ci, cf = C.c_void_p, C.c_int, C.c_float
    L.mcs_last_error.restype = C.c_char_p
    L.mcs_version.restype = C.c_char_p
    L.mcs_extractor_create.restype = vp
    L.mcs_extractor_create.argtypes = [ci, cf, ci, ci, ci, ci, ci, ci, ci]
    L.mcs_extractor_destroy.argtypes = [vp]
    L.mcs_extractor_set_orientation.argtypes = [vp, ci]
    L.mcs_extractor_capacity.argtypes = [vp]
    L.mcs_extract.argtypes = [vp, vp, ci, ci, ci, vp, vp, vp, vp, ci, vp]
    L.mcs_extract_batch.argtypes = [vp, vp, ci, ci, ci, ci, C.c_size_t, vp, vp, vp, ci, vp, vp]
    L.mcs_extractor_get_tables.argtypes = [vp] + [vp] * 6
    L.mcs_extractor_level_size.argtypes = [vp, ci, vp, vp]
    L.mcs_extractor_download_level.argtypes = [vp, ci, ci, ci, vp, ci]
    L.mcs_extractor_download_candidates.argtypes = [vp, ci, ci, vp, ci]
    L.mcs_rig_create.restype = vp
    L.mcs_rig_create.argtypes = [vp, vp]
    L.mcs_rig_destroy.argtypes = [vp]
    L.mcs_rig_feature_capacity.argtypes = [vp]
    L.mcs_rig_row_capacity.argtypes = [vp]
    L.mcs_rig_set_orientation.argtypes = [vp, ci]
    L.mcs_rig_extract.argtypes = [vp, vp, ci]
    L.mcs_rig_extract_device.argtypes = [vp, vp, ci]
    L.mcs_rig_get_features.argtypes = [vp, ci, ci, vp, vp, ci]
    L.mcs_rig_brute_force_match.argtypes = [vp, ci, ci, ci, cf, cf, vp, vp, vp, vp, ci]
    L.mcs_rig_compute_intra_matches.argtypes = [vp, ci, ci, cf, cf]
    L.mcs_rig_get_intra_matches.argtypes = [vp, ci, vp, ci]
    L.mcs_rig_triangulate.argtypes = [vp, ci, vp]
    L.mcs_rig_triangulate_device_masks.argtypes = [vp, ci, vp]
    L.mcs_rig_get_landmarks.argtypes = [vp, ci, vp, ci]
    L.mcs_rig_triangulate_nview.argtypes = [vp, vp, ci, vp, vp, ci, vp]
    L.mcs_rig_triangulate_pair.argtypes = [vp, ci, ci, vp, vp, ci, vp]
    L.mcs_rig_process.argtypes = [vp, vp, vp, ci, ci]
    L.mcs_rig_process_device.argtypes = [vp, vp, vp, ci, ci]
    L.mcs_rig_get_counts.argtypes = [vp, ci, vp, vp, vp]
    L.mcs_rig_stream.restype = vp
    L.mcs_rig_stream.argtypes = [vp]
    L.mcs_rig_synchronize.argtypes = [vp]
    L.mcs_rig_kernel_launches.restype = C.c_longlong
    L.mcs_rig_kernel_launches.argtypes = [vp]
    L.mcs_rig_stage_times.argtypes = [vp, vp]
    L.mcs_rig_set_profiling.argtypes = [vp, ci]
    L.mcs_rig_timer_start.argtypes = [vp]
    L.mcs_rig_set_chunk_frames.argtypes = [vp, ci]
    L.mcs_rig_process_masks_u8.argtypes = [vp, vp, vp, ci, ci]
    L.mcs_rig_process_device_masks_u8.argtypes = [vp, vp, vp, ci, ci]
    L.mcs_rig_set_graph.argtypes = [vp, ci]
    L.mcs_rig_graph_replays.argtypes = [vp]
    L.mcs_rig_graph_replays.restype = C.c_longlong
    L.mcs_rig_last_d2h_bytes.argtypes = [vp]
    L.mcs_rig_last_d2h_bytes.restype = C.c_longlong
    L.mcs_rig_timer_stop.argtypes = [vp, vp]
    L.mcs_rig_download_level.argtypes = [vp, ci, ci, ci, ci, vp, ci]
    L.mcs_rig_level_size.argtypes = [vp, ci, vp, vp]
    L.mcs_matcher_create.restype = vp
    L.mcs_matcher_create.argtypes = [ci, ci, ci, ci]
    L.mcs_matcher_destroy.argtypes = [vp]
    L.mcs_knn2.argtypes = [vp, vp, vp, vp, vp, ci, vp]
    L.mcs_knn2_device.argtypes = [vp, vp, vp, vp, vp, ci, vp]
    L.mcs_knn2_device_static.argtypes = [ci, vp, vp, vp, vp, vp, ci, vp]
    L.mcs_matcher_stream.restype = vp
    L.mcs_matcher_stream.argtypes = [vp]
    L.mcs_matcher_synchronize.argtypes = [vp]
    L.mcs_matcher_kernel_launches.restype = C.c_longlong
    L.mcs_matcher_kernel_launches.argtypes = [vp]
    L.mcs_rig_knn2_images.argtypes = [vp, vp, vp, ci, vp]
    L.mcs_rig_set_features.argtypes = [vp, ci, vp, vp, vp]
    L.mcs_rig_set_undistort.argtypes = [vp, ci]
    L.mcs_vocab_load_text.restype = vp
    L.mcs_vocab_load_text.argtypes = [C.c_char_p, ci]
    L.mcs_vocab_create.restype = vp
    L.mcs_vocab_create.argtypes = [ci, ci, ci, ci, ci, ci, vp, vp, vp, vp]
    L.mcs_vocab_destroy.argtypes = [vp]
    L.mcs_vocab_info.argtypes = [vp, vp, vp, vp, vp]
    L.mcs_vocab_transform.argtypes = [vp, vp, ci, ci, vp, vp, vp]
    L.mcs_rig_vocab_transform.argtypes = [vp, vp, ci, vp, vp, vp]
    L.mcs_rig_bow_best2.argtypes = [vp, ci, ci, vp, vp, vp, vp, vp, vp, cf, vp, vp, vp]
    L.mcs_rig_set_undistort_keypoints.argtypes = [vp, ci]
    L.mcs_rig_get_keypoints_undist.argtypes = [vp, ci, ci, vp, ci]
    L.mcs_rig_process_f32.argtypes = [vp, vp, ci, vp, ci, ci]
    L.mcs_rig_process_device_f32.argtypes = [vp, vp, ci, vp, ci, ci]
    L.mcs_match_subsets.argtypes = [vp, vp, ci, vp, ci, ci, vp, vp, vp, vp, ci, C.c_double, vp, vp, vp, vp]
    L.mcs_device_malloc.restype = vp
    L.mcs_device_malloc.argtypes = [ci, C.c_size_t]
    L.mcs_device_free.argtypes = [ci, vp]
    L.mcs_device_upload.argtypes = [ci, vp, vp, C.c_size_t]
    L.mcs_device_download.argtypes = [ci, vp, vp, C.c_size_t]
    L.mcs_host_alloc_pinned.restype = vp
    L.mcs_host_alloc_pinned.argtypes = [C.c_size_t]
    L.mcs_host_free_pinned.argtypes = [vp]
    _LIB = L
    return L


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _check(rc, what):
    if rc < 0:
        raise McsError("%s failed (%d): %s" % (what, rc, lib().mcs_last_error().decode()))
    return rc


def pinned_empty(shape, dtype):
    """numpy array backed by page-locked host memory (freed with the library's allocator when collected)."""
    n = int(np.prod(shape)) * np.dtype(dtype).itemsize
    ptr = lib().mcs_host_alloc_pinned(max(n, 1))
    if not ptr:
        raise McsError("pinned allocation failed")
    buf = (C.c_char * max(n, 1)).from_address(ptr)
    arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
    return _Pinned(arr, ptr)


class _Pinned:
    def __init__(self, arr, ptr):
        self.arr, self.ptr = arr, ptr

    def __del__(self):
        try:
            lib().mcs_host_free_pinned(self.ptr)
        except Exception:
            pass


class Extractor:
    """Mirror of ORBextractor (MCSlam/include/MCSlam/ORBextractor.h:44-112): `n = ex(image, mask, keypoints..., vLappingArea)`."""

    def __init__(self, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7, device=0, max_width=720, max_height=540,
                 max_images=1):
        L = lib()
        self.h = L.mcs_extractor_create(nfeatures, scale_factor, nlevels, ini_th, min_th, device, max_width, max_height, max_images)
        if not self.h:
            raise McsError("mcs_extractor_create failed: " + L.mcs_last_error().decode())
        self.nlevels = nlevels

    def close(self):
        if getattr(self, "h", None):
            lib().mcs_extractor_destroy(self.h)
            self.h = None

    __del__ = close

    @property
    def capacity(self):
        return lib().mcs_extractor_capacity(self.h)

    def set_orientation(self, on):
        _check(lib().mcs_extractor_set_orientation(self.h, int(on)), "set_orientation")

    def tables(self):
        n = self.nlevels
        sf, inv, s2, is2 = (np.zeros(n, np.float32) for _ in range(4))
        q = np.zeros(n, np.int32)
        nl = C.c_int()
        _check(lib().mcs_extractor_get_tables(self.h, C.byref(nl), _p(sf), _p(inv), _p(s2), _p(is2), _p(q)), "get_tables")
        return dict(nlevels=nl.value, scale=sf, inv_scale=inv, sigma2=s2, inv_sigma2=is2, quota=q)

    def __call__(self, image, mask=None, lapping=(0, 0)):
        """operator()(image, mask, keypoints, descriptors, vLappingArea) -> (monoIndex, keypoints, descriptors)."""
        if image is None or image.size == 0:  # `if(_image.empty()) return -1;` — answered by the library itself
            nt = C.c_int()
            rc = lib().mcs_extract(self.h, None, 0, 0, 0, None, None, None, None, 0, C.byref(nt))
            return rc, np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8)
        assert image.dtype == np.uint8 and image.ndim == 2 and image.strides[1] == 1
        h, w = image.shape
        lap = np.array(lapping, np.int32)
        rc = 0
        for attempt in range(2):  # capacity depends on the (image-size dependent) plan; retry once with the real value
            cap = max(self.capacity, 8)
            kps = np.zeros(cap, KP_DTYPE)
            desc = np.zeros((cap, 32), np.uint8)
            nt = C.c_int()
            rc = lib().mcs_extract(self.h, C.c_void_p(image.ctypes.data), w, h, image.strides[0], None, _p(lap), _p(kps), _p(desc), cap,
                                   C.byref(nt))
            if rc != -2 or self.capacity <= cap:
                break
        _check(rc, "mcs_extract")
        return rc, kps[:nt.value].copy(), desc[:nt.value].copy()

    def extract_batch(self, images, lapping=(0, 0)):
        images = np.ascontiguousarray(images, np.uint8)
        n, h, w = images.shape
        lap = np.array(lapping, np.int32)
        mono = np.zeros(n, np.int32)
        tot = np.zeros(n, np.int32)
        rc = 0
        for attempt in range(2):
            cap = max(self.capacity, 8)
            kps = np.zeros((n, cap), KP_DTYPE)
            desc = np.zeros((n, cap, 32), np.uint8)
            rc = lib().mcs_extract_batch(self.h, _p(images), n, w, h, w, w * h, _p(lap), _p(kps), _p(desc), cap, _p(mono), _p(tot))
            if rc != -2 or self.capacity <= cap:
                break
        _check(rc, "mcs_extract_batch")
        return mono, [kps[i, :tot[i]].copy() for i in range(n)], [desc[i, :tot[i]].copy() for i in range(n)]

    def level(self, image_index, level, kind=0):
        w, h = C.c_int(), C.c_int()
        _check(lib().mcs_extractor_level_size(self.h, level, C.byref(w), C.byref(h)), "level_size")
        pad = 38 if kind == 1 else 0
        out = np.zeros((h.value + pad, w.value + pad), np.uint8)
        _check(lib().mcs_extractor_download_level(self.h, image_index, level, kind, _p(out), out.shape[1]), "download_level")
        return out

    def candidates(self, image_index, level):
        cap = 1 << 20
        out = np.zeros((cap, 3), np.int32)
        n = _check(lib().mcs_extractor_download_candidates(self.h, image_index, level, _p(out), cap), "download_candidates")
        return out[:n].copy()


class RigFrames:
    """Mirror of MultiCameraFrame's extract / intra-match / triangulate entry points for a batch of rig frames."""

    def __init__(self, calib, width=720, height=540, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7, max_frames=1,
                 device=0):
        L = lib()
        self.calib = np.ascontiguousarray(calib, CALIB_DTYPE)
        self.C, self.w, self.h, self.F, self.device = len(self.calib), width, height, max_frames, device
        cfg = _RigConfig(self.C, width, height, nfeatures, scale_factor, nlevels, ini_th, min_th, max_frames, device)
        self.hd = L.mcs_rig_create(C.byref(cfg), _p(self.calib))
        if not self.hd:
            raise McsError("mcs_rig_create failed: " + L.mcs_last_error().decode())
        self.kp_cap = L.mcs_rig_feature_capacity(self.hd)
        self.row_cap = L.mcs_rig_row_capacity(self.hd)
        self.nlevels = nlevels

    def close(self):
        if getattr(self, "hd", None):
            lib().mcs_rig_destroy(self.hd)
            self.hd = None

    __del__ = close

    def set_orientation(self, on):
        _check(lib().mcs_rig_set_orientation(self.hd, int(on)), "rig_set_orientation")

    # --- stage A: extractFeaturesParallel (MultiCameraFrame.cpp:203) -------------------------------------------------
    def extract(self, images):
        images = np.ascontiguousarray(images, np.uint8).reshape(-1, self.C, self.h, self.w)
        _check(lib().mcs_rig_extract(self.hd, _p(images), images.shape[0]), "mcs_rig_extract")
        return images.shape[0]

    def extract_device(self, d_ptr, n_frames):
        _check(lib().mcs_rig_extract_device(self.hd, C.c_void_p(d_ptr), n_frames), "mcs_rig_extract_device")

    def set_undistort(self, on):
        """RECTIFY branch of setData: cv::undistort of images and masks with the calibration's distortion coefficients."""
        _check(lib().mcs_rig_set_undistort(self.hd, int(on)), "rig_set_undistort")

    def set_undistort_keypoints(self, on):
        """RECTIFY = false: image_kps_undist = cv::undistortPoints(image_kps) (UndistortKeyPoints, MultiCameraFrame.cpp:300-347)."""
        _check(lib().mcs_rig_set_undistort_keypoints(self.hd, int(on)), "rig_set_undistort_keypoints")

    def keypoints_undist(self, frame, cam):
        k = np.zeros(self.kp_cap, KP_DTYPE)
        n = _check(lib().mcs_rig_get_keypoints_undist(self.hd, frame, cam, _p(k), self.kp_cap), "rig_get_keypoints_undist")
        return k[:n].copy()

    def set_features(self, kps, descs):
        """Features from elsewhere instead of stage A: kps / descs = per image (frame-major, camera-minor) KP_DTYPE / uint8 [n, 32]."""
        counts = np.array([len(k) for k in kps], np.int32)
        K = np.concatenate([np.ascontiguousarray(k, KP_DTYPE) for k in kps]) if len(kps) else np.zeros(0, KP_DTYPE)
        D = np.concatenate([np.ascontiguousarray(d, np.uint8).reshape(-1, 32) for d in descs])
        _check(lib().mcs_rig_set_features(self.hd, len(kps) // self.C, _p(counts), _p(K), _p(D)), "rig_set_features")

    def features(self, frame, cam):
        kps = np.zeros(self.kp_cap, KP_DTYPE)
        desc = np.zeros((self.kp_cap, 32), np.uint8)
        n = _check(lib().mcs_rig_get_features(self.hd, frame, cam, _p(kps), _p(desc), self.kp_cap), "mcs_rig_get_features")
        return kps[:n].copy(), desc[:n].copy()

    # --- stage B: BruteForceMatch / computeIntraMatches (MultiCameraFrame.cpp:1024, :1100) ---------------------------
    def brute_force_match(self, frame, cam1, cam2, dist_thresh=75.0, ratio=0.85):
        i1 = np.zeros(self.kp_cap, np.uint32)
        i2 = np.zeros(self.kp_cap, np.uint32)
        k1 = np.zeros(self.kp_cap, KP_DTYPE)
        k2 = np.zeros(self.kp_cap, KP_DTYPE)
        n = _check(lib().mcs_rig_brute_force_match(self.hd, frame, cam1, cam2, dist_thresh, ratio, _p(i1), _p(i2), _p(k1), _p(k2),
                                                   self.kp_cap), "mcs_rig_brute_force_match")
        return i1[:n].copy(), i2[:n].copy(), k1[:n].copy(), k2[:n].copy()

    def compute_intra_matches(self, n_frames, old=False, dist_thresh=75.0, ratio=0.85):
        _check(lib().mcs_rig_compute_intra_matches(self.hd, n_frames, int(old), dist_thresh, ratio), "mcs_rig_compute_intra_matches")

    def intra_matches(self, frame):
        rows = np.zeros((self.row_cap, self.C), np.int32)
        m = _check(lib().mcs_rig_get_intra_matches(self.hd, frame, _p(rows), self.row_cap), "mcs_rig_get_intra_matches")
        return rows[:m].copy()

    # --- stage C: obtainLfFeatures triangulation block (FrontEnd.cpp:249-370) ----------------------------------------
    def triangulate(self, n_frames, masks=None):
        if masks is not None:
            masks = np.ascontiguousarray(masks, np.float32).reshape(n_frames, self.C, self.h, self.w)
        _check(lib().mcs_rig_triangulate(self.hd, n_frames, _p(masks)), "mcs_rig_triangulate")

    def landmarks(self, frame):
        cap = self.row_cap
        rows = np.zeros((cap, self.C), np.int32)
        acc = np.zeros(cap, np.uint8)
        nv = np.zeros(cap, np.int32)
        xyz = np.zeros((cap, 3), np.float64)
        uv = np.zeros((cap, 2), np.float32)
        rep = np.zeros(cap, np.int32)
        lm = _Landmarks(rows.ctypes.data, acc.ctypes.data, nv.ctypes.data, xyz.ctypes.data, uv.ctypes.data, rep.ctypes.data)
        m = _check(lib().mcs_rig_get_landmarks(self.hd, frame, C.byref(lm), cap), "mcs_rig_get_landmarks")
        return dict(rows=rows[:m].copy(), status=acc[:m].copy(), nviews=nv[:m].copy(), xyz=xyz[:m].copy(), uv_ref=uv[:m].copy(),
                    rep_cam=rep[:m].copy())

    def triangulate_nview(self, view_inds, xs, ys):
        vi = np.ascontiguousarray(view_inds, np.int32)
        xs = np.ascontiguousarray(xs, np.float64).reshape(-1, len(vi))
        ys = np.ascontiguousarray(ys, np.float64).reshape(-1, len(vi))
        out = np.zeros((xs.shape[0], 3), np.float64)
        _check(lib().mcs_rig_triangulate_nview(self.hd, _p(vi), len(vi), _p(xs), _p(ys), xs.shape[0], _p(out)), "mcs_rig_triangulate_nview")
        return out

    def triangulate_pair(self, cam1, cam2, p1, p2):
        p1 = np.ascontiguousarray(p1, np.float32).reshape(-1, 2)
        p2 = np.ascontiguousarray(p2, np.float32).reshape(-1, 2)
        out = np.zeros((len(p1), 3), np.float32)
        _check(lib().mcs_rig_triangulate_pair(self.hd, cam1, cam2, _p(p1), _p(p2), len(p1), _p(out)), "mcs_rig_triangulate_pair")
        return out

    # --- whole path (FrontEnd::processFrame, FrontEnd.cpp:991-1028) -------------------------------------------------
    def process(self, images, masks=None, old=False, n_frames=None):
        """images: uint8 [F, C, h, w] host array (or raw pointer int with n_frames)."""
        if isinstance(images, int):
            ptr, nf = C.c_void_p(images), n_frames
        else:
            images = np.ascontiguousarray(images, np.uint8).reshape(-1, self.C, self.h, self.w)
            ptr, nf = _p(images), images.shape[0]
        mp = None
        if masks is not None:
            if isinstance(masks, int):
                mp = C.c_void_p(masks)
            else:
                masks = np.ascontiguousarray(masks, np.float32)
                mp = _p(masks)
        _check(lib().mcs_rig_process(self.hd, ptr, mp, nf, int(old)), "mcs_rig_process")
        return nf

    def process_f32(self, images, masks=None, old=False):
        """setData + whole path on float32 images in [0, 1]: [n_frames, C, h, w] (gray) or [n_frames, C, h, w, 3] (BGR)."""
        images = np.asarray(images)
        assert images.dtype == np.float32 and images.flags["C_CONTIGUOUS"]
        ch = 3 if (images.ndim == 5 and images.shape[-1] == 3) else 1
        n = images.size // (self.C * self.h * self.w * ch)
        if masks is not None:
            masks = np.ascontiguousarray(masks, np.float32)
        _check(lib().mcs_rig_process_f32(self.hd, _p(images), ch, _p(masks), n, int(old)), "rig_process_f32")
        return n

    def process_device_f32(self, d_images, channels, n_frames, d_masks=None, old=False):
        _check(lib().mcs_rig_process_device_f32(self.hd, C.c_void_p(d_images), channels, C.c_void_p(d_masks) if d_masks else None, n_frames,
                                                int(old)), "rig_process_device_f32")

    def process_device(self, d_images, n_frames, d_masks=None, old=False):
        _check(lib().mcs_rig_process_device(self.hd, C.c_void_p(d_images), C.c_void_p(d_masks) if d_masks else None, n_frames, int(old)),
               "mcs_rig_process_device")

    def process_masks_u8(self, images, masks_u8, old=False):
        """Host buffers; masks_u8 [n_frames, C, h, w] uint8: 0 = static pixel (the reference's mask < 0.7), non-zero = dynamic."""
        images = np.ascontiguousarray(images, np.uint8)
        masks_u8 = np.ascontiguousarray(masks_u8, np.uint8)
        assert masks_u8.shape == images.shape
        _check(lib().mcs_rig_process_masks_u8(self.hd, _p(images), _p(masks_u8), images.shape[0], int(old)), "mcs_rig_process_masks_u8")

    def process_device_masks_u8(self, d_images, n_frames, d_masks_u8, old=False):
        _check(lib().mcs_rig_process_device_masks_u8(self.hd, C.c_void_p(d_images), C.c_void_p(d_masks_u8), n_frames, int(old)),
               "mcs_rig_process_device_masks_u8")

    def set_graph(self, on):
        """Latency mode: unchunked device-resident calls are captured into a CUDA graph once and replayed."""
        _check(lib().mcs_rig_set_graph(self.hd, int(on)), "mcs_rig_set_graph")

    @property
    def graph_replays(self):
        return lib().mcs_rig_graph_replays(self.hd)

    @property
    def last_d2h_bytes(self):
        """Bytes the last host-buffer process call copied back (counted from the copies it queued)."""
        return lib().mcs_rig_last_d2h_bytes(self.hd)

    def counts(self, frame):
        kc = np.zeros(self.C, np.int32)
        nr, na = C.c_int(), C.c_int()
        _check(lib().mcs_rig_get_counts(self.hd, frame, _p(kc), C.byref(nr), C.byref(na)), "mcs_rig_get_counts")
        return kc, nr.value, na.value

    def synchronize(self):
        _check(lib().mcs_rig_synchronize(self.hd), "mcs_rig_synchronize")

    @property
    def stream(self):
        return lib().mcs_rig_stream(self.hd)

    @property
    def kernel_launches(self):
        return lib().mcs_rig_kernel_launches(self.hd)

    def set_chunk_frames(self, n):
        _check(lib().mcs_rig_set_chunk_frames(self.hd, int(n)), "mcs_rig_set_chunk_frames")

    def timer_start(self):
        _check(lib().mcs_rig_timer_start(self.hd), "mcs_rig_timer_start")

    def timer_stop(self):
        """Device milliseconds on the rig's stream since timer_start() (waits for the stream)."""
        ms = C.c_float()
        _check(lib().mcs_rig_timer_stop(self.hd, C.byref(ms)), "mcs_rig_timer_stop")
        return ms.value

    def set_profiling(self, on):
        lib().mcs_rig_set_profiling(self.hd, int(on))

    def stage_times(self):
        ms = np.zeros(10, np.float32)
        _check(lib().mcs_rig_stage_times(self.hd, _p(ms)), "mcs_rig_stage_times")
        names = ["upload", "pyramid", "fast", "octree", "blur", "describe", "match", "merge", "triangulate", "download"]
        return dict(zip(names, ms.tolist()))

    def knn2_images(self, q_images, t_images):
        """knnMatch(k=2) between resident feature sets (image index = frame * num_cams + cam): the findMatchesMono primitive
        (FrontEnd.cpp:3134) without descriptor traffic.  int32 [n_sets, kp_cap, 4], rows past the query count are -1."""
        qi = np.ascontiguousarray(q_images, np.int32).ravel()
        ti = np.ascontiguousarray(t_images, np.int32).ravel()
        out = np.empty((len(qi), self.kp_cap, 4), np.int32)
        _check(lib().mcs_rig_knn2_images(self.hd, _p(qi), _p(ti), len(qi), _p(out)), "rig_knn2_images")
        return out

    def bow_best2(self, frame, groups, y_gate=50.0):
        """Inner loop of the BoW-guided computeIntraMatches(matches, words_) (MultiCameraFrame.cpp:716-751) on frame `frame`'s resident
        features.  groups: [(cam_a, cam_b, a_feature_indices, b_feature_indices)].  Returns per group (best_j, d1, d2) int32 arrays
        (best_j = position in the group's b list or -1; 1000000000 = no candidate)."""
        ca = np.array([g[0] for g in groups], np.int32)
        cb = np.array([g[1] for g in groups], np.int32)
        ao = np.zeros(len(groups) + 1, np.int32)
        bo = np.zeros(len(groups) + 1, np.int32)
        for k, g in enumerate(groups):
            ao[k + 1] = ao[k] + len(g[2])
            bo[k + 1] = bo[k] + len(g[3])
        ai = np.ascontiguousarray(np.concatenate([np.asarray(g[2], np.uint32) for g in groups]) if groups else np.zeros(0, np.uint32), np.uint32)
        bi = np.ascontiguousarray(np.concatenate([np.asarray(g[3], np.uint32) for g in groups]) if groups else np.zeros(0, np.uint32), np.uint32)
        bj, d1, d2 = (np.empty(max(int(ao[-1]), 1), np.int32) for _ in range(3))
        _check(lib().mcs_rig_bow_best2(self.hd, frame, len(groups), _p(ca), _p(cb), _p(ao), _p(ai), _p(bo), _p(bi), y_gate, _p(bj), _p(d1), _p(d2)),
               "rig_bow_best2")
        return [(bj[ao[k]:ao[k + 1]].copy(), d1[ao[k]:ao[k + 1]].copy(), d2[ao[k]:ao[k + 1]].copy()) for k in range(len(groups))]

    def level(self, frame, cam, level, kind=0):
        w, h = C.c_int(), C.c_int()
        _check(lib().mcs_rig_level_size(self.hd, level, C.byref(w), C.byref(h)), "rig_level_size")
        pad = 38 if kind == 1 else 0
        out = np.zeros((h.value + pad, w.value + pad), np.uint8)
        _check(lib().mcs_rig_download_level(self.hd, frame, cam, level, kind, _p(out), out.shape[1]), "rig_download_level")
        return out


class Matcher:
    """Batched DescriptorMatcher("BruteForce-Hamming").knnMatch(q, t, k=2) (FrontEnd.cpp:3134-3135, :3199-3201)."""

    def __init__(self, max_sets=1, max_query_rows=8192, max_train_rows=8192, device=0):
        self.device, self.max_sets = device, max_sets
        self.hd = lib().mcs_matcher_create(device, max_sets, max_query_rows, max_train_rows)
        if not self.hd:
            raise McsError("mcs_matcher_create failed: " + lib().mcs_last_error().decode())

    def close(self):
        if getattr(self, "hd", None):
            lib().mcs_matcher_destroy(self.hd)
            self.hd = None

    __del__ = close

    def knn2(self, q_sets, t_sets):
        """q_sets / t_sets: lists of uint8 [n_i, 32] arrays (one problem each).  Returns a list of int32 [nq_i, 4] arrays
        (trainIdx0, dist0, trainIdx1, dist1)."""
        qs = [np.ascontiguousarray(q, np.uint8).reshape(-1, 32) for q in q_sets]
        ts = [np.ascontiguousarray(t, np.uint8).reshape(-1, 32) for t in t_sets]
        qc = np.array([len(q) for q in qs], np.int32)
        tc = np.array([len(t) for t in ts], np.int32)
        q = np.concatenate(qs) if qs else np.zeros((0, 32), np.uint8)
        t = np.concatenate(ts) if ts else np.zeros((0, 32), np.uint8)
        out = np.empty((max(len(q), 1), 4), np.int32)
        n = _check(lib().mcs_knn2(self.hd, _p(q), _p(qc), _p(t), _p(tc), len(qs), _p(out)), "knn2")
        assert n == len(q)
        return np.split(out[:n], np.cumsum(qc)[:-1])

    def match_subsets(self, A, B, groups_a, groups_b, th_low=75, max_ratio=0.85):
        """getMatches_distRatio (ORBextractor.cpp:1228-1290) for the candidate-list pairs zip(groups_a, groups_b) in one launch.
        Returns (i_match_A, i_match_B, per-group counts, BookK), matches concatenated in group order."""
        A = np.ascontiguousarray(A, np.uint8).reshape(-1, 32)
        B = np.ascontiguousarray(B, np.uint8).reshape(-1, 32)
        ao = np.zeros(len(groups_a) + 1, np.int32)
        bo = np.zeros(len(groups_b) + 1, np.int32)
        ao[1:] = np.cumsum([len(g) for g in groups_a])
        bo[1:] = np.cumsum([len(g) for g in groups_b])
        ai = np.array([i for g in groups_a for i in g], np.uint32)
        bi = np.array([i for g in groups_b for i in g], np.uint32)
        ma = np.zeros(max(len(ai), 1), np.uint32)
        mb = np.zeros(max(len(ai), 1), np.uint32)
        gc = np.zeros(max(len(groups_a), 1), np.int32)
        bk = C.c_longlong(0)
        n = _check(lib().mcs_match_subsets(self.hd, _p(A), len(A), _p(B), len(B), len(groups_a), _p(ao), _p(ai), _p(bo), _p(bi), th_low,
                                           max_ratio, _p(ma), _p(mb), _p(gc), C.byref(bk)), "match_subsets")
        return ma[:n], mb[:n], gc[:len(groups_a)], bk.value

    def bow_matching(self, A, B, fv_a, fv_b, th_low=75, max_ratio=0.85):
        """Node walk of BowMatching / InterMatchingBow (MultiCameraFrame.cpp:946-1022, FrontEnd.cpp:3791-3965) over two feature
        vectors {node id: [indices]}: nodes present in both, ascending.  Returns (indices_1, indices_2, words, BookK)."""
        nodes = sorted(set(fv_a) & set(fv_b))
        ma, mb, gc, bk = self.match_subsets(A, B, [fv_a[n] for n in nodes], [fv_b[n] for n in nodes], th_low, max_ratio)
        return ma, mb, np.array([n for n, c in zip(nodes, gc) if c >= 1], np.uint32), bk

    def knn2_device(self, d_q, q_counts, d_t, t_counts, d_out):
        qc = np.ascontiguousarray(q_counts, np.int32)
        tc = np.ascontiguousarray(t_counts, np.int32)
        return _check(lib().mcs_knn2_device(self.hd, C.c_void_p(d_q), _p(qc), C.c_void_p(d_t), _p(tc), len(qc), C.c_void_p(d_out)), "knn2_device")

    def synchronize(self):
        _check(lib().mcs_matcher_synchronize(self.hd), "matcher_synchronize")

    @property
    def kernel_launches(self):
        return int(lib().mcs_matcher_kernel_launches(self.hd))


class Vocabulary:
    """DBoW2 vocabulary (loadFromTextFile format, FrontEnd.h:137-138) and its transform (MultiCameraFrame.cpp:257)."""

    def __init__(self, path=None, device=0, k=10, L=None, parent=None, is_leaf=None, node_desc=None, weight=None):
        if path is not None:
            self.hd = lib().mcs_vocab_load_text(str(path).encode(), device)
        else:
            parent = np.ascontiguousarray(parent, np.int32)
            is_leaf = np.ascontiguousarray(is_leaf, np.int32)
            node_desc = np.ascontiguousarray(node_desc, np.uint8).reshape(-1, 32)
            weight = np.ascontiguousarray(weight, np.float64)
            self.hd = lib().mcs_vocab_create(device, k, L, 0, 0, len(parent), _p(parent), _p(is_leaf), _p(node_desc), _p(weight))
        if not self.hd:
            raise McsError("vocabulary: " + lib().mcs_last_error().decode())
        v = [C.c_int() for _ in range(4)]
        _check(lib().mcs_vocab_info(self.hd, *[C.byref(x) for x in v]), "vocab_info")
        self.k, self.L, self.n_nodes, self.n_words = [x.value for x in v]

    def close(self):
        if getattr(self, "hd", None):
            lib().mcs_vocab_destroy(self.hd)
            self.hd = None

    __del__ = close

    def transform(self, feats, levelsup=4):
        feats = np.ascontiguousarray(feats, np.uint8).reshape(-1, 32)
        n = len(feats)
        wid, nid, w = np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.float64)
        _check(lib().mcs_vocab_transform(self.hd, _p(feats), n, levelsup, _p(wid), _p(nid), _p(w)), "vocab_transform")
        return wid[:n], nid[:n], w[:n]

    def transform_rig(self, rig, levelsup=4):
        """Every image of the rig's last batch on the resident descriptors: arrays [n_images, kp_cap]."""
        n_img = rig.C * rig.F
        shape = (n_img, rig.kp_cap)
        wid, nid, w = np.full(shape, -1, np.int32), np.full(shape, -1, np.int32), np.zeros(shape, np.float64)
        _check(lib().mcs_rig_vocab_transform(rig.hd, self.hd, levelsup, _p(wid), _p(nid), _p(w)), "rig_vocab_transform")
        return wid, nid, w


def knn2_device_static(device, d_q, q_counts, d_t, t_counts, d_out, stream=None):
    """mcs_knn2_device_static: the knn2 kernel with its set table in a static __device__ array; allocates nothing."""
    qc = np.ascontiguousarray(q_counts, np.int32)
    tc = np.ascontiguousarray(t_counts, np.int32)
    return _check(lib().mcs_knn2_device_static(device, C.c_void_p(stream), C.c_void_p(d_q), _p(qc), C.c_void_p(d_t), _p(tc), len(qc),
                                               C.c_void_p(d_out)), "knn2_device_static")


def device_malloc(device, nbytes):
    p = lib().mcs_device_malloc(device, nbytes)
    if not p:
        raise McsError("device malloc failed: " + lib().mcs_last_error().decode())
    return p


def device_free(device, p):
    lib().mcs_device_free(device, C.c_void_p(p))


def device_upload(device, dptr, arr):
    arr = np.ascontiguousarray(arr)
    _check(lib().mcs_device_upload(device, C.c_void_p(dptr), _p(arr), arr.nbytes), "device_upload")


def device_download(device, arr, dptr):
    """Copies arr.nbytes from device pointer dptr into the (contiguous) numpy array arr."""
    assert arr.flags["C_CONTIGUOUS"]
    _check(lib().mcs_device_download(device, _p(arr), C.c_void_p(dptr), arr.nbytes), "device_download")


def plan_fast_tables(width, height, nfeatures=1000, scale_factor=1.2, nlevels=8, capacity=65536):
    """Host-only (no GPU): the FAST cell table and strip table of one configuration -> (cells [n, 4] u32, strips [n, 8] u32, smem bytes)."""
    L = lib()
    L.mcs_plan_fast_tables.restype = C.c_int
    L.mcs_plan_fast_tables.argtypes = [C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, C.c_void_p, C.POINTER(C.c_int), C.c_void_p,
                                       C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_int]
    cells = np.zeros((capacity, 4), np.uint32)
    strips = np.zeros((capacity, 8), np.uint32)
    nc, ns, sm = C.c_int(), C.c_int(), C.c_int()
    _check(L.mcs_plan_fast_tables(width, height, nfeatures, scale_factor, nlevels, cells.ctypes.data, C.byref(nc), strips.ctypes.data,
                                  C.byref(ns), C.byref(sm), capacity), "mcs_plan_fast_tables")
    return cells[:nc.value].copy(), strips[:ns.value].copy(), sm.value
