"""CPU: oracle/_ref (the REFERENCE's own ORBextractor.cpp compiled whole, and its MultiCameraFrame / FrontEnd method bodies
extracted by line range — oracle/ref_build.py) against the committed cv2 goldens and against the portable restatement
(oracle/liboracle.so) that the GPU parity tests use.  _ref == liboracle == goldens pins the checker to the reference itself.
Integer outputs bit-exact; f64 points 1e-9 relative (both sides run the same one-sided Jacobi, so they are normally identical)."""
import numpy as np
import pytest

import oracle_lib as O
import ref_lib as R

pytestmark = pytest.mark.skipif(not R.available(), reason="oracle/_ref/libmcsref.so missing and /root/reference not present")

SINGLES = [("single_256x192.npz", 300), ("noise_203x157.npz", 400), ("lowcontrast_320x240.npz", 500)]


@pytest.mark.parametrize("name,nf", SINGLES)
def test_ref_extractor_matches_golden_and_oracle(golden, name, nf):
    g = golden(name)
    e, o = R.Extractor(nf), O.Extractor(nf)
    mono, kps, desc = e(g["image"])
    for a, b in ((e.quota, g["quota"]), (e.scale, g["scale"]), (e.sigma2, g["sigma2"]), (e.inv_scale, g["inv_scale"]), (e.inv_sigma2, g["inv_sigma2"]),
                 (e.umax, o.umax)):
        assert np.array_equal(a, b)
    for l in range(8):
        assert np.array_equal(e.level(l), g["pyr%d" % l]), "mvImagePyramid[%d]" % l
    assert np.array_equal(e.level_padded(3), g["pyr_padded3"])
    assert mono == int(g["mono"]) and len(kps) == len(g["kps"])
    assert (kps == g["kps"]).all(), "keypoints (x, y, size, angle, response, octave, class_id)"
    assert np.array_equal(desc, g["desc"]), "descriptor bits"
    mo, ko, do = o(g["image"])
    assert mo == mono and (ko == kps).all() and np.array_equal(do, desc)
    # vLappingArea: stereo keypoints filled from the back (ORBextractor.cpp:1153-1163)
    m1, k1, d1 = e(g["image"], lapping=(60, 140))
    m2, k2, d2 = o(g["image"], lapping=(60, 140))
    assert m1 == m2 and 0 < m1 < len(k1) and (k1 == k2).all() and np.array_equal(d1, d2)


SWEEP = [(411, 233, 300, 1.5, 5, 20, 7), (640, 480, 800, 2.0, 4, 30, 10), (257, 190, 200, 1.1, 10, 15, 5), (512, 384, 600, 1.3, 6, 40, 40),
         (300, 300, 150, 1.2, 3, 10, 3), (720, 540, 2000, 1.2, 8, 12, 4), (1000, 160, 400, 1.25, 4, 25, 9),
         (640, 480, 2900, 1.2, 1, 12, 4), (500, 400, 4500, 2.0, 2, 10, 5)]


@pytest.mark.parametrize("w,h,nf,sf,nl,ini,mn", SWEEP)
def test_ref_extractor_equals_oracle_over_constructor_parameters(w, h, nf, sf, nl, ini, mn):
    """The same ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST) sweep the GPU test runs
    (tests/test_gpu_extract.py::test_constructor_parameter_sweep): the reference's compiled code and the restatement agree on tables,
    pyramid levels, keypoints and descriptors."""
    import mcslam_b200 as M
    calib = M.synth.make_rig(2, w, h)
    base = M.synth.make_rig_frame(calib, w, h, frame_index=w + h, n_rects=7)[1].astype(np.int32)
    noise = M.synth.noise_image(w, h, 3 * w + nf).astype(np.int32)
    img = np.clip((3 * base + noise) // 4, 0, 255).astype(np.uint8)
    e, o = R.Extractor(nf, sf, nl, ini, mn), O.Extractor(nf, sf, nl, ini, mn)
    mono, kps, desc = e(img)[:3]
    mo, ko, do = o(img)
    assert np.array_equal(e.quota, o.quota) and np.array_equal(e.scale, o.scale) and np.array_equal(e.sigma2, o.sigma2)
    for l in range(nl):
        assert np.array_equal(e.level(l), o.level(l)), l
    assert mono == mo and len(kps) == len(ko) and (kps == ko).all() and np.array_equal(desc, do)


@pytest.mark.parametrize("w,h,nf,seed", [(203, 157, 400, 1), (97, 83, 150, 2), (333, 130, 700, 3), (200, 260, 500, 4), (640, 480, 1000, 5)])
def test_ref_extractor_equals_oracle_on_odd_sizes(w, h, nf, seed):
    """Single-cell levels, levels too small for a cell (nCols = 0), wide images with 3 octree roots, mild portrait, noise worst
    case.  (A portrait image with W/H < 0.5 makes nIni = round(W/H) = 0 and the reference itself dereferences vpIniNodes[0] of an
    empty vector, ORBextractor.cpp:558-586 — undefined there; the restatement's choice is listed in BASELINE.md.)"""
    import mcslam_b200 as M
    rng = np.random.default_rng(seed)
    img = M.synth.noise_image(w, h, seed) if seed % 2 else np.clip(rng.normal(120, 40, (h, w)), 0, 255).astype(np.uint8)
    e, o = R.Extractor(nf), O.Extractor(nf)
    m1, k1, d1 = e(img)
    m2, k2, d2 = o(img)
    assert m1 == m2 and len(k1) == len(k2) and len(k1) > 0
    assert (k1 == k2).all() and np.array_equal(d1, d2)
    for l in range(8):
        a, b = e.level(l), o.level(l)
        assert np.array_equal(a, b)
    # a strided view (ROI of a wider buffer) gives the same result as the tight copy
    wide = np.zeros((h, w + 13), np.uint8)
    wide[:, 5:5 + w] = img
    m3, k3, d3 = e(wide[:, 5:5 + w])
    assert m3 == m1 and (k3 == k1).all() and np.array_equal(d3, d1)


def test_ref_extractor_edge_cases():
    e = R.Extractor(100)
    assert e(None)[0] == -1  # empty image (ORBextractor.cpp:1090-1091)
    mono, k, d = e(np.full((120, 160), 77, np.uint8))
    assert mono == 0 and len(k) == 0
    # fast threshold fallback: a low-contrast image only yields corners at minThFAST
    rng = np.random.default_rng(3)
    low = (100 + rng.integers(0, 12, (200, 260))).astype(np.uint8)
    for ini, mn in ((20, 7), (20, 2)):
        m1, k1, d1 = R.Extractor(300, ini_th=ini, min_th=mn)(low)
        m2, k2, d2 = O.Extractor(300, ini_th=ini, min_th=mn)(low)
        assert m1 == m2 and (k1 == k2).all() and np.array_equal(d1, d2)
    assert len(k1) > 0


@pytest.mark.parametrize("n,N,seed", [(50, 20, 0), (700, 217, 1), (3000, 217, 2), (3000, 1000, 3), (40, 100, 4), (5000, 60, 5)])
def test_ref_octree_equals_oracle(n, N, seed):
    """DistributeOctTree alone (ORBextractor.cpp:554-778) on random candidates with many response ties and duplicate positions:
    the retained keys and their ORDER (std::list push_front order + std::sort tie order) are identical."""
    rng = np.random.default_rng(seed)
    W, H = 688, 508
    xy = np.stack([rng.integers(0, W, n), rng.integers(0, H, n)], 1).astype(np.float32)
    resp = rng.integers(7, 40, n).astype(np.float32)
    xyr = np.concatenate([xy, resp[:, None]], 1)
    ref = R.Extractor(1000).distribute_octtree(xyr, 16, 16 + W, 16, 16 + H, N)
    keep = O.distribute_octtree(xyr, 16, 16 + W, 16, 16 + H, N)
    assert np.array_equal(ref, xyr[keep])


def test_ref_octree_wide_image_two_roots():
    """nIni = round(W / H) = 2 root nodes (ORBextractor.cpp:558-575)."""
    rng = np.random.default_rng(9)
    n, W, H = 1500, 900, 420
    xyr = np.stack([rng.integers(0, W, n), rng.integers(0, H, n), rng.integers(7, 60, n)], 1).astype(np.float32)
    ref = R.Extractor(1000).distribute_octtree(xyr, 16, 16 + W, 16, 16 + H, 300)
    keep = O.distribute_octtree(xyr, 16, 16 + W, 16, 16 + H, 300)
    assert np.array_equal(ref, xyr[keep])


def _lf_from_oracle(K, D, tri, total_feats=3000):
    """currentFrame->intraMatches as the oracle pieces predict it: accepted rows in order, then the mono top-up."""
    acc = tri["status"] == 1
    accepted, mono = O.obtain_lf_features(K, tri, total_feats)
    return acc, accepted, mono


def _check_lf(lf, K, D, tri, nc):
    acc, accepted, mono = _lf_from_oracle(K, D, tri)
    n = lf["intramatch_size"]
    assert n == int(acc.sum()) == tri["accepted"] and lf["mono_size"] == len(mono) and len(lf["xyz"]) == n + len(mono)
    assert np.array_equal(lf["match_index"][:n], tri["rows"][acc]), "accepted rows after the mask filter"
    assert np.array_equal(lf["n_rays"][:n], tri["nviews"][acc]) and not lf["mono"][:n].any()
    assert np.allclose(lf["xyz"][:n], tri["xyz"][acc], rtol=1e-9, atol=1e-12)
    assert np.allclose(lf["uv_ref"][:n], tri["uv_ref"][acc], rtol=1e-6, atol=1e-4)
    rep = np.stack([D[c][r[c]] for c, r in zip(tri["rep_cam"][acc], tri["rows"][acc])]) if n else np.zeros((0, 32), np.uint8)
    assert np.array_equal(lf["desc"][:n], rep), "computeRepresentativeDesc choice"
    got = [(int(np.flatnonzero(r >= 0)[0]), int(r[r >= 0][0])) for r in lf["match_index"][n:]]
    assert got == mono, "mono top-up: argsorte order (unstable std::sort) and the 3000 - intramatch_size cut"
    assert lf["mono"][n:].all() and (lf["n_rays"][n:] == 1).all() and np.isnan(lf["xyz"][n:]).all()
    for (c, j), uv, d in zip(mono, lf["uv_ref"][n:], lf["desc"][n:]):
        assert uv[0] == K[c][j]["x"] and uv[1] == K[c][j]["y"] and np.array_equal(d, D[c][j])


def test_ref_rig_golden(golden):
    """extractFeaturesParallel -> BruteForceMatch -> computeIntraMatches(old = false / true) -> obtainLfFeatures, reference code,
    against the cv2 golden rig and the restatement."""
    g = golden("rig3_480x360.npz")
    calib = g["calib"]
    rig = R.Rig(calib, 480, 360, nfeatures=600)
    rig.extract(g["images"])
    K, D = [], []
    for c in range(3):
        k, u, d = rig.features(c)
        assert (k == g["kps%d" % c]).all() and np.array_equal(d, g["desc%d" % c])
        assert (u == k).all(), "zero distortion: image_kps_undist is a copy (MultiCameraFrame.cpp:302-305)"
        K.append(k)
        D.append(d)
    p = 0
    for i in range(2):
        for j in range(i + 1, 3):
            i1, i2, k1, k2 = rig.bf_match(i, j)
            assert np.array_equal(i1, g["pair%d_i1" % p]) and np.array_equal(i2, g["pair%d_i2" % p]), "pair %d" % p
            assert (k1 == K[i][i1]).all() and (k2 == K[j][i2]).all()
            o1, o2, *_ = O.bf_match(D[i], D[j])
            assert np.array_equal(i1, o1) and np.array_equal(i2, o2)
            p += 1
    rows_old = rig.intra_matches(True)
    ro, _ = O.intra_matches(D, K, calib, g["sigma2"], use_epipolar=True)
    assert np.array_equal(rows_old, ro) and 0 < len(ro) < len(g["rows"]), "epipolar gate (old = true) must remove some matches"
    rows = rig.intra_matches(False)
    assert np.array_equal(rows, g["rows"])
    tri = O.triangulate_rows(calib, K, D, rows)
    _check_lf(rig.obtain_lf_features(), K, D, tri, 3)
    ok = tri["nviews"] >= 2
    assert np.allclose(tri["xyz"][ok], g["xyz"][ok], rtol=1e-6, atol=1e-9)


def test_ref_rig5_masks_and_budget():
    """5 cameras with semantic masks: `segMasks[i].at<float>(p.y, p.x) < 0.7` (FrontEnd.cpp:265-273) removes views, rows drop to mono
    or disappear; 2000 features per camera make the mono top-up hit the 3000 budget (FrontEnd.cpp:513-521)."""
    import mcslam_b200 as M
    calib = M.synth.make_rig(5, 480, 360)
    frame = M.synth.make_rig_frame(calib, 480, 360, frame_index=11, n_rects=10)
    masks = M.synth.make_masks(5, 480, 360, frame_index=11)
    rig = R.Rig(calib, 480, 360, nfeatures=2000)
    rig.extract(frame)
    K, D = [], []
    ex = O.Extractor(2000)
    for c in range(5):
        k, u, d = rig.features(c)
        _, k2, d2 = ex(frame[c])
        assert (k == k2).all() and np.array_equal(d, d2)
        K.append(k)
        D.append(d)
    rows = rig.intra_matches(False)
    ro, _ = O.intra_matches(D, K, calib, ex.sigma2)
    assert np.array_equal(rows, ro) and len(rows) > 300
    tri = O.triangulate_rows(calib, K, D, rows, masks=list(masks))
    assert (tri["rows"] != rows).any()
    lf = rig.obtain_lf_features(masks=list(masks))
    _check_lf(lf, K, D, tri, 5)
    assert len(lf["xyz"]) == 3000, "budget cut reached"


def test_ref_many_to_one_merge_stress():
    """Random descriptors built so that many query features share a train feature: exercises every branch of the sequential row
    merge (MultiCameraFrame.cpp:1227-1266) in the reference's own code against the restatement."""
    rng = np.random.default_rng(21)
    import mcslam_b200 as M
    nc, n = 5, 300
    calib = M.synth.make_rig(nc, 480, 360)
    base = rng.integers(0, 256, (60, 32), dtype=np.uint8)
    D, K = [], []
    for c in range(nc):
        d = base[rng.integers(0, 60, n)].copy()
        flip = rng.integers(0, 256, (n, 32), dtype=np.uint8) & rng.integers(0, 256, (n, 32), dtype=np.uint8) & rng.integers(0, 256, (n, 32), dtype=np.uint8) \
            & rng.integers(0, 256, (n, 32), dtype=np.uint8)
        d ^= flip
        k = np.zeros(n, O.KP_DTYPE)
        k["x"], k["y"] = rng.uniform(20, 460, n), rng.uniform(20, 340, n)
        k["octave"] = rng.integers(0, 8, n)
        k["size"], k["response"], k["class_id"] = 31, rng.integers(7, 90, n), -1
        D.append(d)
        K.append(k)
    rig = R.Rig(calib, 480, 360, nfeatures=n)
    for c in range(nc):
        rig.set_features(c, K[c], D[c])
    for old in (False, True):
        rows = rig.intra_matches(old)
        ro, _ = O.intra_matches(D, K, calib, O.Extractor(n).sigma2, use_epipolar=old)
        assert np.array_equal(rows, ro)
    assert len(ro) > 0 or True
    rows = rig.intra_matches(False)
    assert len(rows) > 100 and (np.sum(rows >= 0, 1) > 2).any()


def test_ref_known_answer_and_dlt():
    """testTriangulateIntraMatches (MultiCameraFrame.cpp:383-409): X = (0, 0, 6) on the 0.165 m rig through the reference's own
    triangulateIntraMatches; N-view and 2-view DLT against the restatement on random points."""
    import mcslam_b200 as M
    calib = M.synth.make_rig(5, 720, 540, spacing=0.165)
    rig = R.Rig(calib, 720, 540)
    X = np.array([0, 0, 6.0, 1.0])
    proj = []
    for c in range(5):
        P = calib[c]["K"].reshape(3, 3) @ np.hstack([calib[c]["R"].reshape(3, 3), calib[c]["t"].reshape(3, 1)])
        proj.append(P)
    xs = [(P @ X)[0] / (P @ X)[2] for P in proj]
    ys = [(P @ X)[1] / (P @ X)[2] for P in proj]
    for views in ([0, 1], [0, 2, 4], [0, 1, 2, 3, 4]):
        got = rig.triangulate_nview(views, [xs[v] for v in views], [ys[v] for v in views])
        assert np.allclose(got, [0, 0, 6], atol=1e-8), (views, got)
        assert np.allclose(got, O.triangulate_pixels(calib, views, [xs[v] for v in views], [ys[v] for v in views]), rtol=1e-9, atol=1e-12)
    rng = np.random.default_rng(5)
    P3 = np.stack([rng.uniform(-3, 3, 200), rng.uniform(-2, 2, 200), rng.uniform(2, 30, 200), np.ones(200)], 1)
    pts = [((P3 @ proj[c].T)[:, :2] / (P3 @ proj[c].T)[:, 2:]).astype(np.float32) for c in (0, 4)]
    got = rig.triangulate_pair(0, 4, pts[0], pts[1])
    ref = O.triangulate_pair_f32(calib, 0, 4, pts[0], pts[1])
    rel = np.abs(got - ref).max(1) / np.abs(ref).max(1)
    assert rel.max() < 1e-3, rel.max()  # f32 4x4 SVD (1e-3 relative)
    assert np.median(np.abs(got - P3[:, :3]).max(1) / np.abs(P3[:, :3]).max(1)) < 2e-2


def test_ref_representative_desc_and_distance():
    rng = np.random.default_rng(8)
    e = R.Extractor(100)
    calib = np.zeros(2, O.CALIB_DTYPE)
    calib["K"] = np.eye(3).reshape(9)
    calib["R"] = np.eye(3).reshape(9)
    rig = R.Rig(calib, 64, 64)
    for n in (2, 3, 4, 5):
        for _ in range(20):
            d = rng.integers(0, 256, (n, 32), dtype=np.uint8)
            d[rng.integers(0, n)] = d[0]  # ties
            dist = np.array([[O.lib().orc_hamming256(O._p(np.ascontiguousarray(a)), O._p(np.ascontiguousarray(b))) for b in d] for a in d])
            assert all(e.descriptor_distance(d[i], d[j]) == dist[i, j] for i in range(n) for j in range(n))
            med = [sorted(dist[i])[int(0.5 * (n - 1))] for i in range(n)]
            assert np.array_equal(rig.representative_desc(d), d[int(np.argmin(med))])


def test_ref_subset_matcher(golden):
    """getMatches_distRatio (ORBextractor.cpp:1228-1290) of the reference against the restatement on the BoW fixture's node groups."""
    g = golden("bow_subsets.npz")
    e = R.Extractor(100)
    fa, fb = {}, {}
    for i, nd in enumerate(g["node_of_A"]):
        fa.setdefault(int(nd), []).append(i)
    for i, nd in enumerate(g["node_of_B"]):
        fb.setdefault(int(nd), []).append(i)
    checked = 0
    for nd in sorted(set(fa) & set(fb)):
        a = e.matches_dist_ratio(g["A"], fa[nd], g["B"], fb[nd])
        b = O.matches_dist_ratio(g["A"], fa[nd], g["B"], fb[nd])
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and a[2] == b[2]
        checked += len(a[0])
    assert checked > 50
    one = e.matches_dist_ratio(g["A"], [0, 1, 2], g["B"], [5])  # single candidate: second best stays 1e9
    assert np.array_equal(one[0], O.matches_dist_ratio(g["A"], [0, 1, 2], g["B"], [5])[0])


def test_ref_undistort_keypoints_branch(golden):
    """UndistortKeyPoints (MultiCameraFrame.cpp:300-347) in the reference's own code: cv::undistortPoints on CV_32F-narrowed K / dist
    (cv2 golden), and the float read of the f64 coefficient that decides the copy branch (:302)."""
    g = golden("undistort_points_cv2.npz")
    pts = g["pts"]
    kps = np.zeros(len(pts), O.KP_DTYPE)
    kps["x"], kps["y"] = pts[:, 0], pts[:, 1]
    for K, dist, want in ((g["K1"], g["d1"], g["out1"]), (g["K2"], g["d2"], g["out2"])):
        calib = np.zeros(1, O.CALIB_DTYPE)
        calib["K"][0] = K.reshape(9)
        calib["R"][0] = np.eye(3).reshape(9)
        calib["dist"][0] = dist
        rig = R.Rig(calib, 640, 480, rectify=False)
        rig.set_features(0, kps, np.zeros((len(pts), 32), np.uint8))
        _, und, _ = rig.features(0)
        assert np.array_equal(np.stack([und["x"], und["y"]], 1), want)
        assert (und == O.undistort_keypoints(kps, K, dist)).all()
    # a first coefficient whose low 32 bits are zero reads as 0.0f => keypoints copied although the lens is distorted
    calib["dist"][0][0] = 0.5
    rig2 = R.Rig(calib, 640, 480, rectify=False)
    rig2.set_features(0, kps, np.zeros((len(pts), 32), np.uint8))
    assert (rig2.features(0)[1] == kps).all()
    assert (O.undistort_keypoints(kps, K, calib["dist"][0]) == kps).all()
    # RECTIFY rigs: image_kps_undist = image_kps without looking at the coefficients (:241-243)
    calib["dist"][0] = g["d1"]
    rig3 = R.Rig(calib, 640, 480, rectify=True)
    rig3.set_features(0, kps, np.zeros((len(pts), 32), np.uint8))
    assert (rig3.features(0)[1] == kps).all()


def test_ref_process_equals_stagewise():
    import mcslam_b200 as M
    calib = M.synth.make_rig(3, 320, 240)
    frame = M.synth.make_rig_frame(calib, 320, 240, frame_index=0, n_rects=6)
    rig = R.Rig(calib, 320, 240, nfeatures=400)
    M1, counts, acc = rig.process(frame)
    ex = [O.Extractor(400) for _ in range(3)]
    M2, c2, a2 = O.process_rig(ex, frame, calib)
    assert M1 == M2 and (counts == c2).all() and acc == a2
