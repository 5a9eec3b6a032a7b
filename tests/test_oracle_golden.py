"""CPU: the portable C++ oracle (oracle/mcs_oracle.cpp) against the committed cv2-generated golden vectors
(tests/golden/*.npz, made by oracle/gen_golden.py with OpenCV 4.13.0).  Integer outputs must be bit-exact."""
import numpy as np
import pytest

import oracle_lib as O

SINGLES = [("single_256x192.npz", 300), ("noise_203x157.npz", 400), ("lowcontrast_320x240.npz", 500)]


@pytest.mark.parametrize("name,nf", SINGLES)
def test_extractor_matches_golden(golden, name, nf):
    g = golden(name)
    e = O.Extractor(nf)
    mono, kps, desc = e(g["image"])
    assert (e.quota == g["quota"]).all() and (e.scale == g["scale"]).all() and (e.sigma2 == g["sigma2"]).all()
    assert (e.inv_scale == g["inv_scale"]).all() and (e.inv_sigma2 == g["inv_sigma2"]).all()
    for l in range(8):
        assert np.array_equal(e.level(l), g["pyr%d" % l]), "pyramid level %d" % l
        assert np.array_equal(e.candidates(l), g["cand%d" % l].reshape(-1, 3)), "FAST candidates level %d" % l
        if "blur%d" % l in g.files:
            assert np.array_equal(e.level(l, blurred=True), g["blur%d" % l]), "blurred level %d" % l
    assert np.array_equal(e.level_padded(3), g["pyr_padded3"])
    assert np.array_equal(e.level_counts(), g["level_counts"])
    assert mono == int(g["mono"]) and len(kps) == len(g["kps"])
    assert (kps == g["kps"]).all(), "keypoints (x, y, size, angle, response, octave, class_id)"
    assert np.array_equal(desc, g["desc"]), "descriptor bits"


def test_rig_golden(golden):
    g = golden("rig3_480x360.npz")
    K, D = [], []
    for c in range(3):
        _, k, d = O.Extractor(600)(g["images"][c])
        assert (k == g["kps%d" % c]).all() and np.array_equal(d, g["desc%d" % c])
        K.append(k)
        D.append(d)
    p = 0
    for i in range(2):
        for j in range(i + 1, 3):
            i1, i2, *_ = O.bf_match(D[i], D[j])
            assert np.array_equal(i1, g["pair%d_i1" % p]) and np.array_equal(i2, g["pair%d_i2" % p]), "pair %d" % p
            p += 1
    rows, pc = O.intra_matches(D, K, g["calib"], g["sigma2"])
    assert np.array_equal(rows, g["rows"])
    tri = O.triangulate_rows(g["calib"], K, D, rows)
    ok = tri["nviews"] >= 2
    # numpy.linalg.svd (golden) vs one-sided Jacobi (oracle): 1e-6 relative
    assert np.allclose(tri["xyz"][ok], g["xyz"][ok], rtol=1e-6, atol=1e-9)
    assert tri["accepted"] == int(((g["xyz"][:, 2] < 40) & (g["xyz"][:, 2] > 0.5) & ok).sum())


def test_fast_atan2_golden(golden):
    g = golden("fast_atan2.npz")
    got = np.array([O.lib().orc_fast_atan2(float(y), float(x)) for y, x in g["yx"]], np.float32)
    assert np.array_equal(got, g["ang"])


def test_known_answer_triangulation():
    """testTriangulateIntraMatches (MultiCameraFrame.cpp:383-409): X = (0, 0, 6) seen by a 0.165 m rig triangulates back."""
    import mcslam_b200
    calib = mcslam_b200.synth.make_rig(5, 720, 540, spacing=0.165)
    X = np.array([0, 0, 6.0, 1.0])
    xs, ys = [], []
    for c in range(5):
        K = calib[c]["K"].reshape(3, 3)
        Rt = np.hstack([calib[c]["R"].reshape(3, 3), calib[c]["t"].reshape(3, 1)])
        p = K @ Rt @ X
        xs.append(p[0] / p[2])
        ys.append(p[1] / p[2])
    for views in ([0, 1], [0, 2, 4], [0, 1, 2, 3, 4]):
        got = O.triangulate_pixels(calib, views, [xs[v] for v in views], [ys[v] for v in views])
        assert np.allclose(got, [0, 0, 6], atol=1e-8), (views, got)
    got32 = O.triangulate_pair_f32(calib, 0, 4, [[xs[0], ys[0]]], [[xs[4], ys[4]]])
    assert np.allclose(got32[0], [0, 0, 6], atol=2e-2)


def test_nview_dlt_matches_cv2_svd(golden):
    """N-view DLT (MultiCameraFrame.cpp:411-441) against OpenCV's own SVD: 200 noisy points over 2..5 views from oracle/gen_golden_dlt.py
    (cv2.SVDecomp of the (3n) x (4+n) design).  1e-10 relative; the product's bar is 1e-6."""
    g = golden("dlt_cv2.npz")
    worst = 0.0
    for v, xs, ys, want in zip(g["views"], g["xs"], g["ys"], g["xyz"]):
        n = int((v >= 0).sum())
        got = O.triangulate_pixels(g["calib"], v[:n].tolist(), xs[:n].tolist(), ys[:n].tolist())
        worst = max(worst, float(np.abs(got - want).max() / np.abs(want).max()))
    assert worst < 1e-10, worst
    got = O.triangulate_pair_f32(g["calib"], 0, 4, g["pair_p1"], g["pair_p2"])  # MultiCameraFrame::triangulate (:481-507), f32 SVD
    rel = np.abs(got - g["pair_xyz"]).max(1) / np.abs(g["pair_xyz"]).max(1)
    assert rel.max() < 1e-5, rel.max()


def test_edge_cases():
    e = O.Extractor(100)
    mono, k, d = e(np.zeros((0, 0), np.uint8).reshape(0, 0)) if False else (-1, [], [])
    assert mono == -1
    flat = np.full((120, 160), 77, np.uint8)  # no corners anywhere
    mono, k, d = e(flat)
    assert mono == 0 and len(k) == 0
    tiny = np.random.default_rng(0).integers(0, 256, (50, 60), dtype=np.uint8)  # smaller than one FAST cell + borders
    mono, k, d = e(tiny)
    assert mono == 0
    # brute-force matcher: fewer than 2 train descriptors -> no matches (reference reads m[1] out of range)
    q = np.random.default_rng(1).integers(0, 256, (5, 32), dtype=np.uint8)
    i1, i2, *_ = O.bf_match(q, q[:1])
    assert len(i1) == 0
    # exact duplicates: best distance 0, ties resolved to the lower train index
    t = np.concatenate([q, q])
    i1, i2, bd, sd, bi = O.bf_match(q, t)
    assert (bd == 0).all() and (bi == np.arange(5)).all() and (sd == 0).all() and len(i1) == 0


def test_lapping_area(golden):
    g = golden("single_256x192.npz")
    e = O.Extractor(300)
    mono0, k0, d0 = e(g["image"])
    mono, k, d = e(g["image"], lapping=(60, 140))
    st = (k0["x"] >= 60) & (k0["x"] <= 140)
    assert mono == int((~st).sum()) and len(k) == len(k0)
    assert (k[:mono] == k0[~st]).all() and (k[mono:] == k0[st][::-1]).all()
    assert np.array_equal(d[:mono], d0[~st]) and np.array_equal(d[mono:], d0[st][::-1])


def test_knn2_matches_cv2_golden(golden):
    """oracle knn2 == cv2.BFMatcher(NORM_HAMMING).knnMatch(k=2) incl. tie order (64 tied rows in the fixture)."""
    import oracle_lib as O
    g = golden("knn2_cv2.npz")
    assert np.array_equal(O.knn2(g["q"], g["t"]), g["knn"])
    one = O.knn2(g["q"][:5], g["t"][:1])
    assert (one[:, 0] == 0).all() and (one[:, 2:] == -1).all()
    assert (O.knn2(g["q"][:5], g["t"][:0]) == -1).all() and O.knn2(g["q"][:0], g["t"]).shape == (0, 4)


def test_inter_frame_host_logic_matches_cv2_restatement(golden):
    """oracle findMatchesMono / findInterMatches (both overloads) == the independent python + cv2 restatement the fixture was
    generated with (oracle/gen_golden_knn2.py)."""
    import oracle_lib as O
    g = golden("inter_frame_cv2.npz")
    assert np.array_equal(O.find_matches_mono(g["d1"], g["d2"], g["xy1"], g["xy2"], g["m1"], g["m2"], g["m1"].shape[1]), g["mono_out"])
    args = (g["d1"], g["d2"], g["mono1"], g["mono2"], g["p1"], g["p2"], g["lids1"], int(g["isz"][0]), int(g["isz"][1]))
    lf, mo = O.find_inter_matches(*args, split_mono=True)
    assert np.array_equal(lf, g["lf_split"]) and np.array_equal(mo, g["mono_split"]) and len(lf) > 50 and len(mo) > 50
    lf, mo = O.find_inter_matches(*args, split_mono=False)
    assert np.array_equal(lf, g["lf_single"]) and len(mo) == 0
    assert len(set(lf[:, 1].tolist())) == len(lf), "train indices are unique after the one-to-one step"


def _bow_fvs(g):
    fa, fb = {}, {}
    for i, n in enumerate(g["node_of_A"]):
        fa.setdefault(int(n), []).append(i)
    for i, n in enumerate(g["node_of_B"]):
        fb.setdefault(int(n), []).append(i)
    return fa, fb


def test_bow_subset_matcher_matches_python_restatement(golden):
    """oracle getMatches_distRatio + BowMatching node walk == the numpy restatement the fixture was generated with
    (oracle/gen_golden_knn2.py): repeated queries claim the same B feature, node ids missing on one side, BookK."""
    import oracle_lib as O
    g = golden("bow_subsets.npz")
    fa, fb = _bow_fvs(g)
    a, b, w, k = O.bow_matching(g["A"], g["B"], fa, fb)
    assert np.array_equal(a, g["indices_1"]) and np.array_equal(b, g["indices_2"]) and np.array_equal(w, g["words"]) and k == int(g["book_k"][0])
    assert len(set(fa) - set(fb)) > 0 and len(a) > 300
    # single candidate in B: second-best stays 1e9, so the ratio test passes (ORBextractor.cpp:1247-1265)
    ma, mb, bk = O.matches_dist_ratio(g["A"], [0, 1], g["A"], [0])
    assert ma.tolist() == [0] and mb.tolist() == [0] and bk == 2
    ma, mb, bk = O.matches_dist_ratio(g["A"], [0, 700], g["B"], [])
    assert len(ma) == 0 and bk == 0


def test_set_data_f32_matches_cv2(golden):
    """oracle setData conversion == cv2 multiply / convertTo(CV_8U) / cvtColor(BGR2GRAY) incl. .5 ties, out-of-range and tiny values."""
    import oracle_lib as O
    g = golden("set_data_f32.npz")
    assert np.array_equal(O.set_data_f32(g["gray"]), g["gray_u8"]) and np.array_equal(O.set_data_f32(g["bgr"]), g["bgr_u8"])


def test_undistort_matches_cv2(golden):
    """oracle cv::undistort restatement (stripe-wise fixed-point maps + bilinear remap, BORDER_CONSTANT) == cv2.undistort bit for bit,
    on u8 images and f32 masks, for a strong barrel calibration and one that samples outside the source image (k3 != 0)."""
    import oracle_lib as O
    g = golden("undistort_cv2.npz")
    for k in ("1", "2"):
        assert np.array_equal(O.undistort(g["img"], g["K" + k], g["d" + k]), g["img_u" + k])
        assert np.array_equal(O.undistort(g["mask"], g["K" + k], g["d" + k]), g["mask_u" + k])
    assert (g["img_u2"] == 0).sum() > 100, "the second calibration must exercise the constant border"


def test_undistort_keypoints_matches_cv2(golden):
    """oracle UndistortKeyPoints == cv2.undistortPoints(pts, K32, dist32, P=K32) bit for bit; the reference's float read of the f64
    coefficient (MultiCameraFrame.cpp:302) makes 0.5 count as 'no distortion'."""
    import oracle_lib as O
    g = golden("undistort_points_cv2.npz")
    kp = np.zeros(len(g["pts"]), O.KP_DTYPE)
    kp["x"], kp["y"] = g["pts"][:, 0], g["pts"][:, 1]
    for k in ("1", "2"):
        o = O.undistort_keypoints(kp, g["K" + k], g["d" + k])
        assert np.array_equal(o["x"], g["out" + k][:, 0]) and np.array_equal(o["y"], g["out" + k][:, 1])
    assert np.array_equal(O.undistort_keypoints(kp, g["K1"], [0.5, 0.1, 0, 0, 0]), kp)
    assert np.array_equal(O.undistort_keypoints(kp, g["K1"], [0, 0.1, 0, 0, 0]), kp)


def _bow_intra_py(D, Y, fvs, th_low=75, ratio=0.85):
    """Second, loop-level restatement of computeIntraMatches(matches, words_) (MultiCameraFrame.cpp:586-943) in Python, written from the
    reference listing independently of oracle/mcs_oracle.cpp; small inputs only."""
    nc = len(D)
    if any(len(fv) == 0 for fv in fvs):
        return [], [], []
    nodes = [sorted(fv) for fv in fvs]
    it = [0] * nc
    end = [len(n) - 1 for n in nodes]
    matches, rays, words = [], [], []
    cur = 0
    pop = lambda a, b: int(np.unpackbits(a ^ b).sum())
    while not all(nodes[c][it[c]] >= nodes[c][end[c]] for c in range(nc)):
        wid = [2 ** 31 - 1 if it[c] == end[c] else nodes[c][it[c]] for c in range(nc)]
        lo = min(min(wid), 2 ** 31 - 2)
        sel = [c for c in range(nc) if wid[c] == lo] if lo < 2 ** 31 - 2 else []
        feats = [fvs[c][nodes[c][it[c]]] for c in sel]
        flags = [[-1] * len(f) for f in feats]
        if len(sel) >= 2:
            for i in range(len(sel) - 1):
                c1 = sel[i]
                for a, fa in enumerate(feats[i]):
                    if flags[i][a] != -1:
                        continue
                    found, once = False, True
                    matches.append([-1] * nc); rays.append(1)
                    matches[-1][c1] = fa
                    flags[i][a] = cur
                    for j in range(i + 1, len(sel)):
                        c2 = sel[j]
                        bj, b1, b2 = -1, 1e9, 1e9
                        for b, fb in enumerate(feats[j]):
                            if abs(np.float32(Y[c1][fa]) - np.float32(Y[c2][fb])) >= 50:
                                continue
                            d = pop(D[c1][fa], D[c2][fb])
                            if d < b1:
                                bj, b2, b1 = b, b1, d
                            elif d < b2:
                                b2 = d
                        if not (b1 <= th_low and b1 / b2 <= ratio):
                            continue
                        own = flags[j][bj]
                        if own == cur:
                            continue
                        if own == -1:
                            matches[cur][c2] = feats[j][bj]; rays[cur] += 1; flags[j][bj] = cur; found = True
                            continue
                        old = matches[own][c1]
                        if old == -1:
                            if not once:
                                continue
                            once = False
                            mine = [t for t in range(nc) if matches[cur][t] != -1]
                            if all(matches[own][t] == -1 for t in mine):
                                for t in mine:
                                    matches[own][t] = matches[cur][t]
                                rays[own] += len(mine)
                                flags[i][a] = own
                                cur = own
                                found = True
                                matches.pop(); rays.pop()
                            continue
                        if b1 < pop(D[c1][old], D[c2][feats[j][bj]]):
                            matches[own][c2] = -1; rays[own] -= 1
                            matches[cur][c2] = feats[j][bj]; rays[cur] += 1; flags[j][bj] = cur; found = True
                    if found:
                        words.append(nodes[sel[0]][it[sel[0]]])
                        cur = len(matches)
                    else:
                        matches.pop(); rays.pop()
                        flags[i][a] = -1
        for c in sel:
            it[c] += 1
        if not sel:
            break
    return matches, rays, words


def test_bow_intra_matches_oracle():
    """oracle intra_matches_bow: hand-checked cases (a three-view row; the never-visited last node; the |dy| >= 50 gate) and agreement
    with the Python restatement on clustered random descriptors, 2 to 5 cameras."""
    import oracle_lib as O
    kp = lambda ys: np.array([(0, y, 31, -1, 0, 0, -1) for y in ys], O.KP_DTYPE)
    d = np.zeros((2, 32), np.uint8)
    d[1] = 255
    fv = [{5: [0], 9: [1]}] * 3
    rows, nr, words = O.intra_matches_bow([d] * 3, [kp([10, 10])] * 3, fv)
    assert rows.tolist() == [[0, 0, 0]] and nr.tolist() == [3] and words.tolist() == [5]  # node 9 is the last node: never visited
    rows, nr, words = O.intra_matches_bow([d] * 3, [kp([10, 10]), kp([70, 10]), kp([20, 10])], fv)
    assert rows.tolist() == [[0, -1, 0]] and nr.tolist() == [2]  # camera 1's feature is 60 rows away from camera 0's
    assert len(O.intra_matches_bow([d] * 3, [kp([10, 10])] * 3, [{5: [0], 9: [1]}, {}, {5: [0]}])[0]) == 0  # an empty FeatureVector: early return
    rng = np.random.default_rng(123)
    for nc in (2, 3, 4, 5):
        centres = rng.integers(0, 256, (40, 32), dtype=np.uint8)
        D, K, fvs = [], [], []
        for c in range(nc):
            n = int(rng.integers(60, 110))
            src = rng.integers(0, 40, n)
            noise = rng.integers(0, 256, (n, 32), dtype=np.uint8) & rng.integers(0, 256, (n, 32), dtype=np.uint8) & rng.integers(0, 256, (n, 32), dtype=np.uint8) \
                & rng.integers(0, 256, (n, 32), dtype=np.uint8)
            D.append(centres[src] ^ noise)
            K.append(kp((src * 7 % 200 + rng.integers(0, 30, n)).astype(np.float32)))
            fv = {}
            for i, s in enumerate(src):
                if rng.random() < 0.9:
                    fv.setdefault(int(s // 4) * 3 + 1, []).append(i)
            fvs.append(fv)
        rows, nr, words = O.intra_matches_bow(D, K, fvs)
        er, en, ew = _bow_intra_py(D, [k["y"] for k in K], fvs)
        assert len(rows) > 10
        assert rows.tolist() == er and nr.tolist() == en and words.tolist() == ew
        if nc >= 3:
            assert len(words) > len(rows) or (nr == 1).any() or (nr >= 3).any()
