"""GPU parity: MultiCameraFrame extract -> BruteForceMatch -> computeIntraMatches -> triangulation on the B200 (through
the C ABI) against the CPU oracle and the cv2 golden rig fixture.  Match indices and IntraMatch rows bit-exact; 3-D points
within 1e-6 relative (f64 N-view DLT) / 1e-3 (f32 2-view)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _oracle_frame(O, imgs, calib, nf, masks=None, old=False):
    ex = [O.Extractor(nf) for _ in range(len(imgs))]
    K, D = [], []
    for c in range(len(imgs)):
        _, k, d = ex[c](imgs[c])
        K.append(k)
        D.append(d)
    rows, pc = O.intra_matches(D, K, calib, ex[0].sigma2, use_epipolar=old)
    tri = O.triangulate_rows(calib, K, D, rows, masks=masks)
    return K, D, rows, pc, tri


def _check_frame(rig, f, K, D, rows, tri, C):
    for c in range(C):
        k, d = rig.features(f, c)
        assert len(k) == len(K[c]) and (k == K[c]).all() and np.array_equal(d, D[c]), "features frame %d cam %d" % (f, c)
    r = rig.intra_matches(f)
    assert r.shape == rows.shape and np.array_equal(r, rows), "IntraMatch rows frame %d: gpu %s oracle %s" % (f, r.shape, rows.shape)
    lm = rig.landmarks(f)
    assert np.array_equal(lm["rows"], tri["rows"]) and np.array_equal(lm["nviews"], tri["nviews"])
    assert np.array_equal(lm["status"], tri["status"]), "depth-gate decisions"
    assert np.array_equal(lm["rep_cam"], tri["rep_cam"]), "representative descriptor choice"
    ok = tri["nviews"] >= 2
    assert np.allclose(lm["xyz"][ok], tri["xyz"][ok], rtol=1e-6, atol=1e-9), np.abs(lm["xyz"][ok] - tri["xyz"][ok]).max()
    acc = tri["status"] == 1
    assert np.allclose(lm["uv_ref"][acc], tri["uv_ref"][acc], rtol=1e-5, atol=1e-3)
    kc, nr, na = rig.counts(f)
    assert nr == len(rows) and na == tri["accepted"] and (kc == [len(k) for k in K]).all()


def test_rig_golden(golden):
    import mcslam_b200 as M
    g = golden("rig3_480x360.npz")
    rig = M.RigFrames(g["calib"], 480, 360, nfeatures=600, max_frames=1)
    rig.extract(g["images"][None])
    for c in range(3):
        k, d = rig.features(0, c)
        assert (k == g["kps%d" % c]).all() and np.array_equal(d, g["desc%d" % c])
    p = 0
    for i in range(2):
        for j in range(i + 1, 3):
            i1, i2, k1, k2 = rig.brute_force_match(0, i, j, 75, 0.85)
            assert np.array_equal(i1, g["pair%d_i1" % p]) and np.array_equal(i2, g["pair%d_i2" % p]), "pair %d" % p
            assert (k1 == g["kps%d" % i][i1]).all() and (k2 == g["kps%d" % j][i2]).all()
            p += 1
    rig.compute_intra_matches(1)
    assert np.array_equal(rig.intra_matches(0), g["rows"])
    rig.triangulate(1)
    lm = rig.landmarks(0)
    ok = lm["nviews"] >= 2
    assert np.allclose(lm["xyz"][ok], g["xyz"][ok], rtol=1e-6, atol=1e-9)


@pytest.mark.parametrize("old", [False, True])
def test_rig5_process_vs_oracle(old):
    """Config 2: 5 cameras 720x540, 1000 features; two frames in one batch through the one-call path."""
    import mcslam_b200 as M
    import oracle_lib as O
    calib = M.synth.make_rig(5, 720, 540)
    frames = np.stack([M.synth.make_rig_frame(calib, 720, 540, frame_index=i) for i in (0, 1)])
    rig = M.RigFrames(calib, 720, 540, nfeatures=1000, max_frames=2)
    rig.process(frames, old=old)
    for f in range(2):
        K, D, rows, pc, tri = _oracle_frame(O, frames[f], calib, 1000, old=old)
        _check_frame(rig, f, K, D, rows, tri, 5)
        assert len(rows) > 300 and tri["accepted"] > 200  # the synthetic scene really produces multi-view landmarks


def test_rig5_masks_2000_features():
    """Config 3: semantic masks (CV_32FC1, keep iff mask < 0.7) applied in the triangulation kernel, 2000 features."""
    import mcslam_b200 as M
    import oracle_lib as O
    calib = M.synth.make_rig(5, 720, 540)
    frame = M.synth.make_rig_frame(calib, 720, 540, frame_index=3)
    masks = M.synth.make_masks(5, 720, 540, frame_index=3)
    rig = M.RigFrames(calib, 720, 540, nfeatures=2000, max_frames=1)
    rig.process(frame[None], masks=masks[None])
    K, D, rows, pc, tri = _oracle_frame(O, frame, calib, 2000, masks=list(masks))
    _check_frame(rig, 0, K, D, rows, tri, 5)
    assert (tri["rows"] != rows).any(), "masks must actually remove some views"


def test_rig8_wide():
    """8 cameras (IntraMatch width 8 is an extension of the reference's hard-coded 5): pairwise matches + rows vs oracle."""
    import mcslam_b200 as M
    import oracle_lib as O
    calib = M.synth.make_rig(8, 480, 360, spacing=0.1)
    frame = M.synth.make_rig_frame(calib, 480, 360, frame_index=7, n_rects=8)
    rig = M.RigFrames(calib, 480, 360, nfeatures=700, max_frames=1)
    rig.process(frame[None])
    K, D, rows, pc, tri = _oracle_frame(O, frame, calib, 700)
    _check_frame(rig, 0, K, D, rows, tri, 8)


def test_match_edge_cases_and_standalone_triangulation():
    import mcslam_b200 as M
    import oracle_lib as O
    calib = M.synth.make_rig(5, 720, 540, spacing=0.165)
    rig = M.RigFrames(calib, 720, 540, nfeatures=1000, max_frames=1)
    # a flat rig frame: no keypoints anywhere -> no rows, nothing crashes
    rig.process(np.full((1, 5, 540, 720), 90, np.uint8))
    assert len(rig.intra_matches(0)) == 0 and len(rig.landmarks(0)["rows"]) == 0
    # known-answer fixture (MultiCameraFrame.cpp:383-409): X = (0,0,6)
    X = np.array([0, 0, 6.0, 1.0])
    xs, ys = [], []
    for c in range(5):
        K = calib[c]["K"].reshape(3, 3)
        Rt = np.hstack([calib[c]["R"].reshape(3, 3), calib[c]["t"].reshape(3, 1)])
        p = K @ Rt @ X
        xs.append(p[0] / p[2])
        ys.append(p[1] / p[2])
    for views in ([0, 1], [0, 2, 4], [0, 1, 2, 3, 4]):
        got = rig.triangulate_nview(views, [[xs[v] for v in views]], [[ys[v] for v in views]])[0]
        assert np.allclose(got, [0, 0, 6], atol=1e-8), (views, got)
        ref = O.triangulate_pixels(calib, views, [xs[v] for v in views], [ys[v] for v in views])
        assert np.allclose(got, ref, rtol=1e-6)
    rng = np.random.default_rng(5)
    P3 = np.stack([rng.uniform(-3, 3, 200), rng.uniform(-2, 2, 200), rng.uniform(2, 30, 200)], 1)
    pts = []
    for c in (0, 4):
        K = calib[c]["K"].reshape(3, 3)
        Xc = P3 @ calib[c]["R"].reshape(3, 3).T + calib[c]["t"]
        uv = Xc @ K.T
        pts.append((uv[:, :2] / uv[:, 2:]).astype(np.float32))
    got = rig.triangulate_pair(0, 4, pts[0], pts[1])
    ref = O.triangulate_pair_f32(calib, 0, 4, pts[0], pts[1])
    rel = np.abs(got - ref).max(1) / np.abs(ref).max(1)
    print("2-view f32 DLT: relative error max %.3g, median %.3g" % (rel.max(), np.median(rel)))
    assert rel.max() < 1e-3  # f32 4x4 SVD: 1e-3 relative, every point (measured on B200: 0, the same f32 Jacobi order)


def test_device_resident_path_matches_host_path():
    import mcslam_b200 as M
    from mcslam_b200 import binding as B
    calib = M.synth.make_rig(5, 720, 540)
    frames = np.stack([M.synth.make_rig_frame(calib, 720, 540, frame_index=i) for i in (4,)])
    rig = M.RigFrames(calib, 720, 540, nfeatures=1000, max_frames=1)
    rig.process(frames)
    ref = [rig.features(0, c) for c in range(5)], rig.intra_matches(0), rig.landmarks(0)
    dptr = B.device_malloc(0, frames.nbytes)
    B.device_upload(0, dptr, frames)
    rig.process_device(dptr, 1)
    rig.synchronize()
    for c in range(5):
        k, d = rig.features(0, c)
        assert (k == ref[0][c][0]).all() and np.array_equal(d, ref[0][c][1])
    assert np.array_equal(rig.intra_matches(0), ref[1])
    assert np.array_equal(rig.landmarks(0)["xyz"], ref[2]["xyz"])
    assert rig.kernel_launches > 0
    B.device_free(0, dptr)


def test_cpp_adaptor_equals_oracle(tmp_path):
    """The C++ host adaptor (mc-slam_b200/host/mcslam_adaptor.hpp: ORBextractor::operator(), MultiCameraFrame::extractFeaturesParallel /
    computeIntraMatches / triangulateIntraMatches) driven from a C++ program, against the oracle on the same rig frame."""
    import subprocess
    import mcslam_b200 as M
    import oracle_lib as O
    from test_host_side import _build_adaptor
    exe = _build_adaptor(tmp_path)
    calib = M.synth.make_rig(3, 320, 240)
    frame = M.synth.make_rig_frame(calib, 320, 240, frame_index=5, n_rects=6)
    (tmp_path / "im.bin").write_bytes(frame.tobytes())
    (tmp_path / "im.bin.calib").write_bytes(np.ascontiguousarray(calib).tobytes())
    out = subprocess.run([exe, str(tmp_path / "im.bin"), "3", "320", "240", str(tmp_path / "o.txt")], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    lines = (tmp_path / "o.txt").read_text().splitlines()
    K, D, rows, pc, tri = _oracle_frame(O, frame, calib, 400)
    head = lines[0].split()
    assert int(head[1]) == len(K[0]) and int(head[3]) == len(K[0]) and int(head[5]) == 8
    kp = [l.split() for l in lines if l.startswith("kp ")]
    assert len(kp) == len(K[0])
    for l, k, d in zip(kp, K[0], D[0]):
        assert np.float32(l[1]) == k["x"] and np.float32(l[2]) == k["y"] and int(l[3]) == k["octave"] and np.float32(l[4]) == k["response"]
        assert int(l[5]) == d[0]
    rl = [list(map(int, l.split()[1:])) for l in lines if l.startswith("row ")]
    assert len(rl) == len(rows) and all(r == [row[0], row[1], row[2]] for r, row in zip(rl, rows))
    lm = [l.split() for l in lines if l.startswith("lm ")]
    acc = tri["status"] == 1
    assert len(lm) == int(acc.sum())
    got = np.array([[float(v) for v in l[2:5]] for l in lm])
    assert np.allclose(got, tri["xyz"][acc], rtol=1e-6, atol=1e-9)
    assert [int(l[1]) for l in lm] == tri["nviews"][acc].tolist()
    # obtainLfFeatures bookkeeping: accepted rows, then mono features by descending response (unstable std::sort order), budget 700
    accepted, mono = O.obtain_lf_features(K, tri, total_feats=700)
    sizes = [l.split() for l in lines if l.startswith("lfsizes ")][0]
    assert int(sizes[1]) == len(accepted) and int(sizes[2]) == len(mono) == 700 - len(accepted) and int(sizes[3]) == 700
    got_mono = [tuple(map(int, l.split()[1:])) for l in lines if l.startswith("lfmono ")]
    assert [(c, j) for c, j, _, _ in got_mono] == mono
    assert all(m == 1 and d0 == D[c][j][0] for c, j, m, d0 in got_mono)


def test_cv_api_equals_oracle(tmp_path):
    """The cv::/Eigen-typed layer (mc-slam_b200/host/mcslam_cv.hpp): ORBextractor::operator()(InputArray, InputArray, vector<KeyPoint>&,
    OutputArray, vector<int>&), MultiCameraFrame with the reference's member types (image_kps, image_descriptors as Mat rows, segMasks,
    intraMatches with cv::Mat matchDesc / point3D), BruteForceMatch / triangulate / triangulateIntraMatches(Mat_<double>, vector<int>,
    Eigen::Vector3d&) / computeRepresentativeDesc with cv containers, driven from tests/cpp/test_cv_api.cpp with the front-end's own
    consumer patterns, against the oracle on the same rig frame (with segmentation masks)."""
    import subprocess
    import mcslam_b200 as M
    import oracle_lib as O
    from test_host_side import _build_cv_api
    exe = _build_cv_api(tmp_path)
    calib = M.synth.make_rig(3, 320, 240)
    frame = M.synth.make_rig_frame(calib, 320, 240, frame_index=5, n_rects=6)
    masks = M.synth.make_masks(3, 320, 240, frame_index=2)
    (tmp_path / "im.bin").write_bytes(frame.tobytes())
    (tmp_path / "im.bin.calib").write_bytes(np.ascontiguousarray(calib).tobytes())
    (tmp_path / "masks.bin").write_bytes(np.ascontiguousarray(masks, np.float32).tobytes())
    out = subprocess.run([exe, str(tmp_path / "im.bin"), "3", "320", "240", str(tmp_path / "o.txt"), str(tmp_path / "masks.bin")],
                         capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    lines = [l.split() for l in (tmp_path / "o.txt").read_text().splitlines()]
    K, D, rows, pc, tri = _oracle_frame(O, frame, calib, 400, masks=masks)
    assert lines[0] == ["orb", str(len(K[0])), str(len(K[0])), str(len(K[0])), "32"]
    assert [l for l in lines if l[0] == "same0"] == [["same0", "1"]]
    for c in range(3):
        kp = [l for l in lines if l[0] == "kp" and int(l[1]) == c]
        assert len(kp) == len(K[c])
        for l, k, d in zip(kp, K[c], D[c]):
            assert np.float32(l[2]) == k["x"] and np.float32(l[3]) == k["y"] and int(l[4]) == k["octave"] and np.float32(l[5]) == k["response"]
            assert int(l[6]) == d[0] and int(l[7]) == d[31] and l[8:] == ["1", "32"]
    rl = [list(map(int, l[1:])) for l in lines if l[0] == "row"]
    assert len(rl) == len(rows) and all(r == list(row) for r, row in zip(rl, rows))
    i1, i2 = O.bf_match(D[0], D[1])[:2]
    bf = [l for l in lines if l[0] == "bf"][0]
    assert int(bf[1]) == len(i1) and bf[2:] == ["%d:%d" % (a, b) for a, b in zip(i1, i2)]
    assert [l for l in lines if l[0] == "bfkps"] == [["bfkps", "1"]]
    # obtainLfFeatures: accepted rows in order with n_rays, point3D, uv_ref and the representative descriptor; then the mono top-up
    acc = np.flatnonzero(tri["status"] == 1)
    accepted, mono = O.obtain_lf_features(K, tri, total_feats=700)
    assert list(acc) == list(accepted) and len(acc) > 20
    sizes = [l for l in lines if l[0] == "lfsizes"][0]
    assert list(map(int, sizes[1:])) == [len(acc), len(mono), len(acc) + len(mono), len(acc) + len(mono)]
    lm = [l for l in lines if l[0] == "lm"]
    assert len(lm) == len(acc)
    for l, r in zip(lm, acc):
        assert int(l[1]) == tri["nviews"][r]
        assert np.allclose([float(v) for v in l[2:5]], tri["xyz"][r], rtol=1e-6, atol=1e-9)
        assert np.float32(l[5]) == tri["uv_ref"][r][0] and np.float32(l[6]) == tri["uv_ref"][r][1]
        rc = tri["rep_cam"][r]
        d = D[rc][tri["rows"][r][rc]]
        assert int(l[7]) == d[0] and int(l[8]) == d[31]
    got_mono = [tuple(map(int, l[1:])) for l in lines if l[0] == "lfmono"]
    assert [(c, j) for c, j, _, _ in got_mono] == mono and all(m == 1 and d0 == D[c][j][0] for c, j, m, d0 in got_mono)
    dif = [l for l in lines if l[0] == "dif"][0]
    assert np.allclose([float(v) for v in dif[1:]], tri["xyz"][acc[0]] - tri["xyz"][acc[1]], rtol=1e-6, atol=1e-9)
    # triangulateIntraMatches(Mat_<double>, view_inds, Eigen::Vector3d&) on the first accepted row; computeRepresentativeDesc on its rows
    r0 = acc[0]
    views = [c for c in range(3) if tri["rows"][r0][c] >= 0]
    ref = O.triangulate_pixels(calib, views, [K[c][tri["rows"][r0][c]]["x"] for c in views], [K[c][tri["rows"][r0][c]]["y"] for c in views])
    t = [l for l in lines if l[0] == "tri"][0]
    assert np.allclose([float(v) for v in t[1:]], ref, rtol=1e-6, atol=1e-9)
    assert [l for l in lines if l[0] == "rep"] == [["rep", "1"]]
    # triangulate(vector<Point2f>&, vector<Point2f>&, vector<Point3f>&, 0, 1) on the BruteForceMatch pairs
    p1 = np.stack([K[0]["x"][i1], K[0]["y"][i1]], 1).astype(np.float32)
    p2 = np.stack([K[1]["x"][i2], K[1]["y"][i2]], 1).astype(np.float32)
    ref3 = O.triangulate_pair_f32(calib, 0, 1, p1, p2)
    p3 = np.array([[float(v) for v in l[1:]] for l in lines if l[0] == "p3"], np.float32)
    assert p3.shape == ref3.shape
    fin = np.isfinite(ref3).all(1)
    assert np.array_equal(np.isfinite(p3).all(1), fin)
    rel = np.abs(p3[fin] - ref3[fin]).max(1) / np.maximum(np.abs(ref3[fin]).max(1), 1e-6)
    assert rel.max() < 1e-3


def test_masks_u8_and_graph_latency_mode():
    """(1) mcs_rig_process_masks_u8 / mcs_rig_process_device_masks_u8: the caller's 1-byte decision (0 = mask < 0.7) gives the rows, mask
    decisions and points of the f32 masks (FrontEnd.cpp:266).  (2) mcs_rig_set_graph: the CUDA-graph replay of an unchunked
    device-resident call equals the stream path, frame counts 1 and 3, with and without masks; changing the input pixels between
    replays changes the results accordingly (the graph holds pointers, not data)."""
    import mcslam_b200 as M
    from mcslam_b200 import binding as B
    calib = M.synth.make_rig(3, 320, 240)
    frames = np.stack([M.synth.make_rig_frame(calib, 320, 240, frame_index=30 + i, n_rects=5) for i in range(3)])
    masks = np.stack([M.synth.make_masks(3, 320, 240, frame_index=i) for i in range(3)])
    masks[0, 0, :8, :8] = 0.7  # exactly 0.7f passes `< 0.7` under the float -> double promotion
    mu8 = np.where(masks.astype(np.float64) < 0.7, 0, 255).astype(np.uint8)
    rig = M.RigFrames(calib, 320, 240, nfeatures=300, max_frames=3)

    def snapshot(n):
        out = []
        for f in range(n):
            feats = [rig.features(f, c) for c in range(3)]
            out.append((feats, rig.intra_matches(f), rig.landmarks(f), rig.counts(f)))
        return out

    def same(a, b):
        assert len(a) == len(b)
        for (fa, ra, la, ca), (fb, rb, lb, cb) in zip(a, b):
            for (ka, da), (kb, db) in zip(fa, fb):
                assert (ka == kb).all() and np.array_equal(da, db)
            assert np.array_equal(ra, rb)
            for k in la:
                assert np.array_equal(la[k], lb[k]), k
            assert (ca[0] == cb[0]).all() and ca[1:] == cb[1:]

    rig.process(frames, masks=masks)
    ref_m = snapshot(3)
    assert sum(int((r[2]["nviews"] < (r[2]["rows"] >= 0).sum(1)).sum() + (r[1] != r[2]["rows"]).any(1).sum()) for r in ref_m) > 0, "masks must drop views"
    rig.process_masks_u8(frames, mu8)
    same(ref_m, snapshot(3))
    d_img = B.device_malloc(0, frames.nbytes)
    B.device_upload(0, d_img, frames)
    d_mu8 = B.device_malloc(0, mu8.nbytes)
    B.device_upload(0, d_mu8, mu8)
    rig.process_device_masks_u8(d_img, 3, d_mu8)
    same(ref_m, snapshot(3))
    rig.process(frames)
    ref_plain = snapshot(3)
    # graph mode
    rig.set_graph(True)
    for n, dm, ref in ((3, None, ref_plain), (1, None, ref_plain[:1]), (3, d_mu8, ref_m)):
        r0 = rig.graph_replays
        for it in range(3):  # capture, replay, replay
            if dm is None:
                rig.process_device(d_img, n)
            else:
                rig.process_device_masks_u8(d_img, n, dm)
            same(ref, snapshot(n))
        assert rig.graph_replays == r0 + 2
    # new pixels behind the same pointer: the replay computes the new frames
    frames2 = np.roll(frames, 16, axis=3)
    rig.set_graph(False)
    B.device_upload(0, d_img, frames2)
    rig.process_device(d_img, 3)
    ref2 = snapshot(3)
    assert any(not np.array_equal(a[1], b[1]) for a, b in zip(ref2, ref_plain))
    rig.set_graph(True)
    B.device_upload(0, d_img, frames)
    rig.process_device(d_img, 3)   # capture on the old pixels
    B.device_upload(0, d_img, frames2)
    r0 = rig.graph_replays
    rig.process_device(d_img, 3)   # replay on the new ones
    assert rig.graph_replays == r0 + 1
    same(ref2, snapshot(3))
    rig.set_graph(False)
    rig.process_device(d_img, 3)
    same(ref2, snapshot(3))
    B.device_free(0, d_img)
    B.device_free(0, d_mu8)


def test_row_download_estimate_and_top_up():
    """mcs_rig_process sizes the downloads of the IntraMatch rows and landmark arrays by the previous call's fullest frame (+ 25 % + 64)
    instead of the table capacity.  A call on nearly flat frames (a handful of rows) followed by a call on textured frames (hundreds)
    must fetch the missing rows afterwards; the results equal those of a fresh handle, and the byte count shrinks once the estimate is
    in place."""
    import mcslam_b200 as M
    calib = M.synth.make_rig(3, 320, 240)
    rich = np.stack([M.synth.make_rig_frame(calib, 320, 240, frame_index=40 + i, n_rects=6) for i in range(3)])
    rng = np.random.default_rng(3)
    flat = np.full_like(rich, 120)
    flat[:, :, 60:120, 80:160] = rng.integers(0, 256, (3, 3, 60, 80), dtype=np.uint8)  # a little texture: few matches

    def snapshot(rig):
        return [([rig.features(f, c) for c in range(3)], rig.intra_matches(f), rig.landmarks(f), rig.counts(f)) for f in range(3)]

    def same(a, b):
        for (fa, ra, la, ca), (fb, rb, lb, cb) in zip(a, b):
            for (ka, da), (kb, db) in zip(fa, fb):
                assert (ka == kb).all() and np.array_equal(da, db)
            assert np.array_equal(ra, rb)
            for k in la:
                assert np.array_equal(la[k], lb[k]), k
            assert (ca[0] == cb[0]).all() and ca[1:] == cb[1:]

    fresh = M.RigFrames(calib, 320, 240, nfeatures=300, max_frames=3)
    fresh.process(rich)
    want_rich = snapshot(fresh)
    full_bytes = fresh.last_d2h_bytes
    rows_rich = max(len(r[1]) for r in want_rich)
    fresh2 = M.RigFrames(calib, 320, 240, nfeatures=300, max_frames=3)
    fresh2.process(flat)
    want_flat = snapshot(fresh2)
    rows_flat = max(len(r[1]) for r in want_flat)
    assert rows_rich > rows_flat + rows_flat // 4 + 64, (rows_rich, rows_flat)  # the rich frames outgrow the estimate left by the flat ones
    rig = M.RigFrames(calib, 320, 240, nfeatures=300, max_frames=3)
    rig.process(flat)
    same(want_flat, snapshot(rig))
    rig.process(rich)            # estimate from the flat call is too small: top-up path
    same(want_rich, snapshot(rig))
    rig.process(rich)            # estimate now fits: trimmed copies only
    same(want_rich, snapshot(rig))
    assert rig.last_d2h_bytes < full_bytes
    rig.process(flat)
    same(want_flat, snapshot(rig))


def test_imma_matcher_experiment_equals_product_matcher(monkeypatch):
    """MCS_MATCH_IMMA=1 (an experiment, not the product path: north_star keeps the matcher on XOR + popc) computes the pair distances as
    popc(q) + popc(t) - 2 <q, t> with the 0 / 1-byte dot product on the u8 integer MMA.  Same packed keys, same top-2, same accept rule:
    rows, landmarks and every pair's match list equal those of hamming_match_kernel, on ragged feature counts (cameras with 1, 0 and
    many features included through a nearly flat image) and with the epipolar gate."""
    import mcslam_b200 as M
    calib = M.synth.make_rig(4, 320, 240)
    frames = np.stack([M.synth.make_rig_frame(calib, 320, 240, frame_index=50 + i, n_rects=6) for i in range(2)])
    frames[1, 2] = 127          # a camera without features
    frames[1, 3] = 127
    frames[1, 3, 100:108, 100:108] = 255   # ... and one with a handful

    def run(old):
        rig = M.RigFrames(calib, 320, 240, nfeatures=400, max_frames=2)
        rig.process(frames, old=old)
        out = []
        for f in range(2):
            pairs = [rig.brute_force_match(f, i, j) for i in range(4) for j in range(i + 1, 4)]
            out.append((rig.intra_matches(f), rig.landmarks(f), pairs, rig.counts(f)))
        return out

    for old in (False, True):
        monkeypatch.delenv("MCS_MATCH_IMMA", raising=False)
        ref = run(old)
        monkeypatch.setenv("MCS_MATCH_IMMA", "1")
        got = run(old)
        monkeypatch.delenv("MCS_MATCH_IMMA")
        assert sum(len(r[0]) for r in ref) > 100
        for (ra, la, pa, ca), (rb, lb, pb, cb) in zip(ref, got):
            assert np.array_equal(ra, rb)
            for k in la:
                assert np.array_equal(la[k], lb[k]), k
            for x, y in zip(pa, pb):
                assert all(np.array_equal(u, v) for u, v in zip(x, y))
            assert (ca[0] == cb[0]).all() and ca[1:] == cb[1:]


@pytest.mark.parametrize("old", [False, True])
def test_small_launch_kernel_variants(monkeypatch, old):
    """Two kernels come in a form for launches of many frames and one for launches of few, picked by call size: merge_kernel<256> /
    merge_kernel<1024> (one query per thread) and tri_solve_kernel (one thread per row) / tri_solve4_kernel (four lanes per row).
    MCS_MERGE_WIDE and MCS_TRI_LANES force either.  All four combinations against the oracle on a 5-camera frame with ~1000 features
    per camera, and the two solvers against each other: they sum a rotation's dot products in the same order (four partial sums), so
    their points are bit-identical and the size of a call does not show in its results."""
    import mcslam_b200 as M
    import oracle_lib as O
    calib = M.synth.make_rig(5, 720, 540)
    frames = np.stack([M.synth.make_rig_frame(calib, 720, 540, frame_index=31)])
    K, D, rows, pc, tri = _oracle_frame(O, frames[0], calib, 1000, old=old)
    assert len(rows) > 500 and (tri["nviews"] >= 3).sum() > 100
    xyz = {}
    for wide in ("0", "1"):
        for lanes in ("1", "4"):
            monkeypatch.setenv("MCS_MERGE_WIDE", wide)
            monkeypatch.setenv("MCS_TRI_LANES", lanes)
            rig = M.RigFrames(calib, 720, 540, nfeatures=1000, max_frames=1)
            rig.process(frames, old=old)
            _check_frame(rig, 0, K, D, rows, tri, 5)
            xyz[lanes] = rig.landmarks(0)["xyz"].copy()
    monkeypatch.delenv("MCS_MERGE_WIDE")
    monkeypatch.delenv("MCS_TRI_LANES")
    ok = tri["nviews"] >= 2
    assert np.array_equal(xyz["1"][ok], xyz["4"][ok])


@pytest.mark.parametrize("ncam,w,h,nf,old,with_masks", [(2, 320, 240, 400, False, True), (6, 256, 192, 250, True, False), (7, 200, 160, 150, False, True)])
def test_rig_camera_count_sweep(ncam, w, h, nf, old, with_masks):
    """Rigs of 2, 6 and 7 cameras (1, 15 and 21 camera pairs; the reference's IntraMatch holds 5, the row table here is [M][num_cams]):
    features, rows, mask / depth-gate decisions and points of a two-frame batch against the oracle."""
    import mcslam_b200 as M
    import oracle_lib as O
    calib = M.synth.make_rig(ncam, w, h, spacing=0.12)
    frames = np.stack([M.synth.make_rig_frame(calib, w, h, frame_index=70 + ncam + i, n_rects=6) for i in range(2)])
    masks = np.stack([M.synth.make_masks(ncam, w, h, frame_index=9 + i) for i in range(2)]) if with_masks else None
    rig = M.RigFrames(calib, w, h, nfeatures=nf, max_frames=2)
    rig.process(frames, masks=masks, old=old)
    for f in range(2):
        K, D, rows, pc, tri = _oracle_frame(O, frames[f], calib, nf, masks=None if masks is None else masks[f], old=old)
        assert len(rows) > 30
        _check_frame(rig, f, K, D, rows, tri, ncam)


def test_repeatability_stress():
    """compute-sanitizer is not available on this pool, so races are hunted by repetition: a batch of 48 rig frames that holds the same
    three frames 16 times over, processed three times (device-resident, six-stream chunk pipeline with chunks of 8).  Every copy of a
    frame, in every pass, must produce identical features, rows and landmarks — a race in any kernel shows up as a difference."""
    import mcslam_b200 as M
    from mcslam_b200 import binding as B
    calib = M.synth.make_rig(5, 720, 540)
    base = np.stack([M.synth.make_rig_frame(calib, 720, 540, frame_index=90 + i) for i in range(3)])
    frames = np.concatenate([base] * 16)
    rig = M.RigFrames(calib, 720, 540, nfeatures=1000, max_frames=48)
    rig.set_chunk_frames(8)
    d_img = B.device_malloc(0, frames.nbytes)
    B.device_upload(0, d_img, frames)
    ref = None
    for rep in range(3):
        rig.process_device(d_img, 48)
        for f in range(48):
            got = ([rig.features(f, c) for c in range(5)], rig.intra_matches(f), rig.landmarks(f))
            if ref is None or len(ref) < 3:
                ref = (ref or []) + [got]
                continue
            want = ref[f % 3]
            for (ka, da), (kb, db) in zip(got[0], want[0]):
                assert (ka == kb).all() and np.array_equal(da, db), (rep, f)
            assert np.array_equal(got[1], want[1]), (rep, f)
            for k in got[2]:
                assert np.array_equal(got[2][k], want[2][k]), (rep, f, k)
    B.device_free(0, d_img)


def test_nview_dlt_matches_cv2_svd_fixture(golden):
    """mcs_rig_triangulate_nview (MultiCameraFrame::triangulateIntraMatches(x, view_inds, pt), :411-441) on the fixture of
    oracle/gen_golden_dlt.py: X from cv2.SVDecomp of the same design matrices, 200 noisy points over 2..5 views, 1e-9 relative."""
    import mcslam_b200 as M
    g = golden("dlt_cv2.npz")
    rig = M.RigFrames(g["calib"], 720, 540, nfeatures=100, max_frames=1)
    worst = 0.0
    for v, xs, ys, want in zip(g["views"], g["xs"], g["ys"], g["xyz"]):
        n = int((v >= 0).sum())
        got = rig.triangulate_nview(v[:n].tolist(), [xs[:n].tolist()], [ys[:n].tolist()])[0]
        worst = max(worst, float(np.abs(got - want).max() / np.abs(want).max()))
    assert worst < 1e-9, worst
    got = rig.triangulate_pair(0, 4, g["pair_p1"], g["pair_p2"])  # :481-507 against cv2.SVDecomp on the f32 system
    rel = np.abs(got - g["pair_xyz"]).max(1) / np.abs(g["pair_xyz"]).max(1)
    assert rel.max() < 1e-5, rel.max()


def test_chunked_pipeline_equals_single_chunk():
    """mcs_rig_process cuts a batch into chunks over three streams; results must not depend on the chunking (7 frames, chunks of 2
    -> 4 chunks incl. a ragged last one, vs one chunk), on the host-buffer and the device-resident path."""
    import mcslam_b200 as M
    from mcslam_b200 import binding as B
    calib = M.synth.make_rig(3, 320, 240)
    base = [M.synth.make_rig_frame(calib, 320, 240, frame_index=20 + i, n_rects=5) for i in range(3)]
    frames = np.stack([np.roll(base[i % 3], 8 * (i // 3), axis=2) for i in range(7)])
    masks = np.stack([M.synth.make_masks(3, 320, 240, frame_index=i) for i in range(7)])
    rig = M.RigFrames(calib, 320, 240, nfeatures=300, max_frames=7)

    def snapshot():
        out = []
        for f in range(7):
            feats = [rig.features(f, c) for c in range(3)]
            lm = rig.landmarks(f)
            out.append((feats, rig.intra_matches(f), lm, rig.counts(f)))
        return out

    def same(a, b):
        for (fa, ra, la, ca), (fb, rb, lb, cb) in zip(a, b):
            for (ka, da), (kb, db) in zip(fa, fb):
                assert (ka == kb).all() and np.array_equal(da, db)
            assert np.array_equal(ra, rb)
            for k in la:
                assert np.array_equal(la[k], lb[k]), k
            assert (ca[0] == cb[0]).all() and ca[1:] == cb[1:]

    rig.set_chunk_frames(0)
    rig.process(frames, masks=masks)
    ref = snapshot()
    assert sum(len(r[1]) for r in ref) > 100
    rig.set_chunk_frames(2)
    rig.process(frames, masks=masks)
    same(ref, snapshot())
    d_img = B.device_malloc(0, frames.nbytes)
    B.device_upload(0, d_img, frames)
    d_msk = B.device_malloc(0, masks.nbytes)
    B.device_upload(0, d_msk, masks)
    rig.set_chunk_frames(3)
    rig.process_device(d_img, 7, d_masks=d_msk)
    same(ref, snapshot())
    B.device_free(0, d_img)
    B.device_free(0, d_msk)


@pytest.mark.parametrize("channels", [1, 3])
def test_process_f32_equals_set_data_then_u8_path(golden, channels):
    """mcs_rig_process_f32 / mcs_rig_process_device_f32 (setData conversion fused in front, MultiCameraFrame.cpp:107-116): level 0 equals
    the oracle's / cv2's converted image bit for bit, and every downstream result equals the u8 entry point fed with that image.
    5 frames in chunks of 2 (ragged last chunk)."""
    import mcslam_b200 as M
    import oracle_lib as O
    from mcslam_b200 import binding as B
    g = golden("set_data_f32.npz")
    calib = M.synth.make_rig(3, 320, 240)
    u8 = np.stack([M.synth.make_rig_frame(calib, 320, 240, frame_index=60 + i, n_rects=6) for i in range(5)])
    rng = np.random.default_rng(21)
    f32 = ((u8.astype(np.float32) + rng.uniform(-0.6, 0.6, u8.shape).astype(np.float32)) / np.float32(255.0)).astype(np.float32)
    if channels == 3:
        f32 = np.stack([f32 * np.float32(0.9), f32, np.clip(f32 * np.float32(1.1), 0, 1.05)], -1).astype(np.float32)
    f32 = np.ascontiguousarray(f32)
    want = O.set_data_f32(f32)
    assert want.shape == u8.shape and (want != u8).any()
    rig = M.RigFrames(calib, 320, 240, nfeatures=300, max_frames=5)
    rig.set_chunk_frames(4)  # host path uses chunks of 2
    rig.process(want)
    ref = [([rig.features(f, c) for c in range(3)], rig.intra_matches(f), rig.landmarks(f)) for f in range(5)]

    def check():
        for f in range(5):
            for c in range(3):
                assert np.array_equal(rig.level(f, c, 0), want[f, c]), (f, c)
                k, d = rig.features(f, c)
                assert (k == ref[f][0][c][0]).all() and np.array_equal(d, ref[f][0][c][1])
            assert np.array_equal(rig.intra_matches(f), ref[f][1])
            lm = rig.landmarks(f)
            for key in lm:
                assert np.array_equal(lm[key], ref[f][2][key]), key

    assert rig.process_f32(f32) == 5
    check()
    dptr = B.device_malloc(0, f32.nbytes)
    B.device_upload(0, dptr, f32)
    rig.process(u8)  # overwrite the handle's state with something else first
    rig.process_device_f32(dptr, channels, 5)
    rig.synchronize()
    check()
    B.device_free(0, dptr)
    # the cv2 fixture straight through the kernel: a 1-camera "rig" of the fixture's size
    img = g["gray"] if channels == 1 else g["bgr"]
    h, w = img.shape[:2]
    one = M.RigFrames(M.synth.make_rig(1, w, h), w, h, nfeatures=50, max_frames=1)
    one.process_f32(np.ascontiguousarray(img[None, None]))
    assert np.array_equal(one.level(0, 0, 0), g["gray_u8"] if channels == 1 else g["bgr_u8"])
    with pytest.raises(M.McsError):
        lib_rc = B.lib().mcs_rig_process_f32(one.hd, B._p(img), 2, None, 1, 0)
        B._check(lib_rc, "process_f32")


@pytest.mark.parametrize("old", [False, True])
def test_merge_many_to_one_stress(old):
    """IntraMatch merge on injected features (mcs_rig_set_features) built to make the order-dependent cases frequent: clusters of
    near-duplicate descriptors, so that many queries of one camera pick the same train feature, rows get attached / overwritten by
    later pairs, and inverse-index entries go stale.  6 frames x 5 cameras, ragged counts (incl. an empty camera); rows, landmarks
    and counts bit-exact against the oracle's sequential walk."""
    import mcslam_b200 as M
    import oracle_lib as O
    C, w, h, F = 5, 320, 240, 6
    calib = M.synth.make_rig(C, w, h)
    rig = M.RigFrames(calib, w, h, nfeatures=400, max_frames=F)
    rng = np.random.default_rng(77 + int(old))
    sigma2 = O.Extractor(400).sigma2
    allK, allD, per_frame = [], [], []
    for f in range(F):
        n_centres = [12, 40, 150, 300, 60, 25][f]
        centres = rng.integers(0, 256, (n_centres, 32), dtype=np.uint8)
        K, D = [], []
        for c in range(C):
            n = int(rng.integers(150, rig.kp_cap - 10)) if not (f == 2 and c == 3) else 0
            which = rng.integers(0, n_centres, n)
            d = centres[which].copy()
            for k in range(n):  # 0..24 flipped bits: within the distance gate, ratio test decided by the cluster structure
                bits = rng.choice(256, int(rng.integers(0, 25)), replace=False)
                np.bitwise_xor.at(d[k], bits // 8, (1 << (bits % 8)).astype(np.uint8))
            k = np.zeros(n, M.KP_DTYPE)
            k["x"] = rng.uniform(20, w - 20, n).astype(np.float32)
            k["y"] = rng.uniform(20, h - 20, n).astype(np.float32)
            k["size"] = 31
            k["response"] = rng.integers(7, 200, n)
            k["octave"] = rng.integers(0, 8, n)
            k["class_id"] = -1
            K.append(k)
            D.append(d)
        per_frame.append((K, D))
        allK += K
        allD += D
    rig.set_features(allK, allD)
    rig.compute_intra_matches(F, old=old)
    rig.triangulate(F)
    many_to_one = 0
    for f, (K, D) in enumerate(per_frame):
        rows, pc = O.intra_matches(D, K, calib, sigma2, use_epipolar=old)
        got = rig.intra_matches(f)
        assert got.shape == rows.shape and np.array_equal(got, rows), "frame %d" % f
        tri = O.triangulate_rows(calib, K, D, rows)
        lm = rig.landmarks(f)
        assert np.array_equal(lm["status"], tri["status"]) and np.array_equal(lm["rep_cam"], tri["rep_cam"])
        a, b = rig.brute_force_match(f, 0, 1)[:2]
        many_to_one += len(b) - len(set(b.tolist()))
    assert old or many_to_one > 50, "the fixture must contain many-to-one matches"


def test_merge_config4_capacity():
    """SURVEY config 4 (8 cameras, 1920 x 1200, 4000 features per camera): the handle must come up (the merge kernel's shared tables
    are sized by camera count x feature capacity) and merge 8 x ~4000 injected features exactly like the oracle."""
    import mcslam_b200 as M
    import oracle_lib as O
    C, w, h = 8, 1920, 1200
    calib = M.synth.make_rig(C, w, h, spacing=0.1)
    rig = M.RigFrames(calib, w, h, nfeatures=4000, max_frames=1)
    assert rig.kp_cap >= 4000
    rng = np.random.default_rng(5)
    centres = rng.integers(0, 256, (3000, 32), dtype=np.uint8)
    K, D = [], []
    for c in range(C):
        n = rig.kp_cap - int(rng.integers(0, 200))
        d = centres[rng.integers(0, len(centres), n)].copy()
        flips = rng.integers(0, 256, (n, 32), dtype=np.uint8) & rng.integers(0, 256, (n, 32), dtype=np.uint8) & rng.integers(0, 256, (n, 32), dtype=np.uint8) & rng.integers(0, 256, (n, 32), dtype=np.uint8)
        d ^= flips  # ~16 flipped bits
        k = np.zeros(n, M.KP_DTYPE)
        k["x"] = rng.uniform(20, w - 20, n).astype(np.float32)
        k["y"] = rng.uniform(20, h - 20, n).astype(np.float32)
        k["size"], k["class_id"] = 31, -1
        K.append(k)
        D.append(d)
    rig.set_features(K, D)
    rig.compute_intra_matches(1)
    rows, pc = O.intra_matches(D, K, calib, O.Extractor(4000).sigma2)
    got = rig.intra_matches(0)
    assert len(rows) > 4000 and got.shape == rows.shape and np.array_equal(got, rows)


def test_undistort_rectify_branch(golden):
    """mcs_rig_set_undistort (RECTIFY branch of setData, MultiCameraFrame.cpp:123-136): level 0 equals cv2.undistort (fixture) / the
    oracle bit for bit, masks are undistorted with the same maps, and everything downstream equals the plain path fed with the
    undistorted images and masks.  Host u8, host f32 and device entry points; 5 frames in chunks of 2."""
    import mcslam_b200 as M
    import oracle_lib as O
    from mcslam_b200 import binding as B
    g = golden("undistort_cv2.npz")
    for k in ("1", "2"):  # the cv2 fixture straight through the kernel: 1-camera rig
        cal = M.synth.make_rig(1, 320, 240).copy()
        cal["K"][0] = g["K" + k].reshape(9)
        cal["dist"][0] = g["d" + k]
        one = M.RigFrames(cal, 320, 240, nfeatures=100, max_frames=1)
        one.set_undistort(True)
        one.process(g["img"][None, None], masks=g["mask"][None, None])
        assert np.array_equal(one.level(0, 0, 0), g["img_u" + k]), k
    C, w, h, F = 3, 320, 240, 5
    calib = M.synth.make_rig(C, w, h).copy()
    dists = [[-0.28, 0.11, 0.0012, -0.0007, 0.0], [0.1, 0.04, -0.002, 0.001, -0.01], [0, 0, 0, 0, 0]]
    for c in range(C):
        calib["dist"][c] = dists[c]
    frames = np.stack([M.synth.make_rig_frame(calib, w, h, frame_index=80 + i, n_rects=6) for i in range(F)])
    masks = np.stack([M.synth.make_masks(C, w, h, frame_index=i) for i in range(F)])
    und = np.stack([[O.undistort(frames[f, c], calib["K"][c], calib["dist"][c]) for c in range(C)] for f in range(F)])
    und_m = np.stack([[O.undistort(masks[f, c], calib["K"][c], calib["dist"][c]) for c in range(C)] for f in range(F)])
    assert np.array_equal(und[:, 2], frames[:, 2]), "zero distortion is the identity"
    assert (und[:, 0] != frames[:, 0]).mean() > 0.3
    rig = M.RigFrames(calib, w, h, nfeatures=300, max_frames=F)
    rig.set_chunk_frames(4)
    rig.process(und, masks=und_m)  # plain path on pre-undistorted inputs = the expectation
    ref = [([rig.features(f, c) for c in range(C)], rig.intra_matches(f), rig.landmarks(f)) for f in range(F)]

    def check():
        for f in range(F):
            for c in range(C):
                assert np.array_equal(rig.level(f, c, 0), und[f, c]), (f, c)
                k, d = rig.features(f, c)
                assert (k == ref[f][0][c][0]).all() and np.array_equal(d, ref[f][0][c][1])
            assert np.array_equal(rig.intra_matches(f), ref[f][1])
            lm = rig.landmarks(f)
            for key in lm:
                assert np.array_equal(lm[key], ref[f][2][key]), key

    rig.set_undistort(True)
    rig.process(frames, masks=masks)
    check()
    rig.process(und)  # scramble the state
    rig.set_undistort(True)
    f32 = np.ascontiguousarray(frames.astype(np.float32) / np.float32(255.0))
    rig.process_f32(f32, masks=masks)
    check()
    dptr, dm = B.device_malloc(0, frames.nbytes), B.device_malloc(0, masks.nbytes)
    B.device_upload(0, dptr, frames)
    B.device_upload(0, dm, masks)
    rig.process_device(dptr, F, d_masks=dm)
    rig.synchronize()
    check()
    rig.extract(frames)           # stage-wise entry points
    rig.compute_intra_matches(F)
    rig.triangulate(F, masks=masks)
    check()
    B.device_free(0, dptr)
    B.device_free(0, dm)
    rig.set_undistort(False)
    rig.process(frames, masks=masks)
    assert np.array_equal(rig.level(0, 0, 0), frames[0, 0])


@pytest.mark.parametrize("old", [False, True])
def test_undistort_keypoints_branch(golden, old):
    """mcs_rig_set_undistort_keypoints (RECTIFY = false, UndistortKeyPoints MultiCameraFrame.cpp:300-347): image_kps_undist equals
    cv2.undistortPoints (fixture, through set_features) / the oracle bit for bit; BruteForceMatch returns the undistorted keypoints,
    the `old` epipolar gate runs on them, triangulation keeps the distorted ones."""
    import mcslam_b200 as M
    import oracle_lib as O
    g = golden("undistort_points_cv2.npz")
    for k in ("1", "2"):
        cal = M.synth.make_rig(1, 320, 240).copy()
        cal["K"][0], cal["dist"][0] = g["K" + k].reshape(9), g["d" + k]
        one = M.RigFrames(cal, 320, 240, nfeatures=2000, max_frames=1)
        one.set_undistort_keypoints(True)
        n = min(len(g["pts"]), one.kp_cap)
        kp = np.zeros(n, M.KP_DTYPE)
        kp["x"], kp["y"] = g["pts"][:n, 0], g["pts"][:n, 1]
        one.set_features([kp], [np.zeros((n, 32), np.uint8)])
        u = one.keypoints_undist(0, 0)
        assert np.array_equal(u["x"], g["out" + k][:n, 0]) and np.array_equal(u["y"], g["out" + k][:n, 1])
    C, w, h = 3, 320, 240
    calib = M.synth.make_rig(C, w, h).copy()
    dists = [[-0.28, 0.11, 0.0012, -0.0007, 0.0], [0.5, 0.04, -0.002, 0.001, 0.0], [0.05, -0.01, 0.001, 0.0005, 0.002]]
    for c in range(C):
        calib["dist"][c] = dists[c]
    frames = np.stack([M.synth.make_rig_frame(calib, w, h, frame_index=90 + i, n_rects=6) for i in range(2)])
    rig = M.RigFrames(calib, w, h, nfeatures=300, max_frames=2)
    rig.set_undistort_keypoints(True)
    rig.process(frames, old=old)
    ex = O.Extractor(300)
    for f in range(2):
        K, D = [], []
        for c in range(C):
            _, k, d = O.Extractor(300)(frames[f, c])
            K.append(k)
            D.append(d)
        KU = [O.undistort_keypoints(K[c], calib["K"][c], calib["dist"][c]) for c in range(C)]
        assert np.array_equal(KU[1], K[1]), "0.5 reads as zero through the reference's float access: copied"
        assert (KU[0]["x"] != K[0]["x"]).any()
        for c in range(C):
            k, _ = rig.features(f, c)
            assert (k == K[c]).all() and (rig.keypoints_undist(f, c) == KU[c]).all()
        rows, _ = O.intra_matches(D, KU, calib, ex.sigma2, use_epipolar=old)
        assert np.array_equal(rig.intra_matches(f), rows)
        tri = O.triangulate_rows(calib, K, D, rows)
        lm = rig.landmarks(f)
        assert np.array_equal(lm["status"], tri["status"])
        ok = tri["nviews"] >= 2
        assert np.allclose(lm["xyz"][ok], tri["xyz"][ok], rtol=1e-6, atol=1e-9)
        i1, i2, k1, k2 = rig.brute_force_match(f, 0, 2)
        assert len(i1) > 5 and (k1 == KU[0][i1]).all() and (k2 == KU[2][i2]).all()


def _write_vocabulary(path, k, L, seed, seeds_desc=None):
    from test_gpu_interframe import _random_vocabulary
    rng = np.random.default_rng(seed)
    parent, leaf, desc, weight = _random_vocabulary(rng, k, L)
    if seeds_desc is not None:  # centre the first-level nodes on real descriptors so that the features spread over the tree
        desc[:k] = seeds_desc[:k]
    with open(path, "w") as f:
        f.write("%d %d 0 0\n" % (k, L))
        for p, l, d, w in zip(parent, leaf, desc, weight):
            f.write("%d %d %s %.17g\n" % (p, l, " ".join(str(int(b)) for b in d), w))
    return parent, leaf, desc, weight


def test_bow_best2_equals_bruteforce():
    """mcs_rig_bow_best2 (the per-node best / second-best search of the BoW-guided computeIntraMatches, MultiCameraFrame.cpp:716-751)
    against a numpy restatement on a rig's resident features: ragged groups, an empty cam2 list, the 50-row gate."""
    import mcslam_b200 as M
    calib = M.synth.make_rig(3, 320, 240)
    rig = M.RigFrames(calib, 320, 240, nfeatures=300, max_frames=2)
    rig.process(np.stack([M.synth.make_rig_frame(calib, 320, 240, frame_index=i, n_rects=6) for i in (3, 4)]))
    rng = np.random.default_rng(5)
    for frame in (0, 1):
        feats = [rig.features(frame, c) for c in range(3)]
        kpu = [rig.keypoints_undist(frame, c) for c in range(3)]
        groups = []
        for g in range(40):
            a, b = sorted(rng.choice(3, 2, replace=False).tolist())
            na, nb = int(rng.integers(1, 70)), int(rng.integers(0, 40)) if g != 7 else 0
            groups.append((a, b, rng.choice(len(feats[a][0]), min(na, len(feats[a][0])), replace=False).astype(np.uint32),
                           rng.choice(len(feats[b][0]), min(nb, len(feats[b][0])), replace=False).astype(np.uint32)))
        got = rig.bow_best2(frame, groups)
        n_gated = 0
        for (a, b, ia, ib), (bj, d1, d2) in zip(groups, got):
            Da, Db = feats[a][1][ia], feats[b][1][ib]
            dist = np.unpackbits(Da[:, None, :] ^ Db[None, :, :], axis=2).sum(2).astype(np.int64) if len(ib) else np.zeros((len(ia), 0), np.int64)
            ok = np.abs(kpu[a]["y"][ia][:, None] - kpu[b]["y"][ib][None, :]) < np.float32(50)
            n_gated += int((~ok).sum())
            dist = np.where(ok, dist, 10 ** 9)
            for i in range(len(ia)):
                row = dist[i]
                if len(row) == 0 or row.min() >= 10 ** 9:
                    assert bj[i] == -1 and d1[i] == 10 ** 9 and d2[i] == 10 ** 9
                    continue
                j = int(np.argmin(row))  # first minimum, as the strict '<' of :735
                rest = np.delete(row, j)
                assert (bj[i], d1[i]) == (j, row[j]) and d2[i] == (rest.min() if len(rest) else 10 ** 9)
        assert n_gated > 0
    assert rig.bow_best2(0, []) == []
    with pytest.raises(M.McsError):
        rig.bow_best2(0, [(0, 1, np.array([100000], np.uint32), np.array([0], np.uint32))])


@pytest.mark.parametrize("levelsup", [1, 2])
def test_cpp_adaptor_bow_intra_matches(tmp_path, levelsup):
    """MultiCameraFrame::computeIntraMatches(matches, words_) — the BoW-guided intra-rig matcher FrontEnd runs by default
    (MultiCameraFrame.cpp:586-943) — through the C++ adaptor (vocabulary transform and per-node searches on the GPU, the row bookkeeping
    on the host) against the oracle's restatement: rows, n_rays and words_ identical."""
    import subprocess
    import mcslam_b200 as M
    import oracle_lib as O
    from test_host_side import _build_adaptor
    exe = _build_adaptor(tmp_path)
    calib = M.synth.make_rig(4, 320, 240)
    frame = M.synth.make_rig_frame(calib, 320, 240, frame_index=9, n_rects=7)
    K, D, _, _, _ = _oracle_frame(O, frame, calib, 400)
    (tmp_path / "im.bin").write_bytes(frame.tobytes())
    (tmp_path / "im.bin.calib").write_bytes(np.ascontiguousarray(calib).tobytes())
    k, L = 6, 3
    pick = np.random.default_rng(2).choice(len(D[0]), k, replace=False)
    parent, leaf, desc, weight = _write_vocabulary(tmp_path / "im.bin.voc.txt", k, L, 77, D[0][pick])
    out = subprocess.run([exe, str(tmp_path / "im.bin"), "4", "320", "240", str(tmp_path / "o.txt"), str(levelsup)], capture_output=True,
                         text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    lines = (tmp_path / "o.txt").read_text().splitlines()
    fvs = []
    for c in range(4):
        _, nid, w = O.vocab_transform(L, parent, leaf, desc, weight, D[c], levelsup)
        fvs.append(O.feature_vector(nid, w))
    rows, n_rays, words = O.intra_matches_bow(D, K, fvs)
    got = np.array([list(map(int, l.split()[1:])) for l in lines if l.startswith("bowrow ")], np.int32).reshape(-1, 5)
    assert len(rows) > 30 and (n_rays >= 3).any()
    assert np.array_equal(got[:, 1:], rows) and np.array_equal(got[:, 0], n_rays)
    assert [int(l.split()[1]) for l in lines if l.startswith("bowword ")] == words.tolist()
    # the pairwise BoW variants on the same FeatureVectors: computeIntraMatches_chained (:1290-1355) and the DMatch overload (:1365-1466)
    chained = np.array([list(map(int, l.split()[1:])) for l in lines if l.startswith("chainrow ")], np.int32).reshape(-1, 4)
    exp_chain = O.intra_matches_chained(D, K, fvs)
    assert len(exp_chain) > 20 and np.array_equal(chained, exp_chain)
    dm = {}
    for l in lines:
        if l.startswith("dm "):
            t = l.split()
            dm.setdefault((int(t[1]), int(t[2])), []).append([float(v) for v in t[3:]])
    exp_dm = O.intra_matches_dmatch(D, K, fvs)
    assert sum(len(v[0]) for v in exp_dm.values()) > 60
    for pair, (i1, i2, p1, p2) in exp_dm.items():
        got_p = np.array(dm.get(pair, []), np.float64).reshape(-1, 6)
        assert got_p[:, 0].tolist() == i1.tolist() and got_p[:, 1].tolist() == i2.tolist(), pair
        assert np.array_equal(got_p[:, 2:4].astype(np.float32), p1) and np.array_equal(got_p[:, 4:6].astype(np.float32), p2)


def test_config4_full_size_rig_frame():
    """SURVEY config 4 at its full size — 8 cameras, 1920 x 1200, nfeatures = 4000, 28 pairs — one rig frame through extract, match, merge
    and triangulation against the oracle (maximum sizes: 5200 FAST cells per camera, ~4000 x 4000 Hamming per pair, IntraMatch width 8)."""
    import mcslam_b200 as M
    import oracle_lib as O
    C, w, h = 8, 1920, 1200
    calib = M.synth.make_rig(C, w, h, spacing=0.1)
    frame = M.synth.make_rig_frame(calib, w, h, frame_index=3, n_rects=10)
    rig = M.RigFrames(calib, w, h, nfeatures=4000, max_frames=1)
    rig.process(frame[None])
    K, D, rows, pc, tri = _oracle_frame(O, frame, calib, 4000)
    assert min(len(k) for k in K) > 3000 and len(rows) > 3000
    _check_frame(rig, 0, K, D, rows, tri, C)
