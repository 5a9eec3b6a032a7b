"""GPU parity for the inter-frame matcher primitive: batched knnMatch(k=2) through the C ABI (mcs_knn2,
mcs_knn2_device, mcs_rig_knn2_images) and the C++ adaptor's findMatchesMono / findInterMatches, against the cv2 golden fixtures and
the oracle.  Indices and distances bit-exact."""
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_knn2_golden(golden):
    import mcslam_b200 as M
    g = golden("knn2_cv2.npz")
    m = M.Matcher(max_sets=1, max_query_rows=1024, max_train_rows=1024)
    out = m.knn2([g["q"]], [g["t"]])[0]
    assert np.array_equal(out, g["knn"])
    assert m.kernel_launches == 1


def test_knn2_ragged_sets_vs_oracle():
    """Many sets in one launch: empty query set, empty / 1-row / 2-row train sets, a train set longer than the shared-memory chunk
    (1024 rows), duplicates (ties), and the capacity errors."""
    import mcslam_b200 as M
    import oracle_lib as O
    rng = np.random.default_rng(5)
    sizes = [(300, 280), (0, 50), (40, 0), (17, 1), (33, 2), (700, 2600), (129, 1024), (128, 1025), (1, 1)]
    qs = [rng.integers(0, 256, (a, 32), dtype=np.uint8) for a, _ in sizes]
    ts = [rng.integers(0, 256, (b, 32), dtype=np.uint8) for _, b in sizes]
    ts[0][100:140] = ts[0][0:40]
    qs[0][:40] = ts[0][:40]
    ts[5][2000:2050] = ts[5][10:60]
    qs[5][:50] = ts[5][10:60]
    m = M.Matcher(max_sets=len(sizes), max_query_rows=2048, max_train_rows=8192)
    outs = m.knn2(qs, ts)
    for (a, b), q, t, o in zip(sizes, qs, ts, outs):
        assert o.shape == (a, 4)
        assert np.array_equal(o, O.knn2(q, t)), (a, b)
    with pytest.raises(M.McsError):
        m.knn2([qs[5]] * 4, [ts[5]] * 4)  # 2800 query rows > 2048
    with pytest.raises(M.McsError):
        M.Matcher(max_sets=1, max_query_rows=8, max_train_rows=70000).knn2([qs[0][:8]], [np.zeros((65536, 32), np.uint8)])


def test_knn2_device_entry():
    import mcslam_b200 as M
    import oracle_lib as O
    from mcslam_b200 import binding as B
    rng = np.random.default_rng(6)
    q = rng.integers(0, 256, (900, 32), dtype=np.uint8)
    t = rng.integers(0, 256, (1500, 32), dtype=np.uint8)
    dq, dt, do = B.device_malloc(0, q.nbytes), B.device_malloc(0, t.nbytes), B.device_malloc(0, 900 * 16)
    B.device_upload(0, dq, q)
    B.device_upload(0, dt, t)
    m = M.Matcher(max_sets=2)
    assert m.knn2_device(dq, [400, 500], dt, [700, 800], do) == 900
    m.synchronize()
    out = np.zeros((900, 4), np.int32)
    B.device_download(0, out, do)
    assert np.array_equal(out[:400], O.knn2(q[:400], t[:700])) and np.array_equal(out[400:], O.knn2(q[400:], t[700:]))
    for p in (dq, dt, do):
        B.device_free(0, p)


def test_knn2_device_back_to_back_different_layouts():
    """mcs_knn2_device is asynchronous and stages its set table through one pinned buffer: a second call with a different layout,
    queued while the first may still be running, must not disturb the first (h_sets reuse)."""
    import mcslam_b200 as M
    import oracle_lib as O
    from mcslam_b200 import binding as B
    rng = np.random.default_rng(16)
    q = rng.integers(0, 256, (4000, 32), dtype=np.uint8)
    t = rng.integers(0, 256, (6000, 32), dtype=np.uint8)
    dq, dt = B.device_malloc(0, q.nbytes), B.device_malloc(0, t.nbytes)
    B.device_upload(0, dq, q)
    B.device_upload(0, dt, t)
    layouts = [([1000, 3000], [2000, 4000]), ([4000], [6000]), ([100, 200, 300, 3400], [3000, 1000, 1000, 1000]), ([2500, 1500], [100, 5900])]
    m = M.Matcher(max_sets=4)
    outs = [B.device_malloc(0, 4000 * 16) for _ in layouts]
    for rep in range(3):
        for (qc, tc), do in zip(layouts, outs):
            assert m.knn2_device(dq, qc, dt, tc, do) == 4000
        m.synchronize()
        for (qc, tc), do in zip(layouts, outs):
            out = np.zeros((4000, 4), np.int32)
            B.device_download(0, out, do)
            qo = to = 0
            for a, b in zip(qc, tc):
                assert np.array_equal(out[qo:qo + a], O.knn2(q[qo:qo + a], t[to:to + b])), (rep, qc, tc)
                qo += a
                to += b
    for p_ in [dq, dt] + outs:
        B.device_free(0, p_)


def test_rig_knn2_images_consecutive_frames():
    """findMatchesMono's knnMatch for every consecutive frame pair of a batch on the descriptors the rig already holds."""
    import mcslam_b200 as M
    import oracle_lib as O
    calib = M.synth.make_rig(3, 320, 240)
    frames = np.stack([M.synth.make_rig_frame(calib, 320, 240, frame_index=40 + i, n_rects=6) for i in range(3)])
    rig = M.RigFrames(calib, 320, 240, nfeatures=400, max_frames=3)
    rig.process(frames)
    qi = [f * 3 + c for f in range(2) for c in range(3)]
    ti = [(f + 1) * 3 + c for f in range(2) for c in range(3)]
    out = rig.knn2_images(qi, ti)
    for s, (a, b) in enumerate(zip(qi, ti)):
        _, da = rig.features(a // 3, a % 3)
        _, db = rig.features(b // 3, b % 3)
        assert len(da) > 50 and len(db) > 50
        assert np.array_equal(out[s, :len(da)], O.knn2(da, db)), s
        assert (out[s, len(da):] == -1).all()
    with pytest.raises(M.McsError):
        rig.knn2_images([0], [99])


def test_bow_subset_matcher(golden):
    """mcs_match_subsets (getMatches_distRatio for all common vocabulary nodes in one launch) against the fixture and the oracle:
    duplicate claims, nodes missing on one side, single-candidate and empty candidate lists, a 1500-query group."""
    import mcslam_b200 as M
    import oracle_lib as O
    from test_oracle_golden import _bow_fvs
    g = golden("bow_subsets.npz")
    fa, fb = _bow_fvs(g)
    m = M.Matcher(max_sets=1, max_query_rows=4096, max_train_rows=4096)
    a, b, w, k = m.bow_matching(g["A"], g["B"], fa, fb)
    assert np.array_equal(a, g["indices_1"]) and np.array_equal(b, g["indices_2"]) and np.array_equal(w, g["words"]) and k == int(g["book_k"][0])
    rng = np.random.default_rng(3)
    A = rng.integers(0, 256, (1500, 32), dtype=np.uint8)
    B = A[rng.integers(0, 1500, 1200)] ^ (rng.integers(0, 256, (1200, 32), dtype=np.uint8) & rng.integers(0, 256, (1200, 32), dtype=np.uint8)
                                          & rng.integers(0, 256, (1200, 32), dtype=np.uint8) & rng.integers(0, 256, (1200, 32), dtype=np.uint8))
    groups_a = [list(range(1500)), [5], [], [7, 8, 9], list(rng.integers(0, 1500, 40))]
    groups_b = [list(range(1200)), [3], [1, 2], [], list(rng.integers(0, 1200, 64))]
    ma, mb, gc, bk = m.match_subsets(A, B, groups_a, groups_b)
    ea, eb, ec, ek = [], [], [], 0
    for ga, gb in zip(groups_a, groups_b):
        x, y, k2 = O.matches_dist_ratio(A, ga, B, gb)
        ea += x.tolist(); eb += y.tolist(); ec.append(len(x)); ek += k2
    assert ma.tolist() == ea and mb.tolist() == eb and gc.tolist() == ec and bk == ek and ec[0] > 300
    with pytest.raises(M.McsError):
        m.match_subsets(A, B, [[1500]], [[0]])


def test_cpp_adaptor_interframe(tmp_path, golden):
    """mcs_b200::InterFrameMatcher::findMatchesMono / findInterMatches (both overloads) and the pairwise
    MultiCameraFrame::triangulateIntraMatches(kps_1, kps_2) from a C++ program: matches equal the cv2 fixture bit for bit, 3-D points
    equal the oracle's f32 2-view DLT (+ camera-0 transform) within 1e-3 relative (f32 SVD, same bar as triangulate())."""
    import mcslam_b200 as M
    import oracle_lib as O
    from test_host_side import _build_adaptor
    exe = _build_adaptor(tmp_path)
    g = golden("inter_frame_cv2.npz")
    d = tmp_path / "inter"
    d.mkdir()
    for k in ("d1", "d2", "xy1", "xy2", "m1", "m2", "mono1", "mono2", "p1", "p2", "lids1"):
        (d / (k + ".bin")).write_bytes(np.ascontiguousarray(g[k]).tobytes())
    C, npair = 4, 25
    calib = M.synth.make_rig(C, 320, 240)
    (d / "calib.bin").write_bytes(np.ascontiguousarray(calib).tobytes())
    rng = np.random.default_rng(9)
    P = [np.asarray(c["K"]).reshape(3, 3) @ np.hstack([np.asarray(c["R"]).reshape(3, 3), np.asarray(c["t"]).reshape(3, 1)]) for c in calib]
    pairs, expect = [], []
    for i in range(C - 1):
        for j in range(i + 1, C):
            X = np.hstack([rng.uniform(-2, 2, (npair, 2)), rng.uniform(4, 15, (npair, 1)), np.ones((npair, 1))])
            u = (P[i] @ X.T).T
            v = (P[j] @ X.T).T
            p1 = (u[:, :2] / u[:, 2:] + rng.normal(0, 0.2, (npair, 2))).astype(np.float32)
            p2 = (v[:, :2] / v[:, 2:] + rng.normal(0, 0.2, (npair, 2))).astype(np.float32)
            pairs.append(np.hstack([p1, p2]))
            pts = O.triangulate_pair_f32(calib, i, j, p1, p2)
            if i != 0:  # MultiCameraFrame.cpp:458-468
                R = np.asarray(calib[i]["R"]).reshape(3, 3).T
                t = -R @ np.asarray(calib[i]["t"])
                pts = (pts @ R.astype(np.float32).T + t.astype(np.float32)).astype(np.float32)
            expect.append(pts)
    (d / "pairs.bin").write_bytes(np.concatenate(pairs).astype(np.float32).tobytes())
    gb = golden("bow_subsets.npz")
    (d / "bowA.bin").write_bytes(np.ascontiguousarray(gb["A"]).tobytes())
    (d / "bowB.bin").write_bytes(np.ascontiguousarray(gb["B"]).tobytes())
    (d / "bow_nodes.bin").write_bytes(np.concatenate([gb["node_of_A"], gb["node_of_B"]]).astype(np.int32).tobytes())
    vrng = np.random.default_rng(32)
    vp, vl, vd, vw = _random_vocabulary(vrng, 5, 3)
    with open(d / "voc.txt", "w") as f:
        f.write("5 3 0 0\n")
        for p_, l_, d_, w_ in zip(vp, vl, vd, vw):
            f.write("%d %d %s %.17g\n" % (p_, l_, " ".join(str(int(b)) for b in d_), w_))
    (d / "meta.txt").write_text("%d %d %d %d %d %d %d\n" % (len(g["d1"]), len(g["d2"]), g["m1"].shape[1], g["isz"][0], g["isz"][1], C, npair))
    out = subprocess.run([exe, "inter", str(d), str(tmp_path / "o.txt")], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    lines = (tmp_path / "o.txt").read_text().splitlines()

    def rows(tag):
        return np.array([list(map(int, l.split()[1:])) for l in lines if l.startswith(tag + " ")], np.int32).reshape(-1, 3)

    assert np.array_equal(rows("mono"), g["mono_out"])
    assert np.array_equal(rows("lf"), g["lf_split"]) and np.array_equal(rows("mo"), g["mono_split"])
    assert np.array_equal(rows("single"), g["lf_single"])
    assert "empty 0" in lines
    bow = np.array([list(map(int, l.split()[1:])) for l in lines if l.startswith("bow ")], np.uint32).reshape(-1, 2)
    assert np.array_equal(bow[:, 0], gb["indices_1"]) and np.array_equal(bow[:, 1], gb["indices_2"])
    assert [int(l.split()[1]) for l in lines if l.startswith("word ")] == sorted(gb["words"].tolist())
    from test_oracle_golden import _bow_fvs
    fa, fb = _bow_fvs(gb)
    n0 = min(fa)
    ea, eb, ek = O.matches_dist_ratio(gb["A"], fa[n0], gb["B"], fb[n0])
    node0 = np.array([list(map(int, l.split()[1:])) for l in lines if l.startswith("node0 ")], np.uint32).reshape(-1, 2)
    assert node0[:, 0].tolist() == ea.tolist() and node0[:, 1].tolist() == eb.tolist() and ("book %d" % ek) in lines
    _, vnid, vwt = O.vocab_transform(3, vp, vl, vd, vw, gb["A"], 2)
    efv = O.feature_vector(vnid, vwt)
    gfv = {int(l.split()[1]): list(map(int, l.split()[2:])) for l in lines if l.startswith("fv ")}
    assert gfv == efv and len(gfv) >= 3
    got = np.array([[float(x) for x in l.split()[1:]] for l in lines if l.startswith("pt ")], np.float32)
    exp = np.concatenate(expect)
    assert got.shape == exp.shape
    assert np.allclose(got, exp, rtol=1e-3, atol=1e-3), np.abs(got - exp).max()


def _random_vocabulary(rng, k, L, p_early_leaf=0.15):
    """A vocabulary in loadFromTextFile layout: nodes in breadth-first file order, children with near-duplicate descriptors (ties),
    some early leaves, some zero weights."""
    parent, leaf, desc, weight = [], [], [], []
    frontier = [(0, 0, rng.integers(0, 256, 32, dtype=np.uint8))]
    while frontier:
        nxt = []
        for pid, depth, centre in frontier:
            for c in range(k):
                d = centre ^ (rng.integers(0, 256, 32, dtype=np.uint8) & rng.integers(0, 256, 32, dtype=np.uint8))
                if c in (3, 4):
                    d = desc[-1].copy()  # duplicate of the previous sibling: distance ties
                nid = len(parent) + 1
                is_leaf = depth + 1 == L or (depth + 1 >= 2 and rng.random() < p_early_leaf)
                parent.append(pid); leaf.append(int(is_leaf)); desc.append(d)
                weight.append(0.0 if (is_leaf and rng.random() < 0.1) else float(rng.uniform(0.1, 9.0)))
                if not is_leaf:
                    nxt.append((nid, depth + 1, d))
        frontier = nxt
    return np.array(parent, np.int32), np.array(leaf, np.int32), np.array(desc, np.uint8), np.array(weight, np.float64)


def test_vocabulary_transform(tmp_path):
    """DBoW2 transform (word id, node id at level L - levelsup, weight) through mcs_vocab_load_text / mcs_vocab_transform /
    mcs_rig_vocab_transform against the oracle, and the FeatureVector-driven BowMatching on top of it."""
    import mcslam_b200 as M
    import oracle_lib as O
    rng = np.random.default_rng(31)
    k, L = 6, 4
    parent, leaf, desc, weight = _random_vocabulary(rng, k, L)
    path = tmp_path / "voc.txt"
    with open(path, "w") as f:
        f.write("%d %d 0 0\n" % (k, L))
        for p, l, d, w in zip(parent, leaf, desc, weight):
            f.write("%d %d %s %.17g\n" % (p, l, " ".join(str(int(b)) for b in d), w))
    voc = M.Vocabulary(path)
    assert (voc.k, voc.L, voc.n_nodes, voc.n_words) == (k, L, len(parent) + 1, int(leaf.sum()))
    feats = desc[rng.integers(0, len(desc), 3000)] ^ (rng.integers(0, 256, (3000, 32), dtype=np.uint8) & rng.integers(0, 256, (3000, 32), dtype=np.uint8)
                                                     & rng.integers(0, 256, (3000, 32), dtype=np.uint8))
    for levelsup in (0, 2, 4, 6):
        got = voc.transform(feats, levelsup)
        exp = O.vocab_transform(L, parent, leaf, desc, weight, feats, levelsup)
        for a, b in zip(got, exp):
            assert np.array_equal(a, b), levelsup
    assert (exp[2] == 0).any() and len(set(got[1].tolist())) == 1  # stopped words exist; levelsup >= L collapses to the root
    voc2 = M.Vocabulary(k=k, L=L, parent=parent, is_leaf=leaf, node_desc=desc, weight=weight)
    assert all(np.array_equal(a, b) for a, b in zip(voc2.transform(feats, 2), O.vocab_transform(L, parent, leaf, desc, weight, feats, 2)))
    # on a rig's resident descriptors, then BowMatching between two cameras through the FeatureVectors
    calib = M.synth.make_rig(3, 320, 240)
    rig = M.RigFrames(calib, 320, 240, nfeatures=300, max_frames=1)
    rig.process(M.synth.make_rig_frame(calib, 320, 240, frame_index=3, n_rects=6)[None])
    wid, nid, w = voc.transform_rig(rig, 2)
    fvs, D = [], []
    for c in range(3):
        _, d = rig.features(0, c)
        e = O.vocab_transform(L, parent, leaf, desc, weight, d, 2)
        n = len(d)
        assert np.array_equal(wid[c, :n], e[0]) and np.array_equal(nid[c, :n], e[1]) and np.array_equal(w[c, :n], e[2])
        assert (wid[c, n:] == -1).all()
        fvs.append(O.feature_vector(nid[c, :n], w[c, :n]))
        D.append(d)
    m = M.Matcher(max_sets=1, max_query_rows=1024, max_train_rows=1024)
    a, b, words, bk = m.bow_matching(D[0], D[1], fvs[0], fvs[1])
    ea, eb, ew, ek = O.bow_matching(D[0], D[1], fvs[0], fvs[1])
    assert np.array_equal(a, ea) and np.array_equal(b, eb) and np.array_equal(words, ew) and bk == ek
    with pytest.raises(M.McsError):
        M.Vocabulary(tmp_path / "missing.txt")
