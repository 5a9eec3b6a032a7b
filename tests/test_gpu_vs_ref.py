"""GPU parity against oracle/_ref — the REFERENCE's own code compiled here (ORBextractor.cpp whole; MultiCameraFrame::
extractFeaturesParallel / BruteForceMatch / computeIntraMatches / FrontEnd::obtainLfFeatures bodies by line range, see
oracle/ref_build.py) — at BASELINE.json's configurations, through the C ABI.  Keypoints, descriptor bits, pair indices and
IntraMatch rows bit-exact; landmarks 1e-6 relative (f64 N-view DLT)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _ref():
    import ref_lib as R
    if not R.available():
        pytest.skip("oracle/_ref/libmcsref.so missing")
    return R


def _check_against_ref(rig, f, ref, masks=None):
    C = ref.nc
    K, D = [], []
    for c in range(C):
        k, d = rig.features(f, c)
        k2, u2, d2 = ref.features(c)
        assert len(k) == len(k2) and (k == k2).all(), "keypoints cam %d" % c
        assert np.array_equal(d, d2), "descriptor bits cam %d" % c
        K.append(k2)
        D.append(d2)
    for i in range(C - 1):
        for j in range(i + 1, C):
            i1, i2, k1, k2 = rig.brute_force_match(f, i, j, 75, 0.85)
            r1, r2, rk1, rk2 = ref.bf_match(i, j)
            assert np.array_equal(i1, r1) and np.array_equal(i2, r2), "BruteForceMatch pair (%d, %d)" % (i, j)
            assert (k1 == rk1).all() and (k2 == rk2).all()
    rows = ref.intra_matches(False)
    got = rig.intra_matches(f)
    assert got.shape == rows.shape and np.array_equal(got, rows), "IntraMatch rows"
    lf = ref.obtain_lf_features(masks=masks)
    lm = rig.landmarks(f)
    acc = lm["status"] == 1
    n = lf["intramatch_size"]
    assert n == int(acc.sum()), "accepted landmarks: gpu %d reference %d" % (int(acc.sum()), n)
    assert np.array_equal(lm["rows"][acc], lf["match_index"][:n]), "rows after the mask filter"
    assert np.array_equal(lm["nviews"][acc], lf["n_rays"][:n])
    assert np.allclose(lm["xyz"][acc], lf["xyz"][:n], rtol=1e-6, atol=1e-9), np.abs(lm["xyz"][acc] - lf["xyz"][:n]).max()
    assert np.allclose(lm["uv_ref"][acc], lf["uv_ref"][:n], rtol=1e-5, atol=1e-3)
    rep = np.stack([D[c][r[c]] for c, r in zip(lm["rep_cam"][acc], lm["rows"][acc])])
    assert np.array_equal(rep, lf["desc"][:n]), "computeRepresentativeDesc choice"
    return K, D, rows, lf


@pytest.mark.parametrize("nf,with_masks", [(1000, False), (2000, True)])
def test_config2_and_3_vs_reference_code(nf, with_masks):
    """Config 2 (5 x 720 x 540, 1000 features) and config 3 (2000 features, f32 semantic masks)."""
    import mcslam_b200 as M
    R = _ref()
    calib = M.synth.make_rig(5, 720, 540)
    frames = np.stack([M.synth.make_rig_frame(calib, 720, 540, frame_index=i) for i in (2, 9)])
    masks = np.stack([M.synth.make_masks(5, 720, 540, frame_index=i) for i in (2, 9)]) if with_masks else None
    rig = M.RigFrames(calib, 720, 540, nfeatures=nf, max_frames=2)
    rig.process(frames, masks=masks)
    for f in range(2):
        ref = R.Rig(calib, 720, 540, nfeatures=nf)
        ref.extract(frames[f])
        K, D, rows, lf = _check_against_ref(rig, f, ref, masks=None if masks is None else list(masks[f]))
        assert len(rows) > 300 and lf["intramatch_size"] > 200
    if with_masks:
        assert (rig.landmarks(0)["rows"] != rig.intra_matches(0)).any(), "masks must actually remove some views"


def test_epipolar_gate_vs_reference_code():
    """computeIntraMatches(matches, old = true): F from the extrinsics and the 3.84 sigma^2 point-line gate (MultiCameraFrame.cpp:1126-1203)."""
    import mcslam_b200 as M
    R = _ref()
    calib = M.synth.make_rig(5, 720, 540)
    frame = M.synth.make_rig_frame(calib, 720, 540, frame_index=5)
    rig = M.RigFrames(calib, 720, 540, nfeatures=1000, max_frames=1)
    rig.process(frame[None], old=True)
    ref = R.Rig(calib, 720, 540, nfeatures=1000)
    ref.extract(frame)
    rows = ref.intra_matches(True)
    assert np.array_equal(rig.intra_matches(0), rows) and len(rows) < len(ref.intra_matches(False))


def test_single_image_extractor_vs_reference_code():
    """ORBextractor::operator() on pure noise (worst case for FAST / octree), a low-contrast image (minThFAST retry) and the lapping area."""
    import mcslam_b200 as M
    R = _ref()
    rng = np.random.default_rng(2)
    imgs = [M.synth.noise_image(720, 540, 3), (100 + rng.integers(0, 14, (540, 720))).astype(np.uint8)]
    ex = M.Extractor(1000, max_width=720, max_height=540)
    ref = R.Extractor(1000)
    for img in imgs:
        for lap in ((0, 0), (200, 420)):
            mono, k, d = ex(img, lapping=lap)
            m2, k2, d2 = ref(img, lapping=lap)
            assert mono == m2 and len(k) == len(k2) and (k == k2).all() and np.array_equal(d, d2)
            assert len(k) > 100
        for l in range(8):
            assert np.array_equal(ex.level(0, l), ref.level(l)), "pyramid level %d" % l
