"""GPU parity: ORBextractor::operator() on the B200 (through the C ABI) against the CPU oracle and the committed cv2 golden
vectors.  Keypoint coordinates, octaves, responses, descriptor bits and their ORDER must be bit-exact."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SINGLES = [("single_256x192.npz", 300), ("noise_203x157.npz", 400), ("lowcontrast_320x240.npz", 500)]


def _cand_sorted(c):
    c = np.asarray(c).reshape(-1, 3).astype(np.int64)
    return c[np.lexsort((c[:, 0], c[:, 1]))]


def _compare_stages(ex, orc, image, nlevels=8, img_index=0):
    """Stage-wise report so that one GPU run localises a bug: pyramid -> candidates -> blurred."""
    for l in range(nlevels):
        a, b = ex.level(img_index, l, 0), orc.level(l)
        assert a.shape == b.shape, ("level size", l, a.shape, b.shape)
        assert np.array_equal(a, b), "pyramid level %d: %d px differ" % (l, int((a != b).sum()))
    for l in range(nlevels):
        a, b = _cand_sorted(ex.candidates(img_index, l)), _cand_sorted(orc.candidates(l))
        assert a.shape == b.shape and np.array_equal(a, b), "FAST candidates level %d: gpu %d vs oracle %d" % (l, len(a), len(b))
    for l in range(nlevels):
        b = orc.level(l, blurred=True)
        if b is None:
            continue
        a = ex.level(img_index, l, 2)
        assert np.array_equal(a, b), "blurred level %d: %d px differ" % (l, int((a != b).sum()))


@pytest.mark.parametrize("name,nf", SINGLES)
def test_extract_vs_golden(golden, name, nf):
    import mcslam_b200 as M
    import oracle_lib as O
    g = golden(name)
    img = g["image"]
    ex = M.Extractor(nf, max_width=img.shape[1], max_height=img.shape[0])
    mono, kps, desc = ex(img)
    orc = O.Extractor(nf)
    orc(img)
    t = ex.tables()
    assert (t["quota"] == g["quota"]).all() and (t["scale"] == g["scale"]).all() and (t["sigma2"] == g["sigma2"]).all()
    assert (t["inv_scale"] == g["inv_scale"]).all() and (t["inv_sigma2"] == g["inv_sigma2"]).all()
    _compare_stages(ex, orc, img)
    assert np.array_equal(ex.level(0, 3, 1), g["pyr_padded3"]), "REFLECT_101 frame of level 3"
    assert mono == int(g["mono"]) and len(kps) == len(g["kps"]), (mono, len(kps), len(g["kps"]))
    bad = np.nonzero(kps != g["kps"])[0]
    assert len(bad) == 0, "keypoint mismatches at %s: gpu %s golden %s" % (bad[:5], kps[bad[:2]], g["kps"][bad[:2]])
    nbits = int(np.unpackbits(desc ^ g["desc"]).sum())
    assert nbits == 0, "mismatched descriptor bits: %d" % nbits


@pytest.mark.parametrize("w,h,nf,seed", [(720, 540, 1000, 11), (720, 540, 2000, 12), (640, 480, 800, 13), (333, 251, 350, 14)])
def test_extract_vs_oracle_synthetic(w, h, nf, seed):
    import mcslam_b200 as M
    import oracle_lib as O
    calib = M.synth.make_rig(2, w, h)
    img = M.synth.make_rig_frame(calib, w, h, frame_index=seed, n_rects=8)[1]
    ex = M.Extractor(nf, max_width=w, max_height=h)
    mono, kps, desc = ex(img)
    orc = O.Extractor(nf)
    m2, k2, d2 = orc(img)
    _compare_stages(ex, orc, img)
    assert mono == m2 and len(kps) == len(k2)
    assert (kps == k2).all()
    assert np.array_equal(desc, d2)


def test_noise_worst_case_720():
    """Pure noise: ~38k raw corners at level 0, exercises candidate capacity and octree depth."""
    import mcslam_b200 as M
    import oracle_lib as O
    img = M.synth.noise_image(720, 540, 99)
    ex = M.Extractor(1000)
    mono, kps, desc = ex(img)
    m2, k2, d2 = O.Extractor(1000)(img)
    assert mono == m2 and (kps == k2).all() and np.array_equal(desc, d2)


def test_one_handle_many_image_sizes():
    """ORBextractor::operator() takes whatever image size it is given (the reference rebuilds mvImagePyramid per call): one handle fed
    320x240, then a larger 640x480 (the workspace grows, plan and tensor maps are rebuilt), a width that is not a multiple of 16
    (level 0 re-pitched), and 320x240 again — each call against the oracle, the first and last identical."""
    import mcslam_b200 as M
    import oracle_lib as O
    ex = M.Extractor(500, max_width=320, max_height=240)
    orc = O.Extractor(500)
    first = None
    for w, h, seed in ((320, 240, 1), (640, 480, 2), (411, 300, 3), (320, 240, 1)):
        calib = M.synth.make_rig(2, w, h)
        img = M.synth.make_rig_frame(calib, w, h, frame_index=seed, n_rects=6)[0]
        mono, kps, desc = ex(img)
        m2, k2, d2 = orc(img)
        _compare_stages(ex, orc, img)
        assert mono == m2 and len(kps) == len(k2) and (kps == k2).all() and np.array_equal(desc, d2), (w, h)
        if (w, h, seed) == (320, 240, 1):
            if first is None:
                first = (kps.copy(), desc.copy())
            else:
                assert (kps == first[0]).all() and np.array_equal(desc, first[1])


FUZZ = [  # w, h, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST
    (411, 233, 300, 1.5, 5, 20, 7),
    (640, 480, 800, 2.0, 4, 30, 10),
    (257, 190, 200, 1.1, 10, 15, 5),
    (512, 384, 600, 1.3, 6, 40, 40),
    (300, 300, 150, 1.2, 3, 10, 3),
    (720, 540, 2000, 1.2, 8, 12, 4),
    (1000, 160, 400, 1.25, 4, 25, 9),   # every level keeps at least one 35 px cell row: below that the reference divides by zero
    (640, 480, 2900, 1.2, 1, 12, 4),    # one level holds every feature: its octree tables exceed shared memory (octree_kernel<true>)
    (500, 400, 4500, 2.0, 2, 10, 5),    # level 0 quota 3000: same
]


@pytest.mark.parametrize("w,h,nf,sf,nl,ini,mn", FUZZ)
def test_constructor_parameter_sweep(w, h, nf, sf, nl, ini, mn):
    """ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST) away from the 1000 / 1.2 / 8 / 20 / 7 every other test uses:
    other scale factors (resize tables, level sizes, quotas), level counts, thresholds (equal thresholds switch the retry off), a wide
    flat image (several octree roots), many features at low thresholds — a textured synthetic view blended with noise, stage by stage
    against the oracle."""
    import mcslam_b200 as M
    import oracle_lib as O
    calib = M.synth.make_rig(2, w, h)
    base = M.synth.make_rig_frame(calib, w, h, frame_index=w + h, n_rects=7)[1].astype(np.int32)
    noise = M.synth.noise_image(w, h, 3 * w + nf).astype(np.int32)
    img = np.clip((3 * base + noise) // 4, 0, 255).astype(np.uint8)
    ex = M.Extractor(nf, scale_factor=sf, nlevels=nl, ini_th=ini, min_th=mn, max_width=w, max_height=h)
    orc = O.Extractor(nf, sf, nl, ini, mn)
    mono, kps, desc = ex(img)
    m2, k2, d2 = orc(img)
    t = ex.tables()
    assert (t["quota"] == orc.quota).all() and (t["scale"] == orc.scale).all() and (t["sigma2"] == orc.sigma2).all()
    _compare_stages(ex, orc, img, nlevels=nl)
    assert mono == m2 and len(kps) == len(k2) and (kps == k2).all() and np.array_equal(desc, d2)
    assert len(kps) > nf // 4


@pytest.mark.parametrize("w,h,seed", [(720, 540, 3), (500, 375, 4), (417, 312, 5)])
def test_threshold_fallback_mixed_cells(w, h, seed):
    """Cells without a corner at iniThFAST are re-run at minThFAST (ORBextractor.cpp:842-845) one by one.  Bands of strong texture,
    weak texture (strengths between the two thresholds) and flat grey of widths unrelated to the cell grid put retried and
    non-retried cells next to each other inside one strip of the FAST kernel, and corners of equal score next to each other."""
    import mcslam_b200 as M
    import oracle_lib as O
    rng = np.random.default_rng(seed)
    img = np.full((h, w), 128, np.uint8)
    x = 0
    while x < w:
        bw = int(rng.integers(9, 90))
        kind = int(rng.integers(0, 4))
        if kind == 0:
            band = rng.integers(0, 256, (h, bw))
        elif kind == 1:
            band = 128 + rng.integers(-9, 10, (h, bw))       # weak: most strengths below 20
        elif kind == 2:
            band = 128 + 12 * rng.integers(0, 2, (h, bw))    # two grey values: many equal scores
        else:
            band = np.full((h, bw), 128)
        y = 0
        while y < h:  # flat rows cut the bands so that cell rows differ too
            bh = int(rng.integers(20, 120))
            if rng.integers(0, 3) == 0:
                band[y:y + bh] = 128
            y += bh
        img[:, x:x + bw] = band[:, :max(0, min(bw, w - x))]
        x += bw
    ex = M.Extractor(1500, max_width=w, max_height=h)
    mono, kps, desc = ex(img)
    orc = O.Extractor(1500)
    m2, k2, d2 = orc(img)
    _compare_stages(ex, orc, img)
    assert mono == m2 and len(kps) == len(k2) and (kps == k2).all() and np.array_equal(desc, d2)


@pytest.mark.parametrize("w,h", [(97, 83), (131, 64), (70, 400), (1000, 71), (752, 480), (1226, 370)])
def test_odd_sizes_single_cell_levels(w, h):
    """Sizes whose upper levels shrink to one cell row / column (cells up to 69 px wide), levels too small for any cell, widths that are
    not a multiple of 16 (re-pitched level 0) and one long strip row: noise images against the oracle, stage by stage."""
    import mcslam_b200 as M
    import oracle_lib as O
    img = M.synth.noise_image(w, h, 7 * w + h)
    ex = M.Extractor(300, max_width=w, max_height=h)
    mono, kps, desc = ex(img)
    orc = O.Extractor(300)
    m2, k2, d2 = orc(img)
    _compare_stages(ex, orc, img)
    assert mono == m2 and len(kps) == len(k2) and (kps == k2).all() and np.array_equal(desc, d2)


def test_batch_equals_single_and_resize_of_handle():
    import mcslam_b200 as M
    import oracle_lib as O
    calib = M.synth.make_rig(3, 400, 300)
    imgs = M.synth.make_rig_frame(calib, 400, 300, frame_index=5, n_rects=6)
    ex = M.Extractor(500, max_width=400, max_height=300, max_images=3)
    mono, K, D = ex.extract_batch(imgs)
    for c in range(3):
        m2, k2, d2 = O.Extractor(500)(imgs[c])
        assert mono[c] == m2 and (K[c] == k2).all() and np.array_equal(D[c], d2), "batched image %d" % c
    # same handle, different image size (the reference's extractor accepts any size per call)
    small = np.ascontiguousarray(imgs[0][:200, :301])
    mono, k, d = ex(small)
    m2, k2, d2 = O.Extractor(500)(small)
    assert mono == m2 and (k == k2).all() and np.array_equal(d, d2)
    # strided input (ROI of a larger image)
    roi = imgs[1][10:250, 20:340]
    mono, k, d = ex(roi)
    m2, k2, d2 = O.Extractor(500)(np.ascontiguousarray(roi))
    assert mono == m2 and (k == k2).all() and np.array_equal(d, d2)


def test_edge_cases_gpu():
    import mcslam_b200 as M
    ex = M.Extractor(100, max_width=160, max_height=120)
    mono, k, d = ex(np.zeros((0, 0), np.uint8))
    assert mono == -1  # ORBextractor.cpp:1090-1091
    mono, k, d = ex(np.full((120, 160), 77, np.uint8))
    assert mono == 0 and len(k) == 0
    mono, k, d = ex(np.random.default_rng(0).integers(0, 256, (50, 60), dtype=np.uint8))
    assert mono == 0 and len(k) == 0


def test_lapping_area_gpu(golden):
    import mcslam_b200 as M
    import oracle_lib as O
    g = golden("single_256x192.npz")
    ex = M.Extractor(300, max_width=256, max_height=192)
    mono, k, d = ex(g["image"], lapping=(60, 140))
    m2, k2, d2 = O.Extractor(300)(g["image"], lapping=(60, 140))
    assert mono == m2 and (k == k2).all() and np.array_equal(d, d2)


def test_orientation_flag_mode(golden):
    """IC_Angle + steered BRIEF (dead code in the shipped reference, ORBextractor.cpp:475): checked against the oracle's
    restatement.  Angles (same fastAtan2 polynomial, same operation order) and descriptor bits are bit-exact."""
    import mcslam_b200 as M
    import oracle_lib as O
    g = golden("single_256x192.npz")
    ex = M.Extractor(300, max_width=256, max_height=192)
    ex.set_orientation(True)
    mono, k, d = ex(g["image"])
    orc = O.Extractor(300)
    orc.set_orientation(True)
    m2, k2, d2 = orc(g["image"])
    assert mono == m2
    for f in ("x", "y", "size", "response", "octave", "class_id"):
        assert (k[f] == k2[f]).all(), f
    dang = np.abs(k["angle"] - k2["angle"])
    dang = np.minimum(dang, 360 - dang)
    nbits = int(np.unpackbits(d ^ d2).sum())
    print("orientation mode: max angle diff %.3g deg, mismatched descriptor bits %d of %d" % (dang.max(), nbits, d.size * 8))
    assert np.array_equal(k["angle"], k2["angle"]), dang.max()
    assert nbits == 0
    ex.set_orientation(False)
    mono, k, d = ex(g["image"])
    assert (k == g["kps"]).all() and np.array_equal(d, g["desc"])
