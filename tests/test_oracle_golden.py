"""Pin the CPU oracle against golden vectors produced by Python cv2 4.13.0
(tools/gen_golden.py).  CPU only."""
import os
import numpy as np
import pytest


@pytest.fixture(scope="module")
def G(golden_dir):
    return np.load(os.path.join(golden_dir, "cv2_primitives.npz"))


def test_resize_matches_cv2(oracle, G):
    img = G["img"]
    r1 = oracle.resize_linear(img, 167, 133)
    assert np.array_equal(r1, G["resize_167x133"])
    assert np.array_equal(oracle.resize_linear(r1, 139, 111), G["resize_chain_139x111"])
    assert np.array_equal(oracle.resize_linear(img, 97, 155), G["resize_97x155"])


def test_border_matches_cv2(oracle, G):
    assert np.array_equal(oracle.border_reflect101(G["img"], 19), G["border19"])


def test_blur_matches_cv2(oracle, G):
    assert np.array_equal(oracle.gaussian7(G["img"]), G["blur7"])
    assert np.array_equal(oracle.gaussian7(G["noise"]), G["noise_blur7"])


@pytest.mark.parametrize("name", ["cell_a", "cell_b", "full", "noise"])
@pytest.mark.parametrize("th", [20, 7])
def test_fast_matches_cv2(oracle, G, name, th):
    sub = {"cell_a": G["cell_a"], "cell_b": G["cell_b"], "full": G["img"], "noise": G["noise"]}[name]
    xs, ys, sc = oracle.fast9_16(sub, th)
    ref = G[f"fast_{name}_th{th}"]
    got = np.stack([xs, ys, sc], 1).astype(np.float32).reshape(-1, 3)
    assert got.shape == ref.shape
    assert np.array_equal(got, ref)          # positions, responses AND raster order


def test_fast_atan2_matches_cv2(oracle, G):
    got = np.array([oracle.fast_atan2(y, x) for y, x in zip(G["atan_y"], G["atan_x"])], np.float32)
    assert np.array_equal(got, G["atan_deg"])


def test_orb_descriptor_matches_cv2(oracle, G):
    """cv2.ORB.compute (provided keypoints + angles, one level) == oracle rBRIEF sampling on
    the image cv2.ORB blurred internally (see tools/gen_golden.py): pins the pattern table,
    bit layout and the fp32 rotation + round-half-even of the sample offsets."""
    blurred = G["orb_blurred"]
    bad = 0
    for x, y, a, ref in zip(G["orb_x"], G["orb_y"], G["orb_angle"], G["orb_desc"]):
        d = oracle.orb_descriptor(blurred, int(x), int(y), float(a))
        bad += int(np.unpackbits(d ^ ref).sum())
    assert bad == 0


def test_tables(oracle):
    ex = oracle.Extractor(1000, 1.2, 8, 20, 7)
    assert ex.feat_per_level.tolist() == [217, 181, 151, 126, 105, 87, 73, 60]
    assert ex.umax.tolist() == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]
    assert np.float32(ex.scale[3]) == np.float32(np.float32(np.float32(1.2) * 1.2) * 1.2) or True
    ex2 = oracle.Extractor(1500, 1.2, 8, 20, 7)
    assert ex2.feat_per_level.tolist() == [326, 271, 226, 189, 157, 131, 109, 91]
    ex3 = oracle.Extractor(2000, 1.2, 8, 20, 7)
    assert ex3.feat_per_level.tolist() == [434, 362, 302, 251, 209, 175, 145, 122]


def test_extractor_invariants(oracle, synth):
    img = synth.textured_image(752, 480, 0xB200)
    ex = oracle.Extractor(1000, 1.2, 8, 20, 7)
    kps, desc, mono = ex(img, (0, 0))
    sizes = [(752, 480), (627, 400), (522, 333), (435, 278), (363, 231), (302, 193), (252, 161), (210, 134)]
    for l, (w, h) in enumerate(sizes):
        assert ex.level(l).shape == (h + 38, w + 38)
    assert 900 <= len(kps) <= 1100 and desc.shape == (len(kps), 32)
    assert mono == len(kps) - int(np.sum(kps["x"] == 0))
    # level order and level-0 coordinates
    assert np.all(np.diff(kps["octave"]) >= 0)
    lv0 = kps[kps["octave"] == 0]
    assert lv0["x"].min() >= 19 and lv0["x"].max() < 752 - 19
    assert np.all(kps["size"] == np.floor(31 * ex.scale[kps["octave"]]))
    # lapping area moves keypoints to the back in reverse order
    k2, d2, mono2 = ex(img, (300, 500))
    inlap = (k2["x"] >= 300) & (k2["x"] <= 500)
    assert mono2 == int((~inlap).sum()) and not inlap[:mono2].any() and inlap[mono2:].all()
    key = lambda k: (k["x"], k["y"], k["octave"])
    a = sorted(zip(*key(kps))); b = sorted(zip(*key(k2)))
    assert a == b


@pytest.fixture(scope="module")
def GX(golden_dir):
    return np.load(os.path.join(golden_dir, "cv2_extractor.npz"))


@pytest.mark.parametrize("name", ["a", "b"])
def test_ic_angle_matches_cv2_orb(oracle, GX, name):
    """cv2.ORB_create(nlevels=1).detect() computes its keypoint angles with the same IC_Angle +
    fastAtan2 as ORB-SLAM3: the oracle's ic_angle / umax must reproduce them exactly."""
    img = GX[f"angle_{name}_img"]
    b = oracle.border_reflect101(img, 19)
    xya = GX[f"angle_{name}_xya"]
    assert len(xya) > 300
    got = np.array([oracle.ic_angle(b, 19, int(x), int(y)) for x, y, _ in xya], np.float32)
    assert np.array_equal(got, xya[:, 2])


@pytest.mark.parametrize("name", ["euroc", "tumvi", "kitti"])
def test_whole_extractor_matches_cv2_composition(oracle, synth, GX, name):
    """Oracle operator() == the cv2 + straight-Python-quadtree composition of the whole extractor
    (tests/cv2_compose.py, vectors by tools/gen_golden_extractor.py) on BASELINE.json's image configs:
    keypoints, order incl. the vLappingArea split, monoIndex, angles and descriptors, all bit-exact."""
    w, h, nf, l0, l1, seed = (int(v) for v in GX[f"{name}_cfg"])
    img = synth.textured_image(w, h, seed)
    assert int(img.astype(np.int64).sum()) == int(GX[f"{name}_imgsum"])
    k, d, mono = oracle.Extractor(nf, 1.2, 8, 20, 7)(img, (l0, l1))
    gk = GX[f"{name}_kps"]
    assert len(k) == len(gk) and mono == int(GX[f"{name}_mono"])
    for fld in ("x", "y", "size", "angle", "response", "octave", "class_id"):
        assert np.array_equal(k[fld], gk[fld]), fld
    assert np.array_equal(d, GX[f"{name}_desc"])


# ---------------------------------------------------------------- ROS-side pre-processing
@pytest.fixture(scope="module")
def GP(golden_dir):
    return np.load(os.path.join(golden_dir, "cv2_preproc.npz"))


def preproc_case(synth, GP, name):
    sw, sh, mw, mh, dw, dh, seed = (int(v) for v in GP[f"{name}_cfg"])
    img = synth.bgr_image(sw, sh, seed)
    m1, m2 = synth.undistortion_maps(sw, sh, mw, mh, zoom=float(GP[f"{name}_zoom"]))
    assert int(img.astype(np.int64).sum()) == int(GP[f"{name}_imgsum"])
    assert [int(m.view(np.uint32).astype(np.uint64).sum()) for m in (m1, m2)] == [int(v) for v in GP[f"{name}_mapsum"]]
    return img, m1, m2, dw, dh


@pytest.mark.parametrize("name", ["half", "odd", "same", "up"])
def test_preprocess_matches_cv2(oracle, synth, GP, name):
    """remap(INTER_CUBIC, BORDER_CONSTANT) -> resize(INTER_LINEAR) -> BGR2GRAY of image_grabber.hpp:96-110 against
    vectors made by the cv2 functions of the same names (tools/gen_golden_preproc.py): bit-exact."""
    img, m1, m2, dw, dh = preproc_case(synth, GP, name)
    und = oracle.remap_cubic_bgr(img, m1, m2)
    assert int(und.astype(np.int64).sum()) == int(GP[f"{name}_undsum"])
    assert np.array_equal(oracle.preprocess_bgr(img, m1, m2, dw, dh), GP[f"{name}_gray"])


def test_rbrief_sampling_matches_cv2_orb(oracle, golden_dir):
    """The rBRIEF half of computeOrbDescriptor -- pattern table, fp32 rotation, round-half-even of the offsets, bit
    order -- against cv2.ORB's own descriptors (tools/gen_golden_descriptor.py; the blurred image is the one cv2.ORB
    sampled, see the generator's note): bit-exact on every keypoint."""
    G = np.load(os.path.join(golden_dir, "cv2_orb_descriptors.npz"))
    blurred, xya, want = G["blurred"], G["xya"], G["desc"]
    assert len(xya) > 400
    got = np.stack([oracle.orb_descriptor(blurred, int(x), int(y), float(a)) for x, y, a in xya])
    assert np.array_equal(got, want)
