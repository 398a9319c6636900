"""Live cross-check of the oracle against cv2 (skipped when cv2 is not importable):
the per-level FAST candidate lists and the pyramid of a full 752x480 extraction are
re-composed from cv2 primitives."""
import math
import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")


def cv2_candidates(level_roi, ini_th, min_th):
    h, w = level_roi.shape
    minB = 16; maxBX = w - 16; maxBY = h - 16
    width = float(maxBX - minB); height = float(maxBY - minB)
    nCols = int(width / 35); nRows = int(height / 35)
    out = []
    if nCols < 1 or nRows < 1:      # reference divides by zero here; oracle defines: no candidates
        return np.zeros((0, 3), np.float32)
    wCell = math.ceil(width / nCols); hCell = math.ceil(height / nRows)
    det = {t: cv2.FastFeatureDetector_create(t, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16) for t in (ini_th, min_th)}
    for i in range(nRows):
        iniY = minB + i * hCell; maxY = iniY + hCell + 6
        if iniY >= maxBY - 3:
            continue
        maxY = min(maxY, maxBY)
        for j in range(nCols):
            iniX = minB + j * wCell; maxX = iniX + wCell + 6
            if iniX >= maxBX - 6:
                continue
            maxX = min(maxX, maxBX)
            cell = np.ascontiguousarray(level_roi[iniY:maxY, iniX:maxX])
            k = det[ini_th].detect(cell)
            if not k:
                k = det[min_th].detect(cell)
            out += [(p.pt[0] + j * wCell, p.pt[1] + i * hCell, p.response) for p in k]
    return np.array(out, np.float32).reshape(-1, 3)


@pytest.mark.parametrize("shape,seed", [((752, 480), 0xB200), ((512, 512), 7), ((640, 200), 11)])
def test_pyramid_and_candidates_vs_cv2(oracle, synth, shape, seed):
    cv2.setNumThreads(1)
    w, h = shape
    img = synth.textured_image(w, h, seed)
    ex = oracle.Extractor(1000, 1.2, 8, 20, 7)
    ex(img)
    prev = img
    for l in range(8):
        lw = int(np.rint(np.float32(w) * ex.inv_scale[l])); lh = int(np.rint(np.float32(h) * ex.inv_scale[l]))
        cur = img if l == 0 else cv2.resize(prev, (lw, lh), interpolation=cv2.INTER_LINEAR)
        bordered = cv2.copyMakeBorder(cur, 19, 19, 19, 19, cv2.BORDER_REFLECT_101)
        assert np.array_equal(ex.level(l), bordered), f"pyramid level {l}"
        ref = cv2_candidates(cur, 20, 7)
        xs, ys, r = ex.candidates(l)
        got = np.stack([xs, ys, r], 1).astype(np.float32).reshape(-1, 3)
        assert got.shape == ref.shape and np.array_equal(got, ref), f"candidates level {l}"
        prev = cur


def test_low_texture_retry_path(oracle, synth):
    """A faint image forces the per-cell minThFAST retry."""
    img = (synth.textured_image(400, 300, 3).astype(np.float32) * 0.12 + 100).astype(np.uint8)
    ex = oracle.Extractor(500, 1.2, 4, 20, 7)
    ex(img)
    ref20 = cv2.FastFeatureDetector_create(20, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16).detect(img)
    ref = cv2_candidates(img, 20, 7)
    xs, ys, r = ex.candidates(0)
    got = np.stack([xs, ys, r], 1).astype(np.float32).reshape(-1, 3)
    assert len(ref) > 3 * max(len(ref20), 1) or len(ref20) == 0
    assert np.array_equal(got, ref)


@pytest.mark.parametrize("w,h,nf,nl,ini,mn,lap,seed,faint", [
    (640, 400, 800, 8, 20, 7, (0, 0), 41, False),
    (400, 300, 500, 4, 20, 7, (120, 260), 42, True),       # low texture: per-cell minThFAST retry + lapping split
    (333, 257, 400, 5, 12, 5, (0, 0), 43, False),          # odd sizes, other thresholds
    (900, 260, 1500, 6, 20, 7, (0, 0), 44, False),         # wide: several initial quadtree nodes
])
def test_whole_extractor_vs_cv2_composition_live(oracle, synth, w, h, nf, nl, ini, mn, lap, seed, faint):
    """the WHOLE extractor re-composed from cv2 primitives and a straight Python
    quadtree with the real std::sort (tests/cv2_compose.py) equals the oracle bit for bit."""
    import os, sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import cv2_compose as cc
    img = synth.textured_image(w, h, seed)
    if faint:
        img = (img.astype(np.float32) * 0.12 + 100).astype(np.uint8)
    ck, cd, cmono = cc.extract(img, nf, 1.2, nl, ini, mn, lap)
    ok, od, omono = oracle.Extractor(nf, 1.2, nl, ini, mn)(img, lap)
    assert len(ck) == len(ok) > 50 and cmono == omono
    for fld in ("x", "y", "size", "angle", "response", "octave", "class_id"):
        assert np.array_equal(ck[fld], ok[fld]), fld
    assert np.array_equal(cd, od)


def test_python_quadtree_vs_oracle_quadtree(oracle):
    """The straight Python DistributeOctTree against the oracle's on random tie-heavy candidate sets
    (few distinct responses, clustered points: many equal (size, UL.x) nodes in the final phase)."""
    import os, sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import cv2_compose as cc
    rng = np.random.default_rng(5)
    for it in range(40):
        W = int(rng.integers(100, 900)); H = int(rng.integers(80, 500)); N = int(rng.integers(5, 400))
        n = int(rng.integers(1, 3000))
        pts = np.unique(np.stack([rng.integers(0, W, n) // rng.integers(1, 4) * 1, rng.integers(0, H, n)], 1), axis=0)
        pts = pts[rng.permutation(len(pts))]
        resp = rng.integers(7, 12, len(pts))
        sel = cc.distribute_octtree([(float(x), float(y), float(r)) for (x, y), r in zip(pts, resp)], 16, 16 + W, 16, 16 + H, N)
        idx = oracle.distribute_octtree(pts[:, 0], pts[:, 1], resp, 16, 16 + W, 16, 16 + H, N)
        got = [(int(x), int(y)) for x, y, _ in sel]
        assert got == [tuple(int(v) for v in pts[i]) for i in idx], it


def test_preprocess_reference_camera_vs_cv2(oracle, synth):
    """The reference node's own shape (mono_inertial_node.cpp:20,59-68: 2048x1536 BGR, fisheye maps, resize_factor 0.3)
    through cv2.remap / resize / cvtColor against the oracle's restatement: bit-exact."""
    import cv2_compose as cc
    cv2.setNumThreads(1)
    img = synth.bgr_image(2048, 1536, 21)
    m1, m2 = cc.fisheye_maps(2048, 1536)
    w, h = int(2048 * 0.3), int(1536 * 0.3)
    assert np.array_equal(oracle.preprocess_bgr(img, m1, m2, w, h), cc.convert_image(img, m1, m2, w, h))
