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
