"""CPU checks of the oracle's matcher half (no GPU): known answers, an independent numpy
restatement of the projection search, cv2.BFMatcher for the kNN, and geometric properties of
the stereo matcher on a synthetic pair with known disparities."""
import numpy as np
import pytest


def _popcount_rows(a, b):
    return np.unpackbits(np.asarray(a, np.uint8) ^ np.asarray(b, np.uint8), axis=-1).sum(-1)


def test_descriptor_distance_known_answers(oracle):
    z = np.zeros(32, np.uint8); o = np.full(32, 255, np.uint8)
    assert oracle.descriptor_distance(z, z) == 0 and oracle.descriptor_distance(z, o) == 256
    for bit in (0, 7, 8, 100, 255):
        v = z.copy(); v[bit // 8] |= 1 << (bit % 8)
        assert oracle.descriptor_distance(z, v) == 1 and oracle.descriptor_distance(o, v) == 255
    rng = np.random.default_rng(0)
    a = rng.integers(0, 256, (20000, 32), dtype=np.uint8); b = rng.integers(0, 256, (20000, 32), dtype=np.uint8)
    d = oracle.hamming_pairs(a, b)
    assert np.array_equal(d, _popcount_rows(a, b))          # the SWAR bit-hack (ORBmatcher.cc:1382-1388) == popcount
    assert np.array_equal(d, oracle.hamming_pairs(b, a))


def test_three_maxima_known_answers(oracle):
    c = np.zeros(30, int); assert oracle.three_maxima(c) == (-1, -1, -1)
    c[4] = 10; assert oracle.three_maxima(c) == (4, -1, -1)
    c[7] = 10; assert oracle.three_maxima(c) == (4, 7, -1)            # strict >: first wins
    c[2] = 1;  assert oracle.three_maxima(c) == (4, 7, 2)            # 1 >= 0.1*10
    c[2] = 0; c[9] = 0; assert oracle.three_maxima(c) == (4, 7, -1)
    c[:] = 0; c[1] = 100; c[3] = 9; c[5] = 50
    assert oracle.three_maxima(c) == (1, 5, -1)                       # 9 < 0.1*100 dropped


def test_knn2_matches_cv2_bfmatcher(oracle, synth):
    cv2 = pytest.importorskip("cv2")
    q, db, _ = synth.descriptor_sets(400, 3000, seed=8)
    d1, i1, d2, i2 = oracle.knn2(q, db, nthreads=4)
    m = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(q, db, k=2)
    cd1 = np.array([x[0].distance for x in m]); cd2 = np.array([x[1].distance for x in m])
    ci1 = np.array([x[0].trainIdx for x in m])
    assert np.array_equal(d1, cd1) and np.array_equal(d2, cd2)
    uniq = d1 < d2
    assert np.array_equal(i1[uniq], ci1[uniq])
    # exhaustive definition
    D = _popcount_rows(q[:, None, :], db[None, :, :])
    o = np.argsort(D, axis=1, kind="stable")[:, :2]
    assert np.array_equal(i1, o[:, 0]) and np.array_equal(i2, o[:, 1])


def _sbp_numpy(F, MP, th, nnRatio, claimed=None):
    """Independent restatement of SearchByProjection (reference src/ORBmatcher.cc:31-123): no grid
    structure -- candidates are gathered by brute force and ordered the way the grid walk of
    GetFeaturesInArea (src/Frame.cc:404-468) would deliver them."""
    f32 = np.float32
    n = len(F["kx"])
    px = np.round((F["kx"] - f32(F["minx"])) * f32(F["gw_inv"])).astype(int)     # np.round is half-even; inputs avoid .5
    py = np.round((F["ky"] - f32(F["miny"])) * f32(F["gh_inv"])).astype(int)
    ok_cell = (py * F["gcols"] + px >= 0) & (py * F["gcols"] + px < F["gcols"] * F["grows"])
    cl = np.zeros(n, bool) if claimed is None else np.asarray(claimed, bool).copy()
    ur = F.get("uright")
    obs = MP.get("has_obs")
    match = np.full(n, -1, np.int32)
    nm = 0
    for i in range(len(MP["x"])):
        if not MP["inview"][i]:
            continue
        lvl = int(MP["level"][i])
        r = f32(2.5) if np.float64(MP["viewcos"][i]) > 0.998 else f32(4.0)
        if th != 1.0:
            r = f32(r * f32(th))
        r = f32(r * F["scale"][lvl])
        x, y = MP["x"][i], MP["y"][i]
        c0 = max(0, int(np.floor(f32(f32(x - f32(F["minx"]) - r) * f32(F["gw_inv"])))))
        c1 = min(F["gcols"] - 1, int(np.ceil(f32(f32(x - f32(F["minx"]) + r) * f32(F["gw_inv"])))))
        r0 = max(0, int(np.floor(f32(f32(y - f32(F["miny"]) - r) * f32(F["gh_inv"])))))
        r1 = min(F["grows"] - 1, int(np.ceil(f32(f32(y - f32(F["miny"]) + r) * f32(F["gh_inv"])))))
        cand = np.nonzero(ok_cell & (px >= c0) & (px <= c1) & (py >= r0) & (py <= r1) &
                          (F["koct"] >= lvl - 1) & (F["koct"] <= lvl) &
                          (np.abs(F["kx"] - x) < r) & (np.abs(F["ky"] - y) < r) & ~cl)[0]
        if ur is not None and len(cand):       # :81-86 stereo gate on keypoints with mvuRight > 0
            er = np.abs(MP["proj_xr"][i] - ur[cand])
            cand = cand[~((ur[cand] > 0) & (er > r))]
        if len(cand) == 0:
            continue
        cand = cand[np.lexsort((cand, py[cand], px[cand]))]
        d = _popcount_rows(MP["desc"][i][None], F["desc"][cand])
        best, best2, bl, bl2, bi = 256, 256, -1, -1, -1
        for k, dist in zip(cand, d):
            if dist < best:
                best2, bl2 = best, bl
                best, bl, bi = dist, F["koct"][k], k
            elif dist < best2:
                best2, bl2 = dist, F["koct"][k]
        if best <= 100 and not (bl == bl2 and f32(best) > f32(nnRatio) * f32(best2)):
            match[bi] = i; nm += 1
            cl[bi] = True if obs is None else bool(obs[i])     # :77-79 only a point with observations blocks
    return nm, match


@pytest.mark.parametrize("th,seed", [(1.0, 1), (3.0, 2), (12.0, 3)])
def test_search_by_projection_vs_numpy(oracle, synth, th, seed):
    F = synth.frame_view(500, seed=seed)
    MP = synth.projection_problem(F, 700, seed=seed + 10)
    on, om = oracle.search_by_projection(F, MP, th, False, 0.0, 0.8)
    nn, nmch = _sbp_numpy(F, MP, th, 0.8)
    assert on == nn and np.array_equal(om, nmch) and on > 50


@pytest.mark.parametrize("th,seed", [(1.0, 4), (3.0, 5), (12.0, 6)])
def test_search_by_projection_stereo_vs_numpy(oracle, synth, th, seed):
    """The mvuRight / mTrackProjXR gate (reference src/ORBmatcher.cc:81-86) and the per-point
    Observations() > 0 rule (:77-79): oracle == independent restatement, and both are live."""
    F = synth.frame_view(500, seed=seed)
    MP = synth.projection_problem(F, 900, seed=seed + 10, dup=0.5)
    Fs, MPs = synth.add_stereo(F, MP, seed=seed + 20)
    on, om = oracle.search_by_projection(Fs, MPs, th, False, 0.0, 0.8)
    nn, nmch = _sbp_numpy(Fs, MPs, th, 0.8)
    assert on == nn and np.array_equal(om, nmch) and on > 50
    # the gate rejects something: mono view of the same problem differs
    MPm = dict(MPs); MPm.pop("has_obs")
    gn, gm = oracle.search_by_projection(Fs, MPm, th, False, 0.0, 0.8)
    mn, mm = oracle.search_by_projection(F, MPm, th, False, 0.0, 0.8)
    assert not np.array_equal(gm, mm)
    # the observations rule is live: a point without observations does not block, and its
    # assignment is counted even when a later point takes the keypoint over
    assert on != gn or not np.array_equal(om, gm)
    assert on >= int((om >= 0).sum())


def test_search_for_triangulation_properties(oracle, synth):
    k1, k2, F12, ep, sf = synth.keyframe_pair(600, 700, 50, seed=5)
    n, pairs = oracle.search_for_triangulation(k1, k2, F12, ep, sf, False, False, False)
    assert n == len(pairs) > 20
    d = _popcount_rows(k1["desc"][pairs[:, 0]], k2["desc"][pairs[:, 1]])
    assert d.max() <= 30                                           # TH_LOW
    assert not k1["has_mp"][pairs[:, 0]].any() and not k2["has_mp"][pairs[:, 1]].any()
    # epipolar gate: |y2 - y1 + 0.3 - 0.002 (x2 - x1)| / sqrt(1 + 0.002^2) < sqrt(3.84)
    x1, y1 = k1["kx"][pairs[:, 0]], k1["ky"][pairs[:, 0]]
    x2, y2 = k2["kx"][pairs[:, 1]], k2["ky"][pairs[:, 1]]
    dist2 = (y2 - y1 + 0.3 - 0.002 * (x2 - x1)) ** 2 / (1 + 0.002 ** 2)
    assert dist2.max() < 3.85
    n_ori, pairs_ori = oracle.search_for_triangulation(k1, k2, F12, ep, sf, False, False, True)
    assert n_ori <= n and set(map(tuple, pairs_ori)) <= set(map(tuple, pairs))
    n_c, _ = oracle.search_for_triangulation(k1, k2, F12, ep, sf, False, True, False)
    assert n_c >= n                                                # bCoarse skips the epipolar gate


def test_stereo_matches_recover_known_disparity(oracle, synth):
    """Right image = left shifted by a per-32-row-band disparity: matched keypoints must report it."""
    w, h = 640, 400
    left, right = synth.stereo_pair(w, h, 31)
    rng = np.random.default_rng(31 ^ 0x5EED)                       # stereo_pair's own generator
    band_d = [int(rng.integers(4, 61)) for _ in range(0, h, 32)]
    exL = oracle.Extractor(800, 1.2, 8, 20, 7); exR = oracle.Extractor(800, 1.2, 8, 20, 7)
    kL, dL, _ = exL(left); kR, dR, _ = exR(right)
    lvL = [exL.level(l) for l in range(8)]; lvR = [exR.level(l) for l in range(8)]
    bf, b = 40.0, 0.1                                              # maxD = 400 px
    ur, dp = oracle.stereo_matches(kL, dL, lvL, kR, dR, lvR, exL.scale, exL.inv_scale, bf, b)
    m = ur >= 0
    assert m.sum() > 0.3 * len(kL)
    disp = kL["x"][m] - ur[m]
    assert np.allclose(dp[m], bf / disp, rtol=1e-6)
    # keypoints whose +-8 px SAD rows stay inside one band report that band's disparity (octave-scaled tolerance)
    yy = kL["y"][m]; sc = exL.scale[kL["octave"][m]]
    inside = ((yy - 8 * sc) // 32 == (yy + 8 * sc) // 32)
    expect = np.array(band_d)[(yy // 32).astype(int)]
    err = np.abs(disp - expect)[inside]
    assert inside.sum() > 50 and np.mean(err < 1.0 * sc[inside]) > 0.9
    assert np.all(ur[~m] == -1) and np.all(dp[~m] == -1)


def _grid_numpy(F):
    """Independent restatement of AssignFeaturesToGrid/PosInGrid (reference src/Frame.cc:157-176, 470-480)."""
    ncell = F["gcols"] * F["grows"]
    cells = [[] for _ in range(ncell)]
    for i, (x, y) in enumerate(zip(F["kx"], F["ky"])):
        # C round() on a float expression: half away from zero
        fx = np.float32(np.float32(x - np.float32(F["minx"])) * np.float32(F["gw_inv"]))
        fy = np.float32(np.float32(y - np.float32(F["miny"])) * np.float32(F["gh_inv"]))
        px = int(np.sign(fx) * np.floor(abs(float(fx)) + 0.5)); py = int(np.sign(fy) * np.floor(abs(float(fy)) + 0.5))
        lin = py * F["gcols"] + px
        if 0 <= lin < ncell:
            cells[lin].append(i)
    return cells


def test_frame_grid_vs_numpy(oracle, synth):
    F = synth.frame_view(700, seed=41)
    F["kx"][:5] = [-3.0, 751.9, 760.0, 0.2, 375.5]          # out-of-image x wraps into a neighbouring row's cell
    off, idx = oracle.assign_features_to_grid(F)
    cells = _grid_numpy(F)
    assert off[-1] == sum(len(c) for c in cells)
    for c, lst in enumerate(cells):
        assert list(idx[off[c]:off[c + 1]]) == lst
    # GetFeaturesInArea: brute-force window + level filter must give the same SET; order = cell order
    rng = np.random.default_rng(3)
    nq = 64
    x = rng.uniform(-20, 780, nq).astype(np.float32); y = rng.uniform(-20, 500, nq).astype(np.float32)
    r = rng.uniform(2, 40, nq).astype(np.float32)
    lo = rng.integers(-1, 5, nq).astype(np.int32); hi = np.where(rng.random(nq) < 0.3, -1, lo + rng.integers(0, 3, nq)).astype(np.int32)
    cnt, got = oracle.features_in_area(F, x, y, r, lo, hi, cap=700)
    ingrid = set(int(i) for i in idx[:off[-1]])
    for q in range(nq):
        check = (lo[q] > 0) or (hi[q] >= 0)
        want = set()
        for i in ingrid:
            if check and (F["koct"][i] < lo[q] or (hi[q] >= 0 and F["koct"][i] > hi[q])):
                continue
            if abs(np.float32(F["kx"][i] - x[q])) < r[q] and abs(np.float32(F["ky"][i] - y[q])) < r[q]:
                # the reference only visits cells within the window's cell range
                want.add(i)
        g = set(int(v) for v in got[q, :cnt[q]])
        assert g <= want
        # a keypoint whose rounded cell lies inside the grid sits in a cell of the window's cell range
        for i in want - g:
            fx = np.float32(np.float32(F["kx"][i] - np.float32(F["minx"])) * np.float32(F["gw_inv"]))
            fy = np.float32(np.float32(F["ky"][i] - np.float32(F["miny"])) * np.float32(F["gh_inv"]))
            px, py = int(np.floor(abs(float(fx)) + 0.5) * np.sign(fx)), int(np.floor(abs(float(fy)) + 0.5) * np.sign(fy))
            assert not (0 <= px < F["gcols"] and 0 <= py < F["grows"])


def test_distinctive_descriptor_known_answers(oracle):
    """MapPoint::ComputeDistinctiveDescriptors (reference src/MapPoint.cc:343-412)."""
    z = np.zeros((1, 32), np.uint8)
    assert oracle.distinctive_descriptor(z) == (0, 0)
    assert oracle.distinctive_descriptor(np.zeros((0, 32), np.uint8))[0] == -1
    # N = 2: median index (int)(0.5 * 1) = 0 -> the self distance 0 for both rows, the first wins
    two = np.zeros((2, 32), np.uint8); two[1] = 255
    assert oracle.distinctive_descriptor(two) == (0, 0)
    # a tight cluster of 4 plus an outlier: the winner is in the cluster, and equals a numpy restatement
    rng = np.random.default_rng(8)
    base = rng.integers(0, 256, 32, dtype=np.uint8)
    d = np.stack([base ^ np.uint8(1 << k) for k in range(4)] + [rng.integers(0, 256, 32, dtype=np.uint8)])
    best, med = oracle.distinctive_descriptor(d)
    D = np.unpackbits(d[:, None, :] ^ d[None, :, :], axis=2).sum(2)
    meds = np.sort(D, axis=1)[:, int(0.5 * (len(d) - 1))]
    assert best == int(np.argmin(meds)) and med == int(meds.min()) and best < 4


def test_search_by_bow_properties(oracle, synth):
    """SearchByBoW (reference src/ORBmatcher.cc:133-327): matches only inside shared nodes, only
    KeyFrame points with a MapPoint, each Frame keypoint matched once, planted twins found."""
    K1, K2, _, _, _ = synth.keyframe_pair(800, 900, 60, 31)
    K1["has_mp"] = (np.random.default_rng(1).random(800) < 0.8).astype(np.uint8)
    n, m = oracle.search_by_bow(K1, K2, 0.7, False)
    assert n == int((m >= 0).sum()) and n > 50
    node1 = np.empty(800, np.int64); node2 = np.empty(900, np.int64)
    for K, node in ((K1, node1), (K2, node2)):
        for j in range(len(K["node_id"])):
            node[K["node_idx"][K["node_off"][j]:K["node_off"][j + 1]]] = K["node_id"][j]
    for i2 in np.nonzero(m >= 0)[0]:
        i1 = m[i2]
        assert K1["has_mp"][i1] and node1[i1] == node2[i2]
        d = int(np.unpackbits(K1["desc"][i1] ^ K2["desc"][i2]).sum())
        assert d <= 30            # TH_LOW (include/ORBmatcher.h:73)
    n2, m2 = oracle.search_by_bow(K1, K2, 0.7, True)
    assert n2 <= n and set(np.nonzero(m2 >= 0)[0]) <= set(np.nonzero(m >= 0)[0])


def test_fuse_search_vs_numpy(oracle, synth):
    """Fuse search core (reference src/ORBmatcher.cc:762-824) against a brute-force restatement."""
    F = synth.frame_view(600, seed=61)
    mp, ur, isg = synth.fuse_problem(F, 500, 62)
    bi, bd = oracle.fuse_search(F, mp, 3.0, 0, ur, isg)
    ncell = F["gcols"] * F["grows"]
    hits = 0
    for i in range(len(mp["u"])):
        if not mp["valid"][i]:
            assert bi[i] == -1
            continue
        lvl = mp["level"][i]
        r = np.float32(3.0) * F["scale"][lvl]
        best, bidx = 256, -1
        # the reference visits the cells of the window's cell range (columns, then rows), keypoints in index order
        lo = lambda c, inv: max(0, int(np.floor(np.float32(np.float32(np.float32(c) - r) * np.float32(inv)))))
        hi = lambda c, inv, m: min(m - 1, int(np.ceil(np.float32(np.float32(np.float32(c) + r) * np.float32(inv)))))
        c0, c1 = lo(mp["u"][i], F["gw_inv"]), hi(mp["u"][i], F["gw_inv"], F["gcols"])
        r0, r1 = lo(mp["v"][i], F["gh_inv"]), hi(mp["v"][i], F["gh_inv"], F["grows"])
        cand = []
        for k in range(len(F["kx"])):
            fx = np.float32(np.float32(F["kx"][k] - np.float32(F["minx"])) * np.float32(F["gw_inv"]))
            fy = np.float32(np.float32(F["ky"][k] - np.float32(F["miny"])) * np.float32(F["gh_inv"]))
            px, py = int(np.floor(abs(float(fx)) + 0.5) * np.sign(fx)), int(np.floor(abs(float(fy)) + 0.5) * np.sign(fy))
            lin = py * F["gcols"] + px
            if not (0 <= lin < ncell):
                continue
            if abs(np.float32(F["kx"][k] - mp["u"][i])) < r and abs(np.float32(F["ky"][k] - mp["v"][i])) < r:
                if c0 <= lin % F["gcols"] <= c1 and r0 <= lin // F["gcols"] <= r1:
                    cand.append((lin % F["gcols"], lin // F["gcols"], k))
        for _, _, k in sorted(cand):
            if F["koct"][k] < lvl - 1 or F["koct"][k] > lvl:
                continue
            ex = np.float32(mp["u"][i] - F["kx"][k]); ey = np.float32(mp["v"][i] - F["ky"][k])
            if ur[k] >= 0:
                er = np.float32(mp["ur"][i] - ur[k])
                e2 = np.float32(np.float32(np.float32(ex * ex) + np.float32(ey * ey)) + np.float32(er * er))
                if float(np.float32(e2 * isg[F["koct"][k]])) > 7.8:
                    continue
            else:
                e2 = np.float32(np.float32(ex * ex) + np.float32(ey * ey))
                if float(np.float32(e2 * isg[F["koct"][k]])) > 5.99:
                    continue
            d = int(np.unpackbits(mp["desc"][i] ^ F["desc"][k]).sum())
            if d < best:
                best, bidx = d, k
        assert bi[i] == (bidx if best <= 30 else -1) and bd[i] == (best if bidx >= 0 else -1)
        hits += bi[i] >= 0
    assert hits > 100


def test_search_for_initialization_properties(oracle, synth):
    """SearchForInitialization (reference src/ORBmatcher.cc:329-439): only level-0 keypoints match,
    inside the window, one-to-one, with the stolen keypoints released."""
    F1, F2 = synth.initialization_pair(800, 800, 71)
    n, m = oracle.search_for_initialization(F1, F2, 100, 0.9, False)
    idx1 = np.nonzero(m >= 0)[0]
    assert n == len(idx1) and n > 50
    assert len(set(m[idx1])) == n                                    # vnMatches21 keeps it one-to-one
    assert (F1["koct"][idx1] == 0).all() and (F2["koct"][m[idx1]] == 0).all()
    assert (np.abs(F1["kx"][idx1] - F2["kx"][m[idx1]]) < 100).all() and (np.abs(F1["ky"][idx1] - F2["ky"][m[idx1]]) < 100).all()
    d = np.unpackbits(F1["desc"][idx1] ^ F2["desc"][m[idx1]], axis=1).sum(1)
    assert (d <= 30).all()
    n2, m2 = oracle.search_for_initialization(F1, F2, 100, 0.9, True)
    assert n2 <= n and set(np.nonzero(m2 >= 0)[0]) <= set(idx1)


def test_forb_mean_value_vs_numpy(oracle):
    """DBoW2 FORB::meanValue (reference Thirdparty/DBoW2/src/DBoW2/FORB.cpp:28-79): per-bit majority with the
    N/2 + N%2 threshold, against a plain numpy vote."""
    rng = np.random.default_rng(3)
    for N in (0, 1, 2, 3, 4, 7, 8, 33, 100, 1001):
        d = rng.integers(0, 256, (N, 32), dtype=np.uint8)
        if N >= 4:
            d[:, 0] = 0xF0; d[: N // 2, 1] = 0xFF; d[N // 2:, 1] = 0x00      # unanimous bits, exact ties
        got = oracle.forb_mean_value(d)
        if N == 0:
            assert not got.any()
            continue
        bits = np.unpackbits(d, axis=1).sum(0)
        exp = np.packbits(bits >= N // 2 + N % 2)
        assert np.array_equal(got, exp), N
