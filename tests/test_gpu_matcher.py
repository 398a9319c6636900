"""GPU parity: CUDA matcher / stereo kernels (through the C ABI) vs the CPU oracle on identical
seeded inputs.  Match indices, distances, counts: bit-exact (BASELINE.json north_star)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def matcher(pkg):
    m = pkg.orbx.ORBmatcher()
    yield m
    m.close()


# ---------------------------------------------------------------- SearchByProjection (map points)
@pytest.mark.parametrize("n,nmp,th,nnRatio,seed", [
    (1000, 1500, 1.0, 0.8, 1), (1000, 2000, 3.0, 0.75, 2), (2000, 3000, 15.0, 0.9, 3), (300, 50, 1.0, 0.8, 4),
    (1, 5, 3.0, 0.8, 5),
])
def test_search_by_projection(matcher, oracle, synth, n, nmp, th, nnRatio, seed):
    F = synth.frame_view(n, seed=seed)
    MP = synth.projection_problem(F, nmp, seed=seed + 100)
    rng = np.random.default_rng(seed)
    claimed = (rng.random(n) < 0.1).astype(np.uint8)
    for far, cl in ((False, None), (True, claimed)):
        on, om = oracle.search_by_projection(F, MP, th, far, 25.0, nnRatio, cl)
        gn, gm = matcher.SearchByProjection(F, MP, th, far, 25.0, nnRatio, cl)
        assert gn == on and np.array_equal(gm, om)
        if n > 100:
            assert on > 0.2 * min(n, nmp)          # the problem is not vacuous


def test_search_by_projection_chained_claims(matcher, oracle, synth):
    """Many map points with IDENTICAL descriptors around the same keypoints: long claim chains
    (map point i takes what i-1 left), the worst case for the fixed-point rounds."""
    F = synth.frame_view(400, seed=9)
    rng = np.random.default_rng(9)
    base = rng.integers(0, 256, 32, dtype=np.uint8)
    F["desc"][:] = base
    F["desc"] ^= np.packbits(rng.random((400, 256)) < 0.02, axis=1, bitorder="little")
    F["kx"][:] = rng.uniform(300, 360, 400).astype(np.float32); F["ky"][:] = rng.uniform(200, 240, 400).astype(np.float32)
    F["koct"][:] = 2
    nmp = 600
    MP = {"inview": np.ones(nmp, np.uint8), "x": rng.uniform(300, 360, nmp).astype(np.float32),
          "y": rng.uniform(200, 240, nmp).astype(np.float32), "level": np.full(nmp, 2, np.int32),
          "viewcos": np.ones(nmp, np.float32), "depth": np.ones(nmp, np.float32),
          "desc": np.tile(base, (nmp, 1))}
    on, om = oracle.search_by_projection(F, MP, 10.0, False, 0.0, 1.0)
    gn, gm = matcher.SearchByProjection(F, MP, 10.0, False, 0.0, 1.0)
    assert gn == on and np.array_equal(gm, om) and on >= 300


@pytest.mark.parametrize("n,nmp,th,seed", [(1000, 1500, 1.0, 11), (1000, 2500, 3.0, 12), (600, 1500, 12.0, 13)])
def test_search_by_projection_stereo_gate(matcher, oracle, synth, n, nmp, th, seed):
    """Stereo frames (mvuRight > 0): the |mTrackProjXR - mvuRight[idx]| <= r gate of reference
    src/ORBmatcher.cc:81-86, and the per-point Observations() > 0 rule of :77-79 (a point without
    observations assigns but does not block; every assignment counts)."""
    F = synth.frame_view(n, seed=seed)
    MP = synth.projection_problem(F, nmp, seed=seed + 100, dup=0.5)
    Fs, MPs = synth.add_stereo(F, MP, seed=seed + 200)
    on, om = oracle.search_by_projection(Fs, MPs, th, False, 0.0, 0.8)
    gn, gm = matcher.SearchByProjection(Fs, MPs, th, False, 0.0, 0.8)
    assert gn == on and np.array_equal(gm, om)
    # gate only (every point has observations)
    MPg = dict(MPs); MPg.pop("has_obs")
    on2, om2 = oracle.search_by_projection(Fs, MPg, th, False, 0.0, 0.8)
    gn2, gm2 = matcher.SearchByProjection(Fs, MPg, th, False, 0.0, 0.8)
    assert gn2 == on2 and np.array_equal(gm2, om2)
    # the gate rejects at least one candidate the mono search accepts; mono results are unchanged
    on3, om3 = oracle.search_by_projection(F, MPg, th, False, 0.0, 0.8)
    gn3, gm3 = matcher.SearchByProjection(F, MPg, th, False, 0.0, 0.8)
    assert gn3 == on3 and np.array_equal(gm3, om3)
    assert not np.array_equal(om2, om3)
    assert not np.array_equal(om, om2) or on != on2           # the observations rule is live too


def test_search_by_projection_takeover_chain(matcher, oracle, synth):
    """Identical descriptors, half of the map points without observations: long chains of
    assignments that later points take over."""
    F = synth.frame_view(300, seed=19)
    rng = np.random.default_rng(19)
    base = rng.integers(0, 256, 32, dtype=np.uint8)
    F["desc"][:] = base
    F["desc"] ^= np.packbits(rng.random((300, 256)) < 0.02, axis=1, bitorder="little")
    F["kx"][:] = rng.uniform(300, 360, 300).astype(np.float32); F["ky"][:] = rng.uniform(200, 240, 300).astype(np.float32)
    F["koct"][:] = 2
    nmp = 700
    MP = {"inview": np.ones(nmp, np.uint8), "x": rng.uniform(300, 360, nmp).astype(np.float32),
          "y": rng.uniform(200, 240, nmp).astype(np.float32), "level": np.full(nmp, 2, np.int32),
          "viewcos": np.ones(nmp, np.float32), "depth": np.ones(nmp, np.float32),
          "desc": np.tile(base, (nmp, 1)), "has_obs": (rng.random(nmp) < 0.5).astype(np.uint8)}
    on, om = oracle.search_by_projection(F, MP, 10.0, False, 0.0, 1.0)
    gn, gm = matcher.SearchByProjection(F, MP, 10.0, False, 0.0, 1.0)
    assert gn == on and np.array_equal(gm, om)
    assert on > int((om >= 0).sum())                          # some assignments were taken over


def test_search_by_projection_resident(pkg, matcher, oracle, synth):
    """Device-resident chaining: extract stereo frames, ComputeStereoMatches, then SearchByProjection (both overloads)
    straight on the extractor's device buffers -- keypoints, descriptors, counts and mvuRight never visit the host.
    Must equal the oracle (and the host-array entry point) on the same data, incl. the mvuRight gate."""
    ox = pkg.orbx
    w, h, B = 640, 400, 6
    sp = [synth.stereo_pair(w, h, 500 + i) for i in range(3)]
    frames = np.stack([im for pr in sp for im in pr])
    ex = ox.ORBextractor(800, 1.2, 8, 20, 7, max_batch=B)
    res = ex.extract_batch(frames)
    lf, rf = np.array([0, 2, 4]), np.array([1, 3, 5])
    ox.stereo_matches_async(ex, ex, lf, rf, 40.0, 0.1)
    ur, _ = ox.stereo_fetch(ex, 3)
    grid = {"minx": 0.0, "miny": 0.0, "gw_inv": float(np.float32(64) / np.float32(w)), "gh_inv": float(np.float32(48) / np.float32(h)),
            "gcols": 64, "grows": 48, "scale": ex.mvScaleFactor}
    rng = np.random.default_rng(17)
    Fs, MPs, cls = [], [], []
    for p, f in enumerate(lf):
        k, d, _ = res[f]
        F = dict(grid, kx=k["x"].copy(), ky=k["y"].copy(), koct=k["octave"].copy(), kangle=k["angle"].copy(), desc=d.copy(),
                 uright=ur[p][:len(k)].copy())
        MP = synth.projection_problem(F, 1200, seed=600 + p, dup=0.4)
        tgt = np.where(F["uright"][MP["src"]] > 0, F["uright"][MP["src"]], F["kx"][MP["src"]] - 20.0)
        far = (rng.random(1200) < 0.25) * rng.uniform(30.0, 120.0, 1200)
        MP["proj_xr"] = (tgt + rng.normal(0, 3.0, 1200) + far).astype(np.float32)
        MP["has_obs"] = (rng.random(1200) < 0.85).astype(np.uint8)
        Fs.append(F); MPs.append(MP)
        c = np.zeros(ex.capacity, np.uint8); c[:len(k)] = rng.random(len(k)) < 0.05
        cls.append(c)
        assert (F["uright"] > 0).sum() > 50
    # stereo frames
    nm, outs = matcher.SearchByProjectionResident(ex, lf, grid, MPs, 3.0, False, 0.0, 0.8, stereo_pairs=[0, 1, 2], claimed=cls)
    assert matcher.last_launches() == 1
    for p in range(3):
        n = len(Fs[p]["kx"])
        on, om = oracle.search_by_projection(Fs[p], MPs[p], 3.0, False, 0.0, 0.8, cls[p][:n])
        assert nm[p] == on and np.array_equal(outs[p][:n], om) and np.all(outs[p][n:] == -1) and on > 100
        hn, hm = matcher.SearchByProjection(Fs[p], MPs[p], 3.0, False, 0.0, 0.8, cls[p][:n])
        assert hn == on and np.array_equal(hm, om)
    # the same frames searched as monocular ones: the gate is off, results differ
    nm2, outs2 = matcher.SearchByProjectionResident(ex, lf, grid, MPs, 3.0, False, 0.0, 0.8)
    for p in range(3):
        n = len(Fs[p]["kx"])
        Fm = dict(Fs[p]); Fm.pop("uright")
        on, om = oracle.search_by_projection(Fm, MPs[p], 3.0, False, 0.0, 0.8)
        assert nm2[p] == on and np.array_equal(outs2[p][:n], om)
    assert any(not np.array_equal(a, b) for a, b in zip(outs, outs2))
    # KeyFrame overload on the right frames
    KPs, Fr = [], []
    for p, f in enumerate(rf):
        k, d, _ = res[f]
        F = dict(grid, kx=k["x"].copy(), ky=k["y"].copy(), koct=k["octave"].copy(), kangle=k["angle"].copy(), desc=d.copy())
        P = synth.projection_problem(F, 1000, seed=700 + p)
        KPs.append({"valid": P["valid"], "u": P["x"], "v": P["y"], "level": P["level"], "angle": P["angle"], "desc": P["desc"]})
        Fr.append(F)
    nm3, outs3 = matcher.SearchByProjectionKFResident(ex, rf, grid, KPs, 15.0, True)
    for p in range(3):
        n = len(Fr[p]["kx"])
        on, om = oracle.search_by_projection_kf(Fr[p], KPs[p], 15.0, True)
        assert nm3[p] == on and np.array_equal(outs3[p][:n], om) and on > 50
    ex.close()


def test_search_by_projection_batched(matcher, oracle, synth):
    Fs = [synth.frame_view(600 + 50 * i, seed=20 + i) for i in range(7)]
    MPs = [synth.projection_problem(F, 900 + 10 * i, seed=40 + i) for i, F in enumerate(Fs)]
    nm, outs = matcher.SearchByProjection(Fs, MPs, 3.0, False, 0.0, 0.8)
    for F, MP, n_, o_ in zip(Fs, MPs, nm, outs):
        on, om = oracle.search_by_projection(F, MP, 3.0, False, 0.0, 0.8)
        assert n_ == on and np.array_equal(o_, om)
    assert matcher.last_launches() == 1


# ---------------------------------------------------------------- SearchByProjection (KeyFrame)
@pytest.mark.parametrize("n,nmp,th,ori,seed", [(1000, 1200, 15.0, True, 1), (1000, 1200, 15.0, False, 2),
                                               (1500, 2500, 40.0, True, 3), (64, 10, 5.0, True, 4)])
def test_search_by_projection_kf(matcher, oracle, synth, n, nmp, th, ori, seed):
    F = synth.frame_view(n, seed=seed + 50)
    P = synth.projection_problem(F, nmp, seed=seed + 150)
    MP = {"valid": P["valid"], "u": P["x"], "v": P["y"], "level": P["level"], "angle": P["angle"], "desc": P["desc"]}
    rng = np.random.default_rng(seed)
    occ = (rng.random(n) < 0.05).astype(np.uint8)
    on, om = oracle.search_by_projection_kf(F, MP, th, ori, occ)
    gn, gm = matcher.SearchByProjectionKF(F, MP, th, ori, occ)
    assert gn == on and np.array_equal(gm, om)
    if n > 100:
        assert on > 50
        if ori:      # the histogram filter removed something
            on2, _ = oracle.search_by_projection_kf(F, MP, th, False, occ)
            assert on2 > on


# ---------------------------------------------------------------- SearchForTriangulation
@pytest.mark.parametrize("stereo,coarse,ori,seed", [(False, False, True, 1), (False, True, True, 2),
                                                    (True, False, False, 3), (False, False, False, 4)])
def test_search_for_triangulation(matcher, oracle, synth, stereo, coarse, ori, seed):
    k1, k2, F12, ep, sf = synth.keyframe_pair(1000, 1100, 90, seed=seed)
    on, op = oracle.search_for_triangulation(k1, k2, F12, ep, sf, stereo, coarse, ori)
    gn, gp = matcher.SearchForTriangulation(k1, k2, F12, ep, sf, stereo, coarse, ori)
    assert gn == on and np.array_equal(gp, op)
    assert len(op) > 20


def test_search_for_triangulation_mono_batched(matcher, oracle, synth):
    probs = [synth.keyframe_pair(500 + 40 * i, 600, 60, seed=30 + i) for i in range(5)]
    for p in probs:             # mono KeyFrames: no mvuRight at all
        p[0]["uright"] = None; p[1]["uright"] = None
    nm, res = matcher.SearchForTriangulation([p[0] for p in probs], [p[1] for p in probs], [p[2] for p in probs],
                                             [p[3] for p in probs], [p[4] for p in probs])
    for p, n_, r_ in zip(probs, nm, res):
        on, op = oracle.search_for_triangulation(p[0], p[1], p[2], p[3], p[4])
        assert n_ == on and np.array_equal(r_, op)


def test_three_maxima(pkg, oracle):
    rng = np.random.default_rng(3)
    cases = [rng.integers(0, 50, 30) for _ in range(50)] + [np.zeros(30, int), np.r_[100, np.zeros(29, int)],
                                                            np.r_[100, 9, 9, np.zeros(27, int)], np.r_[100, 10, 9, np.zeros(27, int)]]
    for c in cases:
        assert pkg.orbx.three_maxima(c) == oracle.three_maxima(c)


def test_forb_mean_value(matcher, oracle):
    """DBoW2 FORB::meanValue batched on the GPU (bit-sliced counters) == the oracle, incl. empty / single / tie sets."""
    rng = np.random.default_rng(8)
    sizes = [0, 1, 2, 3, 4, 5, 8, 31, 32, 33, 64, 100, 257, 1000, 5000, 0, 7]
    sets = [rng.integers(0, 256, (n, 32), dtype=np.uint8) for n in sizes]
    for d in sets:
        if len(d) >= 4:
            d[: len(d) // 2, 1] = 0xFF; d[len(d) // 2:, 1] = 0x00            # exact ties on even N
    sets[13][:, 5] = 0x3C                                                      # unanimous
    got = matcher.MeanValue(sets)
    for s_, d in enumerate(sets):
        assert np.array_equal(got[s_], oracle.forb_mean_value(d)), (s_, len(d))
    assert matcher.last_launches() == 1


# ---------------------------------------------------------------- brute-force 2-NN
@pytest.mark.parametrize("nq,ndb,variant", [(3000, 20000, 1), (3000, 20000, 0), (1, 1, 1), (5, 2, 1), (1030, 257, 1), (7, 0, 1)])
def test_knn2_vs_oracle(matcher, oracle, synth, nq, ndb, variant):
    q, db, _ = synth.descriptor_sets(nq, max(ndb, 1), seed=5)
    db = db[:ndb]
    if ndb > 10:
        db[7] = db[3]; q[0] = db[3]          # exact duplicates: tie -> lowest index
    matcher.set_knn_variant(variant)
    d1, i1, d2, i2, ps = matcher.knn2(q, db, 0.8)
    o1, oi1, o2, oi2 = oracle.knn2(q, db, nthreads=8)
    assert np.array_equal(d1, o1) and np.array_equal(i1, oi1) and np.array_equal(d2, o2) and np.array_equal(i2, oi2)
    assert np.array_equal(ps, d1.astype(np.float32) < np.float32(0.8) * d2.astype(np.float32))
    matcher.set_knn_variant(1)


def test_knn2_large_properties(matcher, synth):
    """Sizes the oracle cannot finish: planted neighbours are found (also across the 4M-row slice
    boundary of the kernel's grid), ties resolve to the lowest index, d1 <= d2."""
    nq, ndb = 4096, (1 << 22) + 5000
    q, db, src = synth.descriptor_sets(nq, ndb, seed=6, planted=0.5, flip_p=0.04)
    q[4000] = db[(1 << 22) + 17]; q[4001] = db[(1 << 22) - 1]
    db[(1 << 22) + 100] = db[50]; q[4002] = db[50]
    d1, i1, d2, i2, ps = matcher.knn2(q, db, 0.8)
    npl = nq // 2
    host_d = np.unpackbits(q[:npl] ^ db[src], axis=1).sum(1)
    assert np.all(d1[:npl] <= host_d) and np.mean(i1[:npl] == src) > 0.999
    assert np.all(d1 <= d2) and np.all(i1 != i2)
    assert (d1[4000], i1[4000]) == (0, (1 << 22) + 17) and (d1[4001], i1[4001]) == (0, (1 << 22) - 1)
    assert (d1[4002], i1[4002], d2[4002], i2[4002]) == (0, 50, 0, (1 << 22) + 100)
    # spot-check 16 queries exhaustively on the host
    for k in range(0, nq, nq // 16):
        d = np.unpackbits(q[k][None] ^ db, axis=1).sum(1)
        o = np.lexsort((np.arange(ndb), d))[:2]
        assert (d1[k], i1[k], d2[k], i2[k]) == (d[o[0]], o[0], d[o[1]], o[1])
    assert ps[:npl].mean() > 0.95 and ps[npl:4000].mean() < 0.05


# ---------------------------------------------------------------- Frame::ComputeStereoMatches
def _stereo_case(pkg, oracle, synth, w, h, nf, seed, same_handle):
    left, right = synth.stereo_pair(w, h, seed)
    ox = pkg.orbx
    if same_handle:
        ex = ox.ORBextractor(nf, 1.2, 8, 20, 7, max_batch=2)
        res = ex.extract_batch(np.stack([left, right]))
        exL = exR = ex
        fL, fR = 0, 1
    else:
        exL = ox.ORBextractor(nf, 1.2, 8, 20, 7); exR = ox.ORBextractor(nf, 1.2, 8, 20, 7)
        res = [exL(left, (0, 0)), exR(right, (0, 0))]
        fL = fR = 0
    (kL, dL, _), (kR, dR, _) = res
    bf, b = 435.2 * 0.11, 0.11          # EuRoC: fx * baseline, baseline
    ur, dp = ox.stereo_matches(exL, exR, [fL], [fR], bf, b)
    lvL = [exL.image_pyramid(l, fL) for l in range(8)]
    lvR = [exR.image_pyramid(l, fR) for l in range(8)]
    our, odp = oracle.stereo_matches(kL, dL, lvL, kR, dR, lvR, exL.mvScaleFactor, exL.mvInvScaleFactor, bf, b)
    n = len(kL)
    assert np.array_equal(ur[0, :n], our) and np.array_equal(dp[0, :n], odp)
    assert np.all(ur[0, n:] == -1)
    nmatch = int((our >= 0).sum())
    assert nmatch > 0.2 * n, (nmatch, n)
    exL.close()
    if exR is not exL:
        exR.close()
    return nmatch


def test_stereo_matches_euroc_same_handle(pkg, oracle, synth):
    _stereo_case(pkg, oracle, synth, 752, 480, 1000, 0xB200, True)


def test_stereo_matches_two_extractors(pkg, oracle, synth):
    _stereo_case(pkg, oracle, synth, 752, 480, 1000, 77, False)


def test_stereo_matches_kitti(pkg, oracle, synth):
    _stereo_case(pkg, oracle, synth, 1241, 376, 2000, 11, True)


def test_stereo_matches_batch(pkg, oracle, synth):
    """4 pairs per launch, frames interleaved L,R,L,R,... in one extractor batch."""
    ox = pkg.orbx
    pairs = [synth.stereo_pair(640, 400, 100 + i) for i in range(4)]
    imgs = np.stack([im for p in pairs for im in p])
    ex = ox.ORBextractor(800, 1.2, 8, 20, 7, max_batch=8)
    res = ex.extract_batch(imgs)
    ur, dp = ox.stereo_matches(ex, ex, [0, 2, 4, 6], [1, 3, 5, 7], 40.0, 0.1)
    for p in range(4):
        kL, dL, _ = res[2 * p]; kR, dR, _ = res[2 * p + 1]
        lvL = [ex.image_pyramid(l, 2 * p) for l in range(8)]; lvR = [ex.image_pyramid(l, 2 * p + 1) for l in range(8)]
        our, odp = oracle.stereo_matches(kL, dL, lvL, kR, dR, lvR, ex.mvScaleFactor, ex.mvInvScaleFactor, 40.0, 0.1)
        assert np.array_equal(ur[p, :len(kL)], our) and np.array_equal(dp[p, :len(kL)], odp)
    ex.close()


# ------------------------------------------------------------------ Frame grid
@pytest.mark.parametrize("n,seed", [(1000, 51), (37, 52), (0, 53), (3000, 54)])
def test_assign_features_to_grid(matcher, oracle, synth, n, seed):
    F = synth.frame_view(n, seed=seed)
    if n >= 5:
        F["kx"][:5] = [-3.0, 751.9, 760.0, 0.2, 375.5]      # the fork range-checks only the LINEAR cell index
    off, idx = matcher.AssignFeaturesToGrid(F)
    o_off, o_idx = oracle.assign_features_to_grid(F)
    assert np.array_equal(off, o_off) and np.array_equal(idx, o_idx)


def test_assign_features_to_grid_batched(matcher, oracle, synth):
    Fs = [synth.frame_view(200 + 97 * i, seed=60 + i, gcols=32 + 8 * (i % 3), grows=24 + 4 * (i % 2)) for i in range(9)]
    for (off, idx), F in zip(matcher.AssignFeaturesToGrid(Fs), Fs):
        o_off, o_idx = oracle.assign_features_to_grid(F)
        assert np.array_equal(off, o_off) and np.array_equal(idx, o_idx)


def test_get_features_in_area(matcher, oracle, synth):
    F = synth.frame_view(1500, seed=71)
    rng = np.random.default_rng(72)
    nq = 500
    x = rng.uniform(-30, 790, nq).astype(np.float32); y = rng.uniform(-30, 510, nq).astype(np.float32)
    r = rng.uniform(1, 60, nq).astype(np.float32)
    lo = rng.integers(-1, 6, nq).astype(np.int32)
    hi = np.where(rng.random(nq) < 0.3, -1, lo + rng.integers(0, 3, nq)).astype(np.int32)
    cnt, idx = matcher.GetFeaturesInArea(F, x, y, r, lo, hi, cap=128)
    o_cnt, o_idx = oracle.features_in_area(F, x, y, r, lo, hi, cap=128)
    assert np.array_equal(cnt, o_cnt) and cnt.max() > 20
    assert np.array_equal(idx, o_idx)          # same candidates in the same (cell, push_back) order


def test_extractor_device_grid(pkg, matcher, oracle, synth):
    """Grid built on the device from the extractor's resident keypoints == oracle grid of the fetched keypoints."""
    W, H, B = 640, 400, 3
    frames = synth.frame_batch(B, W, H, seed=77, distinct=B)
    ex = pkg.orbx.ORBextractor(800, 1.2, 6, 20, 7, max_batch=B).prepare(W, H, B)
    ex.extract_host(frames)
    gcols, grows = 64, 48
    gw_inv, gh_inv = float(np.float32(gcols) / np.float32(W)), float(np.float32(grows) / np.float32(H))
    ex.assign_grid(0.0, 0.0, gw_inv, gh_inv, gcols, grows)
    ex.sync()
    for f in range(B):
        kps, _, _ = ex.fetch(f)
        F = {"kx": kps["x"].copy(), "ky": kps["y"].copy(), "koct": kps["octave"].astype(np.int32),
             "desc": np.zeros((len(kps), 32), np.uint8), "minx": 0.0, "miny": 0.0, "gw_inv": gw_inv, "gh_inv": gh_inv,
             "gcols": gcols, "grows": grows, "scale": np.ones(6, np.float32)}
        off, idx = ex.grid_fetch(f, len(kps))
        o_off, o_idx = oracle.assign_features_to_grid(F)
        assert len(kps) > 300 and np.array_equal(off, o_off) and np.array_equal(idx, o_idx)
    ex.close()


def test_async_entry_points_back_to_back(pkg, synth):
    """The asynchronous entry points (stereo_matches_async, assign_grid, extract with a lapping
    table) stage their small host tables so that a burst of calls with CHANGING arguments, issued
    without any synchronisation in between, leaves exactly the result of the last call (a staging
    buffer rewritten while an earlier call's copy is still queued would tear them)."""
    ox = pkg.orbx
    W, H, B = 640, 400, 32
    sp = [synth.stereo_pair(W, H, 300 + i) for i in range(4)]
    base = [im for pr in sp for im in pr]
    imgs = np.stack([base[i % 8] for i in range(B)])
    ex = ox.ORBextractor(800, 1.2, 8, 20, 7, max_batch=B).prepare(W, H, B)
    lapA = np.tile(np.array([[0, 0]], np.int32), (B, 1)); lapB = np.tile(np.array([[200, 400]], np.int32), (B, 1))
    for k in range(12):                                   # extraction burst, alternating lapping tables
        ex.extract_host(imgs, lapB if k % 2 else lapA)
    ex.sync()
    got = [ex.fetch(f) for f in range(B)]
    ref = ex.extract_batch(imgs, lapB)
    for f in range(B):
        assert np.array_equal(got[f][0], ref[f][0]) and np.array_equal(got[f][1], ref[f][1]) and got[f][2] == ref[f][2]
    lf = np.arange(0, B, 2); rf = lf + 1
    for k in range(40):                                   # stereo burst: the pair list changes every call
        sel = np.roll(np.arange(len(lf)), k)
        ox.stereo_matches_async(ex, ex, lf[sel], rf[sel], 40.0, 0.1)
    ur, dp = ox.stereo_fetch(ex, len(lf))
    sel = np.roll(np.arange(len(lf)), 39)
    ur1, dp1 = ox.stereo_matches(ex, ex, lf[sel], rf[sel], 40.0, 0.1)
    assert np.array_equal(ur, ur1) and np.array_equal(dp, dp1) and (ur >= 0).sum() > 100
    gcols, grows = 64, 48
    gw, gh = float(np.float32(gcols) / np.float32(W)), float(np.float32(grows) / np.float32(H))
    for k in range(40):                                   # grid burst: origin and cell size change every call
        ex.assign_grid(float(k % 3), float(k % 2), gw * (1 + 0.01 * (k % 5)), gh, gcols, grows)
    ex.sync()
    burst = [ex.grid_fetch(f, len(ref[f][0])) for f in range(B)]
    ex.assign_grid(float(39 % 3), float(39 % 2), gw * (1 + 0.01 * (39 % 5)), gh, gcols, grows)
    ex.sync()
    for f in range(B):
        off, idx = ex.grid_fetch(f, len(ref[f][0]))
        assert np.array_equal(off, burst[f][0]) and np.array_equal(idx, burst[f][1])
    ex.close()


# ------------------------------------------------------------------ distinctive descriptors
def test_distinctive_descriptors(matcher, oracle, synth):
    rng = np.random.default_rng(91)
    sizes = [0, 1, 2, 3, 7, 33, 128, 129, 200, 1500] + list(rng.integers(2, 40, 60))
    sets = []
    for n in sizes:
        base = rng.integers(0, 256, 32, dtype=np.uint8)
        d = np.repeat(base[None], n, axis=0)
        flips = rng.random((n, 256)) < rng.uniform(0.02, 0.3)
        d = d ^ np.packbits(flips, axis=1)
        if n > 4:
            d[rng.integers(0, n)] = rng.integers(0, 256, 32, dtype=np.uint8)      # an outlier observation
        sets.append(np.ascontiguousarray(d, np.uint8))
    best, med = matcher.ComputeDistinctiveDescriptors(sets)
    for s, d in enumerate(sets):
        assert (int(best[s]), int(med[s])) == oracle.distinctive_descriptor(d), f"set {s} (N={len(d)})"


# ------------------------------------------------------------------ SearchByBoW
@pytest.mark.parametrize("n1,n2,nnodes,ori,nleft,seed", [
    (1000, 1100, 90, True, -1, 101), (1000, 1100, 90, False, -1, 102), (300, 250, 12, True, -1, 103),
    (1000, 1100, 90, True, 600, 104), (50, 40, 3, True, 20, 105), (0, 10, 2, True, -1, 106),
])
def test_search_by_bow(matcher, oracle, synth, n1, n2, nnodes, ori, nleft, seed):
    if n1 == 0:
        K1, K2, _, _, _ = synth.keyframe_pair(5, n2, nnodes, seed)
        for k in ("kx", "ky", "koct", "kangle", "has_mp", "uright"):
            K1[k] = K1[k][:0]
        K1["desc"] = K1["desc"][:0]; K1["node_id"] = K1["node_id"][:0]; K1["node_off"] = np.zeros(1, np.int32); K1["node_idx"] = K1["node_idx"][:0]
    else:
        K1, K2, _, _, _ = synth.keyframe_pair(n1, n2, nnodes, seed)
        K1["has_mp"] = (np.random.default_rng(seed).random(n1) < 0.8).astype(np.uint8)      # most KeyFrame points carry a MapPoint
    n, m = matcher.SearchByBoW(K1, K2, 0.7, ori, None if nleft == -1 else nleft)
    on, om = oracle.search_by_bow(K1, K2, 0.7, ori, nleft)
    assert n == on and np.array_equal(m, om)
    if n1 >= 300:
        assert n > 20


def test_search_by_bow_batched(matcher, oracle, synth):
    probs = [synth.keyframe_pair(400 + 50 * i, 380 + 30 * i, 20 + 7 * i, 110 + i) for i in range(10)]
    for i, p in enumerate(probs):
        p[0]["has_mp"] = (np.random.default_rng(i).random(len(p[0]["kx"])) < 0.7).astype(np.uint8)
    res = matcher.SearchByBoW([p[0] for p in probs], [p[1] for p in probs], 0.75, True)
    for (n, m), p in zip(res, probs):
        on, om = oracle.search_by_bow(p[0], p[1], 0.75, True)
        assert n == on and np.array_equal(m, om)


# ------------------------------------------------------------------ Fuse search core
@pytest.mark.parametrize("n,nmp,th,variant,stereo,seed", [
    (1000, 1200, 3.0, 0, True, 121), (1000, 1200, 3.0, 0, False, 122), (1000, 1500, 4.0, 1, False, 123),
    (60, 40, 2.5, 0, True, 124), (0, 10, 3.0, 0, True, 125), (500, 0, 3.0, 1, False, 126),
])
def test_fuse_search(matcher, oracle, synth, n, nmp, th, variant, stereo, seed):
    F = synth.frame_view(n, seed=seed)
    mp, ur, isg = synth.fuse_problem(F, nmp, seed + 1) if n > 0 else (
        {"valid": np.ones(nmp, np.uint8), "u": np.full(nmp, 100, np.float32), "v": np.full(nmp, 100, np.float32),
         "ur": np.full(nmp, 90, np.float32), "level": np.zeros(nmp, np.int32), "desc": np.zeros((nmp, 32), np.uint8)},
        np.zeros(0, np.float32), np.ones(8, np.float32))
    if not stereo:
        ur = None
    bi, bd = matcher.FuseSearch(F, mp, th, variant, ur, isg)
    obi, obd = oracle.fuse_search(F, mp, th, variant, ur, isg)
    assert np.array_equal(bi, obi) and np.array_equal(bd, obd)
    if n >= 500 and nmp >= 1000:
        assert (bi >= 0).sum() > 200


def test_fuse_search_batched(matcher, oracle, synth):
    Fs = [synth.frame_view(300 + 111 * i, seed=130 + i) for i in range(7)]
    P = [synth.fuse_problem(F, 200 + 90 * i, 140 + i) for i, F in enumerate(Fs)]
    res = matcher.FuseSearch(Fs, [p[0] for p in P], 3.0, 0, [p[1] for p in P], [p[2] for p in P])
    for (bi, bd), F, p in zip(res, Fs, P):
        obi, obd = oracle.fuse_search(F, p[0], 3.0, 0, p[1], p[2])
        assert np.array_equal(bi, obi) and np.array_equal(bd, obd)


# ------------------------------------------------------------------ SearchBySim3
@pytest.mark.parametrize("n1,n2,th,seed", [(900, 1000, 7.5, 151), (300, 120, 5.0, 152), (40, 50, 10.0, 153)])
def test_search_by_sim3(matcher, oracle, synth, n1, n2, th, seed):
    K1, K2, a, b = synth.sim3_problem(n1, n2, seed)
    n, m = matcher.SearchBySim3(K1, K2, a, b, th)
    on, om = oracle.search_by_sim3(K1, K2, a, b, th)
    assert n == on and np.array_equal(m, om)
    if n1 >= 300:
        assert n > 10


def test_search_by_sim3_batched(matcher, oracle, synth):
    P = [synth.sim3_problem(200 + 70 * i, 260 + 50 * i, 160 + i) for i in range(6)]
    res = matcher.SearchBySim3([p[0] for p in P], [p[1] for p in P], [p[2] for p in P], [p[3] for p in P], 7.5)
    for (n, m), p in zip(res, P):
        on, om = oracle.search_by_sim3(*p, 7.5)
        assert n == on and np.array_equal(m, om)


# ------------------------------------------------------------------ SearchForInitialization
@pytest.mark.parametrize("n1,n2,win,ratio,ori,seed", [
    (1000, 1000, 100, 0.9, True, 171), (1000, 1000, 100, 0.9, False, 172), (400, 700, 50, 0.8, True, 173),
    (30, 25, 100, 0.9, True, 174), (0, 20, 100, 0.9, True, 175), (20, 0, 100, 0.9, True, 176),
])
def test_search_for_initialization(matcher, oracle, synth, n1, n2, win, ratio, ori, seed):
    F1, F2 = synth.initialization_pair(max(n1, 2), max(n2, 2), seed)
    if n1 == 0:
        for k in ("kx", "ky", "koct", "kangle"):
            F1[k] = F1[k][:0]
        F1["desc"] = F1["desc"][:0]
    if n2 == 0:
        for k in ("kx", "ky", "koct", "kangle"):
            F2[k] = F2[k][:0]
        F2["desc"] = F2["desc"][:0]
    n, m = matcher.SearchForInitialization(F1, F2, win, ratio, ori)
    on, om = oracle.search_for_initialization(F1, F2, win, ratio, ori)
    assert n == on and np.array_equal(m, om)
    if n1 >= 400:
        assert n > 30


def test_search_for_initialization_batched(matcher, oracle, synth):
    P = [synth.initialization_pair(300 + 80 * i, 350 + 60 * i, 180 + i) for i in range(6)]
    res = matcher.SearchForInitialization([p[0] for p in P], [p[1] for p in P], 100, 0.9, True)
    for (n, m), p in zip(res, P):
        on, om = oracle.search_for_initialization(p[0], p[1], 100, 0.9, True)
        assert n == on and np.array_equal(m, om)


@pytest.mark.parametrize("win,ratio,ori", [(400, 0.9, True), (150, 1.0, False), (60, 0.6, True), (1000, 2.0, True)])
def test_search_for_initialization_clustered(matcher, oracle, synth, win, ratio, ori):
    """All keypoints on octave 0, descriptors drawn from 6 prototypes with a few flipped bits: every
    window holds far more than 32 candidates within TH_LOW, so the key lists are long, ties are
    everywhere and almost every claim is a steal."""
    rng = np.random.default_rng(win)
    F1, F2 = synth.initialization_pair(700, 800, 190 + win)
    proto = rng.integers(0, 256, (6, 32), dtype=np.uint8)
    for F in (F1, F2):
        n = len(F["kx"])
        F["koct"] = np.zeros(n, np.int32)
        bits = np.unpackbits(proto[rng.integers(0, 6, n)], axis=1)
        bits ^= (rng.random(bits.shape) < 0.03).astype(np.uint8)
        F["desc"] = np.packbits(bits, axis=1)
    n, m = matcher.SearchForInitialization(F1, F2, win, ratio, ori)
    on, om = oracle.search_for_initialization(F1, F2, win, ratio, ori)
    assert n == on and np.array_equal(m, om)
    assert (om >= 0).sum() > 5


# ------------------------------------------------------------------ committed vectors of the Python restatements
import os

import matcher_golden_cases as _G
import matcher_golden_run as _R


@pytest.mark.parametrize("name", list(_G.CASES))
def test_cuda_vs_committed_python_vectors(matcher, synth, golden_dir, name):
    """The CUDA kernels against tests/golden/py_matcher.npz: vectors produced by the straight-Python restatements of
    the reference's matcher functions (tools/gen_golden_matcher.py), i.e. by neither the kernels nor the C++ oracle."""
    _R.check(name, synth, np.load(os.path.join(golden_dir, "py_matcher.npz")), matcher, True)
