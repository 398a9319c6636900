"""Second, independent restatement of the matcher functions that have no OpenCV counterpart: straight sequential Python
written from the reference source (src/ORBmatcher.cc, src/Frame.cc, src/CameraModels/Pinhole.cpp), sharing no code with
oracle/orb_oracle.cpp.  The C++ oracle -- the checker of the CUDA kernels -- must agree with it on seeded problems: two
restatements made separately from the same text are the strongest pin available for logic the reference itself never
tests ("parity unpinned" by the reference).  No GPU."""
import numpy as np
import pytest

from test_oracle_matcher import _grid_numpy, _popcount_rows

f32 = np.float32
TH_LOW, TH_HIGH, HISTO_LENGTH = 30, 100, 30          # include/ORBmatcher.h:73-75 (the fork's values)


def test_thresholds_are_the_reference_values():
    import os
    import re
    hdr = "/root/reference/include/ORBmatcher.h"
    if not os.path.exists(hdr):
        pytest.skip("reference tree not present on this machine")
    src = open(hdr).read()
    got = tuple(int(re.search(name + r"\s*=\s*(\d+)", src).group(1)) for name in ("TH_LOW", "TH_HIGH", "HISTO_LENGTH"))
    assert got == (TH_LOW, TH_HIGH, HISTO_LENGTH)


def c_round(v):
    """C round() of a float expression converted to int: half away from zero."""
    v = float(v)
    return int(np.sign(v) * np.floor(abs(v) + 0.5))


def rot_bin(a1, a2):
    """float rot = a1 - a2; if (rot < 0) rot += 360; bin = round(rot * (1.0f / 30)); if (bin == 30) bin = 0 (e.g. :249-256)."""
    rot = f32(f32(a1) - f32(a2))
    if rot < 0.0:
        rot = f32(rot + f32(360.0))
    b = c_round(f32(rot * f32(f32(1.0) / f32(HISTO_LENGTH))))
    return 0 if b == HISTO_LENGTH else b


def three_maxima(hist):
    """ORBmatcher::ComputeThreeMaxima, src/ORBmatcher.cc:1328-1370, on the bin sizes."""
    max1 = max2 = max3 = 0
    ind1 = ind2 = ind3 = -1
    for i, h in enumerate(hist):
        s = len(h)
        if s > max1:
            max3, max2, max1 = max2, max1, s
            ind3, ind2, ind1 = ind2, ind1, i
        elif s > max2:
            max3, max2 = max2, s
            ind3, ind2 = ind2, i
        elif s > max3:
            max3, ind3 = s, i
    if max2 < f32(0.1) * f32(max1):
        ind2 = ind3 = -1
    elif max3 < f32(0.1) * f32(max1):
        ind3 = -1
    return ind1, ind2, ind3


def features_in_area(F, cells, x, y, r, min_level=-1, max_level=-1):
    """Frame::GetFeaturesInArea, src/Frame.cc:404-468: cell columns outer, rows inner, a cell's keypoints in index order."""
    x, y, r = f32(x), f32(y), f32(r)
    minx, miny, gwi, ghi = f32(F["minx"]), f32(F["miny"]), f32(F["gw_inv"]), f32(F["gh_inv"])
    cols, rows = F["gcols"], F["grows"]
    c0 = max(0, int(np.floor(f32(f32(f32(x - minx) - r) * gwi))))
    if c0 >= cols:
        return []
    c1 = min(cols - 1, int(np.ceil(f32(f32(f32(x - minx) + r) * gwi))))
    if c1 < 0:
        return []
    r0 = max(0, int(np.floor(f32(f32(f32(y - miny) - r) * ghi))))
    if r0 >= rows:
        return []
    r1 = min(rows - 1, int(np.ceil(f32(f32(f32(y - miny) + r) * ghi))))
    if r1 < 0:
        return []
    check = min_level > 0 or max_level >= 0
    out = []
    for ix in range(c0, c1 + 1):
        for iy in range(r0, r1 + 1):
            for j in cells[iy * cols + ix]:
                o = int(F["koct"][j])
                if check and (o < min_level or (max_level >= 0 and o > max_level)):
                    continue
                if abs(f32(F["kx"][j] - x)) < r and abs(f32(F["ky"][j] - y)) < r:
                    out.append(j)
    return out


def dist(a, b):
    return int(_popcount_rows(a, b))


def shared_nodes(A, B):
    """the lockstep walk over two DBoW2 FeatureVectors (ordered maps node id -> feature indices), e.g. :152-157, 296-303"""
    ia = {int(n): k for k, n in enumerate(A["node_id"])}
    out = []
    for k2, n in enumerate(B["node_id"]):
        if int(n) in ia:
            k1 = ia[int(n)]
            out.append((list(A["node_idx"][A["node_off"][k1]:A["node_off"][k1 + 1]]), list(B["node_idx"][B["node_off"][k2]:B["node_off"][k2 + 1]])))
    return out


# ---------------------------------------------------------------- SearchByBoW, src/ORBmatcher.cc:133-327
def py_search_by_bow(KF, F, nn_ratio, check_ori, nleft=-1):
    nF = len(F["kx"])
    match = [-1] * nF                       # vpMapPointMatches as the KeyFrame keypoint whose MapPoint is assigned
    hist = [[] for _ in range(HISTO_LENGTH)]
    nmatches = 0
    for idxK, idxF in shared_nodes(KF, F):
        for rk in idxK:
            if not KF["has_mp"][rk]:
                continue
            b1, bi, b2 = 256, -1, 256
            b1r, bir, b2r = 256, -1, 256
            for rf in idxF:
                if match[rf] >= 0:
                    continue
                d = dist(KF["desc"][rk], F["desc"][rf])
                if nleft == -1 or rf < nleft:
                    if d < b1:
                        b2, b1, bi = b1, d, rf
                    elif d < b2:
                        b2 = d
                else:
                    if d < b1r:
                        b2r, b1r, bir = b1r, d, rf
                    elif d < b2r:
                        b2r = d
            if b1 <= TH_LOW:
                if f32(b1) < f32(nn_ratio) * f32(b2):
                    match[bi] = rk
                    if check_ori:
                        hist[rot_bin(KF["kangle"][rk], F["kangle"][bi])].append(bi)
                    nmatches += 1
                if b1r <= TH_LOW:             # `|| true`: no ratio test on the right camera
                    match[bir] = rk
                    if check_ori:
                        hist[rot_bin(KF["kangle"][rk], F["kangle"][bir])].append(bir)
                    nmatches += 1
    if check_ori:
        keep = three_maxima(hist)
        for i in range(HISTO_LENGTH):
            if i in keep:
                continue
            for j in hist[i]:
                match[j] = -1
                nmatches -= 1
    return nmatches, np.array(match, np.int32)


# ---------------------------------------------------------------- SearchForInitialization, :329-439
def py_search_for_initialization(F1, F2, window, nn_ratio, check_ori):
    n1, n2 = len(F1["kx"]), len(F2["kx"])
    cells2 = _grid_numpy(F2)
    m12 = [-1] * n1
    m21 = [-1] * n2
    mdist = [2 ** 31 - 1] * n2
    hist = [[] for _ in range(HISTO_LENGTH)]
    nmatches = 0
    for i1 in range(n1):
        if F1["koct"][i1] > 0:
            continue
        cand = features_in_area(F2, cells2, F1["kx"][i1], F1["ky"][i1], window, 0, 0)
        if not cand:
            continue
        best, best2, bi = 2 ** 31 - 1, 2 ** 31 - 1, -1
        for i2 in cand:
            d = dist(F1["desc"][i1], F2["desc"][i2])
            if mdist[i2] <= d:
                continue
            if d < best:
                best2, best, bi = best, d, i2
            elif d < best2:
                best2 = d
        if best <= TH_LOW and best < f32(f32(best2) * f32(nn_ratio)):
            if m21[bi] >= 0:
                m12[m21[bi]] = -1
                nmatches -= 1
            m12[i1] = bi; m21[bi] = i1; mdist[bi] = best
            nmatches += 1
            if check_ori:
                hist[rot_bin(F1["kangle"][i1], F2["kangle"][bi])].append(i1)
    if check_ori:
        keep = three_maxima(hist)
        for i in range(HISTO_LENGTH):
            if i in keep:
                continue
            for idx1 in hist[i]:
                if m12[idx1] >= 0:
                    m12[idx1] = -1
                    nmatches -= 1
    return nmatches, np.array(m12, np.int32)


# ---------------------------------------------------------------- SearchForTriangulation, :441-676 + Pinhole.cpp:104-131
def py_search_for_triangulation(K1, K2, F12, ep, scale2, only_stereo, coarse, check_ori):
    F12 = np.asarray(F12, f32).reshape(3, 3)
    n1 = len(K1["kx"])
    m12 = [-1] * n1
    hist = [[] for _ in range(HISTO_LENGTH)]
    nmatches = 0
    ur1, ur2 = K1.get("uright"), K2.get("uright")
    for idx1s, idx2s in shared_nodes(K1, K2):
        for i1 in idx1s:
            if K1["has_mp"][i1]:
                continue
            st1 = ur1 is not None and ur1[i1] >= 0
            if only_stereo and not st1:
                continue
            x1, y1 = f32(K1["kx"][i1]), f32(K1["ky"][i1])
            best, bi = TH_LOW, -1
            for i2 in idx2s:
                if K2["has_mp"][i2]:                 # vbMatched2 is never set in the reference
                    continue
                st2 = ur2 is not None and ur2[i2] >= 0
                if only_stereo and not st2:
                    continue
                d = dist(K1["desc"][i1], K2["desc"][i2])
                if d > TH_LOW or d > best:
                    continue
                x2, y2 = f32(K2["kx"][i2]), f32(K2["ky"][i2])
                if not st1 and not st2:
                    ex, ey = f32(f32(ep[0]) - x2), f32(f32(ep[1]) - y2)
                    if f32(f32(ex * ex) + f32(ey * ey)) < f32(f32(100) * f32(scale2[K2["koct"][i2]])):
                        continue
                a = f32(f32(f32(x1 * F12[0, 0]) + f32(y1 * F12[1, 0])) + F12[2, 0])
                b = f32(f32(f32(x1 * F12[0, 1]) + f32(y1 * F12[1, 1])) + F12[2, 1])
                c = f32(f32(f32(x1 * F12[0, 2]) + f32(y1 * F12[1, 2])) + F12[2, 2])
                num = f32(f32(f32(a * x2) + f32(b * y2)) + c)
                den = f32(f32(a * a) + f32(b * b))
                ok = False
                if den != 0:
                    ok = f32(f32(num * num) / den) < f32(3.84)
                if coarse or ok:
                    bi, best = i2, d
            if bi >= 0:
                m12[i1] = bi
                nmatches += 1
                if check_ori:
                    hist[rot_bin(K1["kangle"][i1], K2["kangle"][bi])].append(i1)
    if check_ori:
        keep = three_maxima(hist)
        for i in range(HISTO_LENGTH):
            if i in keep:
                continue
            for j in hist[i]:
                m12[j] = -1
                nmatches -= 1
    pairs = [(i, m) for i, m in enumerate(m12) if m >= 0]
    return nmatches, np.array(pairs, np.int32).reshape(-1, 2)


# ---------------------------------------------------------------- SearchBySim3 (search core), :1040-1075, 1115-1150, 1153-1169
def py_search_by_sim3(K1, K2, mp1in2, mp2in1, th):
    def one_way(MP, K):
        cells = _grid_numpy(K)
        out = [-1] * len(MP["u"])
        for i in range(len(MP["u"])):
            if not MP["valid"][i]:
                continue
            lvl = int(MP["level"][i])
            cand = features_in_area(K, cells, MP["u"][i], MP["v"][i], f32(f32(th) * f32(K["scale"][lvl])))
            best, bi = 2 ** 31 - 1, -1
            for j in cand:
                if K["koct"][j] < lvl - 1 or K["koct"][j] > lvl:
                    continue
                d = dist(MP["desc"][i], K["desc"][j])
                if d < best:
                    best, bi = d, j
            if best <= TH_HIGH:
                out[i] = bi
        return out
    m1, m2 = one_way(mp1in2, K2), one_way(mp2in1, K1)
    m12 = np.full(len(m1), -1, np.int32)
    found = 0
    for i1, idx2 in enumerate(m1):
        if idx2 >= 0 and m2[idx2] == i1:
            m12[i1] = idx2
            found += 1
    return found, m12


# ---------------------------------------------------------------- SearchByProjection(Frame, KeyFrame), :1202-1326 (search core)
def py_search_by_projection_kf(F, MP, th, check_ori, occupied=None):
    n = len(F["kx"])
    cells = _grid_numpy(F)
    taken = [False] * n if occupied is None else [bool(v) for v in occupied]
    match = [-1] * n
    hist = [[] for _ in range(HISTO_LENGTH)]
    nmatches = 0
    for i in range(len(MP["u"])):
        if not MP["valid"][i]:
            continue
        lvl = int(MP["level"][i])
        cand = features_in_area(F, cells, MP["u"][i], MP["v"][i], f32(f32(th) * f32(F["scale"][lvl])), lvl - 1, lvl + 1)
        best, bi = 256, -1
        for j in cand:
            if taken[j]:
                continue
            d = dist(MP["desc"][i], F["desc"][j])
            if d < best:
                best, bi = d, j
        if best <= TH_HIGH:
            taken[bi] = True
            match[bi] = i
            nmatches += 1
            if check_ori:
                hist[rot_bin(MP["angle"][i], F["kangle"][bi])].append(bi)
    if check_ori:
        keep = three_maxima(hist)
        for i in range(HISTO_LENGTH):
            if i not in keep:
                for j in hist[i]:
                    match[j] = -1
                    nmatches -= 1
    return nmatches, np.array(match, np.int32)


# ================================================================ the C++ oracle against the Python restatement
@pytest.mark.parametrize("n1,n2,nodes,ratio,ori,nleft,seed", [(220, 260, 14, 0.7, True, -1, 1), (300, 240, 9, 0.9, False, -1, 2),
                                                              (250, 250, 20, 0.75, True, 150, 3), (60, 400, 3, 0.6, True, -1, 4)])
def test_search_by_bow(oracle, synth, n1, n2, nodes, ratio, ori, nleft, seed):
    K1, K2, _, _, _ = synth.keyframe_pair(n1, n2, nodes, seed=seed)
    K1["has_mp"] = (np.random.default_rng(seed).random(n1) < 0.8).astype(np.uint8)
    pn, pm = py_search_by_bow(K1, K2, ratio, ori, nleft)
    on, om = oracle.search_by_bow(K1, K2, ratio, ori, nleft)
    assert on == pn and np.array_equal(om, pm) and pn > 5


@pytest.mark.parametrize("n1,n2,win,ratio,ori,seed", [(260, 280, 100, 0.9, True, 11), (300, 200, 40, 0.8, False, 12), (200, 320, 200, 1.0, True, 13)])
def test_search_for_initialization(oracle, synth, n1, n2, win, ratio, ori, seed):
    F1, F2 = synth.initialization_pair(n1, n2, seed, p_level0=0.7)
    pn, pm = py_search_for_initialization(F1, F2, win, ratio, ori)
    on, om = oracle.search_for_initialization(F1, F2, win, ratio, ori)
    assert on == pn and np.array_equal(om, pm) and pn > 10


@pytest.mark.parametrize("stereo,coarse,ori,mono_kf,seed", [(False, False, True, False, 21), (False, True, True, False, 22),
                                                             (True, False, False, False, 23), (False, False, True, True, 24)])
def test_search_for_triangulation(oracle, synth, stereo, coarse, ori, mono_kf, seed):
    K1, K2, F12, ep, sf = synth.keyframe_pair(300, 330, 25, seed=seed)
    K1["has_mp"] = (np.random.default_rng(seed).random(300) < 0.3).astype(np.uint8)
    if mono_kf:
        K1["uright"] = None; K2["uright"] = None
    pn, pp = py_search_for_triangulation(K1, K2, F12, ep, sf, stereo, coarse, ori)
    on, op = oracle.search_for_triangulation(K1, K2, F12, ep, sf, stereo, coarse, ori)
    assert on == pn and np.array_equal(op, pp) and pn > 3


@pytest.mark.parametrize("n1,n2,th,seed", [(250, 280, 7.5, 31), (300, 200, 4.0, 32), (120, 350, 10.0, 33)])
def test_search_by_sim3(oracle, synth, n1, n2, th, seed):
    K1, K2, a, b = synth.sim3_problem(n1, n2, seed)
    pn, pm = py_search_by_sim3(K1, K2, a, b, th)
    on, om = oracle.search_by_sim3(K1, K2, a, b, th)
    assert on == pn and np.array_equal(om, pm) and pn > 5


@pytest.mark.parametrize("n,nmp,th,ori,seed", [(300, 350, 15.0, True, 41), (250, 400, 5.0, False, 42), (300, 300, 40.0, True, 43)])
def test_search_by_projection_keyframe(oracle, synth, n, nmp, th, ori, seed):
    F = synth.frame_view(n, seed=seed)
    P = synth.projection_problem(F, nmp, seed=seed + 1, dup=0.4)
    MP = {"valid": P["valid"], "u": P["x"], "v": P["y"], "level": P["level"], "angle": P["angle"], "desc": P["desc"]}
    occ = (np.random.default_rng(seed).random(n) < 0.08).astype(np.uint8)
    pn, pm = py_search_by_projection_kf(F, MP, th, ori, occ)
    on, om = oracle.search_by_projection_kf(F, MP, th, ori, occ)
    assert on == pn and np.array_equal(om, pm) and pn > 20


def test_distinctive_descriptor(oracle):
    """MapPoint::ComputeDistinctiveDescriptors, src/MapPoint.cc:343-412: all-pairs distances, per row the element at index
    (int)(0.5 * (N - 1)) of the sorted row, the FIRST row with the smallest such median wins."""
    rng = np.random.default_rng(77)
    for n in list(range(1, 12)) + [16, 33, 64] + list(rng.integers(2, 50, 150)):
        base = rng.integers(0, 256, 32, dtype=np.uint8)
        d = base ^ np.packbits(rng.random((int(n), 256)) < rng.uniform(0.01, 0.4), axis=1)
        D = np.unpackbits(d[:, None, :] ^ d[None, :, :], axis=2).sum(2)
        best_median, best_idx = 2 ** 31 - 1, 0
        for i in range(len(d)):
            median = int(sorted(D[i])[int(0.5 * (len(d) - 1))])
            if median < best_median:
                best_median, best_idx = median, i
        assert oracle.distinctive_descriptor(np.ascontiguousarray(d, np.uint8)) == (best_idx, best_median), n


# ================================================================ committed vectors of the Python restatements
import os

import matcher_golden_cases as _G
import matcher_golden_run as _R


@pytest.mark.parametrize("name", list(_G.CASES))
def test_oracle_vs_committed_python_vectors(oracle, synth, golden_dir, name):
    """tests/golden/py_matcher.npz holds what the Python restatements return (tools/gen_golden_matcher.py): the C++ oracle
    must reproduce it.  tests/test_gpu_matcher.py compares the CUDA kernels with the same file."""
    _R.check(name, synth, np.load(os.path.join(golden_dir, "py_matcher.npz")), oracle, False)
