"""Second restatement of Frame::ComputeStereoMatches (upstream ORB-SLAM3 v1.0; absent from the fork) in straight Python,
written from SURVEY.md App. A5 and sharing no code with oracle/orb_oracle.cpp.  The C++ oracle must agree with it on
extracted stereo pairs: mvuRight and mvDepth, bit for bit.  It cannot catch an error of App. A5 itself (recall of
upstream) -- that stays stated as such in DESIGN.md -- but it pins the oracle's transcription of it.  No GPU."""
import numpy as np
import pytest

f32 = np.float32
TH_LOW, TH_HIGH = 30, 100


def c_round(v):
    v = float(v)
    return int(np.sign(v) * np.floor(abs(v) + 0.5))


def py_stereo_matches(kL, dL, lvL, kR, dR, lvR, sf, isf, bf, b, border=19):
    N = len(kL)
    n_rows = lvL[0].shape[0] - 2 * border
    th_orb = (TH_HIGH + TH_LOW) // 2
    rows = [[] for _ in range(n_rows)]
    for iR in range(len(kR)):
        y = f32(kR["y"][iR]); r = f32(f32(2.0) * f32(sf[kR["octave"][iR]]))
        for yi in range(int(np.floor(f32(y - r))), int(np.ceil(f32(y + r))) + 1):
            rows[yi].append(iR)
    min_z = f32(b); min_d = f32(0); max_d = f32(f32(bf) / min_z)
    ur = np.full(N, -1, f32); depth = np.full(N, -1, f32)
    pairs = []
    for iL in range(N):
        lvl = int(kL["octave"][iL]); vL, uL = f32(kL["y"][iL]), f32(kL["x"][iL])
        cand = rows[int(vL)]
        if not cand:
            continue
        min_u, max_u = f32(uL - max_d), f32(uL - min_d)
        if max_u < 0:
            continue
        best, best_r = TH_HIGH, 0
        for iR in cand:
            o = int(kR["octave"][iR])
            if o < lvl - 1 or o > lvl + 1:
                continue
            uR = f32(kR["x"][iR])
            if min_u <= uR <= max_u:
                d = int(np.unpackbits(dL[iL] ^ dR[iR]).sum())
                if d < best:
                    best, best_r = d, iR
        if best >= th_orb:
            continue
        s = f32(isf[lvl])
        suL, svL, suR0 = c_round(f32(uL * s)), c_round(f32(vL * s)), c_round(f32(f32(kR["x"][best_r]) * s))
        w = L = 5
        imL, imR = lvL[lvl], lvR[lvl]
        IL = imL[border + svL - w:border + svL + w + 1, border + suL - w:border + suL + w + 1].astype(np.int64)
        iniu, endu = suR0 + L - w, suR0 + L + w + 1
        if iniu < 0 or endu >= imR.shape[1] - 2 * border:
            continue
        best_sad, best_inc, dists = 2 ** 31 - 1, 0, []
        for inc in range(-L, L + 1):
            IR = imR[border + svL - w:border + svL + w + 1, border + suR0 + inc - w:border + suR0 + inc + w + 1].astype(np.int64)
            sad = int(np.abs(IL - IR).sum())
            if sad < best_sad:
                best_sad, best_inc = sad, inc
            dists.append(sad)
        if best_inc == -L or best_inc == L:
            continue
        d1, d2, d3 = f32(dists[L + best_inc - 1]), f32(dists[L + best_inc]), f32(dists[L + best_inc + 1])
        delta = f32(f32(d1 - d3) / f32(f32(2.0) * f32(f32(d1 + d3) - f32(f32(2.0) * d2))))
        if delta < -1 or delta > 1:
            continue
        best_ur = f32(f32(sf[lvl]) * f32(f32(f32(suR0) + f32(best_inc)) + delta))
        disp = f32(uL - best_ur)
        if min_d <= disp < max_d:
            if disp <= 0:
                disp = f32(0.01); best_ur = f32(uL - f32(0.01))
            depth[iL] = f32(f32(bf) / disp); ur[iL] = best_ur
            pairs.append((best_sad, iL))
    pairs.sort()
    if pairs:
        median = f32(pairs[len(pairs) // 2][0])
        th = f32(f32(f32(1.5) * f32(1.4)) * median)
        for sad, iL in reversed(pairs):
            if sad < th:
                break
            ur[iL] = -1; depth[iL] = -1
    return ur, depth


@pytest.mark.parametrize("w,h,nf,nl,seed,bf,b", [(400, 300, 500, 5, 3, 40.0, 0.1), (752, 480, 600, 8, 0xB200, 47.9, 0.11), (512, 384, 400, 4, 8, 25.0, 0.3)])
def test_oracle_stereo_equals_python_restatement(oracle, synth, w, h, nf, nl, seed, bf, b):
    left, right = synth.stereo_pair(w, h, seed)
    exL, exR = oracle.Extractor(nf, 1.2, nl, 20, 7), oracle.Extractor(nf, 1.2, nl, 20, 7)
    kL, dL, _ = exL(left, (0, 0)); kR, dR, _ = exR(right, (0, 0))
    lvL = [exL.level(l) for l in range(nl)]; lvR = [exR.level(l) for l in range(nl)]
    our, odep = oracle.stereo_matches(kL, dL, lvL, kR, dR, lvR, exL.scale, exL.inv_scale, bf, b)
    pur, pdep = py_stereo_matches(kL, dL, lvL, kR, dR, lvR, exL.scale, exL.inv_scale, bf, b)
    assert (our >= 0).sum() > 50
    assert np.array_equal(our, pur) and np.array_equal(odep, pdep)
