#!/usr/bin/env python3
"""Random matcher problems (sizes, radii, ratios, flags, stereo / mono, claimed / occupied masks) through every search
entry point of the C ABI against the CPU oracle: match indices, counts and distances must be equal.
Run on the GPU box: python tools/fuzz_matcher.py [N] [seed]   (one line per case; exit 1 on any difference)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))


def size(rng, lo, hi):
    """log-uniform integer: small problems as often as large ones"""
    return int(np.exp(rng.uniform(np.log(lo), np.log(hi))))


def case_projection(m, O, sy, rng):
    n, nmp = size(rng, 1, 3000), size(rng, 1, 4000)
    th = float(rng.choice([1.0, 3.0, 7.0, 15.0])); ratio = float(rng.choice([0.6, 0.8, 0.9, 1.0]))
    far = bool(rng.random() < 0.3); stereo = bool(rng.random() < 0.5)
    seed = int(rng.integers(1 << 30))
    F = sy.frame_view(n, seed=seed)
    MP = sy.projection_problem(F, nmp, seed=seed + 1, dup=float(rng.uniform(0, 0.6)), jitter=float(rng.uniform(0.5, 6)))
    if stereo:
        F, MP = sy.add_stereo(F, MP, seed=seed + 2, p_obs=float(rng.uniform(0.3, 1.0)))
    claimed = (rng.random(n) < rng.uniform(0, 0.3)).astype(np.uint8) if rng.random() < 0.5 else None
    on, om = O.search_by_projection(F, MP, th, far, 25.0, ratio, claimed)
    gn, gm = m.SearchByProjection(F, MP, th, far, 25.0, ratio, claimed)
    return gn == on and np.array_equal(gm, om), dict(n=n, nmp=nmp, th=th, ratio=ratio, far=far, stereo=stereo, matches=int(on))


def case_projection_kf(m, O, sy, rng):
    n, nmp = size(rng, 1, 3000), size(rng, 1, 3000)
    th = float(rng.choice([5.0, 15.0, 40.0])); ori = bool(rng.random() < 0.7)
    seed = int(rng.integers(1 << 30))
    F = sy.frame_view(n, seed=seed)
    P = sy.projection_problem(F, nmp, seed=seed + 1, dup=float(rng.uniform(0, 0.5)))
    MP = {"valid": P["valid"], "u": P["x"], "v": P["y"], "level": P["level"], "angle": P["angle"], "desc": P["desc"]}
    occ = (rng.random(n) < rng.uniform(0, 0.2)).astype(np.uint8) if rng.random() < 0.5 else None
    on, om = O.search_by_projection_kf(F, MP, th, ori, occ)
    gn, gm = m.SearchByProjectionKF(F, MP, th, ori, occ)
    return gn == on and np.array_equal(gm, om), dict(n=n, nmp=nmp, th=th, ori=ori, matches=int(on))


def case_triangulation(m, O, sy, rng):
    n1, n2, nodes = size(rng, 2, 2500), size(rng, 2, 2500), size(rng, 1, 200)
    stereo, coarse, ori = (bool(rng.random() < 0.3) for _ in range(3))
    k1, k2, F12, ep, sf = sy.keyframe_pair(n1, n2, nodes, seed=int(rng.integers(1 << 30)))
    k1["has_mp"] = (rng.random(n1) < rng.uniform(0, 0.9)).astype(np.uint8)
    if rng.random() < 0.3:
        k1["uright"] = None; k2["uright"] = None; stereo = False
    on, op = O.search_for_triangulation(k1, k2, F12, ep, sf, stereo, coarse, ori)
    gn, gp = m.SearchForTriangulation(k1, k2, F12, ep, sf, stereo, coarse, ori)
    return gn == on and np.array_equal(gp, op), dict(n1=n1, n2=n2, nodes=nodes, stereo=stereo, coarse=coarse, ori=ori, matches=int(on))


def case_bow(m, O, sy, rng):
    n1, n2, nodes = size(rng, 2, 2500), size(rng, 2, 2500), size(rng, 1, 200)
    ori = bool(rng.random() < 0.7); ratio = float(rng.choice([0.6, 0.7, 0.75, 0.9]))
    K1, K2, _, _, _ = sy.keyframe_pair(n1, n2, nodes, seed=int(rng.integers(1 << 30)))
    K1["has_mp"] = (rng.random(n1) < rng.uniform(0.1, 1.0)).astype(np.uint8)
    nleft = int(rng.integers(1, n2 + 1)) if rng.random() < 0.3 else -1
    n, mm = m.SearchByBoW(K1, K2, ratio, ori, None if nleft == -1 else nleft)
    on, om = O.search_by_bow(K1, K2, ratio, ori, nleft)
    return n == on and np.array_equal(mm, om), dict(n1=n1, n2=n2, nodes=nodes, ratio=ratio, ori=ori, nleft=nleft, matches=int(on))


def case_fuse(m, O, sy, rng):
    n, nmp = size(rng, 1, 3000), size(rng, 1, 3000)
    th = float(rng.choice([2.5, 3.0, 4.0, 6.0])); variant = int(rng.integers(0, 2)); stereo = bool(rng.random() < 0.5)
    seed = int(rng.integers(1 << 30))
    F = sy.frame_view(n, seed=seed)
    mp, ur, isg = sy.fuse_problem(F, nmp, seed + 1)
    if not stereo:
        ur = None
    bi, bd = m.FuseSearch(F, mp, th, variant, ur, isg)
    obi, obd = O.fuse_search(F, mp, th, variant, ur, isg)
    return np.array_equal(bi, obi) and np.array_equal(bd, obd), dict(n=n, nmp=nmp, th=th, variant=variant, stereo=stereo, found=int((obi >= 0).sum()))


def case_sim3(m, O, sy, rng):
    n1, n2 = size(rng, 2, 2500), size(rng, 2, 2500)
    th = float(rng.choice([5.0, 7.5, 10.0]))
    K1, K2, a, b = sy.sim3_problem(n1, n2, int(rng.integers(1 << 30)))
    n, mm = m.SearchBySim3(K1, K2, a, b, th)
    on, om = O.search_by_sim3(K1, K2, a, b, th)
    return n == on and np.array_equal(mm, om), dict(n1=n1, n2=n2, th=th, matches=int(on))


def case_initialization(m, O, sy, rng):
    n1, n2 = size(rng, 2, 2500), size(rng, 2, 2500)
    win = int(rng.choice([30, 60, 100, 200])); ratio = float(rng.choice([0.7, 0.9, 1.0])); ori = bool(rng.random() < 0.7)
    F1, F2 = sy.initialization_pair(n1, n2, int(rng.integers(1 << 30)), motion=float(rng.uniform(2, 60)), p_level0=float(rng.uniform(0.1, 1.0)))
    n, mm = m.SearchForInitialization(F1, F2, win, ratio, ori)
    on, om = O.search_for_initialization(F1, F2, win, ratio, ori)
    return n == on and np.array_equal(mm, om), dict(n1=n1, n2=n2, win=win, ratio=ratio, ori=ori, matches=int(on))


def case_sets(m, O, sy, rng):
    sets = [rng.integers(0, 256, (int(k), 32), dtype=np.uint8) for k in rng.integers(1, 60, int(rng.integers(1, 300)))]
    best, med = m.ComputeDistinctiveDescriptors(sets)
    mean = m.MeanValue(sets)
    ok = all((int(b), int(md)) == O.distinctive_descriptor(s) for b, md, s in zip(best, med, sets))
    ok = ok and all(np.array_equal(mv, O.forb_mean_value(s)) for mv, s in zip(mean, sets))
    return ok, dict(sets=len(sets))


CASES = [case_projection, case_projection_kf, case_triangulation, case_bow, case_fuse, case_sim3, case_initialization, case_sets]

if __name__ == "__main__":
    import orbx_pkg
    pkg = orbx_pkg.load()
    from oracle import oracle as O
    O.build()
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 40
    rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
    m = pkg.orbx.ORBmatcher()
    bad = 0
    for it in range(N):
        fn = CASES[it % len(CASES)]
        try:
            good, cfg = fn(m, O, pkg.synth, rng)
        except Exception as ex:       # noqa: BLE001
            good, cfg = False, {"error": repr(ex)[:300]}
        print("ok " if good else "BAD", fn.__name__[5:], cfg, flush=True)
        bad += not good
    print("failures:", bad, "of", N)
    sys.exit(1 if bad else 0)
