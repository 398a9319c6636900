#!/usr/bin/env python3
"""Edge cases across the C ABI against the oracle that the regular tests do not draw: blank frames inside a batch,
stereo with an empty side, matcher problems with zero keypoints / zero map points on either side, every entry point.
Run on the GPU box: python tools/edge_cases.py  (exit 1 on any difference or exception)."""
import os, sys, traceback
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import orbx_pkg
pkg = orbx_pkg.load(); ox, sy = pkg.orbx, pkg.synth
from oracle import oracle as O
bad = 0


def check(name, fn):
    global bad
    try:
        ok = fn()
    except Exception:       # noqa: BLE001
        ok = False
        traceback.print_exc()
    print("ok " if ok else "BAD", name, flush=True)
    bad += not ok


def eq_frame(g, o):
    gk, gd, gm = g; ok_, od, om = o
    if len(gk) != len(ok_) or gm != om:
        return False
    if len(gk) == 0:
        return True
    return all(np.array_equal(gk[f], ok_[f]) for f in ("x", "y", "octave", "response", "size")) and int((np.unpackbits(gd ^ od, axis=1).sum(1) > 0).sum()) <= 1


W, H = 400, 300
def batch_with_blank_frames():
    frames = np.stack([sy.textured_image(W, H, 1), np.full((H, W), 90, np.uint8), sy.textured_image(W, H, 2), np.zeros((H, W), np.uint8),
                       np.full((H, W), 255, np.uint8)])
    ex = ox.ORBextractor(500, 1.2, 6, 20, 7, max_batch=8)
    res = ex.extract_batch(frames, lapping=[(0, 0), (10, 200), (100, 300), (0, 0), (5, 6)])
    oex = O.Extractor(500, 1.2, 6, 20, 7)
    good = all(eq_frame(res[f], oex(frames[f], lap)) for f, lap in enumerate([(0, 0), (10, 200), (100, 300), (0, 0), (5, 6)]))
    ex.close()
    return good and len(res[1][0]) == 0 and len(res[3][0]) == 0
check("batch with blank frames between textured ones", batch_with_blank_frames)


def stereo_empty_side():
    left, right = sy.stereo_pair(W, H, 3)
    blank = np.full((H, W), 60, np.uint8)
    good = True
    for l_img, r_img in ((left, blank), (blank, right), (blank, blank)):
        exL, exR = ox.ORBextractor(400, 1.2, 5, 20, 7), ox.ORBextractor(400, 1.2, 5, 20, 7)
        kl, dl, _ = exL(l_img, (0, 0)); kr, dr, _ = exR(r_img, (0, 0))
        ur, dep = (a[0][:len(kl)] for a in ox.stereo_matches(exL, exR, [0], [0], 40.0, 0.1))
        good &= bool((ur < 0).all()) and bool((dep < 0).all())
        exL.close(); exR.close()
    return good
check("stereo with an empty left / right / both sides", stereo_empty_side)

def batch_size_sequence_on_one_handle():
    """1 -> 4 -> 1 -> 1 -> 16 -> 32 -> 3 -> 1 frames on one handle (max_batch 32): the latency path captures a graph per
    (pointer, layout, batch), batches >= 32 take the multi-stream path; every call must give per-frame results equal to
    a fresh single-frame extraction."""
    frames = np.stack([sy.textured_image(W, H, 40 + i) for i in range(32)])
    pin = ox.PinnedBuffer((32, H, W)); pin.array[:] = frames
    ex = ox.ORBextractor(400, 1.2, 5, 20, 7, max_batch=32).prepare(W, H, 32)
    one = ox.ORBextractor(400, 1.2, 5, 20, 7)
    ref = [one(frames[f], (0, 0)) for f in range(32)]
    good = True
    for b in (1, 4, 1, 1, 16, 32, 3, 1, 32, 1):
        ex.extract_host(pin.array[:b])
        ex.sync()
        n, _ = ex.counts(b)
        for f in range(b):
            k, d, mono = ex.fetch(f)
            good &= k.tobytes() == ref[f][0].tobytes() and np.array_equal(d, ref[f][1]) and mono == ref[f][2] and n[f] == len(k)
    ex.close(); one.close(); pin.free()
    return good
check("batch sizes 1/4/1/1/16/32/3/1/32/1 on one handle", batch_size_sequence_on_one_handle)

def strided_inputs():
    """Host frames with padded rows and padded frames (a crop of a larger buffer), host and device entry points, against
    the same frames passed densely; the vec_ok / unaligned paths of the level-0 copy are both hit (odd offsets)."""
    B = 3
    big = np.zeros((B, H + 9, W + 37), np.uint8)
    dense = np.stack([sy.textured_image(W, H, 70 + i) for i in range(B)])
    big[:, 4:4 + H, 13:13 + W] = dense
    view = big[:, 4:4 + H, 13:13 + W]
    ex = ox.ORBextractor(400, 1.2, 5, 20, 7, max_batch=B)
    ref = ex.extract_batch(dense)
    ex.extract_host(view); ex.sync()
    good = all(ex.fetch(f)[0].tobytes() == ref[f][0].tobytes() and np.array_equal(ex.fetch(f)[1], ref[f][1]) for f in range(B))
    d = ox.DeviceBuffer(big.nbytes); d.upload(big)
    ox._ck(ox.lib().orbx_extract_device(ex._h, d.ptr.value + 4 * big.strides[1] + 13, big.strides[0], big.strides[1], B, None))
    ex.sync()
    good &= all(ex.fetch(f)[0].tobytes() == ref[f][0].tobytes() and np.array_equal(ex.fetch(f)[1], ref[f][1]) for f in range(B))
    d.free(); ex.close()
    return good
check("padded rows / frames, unaligned origin: host and device entry points", strided_inputs)

m = ox.ORBmatcher()


def empty_frame(F):
    E = dict(F)
    for k in ("kx", "ky", "koct", "kangle"):
        E[k] = F[k][:0]
    E["desc"] = F["desc"][:0]
    if E.get("uright") is not None:
        E["uright"] = E["uright"][:0]
    return E


def empty_pts(MP):
    return {k: (v[:0] if isinstance(v, np.ndarray) else v) for k, v in MP.items()}


F = sy.frame_view(500, seed=11); MP = sy.projection_problem(F, 700, seed=12)
for nameF, FF in (("frame", F), ("empty frame", empty_frame(F))):
    for nameM, MM in (("points", MP), ("no points", empty_pts(MP))):
        def sbp(FF=FF, MM=MM):
            on, om = O.search_by_projection(FF, MM, 3.0, False, 0.0, 0.8)
            gn, gm = m.SearchByProjection(FF, MM, 3.0, False, 0.0, 0.8)
            return gn == on and np.array_equal(gm, om)
        check(f"SearchByProjection: {nameF} x {nameM}", sbp)
        def sbpkf(FF=FF, MM=MM):
            K = {"valid": MM["valid"], "u": MM["x"], "v": MM["y"], "level": MM["level"], "angle": MM["angle"], "desc": MM["desc"]}
            on, om = O.search_by_projection_kf(FF, K, 15.0, True)
            gn, gm = m.SearchByProjectionKF(FF, K, 15.0, True)
            return gn == on and np.array_equal(gm, om)
        check(f"SearchByProjection(KeyFrame): {nameF} x {nameM}", sbpkf)
        def fuse(FF=FF):
            n = len(FF["kx"])
            mp, ur, isg = sy.fuse_problem(F, 300, 13)
            if nameM == "no points":
                mp = empty_pts(mp)
            ur = ur[:n]
            bi, bd = m.FuseSearch(FF, mp, 3.0, 0, ur, isg); obi, obd = O.fuse_search(FF, mp, 3.0, 0, ur, isg)
            return np.array_equal(bi, obi) and np.array_equal(bd, obd)
        check(f"Fuse: {nameF} x {nameM}", fuse)


def tri_no_shared_nodes():
    k1, k2, F12, ep, sf = sy.keyframe_pair(300, 320, 30, seed=14)
    k2 = dict(k2); k2["node_id"] = k2["node_id"] + 100000          # disjoint vocabularies: nothing to compare
    gn, gp = m.SearchForTriangulation(k1, k2, F12, ep, sf, False, False, True); on, op = O.search_for_triangulation(k1, k2, F12, ep, sf, False, False, True)
    n, mm = m.SearchByBoW(k1, k2, 0.7, True); bn, bm = O.search_by_bow(k1, k2, 0.7, True)
    return gn == on == 0 and np.array_equal(gp, op) and n == bn == 0 and np.array_equal(mm, bm)
check("SearchForTriangulation / SearchByBoW: no shared vocabulary node", tri_no_shared_nodes)


def knn_tiny():
    rng = np.random.default_rng(15)
    good = True
    for nq, ndb in ((1, 1), (1, 2), (3, 1), (33, 2), (1, 5000)):
        q = rng.integers(0, 256, (nq, 32), dtype=np.uint8); db = rng.integers(0, 256, (ndb, 32), dtype=np.uint8)
        g = m.knn2(q, db, 0.8); o = O.knn2(q, db)
        good &= all(np.array_equal(a, b) for a, b in zip(g[:4], o))
    return good
check("2-NN: 1 x 1, 1 x 2, 3 x 1, 33 x 2, 1 x 5000", knn_tiny)


def grabber_degenerate_maps():
    rng = np.random.default_rng(16)
    img = rng.integers(0, 256, (40, 60, 3), dtype=np.uint8)
    m1 = rng.uniform(-5, 65, (30, 50)).astype(np.float32); m2 = rng.uniform(-5, 45, (30, 50)).astype(np.float32)
    m1[0, :5] = [np.nan, np.inf, -np.inf, 1e30, -1e30]; m2[1, :3] = [np.nan, np.inf, -1e30]
    g = ox.ImageGrabber((60, 40), m1, m2, 50, 30)
    out = g.ConvertImageToGPU(img[None])[0]
    g.close()
    return np.array_equal(out, O.preprocess_bgr(img, m1, m2, 50, 30))
check("pre-processing: NaN / inf / huge map entries", grabber_degenerate_maps)
m.close()
print("failures:", bad)
sys.exit(1 if bad else 0)
