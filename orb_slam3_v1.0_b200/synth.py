"""Deterministic synthetic inputs: textured u8 frames, stereo pairs,
256-bit descriptor sets.  numpy only; used by tests/ and bench.py (never by the kernels)."""
import numpy as np


def _smooth_noise(rng, h, w, cell):
    gh, gw = h // cell + 3, w // cell + 3
    g = rng.random((gh, gw), dtype=np.float32)
    ys = (np.arange(h, dtype=np.float32) / cell)
    xs = (np.arange(w, dtype=np.float32) / cell)
    y0 = ys.astype(np.int32); x0 = xs.astype(np.int32)
    fy = (ys - y0)[:, None]; fx = (xs - x0)[None, :]
    fy = fy * fy * (3 - 2 * fy); fx = fx * fx * (3 - 2 * fx)
    a = g[y0][:, x0]; b = g[y0][:, x0 + 1]; c = g[y0 + 1][:, x0]; d = g[y0 + 1][:, x0 + 1]
    return (a * (1 - fx) + b * fx) * (1 - fy) + (c * (1 - fx) + d * fx) * fy


def textured_image(width, height, seed=0xB200, nshapes=260):
    """Value-noise octaves + random filled rectangles/discs + +-2 LSB noise -> u8 (h, w)."""
    rng = np.random.default_rng(seed)
    img = np.zeros((height, width), np.float32)
    for cell, amp in ((48, 0.5), (16, 0.3), (5, 0.2)):
        img += amp * _smooth_noise(rng, height, width, cell)
    img = (img - img.min()) / max(float(img.max() - img.min()), 1e-6) * 255.0
    yy, xx = np.mgrid[0:height, 0:width]
    for _ in range(nshapes):
        cx, cy = int(rng.integers(0, width)), int(rng.integers(0, height))
        s = int(rng.integers(4, 40)); t = int(rng.integers(4, 40))
        g = float(rng.integers(0, 256)); alpha = float(rng.uniform(0.55, 1.0))
        if rng.random() < 0.6:
            y0, y1, x0, x1 = max(cy - t, 0), min(cy + t, height), max(cx - s, 0), min(cx + s, width)
            img[y0:y1, x0:x1] = (1 - alpha) * img[y0:y1, x0:x1] + alpha * g
        else:
            y0, y1, x0, x1 = max(cy - s, 0), min(cy + s + 1, height), max(cx - s, 0), min(cx + s + 1, width)
            m = (yy[y0:y1, x0:x1] - cy) ** 2 + (xx[y0:y1, x0:x1] - cx) ** 2 <= s * s
            sub = img[y0:y1, x0:x1]
            sub[m] = (1 - alpha) * sub[m] + alpha * g
    img += rng.integers(-2, 3, size=img.shape).astype(np.float32)
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def stereo_pair(width, height, seed=0xB200):
    """Rectified pair: right = left shifted left by a per-row-band disparity (4..60 px) + noise."""
    rng = np.random.default_rng(seed ^ 0x5EED)
    wide = textured_image(width + 64, height, seed)
    left = wide[:, :width].copy()
    right = np.empty_like(left)
    band = 32
    for y0 in range(0, height, band):
        d = int(rng.integers(4, 61))
        # a point at x in the left image appears at x-d in the right image
        right[y0:y0 + band] = wide[y0:y0 + band, d:d + width]
    noise = rng.integers(-2, 3, size=right.shape)
    right = np.clip(right.astype(np.int32) + noise, 0, 255).astype(np.uint8)
    return left, right


def frame_batch(n, width, height, seed=0xB200, distinct=8):
    """n frames built from `distinct` generated images rolled by a frame-dependent offset."""
    base = [textured_image(width, height, seed + i) for i in range(min(distinct, n))]
    out = np.empty((n, height, width), np.uint8)
    for i in range(n):
        b = base[i % len(base)]
        k = i // len(base)
        out[i] = np.roll(b, (7 * k, 13 * k), axis=(0, 1)) if k else b
    return out


def descriptor_sets(nq, ndb, seed=5, planted=0.5, flip_p=0.06):
    """Config-5 style: uniform-bit DB; `planted` of the queries are DB rows with Binomial flips."""
    rng = np.random.Generator(np.random.PCG64(seed))
    db = rng.integers(0, 256, size=(ndb, 32), dtype=np.uint8)
    q = rng.integers(0, 256, size=(nq, 32), dtype=np.uint8)
    npl = int(nq * planted)
    src = rng.integers(0, ndb, size=npl)
    flips = (rng.random((npl, 256)) < flip_p)
    q[:npl] = db[src] ^ np.packbits(flips, axis=1, bitorder="little")
    return q, db, src


# ---------------------------------------------------------------------------- matcher problems
def _flip_bits(rng, desc, p):
    flips = rng.random((len(desc), 256)) < p
    return desc ^ np.packbits(flips, axis=1, bitorder="little")


def frame_view(n=1000, width=752, height=480, nlevels=8, seed=1, gcols=64, grows=48):
    """A flattened Frame (include/orbx.h orbx_frame_view): random keypoints + descriptors and the
    grid constants Frame::Frame derives (mfGridElementWidthInv = cols / (maxX - minX))."""
    rng = np.random.default_rng(seed)
    sf = np.cumprod(np.r_[1.0, np.full(nlevels - 1, 1.2)]).astype(np.float32)
    return {
        "kx": rng.uniform(0, width, n).astype(np.float32), "ky": rng.uniform(0, height, n).astype(np.float32),
        "koct": rng.integers(0, nlevels, n).astype(np.int32), "kangle": rng.uniform(0, 360, n).astype(np.float32),
        "desc": rng.integers(0, 256, (n, 32), dtype=np.uint8),
        "minx": 0.0, "miny": 0.0, "gw_inv": float(np.float32(gcols) / np.float32(width)),
        "gh_inv": float(np.float32(grows) / np.float32(height)), "gcols": gcols, "grows": grows, "scale": sf,
    }


def projection_problem(F, nmp=1500, seed=2, jitter=3.0, flip_p=0.05, dup=0.3):
    """Map points projected near the frame's keypoints: `dup` of them target a keypoint another map
    point already targets (exercises the first-come claim rule); the rest are fresh or random."""
    rng = np.random.default_rng(seed)
    n = len(F["kx"])
    src = rng.integers(0, n, nmp)
    ndup = int(nmp * dup)
    src[nmp - ndup:] = src[rng.integers(0, nmp - ndup, ndup)]
    rng.shuffle(src)
    desc = _flip_bits(rng, F["desc"][src], flip_p)
    rnd = rng.random(nmp) < 0.15
    desc[rnd] = rng.integers(0, 256, (int(rnd.sum()), 32), dtype=np.uint8)
    lvl = np.clip(F["koct"][src] + rng.integers(0, 2, nmp), 0, len(F["scale"]) - 1).astype(np.int32)
    return {
        "inview": (rng.random(nmp) < 0.9).astype(np.uint8),
        "x": (F["kx"][src] + rng.normal(0, jitter, nmp)).astype(np.float32),
        "y": (F["ky"][src] + rng.normal(0, jitter, nmp)).astype(np.float32),
        "level": lvl, "viewcos": rng.uniform(0.99, 1.0, nmp).astype(np.float32),
        "depth": rng.uniform(0.5, 40.0, nmp).astype(np.float32), "desc": desc,
        # KeyFrame overload names
        "valid": (rng.random(nmp) < 0.9).astype(np.uint8),
        "angle": ((F["kangle"][src] + rng.choice([0.0, 0.0, 0.0, 40.0, 95.0, 200.0], nmp) + rng.normal(0, 4, nmp)) % 360).astype(np.float32),
        "src": src,
    }


def add_stereo(F, MP, seed=7, p_stereo=0.7, p_obs=0.85, xr_sigma=4.0):
    """Turn (frame view, projection problem) into the stereo case of SearchByProjection
    (reference src/ORBmatcher.cc:77-86): F gets mvuRight (> 0 for `p_stereo` of the keypoints, else -1),
    the map points get mTrackProjXR near the uRight of the keypoint they target (sigma `xr_sigma` px,
    20 % far off, so the |mTrackProjXR - mvuRight| <= r gate rejects some candidates) and has_obs
    (Observations() > 0)."""
    rng = np.random.default_rng(seed)
    n, nmp = len(F["kx"]), len(MP["x"])
    disp = rng.uniform(2.0, 60.0, n).astype(np.float32)
    ur = np.where(rng.random(n) < p_stereo, np.maximum(F["kx"] - disp, 0.5), -1.0).astype(np.float32)
    F2, MP2 = dict(F), dict(MP)
    F2["uright"] = ur
    tgt = np.where(ur[MP["src"]] > 0, ur[MP["src"]], F["kx"][MP["src"]] - 20.0)
    far = (rng.random(nmp) < 0.2) * rng.choice([-1.0, 1.0], nmp) * rng.uniform(20.0, 160.0, nmp)    # wrong depth: gate must reject
    MP2["proj_xr"] = (tgt + rng.normal(0, xr_sigma, nmp) + far).astype(np.float32)
    MP2["has_obs"] = (rng.random(nmp) < p_obs).astype(np.uint8)
    return F2, MP2


def keyframe_pair(n1=1000, n2=1100, nnodes=90, seed=3, width=752, height=480, nlevels=8):
    """Two flattened KeyFrames (orbx_kf_view) with DBoW2-style feature vectors (a partition of the
    keypoints over sorted node ids, most nodes shared), a fundamental matrix and an epipole."""
    rng = np.random.default_rng(seed)
    sf = np.cumprod(np.r_[1.0, np.full(nlevels - 1, 1.2)]).astype(np.float32)
    K2 = frame_view(n2, width, height, nlevels, seed + 1)
    src = rng.integers(0, n2, n1)
    k1 = {"kx": (K2["kx"][src] + rng.normal(0, 6, n1)).astype(np.float32),
          "ky": (K2["ky"][src] + rng.normal(0, 1.0, n1)).astype(np.float32),
          "koct": K2["koct"][src].copy(), "kangle": ((K2["kangle"][src] + rng.normal(0, 5, n1)) % 360).astype(np.float32),
          "desc": _flip_bits(rng, K2["desc"][src], 0.03)}
    k2 = {k: K2[k] for k in ("kx", "ky", "koct", "kangle", "desc")}
    ids = np.sort(rng.choice(10000, nnodes + 10, replace=False)).astype(np.int32)
    node_of_2 = rng.integers(0, nnodes, n2)
    node_of_1 = np.where(rng.random(n1) < 0.85, node_of_2[src], rng.integers(0, nnodes + 10, n1))

    def fv(node_of, nid):
        order = np.argsort(node_of, kind="stable")
        used, counts = np.unique(node_of, return_counts=True)
        return {"node_id": nid[used].astype(np.int32), "node_off": np.r_[0, np.cumsum(counts)].astype(np.int32),
                "node_idx": order.astype(np.int32)}
    k1.update(fv(node_of_1, ids)); k2.update(fv(node_of_2, ids))
    k1["has_mp"] = (rng.random(n1) < 0.2).astype(np.uint8); k2["has_mp"] = (rng.random(n2) < 0.2).astype(np.uint8)
    k1["uright"] = np.where(rng.random(n1) < 0.3, k1["kx"] - 5, -1).astype(np.float32)
    k2["uright"] = np.where(rng.random(n2) < 0.3, k2["kx"] - 5, -1).astype(np.float32)
    # rectified-like geometry: horizontal epipolar lines (F = [t]x for t = (1,0,0)) plus a small tilt
    F12 = np.array([[0, 0, 0.002], [0, 0, -1.0], [-0.002, 1.0, 0.3]], np.float32)
    ep = np.array([width * 0.5 + 40, height * 0.5 - 10], np.float32)
    return k1, k2, F12, ep, sf


def fuse_problem(F, nmp=1200, seed=4, jitter=1.5, flip_p=0.04, bf=47.9):
    """Projected map points for ORBmatcher::Fuse against the KeyFrame view F: most land near a
    keypoint with a similar descriptor, some are random; also the KeyFrame's mvuRight and
    mvInvLevelSigma2."""
    rng = np.random.default_rng(seed)
    n = len(F["kx"])
    src = rng.integers(0, n, nmp)
    desc = _flip_bits(rng, F["desc"][src], flip_p)
    rnd = rng.random(nmp) < 0.2
    desc[rnd] = rng.integers(0, 256, (int(rnd.sum()), 32), dtype=np.uint8)
    u = (F["kx"][src] + rng.normal(0, jitter, nmp)).astype(np.float32)
    v = (F["ky"][src] + rng.normal(0, jitter, nmp)).astype(np.float32)
    z = rng.uniform(2, 30, nmp).astype(np.float32)
    uright = np.where(rng.random(n) < 0.5, F["kx"] - np.float32(bf) / rng.uniform(2, 30, n).astype(np.float32), -1).astype(np.float32)
    mp = {"valid": (rng.random(nmp) < 0.9).astype(np.uint8), "u": u, "v": v,
          "ur": (u - np.float32(bf) / z).astype(np.float32),
          "level": np.clip(F["koct"][src] + rng.integers(0, 2, nmp), 0, len(F["scale"]) - 1).astype(np.int32), "desc": desc}
    inv_sigma2 = (1.0 / (F["scale"].astype(np.float32) ** 2)).astype(np.float32)
    return mp, uright, inv_sigma2


def sim3_problem(n1=900, n2=1000, seed=5, jitter=2.0, flip_p=0.05):
    """Two KeyFrame views whose keypoints partly correspond, plus per-keypoint map points projected
    into the other KeyFrame (ORBmatcher::SearchBySim3 after its projection loops)."""
    rng = np.random.default_rng(seed)
    K2 = frame_view(n2, seed=seed + 1)
    src = rng.integers(0, n2, n1)
    K1 = frame_view(n1, seed=seed + 2)
    K1["kx"] = np.clip(K2["kx"][src] + rng.normal(0, 8, n1), 0, 751).astype(np.float32)
    K1["ky"] = np.clip(K2["ky"][src] + rng.normal(0, 8, n1), 0, 479).astype(np.float32)
    K1["koct"] = K2["koct"][src].copy()
    K1["desc"] = _flip_bits(rng, K2["desc"][src], flip_p)
    nl = len(K1["scale"])
    mp1in2 = {"valid": (rng.random(n1) < 0.7).astype(np.uint8),
              "u": (K2["kx"][src] + rng.normal(0, jitter, n1)).astype(np.float32),
              "v": (K2["ky"][src] + rng.normal(0, jitter, n1)).astype(np.float32),
              "level": np.clip(K1["koct"] + rng.integers(0, 2, n1), 0, nl - 1).astype(np.int32), "desc": K1["desc"].copy()}
    inv = np.full(n2, -1, np.int64)
    inv[src] = np.arange(n1)                      # some KF2 keypoints have a twin in KF1
    has = inv >= 0
    tw = np.where(has, inv, 0)
    mp2in1 = {"valid": ((rng.random(n2) < 0.7) & has).astype(np.uint8),
              "u": (K1["kx"][tw] + rng.normal(0, jitter, n2)).astype(np.float32),
              "v": (K1["ky"][tw] + rng.normal(0, jitter, n2)).astype(np.float32),
              "level": np.clip(K2["koct"] + rng.integers(0, 2, n2), 0, nl - 1).astype(np.int32), "desc": K2["desc"].copy()}
    return K1, K2, mp1in2, mp2in1


def initialization_pair(n1=1000, n2=1000, seed=6, motion=25.0, flip_p=0.05, p_level0=0.45):
    """Two consecutive frames for ORBmatcher::SearchForInitialization: F1's keypoints reappear in F2
    displaced by up to `motion` px with slightly changed descriptors and angles; several F1
    keypoints share a twin so the steal / vMatchedDistance logic is exercised."""
    rng = np.random.default_rng(seed)
    F2 = frame_view(n2, seed=seed + 1)
    F2["koct"] = np.where(rng.random(n2) < p_level0, 0, F2["koct"]).astype(np.int32)
    src = rng.integers(0, n2, n1)
    src[n1 // 2:] = src[rng.integers(0, n1 // 2, n1 - n1 // 2)]          # duplicates compete for the same F2 keypoint
    F1 = frame_view(n1, seed=seed + 2)
    F1["kx"] = np.clip(F2["kx"][src] + rng.uniform(-motion, motion, n1), 0, 751).astype(np.float32)
    F1["ky"] = np.clip(F2["ky"][src] + rng.uniform(-motion, motion, n1), 0, 479).astype(np.float32)
    F1["koct"] = np.where(rng.random(n1) < 0.9, F2["koct"][src], rng.integers(0, 8, n1)).astype(np.int32)
    F1["kangle"] = ((F2["kangle"][src] + rng.normal(0, 8, n1)) % 360).astype(np.float32)
    F1["desc"] = _flip_bits(rng, F2["desc"][src], flip_p * rng.uniform(0.2, 2.0, (n1, 1)))
    return F1, F2


def bgr_image(width, height, seed=0xB200):
    """A colour camera frame: three differently textured planes (u8 BGR, interleaved)."""
    return np.ascontiguousarray(np.stack([textured_image(width, height, seed + 101 * c, nshapes=120) for c in range(3)], axis=-1))


def undistortion_maps(src_w, src_h, map_w, map_h, k1=-0.21, k2=0.05, zoom=0.92):
    """Two CV_32FC1 planes like initUndistortRectifyMap's, from a radial polynomial evaluated with + and * only in
    float64 (bit-reproducible everywhere).  zoom < 1 makes the corners of the map fall outside the source, so the
    BORDER_CONSTANT taps are exercised."""
    fx = 0.62 * src_w
    x = (np.arange(map_w, dtype=np.float64) + 0.5) * (src_w / map_w) - 0.5
    y = (np.arange(map_h, dtype=np.float64) + 0.5) * (src_h / map_h) - 0.5
    xn = (x[None, :] - 0.5 * src_w) / (fx * zoom)
    yn = (y[:, None] - 0.5 * src_h) / (fx * zoom)
    r2 = xn * xn + yn * yn
    d = 1.0 + k1 * r2 + k2 * r2 * r2
    mx = xn * d * fx + 0.5 * src_w + 0.0 * yn
    my = yn * d * fx + 0.5 * src_h + 0.0 * xn
    return np.ascontiguousarray(mx, np.float32), np.ascontiguousarray(my, np.float32)
