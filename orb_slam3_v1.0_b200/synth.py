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
