#!/usr/bin/env python3
"""Random extractor configurations (size, features, scale, levels, thresholds, lapping) against the
oracle, including pyramid and blurred levels.  Run on the GPU box: python tools/fuzz_parity.py [N] [seed]"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def run_case(pkg, O, rng):
    """One random configuration; returns (good, cfg, why, keypoints compared, descriptors that differ)."""
    w = int(rng.integers(140, 1000)); h = int(rng.integers(120, 700))
    scale = float(rng.choice([1.1, 1.2, 1.25, 1.3, 1.5, 2.0]))
    nl = int(rng.integers(2, 11))
    while min(w, h) / scale ** (nl - 1) < 60 and nl > 1:
        nl -= 1
    nf = int(rng.integers(50, 3000))
    ini = int(rng.choice([7, 12, 20, 35, 60, 130])); mn = int(rng.integers(2, ini + 1))
    lap = (int(rng.integers(0, w)), int(rng.integers(0, w))) if rng.random() < 0.5 else (0, 0)
    lap = (min(lap), max(lap))
    kind = rng.random()
    if kind < 0.7:
        img = pkg.synth.textured_image(w, h, int(rng.integers(1 << 30)))
    elif kind < 0.85:
        img = rng.integers(0, 256, (h, w), dtype=np.uint8)
    else:
        img = (pkg.synth.textured_image(w, h, int(rng.integers(1 << 30))) // 8 + 100).astype(np.uint8)   # low contrast: minTh retry
    cfg = dict(w=w, h=h, nf=nf, scale=scale, nl=nl, ini=ini, mn=mn, lap=lap, kind=round(kind, 2))
    why = []
    nk = nbad = 0
    try:
        gex = pkg.orbx.ORBextractor(nf, scale, nl, ini, mn)
        oex = O.Extractor(nf, scale, nl, ini, mn)
        gk, gd, gm = gex(img, lap)
        ok, od, om = oex(img, lap)
        good = len(gk) == len(ok) and gm == om
        if not good: why.append(f"count {len(gk)} vs {len(ok)} mono {gm} vs {om}")
        if good:
            for fld in ("x", "y", "size", "response", "octave"):
                good &= bool(np.array_equal(gk[fld], ok[fld]))
            if len(gk):
                da = np.abs(gk["angle"] - ok["angle"]); good &= bool(np.minimum(da, 360 - da).max() <= np.degrees(1e-4))
                bits = np.unpackbits(gd ^ od, axis=1).sum(1)
                nk, nbad = len(gk), int((bits > 0).sum())
                good &= bool(nbad <= max(1, len(gk) // 1000) and bits.max() <= 2)
            if not good and not why: why.append("keypoint fields / angle / descriptors")
        for l in range(nl):
            lv = oex.level(l)
            t = bool(np.array_equal(gex.image_pyramid(l), lv)); good &= t
            if not t: why.append(f"pyramid {l}")
            t = bool(np.array_equal(gex.blurred(l), O.gaussian7(lv[19:-19, 19:-19]))); good &= t
            if not t: why.append(f"blur {l}")
            gx, gy, gr = gex.candidates(l); ox_, oy_, or_ = oex.candidates(l)
            n_ = min(len(gx), len(ox_)) if len(ox_) == 65536 else max(len(gx), len(ox_))   # the oracle's export buffer holds 65536 candidates
            t = bool(np.array_equal(gx[:n_], ox_[:n_]) and np.array_equal(gy[:n_], oy_[:n_]) and np.array_equal(gr[:n_], or_[:n_])); good &= t
            if not t: why.append(f"candidates {l}: {len(gx)} vs {len(ox_)}")
        gex.close()
    except Exception as ex:       # noqa: BLE001
        good = False
        cfg["error"] = str(ex)[:200]
    return good, cfg, why, nk, nbad


if __name__ == "__main__":
    import orbx_pkg
    pkg = orbx_pkg.load()
    from oracle import oracle as O
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 30
    rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
    bad = 0
    for it in range(N):
        good, cfg, why, nk, nbad = run_case(pkg, O, rng)
        print("ok " if good else "BAD", nk, cfg, why, flush=True)
        bad += not good
    print("failures:", bad, "of", N)
    sys.exit(1 if bad else 0)
