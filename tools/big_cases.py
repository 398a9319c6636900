#!/usr/bin/env python3
"""A few large configurations against the oracle (4K frames, many features, extreme aspect ratios): the sizes the fuzz
tool does not draw.  Run on the GPU box: python tools/big_cases.py"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import orbx_pkg
pkg = orbx_pkg.load()
from oracle import oracle as O
CASES = [  # w, h, nfeatures, scale, levels, ini, min, lapping
    (3840, 2160, 5000, 1.2, 8, 20, 7, (0, 0)),
    (4000, 4000, 8000, 1.2, 8, 20, 7, (1000, 3000)),
    (4000, 64, 500, 1.2, 1, 20, 7, (0, 0)),
    (64, 4000, 500, 1.2, 1, 20, 7, (0, 0)),
    (752, 480, 1, 1.2, 8, 20, 7, (0, 0)),           # quotas of 0 and 1 features per level (the quadtree still returns its initial nodes)
    (752, 480, 5, 1.2, 8, 20, 7, (100, 600)),
    (640, 480, 17, 1.5, 6, 20, 7, (0, 0)),
    (100, 64, 300, 1.2, 1, 20, 7, (0, 0)),          # no FAST cell at all: no keypoints, monoIndex 0
    (300, 90, 300, 1.5, 3, 20, 7, (0, 0)),          # upper levels lose their cells
    (2048, 1536, 3000, 1.1, 12, 30, 5, (0, 0)),
    (1280, 720, 12000, 1.2, 8, 12, 3, (0, 0)),
]
bad = 0
for (w, h, nf, sc, nl, ini, mn, lap) in CASES:
    img = pkg.synth.textured_image(w, h, w + h)
    try:
        gk, gd, gm = pkg.orbx.ORBextractor(nf, sc, nl, ini, mn)(img, lap)
    except pkg.orbx.OrbxError as ex:
        print("REJECTED", (w, h, nf, sc, nl), str(ex)[:160]); continue
    ok, od, om = O.Extractor(nf, sc, nl, ini, mn)(img, lap)
    good = len(gk) == len(ok) and gm == om
    if good and len(gk):
        for fld in ("x", "y", "size", "response", "octave"):
            good &= bool(np.array_equal(gk[fld], ok[fld]))
        da = np.abs(gk["angle"] - ok["angle"]); good &= bool(np.minimum(da, 360 - da).max() <= np.degrees(1e-4))
        bits = np.unpackbits(gd ^ od, axis=1).sum(1)
        good &= bool((bits > 0).sum() <= max(1, len(gk) // 1000) and bits.max() <= 2)
    print("ok " if good else "BAD", (w, h, nf, sc, nl, ini, mn, lap), len(gk), len(ok), flush=True)
    bad += not good
print("failures:", bad)
sys.exit(1 if bad else 0)
