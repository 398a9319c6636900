#!/usr/bin/env python3
"""Minimal driver for ncu: N extraction steps of the bench workload (EuRoC shape, batch B) with
device-resident input, nothing else.  Each step = 13 kernel launches (8 pyramid, FAST, quadtree,
order, blur, orient+describe); add --stereo for one k_stereo launch per step and --knn for a
2-NN launch.  Usage under ncu: skip the first (steps-1)*launches_per_step launches."""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import orbx_pkg  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=256)
ap.add_argument("--steps", type=int, default=3)
ap.add_argument("--stereo", action="store_true")
ap.add_argument("--knn", type=int, default=0, help="queries (x 64 database rows each)")
a = ap.parse_args()
pkg = orbx_pkg.load()
ox = pkg.orbx
W, H = 752, 480
if a.stereo:
    pairs = [pkg.synth.stereo_pair(W, H, 0xB200 + i) for i in range(4)]
    base = np.stack([im for p in pairs for im in p])
    frames = np.stack([base[i % 8] for i in range(a.batch)])
else:
    frames = pkg.synth.frame_batch(a.batch, W, H, seed=0xB200, distinct=8)
ex = ox.ORBextractor(1000, 1.2, 8, 20, 7, max_batch=a.batch).prepare(W, H, a.batch)
dev = ox.DeviceBuffer(frames.nbytes)
dev.upload(frames)
for _ in range(a.steps):
    ex.extract_device(dev.ptr, a.batch, W, H)
    if a.stereo:
        lf = np.arange(0, a.batch, 2); rf = lf + 1
        ox.stereo_matches_async(ex, ex, lf, rf, 435.2 * 0.11, 0.11)
ex.sync()
n, _ = ex.counts(a.batch)
print("keypoints/frame", float(n.mean()), "launches/step", ex.last_launches() + int(a.stereo))
if a.knn:
    m = ox.ORBmatcher()
    q, db, _ = pkg.synth.descriptor_sets(a.knn, a.knn * 64, seed=5)
    d1, i1, d2, i2, ps = m.knn2(q, db, 0.8)
    print("knn pass rate", float(ps.mean()))
