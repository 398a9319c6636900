#!/usr/bin/env python3
"""Minimal driver for ncu: N extraction steps of the bench workload (EuRoC shape, batch B) with
device-resident input, nothing else.  With ORBX_LANES=1 each step = 14 kernel launches (8 pyramid,
2 FAST grids, quadtree, order, blur, orient+describe; by default the batch runs as 4 chunks = 56); add --stereo for the two stereo launches per step and --knn for a
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
ap.add_argument("--knn", type=int, default=0, help="queries (x 64 database rows each unless --knn-rows)")
ap.add_argument("--knn-rows", type=int, default=0, help="database rows of the 2-NN launch")
ap.add_argument("--matcher", action="store_true", help="one batched call of every matcher entry point")
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
print("keypoints/frame", float(n.mean()), "launches/step", ex.last_launches() + 2 * int(a.stereo))
if a.knn:
    m = ox.ORBmatcher()
    if a.knn_rows:          # big shapes: uniform random bits are enough for a profile
        rng = np.random.default_rng(5)
        q = rng.integers(0, 256, (a.knn, 32), dtype=np.uint8); db = rng.integers(0, 256, (a.knn_rows, 32), dtype=np.uint8)
    else:
        q, db, _ = pkg.synth.descriptor_sets(a.knn, a.knn * 64, seed=5)
    d1, i1, d2, i2, ps = m.knn2(q, db, 0.8)
    print("knn pass rate", float(ps.mean()))
if a.matcher:
    sy = pkg.synth
    m = ox.ORBmatcher()
    Fs = [sy.frame_view(1000, seed=100 + i) for i in range(8)]
    MPs = [sy.projection_problem(F, 1500, seed=200 + i) for i, F in enumerate(Fs)]
    m.SearchByProjection(Fs, MPs, 3.0, False, 0.0, 0.8)
    kfp = [sy.keyframe_pair(1000, 1100, 90, 300 + i) for i in range(8)]
    for i, p in enumerate(kfp):
        p[0]["has_mp"] = (np.random.default_rng(i).random(1000) < 0.8).astype(np.uint8)
    m.SearchByBoW([p[0] for p in kfp], [p[1] for p in kfp], 0.7, True)
    for p in kfp:
        p[0]["has_mp"] = (np.random.default_rng(9).random(1000) < 0.2).astype(np.uint8)
    m.SearchForTriangulation([p[0] for p in kfp], [p[1] for p in kfp], [p[2] for p in kfp], [p[3] for p in kfp], [p[4] for p in kfp])
    fz = [sy.fuse_problem(F, 1200, 400 + i) for i, F in enumerate(Fs)]
    m.FuseSearch(Fs, [p[0] for p in fz], 3.0, 0, [p[1] for p in fz], [p[2] for p in fz])
    s3 = [sy.sim3_problem(900, 1000, 500 + i) for i in range(8)]
    m.SearchBySim3([p[0] for p in s3], [p[1] for p in s3], [p[2] for p in s3], [p[3] for p in s3], 7.5)
    ini = [sy.initialization_pair(1000, 1000, 600 + i) for i in range(8)]
    m.SearchForInitialization([p[0] for p in ini], [p[1] for p in ini], 100, 0.9, True)
    rng = np.random.default_rng(7)
    m.ComputeDistinctiveDescriptors([rng.integers(0, 256, (int(k), 32), dtype=np.uint8) for k in rng.integers(3, 40, 2000)])
    m.AssignFeaturesToGrid(Fs)
    m.MeanValue([rng.integers(0, 256, (int(k), 32), dtype=np.uint8) for k in rng.integers(3, 40, 2000)])
    # ROS-side pre-processing at the reference node's shape (32 frames of 2048x1536 BGR -> 614x460 gray), both paths
    m1, m2 = sy.undistortion_maps(2048, 1536, 2048, 1536, k1=-0.12, k2=0.02, zoom=0.97)
    bgr = np.stack([sy.bgr_image(2048, 1536, 300 + f) for f in range(2)] * 16)
    for fused in ("1", "0"):
        os.environ["ORBX_PREPROC_FUSED"] = fused
        g = ox.ImageGrabber((2048, 1536), m1, m2, 614, 460, max_batch=32)
        g.ConvertImageToGPU(bgr)
        g.close()
    os.environ.pop("ORBX_PREPROC_FUSED")
    print("matcher entry points done")
