#!/usr/bin/env python3
"""Workload for compute-sanitizer (tools/sanitize.sh): one call of every kernel at small sizes --
batch-1 extraction on the CUDA-graph / per-level dataflow path (three calls: plain, capture, replay), a 40-frame
batch on two internal streams, the stereo pair kernels, the device grid, every matcher entry point, the 2-NN with
database slices (scratch + merge) and its allocation-free variant.  No torch: ctypes + numpy only."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import orbx_pkg  # noqa: E402

pkg = orbx_pkg.load()
ox, sy = pkg.orbx, pkg.synth
w, h = 320, 240
img = sy.textured_image(w, h, 3)
ex = ox.ORBextractor(300, 1.2, 4, 20, 7)
for _ in range(3):
    k, d, mono = ex(img, (0, 0))
print("batch 1:", len(k), "keypoints", flush=True)
ex.close()
B = 40
exb = ox.ORBextractor(300, 1.2, 4, 20, 7, max_batch=B)
frames = sy.frame_batch(B, w, h, seed=5, distinct=4)
res = exb.extract_batch(frames, np.tile(np.array([[50, 200]], np.int32), (B, 1)))
print("batch", B, "on", exb.streams(B), "streams:", sum(len(r[0]) for r in res), "keypoints", flush=True)
sp = [sy.stereo_pair(w, h, 40 + i) for i in range(2)]
st = np.stack([im for pr in sp for im in pr])
exs = ox.ORBextractor(300, 1.2, 4, 20, 7, max_batch=4)
exs.extract_batch(st)
ox.stereo_matches_async(exs, exs, np.array([0, 2]), np.array([1, 3]), 40.0, 0.1)
ur, dp = ox.stereo_fetch(exs, 2)
print("stereo:", int((ur >= 0).sum()), "matches", flush=True)
exs.assign_grid(0.0, 0.0, 64.0 / w, 48.0 / h, 64, 48)
exs.close(); exb.close()

m = ox.ORBmatcher()
F = sy.frame_view(300, seed=1)
MP = sy.projection_problem(F, 400, seed=2, dup=0.5)
Fs, MPs = sy.add_stereo(F, MP, seed=3)
print("sbp", m.SearchByProjection(Fs, MPs, 3.0, False, 0.0, 0.8)[0], flush=True)
P1 = dict(MP); P1["u"], P1["v"] = P1["x"], P1["y"]
print("sbp kf", m.SearchByProjectionKF(F, P1, 15.0, True)[0], flush=True)
k1, k2, F12, ep, sf = sy.keyframe_pair(300, 320, 30, seed=4)
print("tri", m.SearchForTriangulation(k1, k2, F12, ep, sf, False, False, True)[0], flush=True)
kb = dict(k1); kb["has_mp"] = np.ones(300, np.uint8)
print("bow", m.SearchByBoW(kb, k2, 0.7, True)[0], flush=True)
fm, fu, fi = sy.fuse_problem(F, 300, 5)
m.FuseSearch(F, fm, 3.0, 0, fu, fi)
a, b, c, dd = sy.sim3_problem(250, 260, 6)
print("sim3", m.SearchBySim3(a, b, c, dd, 7.5)[0], flush=True)
i1, i2 = sy.initialization_pair(300, 300, 7)
print("init", m.SearchForInitialization(i1, i2, 100, 0.9, True)[0], flush=True)
rng = np.random.default_rng(8)
sets = [rng.integers(0, 256, (int(n), 32), dtype=np.uint8) for n in (0, 1, 5, 40, 1500)]
m.ComputeDistinctiveDescriptors(sets); m.MeanValue(sets)
m.AssignFeaturesToGrid(F)
m.GetFeaturesInArea(F, F["kx"][:50], F["ky"][:50], np.full(50, 20.0), np.zeros(50, np.int32), np.full(50, 7, np.int32))
q, db, _ = sy.descriptor_sets(600, 20000, seed=5)
for v in (0, 1):
    m.set_knn_variant(v)
    m.knn2(q, db, 0.8)
dq = ox.DeviceBuffer(q.nbytes); dq.upload(q)
ddb = ox.DeviceBuffer(db.nbytes); ddb.upload(db)
keys = ox.DeviceBuffer(len(q) * 16)
m.knn2_device_noalloc(dq.ptr, len(q), ddb.ptr, len(db), 0, keys.ptr)
m.synchronize()
print("hamming", ox.lib().orbx_version(), flush=True)
for bfr in (dq, ddb, keys):
    bfr.free()
m.close()
print("sanitize workload done", flush=True)
