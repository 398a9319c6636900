#!/usr/bin/env python3
"""Single-problem (nprob = 1) and batched latency of the matcher entry points, synthetic and real-frame inputs.
usage: python tools/matcher_latency.py [label]"""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import orbx_pkg
pkg = orbx_pkg.load(); ox, sy = pkg.orbx, pkg.synth
def lat(fn, reps=30):
    for _ in range(3): fn()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter(); fn(); ts.append(time.perf_counter() - t0)
    return round(float(np.median(ts) * 1e3), 4)
m = ox.ORBmatcher()
F = sy.frame_view(1000, seed=100); MP = sy.projection_problem(F, 1500, seed=200)
P1 = dict(MP); P1["u"], P1["v"] = P1["x"], P1["y"]
out = {"label": sys.argv[1] if len(sys.argv) > 1 else ""}
out["sbp_synth"] = lat(lambda: m.SearchByProjection(F, MP, 3.0, False, 0.0, 0.8))
out["sbp_kf_synth"] = lat(lambda: m.SearchByProjectionKF(F, P1, 15.0, True))
W, H = 752, 480
ex = ox.ORBextractor(1000, 1.2, 8, 20, 7, max_batch=2)
l, r = sy.stereo_pair(W, H, 0xB200)
res = ex.extract_batch(np.stack([l, r]))
k0, d0, _ = res[0]
grid = {"minx": 0.0, "miny": 0.0, "gw_inv": float(np.float32(64) / np.float32(W)), "gh_inv": float(np.float32(48) / np.float32(H)), "gcols": 64, "grows": 48, "scale": ex.mvScaleFactor}
F0 = dict(grid, kx=k0["x"].copy(), ky=k0["y"].copy(), koct=k0["octave"].copy(), kangle=k0["angle"].copy(), desc=d0.copy())
MP0 = sy.projection_problem(F0, 1500, seed=777)
out["sbp_real"] = lat(lambda: m.SearchByProjection(F0, MP0, 3.0, False, 0.0, 0.8))
out["sbp_real_resident"] = lat(lambda: m.SearchByProjectionResident(ex, [0], grid, [MP0], 3.0, False, 0.0, 0.8))
out["sbp_real_matches"] = int(m.SearchByProjection(F0, MP0, 3.0, False, 0.0, 0.8)[0])
kfp = [sy.keyframe_pair(1000, 1100, 90, 300 + i) for i in range(4)]
for i, p in enumerate(kfp): p[0]["has_mp"] = (np.random.default_rng(i).random(1000) < 0.8).astype(np.uint8)
out["tri"] = lat(lambda: m.SearchForTriangulation(kfp[0][0], kfp[0][1], kfp[0][2], kfp[0][3], kfp[0][4], False, False, True))
out["bow"] = lat(lambda: m.SearchByBoW(kfp[0][0], kfp[0][1], 0.7, True))
ini = sy.initialization_pair(1000, 1000, 600)
out["init"] = lat(lambda: m.SearchForInitialization(ini[0], ini[1], 100, 0.9, True))
Fs = [sy.frame_view(1000, seed=100 + i) for i in range(8)] * 8; MPs = [sy.projection_problem(f, 1500, seed=200 + i) for i, f in enumerate(Fs[:8])] * 8
out["sbp_batch64"] = lat(lambda: m.SearchByProjection(Fs, MPs, 3.0, False, 0.0, 0.8), reps=5)
print(json.dumps(out))
