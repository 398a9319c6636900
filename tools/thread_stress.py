#!/usr/bin/env python3
"""Host-thread stress: T threads with their own handles run extraction / projection search / 2-NN / pre-processing
concurrently for N iterations each; handles are created and destroyed inside the loop every K iterations.
usage: python tools/thread_stress.py [threads] [iters] [ops e.g. espk] [recreate_every] [which handles, e.g. emg]"""
import os, sys, threading, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import orbx_pkg
pkg = orbx_pkg.load(); ox, synth = pkg.orbx, pkg.synth
T = int(sys.argv[1]) if len(sys.argv) > 1 else 6
N = int(sys.argv[2]) if len(sys.argv) > 2 else 50
OPS = sys.argv[3] if len(sys.argv) > 3 else "espk"
K = int(sys.argv[4]) if len(sys.argv) > 4 else 10
WHICH = sys.argv[5] if len(sys.argv) > 5 else "emg"       # which handles are re-created every K iterations
W, H = 400, 300
imgs = [synth.textured_image(W, H, 900 + i) for i in range(4)]
bgr = synth.bgr_image(320, 240, 5); m1, m2 = synth.undistortion_maps(320, 240, 320, 240, zoom=0.8)
F = synth.frame_view(600, seed=1); MP = synth.projection_problem(F, 900, seed=2)
q, db, _ = synth.descriptor_sets(512, 8192, seed=3)
left, right = synth.stereo_pair(W, H, 7)
kfp = synth.keyframe_pair(600, 650, 40, 8); kfp[0]["has_mp"] = (np.random.default_rng(1).random(600) < 0.6).astype(np.uint8)
ini = synth.initialization_pair(500, 500, 9)
fz = synth.fuse_problem(F, 500, 10)
s3 = synth.sim3_problem(400, 450, 11)
errors = []
leaked = []
t0 = time.time()


def work(tid):
    try:
        ex = m = g = None
        for it in range(N):
            if it % K == 0:
                if ex is None or "e" in WHICH:
                    if ex is not None:
                        if os.environ.get("STRESS_LEAK"): leaked.append(ex)
                        else: ex.close()
                    ex = ox.ORBextractor(500, 1.2, 6, 20, 7, max_batch=2)
                if m is None or "m" in WHICH:
                    if m is not None: m.close()
                    m = ox.ORBmatcher()
                if g is None or "g" in WHICH:
                    if g is not None: g.close()
                    g = ox.ImageGrabber((320, 240), m1, m2, 200, 150)
            if "e" in OPS: ex(imgs[(tid + it) % 4], (0, 0))
            if "s" in OPS: m.SearchByProjection(F, MP, 3.0, False, 0.0, 0.8)
            if "k" in OPS: m.knn2(q, db, 0.8)
            if "p" in OPS: g.ConvertImageToGPU(bgr[None])
            if "t" in OPS:          # stereo: both images through this thread's extractor, then the row search on the resident results
                ex.extract_batch(np.stack([left, right]))
                ox.stereo_matches(ex, ex, [0], [1], 40.0, 0.1)
            if "b" in OPS:          # the other matcher entry points, two problems per call
                m.SearchByBoW([kfp[0]] * 2, [kfp[1]] * 2, 0.7, True)
                m.SearchForTriangulation([kfp[0]] * 2, [kfp[1]] * 2, [kfp[2]] * 2, [kfp[3]] * 2, [kfp[4]] * 2)
                m.SearchForInitialization([ini[0]] * 2, [ini[1]] * 2, 100, 0.9, True)
                m.FuseSearch([F] * 2, [fz[0]] * 2, 3.0, 0, [fz[1]] * 2, [fz[2]] * 2)
                m.SearchBySim3([s3[0]] * 2, [s3[1]] * 2, [s3[2]] * 2, [s3[3]] * 2, 7.5)
    except Exception as e:       # noqa: BLE001
        errors.append((round(time.time() - t0, 4), tid, it, repr(e)[:300]))


th = [threading.Thread(target=work, args=(t,)) for t in range(T)]
for t in th: t.start()
for t in th: t.join()
print(f"threads {T} iters {N} ops {OPS} recreate {K}: {len(errors)} errors in {time.time() - t0:.1f}s")
for e in sorted(errors)[:6]:
    print("  ", e)
sys.exit(1 if errors else 0)
