#!/usr/bin/env python3
"""Run-to-run determinism under load: the same 256-frame batch extracted REPS times on two rotating handles (kernels of
consecutive calls overlap on the device), the same stereo / matcher problems repeated; every repetition must return
bit-identical results (the kernels use atomics for unordered scratch only; the outputs are ordered afterwards).
Run on the GPU box: python tools/determinism_soak.py [reps]"""
import hashlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import orbx_pkg
pkg = orbx_pkg.load(); ox, sy = pkg.orbx, pkg.synth
REPS = int(sys.argv[1]) if len(sys.argv) > 1 else 200
B, W, H = 256, 752, 480
frames = sy.frame_batch(B, W, H, seed=0xB200, distinct=16)
pin = ox.PinnedBuffer((B, H, W)); pin.array[:] = frames
exs = [ox.ORBextractor(1000, 1.2, 8, 20, 7, max_batch=B).prepare(W, H, B) for _ in range(2)]


def digest(ex):
    ex.sync()
    h = hashlib.sha256()
    n, mono = ex.counts(B)
    h.update(n.tobytes()); h.update(mono.tobytes())
    for f in range(0, B, 7):
        k, d, _ = ex.fetch(f)
        h.update(k.tobytes()); h.update(d.tobytes())
    ur, dp = ox.stereo_matches(ex, ex, list(range(0, B, 2)), list(range(1, B, 2)), 47.9, 0.11)
    h.update(ur.tobytes()); h.update(dp.tobytes())
    return h.hexdigest()


exs[0].extract_host(pin.array); exs[1].extract_host(pin.array)
ref = digest(exs[0])
bad = 0
for r in range(REPS):
    e = exs[r & 1]
    exs[(r + 1) & 1].extract_host(pin.array)          # the other handle's kernels run while this one is read back
    d = digest(e)
    if d != ref:
        bad += 1
        print("MISMATCH at repetition", r, flush=True)
    e.extract_host(pin.array)
m = ox.ORBmatcher()
F = sy.frame_view(1000, seed=1); MP = sy.projection_problem(F, 1500, seed=2, dup=0.5)
kfp = sy.keyframe_pair(1000, 1100, 90, 3); kfp[0]["has_mp"] = (np.random.default_rng(1).random(1000) < 0.7).astype(np.uint8)
ini = sy.initialization_pair(1000, 1000, 4)


def mdigest():
    h = hashlib.sha256()
    for out in (m.SearchByProjection([F] * 8, [MP] * 8, 3.0, False, 0.0, 0.8), m.SearchByBoW([kfp[0]] * 8, [kfp[1]] * 8, 0.7, True),
                m.SearchForInitialization([ini[0]] * 8, [ini[1]] * 8, 100, 0.9, True)):
        for item in out:
            for a in item:
                h.update(np.asarray(a).tobytes())
    return h.hexdigest()


mref = mdigest()
for r in range(REPS):
    if mdigest() != mref:
        bad += 1
        print("MATCHER MISMATCH at repetition", r, flush=True)
print(f"extraction + stereo: {REPS} repetitions of {B} frames; matchers: {REPS} repetitions of 3 x 8 problems; mismatches: {bad}")
sys.exit(1 if bad else 0)
