#!/usr/bin/env python3
"""Wall-clock latency of one extract_host + sync (+ fetch) at batch 1..8, as shipped (no profiling)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import orbx_pkg
pkg = orbx_pkg.load(); ox = pkg.orbx
W, H = 752, 480
for B in (1, 2, 8):
    frames = pkg.synth.frame_batch(B, W, H, seed=0xB200, distinct=min(B, 8))
    ex = ox.ORBextractor(1000, 1.2, 8, 20, 7, max_batch=B).prepare(W, H, B)
    pin = ox.PinnedBuffer((B, H, W)); pin.array[:] = frames
    wall = []; wall2 = []
    for i in range(300):
        t0 = time.perf_counter()
        ex.extract_host(pin.array); ex.sync()
        t1 = time.perf_counter()
        ex.fetch(0)
        wall.append(t1 - t0); wall2.append(time.perf_counter() - t0)
    print(f"batch {B}: extract+sync median {np.median(wall[50:])*1e3:.3f} ms  p95 {np.percentile(wall[50:],95)*1e3:.3f}; +fetch {np.median(wall2[50:])*1e3:.3f} ms", flush=True)
    ex.close()
