#!/usr/bin/env python3
"""Per-stage CUDA-event times of the extractor at small batch sizes (latency view)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import orbx_pkg
pkg = orbx_pkg.load(); ox = pkg.orbx
W, H = 752, 480
for B in (1, 2, 8, 32):
    frames = pkg.synth.frame_batch(B, W, H, seed=0xB200, distinct=min(B, 8))
    ex = ox.ORBextractor(1000, 1.2, 8, 20, 7, max_batch=B).prepare(W, H, B)
    pin = ox.PinnedBuffer((B, H, W)); pin.array[:] = frames
    ex.set_profiling(True)
    acc = np.zeros(6); wall = []
    for i in range(60):
        t0 = time.perf_counter()
        ex.extract_host(pin.array); ex.sync()
        wall.append(time.perf_counter() - t0)
        if i >= 10: acc += ex.stage_ms()
    acc /= 50
    print(f"batch {B}: wall median {np.median(wall[10:])*1e3:.3f} ms; stages(us) pyr {acc[0]*1e3:.0f} fast {acc[1]*1e3:.0f} oct {acc[2]*1e3:.0f} order {acc[3]*1e3:.0f} blur {acc[4]*1e3:.0f} desc {acc[5]*1e3:.0f} sum {acc.sum()*1e3:.0f}")
    ex.close()
