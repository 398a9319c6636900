#!/usr/bin/env python3
"""Per-level k_octree launches (dataflow path, batch 8) a few times: run under
ncu --metrics gpu__time_duration.sum -k regex:k_octree to read the latency of one (frame, level) CTA per level."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import orbx_pkg
pkg = orbx_pkg.load(); ox = pkg.orbx
W, H, B = 752, 480, 8
frames = pkg.synth.frame_batch(B, W, H, seed=0xB200, distinct=B)
ex = ox.ORBextractor(1000, 1.2, 8, 20, 7, max_batch=B).prepare(W, H, B)
pin = ox.PinnedBuffer((B, H, W)); pin.array[:] = frames
for i in range(4):
    ex.extract_host(pin.array); ex.sync()
print("done")
