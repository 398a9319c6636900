#!/usr/bin/env python3
"""Device time of Frame::ComputeStereoMatches on 128 EuRoC-shape pairs (per-kernel via CUDA events)."""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import orbx_pkg
pkg = orbx_pkg.load(); ox = pkg.orbx
W, H, B = 752, 480, 256
pairs = [pkg.synth.stereo_pair(W, H, 0xB200 + i) for i in range(4)]
base = np.stack([im for p in pairs for im in p])
frames = np.stack([base[i % 8] for i in range(B)])
ex = ox.ORBextractor(1000, 1.2, 8, 20, 7, max_batch=B).prepare(W, H, B)
ex.extract_host(frames); ex.sync()
lf = np.arange(0, B, 2); rf = lf + 1
st = torch.cuda.ExternalStream(ex.stream())
for _ in range(3):
    ox.stereo_matches_async(ex, ex, lf, rf, 435.2 * 0.11, 0.11)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(st)
for _ in range(20):
    ox.stereo_matches_async(ex, ex, lf, rf, 435.2 * 0.11, 0.11)
e1.record(st); torch.cuda.synchronize()
ur, _ = ox.stereo_fetch(ex, len(lf))
print(f"stereo: {e0.elapsed_time(e1) / 20 * 1e3:.1f} us per 128 pairs, matched/pair {(ur >= 0).sum() / len(lf):.1f}")
