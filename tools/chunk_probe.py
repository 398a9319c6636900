#!/usr/bin/env python3
"""Device-resident throughput when a 256-frame step is split over G handles (streams) of 256/G frames.
Run on the GPU box: python tools/chunk_probe.py"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import orbx_pkg
pkg = orbx_pkg.load()
orbx = pkg.orbx
import bench
W, H, B = bench.W, bench.H, 256
frames = bench.make_frames(B, seed=0xB200)
dev = [torch.from_numpy(np.roll(frames, k, axis=0)).cuda() for k in range(2)]
for G in (1, 2, 4, 8, 16):
    b = B // G
    exs = [orbx.ORBextractor(bench.NFEAT, bench.SCALE, bench.NLEVELS, bench.INI_TH, bench.MIN_TH, device=0, max_batch=b).prepare(W, H, b) for _ in range(G)]
    def step(i):
        base = dev[i & 1].data_ptr()
        for g, ex in enumerate(exs):
            ex.extract_device(base + g * b * W * H, b, W, H)
    for i in range(4): step(i)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    N = 30
    for i in range(N): step(i)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / N
    print(f"G={G:2d} batch={b:3d}: {dt*1e3:.3f} ms/step  {B/dt:,.0f} frames/s", flush=True)
    for ex in exs: ex.close()
