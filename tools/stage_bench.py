#!/usr/bin/env python3
"""Quick A/B number for kernel work: device-resident frames/s of the bench workload (EuRoC shape, batch 256, 2 alternating
input batches) and the per-stage CUDA-event times, nothing else.  ORBX_LIB_DIR / ORBX_* tuning variables select the
variant.  usage: python tools/stage_bench.py [steps] [label]"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import orbx_pkg  # noqa: E402

pkg = orbx_pkg.load()
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
W, H, B = 752, 480, 256
ex = pkg.orbx.ORBextractor(1000, 1.2, 8, 20, 7, max_batch=B).prepare(W, H, B)
fr = pkg.synth.frame_batch(B, W, H, seed=0xB200, distinct=8)
dev = [torch.from_numpy(np.roll(fr, k, axis=0)).cuda() for k in range(2)]
st = torch.cuda.ExternalStream(ex.stream())
for i in range(4):
    ex.extract_device(dev[i & 1].data_ptr(), B, W, H)
torch.cuda.synchronize()
best = 1e9
for rep in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for i in range(steps):
        ex.extract_device(dev[i & 1].data_ptr(), B, W, H)
    e1.record(st)
    torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1) / steps)
ex.set_profiling(True)
stage = np.zeros(6)
for i in range(steps):
    ex.extract_device(dev[i & 1].data_ptr(), B, W, H)
    stage += ex.stage_ms()
stage /= steps
ex.sync()
n, _ = ex.counts(B)
print(json.dumps({"label": sys.argv[2] if len(sys.argv) > 2 else "", "frames_per_s": round(B / best * 1e3), "ms_per_step": round(best, 4),
                  "stage_ms": dict(zip(["pyramid", "fast", "octree", "order", "blur", "describe"], [round(float(v), 4) for v in stage])),
                  "sum_stages": round(float(stage.sum()), 4), "kp": float(np.mean(n))}))
