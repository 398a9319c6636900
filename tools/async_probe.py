#!/usr/bin/env python3
"""Do the asynchronous entry points return while the stream is still busy?  Host time per call of back-to-back
orbx_extract_device calls WITH a lapping table (staged from pageable memory) and without, against the device time."""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import orbx_pkg
pkg = orbx_pkg.load()
W, H, B = 752, 480, 256
ex = pkg.orbx.ORBextractor(1000, 1.2, 8, 20, 7, max_batch=B).prepare(W, H, B)
fr = torch.from_numpy(pkg.synth.frame_batch(B, W, H, seed=1, distinct=8)).cuda()
laps = np.tile(np.array([[100, 300]], np.int32), (B, 1))
st = torch.cuda.ExternalStream(ex.stream())
for lap in (None, laps):
    for _ in range(3):
        ex.extract_device(fr.data_ptr(), B, W, H, lapping=lap)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    host = []
    for _ in range(10):
        t0 = time.perf_counter(); ex.extract_device(fr.data_ptr(), B, W, H, lapping=lap); host.append(time.perf_counter() - t0)
    e1.record(st); torch.cuda.synchronize()
    print("lapping" if lap is not None else "no lapping", "host ms/call", [round(h * 1e3, 3) for h in host], "device ms/call", round(e0.elapsed_time(e1) / 10, 3))
