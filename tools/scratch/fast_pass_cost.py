import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import orbx_pkg
pkg = orbx_pkg.load(); ox = pkg.orbx
W, H, B = 752, 480, 256
frames = pkg.synth.frame_batch(B, W, H, seed=0xB200, distinct=8)
for ini, mn in ((20, 7), (20, 20), (7, 7)):
    ex = ox.ORBextractor(1000, 1.2, 8, ini, mn, max_batch=B).prepare(W, H, B)
    dev = ox.DeviceBuffer(frames.nbytes); dev.upload(frames)
    ex.set_profiling(True)
    st = np.zeros(6)
    for i in range(8):
        ex.extract_device(dev.ptr, B, W, H)
        if i >= 3: st += ex.stage_ms()
    print(ini, mn, "fast ms", st[1] / 5, "octree", st[2] / 5)
    ex.close()
