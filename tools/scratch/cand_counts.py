import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import orbx_pkg
pkg = orbx_pkg.load(); ox = pkg.orbx
W, H, B = 752, 480, 8
frames = pkg.synth.frame_batch(B, W, H, seed=0xB200, distinct=8)
ex = ox.ORBextractor(1000, 1.2, 8, 20, 7, max_batch=B).prepare(W, H, B)
ex.extract_host(frames); ex.sync()
for f in range(4):
    print(f, [len(ex.candidates(l, f)[0]) for l in range(8)])
