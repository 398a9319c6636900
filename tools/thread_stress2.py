#!/usr/bin/env python3
"""Variant of thread_stress.py: T worker threads run extraction + 2-NN on PERSISTENT handles while one extra thread only
creates (and closes) extractor handles, never using them.  usage: thread_stress2.py [workers] [iters] [creator mode: create|use|none]"""
import os, sys, threading, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import orbx_pkg
pkg = orbx_pkg.load(); ox, synth = pkg.orbx, pkg.synth
T = int(sys.argv[1]) if len(sys.argv) > 1 else 5
N = int(sys.argv[2]) if len(sys.argv) > 2 else 150
MODE = sys.argv[3] if len(sys.argv) > 3 else "create"
OPS = sys.argv[4] if len(sys.argv) > 4 else "ek"
imgs = [synth.textured_image(400, 300, 900 + i) for i in range(4)]
q, db, _ = synth.descriptor_sets(512, 8192, seed=3)
errors = []
stop = False


def worker(tid):
    try:
        ex = ox.ORBextractor(500, 1.2, 6, 20, 7); m = ox.ORBmatcher()
        for it in range(N):
            if "e" in OPS: ex(imgs[(tid + it) % 4], (0, 0))
            if "k" in OPS: m.knn2(q, db, 0.8)
    except Exception as e:       # noqa: BLE001
        errors.append(("worker", tid, it, repr(e)[:200]))


def creator():
    n = 0
    try:
        while not stop:
            if MODE == "none":
                time.sleep(0.01); continue
            ex = ox.ORBextractor(500, 1.2, 6, 20, 7)
            ex.prepare(400, 300, 1)
            if MODE == "use":
                ex(imgs[0], (0, 0))
            ex.close(); n += 1
    except Exception as e:       # noqa: BLE001
        errors.append(("creator", n, repr(e)[:200]))
    print("creator made", n, "handles")


th = [threading.Thread(target=worker, args=(t,)) for t in range(T)]
c = threading.Thread(target=creator)
for t in th: t.start()
c.start()
for t in th: t.join()
stop = True
c.join()
print(f"workers {T} iters {N} creator {MODE} ops {OPS}: {len(errors)} errors")
for e in errors[:4]:
    print("  ", e)
sys.exit(1 if errors else 0)
