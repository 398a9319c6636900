"""Several host threads, each with its own handles, calling into the library at the same time (ctypes releases the GIL
for the duration of a call): extraction, stereo, projection search, 2-NN and pre-processing.  Every thread must get the
results a single thread gets; first-use paths (shared-memory opt-in map, tensor-map encoder lookup, static scratch
address) are hit concurrently because the handles are created inside the threads."""
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_concurrent_threads_own_handles(pkg, synth):
    ox = pkg.orbx
    W, H = 400, 300
    imgs = [synth.textured_image(W, H, 900 + i) for i in range(4)]
    bgr = synth.bgr_image(320, 240, 5); m1, m2 = synth.undistortion_maps(320, 240, 320, 240, zoom=0.8)
    F = synth.frame_view(600, seed=1); MP = synth.projection_problem(F, 900, seed=2)
    q, db, _ = synth.descriptor_sets(512, 8192, seed=3)

    def work(tid, out, iters):
        try:
            ex = ox.ORBextractor(500, 1.2, 6, 20, 7)
            m = ox.ORBmatcher()
            g = ox.ImageGrabber((320, 240), m1, m2, 200, 150)
            res = []
            for it in range(iters):
                k, d, mono = ex(imgs[(tid + it) % 4], (0, 0))
                n, mm = m.SearchByProjection(F, MP, 3.0, False, 0.0, 0.8)
                kn = m.knn2(q, db, 0.8)
                gray = g.ConvertImageToGPU(bgr[None])[0]
                res.append((k.tobytes(), d.tobytes(), mono, n, mm.tobytes(), kn[0].tobytes(), kn[1].tobytes(), gray.tobytes()))
            ex.close(); m.close(); g.close()
            out[tid] = res
        except Exception as e:       # noqa: BLE001
            out[tid] = e

    ref = {}
    work(0, ref, 4)
    assert not isinstance(ref[0], Exception), ref[0]
    by_img = {i: ref[0][i] for i in range(4)}           # thread 0, iteration i used image i
    out = {}
    threads = [threading.Thread(target=work, args=(t, out, 12)) for t in range(6)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    for tid in range(6):
        assert not isinstance(out[tid], Exception), out[tid]
        for it, r in enumerate(out[tid]):
            assert r == by_img[(tid + it) % 4], (tid, it)


@pytest.mark.parametrize("ops", ["ek", "espk"])
def test_handles_created_and_used_while_other_threads_run(ops):
    """Six threads re-create their extractor / matcher / pre-processing handles every 10 iterations and use them at
    once, while the other threads' kernels and copies are in flight (tools/thread_stress.py).  Regression test: the
    create-time uploads (tensor maps, FAST unit table, remap tables) were synchronous copies from pageable memory,
    whose DMA may still be in flight on the legacy default stream when the call returns; the first kernels on the
    handle's non-blocking stream then read a tensor map that had not arrived ("illegal instruction" from TMA, in about
    two of three runs of this test)."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for _ in range(3):
        r = subprocess.run([sys.executable, os.path.join(root, "tools", "thread_stress.py"), "6", "120", ops, "10"], capture_output=True,
                           text=True, timeout=300)
        assert r.returncode == 0 and " 0 errors" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def test_mbarrier_wait_is_bounded(pkg):
    """The mbarrier wait every TMA / bulk-copy consumer of the library uses (csrc/orbx_internal.h) returns on a completed
    barrier and ends in a launch failure -- not in a kernel that polls for ever -- on a barrier nothing arrives on
    (tools/mbar_timeout_probe.cu, built by csrc/Makefile with the bound shortened to ~0.5 s)."""
    import os
    import subprocess
    exe = os.path.join(os.path.dirname(pkg.orbx.build()), "mbar_probe")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "completed barrier: no error, flag 1" in r.stdout and "barrier nothing arrives on:" in r.stdout and "no error after" not in r.stdout
