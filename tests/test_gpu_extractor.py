"""GPU parity: CUDA extractor (through the C ABI) vs the CPU oracle on identical seeded
synthetic frames.  Pyramid, FAST candidates, keypoint coordinates/octaves/responses and
order: bit-exact.  Orientation: <= 1e-4 rad.  Descriptors: bit-exact except a documented
cos/sin ULP allowance (<= 0.1% of descriptors, <= 2 bits each) -- BASELINE.json north_star."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

CONFIGS = [
    # (w, h, nfeatures, nlevels, seed)
    (752, 480, 1000, 8, 0xB200),      # EuRoC
    (512, 512, 1500, 8, 7),           # TUM-VI
    (1241, 376, 2000, 8, 11),         # KITTI
    (320, 240, 300, 4, 3),
    (276, 200, 300, 3, 13),           # 6 cells of 41 px per row: widest FAST unit strip (TMA box limit)
]


POOL = {"kps": 0, "bad": 0}     # every keypoint compared in this session: the <= 0.1 % bound is also asserted pooled


def _assert_pooled(nk, nbad, what):
    """north_star: <= 0.1 % of descriptors may differ.  Per frame the test has to allow 1 (a frame has
    ~1000 keypoints); over the pooled keypoints of a multi-frame test the bound is applied as stated, so
    a systematic one-per-frame flip cannot hide."""
    assert nbad <= 0.001 * nk, f"{what}: {nbad} of {nk} descriptors differ (> 0.1 %)"


def _compare_frame(gex, oex, img, lapping, f=0, got=None, ref=None):
    ok, od, omono = ref if ref is not None else oex(img, lapping)
    gk, gd, gmono = got
    assert len(gk) == len(ok), (len(gk), len(ok))
    assert gmono == omono
    for fld in ("x", "y", "size", "response", "octave", "class_id"):
        assert np.array_equal(gk[fld], ok[fld]), fld
    if len(gk) == 0:
        return 0
    dang = np.abs(gk["angle"] - ok["angle"])
    dang = np.minimum(dang, 360 - dang)
    assert dang.max() <= np.degrees(1e-4)
    bits = np.unpackbits(gd ^ od, axis=1).sum(1)
    nbad = int((bits > 0).sum())
    assert nbad <= max(1, int(0.001 * len(gk))) and bits.max() <= 2, (nbad, bits.max())
    POOL["kps"] += len(gk); POOL["bad"] += nbad
    return nbad


@pytest.mark.parametrize("w,h,nf,nl,seed", CONFIGS)
def test_full_extraction_parity(pkg, oracle, synth, w, h, nf, nl, seed):
    img = synth.textured_image(w, h, seed)
    gex = pkg.orbx.ORBextractor(nf, 1.2, nl, 20, 7)
    oex = oracle.Extractor(nf, 1.2, nl, 20, 7)
    got = gex(img, (0, 0))
    # tables
    assert np.array_equal(gex.mvScaleFactor, oex.scale) and np.array_equal(gex.mvInvScaleFactor, oex.inv_scale)
    assert np.array_equal(gex.mvLevelSigma2, oex.sigma2) and np.array_equal(gex.mvInvLevelSigma2, oex.inv_sigma2)
    assert np.array_equal(gex.mnFeaturesPerLevel, oex.feat_per_level) and np.array_equal(gex.umax, oex.umax)
    _compare_frame(gex, oex, img, (0, 0), got=got)
    # stage-wise: pyramid (bordered), FAST candidates in reference order, blur
    for l in range(nl):
        assert np.array_equal(gex.image_pyramid(l), oex.level(l)), f"pyramid level {l}"
        gx, gy, gr = gex.candidates(l)
        ox, oy, orr = oex.candidates(l)
        assert np.array_equal(gx, ox) and np.array_equal(gy, oy) and np.array_equal(gr, orr), f"candidates level {l}"
        lv = oex.level(l)
        assert np.array_equal(gex.blurred(l), oracle.gaussian7(lv[19:-19, 19:-19])), f"blur level {l}"
    gex.close()


def test_lapping_area_and_batch(pkg, oracle, synth):
    """TUM-VI style: vLappingArea split, several frames per launch, per-frame lapping."""
    w, h, B = 512, 512, 6
    imgs = synth.frame_batch(B, w, h, seed=21, distinct=3)
    laps = [(0, 0), (256, 511), (0, 255), (100, 300), (0, 1000), (511, 511)]
    gex = pkg.orbx.ORBextractor(1500, 1.2, 8, 20, 7, max_batch=B)
    oex = oracle.Extractor(1500, 1.2, 8, 20, 7)
    res = gex.extract_batch(imgs, laps)
    nbad = sum(_compare_frame(gex, oex, imgs[f], laps[f], got=res[f]) for f in range(B))
    _assert_pooled(sum(len(r[0]) for r in res), nbad, "lapping batch")
    gex.close()


def test_low_texture_retry_and_blank(pkg, oracle, synth):
    """Faint texture forces the per-cell minThFAST retry; a constant image yields zero keypoints."""
    img = (synth.textured_image(400, 300, 3).astype(np.float32) * 0.12 + 100).astype(np.uint8)
    gex = pkg.orbx.ORBextractor(500, 1.2, 4, 20, 7)
    oex = oracle.Extractor(500, 1.2, 4, 20, 7)
    _compare_frame(gex, oex, img, (0, 0), got=gex(img, (0, 0)))
    blank = np.full((300, 400), 77, np.uint8)
    k, d, mono = gex(blank, (0, 0))
    assert len(k) == 0 and mono == 0
    k, d, mono = gex(np.zeros((0, 0), np.uint8))
    assert mono == -1 and k is None
    gex.close()


def test_noise_image_dense_candidates(pkg, oracle):
    """Uniform noise: the densest FAST response the detector can produce (cell slot capacity)."""
    rng = np.random.default_rng(99)
    img = rng.integers(0, 256, size=(240, 320), dtype=np.uint8)
    gex = pkg.orbx.ORBextractor(800, 1.2, 4, 20, 7)
    oex = oracle.Extractor(800, 1.2, 4, 20, 7)
    _compare_frame(gex, oex, img, (0, 0), got=gex(img, (0, 0)))
    gex.close()


def test_fast_nms_scan_path(pkg, oracle, synth):
    """A tiny FAST corner list forces every region through the score-map scan variant of the
    non-maximum suppression; results must not change."""
    img = synth.textured_image(640, 400, seed=31)
    gex = pkg.orbx.ORBextractor(900, 1.2, 6, 20, 7).prepare(640, 400, 1)
    gex.set_fast_corner_list(4)
    oex = oracle.Extractor(900, 1.2, 6, 20, 7)
    _compare_frame(gex, oex, img, (0, 0), got=gex(img, (0, 0)))
    gex.close()


def test_many_frames_property(pkg, oracle, synth):
    """Larger batch: every frame equals the oracle; repeated launches are deterministic."""
    w, h, B = 752, 480, 24
    imgs = synth.frame_batch(B, w, h, seed=0xB200, distinct=6)
    gex = pkg.orbx.ORBextractor(1000, 1.2, 8, 20, 7, max_batch=B)
    oex = oracle.Extractor(1000, 1.2, 8, 20, 7)
    r1 = gex.extract_batch(imgs)
    r2 = gex.extract_batch(imgs)
    for f in range(B):
        assert np.array_equal(r1[f][0], r2[f][0]) and np.array_equal(r1[f][1], r2[f][1])
    nbad = sum(_compare_frame(gex, oex, imgs[f], (0, 0), got=r1[f]) for f in range(B))
    _assert_pooled(sum(len(r[0]) for r in r1), nbad, "24-frame batch")
    gex.close()


@pytest.mark.parametrize("lanes", [None, "4"])
def test_chunked_batch_streams(pkg, oracle, synth, monkeypatch, lanes):
    """A batch large enough to run as two chunks on the handle's internal streams: every frame
    equals the single-stream result and (sampled) the oracle; the stereo search on the
    device-resident output of such a batch equals the one on a single-stream batch."""
    w, h, B = 640, 400, 72
    sp = [synth.stereo_pair(w, h, 0xB200 + 77 + i) for i in range(4)]           # frames (2p, 2p+1) = a rectified pair
    base = [im for pr in sp for im in pr]
    imgs = np.stack([np.roll(base[i % 8], 5 * (i // 8), axis=1) for i in range(B)])
    if lanes:
        monkeypatch.setenv("ORBX_LANES", lanes)      # read when the handle is created
    gex = pkg.orbx.ORBextractor(800, 1.2, 8, 20, 7, max_batch=B)
    assert gex.prepare(w, h, B).streams(B) == int(lanes or 2) and gex.streams(8) == 1
    monkeypatch.delenv("ORBX_LANES", raising=False)
    lap = np.tile(np.array([[100, 300]], np.int32), (B, 1))
    r = gex.extract_batch(imgs, lap)
    lf = np.arange(0, B, 2); rf = lf + 1
    pkg.orbx.stereo_matches_async(gex, gex, lf, rf, 40.0, 0.1)
    ur, dp = pkg.orbx.stereo_fetch(gex, len(lf))
    one = pkg.orbx.ORBextractor(800, 1.2, 8, 20, 7, max_batch=8)
    oex = oracle.Extractor(800, 1.2, 8, 20, 7)
    for f0 in range(0, B, 8):
        q = one.extract_batch(imgs[f0:f0 + 8], lap[f0:f0 + 8])
        for k in range(8):
            assert np.array_equal(r[f0 + k][0], q[k][0]) and np.array_equal(r[f0 + k][1], q[k][1]) and r[f0 + k][2] == q[k][2], f0 + k
        pkg.orbx.stereo_matches_async(one, one, np.arange(0, 8, 2), np.arange(1, 8, 2), 40.0, 0.1)
        ur1, dp1 = pkg.orbx.stereo_fetch(one, 4)
        for k in range(4):
            n = len(q[2 * k][0])
            assert np.array_equal(ur[f0 // 2 + k][:n], ur1[k][:n]) and np.array_equal(dp[f0 // 2 + k][:n], dp1[k][:n])
    for f in (0, 17, 35, 54, 71):
        _compare_frame(gex, oex, imgs[f], (100, 300), got=r[f])
    gex.close(); one.close()


def test_handles_of_different_size_coexist(pkg, oracle, synth):
    """Shared-memory opt-ins are per kernel, not per handle: a small handle created after a large one
    must not shrink the limit the large one launches with."""
    img = synth.textured_image(1241, 376, 901)
    big = pkg.orbx.ORBextractor(5000, 1.2, 8, 20, 7)
    rb = big(img, (0, 0))
    small = pkg.orbx.ORBextractor(200, 1.2, 4, 20, 7)
    small(synth.textured_image(320, 240, 902), (0, 0))
    rb2 = big(img, (0, 0))
    assert np.array_equal(rb[0], rb2[0]) and np.array_equal(rb[1], rb2[1]) and len(rb[0]) > 4000
    _compare_frame(big, oracle.Extractor(5000, 1.2, 8, 20, 7), img, (0, 0), got=rb2)
    big.close(); small.close()


def test_hamming_pairs(pkg, oracle):
    rng = np.random.default_rng(5)
    a = rng.integers(0, 256, size=(5000, 32), dtype=np.uint8)
    b = rng.integers(0, 256, size=(5000, 32), dtype=np.uint8)
    b[:10] = a[:10]; b[10] = ~a[10]
    got = pkg.orbx.hamming_pairs(a, b)
    assert np.array_equal(got, oracle.hamming_pairs(a, b))
    assert got[:10].max() == 0 and got[10] == 256


def test_cpp_shim_matches_oracle(pkg, oracle, synth, tmp_path):
    """include/orbx/ORBextractor.h (C++ drop-in) -> same keypoints/descriptors/pyramid as the oracle."""
    from test_abi import build_shim_test
    exe = build_shim_test(pkg.orbx.build(), tmp_path)
    import subprocess
    w, h, nf = 640, 400, 700
    img = synth.textured_image(w, h, 5)
    (tmp_path / "img.raw").write_bytes(img.tobytes())
    out = tmp_path / "out.bin"
    r = subprocess.run([exe, str(tmp_path / "img.raw"), str(w), str(h), str(nf), str(out)], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.startswith("OK"), r.stderr
    raw = out.read_bytes()
    n, mono, l1w, l1h, d01, nsf = np.frombuffer(raw[:24], np.int32)
    kps = np.frombuffer(raw[24:24 + 28 * n], pkg.orbx.KP_DTYPE)
    desc = np.frombuffer(raw[24 + 28 * n:24 + 60 * n], np.uint8).reshape(n, 32)
    lvl1 = np.frombuffer(raw[24 + 60 * n:], np.uint8).reshape(l1h + 38, l1w + 38)
    oex = oracle.Extractor(nf, 1.2, 8, 20, 7)
    _compare_frame(None, oex, img, (0, 0), got=(kps, desc, int(mono)))
    assert np.array_equal(lvl1, oex.level(1)) and nsf == 8
    assert d01 == oracle.descriptor_distance(desc[0], desc[1])


@pytest.mark.parametrize("w,h,nf,scale,nl,ini,mn,seed", [
    (640, 480, 1200, 1.1, 12, 20, 7, 201),        # many shallow levels
    (640, 480, 800, 1.5, 4, 20, 7, 202),          # few steep levels
    (512, 384, 500, 1.2, 8, 40, 40, 203),         # iniThFAST == minThFAST: no retry pass
    (512, 384, 3000, 1.2, 8, 12, 5, 204),         # more features than most levels can supply
    (400, 300, 600, 1.25, 6, 130, 20, 205),       # threshold above 127 (other branch of the packed compare)
    (400, 300, 600, 1.25, 6, 200, 140, 208),      # both thresholds above 127: the retry pass takes that branch too
    (333, 257, 400, 1.2, 5, 20, 7, 206),          # odd sizes
    (1920, 1080, 2000, 1.2, 8, 20, 7, 207),       # full HD
])
def test_parameter_sweep(pkg, oracle, synth, w, h, nf, scale, nl, ini, mn, seed):
    img = synth.textured_image(w, h, seed)
    gex = pkg.orbx.ORBextractor(nf, scale, nl, ini, mn)
    oex = oracle.Extractor(nf, scale, nl, ini, mn)
    _compare_frame(gex, oex, img, (0, 0), got=gex(img, (0, 0)))
    for l in range(nl):
        assert np.array_equal(gex.image_pyramid(l), oex.level(l)), f"pyramid level {l}"
        gx, gy, gr = gex.candidates(l)
        ox, oy, orr = oex.candidates(l)
        assert np.array_equal(gx, ox) and np.array_equal(gy, oy) and np.array_equal(gr, orr), f"candidates level {l}"
    gex.close()


def test_async_entry_points_return_while_the_stream_is_busy(pkg, synth):
    """orbx_extract_device with a lapping table stages it from pageable host memory: the call must still
    return while the device is busy -- host time per call far below the device time per call -- or the rotating-handle
    pipeline would serialise.  (Measured: 0.135 ms host against 1.27 ms device per 256-frame call.)"""
    import time
    torch = pytest.importorskip("torch")
    w, h, B = 752, 480, 128
    ex = pkg.orbx.ORBextractor(1000, 1.2, 8, 20, 7, max_batch=B).prepare(w, h, B)
    fr = torch.from_numpy(synth.frame_batch(B, w, h, seed=1, distinct=4)).cuda()
    laps = np.tile(np.array([[100, 300]], np.int32), (B, 1))
    st = torch.cuda.ExternalStream(ex.stream())
    for _ in range(3):
        ex.extract_device(fr.data_ptr(), B, w, h, lapping=laps)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    t0 = time.perf_counter()
    for _ in range(8):
        ex.extract_device(fr.data_ptr(), B, w, h, lapping=laps)
    host_ms = (time.perf_counter() - t0) * 1e3 / 8
    e1.record(st); torch.cuda.synchronize()
    dev_ms = e0.elapsed_time(e1) / 8
    ex.close()
    assert host_ms < 0.5 * dev_ms, (host_ms, dev_ms)


GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cv2_extractor.npz")


@pytest.mark.parametrize("name", ["euroc", "tumvi", "kitti"])
def test_cuda_vs_cv2_composition_golden(pkg, synth, name):
    """The CUDA extractor against the whole-extractor composition from cv2 primitives + a straight Python
    quadtree (tests/cv2_compose.py, vectors by tools/gen_golden_extractor.py) -- no oracle in between."""
    G = np.load(GOLDEN)
    w, h, nf, l0, l1, seed = (int(v) for v in G[f"{name}_cfg"])
    img = synth.textured_image(w, h, seed)
    assert int(img.astype(np.int64).sum()) == int(G[f"{name}_imgsum"])
    gex = pkg.orbx.ORBextractor(nf, 1.2, 8, 20, 7)
    got = gex(img, (l0, l1))
    nbad = _compare_frame(gex, None, img, (l0, l1), got=got, ref=(G[f"{name}_kps"], G[f"{name}_desc"], int(G[f"{name}_mono"])))
    assert nbad <= 1
    gex.close()


def test_fuzz_parity_50(pkg, oracle):
    """tools/fuzz_parity.py as a test: 50 random configurations (size, features, scale, levels,
    thresholds, lapping, image kind) incl. pyramid, blurred levels and candidate lists."""
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import fuzz_parity
    rng = np.random.default_rng(2)
    nk = nbad = 0
    for it in range(50):
        good, cfg, why, k, b = fuzz_parity.run_case(pkg, oracle, rng)
        assert good, (it, cfg, why)
        nk += k; nbad += b
    _assert_pooled(nk, nbad, "fuzz")


def test_pooled_descriptor_bound():
    """Runs last in this file: the <= 0.1 % descriptor bound over every keypoint the tests above compared."""
    if POOL["kps"] < 20000:
        pytest.skip("run with the whole file to pool enough keypoints")
    _assert_pooled(POOL["kps"], POOL["bad"], "all extractor tests")


def test_device_views_of_pyramid_and_results(pkg, synth):
    """orbx_extractor_level_device / orbx_extractor_results_device: the in-place device views equal what the download
    entry points return (bordered level incl. the 19-px border around the ROI; keypoints, descriptors, counts)."""
    ox = pkg.orbx
    B, w, h = 3, 512, 384
    frames = synth.frame_batch(B, w, h, seed=77, distinct=3)
    ex = ox.ORBextractor(800, 1.2, 6, 20, 7, max_batch=4)
    res = ex.extract_batch(frames)
    L = ox.lib()
    for level in (0, 3):
        ptr, lw, lh, pitch, fs = ex.level_device(level)
        assert (lw, lh) == ex.level_size(level) and pitch >= lw + 38
        for f in (0, 2):
            rows = np.empty((lh + 38, pitch), np.uint8)
            ox._ck(L.orbx_memcpy_d2h(rows.ctypes.data, ptr + f * fs - 19 * pitch - 32, rows.nbytes))
            assert np.array_equal(rows[:, 32 - 19:32 + lw + 19], ex.image_pyramid(level, f))
    dk, dd, dn, cap = ex.results_device()
    assert cap == ex.capacity
    cnt = np.empty(2 * 4, np.int32)
    ox._ck(L.orbx_memcpy_d2h(cnt.ctypes.data, dn, cnt.nbytes))
    for f in range(B):
        k, d, mono = res[f]
        assert cnt[f] == len(k) and cnt[4 + f] == mono
        gk = np.empty(len(k), ox.KP_DTYPE); gd = np.empty((len(k), 32), np.uint8)
        ox._ck(L.orbx_memcpy_d2h(gk.ctypes.data, dk + f * cap * 28, gk.nbytes))
        ox._ck(L.orbx_memcpy_d2h(gd.ctypes.data, dd + f * cap * 32, gd.nbytes))
        assert gk.tobytes() == k.tobytes() and np.array_equal(gd, d)
    ex.close()


@pytest.mark.parametrize("w,h,nf,scale,nl", [(100, 64, 300, 1.2, 1), (4000, 64, 500, 1.2, 1), (64, 4000, 500, 1.2, 1), (300, 90, 300, 1.5, 3)])
def test_levels_without_fast_cells(pkg, oracle, synth, w, h, nf, scale, nl):
    """Detection areas under 35 px in one direction have no FAST cell (upstream's nCols / nRows == 0; the oracle
    defines: no candidates): the handle is created, such levels are skipped, the others match the oracle."""
    img = synth.textured_image(w, h, w * 7 + h)
    gex = pkg.orbx.ORBextractor(nf, scale, nl, 20, 7)
    oex = oracle.Extractor(nf, scale, nl, 20, 7)
    _compare_frame(gex, oex, img, (0, 0), got=gex(img, (0, 0)))
    gex.close()
