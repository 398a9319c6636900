"""GPU parity of the ROS-side pre-processing (image_grabber.hpp:96-110): the CUDA kernels of
csrc/preproc.cu through the C ABI against the cv2 golden vectors, the CPU oracle and -- when cv2 is importable -- the
cv2 functions themselves.  u8 results: bit-exact."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def GP(golden_dir):
    return np.load(os.path.join(golden_dir, "cv2_preproc.npz"))


def _case(synth, GP, name):
    sw, sh, mw, mh, dw, dh, seed = (int(v) for v in GP[f"{name}_cfg"])
    img = synth.bgr_image(sw, sh, seed)
    m1, m2 = synth.undistortion_maps(sw, sh, mw, mh, zoom=float(GP[f"{name}_zoom"]))
    return img, m1, m2, dw, dh


MODES = {"auto": {}, "two_kernels": {"ORBX_PREPROC_FUSED": "0"}, "fused": {"ORBX_PREPROC_FUSED": "1"}}


@pytest.fixture(params=list(MODES))
def mode(request, monkeypatch):
    """Both kernel paths (remap -> resize+gray, and the one-pass kernel) on every case: the library picks one by the
    resize ratio at create time, ORBX_PREPROC_FUSED forces it."""
    monkeypatch.delenv("ORBX_PREPROC_FUSED", raising=False)
    for k, v in MODES[request.param].items():
        monkeypatch.setenv(k, v)
    return request.param


@pytest.mark.parametrize("name", ["half", "odd", "same", "up"])
def test_convert_image_vs_cv2_golden(pkg, synth, GP, name, mode):
    img, m1, m2, dw, dh = _case(synth, GP, name)
    g = pkg.orbx.ImageGrabber((img.shape[1], img.shape[0]), m1, m2, dw, dh, max_batch=1)
    out = g.ConvertImageToGPU(img[None])
    g.close()
    assert np.array_equal(out[0], GP[f"{name}_gray"])


def test_convert_batch_vs_oracle(pkg, oracle, synth, mode):
    """A batch of different frames through one launch (the weights of a map pixel are reused across the batch), rows of
    the host buffer padded: every frame equals the oracle's."""
    sw, sh, dw, dh, B = 400, 300, 211, 157, 5
    m1, m2 = synth.undistortion_maps(sw, sh, sw, sh, zoom=0.75)
    padded = np.zeros((B, sh, sw + 7, 3), np.uint8)
    for f in range(B):
        padded[f, :, :sw] = synth.bgr_image(sw, sh, 40 + f)
    frames = padded[:, :, :sw]
    g = pkg.orbx.ImageGrabber((sw, sh), m1, m2, dw, dh, max_batch=8)
    out = g.ConvertImageToGPU(frames)
    for f in range(B):
        assert np.array_equal(out[f], oracle.preprocess_bgr(np.ascontiguousarray(frames[f]), m1, m2, dw, dh)), f
    with pytest.raises(pkg.orbx.OrbxError):
        g.ConvertImageToGPU(np.zeros((9, sh, sw, 3), np.uint8))           # beyond max_batch: loud, no fallback
    g.close()


def test_map_edge_cases_vs_oracle(pkg, oracle, synth, mode):
    """Maps that point exactly at the borders, half a pixel outside, far outside, and at +-1/64 px rounding ties."""
    sw, sh = 64, 48
    rng = np.random.default_rng(5)
    img = rng.integers(0, 256, (sh, sw, 3), dtype=np.uint8)
    xs = np.concatenate([np.array([-40, -3, -2, -1.5, -1, -0.5, 0, 0.015625, 0.5, 1, 2, 2.984375]),
                         sw - np.array([5, 4, 3.5, 3, 2, 1.5, 1, 0.5, 0, -1, -2, -3, -50]), rng.uniform(-4, sw + 4, 71)]).astype(np.float32)
    ys = np.concatenate([np.array([-30, -3, -2, -1, -0.5, 0, 0.5, 1]), sh - np.array([4, 3, 2, 1, 0.5, 0, -1, -2, -3, -20]),
                         rng.uniform(-4, sh + 4, 46)]).astype(np.float32)
    m1 = np.ascontiguousarray(np.broadcast_to(xs[None, :], (len(ys), len(xs))), np.float32)
    m2 = np.ascontiguousarray(np.broadcast_to(ys[:, None], (len(ys), len(xs))), np.float32)
    g = pkg.orbx.ImageGrabber((sw, sh), m1, m2, m1.shape[1], m1.shape[0])
    out = g.ConvertImageToGPU(img[None])[0]
    g.close()
    assert np.array_equal(out, oracle.preprocess_bgr(img, m1, m2, m1.shape[1], m1.shape[0]))


def test_reference_camera_live_cv2(pkg, synth, mode):
    """The reference node's shape: 2048x1536 BGR, fisheye maps, resize_factor 0.3 (mono_inertial_node.cpp:20,59-68)."""
    cv2 = pytest.importorskip("cv2")
    import cv2_compose as cc
    cv2.setNumThreads(1)
    m1, m2 = cc.fisheye_maps(2048, 1536)
    w, h = int(2048 * 0.3), int(1536 * 0.3)
    frames = np.stack([synth.bgr_image(2048, 1536, 21 + f) for f in range(2)])
    g = pkg.orbx.ImageGrabber((2048, 1536), m1, m2, w, h, max_batch=2)
    out = g.ConvertImageToGPU(frames)
    g.close()
    for f in range(2):
        assert np.array_equal(out[f], cc.convert_image(frames[f], m1, m2, w, h)), f


def test_convert_and_extract_chain(pkg, oracle, synth):
    """ConvertImageToGPU -> ORBextractor::operator() chained on the device equals pre-processing on the oracle and
    extracting its gray image on the GPU from the host (the extractor's own parity is tests/test_gpu_extractor.py)."""
    sw, sh, dw, dh, B = 1024, 768, 640, 480, 3
    m1, m2 = synth.undistortion_maps(sw, sh, sw, sh, zoom=0.9)
    frames = np.stack([synth.bgr_image(sw, sh, 60 + f) for f in range(B)])
    g = pkg.orbx.ImageGrabber((sw, sh), m1, m2, dw, dh, max_batch=B)
    ex = pkg.orbx.ORBextractor(1000, 1.2, 8, 20, 7, max_batch=B)
    g.convert_and_extract(frames, ex)
    ex.sync()
    got = [ex.fetch(f) for f in range(B)]
    gray = np.stack([oracle.preprocess_bgr(frames[f], m1, m2, dw, dh) for f in range(B)])
    want = ex.extract_batch(gray)
    for f in range(B):
        assert len(got[f][0]) > 500
        assert np.array_equal(got[f][0], want[f][0]) and np.array_equal(got[f][1], want[f][1]) and got[f][2] == want[f][2]
    g.close(); ex.close()


def test_convert_device_resident(pkg, oracle, synth):
    sw, sh, dw, dh = 320, 240, 200, 150
    m1, m2 = synth.undistortion_maps(sw, sh, sw, sh, zoom=0.85)
    img = synth.bgr_image(sw, sh, 77)
    g = pkg.orbx.ImageGrabber((sw, sh), m1, m2, dw, dh)
    d_in = pkg.orbx.DeviceBuffer(img.nbytes); d_out = pkg.orbx.DeviceBuffer(dw * dh)
    d_in.upload(img)
    g.convert_device(d_in.ptr, 1, d_out.ptr)
    out = np.empty((dh, dw), np.uint8)
    pkg.orbx._ck(pkg.orbx.lib().orbx_device_synchronize())
    pkg.orbx._ck(pkg.orbx.lib().orbx_memcpy_d2h(out.ctypes.data, d_out.ptr, out.nbytes))
    assert np.array_equal(out, oracle.preprocess_bgr(img, m1, m2, dw, dh))
    d_in.free(); d_out.free(); g.close()


def test_one_grabber_feeds_rotating_extractors(pkg, oracle, synth):
    """One pre-processing handle in front of three extractor handles (three streams), calls issued back to back without
    a sync in between: the handle's staging buffers are reused by every call, so each call must be ordered after the
    previous one's last use on the device.  Every extractor ends up with its own batch's result."""
    sw, sh, dw, dh, B = 800, 600, 480, 360, 4
    m1, m2 = synth.undistortion_maps(sw, sh, sw, sh, zoom=0.9)
    g = pkg.orbx.ImageGrabber((sw, sh), m1, m2, dw, dh, max_batch=B)
    exs = [pkg.orbx.ORBextractor(700, 1.2, 6, 20, 7, max_batch=B) for _ in range(3)]
    batches = [np.stack([synth.bgr_image(sw, sh, 500 + 10 * k + f) for f in range(B)]) for k in range(3)]
    for rnd in range(2):
        for k in range(3):
            g.convert_and_extract(batches[(k + rnd) % 3], exs[k])
    got = []
    for k in range(3):
        exs[k].sync()
        got.append([exs[k].fetch(f) for f in range(B)])
    ref = pkg.orbx.ORBextractor(700, 1.2, 6, 20, 7, max_batch=B)
    for k in range(3):
        gray = np.stack([oracle.preprocess_bgr(batches[(k + 1) % 3][f], m1, m2, dw, dh) for f in range(B)])
        want = ref.extract_batch(gray)
        for f in range(B):
            assert np.array_equal(got[k][f][0], want[f][0]) and np.array_equal(got[k][f][1], want[f][1]), (k, f)
    g.close(); ref.close()
    for e in exs:
        e.close()
