"""The C++ boundary, executed: tests/cpp/test_ref_api.cpp calls every wrapper of include/orbx/ORBmatcher.h
(flattened views) and every reference-signature template of include/orbx/ORBmatcher_ref.h (the reference's static
ORBmatcher functions over shared_ptr<Frame/KeyFrame/MapPoint>, instantiated with mock SLAM types) on the GPU; the
results must equal the CPU oracle.  tests/cpp/test_shim_opencv.cpp compiles the cv::Mat / cv::KeyPoint overload of
ORBextractor against a minimal OpenCV stub and runs it, with ComputeStereoMatches and mvImagePyramid."""
import os
import struct
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def write_blob(path, arrays):
    with open(path, "wb") as f:
        for k, v in arrays.items():
            v = np.ascontiguousarray(v)
            f.write(struct.pack("<I", len(k))); f.write(k.encode()); f.write(struct.pack("<Q", v.nbytes)); f.write(v.tobytes())


def read_blob(path):
    out, b = {}, open(path, "rb").read()
    o = 0
    while o < len(b):
        nl, = struct.unpack_from("<I", b, o); o += 4
        name = b[o:o + nl].decode(); o += nl
        nb, = struct.unpack_from("<Q", b, o); o += 8
        out[name] = b[o:o + nb]; o += nb
    return out


def _frame(A, p, F, uright=False):
    A[p + "kx"] = F["kx"].astype(np.float32); A[p + "ky"] = F["ky"].astype(np.float32); A[p + "koct"] = F["koct"].astype(np.int32)
    if F.get("kangle") is not None:
        A[p + "kangle"] = F["kangle"].astype(np.float32)
    A[p + "desc"] = F["desc"]
    A[p + "grid"] = np.array([F["minx"], F["miny"], F["gw_inv"], F["gh_inv"]], np.float32)
    A[p + "gridi"] = np.array([F["gcols"], F["grows"]], np.int32)
    A[p + "scale"] = F["scale"].astype(np.float32)
    if uright and F.get("uright") is not None:
        A[p + "uright"] = F["uright"].astype(np.float32)


def _kf(A, p, K):
    for k, t in (("kx", np.float32), ("ky", np.float32), ("koct", np.int32), ("kangle", np.float32), ("desc", np.uint8),
                 ("has_mp", np.uint8), ("node_id", np.int32), ("node_off", np.int32), ("node_idx", np.int32)):
        A[p + k] = np.asarray(K[k]).astype(t)
    if K.get("uright") is not None:
        A[p + "uright"] = K["uright"].astype(np.float32)


def _fuse_pts(A, p, M):
    for k, t in (("valid", np.uint8), ("u", np.float32), ("v", np.float32), ("level", np.int32), ("desc", np.uint8)):
        A[p + k] = np.asarray(M[k]).astype(t)
    if M.get("ur") is not None:
        A[p + "ur"] = M["ur"].astype(np.float32)


def _i32(b):
    return np.frombuffer(b, np.int32)


@pytest.fixture(scope="module")
def exe(pkg, tmp_path_factory):
    so = pkg.orbx.build()
    out = str(tmp_path_factory.mktemp("cpp") / "test_ref_api")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-ffp-contract=off", "-Wall", "-Werror", "-o", out,
                           os.path.join(ROOT, "tests", "cpp", "test_ref_api.cpp"), "-L" + os.path.dirname(so), "-lorbx",
                           "-Wl,-rpath," + os.path.dirname(so)])
    return out


def test_every_cpp_wrapper_and_reference_signature(exe, oracle, synth, tmp_path):
    rng = np.random.default_rng(77)
    A = {}
    # DescriptorDistance
    a = rng.integers(0, 256, (500, 32), dtype=np.uint8); b = rng.integers(0, 256, (500, 32), dtype=np.uint8)
    b[:5] = a[:5]; b[5] = ~a[5]
    A["hd.a"], A["hd.b"] = a, b
    # SearchByProjection(Frame, MapPoints): stereo frame, per-point observations, pre-claimed keypoints
    F = synth.frame_view(800, seed=301)
    MP = synth.projection_problem(F, 1400, seed=302, dup=0.5)
    Fs, MPs = synth.add_stereo(F, MP, seed=303)
    claimed = (rng.random(800) < 0.08).astype(np.uint8)
    _frame(A, "sbp.F.", Fs, uright=True)
    for k, t in (("inview", np.uint8), ("x", np.float32), ("y", np.float32), ("level", np.int32), ("viewcos", np.float32),
                 ("depth", np.float32), ("desc", np.uint8), ("proj_xr", np.float32), ("has_obs", np.uint8)):
        A["sbp.MP." + k] = MPs[k].astype(t)
    A["sbp.claimed"] = claimed
    A["sbp.params"] = np.array([3.0, 1.0, 25.0, 0.8], np.float32)
    # SearchByProjection(Frame, KeyFrame)
    F2 = synth.frame_view(900, seed=311)
    P = synth.projection_problem(F2, 1300, seed=312)
    KP = {"valid": (P["valid"].astype(bool) & (P["x"] >= 0) & (P["y"] >= 0)).astype(np.uint8), "u": P["x"], "v": P["y"],
          "level": P["level"], "angle": P["angle"], "desc": P["desc"]}
    occ = (rng.random(900) < 0.05).astype(np.uint8)
    _frame(A, "sbpkf.F.", F2)
    for k, t in (("valid", np.uint8), ("u", np.float32), ("v", np.float32), ("level", np.int32), ("angle", np.float32), ("desc", np.uint8)):
        A["sbpkf.MP." + k] = KP[k].astype(t)
    A["sbpkf.occupied"] = occ
    A["sbpkf.params"] = np.array([15.0, 1.0], np.float32)
    # SearchByBoW / SearchForTriangulation
    k1, k2, F12, ep, sf = synth.keyframe_pair(700, 800, 60, seed=321)
    kb = dict(k1); kb["has_mp"] = (rng.random(700) < 0.8).astype(np.uint8); kb["uright"] = None
    fb = dict(k2); fb["uright"] = None
    _kf(A, "bow.KF.", kb); _kf(A, "bow.F.", fb)
    A["bow.params"] = np.array([0.7, 1.0], np.float32)
    _kf(A, "tri.K1.", k1); _kf(A, "tri.K2.", k2)
    A["tri.F12"], A["tri.ep"], A["tri.scale2"] = F12.astype(np.float32).reshape(-1), ep.astype(np.float32), sf.astype(np.float32)
    A["tri.params"] = np.array([0, 0, 1], np.int32)
    A["tri.geo"] = np.array([1.0, 0.002, 0.0, -0.3, ep[0], ep[1]], np.float32)      # t12, cy of camera 2, epipole
    # SearchForInitialization
    I1, I2 = synth.initialization_pair(700, 750, seed=331)
    _frame(A, "init.F1.", I1); _frame(A, "init.F2.", I2)
    A["init.params"] = np.array([100, 0.9, 1.0], np.float32)
    # Fuse
    KF = synth.frame_view(800, seed=341)
    fmp, fur, fis = synth.fuse_problem(KF, 1000, seed=342)
    bf = np.float32(47.9)
    fmp["valid"] = (fmp["valid"].astype(bool) & (fmp["u"] >= 0) & (fmp["v"] >= 0)).astype(np.uint8)
    fmp["ur"] = (fmp["u"] - np.float32(bf * np.float32(1.0))).astype(np.float32)       # mock geometry: every point at z = 1
    _frame(A, "fuse.KF.", KF); _fuse_pts(A, "fuse.MP.", fmp)
    A["fuse.uright"], A["fuse.inv_sigma2"] = fur.astype(np.float32), fis.astype(np.float32)
    A["fuse.params"] = np.array([3.0, bf], np.float32)
    kf_obs = np.where(rng.random(800) < 0.5, rng.integers(1, 6, 800), 0).astype(np.uint8)
    mp_obs = rng.integers(1, 6, 1000).astype(np.uint8)
    A["fuse.kf_obs"], A["fuse.mp_obs"] = kf_obs, mp_obs
    # SearchBySim3, distinctive descriptor, grid
    A["fuse2.params"] = np.array([3.0], np.float32)
    S1, S2, s12, s21 = synth.sim3_problem(600, 650, seed=351)
    for sp in (s12, s21):
        sp["valid"] = (sp["valid"].astype(bool) & (sp["u"] >= 0) & (sp["v"] >= 0)).astype(np.uint8)
    _frame(A, "sim3.KF1.", S1); _frame(A, "sim3.KF2.", S2); _fuse_pts(A, "sim3.MP1in2.", s12); _fuse_pts(A, "sim3.MP2in1.", s21)
    A["sim3.params"] = np.array([7.5], np.float32)
    dd = rng.integers(0, 256, (23, 32), dtype=np.uint8); dd[7] = dd[3]
    A["dd.desc"] = dd
    _frame(A, "grid.F.", F)

    inp, outp = str(tmp_path / "in.blob"), str(tmp_path / "out.blob")
    write_blob(inp, A)
    r = subprocess.run([exe, inp, outp], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and r.stdout.startswith("OK"), r.stdout + r.stderr
    R = read_blob(outp)
    assert _i32(R["ran"])[0] == 11

    ref = oracle.hamming_pairs(a, b)
    assert np.array_equal(_i32(R["hd.one"]), ref) and np.array_equal(_i32(R["hd.bulk"]), ref) and _i32(R["hd.ref0"])[0] == ref[0]
    on, om = oracle.search_by_projection(Fs, MPs, 3.0, True, 25.0, 0.8, claimed)
    for w in ("flat", "ref"):
        assert _i32(R[f"sbp.{w}.n"])[0] == on and np.array_equal(_i32(R[f"sbp.{w}.match"]), om), w
    assert on > 100
    on, om = oracle.search_by_projection_kf(F2, KP, 15.0, True, occ)
    for w in ("flat", "ref"):
        assert _i32(R[f"sbpkf.{w}.n"])[0] == on and np.array_equal(_i32(R[f"sbpkf.{w}.match"]), om), w
    assert on > 50
    on, om = oracle.search_by_bow(kb, fb, 0.7, True)
    for w in ("flat", "ref"):
        assert _i32(R[f"bow.{w}.n"])[0] == on and np.array_equal(_i32(R[f"bow.{w}.match"]), om), w
    on, op = oracle.search_for_triangulation(k1, k2, F12, ep, sf, False, False, True)
    assert _i32(R["tri.flat.n"])[0] == on and np.array_equal(_i32(R["tri.flat.pairs"]).reshape(-1, 2), op) and on > 20
    cF12 = np.frombuffer(R["tri.ref.F12"], np.float32).reshape(3, 3); cep = np.frombuffer(R["tri.ref.ep"], np.float32)
    assert np.allclose(cF12, F12, atol=1e-6) and np.allclose(cep, ep, atol=1e-3)     # the mock geometry reproduces the synthetic F12 / epipole
    on2, op2 = oracle.search_for_triangulation(k1, k2, cF12, cep, sf, False, False, True)
    assert _i32(R["tri.ref.n"])[0] == on2 and np.array_equal(_i32(R["tri.ref.pairs"]).reshape(-1, 2), op2)
    on, om = oracle.search_for_initialization(I1, I2, 100, 0.9, True)
    for w in ("flat", "ref"):
        assert _i32(R[f"init.{w}.n"])[0] == on and np.array_equal(_i32(R[f"init.{w}.match"]), om), w
    bi, _ = oracle.fuse_search(KF, fmp, 3.0, 0, fur, fis)
    assert np.array_equal(_i32(R["fuse.flat.best"]), bi) and (bi >= 0).sum() > 100
    # replay of the graph updates of reference src/ORBmatcher.cc:826-848 on the oracle's search result
    ev = np.zeros(1000, np.int32); at = np.full(1000, -1, np.int32)
    pt = {k: ["kf", int(kf_obs[k]), False] for k in range(800) if kf_obs[k]}
    nf = 0
    for i in range(1000):
        k = int(bi[i])
        if k < 0:
            continue
        nf += 1
        cur = pt.get(k)
        if cur is None:
            ev[i], at[i] = 1, k
            pt[k] = ["mp", int(mp_obs[i]) + 1, False]
        elif not cur[2]:
            if cur[1] > int(mp_obs[i]):
                ev[i] = 2
            else:
                cur[2] = True
                if cur[0] == "kf":
                    ev[i], at[i] = 3, k
    assert _i32(R["fuse.ref.n"])[0] == nf
    assert np.array_equal(_i32(R["fuse.ref.event"]), ev) and np.array_equal(_i32(R["fuse.ref.at"]), at)
    assert {1, 2, 3} <= set(ev.tolist())
    on, om = oracle.search_by_sim3(S1, S2, s12, s21, 7.5)
    for w in ("flat", "ref"):
        assert _i32(R[f"sim3.{w}.n"])[0] == on and np.array_equal(_i32(R[f"sim3.{w}.match"]), om), w
    assert on > 30
    # Fuse(pKF, Scw, ...): search (variant 1) + the reference's bookkeeping (:958-968) replayed on the oracle's result
    fmp1 = dict(fmp); fmp1.pop("ur")
    b2, _ = oracle.fuse_search(KF, fmp1, 3.0, 1)
    assert np.array_equal(_i32(R["fuse2.flat.best"]), b2)
    rid = np.full(1000, -1, np.int32); at2 = np.full(1000, -1, np.int32)
    held = {k: 1000000 + k for k in range(800) if kf_obs[k]}
    nf2 = 0
    for i in range(1000):
        k = int(b2[i])
        if k < 0:
            continue
        nf2 += 1
        if k in held:
            rid[i] = held[k]
        else:
            at2[i] = k; held[k] = i
    assert _i32(R["fuse2.ref.n"])[0] == nf2 and nf2 > 100
    assert np.array_equal(_i32(R["fuse2.ref.replace"]), rid) and np.array_equal(_i32(R["fuse2.ref.at"]), at2)
    assert _i32(R["dd.best"])[0] == oracle.distinctive_descriptor(dd)[0]
    off, idx = oracle.assign_features_to_grid(F)
    assert np.array_equal(_i32(R["grid.off"]), off) and np.array_equal(_i32(R["grid.idx"]), idx)


def test_opencv_overload_two_threads_and_stereo(pkg, oracle, synth, tmp_path):
    """ORBextractor's cv::Mat / cv::KeyPoint operator() (compiled against tests/cpp/opencv_stub), called for the
    left and right image from two std::threads as upstream's stereo Frame constructor does, then
    ORB_SLAM3::ComputeStereoMatches and mvImagePyramid[l] as a cv::Mat: everything equals the oracle."""
    so = pkg.orbx.build()
    exe = str(tmp_path / "test_shim_opencv")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-DORBX_WITH_OPENCV", "-I" + os.path.join(ROOT, "tests", "cpp", "opencv_stub"),
                           "-pthread", "-o", exe, os.path.join(ROOT, "tests", "cpp", "test_shim_opencv.cpp"), "-L" + os.path.dirname(so),
                           "-lorbx", "-Wl,-rpath," + os.path.dirname(so)])
    w, h, nf, bf, b = 752, 480, 1000, 435.2 * 0.11, 0.11
    left, right = synth.stereo_pair(w, h, 0xB200)
    (tmp_path / "l.raw").write_bytes(left.tobytes()); (tmp_path / "r.raw").write_bytes(right.tobytes())
    outp = str(tmp_path / "o.blob")
    r = subprocess.run([exe, str(tmp_path / "l.raw"), str(tmp_path / "r.raw"), str(w), str(h), str(nf), repr(bf), repr(b), outp],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and r.stdout.startswith("OK"), r.stdout + r.stderr
    R = read_blob(outp)
    dt = np.dtype([("x", "f4"), ("y", "f4"), ("size", "f4"), ("angle", "f4"), ("response", "f4"), ("octave", "i4"), ("class_id", "i4")])
    exL, exR = oracle.Extractor(nf, 1.2, 8, 20, 7), oracle.Extractor(nf, 1.2, 8, 20, 7)
    sides = {}
    for tag, ex, img in (("L", exL, left), ("R", exR, right)):
        ok, od, om = ex(img, (0, 0))
        gk = np.frombuffer(R["k" + tag], dt); gd = np.frombuffer(R["d" + tag], np.uint8).reshape(-1, 32)
        assert len(gk) == len(ok) and _i32(R["mono" + tag])[0] == om
        for fld in ("x", "y", "size", "response", "octave", "class_id"):
            assert np.array_equal(gk[fld], ok[fld]), (tag, fld)
        assert np.abs(gk["angle"] - ok["angle"]).max() <= np.degrees(1e-4)
        bits = np.unpackbits(gd ^ od, axis=1).sum(1)
        assert (bits > 0).sum() <= 1 and bits.max() <= 2
        sides[tag] = (ok, od)
    # stereo on the oracle's own keypoints (equal to the GPU's up to the descriptor allowance above)
    lvL = [exL.level(l) for l in range(8)]; lvR = [exR.level(l) for l in range(8)]
    gdL = np.frombuffer(R["dL"], np.uint8).reshape(-1, 32); gdR = np.frombuffer(R["dR"], np.uint8).reshape(-1, 32)
    ur, dep = oracle.stereo_matches(sides["L"][0], gdL, lvL, sides["R"][0], gdR, lvR, exL.scale, exL.inv_scale, bf, b)
    gur, gdep = np.frombuffer(R["uR"], np.float32), np.frombuffer(R["depth"], np.float32)
    assert np.array_equal(gur, ur) and np.array_equal(gdep, dep) and (ur >= 0).sum() > 100
    rows, cols, step, psum, border_px, nlv = _i32(R["L1"])
    l1 = exL.level(1)
    assert (rows, cols, step, nlv) == (l1.shape[0] - 38, l1.shape[1] - 38, l1.shape[1], 8)
    assert np.array_equal(np.frombuffer(R["L1roi"], np.uint8).reshape(rows, cols), l1[19:-19, 19:-19])
    assert psum == int(l1[19 + 20:19 + 31, 19 + 30:19 + 41].astype(np.int64).sum()) and border_px == int(l1[0, 0])


def test_image_grabber_cpp(pkg, oracle, synth, tmp_path):
    """ros2_orbslam3::ImageGrabber (include/orbx/ImageGrabber.h, cv::Mat form against the OpenCV stub):
    ConvertImageToGPU equals the oracle's remap -> resize -> gray bit for bit, and the chained ConvertAndExtract equals
    extracting that gray image through the C ABI."""
    so = pkg.orbx.build()
    exe = str(tmp_path / "test_shim_opencv")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-DORBX_WITH_OPENCV", "-I" + os.path.join(ROOT, "tests", "cpp", "opencv_stub"),
                           "-pthread", "-o", exe, os.path.join(ROOT, "tests", "cpp", "test_shim_opencv.cpp"), "-L" + os.path.dirname(so),
                           "-lorbx", "-Wl,-rpath," + os.path.dirname(so)])
    w, h, nf, sw, sh = 640, 480, 1000, 1024, 768
    left, right = synth.stereo_pair(w, h, 5)
    bgr = synth.bgr_image(sw, sh, 91)
    m1, m2 = synth.undistortion_maps(sw, sh, sw, sh, zoom=0.9)
    for name, arr in (("l.raw", left), ("r.raw", right), ("bgr.raw", bgr), ("m1.raw", m1), ("m2.raw", m2)):
        (tmp_path / name).write_bytes(np.ascontiguousarray(arr).tobytes())
    o1, o2 = str(tmp_path / "o1.blob"), str(tmp_path / "o2.blob")
    r = subprocess.run([exe, str(tmp_path / "l.raw"), str(tmp_path / "r.raw"), str(w), str(h), str(nf), "47.9", "0.11", o1,
                        str(tmp_path / "bgr.raw"), str(tmp_path / "m1.raw"), str(tmp_path / "m2.raw"), str(sw), str(sh), o2],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and r.stdout.startswith("OK"), r.stdout + r.stderr
    R = read_blob(o2)
    grey = oracle.preprocess_bgr(bgr, m1, m2, w, h)
    assert np.array_equal(np.frombuffer(R["grey"], np.uint8).reshape(h, w), grey)
    ex = pkg.orbx.ORBextractor(nf, 1.2, 8, 20, 7)
    k, d, mono = ex(grey, (0, 0))
    ex.close()
    dt = np.dtype([("x", "f4"), ("y", "f4"), ("size", "f4"), ("angle", "f4"), ("response", "f4"), ("octave", "i4"), ("class_id", "i4")])
    gk = np.frombuffer(R["k"], dt)
    assert len(gk) == len(k) > 500 and _i32(R["mono"])[0] == mono
    assert gk.tobytes() == k.tobytes() and np.array_equal(np.frombuffer(R["d"], np.uint8).reshape(-1, 32), d)
