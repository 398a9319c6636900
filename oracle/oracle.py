"""ctypes binding of the CPU oracle (oracle/_build/liborb_oracle.so).

TEST INFRASTRUCTURE ONLY: importable from tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  The product path never imports this."""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liborb_oracle.so")

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])


def build(force=False):
    src = [os.path.join(_HERE, f) for f in ("orb_oracle.cpp", "orb_oracle.h")]
    if force or not os.path.exists(_SO) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in src):
        subprocess.check_call(["make", "-s", "-C", _HERE])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.orc_fast_atan2.restype = C.c_float
        _lib.orc_fast_atan2.argtypes = [C.c_float, C.c_float]
        _lib.orc_extractor_create.restype = C.c_void_p
        _lib.orc_extractor_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        _lib.orc_ic_angle.restype = C.c_float
        _lib.orc_bench_extract.restype = C.c_double
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _u8(a):
    return np.ascontiguousarray(a, dtype=np.uint8)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def resize_linear(src, dw, dh):
    src = _u8(src)
    dst = np.empty((dh, dw), np.uint8)
    lib().orc_resize_linear_u8(_p(src), src.shape[1], src.shape[0], src.strides[0], _p(dst), dw, dh, dw)
    return dst


def border_reflect101(img, b):
    img = _u8(img)
    h, w = img.shape
    buf = np.zeros((h + 2 * b, w + 2 * b), np.uint8)
    buf[b:b + h, b:b + w] = img
    lib().orc_border_reflect101(_p(buf), w, h, w + 2 * b, b)
    return buf


def gaussian7(img):
    img = _u8(img)
    dst = np.empty_like(img)
    lib().orc_gaussian7_s2_u8(_p(img), img.shape[1], img.shape[0], img.strides[0], _p(dst), img.shape[1])
    return dst


def remap_cubic_bgr(src, mapx, mapy):
    src = _u8(src); mx, my = _f32(mapx), _f32(mapy)
    out = np.empty(mx.shape + (3,), np.uint8)
    lib().orc_remap_cubic_bgr(_p(src), src.shape[1], src.shape[0], _p(mx), _p(my), mx.shape[1], mx.shape[0], _p(out))
    return out


def preprocess_bgr(src, mapx, mapy, dw, dh):
    """remap(INTER_CUBIC, BORDER_CONSTANT) -> resize(INTER_LINEAR) -> BGR2GRAY (image_grabber.hpp:96-110)."""
    src = _u8(src); mx, my = _f32(mapx), _f32(mapy)
    out = np.empty((dh, dw), np.uint8)
    lib().orc_preprocess_bgr(_p(src), src.shape[1], src.shape[0], _p(mx), _p(my), mx.shape[1], mx.shape[0], int(dw), int(dh), _p(out))
    return out


def fast9_16(img, th, max_out=1 << 16):
    img = _u8(img)
    xs = np.empty(max_out, np.int32); ys = np.empty(max_out, np.int32); sc = np.empty(max_out, np.int32)
    n = lib().orc_fast9_16(_p(img), img.shape[1], img.shape[0], img.strides[0], int(th), _p(xs), _p(ys), _p(sc), max_out)
    assert n <= max_out
    return xs[:n].copy(), ys[:n].copy(), sc[:n].copy()


def fast_atan2(y, x):
    return float(lib().orc_fast_atan2(float(y), float(x)))


def distribute_octtree(xs, ys, resp, minX, maxX, minY, maxY, N):
    xs, ys, resp = _i32(xs), _i32(ys), _i32(resp)
    out = np.empty(max(len(xs), 1), np.int32)
    n = lib().orc_distribute_octtree(len(xs), _p(xs), _p(ys), _p(resp), minX, maxX, minY, maxY, N, _p(out), len(out))
    return out[:n].copy()


def ic_angle(bordered, border, x, y):
    """bordered: (h+2b, w+2b) u8; (x, y) ROI coordinates."""
    bordered = _u8(bordered)
    roi = bordered.ctypes.data + border * bordered.strides[0] + border
    return float(lib().orc_ic_angle(C.c_void_p(roi), bordered.strides[0], int(x), int(y)))


def orb_descriptor(blurred, x, y, angle_deg):
    blurred = _u8(blurred)
    d = np.empty(32, np.uint8)
    lib().orc_orb_descriptor(_p(blurred), blurred.strides[0], int(x), int(y), C.c_float(angle_deg), _p(d))
    return d


class Extractor:
    """Mirror of upstream ORBextractor (5-arg ctor, operator())."""

    def __init__(self, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7):
        self.nfeatures, self.nlevels = nfeatures, nlevels
        self._h = C.c_void_p(lib().orc_extractor_create(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST))
        t = [np.empty(nlevels, np.float32) for _ in range(4)]
        self.feat_per_level = np.empty(nlevels, np.int32)
        self.umax = np.empty(16, np.int32)
        lib().orc_extractor_tables(self._h, _p(t[0]), _p(t[1]), _p(t[2]), _p(t[3]), _p(self.feat_per_level), _p(self.umax))
        self.scale, self.inv_scale, self.sigma2, self.inv_sigma2 = t

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_extractor_destroy(self._h)
            self._h = None

    def __call__(self, image, lapping=(0, 0)):
        image = _u8(image)
        h, w = image.shape
        cap = self.nfeatures * 2 + 256
        kps = np.zeros(cap, KP_DTYPE); desc = np.zeros((cap, 32), np.uint8)
        lap = np.asarray(lapping, np.int32)
        mono = C.c_int(0)
        n = lib().orc_extract(self._h, _p(image), w, h, image.strides[0], _p(lap), _p(kps), _p(desc), cap, C.byref(mono))
        assert n <= cap
        if n < 0:
            return None, None, -1
        return kps[:n].copy(), desc[:n].copy(), mono.value

    def level(self, l):
        w, h = C.c_int(), C.c_int()
        lib().orc_level_size(self._h, l, C.byref(w), C.byref(h))
        buf = np.empty((h.value + 38, w.value + 38), np.uint8)
        lib().orc_level_bordered(self._h, l, _p(buf))
        return buf

    def candidates(self, l, cap=1 << 16):
        xs = np.empty(cap, np.int32); ys = np.empty(cap, np.int32); r = np.empty(cap, np.int32)
        n = lib().orc_level_candidates(self._h, l, _p(xs), _p(ys), _p(r), cap)
        return xs[:n].copy(), ys[:n].copy(), r[:n].copy()

    def level_keypoints(self, l, cap=1 << 14):
        k = np.zeros(cap, KP_DTYPE)
        n = lib().orc_level_keypoints(self._h, l, _p(k), cap)
        return k[:n].copy()


def descriptor_distance(a, b):
    return int(lib().orc_descriptor_distance(_p(_u8(a)), _p(_u8(b))))


def hamming_pairs(a, b):
    a, b = _u8(a), _u8(b)
    out = np.empty(len(a), np.int32)
    lib().orc_hamming_pairs(_p(a), _p(b), len(a), _p(out))
    return out


def three_maxima(counts):
    counts = _i32(counts)
    i = [C.c_int(), C.c_int(), C.c_int()]
    lib().orc_three_maxima(_p(counts), len(counts), C.byref(i[0]), C.byref(i[1]), C.byref(i[2]))
    return tuple(v.value for v in i)


class _FrameView(C.Structure):
    _fields_ = [("n", C.c_int), ("kx", C.c_void_p), ("ky", C.c_void_p), ("koct", C.c_void_p),
                ("desc", C.c_void_p), ("uright", C.c_void_p), ("minx", C.c_float), ("miny", C.c_float),
                ("gw_inv", C.c_float), ("gh_inv", C.c_float), ("gcols", C.c_int), ("grows", C.c_int),
                ("scale", C.c_void_p), ("nlevels", C.c_int)]


def _frame_view(F, keep):
    kx, ky, ko, d, sc = _f32(F["kx"]), _f32(F["ky"]), _i32(F["koct"]), _u8(F["desc"]), _f32(F["scale"])
    ur = _f32(F["uright"]) if F.get("uright") is not None else None
    keep += [kx, ky, ko, d, sc, ur]
    return _FrameView(len(kx), kx.ctypes.data, ky.ctypes.data, ko.ctypes.data, d.ctypes.data,
                      ur.ctypes.data if ur is not None else None,
                      F["minx"], F["miny"], F["gw_inv"], F["gh_inv"], F["gcols"], F["grows"],
                      sc.ctypes.data, len(sc))


def distinctive_descriptor(desc):
    d = _u8(desc)
    med = C.c_int()
    return lib().orc_distinctive_descriptor(_p(d), len(d), C.byref(med)), med.value


def forb_mean_value(desc):
    d = _u8(np.asarray(desc, np.uint8).reshape(-1, 32))
    out = np.zeros(32, np.uint8)
    lib().orc_forb_mean_value(_p(d), len(d), _p(out))
    return out


def assign_features_to_grid(F):
    keep = []
    fv = _frame_view(F, keep)
    off = np.empty(F["gcols"] * F["grows"] + 1, np.int32)
    idx = np.empty(max(fv.n, 1), np.int32)
    lib().orc_assign_features_to_grid(C.byref(fv), _p(off), _p(idx))
    return off, idx[:fv.n]


def features_in_area(F, x, y, r, minLevel, maxLevel, cap=256):
    keep = []
    fv = _frame_view(F, keep)
    x, y, r = _f32(x), _f32(y), _f32(r)
    a, b = _i32(minLevel), _i32(maxLevel)
    cnt = np.empty(len(x), np.int32)
    idx = np.full((len(x), cap), -1, np.int32)
    lib().orc_features_in_area(C.byref(fv), len(x), _p(x), _p(y), _p(r), _p(a), _p(b), _p(cnt), _p(idx), cap)
    return cnt, idx


def search_by_projection(F, mp, th, bFarPoints=False, thFarPoints=0.0, nnRatio=0.8, claimed=None):
    keep = []
    fv = _frame_view(F, keep)
    inview, x, y = _u8(mp["inview"]), _f32(mp["x"]), _f32(mp["y"])
    lvl, vc, dep, d = _i32(mp["level"]), _f32(mp["viewcos"]), _f32(mp["depth"]), _u8(mp["desc"])
    cl = _u8(claimed) if claimed is not None else None
    xr = _f32(mp["proj_xr"]) if mp.get("proj_xr") is not None else None
    ho = _u8(mp["has_obs"]) if mp.get("has_obs") is not None else None
    assert fv.uright is None or xr is not None, "a stereo frame view needs the map points' proj_xr"
    out = np.empty(fv.n, np.int32)
    n = lib().orc_search_by_projection(C.byref(fv), len(x), _p(inview), _p(x), _p(y), _p(lvl), _p(vc), _p(dep),
                                       _p(d), _p(xr), _p(ho), _p(cl), C.c_float(th), int(bFarPoints), C.c_float(thFarPoints),
                                       C.c_float(nnRatio), _p(out))
    return n, out


def search_by_projection_kf(F, mp, th, checkOrientation=True, occupied=None):
    keep = []
    fv = _frame_view(F, keep)
    fa = _f32(F["kangle"])
    valid, u, v = _u8(mp["valid"]), _f32(mp["u"]), _f32(mp["v"])
    lvl, ang, d = _i32(mp["level"]), _f32(mp["angle"]), _u8(mp["desc"])
    oc = _u8(occupied) if occupied is not None else None
    out = np.empty(fv.n, np.int32)
    n = lib().orc_search_by_projection_kf(C.byref(fv), _p(fa), len(u), _p(valid), _p(u), _p(v), _p(lvl), _p(ang),
                                          _p(d), _p(oc), C.c_float(th), int(checkOrientation), _p(out))
    return n, out


class _KFView(C.Structure):
    _fields_ = [("n", C.c_int), ("kx", C.c_void_p), ("ky", C.c_void_p), ("koct", C.c_void_p),
                ("kangle", C.c_void_p), ("desc", C.c_void_p), ("has_mp", C.c_void_p), ("uright", C.c_void_p),
                ("nnodes", C.c_int), ("node_id", C.c_void_p), ("node_off", C.c_void_p), ("node_idx", C.c_void_p)]


def _kf_view(K, keep):
    a = [_f32(K["kx"]), _f32(K["ky"]), _i32(K["koct"]), _f32(K["kangle"]), _u8(K["desc"]), _u8(K["has_mp"]),
         _i32(K["node_id"]), _i32(K["node_off"]), _i32(K["node_idx"])]
    ur = _f32(K["uright"]) if K.get("uright") is not None else None
    keep += a + [ur]
    return _KFView(len(a[0]), a[0].ctypes.data, a[1].ctypes.data, a[2].ctypes.data, a[3].ctypes.data,
                   a[4].ctypes.data, a[5].ctypes.data, ur.ctypes.data if ur is not None else None,
                   len(a[6]), a[6].ctypes.data, a[7].ctypes.data, a[8].ctypes.data)


def fuse_search(KF, mp, th=3.0, variant=0, uright=None, inv_sigma2=None):
    keep = []
    fv = _frame_view(KF, keep)
    valid, u, v, lvl, d = _u8(mp["valid"]), _f32(mp["u"]), _f32(mp["v"]), _i32(mp["level"]), _u8(mp["desc"])
    ur = _f32(mp["ur"]) if mp.get("ur") is not None else None
    kur = _f32(uright) if uright is not None else None
    isg = _f32(inv_sigma2) if inv_sigma2 is not None else None
    bi = np.empty(max(len(u), 1), np.int32); bd = np.empty(max(len(u), 1), np.int32)
    lib().orc_fuse_search(C.byref(fv), _p(kur), _p(isg), len(u), _p(valid), _p(u), _p(v), _p(ur), _p(lvl), _p(d),
                          C.c_float(th), int(variant), _p(bi), _p(bd))
    return bi[:len(u)], bd[:len(u)]


def search_by_sim3(KF1, KF2, mp1in2, mp2in1, th=7.5):
    keep = []
    a, b = _frame_view(KF1, keep), _frame_view(KF2, keep)
    args = []
    for mp in (mp1in2, mp2in1):
        arrs = [_u8(mp["valid"]), _f32(mp["u"]), _f32(mp["v"]), _i32(mp["level"]), _u8(mp["desc"])]
        keep += arrs
        args += [_p(x) for x in arrs]
    out = np.empty(max(a.n, 1), np.int32)
    n = lib().orc_search_by_sim3(C.byref(a), C.byref(b), *args, C.c_float(th), _p(out))
    return n, out[:a.n]


def search_for_initialization(F1, F2, windowSize=100, nnRatio=0.9, checkOrientation=True):
    keep = []
    a, b = _frame_view(F1, keep), _frame_view(F2, keep)
    a1, a2 = _f32(F1["kangle"]), _f32(F2["kangle"])
    out = np.empty(max(a.n, 1), np.int32)
    n = lib().orc_search_for_initialization(C.byref(a), _p(a1), C.byref(b), _p(a2), int(windowSize), C.c_float(nnRatio),
                                            int(checkOrientation), _p(out))
    return n, out[:a.n]


def search_by_bow(KF, F, nnRatio=0.7, checkOrientation=True, nleft=-1):
    keep = []
    a, b = _kf_view(KF, keep), _kf_view(F, keep)
    out = np.empty(max(b.n, 1), np.int32)
    n = lib().orc_search_by_bow(C.byref(a), C.byref(b), int(nleft), C.c_float(nnRatio), int(checkOrientation), _p(out))
    return n, out[:b.n]


def search_for_triangulation(K1, K2, F12, ep, scale2, bOnlyStereo=False, bCoarse=False, checkOrientation=True):
    keep = []
    v1, v2 = _kf_view(K1, keep), _kf_view(K2, keep)
    F12 = _f32(F12).reshape(9); sc = _f32(scale2)
    pairs = np.empty((max(v1.n, 1), 2), np.int32)
    npairs = C.c_int(0)
    n = lib().orc_search_for_triangulation(C.byref(v1), C.byref(v2), _p(F12), C.c_float(ep[0]), C.c_float(ep[1]),
                                           _p(sc), int(bOnlyStereo), int(bCoarse), int(checkOrientation),
                                           _p(pairs), C.byref(npairs))
    return n, pairs[:npairs.value].copy()


class _StereoSide(C.Structure):
    _fields_ = [("n", C.c_int), ("kx", C.c_void_p), ("ky", C.c_void_p), ("koct", C.c_void_p), ("desc", C.c_void_p),
                ("nlevels", C.c_int), ("lvl", C.c_void_p), ("lvl_stride", C.c_void_p), ("lvl_w", C.c_void_p),
                ("lvl_h", C.c_void_p)]


def _stereo_side(kps, desc, levels, border, keep):
    """levels: list of bordered (h+2b, w+2b) u8 arrays."""
    kx, ky, ko, d = _f32(kps["x"]), _f32(kps["y"]), _i32(kps["octave"]), _u8(desc)
    lv = [_u8(l) for l in levels]
    ptrs = (C.c_void_p * len(lv))(*[l.ctypes.data + border * l.strides[0] + border for l in lv])
    st = _i32([l.strides[0] for l in lv]); w = _i32([l.shape[1] - 2 * border for l in lv])
    h = _i32([l.shape[0] - 2 * border for l in lv])
    keep += [kx, ky, ko, d, lv, ptrs, st, w, h]
    return _StereoSide(len(kx), kx.ctypes.data, ky.ctypes.data, ko.ctypes.data, d.ctypes.data, len(lv),
                       C.cast(ptrs, C.c_void_p), st.ctypes.data, w.ctypes.data, h.ctypes.data)


def stereo_matches(kpsL, descL, levelsL, kpsR, descR, levelsR, scale, inv_scale, bf, b, border=19):
    keep = []
    L = _stereo_side(kpsL, descL, levelsL, border, keep)
    R = _stereo_side(kpsR, descR, levelsR, border, keep)
    sc, isc = _f32(scale), _f32(inv_scale)
    ur = np.empty(L.n, np.float32); dep = np.empty(L.n, np.float32)
    lib().orc_stereo_matches(C.byref(L), C.byref(R), _p(sc), _p(isc), C.c_float(bf), C.c_float(b), _p(ur), _p(dep))
    return ur, dep


def knn2(q, db, nthreads=1):
    q, db = _u8(q), _u8(db)
    o = [np.empty(len(q), np.int32) for _ in range(4)]
    lib().orc_knn2(_p(q), len(q), _p(db), len(db), _p(o[0]), _p(o[1]), _p(o[2]), _p(o[3]), int(nthreads))
    return o  # d1, i1, d2, i2


def bench_extract(images, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniTh=20, minTh=7, iters_per_thread=4, nthreads=1):
    images = _u8(images)
    n, h, w = images.shape
    tot = C.c_long(0)
    lib().orc_bench_extract.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int,
                                        C.c_int, C.c_int, C.c_int, C.c_void_p]
    s = lib().orc_bench_extract(nfeatures, scaleFactor, nlevels, iniTh, minTh, _p(images), n, w, h,
                                iters_per_thread, nthreads, C.byref(tot))
    return s, tot.value
