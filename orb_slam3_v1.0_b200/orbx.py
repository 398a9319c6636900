"""ctypes binding of liborbx.so (include/orbx.h) -- the host-side mirror used by tests/ and
bench.py.  Mirrors the reference interface for the path: `ORBextractor(nfeatures, scaleFactor,
nlevels, iniThFAST, minThFAST)` + `__call__(image, lapping)` (operator()), `mvImagePyramid`,
`mvScaleFactor`, `mvInvScaleFactor` (reference include/ORBextractor.h:52-129, upstream form).

No CPU fallback: a missing library or device raises OrbxError."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "lib", "liborbx.so")

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])


class OrbxError(RuntimeError):
    pass


class _Params(C.Structure):
    _fields_ = [("nfeatures", C.c_int), ("scale_factor", C.c_float), ("nlevels", C.c_int),
                ("ini_th_fast", C.c_int), ("min_th_fast", C.c_int), ("width", C.c_int),
                ("height", C.c_int), ("max_batch", C.c_int), ("device", C.c_int)]


def build(force=False):
    """Compile liborbx.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
    csrc = os.path.join(_HERE, "csrc")
    if force:
        subprocess.check_call(["make", "-s", "-C", csrc, "clean"])
    subprocess.check_call(["make", "-s", "-C", csrc])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            raise OrbxError(f"{_SO} is missing: run __graft_entry__.build() (there is no CPU fallback)")
        L = C.CDLL(_SO)
        L.orbx_last_error.restype = C.c_char_p
        L.orbx_extractor_stream.restype = C.c_void_p
        L.orbx_extractor_stream.argtypes = [C.c_void_p]
        L.orbx_extractor_create.argtypes = [C.POINTER(_Params), C.POINTER(C.c_void_p)]
        L.orbx_extractor_destroy.argtypes = [C.c_void_p]
        L.orbx_extract_device.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p]
        L.orbx_extract_host.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p]
        L.orbx_extractor_sync.argtypes = [C.c_void_p]
        L.orbx_extractor_fetch.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.orbx_extractor_counts.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.orbx_extractor_tables.argtypes = [C.c_void_p] + [C.c_void_p] * 6
        L.orbx_extractor_capacity.argtypes = [C.c_void_p]
        L.orbx_extractor_level_size.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.orbx_extractor_level_download.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.orbx_extractor_blurred_download.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.orbx_extractor_candidates.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.orbx_extractor_last_launches.argtypes = [C.c_void_p]
        L.orbx_extractor_set_profiling.argtypes = [C.c_void_p, C.c_int]
        L.orbx_extractor_stage_ms.argtypes = [C.c_void_p, C.c_void_p]
        L.orbx_host_alloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t]
        L.orbx_host_free.argtypes = [C.c_void_p]
        L.orbx_device_alloc.argtypes = [C.c_int, C.POINTER(C.c_void_p), C.c_size_t]
        L.orbx_device_free.argtypes = [C.c_void_p]
        L.orbx_memcpy_h2d.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.orbx_memcpy_d2h.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.orbx_hamming_pairs.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        _lib = L
    return _lib


def _ck(st):
    if st != 0:
        raise OrbxError(f"orbx status {st}: {lib().orbx_last_error().decode()}")


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class PinnedBuffer:
    """Page-locked host memory exposed as a numpy array."""

    def __init__(self, shape, dtype=np.uint8):
        self.nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        self._ptr = C.c_void_p()
        _ck(lib().orbx_host_alloc(C.byref(self._ptr), self.nbytes))
        buf = (C.c_uint8 * self.nbytes).from_address(self._ptr.value)
        self.array = np.frombuffer(buf, dtype=dtype).reshape(shape)

    def free(self):
        if self._ptr:
            self.array = None
            lib().orbx_host_free(self._ptr)
            self._ptr = None


class DeviceBuffer:
    def __init__(self, nbytes, device=0):
        self.nbytes = nbytes
        self.ptr = C.c_void_p()
        _ck(lib().orbx_device_alloc(device, C.byref(self.ptr), nbytes))

    def upload(self, arr):
        arr = np.ascontiguousarray(arr)
        assert arr.nbytes <= self.nbytes
        _ck(lib().orbx_memcpy_h2d(self.ptr, _p(arr), arr.nbytes))

    def free(self):
        if self.ptr:
            lib().orbx_device_free(self.ptr)
            self.ptr = None


class ORBextractor:
    """Batched drop-in for ORBextractor.  The C handle is built for one image size; it is
    (re)created lazily when the size or batch changes, as the 5-arg reference ctor has no size."""

    def __init__(self, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7,
                 device=0, max_batch=1):
        self.nfeatures, self.scaleFactor, self.nlevels = nfeatures, scaleFactor, nlevels
        self.iniThFAST, self.minThFAST = iniThFAST, minThFAST
        self.device, self.max_batch = device, max_batch
        self._h = None
        self._size = None

    # ---- lifecycle
    def _ensure(self, w, h, batch):
        if self._h is not None and self._size == (w, h) and batch <= self.max_batch:
            return
        self.close()
        self.max_batch = max(self.max_batch, batch)
        p = _Params(self.nfeatures, self.scaleFactor, self.nlevels, self.iniThFAST, self.minThFAST,
                    w, h, self.max_batch, self.device)
        hnd = C.c_void_p()
        _ck(lib().orbx_extractor_create(C.byref(p), C.byref(hnd)))
        self._h, self._size = hnd, (w, h)
        nl = self.nlevels
        t = [np.empty(nl, np.float32) for _ in range(4)]
        self.mnFeaturesPerLevel = np.empty(nl, np.int32)
        self.umax = np.empty(16, np.int32)
        _ck(lib().orbx_extractor_tables(hnd, _p(t[0]), _p(t[1]), _p(t[2]), _p(t[3]), _p(self.mnFeaturesPerLevel), _p(self.umax)))
        self.mvScaleFactor, self.mvInvScaleFactor, self.mvLevelSigma2, self.mvInvLevelSigma2 = t
        self.capacity = lib().orbx_extractor_capacity(hnd)

    def prepare(self, width, height, batch=1):
        self._ensure(width, height, batch)
        return self

    def close(self):
        if self._h is not None:
            lib().orbx_extractor_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- extraction
    def extract_host(self, images, lapping=None):
        """images: (B,H,W) u8 host array (pinned for async copies). Enqueues H2D + extraction."""
        b, h, w = images.shape
        self._ensure(w, h, b)
        lap = np.ascontiguousarray(lapping, np.int32).reshape(b, 2) if lapping is not None else None
        _ck(lib().orbx_extract_host(self._h, images.ctypes.data, images.strides[0], images.strides[1], b, _p(lap)))

    def extract_device(self, dptr, batch, width, height, lapping=None):
        self._ensure(width, height, batch)
        lap = np.ascontiguousarray(lapping, np.int32).reshape(batch, 2) if lapping is not None else None
        _ck(lib().orbx_extract_device(self._h, dptr, width * height, width, batch, _p(lap)))

    def sync(self):
        _ck(lib().orbx_extractor_sync(self._h))

    def counts(self, batch):
        n = np.empty(batch, np.int32); m = np.empty(batch, np.int32)
        _ck(lib().orbx_extractor_counts(self._h, _p(n), _p(m)))
        return n, m

    def fetch(self, f):
        kps = np.empty(self.capacity, KP_DTYPE); desc = np.empty((self.capacity, 32), np.uint8)
        n, mono = C.c_int(), C.c_int()
        _ck(lib().orbx_extractor_fetch(self._h, f, _p(kps), _p(desc), self.capacity, C.byref(n), C.byref(mono)))
        return kps[:n.value].copy(), desc[:n.value].copy(), mono.value

    def __call__(self, image, lapping=(0, 0)):
        """operator()(image, mask, keypoints, descriptors, vLappingArea) for one frame:
        returns (keypoints, descriptors, monoIndex); (None, None, -1) on an empty image."""
        if image is None or image.size == 0:
            return None, None, -1
        image = np.ascontiguousarray(image, np.uint8)
        self.extract_host(image[None], [lapping])
        self.sync()
        return self.fetch(0)

    def extract_batch(self, images, lapping=None):
        images = np.ascontiguousarray(images, np.uint8)
        self.extract_host(images, lapping)
        self.sync()
        return [self.fetch(f) for f in range(images.shape[0])]

    # ---- exposed state
    def level_size(self, level):
        w, h = C.c_int(), C.c_int()
        _ck(lib().orbx_extractor_level_size(self._h, level, C.byref(w), C.byref(h)))
        return w.value, h.value

    def image_pyramid(self, level, f=0):
        """mvImagePyramid[level] with its 19-px border: (h+38, w+38) u8."""
        w, h = self.level_size(level)
        out = np.empty((h + 38, w + 38), np.uint8)
        _ck(lib().orbx_extractor_level_download(self._h, f, level, _p(out)))
        return out

    def blurred(self, level, f=0):
        w, h = self.level_size(level)
        out = np.empty((h, w), np.uint8)
        _ck(lib().orbx_extractor_blurred_download(self._h, f, level, _p(out)))
        return out

    def candidates(self, level, f=0, cap=1 << 18):
        xs = np.empty(cap, np.int32); ys = np.empty(cap, np.int32); r = np.empty(cap, np.int32)
        n = C.c_int()
        _ck(lib().orbx_extractor_candidates(self._h, f, level, _p(xs), _p(ys), _p(r), cap, C.byref(n)))
        return xs[:n.value].copy(), ys[:n.value].copy(), r[:n.value].copy()

    def last_launches(self):
        return lib().orbx_extractor_last_launches(self._h)

    def set_profiling(self, on=True):
        _ck(lib().orbx_extractor_set_profiling(self._h, int(on)))

    def stage_ms(self):
        ms = np.empty(6, np.float32)
        _ck(lib().orbx_extractor_stage_ms(self._h, _p(ms)))
        return ms

    def stream(self):
        return lib().orbx_extractor_stream(self._h)


def hamming_pairs(a, b):
    """ORBmatcher::DescriptorDistance over n pairs (reference src/ORBmatcher.cc:1375-1391)."""
    a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
    out = np.empty(len(a), np.int32)
    _ck(lib().orbx_hamming_pairs(_p(a), _p(b), len(a), _p(out)))
    return out
