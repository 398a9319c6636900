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
# ORBX_LIB_DIR: load another build of the SAME library (tools/checked_run.sh: the -DORBX_CHECKS variant); default = the shipped one
_SO = os.path.join(os.environ.get("ORBX_LIB_DIR") or os.path.join(_HERE, "lib"), "liborbx.so")

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
    subprocess.check_call(["make", "-s", "-j4", "-C", csrc])
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
        L.orbx_extractor_streams.argtypes = [C.c_void_p, C.c_int]
        L.orbx_extractor_set_profiling.argtypes = [C.c_void_p, C.c_int]
        L.orbx_extractor_set_fast_corner_list.argtypes = [C.c_void_p, C.c_int]
        L.orbx_extractor_assign_grid.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int]
        L.orbx_extractor_grid_fetch.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
        L.orbx_fuse_search.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_void_p, C.c_void_p]
        L.orbx_search_by_sim3.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 4 + [C.c_float, C.c_void_p, C.c_void_p]
        L.orbx_search_for_initialization.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_int, C.c_void_p, C.c_void_p]
        L.orbx_search_by_bow.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_void_p, C.c_void_p]
        L.orbx_distinctive_descriptors.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 4
        L.orbx_mean_value.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 3
        L.orbx_assign_features_to_grid.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orbx_features_in_area.argtypes = [C.c_void_p, C.c_void_p, C.c_int] + [C.c_void_p] * 7 + [C.c_int]
        L.orbx_extractor_stage_ms.argtypes = [C.c_void_p, C.c_void_p]
        L.orbx_host_alloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t]
        L.orbx_host_alloc_wc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t]
        L.orbx_host_free.argtypes = [C.c_void_p]
        L.orbx_device_alloc.argtypes = [C.c_int, C.POINTER(C.c_void_p), C.c_size_t]
        L.orbx_device_free.argtypes = [C.c_void_p]
        L.orbx_memcpy_h2d.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.orbx_memcpy_d2h.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.orbx_hamming_pairs.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.orbx_three_maxima.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orbx_matcher_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
        L.orbx_matcher_destroy.argtypes = [C.c_void_p]
        L.orbx_matcher_stream.restype = C.c_void_p
        L.orbx_matcher_stream.argtypes = [C.c_void_p]
        L.orbx_matcher_last_launches.argtypes = [C.c_void_p]
        L.orbx_matcher_set_knn_variant.argtypes = [C.c_void_p, C.c_int]
        L.orbx_matcher_synchronize.argtypes = [C.c_void_p]
        L.orbx_search_by_projection.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float,
                                                C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p]
        L.orbx_search_by_projection_resident.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                                         C.c_void_p, C.c_float, C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p]
        L.orbx_search_by_projection_kf_resident.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                                            C.c_void_p, C.c_float, C.c_int, C.c_void_p, C.c_void_p]
        L.orbx_search_by_projection_kf.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float,
                                                   C.c_int, C.c_void_p, C.c_void_p]
        L.orbx_search_for_triangulation.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                                    C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                                    C.c_void_p]
        L.orbx_stereo_matches.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_float, C.c_float,
                                          C.c_void_p, C.c_void_p, C.c_int]
        L.orbx_stereo_matches_async.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_float,
                                                C.c_float]
        L.orbx_stereo_fetch.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
        L.orbx_knn2.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_float] + [C.c_void_p] * 5
        L.orbx_knn2_device.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_longlong, C.c_void_p]
        L.orbx_knn2_device_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_longlong, C.c_void_p]
        L.orbx_knn2_noalloc_scratch_bytes.restype = C.c_size_t
        L.orbx_knn2_merge_device.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float] + [C.c_void_p] * 5
        L.orbx_extractor_level_device.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 5
        L.orbx_extractor_results_device.argtypes = [C.c_void_p] + [C.c_void_p] * 4
        L.orbx_preproc_create.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                          C.POINTER(C.c_void_p)]
        L.orbx_preproc_destroy.argtypes = [C.c_void_p]
        L.orbx_preproc_stream.restype = C.c_void_p
        L.orbx_preproc_stream.argtypes = [C.c_void_p]
        L.orbx_preproc_device.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.orbx_preproc_host.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p]
        L.orbx_preproc_extract_host.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p]
        _lib = L
    return _lib


def _ck(st):
    if st != 0:
        raise OrbxError(f"orbx status {st}: {lib().orbx_last_error().decode()}")


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class PinnedBuffer:
    """Page-locked host memory exposed as a numpy array."""

    def __init__(self, shape, dtype=np.uint8, write_combined=False):
        self.nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        self._ptr = C.c_void_p()
        _ck((lib().orbx_host_alloc_wc if write_combined else lib().orbx_host_alloc)(C.byref(self._ptr), self.nbytes))
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

    def level_device(self, level):
        """(device pointer of ROI pixel (0,0) of frame 0, w, h, pitch, frame_stride) of mvImagePyramid[level] in place."""
        p = C.c_void_p(); w, h, pitch = C.c_int(), C.c_int(), C.c_int(); fs = C.c_size_t()
        _ck(lib().orbx_extractor_level_device(self._h, level, C.byref(p), C.byref(w), C.byref(h), C.byref(pitch), C.byref(fs)))
        return p.value, w.value, h.value, pitch.value, fs.value

    def results_device(self):
        """(d_kps, d_desc, d_counts, capacity): the last extract's outputs in place on the device."""
        k, d, n = C.c_void_p(), C.c_void_p(), C.c_void_p(); cap = C.c_int()
        _ck(lib().orbx_extractor_results_device(self._h, C.byref(k), C.byref(d), C.byref(n), C.byref(cap)))
        return k.value, d.value, n.value, cap.value

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

    def streams(self, batch):
        """internal streams a batch of this size is spread over"""
        return lib().orbx_extractor_streams(self._h, int(batch))

    def set_profiling(self, on=True):
        _ck(lib().orbx_extractor_set_profiling(self._h, int(on)))

    def assign_grid(self, minx, miny, gw_inv, gh_inv, gcols, grows):
        """Device-side Frame::AssignFeaturesToGrid for every frame of the last batch (enqueued)."""
        _ck(lib().orbx_extractor_assign_grid(self._h, C.c_float(minx), C.c_float(miny), C.c_float(gw_inv), C.c_float(gh_inv),
                                             int(gcols), int(grows)))
        self._grid_cells = int(gcols) * int(grows)

    def grid_fetch(self, f, n):
        off = np.empty(self._grid_cells + 1, np.int32)
        idx = np.empty(max(n, 1), np.int32)
        _ck(lib().orbx_extractor_grid_fetch(self._h, int(f), _p(off), _p(idx), int(n)))
        return off, idx[:n]

    def set_fast_corner_list(self, entries):
        _ck(lib().orbx_extractor_set_fast_corner_list(self._h, int(entries)))

    def stage_ms(self):
        ms = np.empty(6, np.float32)
        _ck(lib().orbx_extractor_stage_ms(self._h, _p(ms)))
        return ms

    def stream(self):
        return lib().orbx_extractor_stream(self._h)


class ImageGrabber:
    """The image half of ros2_orbslam3::ImageGrabber (ros2_ws/src/mono-inertial/include/image_grabber.hpp:38-49, 96-110):
    holds the two undistortion maps and the output size; ConvertImageToGPU = remap(INTER_CUBIC, BORDER_CONSTANT) ->
    resize(INTER_LINEAR) -> BGR2GRAY, batched.  No CPU fallback."""

    def __init__(self, src_size, undistortion_map_1, undistortion_map_2, width, height, max_batch=1, device=0):
        m1 = np.ascontiguousarray(undistortion_map_1, np.float32); m2 = np.ascontiguousarray(undistortion_map_2, np.float32)
        if m1.ndim != 2 or m1.shape != m2.shape:
            raise ValueError("the undistortion maps are two CV_32FC1 planes of one size")
        self.src_w, self.src_h = int(src_size[0]), int(src_size[1])
        self.m_width, self.m_height, self.max_batch, self.device = int(width), int(height), int(max_batch), device
        h = C.c_void_p()
        _ck(lib().orbx_preproc_create(self.src_w, self.src_h, _p(m1), _p(m2), m1.shape[1], m1.shape[0], self.m_width, self.m_height,
                                      self.max_batch, device, C.byref(h)))
        self._h = h

    def close(self):
        if getattr(self, "_h", None) is not None:
            lib().orbx_preproc_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def stream(self):
        return lib().orbx_preproc_stream(self._h)

    def _batch(self, images):
        images = np.asarray(images)
        if images.dtype != np.uint8 or images.ndim != 4 or images.shape[1:] != (self.src_h, self.src_w, 3) or images.strides[2:] != (3, 1):
            raise ValueError(f"expected (B, {self.src_h}, {self.src_w}, 3) u8 BGR frames with dense pixels")
        return images

    def ConvertImageToGPU(self, images):
        """(B, H, W, 3) u8 BGR host frames -> (B, m_height, m_width) u8 gray."""
        images = self._batch(images)
        out = np.empty((images.shape[0], self.m_height, self.m_width), np.uint8)
        _ck(lib().orbx_preproc_host(self._h, images.ctypes.data, images.strides[0], images.strides[1], images.shape[0], _p(out)))
        return out

    def convert_device(self, d_bgr, batch, d_gray, stream=None):
        """device pointers (ints); dense frames; asynchronous on `stream` (None = the handle's)."""
        _ck(lib().orbx_preproc_device(self._h, d_bgr, self.src_w * self.src_h * 3, self.src_w * 3, batch, d_gray, stream))

    def convert_and_extract(self, images, extractor, lapping=None):
        """ConvertImageToGPU -> TrackMonocular -> ORBextractor::operator() chained on the device (asynchronous:
        extractor.sync() / fetch() afterwards)."""
        images = self._batch(images)
        b = images.shape[0]
        extractor._ensure(self.m_width, self.m_height, b)
        lap = np.ascontiguousarray(lapping, np.int32).reshape(b, 2) if lapping is not None else None
        _ck(lib().orbx_preproc_extract_host(self._h, extractor._h, images.ctypes.data, images.strides[0], images.strides[1], b, _p(lap)))


def hamming_pairs(a, b):
    """ORBmatcher::DescriptorDistance over n pairs (reference src/ORBmatcher.cc:1375-1391)."""
    a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
    out = np.empty(len(a), np.int32)
    _ck(lib().orbx_hamming_pairs(_p(a), _p(b), len(a), _p(out)))
    return out


def three_maxima(counts):
    """ORBmatcher::ComputeThreeMaxima (reference src/ORBmatcher.cc:1328-1370)."""
    counts = np.ascontiguousarray(counts, np.int32)
    i = [C.c_int(), C.c_int(), C.c_int()]
    _ck(lib().orbx_three_maxima(_p(counts), len(counts), C.byref(i[0]), C.byref(i[1]), C.byref(i[2])))
    return tuple(v.value for v in i)


# ------------------------------------------------------------------ stereo (Frame::ComputeStereoMatches)
def stereo_matches(left, right, left_frames, right_frames, bf, b):
    """mvuRight / mvDepth of `left`'s frames left_frames[p] against `right`'s frames
    right_frames[p] (device-resident results of the last extract of both handles)."""
    lf = np.ascontiguousarray(left_frames, np.int32); rf = np.ascontiguousarray(right_frames, np.int32)
    n = len(lf)
    ur = np.empty((n, left.capacity), np.float32); dp = np.empty((n, left.capacity), np.float32)
    _ck(lib().orbx_stereo_matches(left._h, right._h, n, _p(lf), _p(rf), bf, b, _p(ur), _p(dp), left.capacity))
    return ur, dp


def stereo_matches_async(left, right, left_frames, right_frames, bf, b):
    lf = np.ascontiguousarray(left_frames, np.int32); rf = np.ascontiguousarray(right_frames, np.int32)
    _ck(lib().orbx_stereo_matches_async(left._h, right._h, len(lf), _p(lf), _p(rf), bf, b))


def stereo_fetch(left, npairs):
    ur = np.empty((npairs, left.capacity), np.float32); dp = np.empty((npairs, left.capacity), np.float32)
    _ck(lib().orbx_stereo_fetch(left._h, npairs, _p(ur), _p(dp), left.capacity))
    return ur, dp


# ------------------------------------------------------------------ matcher
class _FrameView(C.Structure):
    _fields_ = [("n", C.c_int), ("kx", C.c_void_p), ("ky", C.c_void_p), ("koct", C.c_void_p), ("kangle", C.c_void_p),
                ("desc", C.c_void_p), ("minx", C.c_float), ("miny", C.c_float), ("gw_inv", C.c_float),
                ("gh_inv", C.c_float), ("gcols", C.c_int), ("grows", C.c_int), ("scale", C.c_void_p),
                ("nlevels", C.c_int), ("uright", C.c_void_p)]


class _MapPoints(C.Structure):
    _fields_ = [("n", C.c_int), ("inview", C.c_void_p), ("x", C.c_void_p), ("y", C.c_void_p), ("level", C.c_void_p),
                ("viewcos", C.c_void_p), ("depth", C.c_void_p), ("desc", C.c_void_p), ("proj_xr", C.c_void_p),
                ("has_obs", C.c_void_p)]


class _KFPoints(C.Structure):
    _fields_ = [("n", C.c_int), ("valid", C.c_void_p), ("u", C.c_void_p), ("v", C.c_void_p), ("level", C.c_void_p),
                ("angle", C.c_void_p), ("desc", C.c_void_p)]


class _FusePoints(C.Structure):
    _fields_ = [("n", C.c_int), ("valid", C.c_void_p), ("u", C.c_void_p), ("v", C.c_void_p), ("ur", C.c_void_p),
                ("level", C.c_void_p), ("desc", C.c_void_p)]


class _KFView(C.Structure):
    _fields_ = [("n", C.c_int), ("kx", C.c_void_p), ("ky", C.c_void_p), ("koct", C.c_void_p), ("kangle", C.c_void_p),
                ("desc", C.c_void_p), ("has_mp", C.c_void_p), ("uright", C.c_void_p), ("nnodes", C.c_int),
                ("node_id", C.c_void_p), ("node_off", C.c_void_p), ("node_idx", C.c_void_p)]


def _arr(a, dt, keep):
    if a is None:
        return None
    a = np.ascontiguousarray(a, dt)
    keep.append(a)
    return a.ctypes.data


def _kf_view(K, keep):
    """K: dict with kx, ky, koct, [kangle], desc, [has_mp], [uright], node_id, node_off, node_idx
    (include/orbx.h orbx_kf_view)."""
    return _KFView(len(K["desc"]), _arr(K.get("kx"), np.float32, keep), _arr(K.get("ky"), np.float32, keep),
                   _arr(K.get("koct"), np.int32, keep), _arr(K.get("kangle"), np.float32, keep),
                   _arr(K["desc"], np.uint8, keep), _arr(K.get("has_mp"), np.uint8, keep),
                   _arr(K.get("uright"), np.float32, keep), len(K["node_id"]), _arr(K["node_id"], np.int32, keep),
                   _arr(K["node_off"], np.int32, keep), _arr(K["node_idx"], np.int32, keep))


def _frame_view(F, keep):
    """F: dict with kx, ky, koct, [kangle], desc, minx, miny, gw_inv, gh_inv, gcols, grows, scale
    (the Frame members named in include/orbx.h)."""
    return _FrameView(len(F["kx"]), _arr(F["kx"], np.float32, keep), _arr(F["ky"], np.float32, keep),
                      _arr(F["koct"], np.int32, keep), _arr(F.get("kangle"), np.float32, keep),
                      _arr(F["desc"], np.uint8, keep), F["minx"], F["miny"], F["gw_inv"], F["gh_inv"],
                      F["gcols"], F["grows"], _arr(F["scale"], np.float32, keep), len(F["scale"]),
                      _arr(F.get("uright"), np.float32, keep))


def _ptr_array(arrs, keep):
    pa = (C.c_void_p * len(arrs))(*[a if a is None else a.ctypes.data for a in arrs])
    keep.append(pa)
    return pa


class ORBmatcher:
    """Batched device matcher behind the reference's static ORBmatcher functions
    (include/ORBmatcher.h:40-70).  Every method takes a LIST of problems (one CTA each); a single
    dict is accepted as a batch of one and returns un-listed results."""
    TH_LOW, TH_HIGH, HISTO_LENGTH = 30, 100, 30

    def __init__(self, device=0):
        self._h = C.c_void_p()
        _ck(lib().orbx_matcher_create(device, C.byref(self._h)))

    def close(self):
        if self._h:
            lib().orbx_matcher_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def stream(self):
        return lib().orbx_matcher_stream(self._h)

    def last_launches(self):
        return lib().orbx_matcher_last_launches(self._h)

    def synchronize(self):
        _ck(lib().orbx_matcher_synchronize(self._h))

    def set_knn_variant(self, v):
        _ck(lib().orbx_matcher_set_knn_variant(self._h, v))

    @staticmethod
    def DescriptorDistance(a, b):
        return int(hamming_pairs(np.asarray(a, np.uint8).reshape(1, 32), np.asarray(b, np.uint8).reshape(1, 32))[0])

    def SearchByBoW(self, KFs, Fs, nnRatio=0.7, checkOrientation=True, nleft=None):
        """ORBmatcher::SearchByBoW(KeyFrame, Frame): list of problems -> [(nmatches, match_kf)]."""
        single = isinstance(KFs, dict)
        if single:
            KFs, Fs = [KFs], [Fs]
            nleft = None if nleft is None else [nleft]
        keep = []
        a = (_KFView * len(KFs))(*[_kf_view(k, keep) for k in KFs])
        b = (_KFView * len(Fs))(*[_kf_view(f, keep) for f in Fs])
        outs = [np.empty(max(len(f["desc"]), 1), np.int32) for f in Fs]
        po = (C.c_void_p * len(Fs))(*[o.ctypes.data for o in outs])
        nm = np.empty(len(Fs), np.int32)
        nl = None if nleft is None else np.ascontiguousarray(nleft, np.int32)
        _ck(lib().orbx_search_by_bow(self._h, len(KFs), a, b, _p(nl), C.c_float(nnRatio), int(checkOrientation), po, _p(nm)))
        res = [(int(nm[p]), outs[p][:len(Fs[p]["desc"])]) for p in range(len(Fs))]
        return res[0] if single else res

    def FuseSearch(self, KFs, MPs, th=3.0, variant=0, uright=None, inv_sigma2=None):
        """Search core of ORBmatcher::Fuse (variant 0: KeyFrame/MapPoints overload with the
        chi-square gates; 1: Sim3 overload) -> [(best_idx, best_dist)] per problem."""
        single = isinstance(KFs, dict)
        if single:
            KFs, MPs = [KFs], [MPs]
            uright = None if uright is None else [uright]
            inv_sigma2 = None if inv_sigma2 is None else [inv_sigma2]
        keep = []
        fv = (_FrameView * len(KFs))(*[_frame_view(f, keep) for f in KFs])
        mp = (_FusePoints * len(MPs))(*[_FusePoints(len(M["u"]), _arr(M["valid"], np.uint8, keep), _arr(M["u"], np.float32, keep),
                                                    _arr(M["v"], np.float32, keep), _arr(M.get("ur"), np.float32, keep),
                                                    _arr(M["level"], np.int32, keep), _arr(M["desc"], np.uint8, keep)) for M in MPs])
        bi = [np.empty(max(len(M["u"]), 1), np.int32) for M in MPs]
        bd = [np.empty(max(len(M["u"]), 1), np.int32) for M in MPs]
        pur = None if uright is None else (C.c_void_p * len(KFs))(*[_arr(a, np.float32, keep) for a in uright])
        pis = None if inv_sigma2 is None else (C.c_void_p * len(KFs))(*[_arr(a, np.float32, keep) for a in inv_sigma2])
        _ck(lib().orbx_fuse_search(self._h, len(KFs), fv, pur, pis, mp, C.c_float(th), int(variant),
                                   (C.c_void_p * len(MPs))(*[a.ctypes.data for a in bi]),
                                   (C.c_void_p * len(MPs))(*[a.ctypes.data for a in bd])))
        res = [(bi[p][:len(MPs[p]["u"])], bd[p][:len(MPs[p]["u"])]) for p in range(len(MPs))]
        return res[0] if single else res

    def SearchBySim3(self, KF1s, KF2s, MP1in2, MP2in1, th=7.5):
        """ORBmatcher::SearchBySim3 after the host-side projections -> [(nfound, match12)]."""
        single = isinstance(KF1s, dict)
        if single:
            KF1s, KF2s, MP1in2, MP2in1 = [KF1s], [KF2s], [MP1in2], [MP2in1]
        keep = []
        v1 = (_FrameView * len(KF1s))(*[_frame_view(f, keep) for f in KF1s])
        v2 = (_FrameView * len(KF2s))(*[_frame_view(f, keep) for f in KF2s])

        def pts(Ms):
            return (_FusePoints * len(Ms))(*[_FusePoints(len(M["u"]), _arr(M["valid"], np.uint8, keep), _arr(M["u"], np.float32, keep),
                                                         _arr(M["v"], np.float32, keep), None, _arr(M["level"], np.int32, keep),
                                                         _arr(M["desc"], np.uint8, keep)) for M in Ms])
        outs = [np.empty(max(len(f["kx"]), 1), np.int32) for f in KF1s]
        nf = np.empty(len(KF1s), np.int32)
        _ck(lib().orbx_search_by_sim3(self._h, len(KF1s), v1, v2, pts(MP1in2), pts(MP2in1), C.c_float(th),
                                      (C.c_void_p * len(outs))(*[o.ctypes.data for o in outs]), _p(nf)))
        res = [(int(nf[p]), outs[p][:len(KF1s[p]["kx"])]) for p in range(len(KF1s))]
        return res[0] if single else res

    def SearchForInitialization(self, F1s, F2s, windowSize=100, nnRatio=0.9, checkOrientation=True):
        """ORBmatcher::SearchForInitialization -> [(nmatches, vnMatches12)]."""
        single = isinstance(F1s, dict)
        if single:
            F1s, F2s = [F1s], [F2s]
        keep = []
        v1 = (_FrameView * len(F1s))(*[_frame_view(f, keep) for f in F1s])
        v2 = (_FrameView * len(F2s))(*[_frame_view(f, keep) for f in F2s])
        outs = [np.empty(max(len(f["kx"]), 1), np.int32) for f in F1s]
        nm = np.empty(len(F1s), np.int32)
        _ck(lib().orbx_search_for_initialization(self._h, len(F1s), v1, v2, int(windowSize), C.c_float(nnRatio), int(checkOrientation),
                                                 (C.c_void_p * len(outs))(*[o.ctypes.data for o in outs]), _p(nm)))
        res = [(int(nm[p]), outs[p][:len(F1s[p]["kx"])]) for p in range(len(F1s))]
        return res[0] if single else res

    def ComputeDistinctiveDescriptors(self, sets):
        """MapPoint::ComputeDistinctiveDescriptors for a list of (N_s, 32) u8 descriptor sets ->
        (best row per set, its median distance)."""
        off = np.zeros(len(sets) + 1, np.int32)
        off[1:] = np.cumsum([len(d) for d in sets])
        desc = np.ascontiguousarray(np.concatenate([np.asarray(d, np.uint8).reshape(-1, 32) for d in sets]) if len(sets)
                                    else np.zeros((0, 32), np.uint8))
        best = np.empty(len(sets), np.int32); med = np.empty(len(sets), np.int32)
        _ck(lib().orbx_distinctive_descriptors(self._h, len(sets), _p(off), _p(desc), _p(best), _p(med)))
        return best, med

    def MeanValue(self, sets):
        """DBoW2 FORB::meanValue for a list of (N_s, 32) descriptor sets -> (nsets, 32) u8."""
        off = np.zeros(len(sets) + 1, np.int32)
        off[1:] = np.cumsum([len(d) for d in sets])
        desc = np.ascontiguousarray(np.concatenate([np.asarray(d, np.uint8).reshape(-1, 32) for d in sets]) if len(sets)
                                    else np.zeros((0, 32), np.uint8))
        mean = np.zeros((len(sets), 32), np.uint8)
        _ck(lib().orbx_mean_value(self._h, len(sets), _p(off), _p(desc), _p(mean)))
        return mean

    def AssignFeaturesToGrid(self, Fs):
        """Frame::AssignFeaturesToGrid for a list of frame views -> [(cell_off, cell_idx)]."""
        single = isinstance(Fs, dict)
        Fs = [Fs] if single else list(Fs)
        keep = []
        fv = (_FrameView * len(Fs))(*[_frame_view(f, keep) for f in Fs])
        offs = [np.empty(f["gcols"] * f["grows"] + 1, np.int32) for f in Fs]
        idxs = [np.empty(max(len(f["kx"]), 1), np.int32) for f in Fs]
        po = (C.c_void_p * len(Fs))(*[o.ctypes.data for o in offs])
        pi = (C.c_void_p * len(Fs))(*[i.ctypes.data for i in idxs])
        _ck(lib().orbx_assign_features_to_grid(self._h, len(Fs), fv, po, pi))
        out = [(o, i[:len(f["kx"])]) for o, i, f in zip(offs, idxs, Fs)]
        return out[0] if single else out

    def GetFeaturesInArea(self, F, x, y, r, minLevel, maxLevel, cap=256):
        """Frame::GetFeaturesInArea for arrays of queries -> (counts, indices[nq, cap])."""
        keep = []
        fv = _frame_view(F, keep)
        x, y, r = (np.ascontiguousarray(v, np.float32) for v in (x, y, r))
        a, b = (np.ascontiguousarray(v, np.int32) for v in (minLevel, maxLevel))
        cnt = np.empty(len(x), np.int32)
        idx = np.full((len(x), cap), -1, np.int32)
        _ck(lib().orbx_features_in_area(self._h, C.byref(fv), len(x), _p(x), _p(y), _p(r), _p(a), _p(b), _p(cnt), _p(idx), cap))
        for q in range(len(x)):
            idx[q, min(cnt[q], cap):] = -1
        return cnt, idx

    def SearchByProjection(self, F, MP, th, bFarPoints=False, thFarPoints=0.0, nnRatio=0.8, claimed=None):
        """SearchByProjection(Frame, vpMapPoints, th, bFarPoints, thFarPoints, nnRatio, ...)
        reference src/ORBmatcher.cc:31-123 -> (nmatches, match_mp[n])."""
        single = isinstance(F, dict)
        Fs, MPs = ([F], [MP]) if single else (F, MP)
        cl = [claimed] if single else (claimed if claimed is not None else [None] * len(Fs))
        keep = []
        fv = (_FrameView * len(Fs))(*[_frame_view(f, keep) for f in Fs])
        mp = (_MapPoints * len(Fs))(*[_MapPoints(len(m["x"]), _arr(m["inview"], np.uint8, keep), _arr(m["x"], np.float32, keep),
                                                 _arr(m["y"], np.float32, keep), _arr(m["level"], np.int32, keep),
                                                 _arr(m["viewcos"], np.float32, keep), _arr(m.get("depth"), np.float32, keep),
                                                 _arr(m["desc"], np.uint8, keep), _arr(m.get("proj_xr"), np.float32, keep),
                                                 _arr(m.get("has_obs"), np.uint8, keep)) for m in MPs])
        cla = [None if c is None else np.ascontiguousarray(c, np.uint8) for c in cl]
        keep.append(cla)
        outs = [np.empty(len(f["kx"]), np.int32) for f in Fs]
        nm = np.empty(len(Fs), np.int32)
        _ck(lib().orbx_search_by_projection(self._h, len(Fs), fv, mp, _ptr_array(cla, keep), th, int(bFarPoints),
                                            thFarPoints, nnRatio, _ptr_array(outs, keep), _p(nm)))
        return (int(nm[0]), outs[0]) if single else (nm, outs)

    def SearchByProjectionResident(self, ex, frames, grid, MPs, th, bFarPoints=False, thFarPoints=0.0, nnRatio=0.8,
                                   stereo_pairs=None, claimed=None):
        """orbx_search_by_projection_resident: SearchByProjection(Frame, vpMapPoints, ...) for frames of the extractor's
        last batch, the Frame side read in place on the device.  grid: dict with minx, miny, gw_inv, gh_inv, gcols, grows,
        scale.  -> (nmatches[nprob], [match_mp (capacity entries)])."""
        frames = np.ascontiguousarray(frames, np.int32)
        keep = []
        g = dict(grid); g.update(kx=np.zeros(0, np.float32), ky=np.zeros(0, np.float32), koct=np.zeros(0, np.int32), desc=np.zeros((0, 32), np.uint8))
        fv = _frame_view(g, keep)
        mp = (_MapPoints * len(frames))(*[_MapPoints(len(m["x"]), _arr(m["inview"], np.uint8, keep), _arr(m["x"], np.float32, keep),
                                                     _arr(m["y"], np.float32, keep), _arr(m["level"], np.int32, keep),
                                                     _arr(m["viewcos"], np.float32, keep), _arr(m.get("depth"), np.float32, keep),
                                                     _arr(m["desc"], np.uint8, keep), _arr(m.get("proj_xr"), np.float32, keep),
                                                     _arr(m.get("has_obs"), np.uint8, keep)) for m in MPs])
        cap = ex.capacity
        cla = [None if c is None else np.ascontiguousarray(c, np.uint8) for c in (claimed or [None] * len(frames))]
        assert all(c is None or len(c) == cap for c in cla)
        keep.append(cla)
        sp = None if stereo_pairs is None else np.ascontiguousarray(stereo_pairs, np.int32)
        outs = [np.empty(cap, np.int32) for _ in frames]
        nm = np.empty(len(frames), np.int32)
        _ck(lib().orbx_search_by_projection_resident(self._h, ex._h, len(frames), _p(frames), _p(sp), C.byref(fv), mp,
                                                     _ptr_array(cla, keep) if claimed is not None else None, th, int(bFarPoints),
                                                     thFarPoints, nnRatio, _ptr_array(outs, keep), _p(nm)))
        return nm, outs

    def SearchByProjectionKFResident(self, ex, frames, grid, MPs, th, checkOrientation=True, occupied=None):
        """orbx_search_by_projection_kf_resident -> (nmatches[nprob], [match_mp (capacity entries)])."""
        frames = np.ascontiguousarray(frames, np.int32)
        keep = []
        g = dict(grid); g.update(kx=np.zeros(0, np.float32), ky=np.zeros(0, np.float32), koct=np.zeros(0, np.int32), desc=np.zeros((0, 32), np.uint8))
        fv = _frame_view(g, keep)
        mp = (_KFPoints * len(frames))(*[_KFPoints(len(m["u"]), _arr(m["valid"], np.uint8, keep), _arr(m["u"], np.float32, keep),
                                                   _arr(m["v"], np.float32, keep), _arr(m["level"], np.int32, keep),
                                                   _arr(m.get("angle"), np.float32, keep), _arr(m["desc"], np.uint8, keep)) for m in MPs])
        cap = ex.capacity
        oca = [None if c is None else np.ascontiguousarray(c, np.uint8) for c in (occupied or [None] * len(frames))]
        keep.append(oca)
        outs = [np.empty(cap, np.int32) for _ in frames]
        nm = np.empty(len(frames), np.int32)
        _ck(lib().orbx_search_by_projection_kf_resident(self._h, ex._h, len(frames), _p(frames), C.byref(fv), mp,
                                                        _ptr_array(oca, keep) if occupied is not None else None, th,
                                                        int(checkOrientation), _ptr_array(outs, keep), _p(nm)))
        return nm, outs

    def SearchByProjectionKF(self, F, MP, th, checkOrientation=True, occupied=None):
        """SearchByProjection(Frame, KeyFrame, sAlreadyFound, th, checkOrientation)
        reference src/ORBmatcher.cc:1202-1326 -> (nmatches, match_mp[n])."""
        single = isinstance(F, dict)
        Fs, MPs = ([F], [MP]) if single else (F, MP)
        oc = [occupied] if single else (occupied if occupied is not None else [None] * len(Fs))
        keep = []
        fv = (_FrameView * len(Fs))(*[_frame_view(f, keep) for f in Fs])
        mp = (_KFPoints * len(Fs))(*[_KFPoints(len(m["u"]), _arr(m["valid"], np.uint8, keep), _arr(m["u"], np.float32, keep),
                                               _arr(m["v"], np.float32, keep), _arr(m["level"], np.int32, keep),
                                               _arr(m.get("angle"), np.float32, keep), _arr(m["desc"], np.uint8, keep))
                                     for m in MPs])
        oca = [None if c is None else np.ascontiguousarray(c, np.uint8) for c in oc]
        keep.append(oca)
        outs = [np.empty(len(f["kx"]), np.int32) for f in Fs]
        nm = np.empty(len(Fs), np.int32)
        _ck(lib().orbx_search_by_projection_kf(self._h, len(Fs), fv, mp, _ptr_array(oca, keep), th, int(checkOrientation),
                                               _ptr_array(outs, keep), _p(nm)))
        return (int(nm[0]), outs[0]) if single else (nm, outs)

    def SearchForTriangulation(self, K1, K2, F12, ep, scale2, bOnlyStereo=False, bCoarse=False, checkOrientation=True):
        """reference src/ORBmatcher.cc:441-676 -> (nmatches, pairs[npairs, 2])."""
        single = isinstance(K1, dict)
        if single:
            K1, K2, F12, ep, scale2 = [K1], [K2], [F12], [ep], [scale2]
        keep = []

        def kv(K):
            return _KFView(len(K["kx"]), _arr(K["kx"], np.float32, keep), _arr(K["ky"], np.float32, keep),
                           _arr(K["koct"], np.int32, keep), _arr(K.get("kangle"), np.float32, keep),
                           _arr(K["desc"], np.uint8, keep), _arr(K.get("has_mp"), np.uint8, keep),
                           _arr(K.get("uright"), np.float32, keep), len(K["node_id"]), _arr(K["node_id"], np.int32, keep),
                           _arr(K["node_off"], np.int32, keep), _arr(K["node_idx"], np.int32, keep))
        v1 = (_KFView * len(K1))(*[kv(k) for k in K1])
        v2 = (_KFView * len(K2))(*[kv(k) for k in K2])
        Fm = np.ascontiguousarray(np.stack([np.asarray(f, np.float32).reshape(9) for f in F12]))
        E = np.ascontiguousarray(np.asarray(ep, np.float32).reshape(len(K1), 2))
        sc = [np.ascontiguousarray(s_, np.float32) for s_ in scale2]
        keep.append(sc)
        pairs = [np.empty((max(len(k["kx"]), 1), 2), np.int32) for k in K1]
        npairs = np.empty(len(K1), np.int32); nm = np.empty(len(K1), np.int32)
        _ck(lib().orbx_search_for_triangulation(self._h, len(K1), v1, v2, _p(Fm), _p(E), _ptr_array(sc, keep),
                                                int(bOnlyStereo), int(bCoarse), int(checkOrientation),
                                                _ptr_array(pairs, keep), _p(npairs), _p(nm)))
        res = [p[:k].copy() for p, k in zip(pairs, npairs)]
        return (int(nm[0]), res[0]) if single else (nm, res)

    def knn2(self, q, db, ratio=0.8):
        """brute-force 2-NN + ratio test -> d1, i1, d2, i2, pass."""
        q = np.ascontiguousarray(q, np.uint8); db = np.ascontiguousarray(db, np.uint8)
        o = [np.empty(len(q), np.int32) for _ in range(4)]
        ps = np.empty(len(q), np.uint8)
        _ck(lib().orbx_knn2(self._h, _p(q), len(q), _p(db), len(db), ratio, _p(o[0]), _p(o[1]), _p(o[2]), _p(o[3]), _p(ps)))
        return o[0], o[1], o[2], o[3], ps.astype(bool)

    def knn2_device(self, d_q, nq, d_db, ndb, index_base, d_keys):
        """device pointers (ints); asynchronous on the matcher's stream."""
        _ck(lib().orbx_knn2_device(self._h, d_q, nq, d_db, ndb, index_base, d_keys))

    def knn2_device_noalloc(self, d_q, nq, d_db, ndb, index_base, d_keys):
        """orbx_knn2_device_noalloc: as knn2_device, partial keys in the library's static device array."""
        _ck(lib().orbx_knn2_device_noalloc(self._h, d_q, nq, d_db, ndb, index_base, d_keys))

    def knn2_merge_device(self, d_parts, nparts, nq, ratio, d_d1=None, d_i1=None, d_d2=None, d_i2=None, d_pass=None):
        _ck(lib().orbx_knn2_merge_device(self._h, d_parts, nparts, nq, ratio, d_d1, d_i1, d_d2, d_i2, d_pass))
