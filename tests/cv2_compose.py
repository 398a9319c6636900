"""The WHOLE upstream ORBextractor::operator() re-composed in Python from cv2 primitives
(SURVEY.md 7.1 step 2, Appendix A): cv2.resize / copyMakeBorder / FastFeatureDetector /
GaussianBlur / fastAtan2 do the pixel work, a straight Python transliteration of
DistributeOctTree (reference src/ORBextractor.cc:151-431, :514-533) with a real std::list-like
Python list and the REAL libstdc++ std::sort (tests/cpp/stdsort_shim.cpp) does the culling, numpy
does the intensity-centroid moments and the rBRIEF sampling.  Nothing here calls the oracle or the
CUDA library: it is the independent second implementation both are checked against
(tests/test_oracle_cv2_live.py).  Test infrastructure only."""
import ctypes as C
import math
import os
import re
import subprocess
import tempfile

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
f32 = np.float32
EDGE, PATCH, HALF = 19, 31, 15

_sort_lib = None


def _stdsort():
    global _sort_lib
    if _sort_lib is None:
        d = tempfile.mkdtemp(prefix="stdsort_shim_")
        so = os.path.join(d, "libstdsort_shim.so")
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-o", so,
                               os.path.join(ROOT, "tests", "cpp", "stdsort_shim.cpp")])
        _sort_lib = C.CDLL(so)
    return _sort_lib


def std_sort_nodes(sizes, ulx):
    n = len(sizes)
    a = np.ascontiguousarray(sizes, np.int32); b = np.ascontiguousarray(ulx, np.int32)
    out = np.empty(max(n, 1), np.int32)
    _stdsort().sort_nodes(n, a.ctypes.data_as(C.c_void_p), b.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p))
    return out[:n]


def pattern31():
    """bit_pattern_31_ as 512 (x, y) points (reference src/cuda/Orb_gpu.cu:52-309; include/orbx_pattern31.inc)."""
    txt = open(os.path.join(ROOT, "include", "orbx_pattern31.inc")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    v = np.array([float(t.rstrip("f")) for t in re.findall(r"-?\d+(?:\.\d*)?f?", txt)], np.float32)
    assert v.size == 1024, v.size
    return v.reshape(512, 2)


def umax_table():
    """reference src/ORBextractor.cc:128-143"""
    umax = [0] * (HALF + 1)
    vmax = int(math.floor(HALF * math.sqrt(2.0) / 2 + 1)); vmin = int(math.ceil(HALF * math.sqrt(2.0) / 2))
    hp2 = HALF * HALF
    for v in range(vmax + 1):
        umax[v] = int(np.rint(math.sqrt(hp2 - v * v)))
    v0 = 0
    for v in range(HALF, vmin - 1, -1):
        while umax[v0] == umax[v0 + 1]:
            v0 += 1
        umax[v] = v0
        v0 += 1
    return umax


class _Node:
    __slots__ = ("ULx", "ULy", "URx", "BRy", "keys", "noMore")

    def __init__(self, ULx, ULy, URx, BRy):
        self.ULx, self.ULy, self.URx, self.BRy = ULx, ULy, URx, BRy
        self.keys = []
        self.noMore = False

    def divide(self):
        """ExtractorNode::DivideNode, reference :151-207"""
        halfX = int(math.ceil(f32(self.URx - self.ULx) / f32(2))); halfY = int(math.ceil(f32(self.BRy - self.ULy) / f32(2)))
        mx, my = self.ULx + halfX, self.ULy + halfY
        n1 = _Node(self.ULx, self.ULy, mx, my); n2 = _Node(mx, self.ULy, self.URx, my)
        n3 = _Node(self.ULx, my, mx, self.BRy); n4 = _Node(mx, my, self.URx, self.BRy)
        for kp in self.keys:
            if kp[0] < mx:
                (n1 if kp[1] < my else n3).keys.append(kp)
            elif kp[1] < my:
                n2.keys.append(kp)
            else:
                n4.keys.append(kp)
        for n in (n1, n2, n3, n4):
            if len(n.keys) == 1:
                n.noMore = True
        return n1, n2, n3, n4


def distribute_octtree(cands, minX, maxX, minY, maxY, N):
    """DistributeOctTree + best-per-node, reference src/ORBextractor.cc:226-431 and :514-533.
    cands: list of (x, y, response) with coordinates relative to (minX, minY), reference order.
    Returns the selected candidates in list order."""
    nIni = int(math.floor(float(f32(maxX - minX) / f32(maxY - minY)) + 0.5))      # C round() of a positive float
    if nIni < 1 or not cands:
        return []
    hX = f32(f32(maxX - minX) / f32(nIni))
    lnodes = []
    ini = []
    for i in range(nIni):
        n = _Node(int(hX * f32(i)), 0, int(hX * f32(i + 1)), maxY - minY)
        lnodes.append(n); ini.append(n)
    for kp in cands:
        ini[int(f32(kp[0]) / hX)].keys.append(kp)
    kept = []
    for n in lnodes:
        if len(n.keys) == 1:
            n.noMore = True; kept.append(n)
        elif n.keys:
            kept.append(n)
    lnodes = kept
    finish = False
    vsp = []

    def add_children(children, count_expand):
        added = 0
        for c in children:
            if c.keys:
                lnodes.insert(0, c); added += 1          # push_front
                if len(c.keys) > 1:
                    count_expand[0] += 1
                    vsp.append((len(c.keys), c))
        return added

    while not finish:
        prevSize = len(lnodes)
        nexp = [0]
        vsp = []
        i = 0
        while i < len(lnodes):
            nd = lnodes[i]
            if nd.noMore:
                i += 1
                continue
            i += add_children(nd.divide(), nexp)        # the iterator still points at nd
            del lnodes[i]                                # lit = erase(lit)
        if len(lnodes) >= N or len(lnodes) == prevSize:
            finish = True
        elif len(lnodes) + nexp[0] * 3 > N:
            while not finish:
                prevSize = len(lnodes)
                prev = vsp
                vsp = []
                order = std_sort_nodes([p[0] for p in prev], [p[1].ULx for p in prev])
                prev = [prev[o] for o in order]
                for j in range(len(prev) - 1, -1, -1):
                    nd = prev[j][1]
                    add_children(nd.divide(), [0])
                    for k, x in enumerate(lnodes):       # lNodes.erase(node->lit)
                        if x is nd:
                            del lnodes[k]
                            break
                    if len(lnodes) >= N:
                        break
                if len(lnodes) >= N or len(lnodes) == prevSize:
                    finish = True
    out = []
    for n in lnodes:
        best = n.keys[0]
        for kp in n.keys[1:]:
            if kp[2] > best[2]:
                best = kp
        out.append(best)
    return out


def level_candidates(level_roi, ini_th, min_th):
    """ComputeKeyPointsOctTree's per-cell FAST loop with cv2.FastFeatureDetector."""
    h, w = level_roi.shape
    minB = EDGE - 3; maxBX = w - EDGE + 3; maxBY = h - EDGE + 3
    width = float(maxBX - minB); height = float(maxBY - minB)
    nCols = int(width / 35); nRows = int(height / 35)
    out = []
    if nCols < 1 or nRows < 1:
        return out
    wCell = math.ceil(width / nCols); hCell = math.ceil(height / nRows)
    det = {t: cv2.FastFeatureDetector_create(t, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16) for t in (ini_th, min_th)}
    for i in range(nRows):
        iniY = minB + i * hCell; maxY = iniY + hCell + 6
        if iniY >= maxBY - 3:
            continue
        maxY = min(maxY, maxBY)
        for j in range(nCols):
            iniX = minB + j * wCell; maxX = iniX + wCell + 6
            if iniX >= maxBX - 6:
                continue
            maxX = min(maxX, maxBX)
            cell = np.ascontiguousarray(level_roi[iniY:maxY, iniX:maxX])
            k = det[ini_th].detect(cell)
            if not k:
                k = det[min_th].detect(cell)
            out += [(p.pt[0] + j * wCell, p.pt[1] + i * hCell, p.response) for p in k]
    return out


def ic_angle(bordered, x, y, umax):
    """IC_Angle (moment sums of reference src/cuda/Angle_gpu.cu:39-69) + cv2.fastAtan2."""
    cy, cx = y + EDGE, x + EDGE
    img = bordered.astype(np.int64)
    u = np.arange(-HALF, HALF + 1)
    m10 = int((u * img[cy, cx - HALF:cx + HALF + 1]).sum())
    m01 = 0
    for v in range(1, HALF + 1):
        d = umax[v]
        uu = np.arange(-d, d + 1)
        lo = img[cy + v, cx - d:cx + d + 1]; up = img[cy - v, cx - d:cx + d + 1]
        m10 += int((uu * (lo + up)).sum())
        m01 += v * int((lo - up).sum())
    return cv2.fastAtan2(float(m01), float(m10))


def orb_descriptor(blurred, x, y, angle_deg, pat):
    """computeOrbDescriptor (bit layout of reference src/cuda/Orb_gpu.cu:324-349):
    fp32 rotation, cvRound (half-even) of the sample offsets."""
    ang = f32(angle_deg) * f32(math.pi / 180.0)
    a = f32(math.cos(float(ang))); b = f32(math.sin(float(ang)))
    px, py = pat[:, 0], pat[:, 1]
    iy = np.rint(px * b + py * a).astype(np.int64)        # float32 products and sum, no contraction
    ix = np.rint(px * a - py * b).astype(np.int64)
    s = blurred[y + iy, x + ix].reshape(32, 8, 2)
    bits = (s[:, :, 0] < s[:, :, 1]).astype(np.uint8)
    return (bits << np.arange(8, dtype=np.uint8)).sum(1).astype(np.uint8)


def extract(image, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7, lapping=(0, 0)):
    """ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)(image, mask, kps, desc, vLappingArea)
    -> (keypoints structured array, descriptors (K, 32) u8, monoIndex), SURVEY App. A1-A4."""
    cv2.setNumThreads(1)
    # tables, reference src/ORBextractor.cc:92-124 (float / double mix as written there)
    scaleFactor = float(f32(scaleFactor))            # `float _scaleFactor` stored in `double scaleFactor` (include/ORBextractor.h:55,109)
    sf = [f32(1.0)]
    for _ in range(1, nlevels):
        sf.append(f32(float(sf[-1]) * scaleFactor))
    inv = [f32(f32(1.0) / s) for s in sf]
    factor = f32(1.0 / scaleFactor)
    nDesired = f32(f32(nfeatures) * (f32(1) - factor) / (f32(1) - f32(math.pow(float(factor), float(nlevels)))))
    nfeat, tot = [], 0
    for _ in range(nlevels - 1):
        nfeat.append(int(np.rint(nDesired))); tot += nfeat[-1]
        nDesired = f32(nDesired * factor)
    nfeat.append(max(nfeatures - tot, 0))
    umax = umax_table(); pat = pattern31()
    h0, w0 = image.shape
    per_level = []
    prev = None
    for l in range(nlevels):
        lw = int(np.rint(f32(w0) * inv[l])); lh = int(np.rint(f32(h0) * inv[l]))
        cur = image if l == 0 else cv2.resize(prev, (lw, lh), interpolation=cv2.INTER_LINEAR)
        bordered = cv2.copyMakeBorder(cur, EDGE, EDGE, EDGE, EDGE, cv2.BORDER_REFLECT_101)
        cands = level_candidates(cur, iniThFAST, minThFAST)
        sel = distribute_octtree(cands, EDGE - 3, lw - EDGE + 3, EDGE - 3, lh - EDGE + 3, nfeat[l])
        kps = [(int(x) + EDGE - 3, int(y) + EDGE - 3, r) for x, y, r in sel]
        angles = [ic_angle(bordered, x, y, umax) for x, y, _ in kps]
        per_level.append((cur, kps, angles))
        prev = cur
    n = sum(len(k) for _, k, _ in per_level)
    dt = np.dtype([("x", "f4"), ("y", "f4"), ("size", "f4"), ("angle", "f4"), ("response", "f4"), ("octave", "i4"), ("class_id", "i4")])
    out_k = np.zeros(n, dt); out_d = np.zeros((n, 32), np.uint8)
    mono, stereo = 0, n - 1
    for l, (cur, kps, angles) in enumerate(per_level):
        if not kps:
            continue
        blurred = cv2.GaussianBlur(cur.copy(), (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
        blurred = cv2.copyMakeBorder(blurred, EDGE, EDGE, EDGE, EDGE, cv2.BORDER_CONSTANT)      # offsets <= 18 < 19 stay inside: padding only for indexing
        size = f32(int(f32(PATCH) * sf[l]))
        for (x, y, r), ang in zip(kps, angles):
            d = orb_descriptor(blurred, x + EDGE, y + EDGE, ang, pat)
            fx, fy = f32(x), f32(y)
            if l != 0:
                fx, fy = f32(fx * sf[l]), f32(fy * sf[l])
            if fx >= lapping[0] and fx <= lapping[1]:
                at = stereo; stereo -= 1
            else:
                at = mono; mono += 1
            out_k[at] = (fx, fy, size, ang, r, l, -1)
            out_d[at] = d
    return out_k, out_d, mono


# ---------------------------------------------------------------- ROS-side pre-processing
def fisheye_maps(width=2048, height=1536):
    """The two CV_32FC1 undistortion maps the way the reference node makes them (cv::fisheye::initUndistortRectifyMap
    with R = I and P = K, ros2_ws/src/mono-inertial/src/mono_inertial_node.cpp:61-68) for a 2048x1536-class camera."""
    K = np.array([[0.653 * width, 0, 0.4985 * width], [0, 0.869 * height, 0.5125 * height], [0, 0, 1]], np.float64)
    D = np.array([-0.008, 0.0615, -0.072, 0.0356], np.float64)
    return cv2.fisheye.initUndistortRectifyMap(K, D, np.eye(3), K, (width, height), cv2.CV_32FC1)


def convert_image(bgr, map1, map2, width, height):
    """ImageGrabber::ConvertImageToGPU (image_grabber.hpp:96-110) from the CPU cv2 functions of the same names."""
    und = cv2.remap(bgr, map1, map2, cv2.INTER_CUBIC, borderMode=cv2.BORDER_CONSTANT, borderValue=0)
    return cv2.cvtColor(cv2.resize(und, (width, height), interpolation=cv2.INTER_LINEAR), cv2.COLOR_BGR2GRAY)
