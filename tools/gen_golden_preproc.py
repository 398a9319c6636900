#!/usr/bin/env python3
"""Generate tests/golden/cv2_preproc.npz with Python cv2 (4.13.0 in the build container): the ROS-side pre-processing
of ros2_ws/src/mono-inertial/include/image_grabber.hpp:96-110 composed from the CPU functions of the same names
    cv2.remap(INTER_CUBIC, BORDER_CONSTANT 0) -> cv2.resize(INTER_LINEAR) -> cv2.cvtColor(BGR2GRAY)
on seeded synthetic BGR frames and seeded radial maps (synth.bgr_image / synth.undistortion_maps): pins the oracle's
restatement (oracle/orb_oracle.cpp orc_preprocess_bgr) and, through it and directly, the CUDA kernels."""
import os, sys
import numpy as np
import cv2
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import orbx_pkg
synth = orbx_pkg.load().synth
cv2.setNumThreads(1)

CASES = {  # name: (src_w, src_h, map_w, map_h, dst_w, dst_h, seed, zoom)
    "half": (320, 240, 320, 240, 160, 120, 3, 0.7),        # the reference's shape: map = source size, resize_factor 0.5
    "odd": (301, 203, 280, 190, 217, 131, 4, 0.8),         # ragged sizes, map != source, non-uniform resize
    "same": (256, 192, 256, 192, 256, 192, 5, 1.1),        # no resize (dsize == ssize is a copy)
    "up": (200, 160, 200, 160, 260, 208, 6, 0.95),         # enlarging resize
}
out = {"cv2_version": np.array(cv2.__version__)}
for name, (sw, sh, mw, mh, dw, dh, seed, zoom) in CASES.items():
    img = synth.bgr_image(sw, sh, seed)
    m1, m2 = synth.undistortion_maps(sw, sh, mw, mh, zoom=zoom)
    und = cv2.remap(img, m1, m2, cv2.INTER_CUBIC, borderMode=cv2.BORDER_CONSTANT, borderValue=0)
    gray = cv2.cvtColor(cv2.resize(und, (dw, dh), interpolation=cv2.INTER_LINEAR), cv2.COLOR_BGR2GRAY)
    out[f"{name}_cfg"] = np.array([sw, sh, mw, mh, dw, dh, seed], np.int64)
    out[f"{name}_zoom"] = np.array(zoom)
    out[f"{name}_imgsum"] = np.array(int(img.astype(np.int64).sum()))
    out[f"{name}_mapsum"] = np.array([int(m1.view(np.uint32).astype(np.uint64).sum()), int(m2.view(np.uint32).astype(np.uint64).sum())], np.uint64)
    out[f"{name}_undsum"] = np.array(int(und.astype(np.int64).sum()))
    out[f"{name}_outside"] = np.array(int(((m1 < -2) | (m2 < -2) | (m1 > sw + 1) | (m2 > sh + 1)).sum()))
    out[f"{name}_gray"] = gray
    print(name, gray.shape, int(out[f"{name}_outside"]), "map pixels fully outside")
path = os.path.join(ROOT, "tests", "golden", "cv2_preproc.npz")
np.savez_compressed(path, **out)
print("wrote", path, os.path.getsize(path), "bytes")
