#!/usr/bin/env python3
"""Generate tests/golden/cv2_extractor.npz with Python cv2 (4.13.0 in the build container):
  * the WHOLE upstream ORBextractor::operator() re-composed from cv2 primitives + a straight Python
    quadtree (tests/cv2_compose.py) on BASELINE.json's three image configs:
    keypoints, descriptors, monoIndex -- pins the oracle's (and the CUDA path's) composition:
    cell loop -> quadtree -> orientation -> blur -> descriptor -> operator() tail;
  * IC_Angle pins: cv2.ORB_create(nlevels=1).detect() keypoints with the angles OpenCV computed
    (same IC_Angle + fastAtan2 as ORB-SLAM3), on two crops.
The reference repo holds no golden vectors for this path."""
import os, sys
import numpy as np
import cv2
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import orbx_pkg
import cv2_compose as cc
synth = orbx_pkg.load().synth
cv2.setNumThreads(1)

CONFIGS = {  # name: (w, h, nfeatures, lapping, image seed) -- BASELINE.json configs[0..3]
    "euroc": (752, 480, 1000, (0, 0), 0xB200),
    "tumvi": (512, 512, 1500, (256, 511), 7),
    "kitti": (1241, 376, 2000, (0, 0), 11),
}
out = {"cv2_version": np.array(cv2.__version__)}
for name, (w, h, nf, lap, seed) in CONFIGS.items():
    img = synth.textured_image(w, h, seed)
    k, d, mono = cc.extract(img, nf, 1.2, 8, 20, 7, lap)
    out[f"{name}_cfg"] = np.array([w, h, nf, lap[0], lap[1], seed], np.int64)
    out[f"{name}_kps"] = k
    out[f"{name}_desc"] = d
    out[f"{name}_mono"] = np.array(mono)
    out[f"{name}_imgsum"] = np.array(int(img.astype(np.int64).sum()))     # guards the synthetic generator
    print(name, len(k), mono)
full = synth.textured_image(752, 480, 0xB200)
for name, crop, th in (("a", full[40:340, 100:500], 20), ("b", full[150:470, 300:740], 7)):
    crop = np.ascontiguousarray(crop)
    orb = cv2.ORB_create(nfeatures=3000, scaleFactor=1.2, nlevels=1, edgeThreshold=31, firstLevel=0, WTA_K=2,
                         patchSize=31, fastThreshold=th)
    kps = orb.detect(crop)
    arr = np.array([[k.pt[0], k.pt[1], k.angle] for k in kps], np.float32)
    assert np.all(arr[:, :2] == np.floor(arr[:, :2]))
    out[f"angle_{name}_img"] = crop
    out[f"angle_{name}_xya"] = arr
    print("angles", name, len(arr))
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "cv2_extractor.npz"), **out)
