#!/usr/bin/env python3
"""Generate tests/golden/cv2_orb_descriptors.npz with Python cv2 (4.13.0 in the build container): keypoints, angles and
256-bit descriptors of cv2.ORB (nlevels=1, WTA_K=2, patch 31) on a seeded synthetic image, with the blurred image the
descriptors were sampled from.  Pins the rBRIEF half of computeOrbDescriptor -- the 512-point pattern table
(include/orbx_pattern31.inc vs OpenCV's bit_pattern_31_), the fp32 rotation, the round-half-even of the sample offsets
and the bit order (reference src/cuda/Orb_gpu.cu:52-350) -- against OpenCV's own ORB.
Note on the blur: cv2.ORB blurs a SUBMATRIX of its pyramid buffer, for which cv::GaussianBlur leaves its bit-exact
fixed-point path and falls back to sepFilter2D with the float kernel (integer taps 18,34,49,55 / 256); ORB-SLAM3 blurs
a clone() of the level (fixed-point path, taps 18,34,48,56), which is what the oracle and the CUDA kernel implement and
what tests/test_oracle_golden.py pins against cv2.GaussianBlur.  So the blurred image stored here is cv2.sepFilter2D's,
and only the sampling is under test."""
import os, sys
import numpy as np
import cv2
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import orbx_pkg
synth = orbx_pkg.load().synth
cv2.setNumThreads(1)
W, H, SEED = 400, 300, 99
img = synth.textured_image(W, H, SEED)
orb = cv2.ORB_create(nfeatures=2000, scaleFactor=1.2, nlevels=1, edgeThreshold=31, firstLevel=0, WTA_K=2, patchSize=31, fastThreshold=20)
kps, desc = orb.detectAndCompute(img, None)
k1 = cv2.getGaussianKernel(7, 2, cv2.CV_32F)
blur = cv2.sepFilter2D(img, cv2.CV_8U, k1, k1, borderType=cv2.BORDER_REFLECT_101)
xya = np.array([(k.pt[0], k.pt[1], k.angle) for k in kps], np.float32)
assert np.array_equal(xya[:, :2], np.round(xya[:, :2]))
path = os.path.join(ROOT, "tests", "golden", "cv2_orb_descriptors.npz")
np.savez_compressed(path, cv2_version=np.array(cv2.__version__), cfg=np.array([W, H, SEED], np.int64), imgsum=np.array(int(img.astype(np.int64).sum())),
                    blurred=blur, xya=xya, desc=desc)
print(len(kps), "keypoints;", os.path.getsize(path), "bytes")
