#!/usr/bin/env python3
"""Generate tests/golden/cv2_primitives.npz with Python cv2 (4.13.0 in the build container).

These vectors PIN the oracle's restatement of every OpenCV primitive on the hot path
(SURVEY.md Appendix B): resize INTER_LINEAR, copyMakeBorder REFLECT_101, GaussianBlur 7x7
sigma 2, FAST-9/16 with NMS, fastAtan2 and the rBRIEF sampling/rotation of cv2.ORB.compute.
The reference repo holds no golden vectors for this path."""
import os, sys
import numpy as np
import cv2
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import orbx_pkg
synth = orbx_pkg.load().synth

cv2.setNumThreads(1)
rng = np.random.default_rng(1234)
full = synth.textured_image(752, 480, 0xB200)
img = np.ascontiguousarray(full[100:260, 200:400])          # 160 x 200
out = {"cv2_version": np.array(cv2.__version__), "img": img}
# B1 resize: level chain sizes as ORBextractor computes them (cvRound(w/1.2^l))
r1 = cv2.resize(img, (167, 133), interpolation=cv2.INTER_LINEAR)
r2 = cv2.resize(r1, (139, 111), interpolation=cv2.INTER_LINEAR)
r3 = cv2.resize(img, (97, 155), interpolation=cv2.INTER_LINEAR)       # anisotropic, upscale in y is not used by ORB but pins clamps
out.update(resize_167x133=r1, resize_chain_139x111=r2, resize_97x155=r3)
# B5 border
out["border19"] = cv2.copyMakeBorder(img, 19, 19, 19, 19, cv2.BORDER_REFLECT_101)
# B2 blur
out["blur7"] = cv2.GaussianBlur(img, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
noise = rng.integers(0, 256, size=(64, 80), dtype=np.uint8)
out["noise"] = noise
out["noise_blur7"] = cv2.GaussianBlur(noise, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
# B4 FAST on cell-sized sub-images and the full crop
for name, sub in (("cell_a", img[10:53, 20:61]), ("cell_b", img[80:123, 120:161]), ("full", img), ("noise", noise)):
    for th in (20, 7):
        kps = cv2.FastFeatureDetector_create(th, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16).detect(np.ascontiguousarray(sub))
        arr = np.array([[k.pt[0], k.pt[1], k.response] for k in kps], np.float32).reshape(-1, 3)
        out[f"fast_{name}_th{th}"] = arr
out["cell_a"] = np.ascontiguousarray(img[10:53, 20:61]); out["cell_b"] = np.ascontiguousarray(img[80:123, 120:161])
# B3 fastAtan2 on integer moment-like pairs
yy = rng.integers(-200000, 200000, size=4000).astype(np.float32)
xx = rng.integers(-200000, 200000, size=4000).astype(np.float32)
yy[:8] = [0, 0, 1, -1, 5, -5, 0, 7]; xx[:8] = [0, 1, 0, 0, 5, 5, -3, -7]
out["atan_y"], out["atan_x"] = yy, xx
out["atan_deg"] = np.array([cv2.fastAtan2(float(a), float(b)) for a, b in zip(yy, xx)], np.float32)
# A4 rBRIEF: cv2.ORB.compute with provided keypoints/angles on a single level
big = np.ascontiguousarray(full[40:340, 100:500])             # 300 x 400
orb = cv2.ORB_create(nfeatures=500, scaleFactor=1.2, nlevels=1, edgeThreshold=31, firstLevel=0, WTA_K=2,
                     patchSize=31)
n = 300
xs = rng.integers(40, 360, size=n); ys = rng.integers(40, 260, size=n)
ang = rng.uniform(0, 360, size=n).astype(np.float32)
ang[:6] = [0, 90, 180, 270, 45, 359.99]
kps = [cv2.KeyPoint(float(x), float(y), 31.0, float(a), 1.0, 0) for x, y, a in zip(xs, ys, ang)]
kps2, desc = orb.compute(big, kps)
assert len(kps2) == n, len(kps2)
for k, x, y, a in zip(kps2, xs, ys, ang):
    assert k.pt == (float(x), float(y)) and np.float32(k.angle) == a
# cv2.ORB blurs a SUBMATRIX of its bordered pyramid buffer, which takes cv::GaussianBlur's
# sepFilter2D path (taps round(k*256) = 18,34,49,55,...), not the fixed-point path
# (18,34,48,56,...) that ORB-SLAM3's `clone()`d level takes (pinned separately above, B2).
# Store the image cv2.ORB actually sampled so the test pins sampling/rotation/rounding only.
gk = cv2.getGaussianKernel(7, 2, cv2.CV_32F)
out["orb_blurred"] = cv2.sepFilter2D(big, cv2.CV_8U, gk, gk, borderType=cv2.BORDER_REFLECT_101)
out.update(orb_img=big, orb_x=xs.astype(np.int32), orb_y=ys.astype(np.int32), orb_angle=ang, orb_desc=desc)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "cv2_primitives.npz"), **out)
print("wrote", {k: getattr(v, "shape", None) for k, v in out.items()})
