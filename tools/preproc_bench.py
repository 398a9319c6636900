#!/usr/bin/env python3
"""Device time of the ROS-side pre-processing (csrc/preproc.cu) at the reference node's shape: 2048x1536 BGR ->
fisheye-class undistortion -> 614x460 gray (mono_inertial_node.cpp:20,59-68), device-resident, CUDA events on the
handle's stream; plus the chained convert+extract step and, with --cpu, cv2 on one host thread."""
import argparse, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import orbx_pkg
pkg = orbx_pkg.load()
orbx, synth = pkg.orbx, pkg.synth

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=32)
ap.add_argument("--iters", type=int, default=20)
ap.add_argument("--cpu", action="store_true")
a = ap.parse_args()
SW, SH = 2048, 1536
DW, DH = int(SW * 0.3), int(SH * 0.3)
m1, m2 = synth.undistortion_maps(SW, SH, SW, SH, k1=-0.12, k2=0.02, zoom=0.97)
B = a.batch
frames = np.stack([synth.bgr_image(SW, SH, 300 + f) for f in range(min(B, 4))])
frames = np.ascontiguousarray(np.concatenate([frames] * ((B + len(frames) - 1) // len(frames)))[:B])
g = orbx.ImageGrabber((SW, SH), m1, m2, DW, DH, max_batch=B)
d_in = torch.from_numpy(frames).cuda()
d_out = torch.empty((B, DH, DW), dtype=torch.uint8, device="cuda")
st = torch.cuda.ExternalStream(g.stream())
torch.cuda.synchronize()
res = {"workload": f"{SW}x{SH} BGR -> remap cubic -> {DW}x{DH} gray", "batch": B}
with torch.cuda.stream(st):
    for _ in range(3):
        g.convert_device(d_in.data_ptr(), B, d_out.data_ptr())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(a.iters):
        g.convert_device(d_in.data_ptr(), B, d_out.data_ptr())
    e1.record(st)
    e1.synchronize()
ms = e0.elapsed_time(e1) / a.iters
res["device_ms_per_batch"] = round(ms, 4)
res["device_frames_per_s"] = round(B / ms * 1e3, 1)
alg = B * (SW * SH * 3 + DW * DH)
res["algorithmic_GBps"] = round(alg / ms / 1e6, 1)
# host entry point (pageable numpy in, gray out)
t0 = time.perf_counter()
for _ in range(3):
    out = g.ConvertImageToGPU(frames)
res["host_frames_per_s"] = round(3 * B / (time.perf_counter() - t0), 1)
assert np.array_equal(out, d_out.cpu().numpy())
if a.cpu:
    import cv2, cv2_compose as cc
    cv2.setNumThreads(1)
    t0 = time.perf_counter(); n = 0
    while time.perf_counter() - t0 < 5:
        ref = cc.convert_image(frames[n % B], m1, m2, DW, DH); n += 1
    res["cv2_1thread_frames_per_s"] = round(n / (time.perf_counter() - t0), 2)
    assert np.array_equal(ref, out[(n - 1) % B])
print(json.dumps(res))
