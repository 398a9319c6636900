#!/usr/bin/env python3
"""bench.py -- ORB front-end throughput on B200 (contract: see DESIGN.md "Measurement").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

A "step" is one pass of the hot path (ORBextractor over one batch of EuRoC-shape frames:
752x480, 1000 features, 8 levels, scale 1.2, FAST 20/7) on every rank.  Prints ONE JSON line.
  value      frames/s, whole job, inputs already resident in HBM (device-timed, max over ranks)
  e2e        frames/s through the C ABI with pinned HOST buffers: H2D of the frames and D2H of
             keypoints+descriptors inside the timed region
  roofline   dominant kernel, algorithmic bytes / CUDA-event time vs MEASURED_PEAKS.json
  cpu_baseline  the CPU oracle (a port of the reference semantics) on the host cores
`--impl reference` times the CPU oracle on all host cores (the reference itself needs
OpenCV-CUDA and cannot be built here)."""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H, NFEAT, NLEVELS, SCALE, INI_TH, MIN_TH = 752, 480, 1000, 8, 1.2, 20, 7
BATCH = 256                      # frames per launch per GPU (128 stereo pairs)
TRAFFIC_FILE = "r2_traffic.json"   # ncu --set full capture the roofline's DRAM-traffic figure is read from (not measured in the run)
# what each stage's kernel is limited by according to that capture (the HBM roofline is the contract's denominator)
LIMITER = {"k_fast_units": "instruction issue + shared-memory occupancy", "k_resize_band": "instruction issue + bulk-copy wait",
           "k_blur7": "instruction issue (DRAM ~50 % of peak)", "k_orient_describe": "DRAM sectors (2x over-fetch) + shared-memory gathers",
           "k_octree": "dependent-step latency"}
METRIC = "orb_frames_per_sec"
UNIT = "frames/s"
WORKLOAD = "euroc_752x480_nf1000_l8_sf1.2_fast20-7_batch256"


def workload_config():
    """The `config` object both arms print (identical by construction: it names the workload, not the run)."""
    return {"workload": WORKLOAD, "image": f"{W}x{H} u8 synthetic textured", "nfeatures": NFEAT, "nlevels": NLEVELS,
            "scale_factor": SCALE, "fast_thresholds": [INI_TH, MIN_TH], "frames_per_step_per_gpu": BATCH,
            "l2_policy": "inputs larger than L2: 2 alternating 92 MB input batches + 360 MB pyramid working set per step",
            "sharding": "independent frames per rank, no collective"}


def level_sizes():
    sf = [np.float32(1.0)]
    for _ in range(1, NLEVELS):
        sf.append(np.float32(np.float64(sf[-1]) * np.float64(np.float32(SCALE))))
    inv = [np.float32(1.0) / s for s in sf]
    return [(int(np.rint(np.float32(W) * i)), int(np.rint(np.float32(H) * i))) for i in inv]


def algorithmic_bytes():
    """stage-wise compulsory traffic per frame."""
    P = [w * h for w, h in level_sizes()]
    pyr = 2 * P[0] + sum(P[:-1]) + sum(P[1:])
    fast = sum(P)
    blur = 2 * sum(P)
    kp = NFEAT * (709 + 512 + 32 + 28)
    return {"pyramid": pyr, "fast": fast, "octree": 0, "order": 0, "blur": blur, "describe": kp}


class ClockSampler(threading.Thread):
    """Samples SM clock + throttle reasons of one GPU during the timed region (pynvml)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag = index, [], set(), False
        self.max_mhz = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40,
                 "sw_thermal_slowdown": 0x20, "hw_power_brake_slowdown": 0x80}
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.05)

    def result(self):
        self.stop_flag = True
        if self.nv is None or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unavailable"]}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def make_frames(n, seed=0xB200):
    import orbx_pkg
    synth = orbx_pkg.load().synth
    return synth.frame_batch(n, W, H, seed=seed, distinct=8)


def cpu_baseline(nthreads, seconds_target=12.0, extended=False):
    """Times the CPU oracle (kind 'port') on a bounded sample of the same workload."""
    from oracle import oracle as O
    frames = make_frames(16)
    t1, _ = O.bench_extract(frames, NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH, iters_per_thread=2, nthreads=1)
    per_frame = t1 / 2
    iters = max(2, int(seconds_target / max(per_frame, 1e-3)))
    s, kps = O.bench_extract(frames, NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH, iters_per_thread=iters, nthreads=nthreads)
    n = iters * nthreads
    out = {"value": n / s, "unit": UNIT, "cores": nthreads, "kind": "port",
           "sample": f"{n} EuRoC-shape frames ({iters}/thread x {nthreads} threads, frames-parallel), "
                     f"single-thread latency {per_frame * 1e3:.1f} ms/frame",
           "single_thread_fps": 1.0 / per_frame, "flags": "-O2 -ffp-contract=off (the parity build)"}
    if extended:
        # C1b (-O3 -march=native), C3 (cv2 SIMD composition), C4 (matcher loops), C5 (brute force)
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import cpu_baselines as cb
        import orbx_pkg
        synth = orbx_pkg.load().synth
        params = (NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH)
        out["extraction_o3_native"] = cb.oracle_o3_native(frames[:8], params, nthreads)
        out["extraction_cv2_composition"] = cb.cv2_composition(frames[:4], params, nthreads)
        out["matching"] = cb.matcher_loops(synth, nthreads)
        out["knn2"] = cb.brute_force(synth, nthreads)
    return out


def single_frame_latency(pkg, local, frame):
    """operator() as the tracking thread calls it: ONE frame from a host buffer, keypoints and
    descriptors back on the host (orbx_extract_host + sync + fetch), wall clock per call."""
    ox = pkg.orbx
    ex1 = ox.ORBextractor(NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH, device=local, max_batch=1).prepare(W, H, 1)
    pin = ox.PinnedBuffer((1, H, W))
    pin.array[0] = frame
    for _ in range(20):
        ex1(pin.array[0], (0, 0))
    ts = []
    for _ in range(200):
        t0 = time.perf_counter()
        ex1(pin.array[0], (0, 0))
        ts.append(time.perf_counter() - t0)
    ex1.close()
    ts = np.array(ts) * 1e3
    return {"median_ms": float(np.median(ts)), "p95_ms": float(np.percentile(ts, 95)), "calls": 200,
            "note": "batch 1 through the C ABI with host buffers: H2D + one CUDA-graph launch (23 kernels, per-level branches) + D2H + sync + fetch per call"}


def cpp_operator_latency(pkg, frame):
    """The drop-in C++ class itself: tools/shim_latency.cpp (include/orbx/ORBextractor.h over the C ABI)
    compiled with g++ and run on one frame in pageable host memory.  None when g++ is unavailable."""
    import subprocess
    import tempfile
    try:
        so = pkg.orbx.build()
        with tempfile.TemporaryDirectory() as td:
            exe = os.path.join(td, "shim_latency")
            subprocess.check_call(["g++", "-std=c++17", "-O2", "-o", exe, os.path.join(ROOT, "tools", "shim_latency.cpp"),
                                   "-L" + os.path.dirname(so), "-lorbx", "-Wl,-rpath," + os.path.dirname(so)],
                                  stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, timeout=120)
            raw = os.path.join(td, "img.raw")
            with open(raw, "wb") as fh:
                fh.write(np.ascontiguousarray(frame).tobytes())
            r = subprocess.run([exe, raw, str(W), str(H), str(NFEAT), "300"], capture_output=True, text=True, timeout=120)
            out = json.loads(r.stdout.strip().splitlines()[-1])
            out["note"] = "ORB_SLAM3::ORBextractor::operator() of include/orbx/ORBextractor.h (C++), frame in pageable host memory, keypoints + descriptors in std::vectors"
            return out
    except Exception as ex:     # noqa: BLE001 - a secondary leg must not take the line down
        return {"unavailable": str(ex)[:160]}


def cpp_matcher_latency(pkg):
    """One SearchByProjection call (nprob = 1) through the C++ layers: tools/matcher_latency.cpp times the flattened-view
    wrapper of include/orbx/ORBmatcher.h and the reference-signature template of include/orbx/ORBmatcher_ref.h (object graphs
    of mock SLAM types; flattening and write-back included) on 1000 keypoints x 1500 map points in ordinary host memory."""
    import struct
    import subprocess
    import tempfile
    try:
        so = pkg.orbx.build()
        F = pkg.synth.frame_view(1000, seed=1)
        MP = pkg.synth.projection_problem(F, 1500, seed=101)
        A = {"sbp.F.kx": F["kx"].astype(np.float32), "sbp.F.ky": F["ky"].astype(np.float32), "sbp.F.koct": F["koct"].astype(np.int32),
             "sbp.F.desc": F["desc"], "sbp.F.grid": np.array([F["minx"], F["miny"], F["gw_inv"], F["gh_inv"]], np.float32),
             "sbp.F.gridi": np.array([F["gcols"], F["grows"]], np.int32), "sbp.F.scale": F["scale"].astype(np.float32),
             "sbp.params": np.array([3.0, 0.0, 0.0, 0.8], np.float32)}
        for k, t in (("inview", np.uint8), ("x", np.float32), ("y", np.float32), ("level", np.int32), ("viewcos", np.float32),
                     ("depth", np.float32), ("desc", np.uint8)):
            A["sbp.MP." + k] = MP[k].astype(t)
        with tempfile.TemporaryDirectory() as td:
            exe, blob = os.path.join(td, "matcher_latency"), os.path.join(td, "in.blob")
            subprocess.check_call(["g++", "-std=c++17", "-O2", "-o", exe, os.path.join(ROOT, "tools", "matcher_latency.cpp"),
                                   "-L" + os.path.dirname(so), "-lorbx", "-Wl,-rpath," + os.path.dirname(so)],
                                  stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, timeout=180)
            with open(blob, "wb") as fh:
                for k, v in A.items():
                    v = np.ascontiguousarray(v)
                    fh.write(struct.pack("<I", len(k))); fh.write(k.encode()); fh.write(struct.pack("<Q", v.nbytes)); fh.write(v.tobytes())
            r = subprocess.run([exe, blob, "300"], capture_output=True, text=True, timeout=120)
            out = json.loads(r.stdout.strip().splitlines()[-1])
            out["note"] = ("C++ callers, median wall-clock ms incl. packing, H2D, kernel, D2H, sync: nprob = 1 (compare "
                           "cpu_baseline.matching.search_by_projection.ms_per_problem_1thread) and one batched call of 64 problems "
                           "(compare ...problems_per_s_all_cores; matching.search_by_projection is the same call through the Python mirror)")
            return out
    except Exception as ex:     # noqa: BLE001 - a secondary leg must not take the line down
        return {"unavailable": str(ex)[:160]}


def _dbg(torch, tag):
    if os.environ.get("ORBX_BENCH_DEBUG_SYNC"):
        torch.cuda.synchronize()
        print("[phase ok]", tag, file=sys.stderr, flush=True)


def matching_bench(pkg, torch, local, ex, B):
    """Secondary numbers of BASELINE.json's metric ("Hamming matches/sec"): device-timed stereo
    row search on the extractor's resident output, batched projection search through the C ABI
    (host buffers, copies included), and brute-force 2-NN descriptor pairs/s (both kernel variants)."""
    ox = pkg.orbx
    out = {}
    dev = torch.device("cuda", local)
    # ---- Frame::ComputeStereoMatches: rectified synthetic pairs (4 distinct, tiled) extracted as
    # frames (2p, 2p+1) of one batch; the search runs on the extractor's device-resident output
    st = torch.cuda.ExternalStream(ex.stream(), device=dev)
    sp = [pkg.synth.stereo_pair(W, H, 0xB200 + i) for i in range(4)]
    sbase = np.stack([im for pr in sp for im in pr])
    _dbg(torch, "before matching")
    spin = ox.PinnedBuffer((B, H, W))          # pinned staging, alive until the function returns
    for i in range(B):
        spin.array[i] = sbase[i % 8]
    ex.extract_host(spin.array); ex.sync()
    _dbg(torch, "matching: extraction")
    lf = np.arange(0, B, 2); rf = lf + 1
    for _ in range(2):
        ox.stereo_matches_async(ex, ex, lf, rf, 435.2 * 0.11, 0.11)
        _dbg(torch, "matching: stereo warm-up call")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 10
    e0.record(st)
    for _ in range(reps):
        ox.stereo_matches_async(ex, ex, lf, rf, 435.2 * 0.11, 0.11)
    e1.record(st)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    ur, _ = ox.stereo_fetch(ex, len(lf))
    out["stereo"] = {"pairs_per_s": len(lf) / (ms * 1e-3), "ms_per_launch": ms, "pairs_per_launch": int(len(lf)),
                     "matched_per_pair": float((ur >= 0).sum() / len(lf)),
                     "note": "rectified synthetic stereo pairs; device time of rows + match/SAD/median-cut kernels on resident extractor output"}
    # ---- Frame::AssignFeaturesToGrid on the device-resident keypoints of the whole batch
    gw_inv, gh_inv = float(np.float32(64) / np.float32(W)), float(np.float32(48) / np.float32(H))
    ex.assign_grid(0.0, 0.0, gw_inv, gh_inv, 64, 48)
    e0.record(st)
    for _ in range(reps):
        ex.assign_grid(0.0, 0.0, gw_inv, gh_inv, 64, 48)
    e1.record(st)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    out["frame_grid"] = {"frames_per_s": B / (ms * 1e-3), "ms_per_launch": ms, "frames_per_launch": int(B),
                         "note": "64x48 CSR grid of every frame of the batch from its device-resident keypoints"}
    # ---- SearchByProjection, 64 problems per call, 1000 keypoints x 1500 map points, th = 3
    m = ox.ORBmatcher(local)
    Fs = [pkg.synth.frame_view(1000, seed=100 + i) for i in range(8)]
    MPs = [pkg.synth.projection_problem(F, 1500, seed=200 + i) for i, F in enumerate(Fs)]
    Fs, MPs = Fs * 8, MPs * 8
    m.SearchByProjection(Fs, MPs, 3.0, False, 0.0, 0.8)
    t0 = time.perf_counter()
    reps = 5
    for _ in range(reps):
        nm, _ = m.SearchByProjection(Fs, MPs, 3.0, False, 0.0, 0.8)
    dt = (time.perf_counter() - t0) / reps
    out["search_by_projection"] = {"problems_per_s": len(Fs) / dt, "matches_per_s": float(nm.sum()) / dt,
                                   "ms_per_call": dt * 1e3, "problems_per_call": len(Fs),
                                   "note": "through the C ABI with host buffers (packing + H2D + kernel + D2H)"}
    # ---- ONE problem per call: the reference's call pattern (src/Tracking.cc:607,835,1115 call these one at a
    # time from the tracking thread); wall clock per call through the C ABI, host buffers in and out
    def lat(fn, reps=30):
        for _ in range(3):
            fn()
        ts = []
        for _ in range(reps):
            t0 = time.perf_counter(); fn(); ts.append(time.perf_counter() - t0)
        return float(np.median(ts) * 1e3)
    P1 = dict(MPs[0]); P1["u"], P1["v"] = P1["x"], P1["y"]
    one = {"search_by_projection_ms": lat(lambda: m.SearchByProjection(Fs[0], MPs[0], 3.0, False, 0.0, 0.8)),
           "search_by_projection_kf_ms": lat(lambda: m.SearchByProjectionKF(Fs[0], P1, 15.0, True))}
    # resident form: the Frame side stays on the device (frame 0 of the batch extracted above), only the map
    # points go up and the matches come down; beside it the host-array call on the same frame
    k0, d0, _ = ex.fetch(0)
    grid = {"minx": 0.0, "miny": 0.0, "gw_inv": gw_inv, "gh_inv": gh_inv, "gcols": 64, "grows": 48, "scale": ex.mvScaleFactor}
    F0 = dict(grid, kx=k0["x"].copy(), ky=k0["y"].copy(), koct=k0["octave"].copy(), kangle=k0["angle"].copy(), desc=d0.copy())
    MP0 = pkg.synth.projection_problem(F0, 1500, seed=777)
    one["search_by_projection_resident_ms"] = lat(lambda: m.SearchByProjectionResident(ex, [0], grid, [MP0], 3.0, False, 0.0, 0.8))
    one["search_by_projection_same_frame_host_arrays_ms"] = lat(lambda: m.SearchByProjection(F0, MP0, 3.0, False, 0.0, 0.8))
    # ---- the widened matcher entry points, 32 problems per call through the C ABI
    def timed(fn, nprob, reps=5):
        fn()
        t0 = time.perf_counter()
        for _ in range(reps):
            fn()
        dt = (time.perf_counter() - t0) / reps
        return {"problems_per_s": nprob / dt, "ms_per_call": dt * 1e3, "problems_per_call": nprob}
    sy = pkg.synth
    kfp = [sy.keyframe_pair(1000, 1100, 90, 300 + i) for i in range(4)]
    for i, p in enumerate(kfp):
        p[0]["has_mp"] = (np.random.default_rng(i).random(1000) < 0.8).astype(np.uint8)
    K1s, K2s = [p[0] for p in kfp] * 8, [p[1] for p in kfp] * 8
    out["search_by_bow"] = timed(lambda: m.SearchByBoW(K1s, K2s, 0.7, True), len(K1s))
    fz = [sy.fuse_problem(F, 1200, 400 + i) for i, F in enumerate(Fs[:4])]
    FF, FM = Fs[:4] * 8, [p[0] for p in fz] * 8
    FU, FI = [p[1] for p in fz] * 8, [p[2] for p in fz] * 8
    out["fuse_search"] = timed(lambda: m.FuseSearch(FF, FM, 3.0, 0, FU, FI), len(FF))
    s3 = [sy.sim3_problem(900, 1000, 500 + i) for i in range(4)] * 8
    out["search_by_sim3"] = timed(lambda: m.SearchBySim3([p[0] for p in s3], [p[1] for p in s3], [p[2] for p in s3], [p[3] for p in s3], 7.5), len(s3))
    ini = [sy.initialization_pair(1000, 1000, 600 + i) for i in range(4)] * 8
    out["search_for_initialization"] = timed(lambda: m.SearchForInitialization([p[0] for p in ini], [p[1] for p in ini], 100, 0.9, True), len(ini))
    rng = np.random.default_rng(7)
    sets = [rng.integers(0, 256, (int(k), 32), dtype=np.uint8) for k in rng.integers(3, 40, 2000)]
    out["distinctive_descriptors"] = timed(lambda: m.ComputeDistinctiveDescriptors(sets), len(sets))
    out["search_by_bow"]["note"] = "all through the C ABI with host buffers; Python-side view construction included"
    F12s, eps, sfs = [p[2] for p in kfp], [p[3] for p in kfp], [p[4] for p in kfp]
    one["search_for_triangulation_ms"] = lat(lambda: m.SearchForTriangulation(kfp[0][0], kfp[0][1], F12s[0], eps[0], sfs[0], False, False, True))
    one["search_by_bow_ms"] = lat(lambda: m.SearchByBoW(kfp[0][0], kfp[0][1], 0.7, True))
    one["search_for_initialization_ms"] = lat(lambda: m.SearchForInitialization(ini[0][0], ini[0][1], 100, 0.9, True))
    one["note"] = ("nprob = 1, median wall-clock ms per call incl. Python view construction, packing, H2D, kernel, D2H, sync; "
                   "compare cpu_baseline.matching.*.ms_per_problem_1thread")
    out["single_problem_latency"] = one
    # ---- brute-force 2-NN kernel variants: 65536 queries x 1M database rows, device resident (the full
    # config-5 shape, 1 M queries x 2 M rows per GPU, is the knn2_sharded leg)
    nq, ndb = 65536, 1 << 20
    g = torch.Generator(device=dev); g.manual_seed(5)
    q = torch.randint(0, 256, (nq, 32), dtype=torch.uint8, device=dev, generator=g)
    db = torch.randint(0, 256, (ndb, 32), dtype=torch.uint8, device=dev, generator=g)
    keys = torch.empty((nq, 2), dtype=torch.int64, device=dev)
    ms_ = torch.cuda.ExternalStream(m.stream(), device=dev)
    torch.cuda.synchronize()
    knn = {}
    for variant in (0, 1):
        m.set_knn_variant(variant)
        m.knn2_device(q.data_ptr(), nq, db.data_ptr(), ndb, 0, keys.data_ptr())
        m.synchronize()
        e0.record(ms_)
        for _ in range(3):
            m.knn2_device(q.data_ptr(), nq, db.data_ptr(), ndb, 0, keys.data_ptr())
        e1.record(ms_)
        m.synchronize()
        t = e0.elapsed_time(e1) / 3
        knn["popc8" if variant == 0 else "csa_popc4"] = {"pairs_per_s": nq * ndb / (t * 1e-3), "ms": t}
    m.set_knn_variant(1)
    # measured integer-pipe ceilings (tools/popc_peak.cu): POPC-only, LOP3-only and the executed mix
    # (16 LOP3 + 4 POPC + 3 add + 1 compare per pair, rows from shared memory)
    pk = popc_peak(pkg)
    knn.update({"queries": nq, "database_rows": ndb, "int_pipe_peaks": pk})
    if "popc32_per_s" in pk:
        csa, p8 = knn["csa_popc4"]["pairs_per_s"], knn["popc8"]["pairs_per_s"]
        knn["roofline"] = {"bound": "integer pipe (POPC + LOP3 issue)", "formulation": "csa_popc4: 16 LOP3 + 4 POPC per pair",
                           "achieved_pairs_per_s": csa, "peak_pairs_per_s": pk["knn_mix_pairs_per_s"],
                           "frac_of_measured_mix_peak": csa / pk["knn_mix_pairs_per_s"],
                           "popc_per_s_executed": csa * 4, "frac_of_measured_popc_peak": csa * 4 / pk["popc32_per_s"],
                           "popc8_variant_frac_of_measured_popc_peak": p8 * 8 / pk["popc32_per_s"]}
    out["knn2"] = knn
    m.close()
    return out


def popc_peak(pkg):
    """Measured POPC / LOP3 / 2-NN-mix ceilings of this GPU: tools/popc_peak.cu, built beside liborbx.so."""
    import subprocess
    exe = os.path.join(os.path.dirname(pkg.orbx.build()), "popc_peak")
    try:
        r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
        return json.loads(r.stdout.strip().splitlines()[-1])
    except Exception as ex:     # noqa: BLE001
        return {"unavailable": str(ex)[:160]}


KNN_NQ, KNN_SHARD_ROWS = 1 << 20, 1 << 21       # BASELINE.json configs[4]: 1 M queries, 16 M rows over 8 GPUs = 2 M rows per rank


def knn2_sharded_bench(pkg, torch, dist, rank, world, local):
    """BASELINE.json configs[4]: brute-force 2-NN of 1 M queries against a database sharded over the
    ranks (2 M rows per rank: 16 M rows at 8 GPUs), local search -> NCCL all-gather of the packed
    top-2 keys (16 B per query per rank) -> merge + ratio test.  Every rank calls this; device-timed
    on the matcher's stream (local search, merge) and torch's stream (all-gather), max over ranks."""
    ox = pkg.orbx
    dev = torch.device("cuda", local)
    nq, rows = KNN_NQ, KNN_SHARD_ROWS
    base = rank * rows
    g = torch.Generator(device=dev); g.manual_seed(5)
    q = torch.randint(0, 256, (nq, 32), dtype=torch.uint8, device=dev, generator=g)        # same on every rank
    gd = torch.Generator(device=dev); gd.manual_seed(500 + rank)
    db = torch.randint(0, 256, (rows, 32), dtype=torch.uint8, device=dev, generator=gd)
    # planted neighbours: query k*1024 is database row (k * 7919) % rows of rank k % world, a few bits flipped
    nplant = 64
    plant = []
    for k in range(nplant):
        r_, row = k % world, (k * 7919) % rows
        plant.append((k * 1024, r_, row))
    own = torch.zeros((nplant, 32), dtype=torch.uint8, device=dev)
    for j, (qi, r_, row) in enumerate(plant):
        if r_ == rank:
            own[j] = db[row]
    if dist is not None:
        dist.all_reduce(own, op=dist.ReduceOp.SUM)       # one rank contributes each row
    own[:, 0] ^= 0x05                                     # 2 bits flipped: distance 2 to the planted row
    for j, (qi, r_, row) in enumerate(plant):
        q[qi] = own[j]
    m = ox.ORBmatcher(local)
    ms_ = torch.cuda.ExternalStream(m.stream(), device=dev)
    cur = torch.cuda.current_stream(dev)
    keys = torch.empty((nq, 2), dtype=torch.int64, device=dev)
    parts = torch.empty((world, nq, 2), dtype=torch.int64, device=dev) if world > 1 else keys
    out = [torch.empty(nq, dtype=torch.int32, device=dev) for _ in range(4)]
    ps = torch.empty(nq, dtype=torch.uint8, device=dev)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]

    def run(n):
        ms_.wait_stream(cur)
        ev[0].record(ms_)
        m.knn2_device(q.data_ptr(), n, db.data_ptr(), rows, base, keys.data_ptr())
        ev[1].record(ms_)
        if world > 1:
            cur.wait_stream(ms_)
            dist.all_gather_into_tensor(parts, keys)
            ev[2].record(cur)
            ms_.wait_stream(cur)
        else:
            ev[2].record(ms_)
        m.knn2_merge_device(parts.data_ptr(), world, n, 0.8, out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(),
                            out[3].data_ptr(), ps.data_ptr())
        ev[3].record(ms_)
        cur.wait_stream(ms_)

    run(nq)                               # warm-up at full size (NCCL channels, kernel load)
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    run(nq)
    torch.cuda.synchronize()
    t = torch.tensor([ev[0].elapsed_time(ev[3]), ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2]), ev[2].elapsed_time(ev[3])],
                     device=dev, dtype=torch.float64)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total, search, gather, merge = (float(v) for v in t)
    # correctness spot checks: planted neighbours found with their GLOBAL index, 16 queries exhaustively on this rank's shard
    d1, i1, d2 = out[0].cpu().numpy(), out[1].cpu().numpy(), out[2].cpu().numpy()
    ok_plant = all(d1[qi] <= 2 and (d1[qi] < 2 or i1[qi] == r_ * rows + row) for qi, r_, row in plant)
    ok_local = True
    for k in range(1, nq, nq // 16):
        dd = (q[k][None] ^ db).to(torch.int16)            # popcount through a byte table
        tab = torch.tensor([bin(v).count("1") for v in range(256)], dtype=torch.int16, device=dev)
        dist_k = tab[dd.long()].sum(1)
        ok_local &= bool(int(dist_k.min()) >= int(d1[k]))     # nothing on this shard beats the merged best
        if world == 1:
            srt = torch.sort(dist_k.to(torch.int32) * (1 << 22) + torch.arange(rows, device=dev, dtype=torch.int32) % (1 << 22))[0][:2]
            ok_local &= bool(int(srt[0]) >> 22 == int(d1[k]) and int(srt[1]) >> 22 == int(d2[k]))
    flags = torch.tensor([int(ok_plant), int(ok_local)], device=dev)
    if dist is not None:
        dist.all_reduce(flags, op=dist.ReduceOp.MIN)
    m.close()
    pairs = float(nq) * rows * world
    return {"queries": nq, "database_rows": rows * world, "rows_per_rank": rows, "ranks": world,
            "pairs_per_s": pairs / (total * 1e-3), "ms_total": total, "ms_local_search": search,
            "ms_all_gather": gather, "all_gather_bytes_per_rank": 16 * nq if world > 1 else 0, "ms_merge": merge,
            "ratio_pass_fraction": float(ps.float().mean()), "planted_found": bool(flags[0]), "exhaustive_spot_check": bool(flags[1]),
            "bound": "the popcount loop (local search) - the all-gather moves 16 B per query per rank",
            "note": "device-resident; database generated on the device (uniform bits, seed 500 + rank), queries seed 5 with 64 planted near-duplicates"}


def other_shapes_bench(pkg, torch, local):
    """BASELINE.json configs 3 and 4 as secondary numbers: device-resident frames/s of the same
    extraction path on the TUM-VI shape (512x512, 1500 features, vLappingArea split) and the
    KITTI shape (1241x376, 2000 features), batch 256, synthetic frames, 3 warm-ups + 10 steps."""
    ox = pkg.orbx
    out = {}
    dev = torch.device("cuda", local)
    for name, (w, h, nf, lap) in {"tumvi_512x512_nf1500_lapping": (512, 512, 1500, (0, 255)),
                                  "kitti_1241x376_nf2000": (1241, 376, 2000, None)}.items():
        B = 256
        ex = ox.ORBextractor(nf, SCALE, NLEVELS, INI_TH, MIN_TH, device=local, max_batch=B).prepare(w, h, B)
        frames = pkg.synth.frame_batch(B, w, h, seed=0xB300, distinct=8)
        bufs = [torch.from_numpy(np.roll(frames, k, axis=0)).to(dev) for k in range(2)]
        laps = None if lap is None else np.tile(np.array(lap, np.int32), (B, 1))
        st = torch.cuda.ExternalStream(ex.stream(), device=dev)
        for i in range(3):
            ex.extract_device(bufs[i & 1].data_ptr(), B, w, h, lapping=laps)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for i in range(10):
            ex.extract_device(bufs[i & 1].data_ptr(), B, w, h, lapping=laps)
        e1.record(st)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        ex.sync()
        n, _ = ex.counts(B)
        out[name] = {"frames_per_s": B / (ms * 1e-3), "ms_per_step": ms, "frames_per_step": B,
                     "keypoints_per_frame": float(np.mean(n))}
        ex.close()
    return out


def preproc_bench(pkg, torch, local):
    """ImageGrabber::ConvertImageToGPU (image_grabber.hpp:96-110) at the reference node's own shape --
    2048x1536 BGR, undistortion remap (cubic), resize_factor 0.3 -> 614x460 gray (mono_inertial_node.cpp:20,59-68) --
    32 frames per step.  device: frames resident in HBM, CUDA events on the handle's stream; chained: pinned host BGR ->
    H2D -> pre-process -> extract (1000 features) -> keypoints + descriptors on the host, wall clock; cpu_port: the
    oracle's restatement of cv::remap / resize / cvtColor on one host thread, 2 frames."""
    ox, synth = pkg.orbx, pkg.synth
    dev = torch.device("cuda", local)
    SW, SH, B = 2048, 1536, 32
    DW, DH = int(SW * 0.3), int(SH * 0.3)
    m1, m2 = synth.undistortion_maps(SW, SH, SW, SH, k1=-0.12, k2=0.02, zoom=0.97)
    base = np.stack([synth.bgr_image(SW, SH, 300 + f) for f in range(2)])
    pin = ox.PinnedBuffer((B, SH, SW, 3))
    for f in range(B):
        pin.array[f] = np.roll(base[f & 1], 17 * f, axis=1)
    g = ox.ImageGrabber((SW, SH), m1, m2, DW, DH, max_batch=B, device=local)
    d_in = torch.from_numpy(pin.array).to(dev)
    d_out = torch.empty((B, DH, DW), dtype=torch.uint8, device=dev)
    st = torch.cuda.ExternalStream(g.stream(), device=dev)
    torch.cuda.synchronize()
    for _ in range(3):
        g.convert_device(d_in.data_ptr(), B, d_out.data_ptr())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(10):
        g.convert_device(d_in.data_ptr(), B, d_out.data_ptr())
    e1.record(st)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    from oracle import oracle as O
    t0 = time.perf_counter()
    ref = [O.preprocess_bgr(pin.array[f], m1, m2, DW, DH) for f in range(2)]
    cpu_s = (time.perf_counter() - t0) / 2
    got = d_out[:2].cpu().numpy()
    assert np.array_equal(got[0], ref[0]) and np.array_equal(got[1], ref[1]), "pre-processing differs from the oracle"
    ex = ox.ORBextractor(NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH, device=local, max_batch=B).prepare(DW, DH, B)
    for _ in range(2):
        g.convert_and_extract(pin.array, ex); ex.sync()
    t0 = time.perf_counter()
    for _ in range(5):
        g.convert_and_extract(pin.array, ex); ex.sync()
    wall = (time.perf_counter() - t0) / 5
    n, _ = ex.counts(B)
    ex.close(); g.close(); pin.free()
    alg = SW * SH * 3 + DW * DH
    return {"workload": f"{SW}x{SH} BGR -> remap INTER_CUBIC -> resize {DW}x{DH} INTER_LINEAR -> BGR2GRAY, {B} frames per step",
            "kernel": "k_remap_resize_gray", "device": {"frames_per_s": B / (ms * 1e-3), "ms_per_step": ms,
                                                       "algorithmic_bytes_per_frame": alg, "gbs": alg * B / (ms * 1e-3) / 1e9},
            "chained_with_extract_e2e": {"frames_per_s": B / wall, "ms_per_step": wall * 1e3, "h2d_bytes_per_step": B * SW * SH * 3,
                                         "keypoints_per_frame": float(np.mean(n))},
            "cpu_port": {"frames_per_s": 1.0 / cpu_s, "cores": 1, "sample": "2 frames"}, "parity": "bit-exact on the 2 sampled frames"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    ncores = os.cpu_count() or 1
    vals, secs = [], []
    base = None
    for _ in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        base = cpu_baseline(ncores, seconds_target=3.0)
        secs.append(time.perf_counter() - t0)
        vals.append(base["value"])
    v = float(np.mean(vals[args.warmup:]))
    base["value"] = v
    if args.steps >= 2:          # the driver's run: the matching-side CPU numbers ride along once
        base.update({k: v_ for k, v_ in cpu_baseline(ncores, seconds_target=1.0, extended=True).items()
                     if k in ("extraction_o3_native", "extraction_cv2_composition", "matching", "knn2")})
    line = {"metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": float(np.mean(secs[args.warmup:])) * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
            "data": "synthetic", "impl": "reference",
            "config": workload_config(),
            "run": {"note": "CPU oracle (port of the reference semantics) on all host cores, each step a bounded sample of the "
                            "workload (cpu_baseline.sample; ms_per_step = wall time of one such sample incl. its single-thread probe); "
                            "the reference's own path needs OpenCV-CUDA and cannot be built in this image"},
            "cpu_baseline": base,
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def bind_to_gpu_numa(torch, local):
    """Pin this rank to the CPUs of its GPU's NUMA node before the pinned staging buffers are
    allocated (first touch places them next to the GPU's PCIe root).  Returns (node, ncpus, saved
    affinity) or None when the topology is not exposed."""
    try:
        pr = torch.cuda.get_device_properties(local)
        path = f"/sys/bus/pci/devices/{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0/numa_node"
        node = int(open(path).read())
        if node < 0:
            return None
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        saved = os.sched_getaffinity(0)
        cpus &= saved
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return node, len(cpus), saved
    except (OSError, ValueError, AttributeError):
        return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="orbx")
    ap.add_argument("--batch", type=int, default=BATCH)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-matching", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl != "reference" else args.warmup
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import orbx_pkg
    pkg = orbx_pkg.load()
    orbx = pkg.orbx
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: liborbx has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    B = args.batch
    numa = bind_to_gpu_numa(torch, local)
    ex = orbx.ORBextractor(NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH, device=local, max_batch=B).prepare(W, H, B)
    stream = torch.cuda.ExternalStream(ex.stream(), device=torch.device("cuda", local))

    # inputs: two distinct device-resident batches (2 x 92 MB) + the 360 MB pyramid working set
    # per step exceed the 126 MB L2, so no step finds its input in cache
    frames = make_frames(B, seed=0xB200 + 1000 * rank)
    NH = int(os.environ.get("ORBX_BENCH_NH", "3"))                                     # batches in flight on the end-to-end path
    WC = os.environ.get("ORBX_BENCH_WC", "0") == "1"                       # write-combined staging (A/B knob)
    pin = [orbx.PinnedBuffer((B, H, W), write_combined=WC) for _ in range(NH)]
    for k in range(NH):
        pin[k].array[:] = np.roll(frames, k, axis=0)
    dev = [torch.from_numpy(p.array).to(f"cuda:{local}") for p in pin[:2]]
    torch.cuda.synchronize()

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def step_device(i):
        ex.extract_device(dev[i & 1].data_ptr(), B, W, H)

    # e2e: three handles (three streams) rotate so the H2D copy of batch i+1 and the D2H copy of
    # batch i-1 overlap the kernels of batch i; every batch's results are still read back (into the
    # handle's pinned buffers) inside the timed region
    exs = [ex] + [orbx.ORBextractor(NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH, device=local, max_batch=B).prepare(W, H, B)
                  for _ in range(NH - 1)]

    def step_host(i, first=False):
        exs[i % NH].extract_host(pin[i % NH].array)
        if i >= NH - 1:
            exs[(i - (NH - 1)) % NH].sync()

    def drain_host(n):
        for i in range(max(n - (NH - 1), 0), n):
            exs[i % NH].sync()

    # ---- device-resident throughput
    for i in range(args.warmup):
        step_device(i)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for i in range(args.steps):
        step_device(i)
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    launches = ex.last_launches() * args.steps
    # ---- per-stage times (CUDA events between stages on the launching stream)
    ex.set_profiling(True)
    stage = np.zeros(6)
    for i in range(args.steps):
        step_device(i)
        stage += ex.stage_ms()
    stage /= args.steps
    ex.set_profiling(False)
    barrier()
    # ---- end to end through the C ABI with host buffers
    for i in range(NH):
        step_host(i)
    drain_host(NH)
    barrier()
    t0 = time.perf_counter()
    h0, h1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    h0.record(stream)
    for i in range(args.steps):
        step_host(i)
    drain_host(args.steps)
    h1.record(stream)
    barrier()
    e2e_ms = max(h0.elapsed_time(h1), (time.perf_counter() - t0) * 1e3 * 0.0)
    e2e_wall_ms = (time.perf_counter() - t0) * 1e3
    clocks = sampler.result()
    n_kp, _ = ex.counts(B)
    # ---- copy-only ceiling of the same end-to-end pattern: the same pinned buffers, the same three streams,
    # H2D of the frames and D2H of keypoints + descriptors, NO kernels (what the host side of the box can deliver)
    d2h_bytes = int(B * ex.capacity * (28 + 32) + 8 * B)
    cstreams = [torch.cuda.ExternalStream(x.stream(), device=torch.device("cuda", local)) for x in exs]
    din = [torch.empty(B * H * W, dtype=torch.uint8, device=f"cuda:{local}") for _ in range(NH)]
    dout = [torch.empty(d2h_bytes, dtype=torch.uint8, device=f"cuda:{local}") for _ in range(NH)]
    hin = [torch.from_numpy(p.array.reshape(-1)) for p in pin]
    hout = [torch.empty(d2h_bytes, dtype=torch.uint8).pin_memory() for _ in range(NH)]

    def copy_step(i):
        k = i % NH
        with torch.cuda.stream(cstreams[k]):
            din[k].copy_(hin[k], non_blocking=True)
            hout[k].copy_(dout[k], non_blocking=True)
        if i >= NH - 1:
            cstreams[(i - (NH - 1)) % NH].synchronize()
    for i in range(NH):
        copy_step(i)
    torch.cuda.synchronize()
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        copy_step(i)
    torch.cuda.synchronize()
    copy_wall_ms = (time.perf_counter() - t0) * 1e3
    barrier()
    del din, dout, hout
    # ---- BASELINE.json configs[1] end to end: host L+R frames in -> extract both -> ComputeStereoMatches ->
    # keypoints + descriptors + mvuRight/mvDepth out, copies in the timed region; frames (2p, 2p+1) = a pair
    sp = [pkg.synth.stereo_pair(W, H, 0xB200 + 31 * rank + i) for i in range(4)]
    sbase = np.stack([im for pr in sp for im in pr])
    spin = [orbx.PinnedBuffer((B, H, W)) for _ in range(NH)]
    for k in range(NH):
        for i in range(B):
            spin[k].array[i] = sbase[(i + 2 * k) % 8]
    lf = np.arange(0, B, 2); rf = lf + 1

    def stereo_step(i):
        k = i % NH
        exs[k].extract_host(spin[k].array)
        orbx.stereo_matches_async(exs[k], exs[k], lf, rf, 435.2 * 0.11, 0.11)
        if i >= NH - 1:
            j = (i - (NH - 1)) % NH
            exs[j].sync(); orbx.stereo_fetch(exs[j], len(lf))
    for i in range(NH):
        stereo_step(i)
    for i in range(max(NH - (NH - 1), 0), NH):
        exs[i % NH].sync(); orbx.stereo_fetch(exs[i % NH], len(lf))
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        stereo_step(i)
    for i in range(max(args.steps - (NH - 1), 0), args.steps):
        exs[i % NH].sync(); orbx.stereo_fetch(exs[i % NH], len(lf))
    torch.cuda.synchronize()
    stereo_wall_ms = (time.perf_counter() - t0) * 1e3
    barrier()
    del spin
    knn_sharded = None
    if not args.no_matching:
        knn_sharded = knn2_sharded_bench(pkg, torch, dist, rank, world, local)

    if dist is not None:
        t = torch.tensor([ms, e2e_wall_ms, copy_wall_ms, stereo_wall_ms], device=f"cuda:{local}", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, e2e_wall_ms, copy_wall_ms, stereo_wall_ms = (float(v) for v in t)
    total_frames = B * args.steps * world
    value = total_frames / (ms * 1e-3)
    e2e_value = total_frames / (e2e_wall_ms * 1e-3)

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        alg = algorithmic_bytes()
        names = ["pyramid", "fast", "octree", "order", "blur", "describe"]
        dom = int(np.argmax(stage))
        stages = {}
        for i, n in enumerate(names):
            gbs = alg[n] * B / (stage[i] * 1e-3) / 1e9 if stage[i] > 0 else 0.0
            stages[n] = {"ms": round(float(stage[i]), 4), "alg_bytes_per_frame": alg[n], "gbs": round(gbs, 1),
                         "frac_hbm": round(gbs / hbm_peak, 4)}
        dom_gbs = stages[names[dom]]["gbs"]
        stage_kernels = {"pyramid": ["k_level0", "k_resize_band"], "fast": ["k_fast_units"], "octree": ["k_octree"], "order": ["k_order"],
                         "blur": ["k_blur7"], "describe": ["k_orient_describe"]}
        kern = stage_kernels[names[dom]][-1]
        # issue-slot view of every stage: warp instructions of one 256-frame step from the committed ncu capture (they do not
        # depend on timing), over the stage time measured in THIS run, against one instruction per clock per scheduler
        # (4 schedulers per SM) at the SM clock sampled in this run
        try:
            cap = json.load(open(os.path.join(ROOT, "profiles", TRAFFIC_FILE)))["kernels"]
            sm_mhz = float(clocks.get("sm_mhz") or 0) or 1965.0
            issue_peak = 148 * 4 * sm_mhz * 1e6
            for n in names:
                wi = sum(cap[k]["warp_instructions"] for k in stage_kernels[n]) * B / 256
                stages[n]["warp_instr"] = int(wi)
                stages[n]["issue_frac"] = round(wi / (stage[names.index(n)] * 1e-3) / issue_peak, 3) if stage[names.index(n)] > 0 else None
            step_wi = sum(stages[n]["warp_instr"] for n in names)
            issue = {"warp_instr_per_step": int(step_wi), "peak_warp_instr_per_s": issue_peak, "sm_mhz_used": sm_mhz,
                     "floor_ms_per_step": round(step_wi / issue_peak * 1e3, 4), "frac_of_issue_ceiling": round(step_wi / issue_peak * 1e3 / (ms / args.steps), 3),
                     "source": TRAFFIC_FILE + " (instruction counts) x this run's times"}
        except Exception:
            issue = None
        traffic = None          # DRAM bytes per launch of the dominant kernel, from the committed ncu --set full capture
        try:
            tr = json.load(open(os.path.join(ROOT, "profiles", TRAFFIC_FILE)))["kernels"][kern]
            traffic = int(tr["dram_bytes"] * B / 256)    # every launch of the kernel in ONE step (the capture holds one step)
        except Exception:
            pass
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": dict(workload_config(), frames_per_step_per_gpu=B),
            "run": {"keypoints_per_frame": float(np.mean(n_kp)), "e2e_batches_in_flight": NH, "internal_streams": ex.streams(B),
                    "numa": None if numa is None else {"node": numa[0], "cpus": numa[1]}},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(B * W * H),
                    "d2h_bytes_per_step": d2h_bytes, "ms_per_step": e2e_wall_ms / args.steps,
                    "device_ms_per_step": e2e_ms / args.steps,
                    "copy_only": {"value": total_frames / (copy_wall_ms * 1e-3), "unit": UNIT, "ms_per_step": copy_wall_ms / args.steps,
                                  "h2d_gb_per_s_all_ranks": total_frames * W * H / (copy_wall_ms * 1e-3) / 1e9,
                                  "frac_of_copy_ceiling": (total_frames / (e2e_wall_ms * 1e-3)) / (total_frames / (copy_wall_ms * 1e-3)),
                                  "note": "the same pinned buffers and streams, H2D + D2H only, no kernels: what the host side of this box delivers"},
                    "stereo_pairs": {"value": (total_frames / 2) / (stereo_wall_ms * 1e-3), "unit": "stereo pairs/s", "ms_per_step": stereo_wall_ms / args.steps,
                                     "pairs_per_step_per_gpu": B // 2,
                                     "note": "BASELINE.json configs[1]: host L+R frames -> extract both -> ComputeStereoMatches -> keypoints, descriptors, "
                                             "mvuRight, mvDepth on the host; H2D/D2H inside the timed region"}},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "hbm", "limiter_measured": LIMITER.get(kern, "issue"), "kernel": kern, "stage": names[dom], "achieved": dom_gbs, "peak": hbm_peak, "unit": "GB/s",
                         "frac": round(dom_gbs / hbm_peak, 4), "traffic": traffic, "traffic_source": TRAFFIC_FILE if traffic else None,
                         "algorithmic_bytes_per_launch": int(alg[names[dom]] * B), "launch_ms": round(float(stage[dom]), 4),
                         "launch_scope": "all launches of the stage for one 256-frame step on one stream (FAST: 2 grids, short and tall strips)",
                         "peak_source": "measured" if peaks else "fallback"},
            "stages": stages,
            "issue_roofline": issue,
        }
        if not args.no_matching:
            line["matching"] = matching_bench(pkg, torch, local, ex, B)
            line["matching"]["knn2_sharded"] = knn_sharded
            line["other_shapes"] = other_shapes_bench(pkg, torch, local)
            line["preprocessing"] = preproc_bench(pkg, torch, local)
            line["single_frame_latency"] = single_frame_latency(pkg, local, frames[0])
            line["cpp_operator_latency"] = cpp_operator_latency(pkg, frames[0])
            line["cpp_matcher_latency"] = cpp_matcher_latency(pkg)
        if numa is not None:
            os.sched_setaffinity(0, numa[2])          # the CPU baseline gets every core back
        if not args.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = cpu_baseline(os.cpu_count() or 1, extended=not args.no_matching)
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
