#!/usr/bin/env python3
"""Generate tests/golden/py_matcher.npz: outputs of the straight-Python restatements of the matcher functions
(tests/test_oracle_matcher_py.py, tests/test_oracle_matcher.py -- written from the reference text, independent of
oracle/orb_oracle.cpp) on seeded synthetic problems.  Committed so that the C++ oracle (CPU suite) and the CUDA kernels
(GPU suite) are compared with vectors neither of them produced.  Inputs are regenerated from seeds by synth.py; their
checksums are stored to catch a drifting generator."""
import os, sys, zlib
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import orbx_pkg
import matcher_golden_cases as G
import test_oracle_matcher_py as P
import test_oracle_matcher as N

synth = orbx_pkg.load().synth
out = {}
for name, build in G.CASES.items():
    prob = build(synth)
    out[name + "_crc"] = np.array(G.checksum(prob), np.uint32)
    kind = name.split("_")[0]
    if kind == "sbp":
        n, m = N._sbp_numpy(prob["F"], prob["MP"], prob["th"], prob["ratio"], prob.get("claimed"))
    elif kind == "sbpkf":
        n, m = P.py_search_by_projection_kf(prob["F"], prob["MP"], prob["th"], prob["ori"], prob.get("occupied"))
    elif kind == "bow":
        n, m = P.py_search_by_bow(prob["K1"], prob["K2"], prob["ratio"], prob["ori"], prob["nleft"])
    elif kind == "init":
        n, m = P.py_search_for_initialization(prob["F1"], prob["F2"], prob["win"], prob["ratio"], prob["ori"])
    elif kind == "tri":
        n, m = P.py_search_for_triangulation(prob["K1"], prob["K2"], prob["F12"], prob["ep"], prob["sf"], prob["stereo"], prob["coarse"], prob["ori"])
    elif kind == "sim3":
        n, m = P.py_search_by_sim3(prob["K1"], prob["K2"], prob["a"], prob["b"], prob["th"])
    else:
        raise SystemExit(name)
    out[name + "_n"] = np.array(n, np.int32); out[name + "_m"] = np.asarray(m, np.int32)
    print(name, int(n), np.asarray(m).shape)
path = os.path.join(ROOT, "tests", "golden", "py_matcher.npz")
np.savez_compressed(path, **out)
print("wrote", path, os.path.getsize(path), "bytes")
