"""Run one golden matcher case (tests/matcher_golden_cases.py) through an implementation that has the oracle's / the
Python mirror's method names, and compare with tests/golden/py_matcher.npz."""
import numpy as np

import matcher_golden_cases as G


def run_case(name, prob, impl, is_gpu):
    kind = name.split("_")[0]
    if kind == "sbp":
        f = impl.SearchByProjection if is_gpu else impl.search_by_projection
        return f(prob["F"], prob["MP"], prob["th"], False, 0.0, prob["ratio"], prob["claimed"])
    if kind == "sbpkf":
        f = impl.SearchByProjectionKF if is_gpu else impl.search_by_projection_kf
        return f(prob["F"], prob["MP"], prob["th"], prob["ori"], prob["occupied"])
    if kind == "bow":
        if is_gpu:
            return impl.SearchByBoW(prob["K1"], prob["K2"], prob["ratio"], prob["ori"], None if prob["nleft"] == -1 else prob["nleft"])
        return impl.search_by_bow(prob["K1"], prob["K2"], prob["ratio"], prob["ori"], prob["nleft"])
    if kind == "init":
        f = impl.SearchForInitialization if is_gpu else impl.search_for_initialization
        return f(prob["F1"], prob["F2"], prob["win"], prob["ratio"], prob["ori"])
    if kind == "tri":
        f = impl.SearchForTriangulation if is_gpu else impl.search_for_triangulation
        return f(prob["K1"], prob["K2"], prob["F12"], prob["ep"], prob["sf"], prob["stereo"], prob["coarse"], prob["ori"])
    if kind == "sim3":
        f = impl.SearchBySim3 if is_gpu else impl.search_by_sim3
        return f(prob["K1"], prob["K2"], prob["a"], prob["b"], prob["th"])
    raise ValueError(name)


def check(name, synth, golden, impl, is_gpu):
    prob = G.CASES[name](synth)
    assert G.checksum(prob) == int(golden[name + "_crc"]), "the synthetic generator changed: regenerate with tools/gen_golden_matcher.py"
    n, m = run_case(name, prob, impl, is_gpu)
    assert int(n) == int(golden[name + "_n"]), (name, int(n), int(golden[name + "_n"]))
    assert np.array_equal(np.asarray(m, np.int32).reshape(golden[name + "_m"].shape), golden[name + "_m"]), name
