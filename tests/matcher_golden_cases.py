"""Seeded matcher problems shared by tools/gen_golden_matcher.py (which writes tests/golden/py_matcher.npz with the
Python restatements' outputs) and the CPU / GPU tests that compare the oracle and the CUDA kernels with those vectors."""
import zlib

import numpy as np


def checksum(prob):
    c = 0
    def walk(v):
        nonlocal c
        if isinstance(v, dict):
            for k in sorted(v):
                walk(v[k])
        elif isinstance(v, np.ndarray):
            c = zlib.crc32(np.ascontiguousarray(v).tobytes(), c)
        elif isinstance(v, (list, tuple)):
            for x in v:
                walk(x)
        elif v is not None:
            c = zlib.crc32(repr(v).encode(), c)
    walk(prob)
    return c


def _sbp(synth, n, nmp, th, ratio, seed, stereo, claimed):
    F = synth.frame_view(n, seed=seed)
    MP = synth.projection_problem(F, nmp, seed=seed + 1, dup=0.4)
    if stereo:
        F, MP = synth.add_stereo(F, MP, seed=seed + 2)
    cl = (np.random.default_rng(seed).random(n) < 0.1).astype(np.uint8) if claimed else None
    return {"F": F, "MP": MP, "th": th, "ratio": ratio, "claimed": cl}


def _sbpkf(synth, n, nmp, th, ori, seed):
    F = synth.frame_view(n, seed=seed)
    P = synth.projection_problem(F, nmp, seed=seed + 1, dup=0.4)
    MP = {"valid": P["valid"], "u": P["x"], "v": P["y"], "level": P["level"], "angle": P["angle"], "desc": P["desc"]}
    return {"F": F, "MP": MP, "th": th, "ori": ori, "occupied": (np.random.default_rng(seed).random(n) < 0.06).astype(np.uint8)}


def _kf(synth, n1, n2, nodes, seed, p_mp):
    K1, K2, F12, ep, sf = synth.keyframe_pair(n1, n2, nodes, seed=seed)
    K1["has_mp"] = (np.random.default_rng(seed).random(n1) < p_mp).astype(np.uint8)
    return K1, K2, F12, ep, sf


def _bow(synth, n1, n2, nodes, ratio, ori, nleft, seed):
    K1, K2, _, _, _ = _kf(synth, n1, n2, nodes, seed, 0.8)
    return {"K1": K1, "K2": K2, "ratio": ratio, "ori": ori, "nleft": nleft}


def _tri(synth, n1, n2, nodes, stereo, coarse, ori, seed):
    K1, K2, F12, ep, sf = _kf(synth, n1, n2, nodes, seed, 0.3)
    return {"K1": K1, "K2": K2, "F12": F12, "ep": ep, "sf": sf, "stereo": stereo, "coarse": coarse, "ori": ori}


def _init(synth, n1, n2, win, ratio, ori, seed):
    F1, F2 = synth.initialization_pair(n1, n2, seed, p_level0=0.7)
    return {"F1": F1, "F2": F2, "win": win, "ratio": ratio, "ori": ori}


def _sim3(synth, n1, n2, th, seed):
    K1, K2, a, b = synth.sim3_problem(n1, n2, seed)
    return {"K1": K1, "K2": K2, "a": a, "b": b, "th": th}


CASES = {
    "sbp_mono": lambda s: _sbp(s, 400, 600, 3.0, 0.8, 501, False, True),
    "sbp_stereo": lambda s: _sbp(s, 400, 700, 7.0, 0.9, 502, True, False),
    "sbpkf_ori": lambda s: _sbpkf(s, 350, 500, 15.0, True, 511),
    "sbpkf_plain": lambda s: _sbpkf(s, 300, 450, 5.0, False, 512),
    "bow_left": lambda s: _bow(s, 300, 330, 18, 0.7, True, -1, 521),
    "bow_right": lambda s: _bow(s, 280, 300, 12, 0.75, True, 170, 522),
    "init_a": lambda s: _init(s, 300, 320, 100, 0.9, True, 531),
    "init_b": lambda s: _init(s, 260, 240, 40, 0.8, False, 532),
    "tri_mono": lambda s: _tri(s, 320, 340, 24, False, False, True, 541),
    "tri_coarse": lambda s: _tri(s, 300, 300, 20, False, True, True, 542),
    "tri_stereo": lambda s: _tri(s, 300, 320, 22, True, False, False, 543),
    "sim3_a": lambda s: _sim3(s, 280, 300, 7.5, 551),
    "sim3_b": lambda s: _sim3(s, 200, 340, 4.0, 552),
}
