#!/usr/bin/env python3
"""Host-call latency of the single-frame searches (the reference's per-frame use): SearchByProjection(F, 5000 map points),
SearchByProjectionWindow (frame-to-frame), SearchByBoW.   python tools/match_latency.py"""
import importlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
orbx = importlib.import_module("3_orb_slam3_selfnote_b200")
synth = importlib.import_module("3_orb_slam3_selfnote_b200.synth")
from bowsynth import feature_vector  # noqa: E402

W, H = 752, 480


def extract(img, nfeat):
    ex = orbx.ORBextractor(nfeat, max_batch=1)
    mono, k, d = ex(img, None, (0, 0))
    sf = ex.GetScaleFactors()
    ex.close()
    return k, d, np.asarray(sf, np.float32)


def bench(fn, reps=30):
    for _ in range(3):
        fn()
    t0 = time.perf_counter()
    for _ in range(reps):
        r = fn()
    return (time.perf_counter() - t0) / reps * 1e3, r


k, d, sf = extract(synth.frame(W, H, 31, 0, False), 1500)
kj, dj, _ = extract(synth.frame(W, H, 31, 0, False, True), 1500)
rng = np.random.default_rng(0)
nmp = 5000
pad = nmp - len(kj)
px = np.concatenate([kj["x"], rng.uniform(0, W, pad)]).astype(np.float32)
py = np.concatenate([kj["y"], rng.uniform(0, H, pad)]).astype(np.float32)
lv = np.concatenate([kj["octave"], rng.integers(0, 8, pad)]).astype(np.int32)
md = np.concatenate([dj, synth.random_descriptors(pad, 1)])
vc = np.where(np.arange(nmp) % 2 == 0, 1.0, 0.9).astype(np.float32)
F = orbx.FrameView(k, d, sf, W, H)
m = orbx.ORBmatcher(0.8, True)
for th in (1.0, 3.0):
    ms, (nm, _) = bench(lambda: m.SearchByProjection(F, np.ones(nmp, np.uint8), px, py, lv, vc, md, th))
    print("SearchByProjection(F, %d map points, th=%g): %.3f ms per call, %d matches" % (nmp, th, ms, nm))
n = len(kj)
ms, (nm, _) = bench(lambda: m.SearchByProjectionWindow(F, np.ones(n, np.uint8), kj["x"], kj["y"], kj["octave"].astype(np.int32), kj["angle"], dj, 15.0))
print("SearchByProjection(Cur, Last) window search, %d points, th=15: %.3f ms per call, %d matches" % (n, ms, nm))
fv1, fv2 = feature_vector(dj, 100, 3), feature_vector(d, 100, 3)
ms, (nm, _) = bench(lambda: m.SearchByBoW(kj, dj, fv1, np.ones(n, np.uint8), k, d, fv2, None, 50))
print("SearchByBoW(KF, F), %d x %d features, 100 nodes: %.3f ms per call, %d matches" % (n, len(k), ms, nm))
