#!/usr/bin/env python3
"""Ad-hoc GPU check (no torch): the CUDA extractor against the CPU oracle on seeded random geometries (the generator of
tests/test_oracle_ref.py::test_random_geometries_equal_reference_source).  python tools/random_geometry_check.py [ncases]"""
import importlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
orbx = importlib.import_module("3_orb_slam3_selfnote_b200")
synth = importlib.import_module("3_orb_slam3_selfnote_b200.synth")
import oracle  # noqa: E402  (checker)

oracle.build()
rng = np.random.default_rng(2024)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 16
t0 = time.time()
ran = bad = 0
for t in range(n):
    w, h = int(rng.integers(80, 900)), int(rng.integers(80, 700))
    if w < 0.55 * h:
        w = int(0.6 * h)
    nfeat = int(rng.integers(50, 2500)); scale = float(np.float32(rng.uniform(1.05, 1.6)))
    nlev = int(rng.integers(2, 11)); ini = int(rng.integers(8, 40)); mn = int(rng.integers(2, ini + 1))
    lap = (0, 0) if t % 3 else (int(rng.integers(0, w // 2)), int(rng.integers(w // 2, w)))
    if min(w, h) / scale ** (nlev - 1) < 60:
        nlev = max(2, int(np.log(min(w, h) / 60) / np.log(scale)))
    if any(round((round(w / scale ** l) - 32) / max(1, round(h / scale ** l) - 32)) < 1 for l in range(nlev)):
        continue
    img = synth.frame(w, h, 500 + t, t)
    ex = oracle.Extractor(nfeat, scale, nlev, ini, mn, sincos_mode=oracle.SINCOS_PORTABLE)
    mo, ko, do = ex(img, lap)
    try:
        gx = orbx.ORBextractor(nfeat, scale, nlev, ini, mn)
        mg, kg, dg = gx(img, None, lap)
        ok = mg == mo and len(kg) == len(ko) and kg.tobytes() == ko.tobytes() and np.array_equal(dg, do)
        what = "" if ok else "%d vs %d keypoints" % (len(kg), len(ko))
    except Exception as e:   # noqa: BLE001
        ok, what = False, repr(e)
    ran += 1
    if not ok:
        bad += 1
        print("MISMATCH case %d: %dx%d nfeat %d scale %.4f levels %d fast %d/%d lap %s: %s" % (t, w, h, nfeat, scale, nlev, ini, mn, lap, what), flush=True)
print("random geometries: %d cases, %d mismatches, %.1f s" % (ran, bad, time.time() - t0), flush=True)
