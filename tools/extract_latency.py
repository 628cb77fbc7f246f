#!/usr/bin/env python3
"""Host-call latency of ORBextractor::operator() on ONE frame (BASELINE.json configs[0]: 752x480, nFeatures 1000, EuRoC mono
settings) -- the reference's per-frame use -- through orbx_extract (host image in, keypoints / descriptors out).  The CPU side
of the comparison is bench.py's cpu_baseline (`two_thread_value`: one stereo frame at a time, one thread per image).
    python tools/extract_latency.py [reps=200]"""
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
orbx = importlib.import_module("3_orb_slam3_selfnote_b200")
synth = importlib.import_module("3_orb_slam3_selfnote_b200.synth")

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
out = {}
for (w, h, nf, lap, name) in ((752, 480, 1000, (0, 1000), "euroc_mono_752x480_nfeat1000"), (512, 512, 1500, (0, 0), "tumvi_512x512_nfeat1500"),
                              (1280, 720, 2000, (0, 0), "hd720_nfeat2000")):
    imgs = [synth.frame(w, h, 77, i) for i in range(8)]
    gx = orbx.ORBextractor(nf, max_batch=1)
    for i in range(10):
        gx(imgs[i % 8], None, lap)
    ts = []
    for i in range(reps):
        t0 = time.perf_counter()
        mono, k, d = gx(imgs[i % 8], None, lap)
        ts.append(time.perf_counter() - t0)
    ts = np.array(ts) * 1e3
    out[name] = {"gpu_call_ms_median": float(np.median(ts)), "gpu_call_ms_p95": float(np.percentile(ts, 95)), "keypoints": int(len(k))}
    gx.close()
print(json.dumps(out, indent=1))
