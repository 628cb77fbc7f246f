#!/usr/bin/env python3
"""Development: run orbx_stereo_frames_host a few times with ORBX_TRACE=1 and print the device timeline of the last call."""
import ctypes as C
import importlib
import os
import sys
import time

import numpy as np

os.environ["ORBX_TRACE"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
orbx = importlib.import_module("3_orb_slam3_selfnote_b200")
synth = importlib.import_module("3_orb_slam3_selfnote_b200.synth")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
chunks = int(sys.argv[2]) if len(sys.argv) > 2 else 4
W, H = 752, 480
L = np.concatenate([synth.frames(W, H, 1000, range(32), False)] * (B // 32))
R = np.concatenate([synth.frames(W, H, 1000, range(32), True)] * (B // 32))
eL, eR, em = orbx.ORBextractor(1200, max_batch=B // chunks), orbx.ORBextractor(1200, max_batch=B // chunks), orbx.ORBmatcher()
cap = eL.max_keypoints(W, H)
hb = orbx.HostBuffer
hL, hR = hb((B, H, W)), hb((B, H, W))
hL.array[...] = L
hR.array[...] = R
out = [hb((B, cap, 28)), hb((B, cap, 32)), hb((B,), np.int32), hb((B, cap, 28)), hb((B, cap, 32)), hb((B,), np.int32), hb((B, cap), np.float32),
       hb((B, cap), np.float32)]
mbf = np.float32(47.90639384423901)
mb = np.float32(mbf / np.float32(435.2046959714599))
for it in range(4):
    sys.stderr.write("---- call %d\n" % it)
    t0 = time.perf_counter()
    rc = orbx.lib().orbx_stereo_frames_host(em._m, eL._h, eR._h, hL.ptr, hR.ptr, B, W, H, W, C.c_size_t(W * H), C.c_float(mb), C.c_float(mbf),
                                            *[o.ptr for o in out], cap)
    sys.stderr.write("rc %d  host wall %.3f ms\n" % (rc, (time.perf_counter() - t0) * 1e3))
