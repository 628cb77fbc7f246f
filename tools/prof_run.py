#!/usr/bin/env python3
"""Minimal device-resident run of the bench workload for ncu.

    python tools/prof_run.py [B=128] [NB=3] [mono|stereo]

NB batches of B frames (752x480).  `stereo` = bench.py's step (extract L, extract R at nFeatures 1200, ComputeStereoMatches);
`mono` = extraction only at nFeatures 1000.
"""
import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
orbx = importlib.import_module("3_orb_slam3_selfnote_b200")
synth = importlib.import_module("3_orb_slam3_selfnote_b200.synth")

B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
NB = int(sys.argv[2]) if len(sys.argv) > 2 else 3
MODE = sys.argv[3] if len(sys.argv) > 3 else "stereo"
W, H = 752, 480
tile = lambda a: np.concatenate([a] * ((B + 31) // 32))[:B]
devL = torch.from_numpy(tile(synth.frames(W, H, 1000, range(32), False))).cuda()
st = torch.cuda.Stream()
if MODE == "mono":
    gx = orbx.ORBextractor(1000, max_batch=B)
    for _ in range(NB):
        gx.extract_batch_device(devL.data_ptr(), B, W, H, W, W * H, (0, 0), st.cuda_stream)
    st.synchronize()
    counts, mono, kps, desc = gx.results_host(B)
    print("ok", int(counts.sum()))
else:
    devR = torch.from_numpy(tile(synth.frames(W, H, 1000, range(32), True))).cuda()
    exL, exR, mt = orbx.ORBextractor(1200, max_batch=B), orbx.ORBextractor(1200, max_batch=B), orbx.ORBmatcher()
    fx, bf = 435.2046959714599, 47.90639384423901
    for _ in range(NB):
        exL.extract_batch_device(devL.data_ptr(), B, W, H, W, W * H, (0, 0), st.cuda_stream)
        exR.extract_batch_device(devR.data_ptr(), B, W, H, W, W * H, (0, 0), st.cuda_stream)
        mt.stereo_match_batch(exL, exR, float(np.float32(bf) / np.float32(fx)), bf, st.cuda_stream)
    st.synchronize()
    uR, depth, _, _ = mt.stereo_results(B)
    print("ok", int((uR >= 0).sum()))
