#!/usr/bin/env python3
"""Minimal device-resident run of the bench workload for ncu (2 warm-up batches + 2 batches of 256 frames, 752x480)."""
import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
orbx = importlib.import_module("3_orb_slam3_selfnote_b200")
synth = importlib.import_module("3_orb_slam3_selfnote_b200.synth")

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
NB = int(sys.argv[2]) if len(sys.argv) > 2 else 4
base = synth.frames(752, 480, 1000, range(32))
host = np.concatenate([base] * ((B + 31) // 32))[:B]
dev = torch.from_numpy(host).cuda()
gx = orbx.ORBextractor(1000, max_batch=B)
st = torch.cuda.Stream()
for _ in range(NB):
    gx.extract_batch_device(dev.data_ptr(), B, 752, 480, 752, 752 * 480, (0, 0), st.cuda_stream)
st.synchronize()
counts, mono, kps, desc = gx.results_host(B)
print("ok", int(counts.sum()))
