#!/usr/bin/env python3
"""A/B timing of the extraction stages with alternative builds of liborbx.so:  python tools/fast_ab.py lib1.so [lib2.so ...]
Each library runs in its own process (one CDLL per process); prints the per-stage CUDA-event times of 256 images of 752x480."""
import importlib
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def one(lib):
    import numpy as np
    import torch
    sys.path.insert(0, ROOT)
    orbx = importlib.import_module("3_orb_slam3_selfnote_b200")
    synth = importlib.import_module("3_orb_slam3_selfnote_b200.synth")
    orbx.LIB_PATH = os.path.abspath(lib)
    B, W, H = 256, 752, 480
    imgs = synth.frames(W, H, 1000, range(32), False)
    dev = torch.from_numpy(np.concatenate([imgs] * (B // 32))).cuda()
    ex = orbx.ORBextractor(1200, max_batch=B)
    st = torch.cuda.Stream()
    for _ in range(3):
        ex.extract_batch_device(dev.data_ptr(), B, W, H, W, W * H, (0, 0), st.cuda_stream)
    st.synchronize()
    ex.set_profiling(True)
    acc = None
    N = 10
    for _ in range(N):
        ex.extract_batch_device(dev.data_ptr(), B, W, H, W, W * H, (0, 0), st.cuda_stream)
        torch.cuda.synchronize()
        t = ex.stage_times()
        acc = list(t) if acc is None else [(n, a + b) for (n, a), (_, b) in zip(acc, t)]
    counts, _, kps, desc = ex.results_host(B)
    import zlib
    print(os.path.basename(lib), " ".join(f"{n}={a / N:.4f}" for n, a in acc), "crc", zlib.crc32(desc.tobytes()) ^ zlib.crc32(kps.tobytes()), flush=True)


if __name__ == "__main__":
    if len(sys.argv) == 3 and sys.argv[1] == "--one":
        one(sys.argv[2])
    else:
        for lib in sys.argv[1:]:
            subprocess.run([sys.executable, __file__, "--one", lib])
