#!/usr/bin/env python3
"""Device-resident timing of the pre-processing kernels against their HBM roofline.

    python tools/preproc_bench.py [B=256] [iters=20]

Algorithmic bytes per pixel: cvtColor C+1 (read C channels, write gray); remap 4 (table) + 1 (read) + 1 (write);
CLAHE 1 (histogram read) + 1 (apply read) + 1 (write).  CUDA events on the launching stream; the working set
(B x 752x480 x (C+1)) exceeds L2 for B >= 128.
"""
import importlib
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
pre_mod = importlib.import_module("3_orb_slam3_selfnote_b200.preproc")
import preproc_cases as pc  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
IT = int(sys.argv[2]) if len(sys.argv) > 2 else 20
peak = 6552.0
try:
    peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass
pre = pre_mod.Preprocessor()
st = torch.cuda.Stream()
out = {}


def timed(fn):
    for _ in range(3):
        fn()
    st.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(st):
        e0.record(st)
        for _ in range(IT):
            fn()
        e1.record(st)
    st.synchronize()
    return e0.elapsed_time(e1) / IT


for name, (h, w) in (("euroc", (480, 752)), ("tumvi", (512, 512))):
    px = B * h * w
    g = torch.from_numpy(np.stack([pc.gray_image(name)] * 4)).cuda().repeat((B + 3) // 4, 1, 1)[:B].contiguous()
    dst = torch.empty_like(g)
    for ch in (3, 4):
        col = torch.from_numpy(pc.colour_image(name, ch)).cuda().unsqueeze(0).repeat(B, 1, 1, 1).contiguous()
        ms = timed(lambda: pre.cvt_gray_device(col.data_ptr(), B, w, h, w * ch, h * w * ch, ch, False, dst.data_ptr(), w, h * w, st.cuda_stream))
        gbs = px * (ch + 1) / ms / 1e6
        out["cvt_gray_%s_c%d" % (name, ch)] = {"ms": ms, "GB/s": gbs, "frac": gbs / peak, "frames/s": B / ms * 1e3}
    mx, my = pc.rectify_maps(name)
    t = pre.remap_table(mx, my, (w, h))
    ms = timed(lambda: pre.remap_device(t, g.data_ptr(), B, w, h * w, dst.data_ptr(), w, h * w, st.cuda_stream))
    gbs = (px * 2 + px * 4) / ms / 1e6
    out["remap_%s" % name] = {"ms": ms, "GB/s": gbs, "frac": gbs / peak, "frames/s": B / ms * 1e3,
                              "note": "table re-read per frame (4 B/px, L2-resident: %d KB)" % (h * w * 4 // 1024)}
    ms = timed(lambda: pre.clahe_device(g.data_ptr(), B, w, h, w, h * w, dst.data_ptr(), w, h * w, 3.0, (8, 8), st.cuda_stream))
    gbs = px * 3 / ms / 1e6
    out["clahe_%s" % name] = {"ms": ms, "GB/s": gbs, "frac": gbs / peak, "frames/s": B / ms * 1e3}
print(json.dumps({"batch": B, "iters": IT, "hbm_peak_gbs": peak, "kernels": out}, indent=1))
