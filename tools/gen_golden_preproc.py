#!/usr/bin/env python3
"""Generates tests/golden/preproc/preproc_cv2.npz: outputs of the REAL OpenCV (python cv2, version recorded in the file) for the
three pre-processing operations of SURVEY.md §8(f) rank 3 on seeded inputs.  Unlike the extractor fixtures these are
reference-library outputs, so they pin the oracle (and through it the CUDA kernels) without cv2 at test time.

Inputs are regenerated from the seed by tests/preproc_cases.py; only the outputs' CRC32 and a 64x64 corner are stored.
"""
import os
import sys
import zlib

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import preproc_cases as pc  # noqa: E402

out = {"cv2_version": np.array(cv2.__version__)}
for name, (h, w) in pc.SHAPES.items():
    col = pc.colour_image(name, 4)
    for code, cvcode in (("bgr", cv2.COLOR_BGR2GRAY), ("rgb", cv2.COLOR_RGB2GRAY), ("bgra", cv2.COLOR_BGRA2GRAY), ("rgba", cv2.COLOR_RGBA2GRAY)):
        g = cv2.cvtColor(np.ascontiguousarray(col[:, :, :len(code)]), cvcode)
        out["gray_%s_%s_crc" % (name, code)] = np.uint32(zlib.crc32(g.tobytes()))
        out["gray_%s_%s_corner" % (name, code)] = g[:64, :64].copy()
    img = pc.gray_image(name)
    mx, my = pc.rectify_maps(name)
    r = cv2.remap(img, mx, my, cv2.INTER_LINEAR)
    out["remap_%s_crc" % name] = np.uint32(zlib.crc32(r.tobytes()))
    out["remap_%s_corner" % name] = r[:64, :64].copy()
    for clip, tiles in pc.CLAHE_PARAMS:
        c = cv2.createCLAHE(clip, tiles).apply(img)
        key = "clahe_%s_%g_%dx%d" % (name, clip, tiles[0], tiles[1])
        out[key + "_crc"] = np.uint32(zlib.crc32(c.tobytes()))
        out[key + "_corner"] = c[:64, :64].copy()
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "preproc", "preproc_cv2.npz"), **out)
print("wrote", len(out), "entries, cv2", cv2.__version__)
