#!/usr/bin/env python3
"""Generates tests/golden/*.npz (dev-time tool, run in the build container; the fixtures are committed).

Each fixture: a seeded synthetic frame spec, per-level CRC32 of the pyramid, the keypoints and descriptors returned by the
REFERENCE'S OWN ORBextractor (oracle/_ref: /root/reference/src/ORBextractor.cc compiled unmodified against the stand-in
cv:: headers, blur taps OCV346) -- `source` = "reference".  The restated oracle must reproduce them byte for byte in both
sin/cos modes, else this script stops.  Without /root/reference the restatement is used and `source` says "oracle".
"""
import importlib
import os
import sys
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
from oracle import ref as oref  # noqa: E402

synth = importlib.import_module("3_orb_slam3_selfnote_b200.synth")
OUT = os.path.join(ROOT, "tests", "golden")

CASES = [
    # name, w, h, nfeatures, seed, id, right, lapping
    ("euroc_mono", 752, 480, 1000, 1, 0, False, (0, 1000)),
    ("euroc_stereo_left", 752, 480, 1200, 2, 3, False, (0, 0)),
    ("euroc_stereo_right", 752, 480, 1200, 2, 3, True, (0, 0)),
    ("tumvi_512", 512, 512, 1500, 3, 0, False, (0, 0)),
    ("tumvi_lapping", 512, 512, 1000, 3, 1, False, (0, 255)),
    ("hd_720p", 1280, 720, 2000, 5, 0, False, (0, 0)),
]

if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    for name, w, h, nfeat, seed, fid, right, lap in CASES:
        img = synth.frame(w, h, seed, fid, right)
        ex = oracle.Extractor(nfeat, blur_taps=oracle.TAPS_OCV346, sincos_mode=oracle.SINCOS_PORTABLE)
        mono, kps, desc = ex(img, lap)
        crcs = np.array([zlib.crc32(ex.level(l).tobytes()) for l in range(8)], np.uint32)
        source = "oracle"
        if oref.available():
            rx = oref.RefExtractor(nfeat, blur_taps=oracle.TAPS_OCV346)
            mono_r, kps_r, desc_r = rx(img, lap)
            crcs_r = np.array([zlib.crc32(rx.level(l).tobytes()) for l in range(8)], np.uint32)
            assert mono_r == mono and kps_r.tobytes() == kps.tobytes() and np.array_equal(desc_r, desc) and np.array_equal(crcs, crcs_r), \
                "restatement differs from the reference source on " + name
            mono, kps, desc, crcs, source = mono_r, kps_r, desc_r, crcs_r, "reference"
        ex2 = oracle.Extractor(nfeat, blur_taps=oracle.TAPS_OCV346, sincos_mode=oracle.SINCOS_LIBM_F32)
        mono2, kps2, desc2 = ex2(img, lap)
        assert mono2 == mono and np.array_equal(kps, kps2)
        ndiff = int((desc != desc2).any(axis=1).sum())
        np.savez_compressed(os.path.join(OUT, name + ".npz"), spec=np.array([w, h, nfeat, seed, fid, int(right), lap[0], lap[1]], np.int64),
                            image_crc=np.uint32(zlib.crc32(img.tobytes())), pyramid_crc=crcs, mono=np.int32(mono), keypoints=kps,
                            descriptors=desc, libm_desc_rows_differing=np.int32(ndiff), source=np.array(source))
        print(name, "source", source, "n", len(kps), "mono", mono, "libm-mode descriptor rows differing", ndiff)
