#!/usr/bin/env python3
"""Stage-by-stage GPU-vs-oracle diagnosis of the extractor (dev tool; prints mismatch statistics, never asserts)."""
import importlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402

orbx = importlib.import_module("3_orb_slam3_selfnote_b200")
synth = importlib.import_module("3_orb_slam3_selfnote_b200.synth")


def diag(w, h, nfeat, seed, fid, taps, lap):
    img = synth.frame(w, h, seed, fid)
    ex = oracle.Extractor(nfeat, blur_taps=taps, sincos_mode=oracle.SINCOS_PORTABLE)
    mono_o, k_o, d_o = ex(img, lap)
    gx = orbx.ORBextractor(nfeat, blur_taps=taps, max_batch=2)
    t = time.time()
    mono_g, k_g, d_g = gx(img, None, lap)
    print("== %dx%d nfeat %d seed %d id %d taps %d lap %s: gpu call %.3fs" % (w, h, nfeat, seed, fid, taps, lap, time.time() - t))
    print("  n oracle %d gpu %d   mono oracle %d gpu %d" % (len(k_o), len(k_g), mono_o, mono_g))
    for l in range(8):
        po, pg = ex.level(l), gx.pyramid_level(l)
        bad = -1 if po.shape != pg.shape else int((po != pg).sum())
        bo = oracle.blur7(po, taps)
        bg = gx.debug_blurred_level(l)
        badb = -1 if bo.shape != bg.shape else int((bo != bg).sum())
        co = ex.level_candidates(l)
        cg = gx.debug_candidates(l)
        so = set(map(tuple, co.tolist())); sg = set(map(tuple, cg.tolist()))
        ko = ex.level_keypoints(l); kg = gx.debug_level_keypoints(l)
        same_kp = len(ko) == len(kg) and bool(np.all(ko["x"] == kg["x"]) and np.all(ko["y"] == kg["y"]) and np.all(ko["response"] == kg["response"]))
        ang = float(np.abs(ko["angle"] - kg["angle"]).max()) if same_kp and len(ko) else -1
        print("  L%d %s pyr_bad %d blur_bad %d cand o/g %d/%d only_o %d only_g %d  kps o/g %d/%d same %s max|dangle| %g"
              % (l, po.shape, bad, badb, len(so), len(sg), len(so - sg), len(sg - so), len(ko), len(kg), same_kp, ang))
        if so != sg:
            print("     only_o", sorted(so - sg)[:8], "only_g", sorted(sg - so)[:8])
        if not same_kp and len(so - sg) == 0 and len(sg - so) == 0:
            n = min(len(ko), len(kg))
            diff = [i for i in range(n) if (ko["x"][i], ko["y"][i]) != (kg["x"][i], kg["y"][i])]
            print("     first kp diffs at", diff[:10], "set equal:", set(zip(ko["x"], ko["y"])) == set(zip(kg["x"], kg["y"])))
    if len(k_o) == len(k_g):
        same = all(np.array_equal(k_o[f], k_g[f]) for f in ("x", "y", "size", "response", "octave", "class_id"))
        print("  final kps equal (excluding angle):", same, " max|dangle|", float(np.abs(k_o["angle"] - k_g["angle"]).max()),
              " angle bit-equal:", bool(np.array_equal(k_o["angle"], k_g["angle"])), " desc rows differing:", int((d_o != d_g).any(axis=1).sum()))
    gx.close()


if __name__ == "__main__":
    print(orbx.lib().orbx_version())
    diag(752, 480, 1000, 1, 0, oracle.TAPS_OCV346, (0, 0))
    diag(752, 480, 1000, 1, 1, oracle.TAPS_OCV4, (0, 1000))
    diag(752, 480, 1200, 2, 5, oracle.TAPS_OCV346, (300, 500))
    diag(512, 512, 1500, 3, 0, oracle.TAPS_OCV346, (0, 0))
    diag(640, 480, 1000, 4, 0, oracle.TAPS_OCV346, (0, 0))
    diag(1280, 720, 2000, 5, 0, oracle.TAPS_OCV346, (0, 0))
    # batch
    imgs = synth.frames(752, 480, 7, range(6))
    gx = orbx.ORBextractor(1000, max_batch=4)
    counts, mono, kps, desc = gx.extract_batch(imgs, (0, 0))
    ex = oracle.Extractor(1000, sincos_mode=oracle.SINCOS_PORTABLE)
    for f in range(6):
        m, k, d = ex(imgs[f], (0, 0))
        ok = counts[f] == len(k) and all(np.array_equal(k[n], kps[f, :len(k)][n]) for n in ("x", "y", "response", "octave", "angle")) \
            and np.array_equal(d, desc[f, :len(k)])
        print("batch frame", f, "count", counts[f], len(k), "mono", mono[f], m, "equal", ok)
