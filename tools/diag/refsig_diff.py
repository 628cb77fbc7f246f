"""Diagnostic: where the drop-in and the reference differ on a KB8 frame-frame search, with the geometric margins."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "..", "tests")); sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import numpy as np
import world_cases as wc
from oracle import refmatch as rm
W = wc.World(seed=6, n_pts=1800, cam=wc.KB8, n_dup=200)
Ra, ta = W.pose(seed=10); Rb, tb = W.pose(seed=11)
A = W.observe(Ra, ta, seed=20)
pts = W.points(rm)
for dt in ((0, 0, 0.6), (0, 0, -0.6), (0, 0, 0)):
    B = W.observe(Rb, tb - np.array(dt), seed=23)
    pre = np.where(np.random.default_rng(4).random(len(B["kps"])) < 0.1, B["assoc"], -1).astype(np.int32)
    for assoc in (np.full(len(B["kps"]), -1, np.int32), pre):
        cur = W.view(rm, B, assoc=assoc); last = W.view(rm, A, assoc=A["assoc"], outlier=A["outlier"])
        for ori in (True, False):
            r = rm.search_by_projection_frame(cur, last, pts, 7.0, False, 0.9, ori)
            with rm.twin():
                g = rm.search_by_projection_frame(cur, last, pts, 7.0, False, 0.9, ori)
            bad = np.nonzero(r[1] != g[1])[0]
            print("dt", dt, "pre", (assoc >= 0).sum(), "ori", ori, "nm", r[0], g[0], "diff at", bad)
            for k in bad:
                print("  kp", k, B["kps"][k], "ref mp", r[1][k], "got mp", g[1][k], "assoc", assoc[k])
                for mp in (r[1][k], g[1][k]):
                    if mp < 0: continue
                    # which last-frame keypoints hold this map point
                    src = np.nonzero(A["assoc"] == mp)[0]
                    Pc = W.P[mp].astype(np.float64) @ B["Tcw"][:3, :3].astype(np.float64).T + B["Tcw"][:3, 3]
                    u, v = wc.project(W.cam, Pc[None])
                    d = np.unpackbits(B["desc"][k] ^ W.desc[mp]).sum()
                    print("    mp", mp, "src kps", src, "oct", A["kps"]["octave"][src], "proj", u, v, "ham", d, "dx,dy", B["kps"]["x"][k] - u, B["kps"]["y"][k] - v,
                          "radius", 7.0 * W.view(rm, B).scale[A["kps"]["octave"][src]])
                    # other keypoints claimed by this mp
                    print("    ref holders", np.nonzero(r[1] == mp)[0], "got holders", np.nonzero(g[1] == mp)[0])
