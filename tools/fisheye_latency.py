#!/usr/bin/env python3
"""Latency of orbx_search_by_projection_mappoints_stereo (SearchByProjection(Frame, MapPoints) on a fisheye-stereo frame, host arrays in /
out) with the two resolve kernels: every map point observed -> k_resolve_fisheye_par (cluster fixed point); one participating point
without observations -> k_resolve_fisheye (one-warp replay of the reference's order).  Same world, results compared where they must agree."""
import importlib, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import world_cases as wc
from oracle import refmatch as rm
orbx = importlib.import_module("3_orb_slam3_selfnote_b200")
out = {}
for n_pts in (2200, 6000):
    W = wc.World(seed=8, n_pts=n_pts, cam=wc.KB8, n_dup=300)
    R, t = W.pose(seed=12)
    obs = W.observe_two_cameras(R, t, seed=30)
    view = W.view_two_cameras(rm, obs)
    tr = rm.is_in_frustum(view, W.points(rm), 0.5, right=True)
    left = dict(in_view=tr["in_view"], proj_x=tr["proj_x"], proj_y=tr["proj_y"], level=tr["level"], view_cos=tr["view_cos"])
    right = dict(in_view=tr["in_view_r"], proj_x=tr["proj_xr_r"], proj_y=tr["proj_y_r"], level=tr["level_r"], view_cos=tr["view_cos_r"])
    nl = obs["n_left"]
    sf = view.scale
    w, h = W.cam["size"]
    FL = orbx.FrameView(obs["kps"], obs["desc"][:nl], sf, w, h)
    FR = orbx.FrameView(obs["kps_right"], obs["desc"][nl:], sf, w, h)
    m = orbx.ORBmatcher(0.8)
    res = {}
    for name in ("all_observed_parallel", "one_unobserved_sequential"):
        ob = np.ones(W.n, np.uint8)
        if name.startswith("one"):
            ob[np.nonzero(tr["in_view"])[0][-1]] = 0
        for _ in range(3):
            r = m.SearchByProjectionTwoCameras(FL, FR, left, right, W.desc, obs["l2r"], obs["r2l"], 3.0, observed=ob)
        ts = []
        for _ in range(30):
            t0 = time.perf_counter(); r = m.SearchByProjectionTwoCameras(FL, FR, left, right, W.desc, obs["l2r"], obs["r2l"], 3.0, observed=ob); ts.append(time.perf_counter() - t0)
        res[name] = {"ms_median": round(float(np.median(ts) * 1e3), 3), "nmatches": int(r[0])}
    res["map_points_in_view"] = [int(tr["in_view"].sum()), int(tr["in_view_r"].sum())]
    res["keypoints"] = [int(nl), int(len(obs["kps_right"]))]
    out["map_points_%d" % n_pts] = res
    m.close()
print(json.dumps(out, indent=1))
