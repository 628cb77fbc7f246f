"""GPU parity tests of the SURVEY.md §8(f) rank-1 entry points (SearchByBoW x2, SearchForTriangulation, the Fuse /
SearchBySim3 search loops, SearchByProjection(KeyFrame*, Scw, ...)) against the CPU oracle: bit-exact indices / counts."""
import numpy as np
import pytest

from bowsynth import feature_vector, triangulation_geometry

pytestmark = pytest.mark.gpu

W, H = 752, 480


@pytest.fixture(scope="module")
def frames(oracle, synth):
    out = {}
    ex = oracle.Extractor(1200, sincos_mode=oracle.SINCOS_PORTABLE)
    for name, fid, right, jitter in (("a", 0, False, False), ("a_right", 0, True, False), ("a_jit", 0, False, True), ("b", 1, False, False)):
        img = synth.frame(W, H, 31, fid, right, jitter)
        _, k, d = ex(img, (0, 0))
        out[name] = (k, d)
    out["scale"] = ex.scale.copy()
    out["sigma2"] = (ex.scale * ex.scale).astype(np.float32)
    out["inv_sigma2"] = (np.float32(1.0) / out["sigma2"]).astype(np.float32)
    return out


@pytest.mark.parametrize("pair,kfkf,ratio,ori", [(("a", "a_jit"), False, 0.7, True), (("a", "a_jit"), True, 0.75, True),
                                                 (("a", "a_right"), False, 0.9, False), (("a", "b"), True, 0.9, True)])
def test_search_by_bow(orbx, oracle, frames, pair, kfkf, ratio, ori):
    (k1, d1), (k2, d2) = frames[pair[0]], frames[pair[1]]
    rng = np.random.default_rng(3)
    fv1, fv2 = feature_vector(d1, 96, 11), feature_vector(d2, 96, 11)
    elig1 = (rng.uniform(size=len(k1)) < 0.8).astype(np.uint8)          # KF features holding a good map point
    elig2 = (rng.uniform(size=len(k2)) < 0.85).astype(np.uint8) if kfkf else None
    max_dist = 49 if kfkf else 50
    ref_n, ref = oracle.search_by_bow(k1, d1, fv1, elig1, k2, d2, fv2, elig2, max_dist, ratio, ori)
    m = orbx.ORBmatcher(ratio, ori)
    n, got = m.SearchByBoW(k1, d1, fv1, elig1, k2, d2, fv2, elig2, max_dist)
    assert n == ref_n and np.array_equal(got, ref)
    if pair[1] != "b":
        assert ref_n > 20
    # claimed side-2 features are unique and share the vocabulary node of their side-1 feature
    mm = got[got >= 0]
    assert len(np.unique(mm)) == len(mm)
    m.close()


def test_search_by_bow_edge_cases(orbx, oracle, frames):
    (k1, d1), (k2, d2) = frames["a"], frames["a_jit"]
    m = orbx.ORBmatcher(0.7, True)
    fv1, fv2 = feature_vector(d1, 8, 1), feature_vector(d2, 8, 1)   # few, very long node lists (> 32 entries per warp pass)
    e1 = np.ones(len(k1), np.uint8)
    ref = oracle.search_by_bow(k1, d1, fv1, e1, k2, d2, fv2, None, 50, 0.7, True)
    got = m.SearchByBoW(k1, d1, fv1, e1, k2, d2, fv2, None, 50)
    assert got[0] == ref[0] and np.array_equal(got[1], ref[1])
    # disjoint vocabularies -> nothing; empty sides -> nothing
    fv2d = (fv2[0] + 100000, fv2[1], fv2[2])
    assert m.SearchByBoW(k1, d1, fv1, e1, k2, d2, fv2d, None, 50)[0] == 0
    empty = (np.zeros(0, np.uint32), np.zeros(1, np.int32), np.zeros(0, np.uint32))
    n, mm = m.SearchByBoW(k1, d1, fv1, e1, k2[:0], d2[:0], empty, None, 50)
    assert n == 0 and (mm == -1).all()
    # duplicated descriptors: ties keep the first entry of the node list, second-best equals best -> the ratio test rejects
    k2d = np.concatenate([k2, k2]); d2d = np.concatenate([d2, d2])
    fvd = feature_vector(d2d, 8, 1)
    ref = oracle.search_by_bow(k1, d1, fv1, e1, k2d, d2d, fvd, None, 50, 0.7, False)
    mo = orbx.ORBmatcher(0.7, False)
    got = mo.SearchByBoW(k1, d1, fv1, e1, k2d, d2d, fvd, None, 50)
    assert got[0] == ref[0] and np.array_equal(got[1], ref[1])
    m.close(); mo.close()


@pytest.mark.parametrize("only_stereo,coarse,ori", [(False, False, True), (True, False, True), (False, True, False)])
def test_search_for_triangulation(orbx, oracle, frames, only_stereo, coarse, ori):
    (k1, d1), (k2, d2) = frames["a"], frames["a_right"]
    rng = np.random.default_rng(4)
    fv1, fv2 = feature_vector(d1, 64, 5), feature_vector(d2, 64, 5)
    F12, ep = triangulation_geometry(W, H, 7)
    stereo1 = (rng.uniform(size=len(k1)) < 0.5).astype(np.uint8)
    stereo2 = (rng.uniform(size=len(k2)) < 0.5).astype(np.uint8)
    nomp1 = rng.uniform(size=len(k1)) < 0.7
    nomp2 = rng.uniform(size=len(k2)) < 0.7
    elig1 = (nomp1 & ((stereo1 > 0) | (not only_stereo))).astype(np.uint8)
    elig2 = (nomp2 & ((stereo2 > 0) | (not only_stereo))).astype(np.uint8)
    ref_n, ref = oracle.search_for_triangulation(k1, d1, fv1, elig1, stereo1, k2, d2, fv2, elig2, stereo2, F12, ep, frames["scale"],
                                                 frames["sigma2"], coarse, ori)
    m = orbx.ORBmatcher(0.6, ori)
    n, got = m.SearchForTriangulation(k1, d1, fv1, elig1, stereo1, k2, d2, fv2, elig2, stereo2, F12, ep, frames["scale"],
                                      frames["sigma2"], coarse)
    assert n == ref_n and np.array_equal(got, ref)
    assert ref_n > 3
    # monocular keyframes (no stereo flags at all)
    ref_n, ref = oracle.search_for_triangulation(k1, d1, fv1, nomp1.astype(np.uint8), None, k2, d2, fv2, nomp2.astype(np.uint8), None,
                                                 F12, ep, frames["scale"], frames["sigma2"], coarse, ori)
    n, got = m.SearchForTriangulation(k1, d1, fv1, nomp1.astype(np.uint8), None, k2, d2, fv2, nomp2.astype(np.uint8), None, F12, ep,
                                      frames["scale"], frames["sigma2"], coarse)
    assert n == ref_n and np.array_equal(got, ref)
    m.close()


def _projected_points(frames, synth, name, n, seed):
    """map points 'seen' from another keyframe: keypoints of `name` with sub-pixel noise + random padding"""
    k, d = frames[name]
    rng = np.random.default_rng(seed)
    pad = n - len(k)
    u = np.concatenate([k["x"] + rng.normal(0, 0.7, len(k)), rng.uniform(-20, W + 20, pad)]).astype(np.float32)
    v = np.concatenate([k["y"] + rng.normal(0, 0.7, len(k)), rng.uniform(-20, H + 20, pad)]).astype(np.float32)
    level = np.concatenate([np.clip(k["octave"] + rng.integers(0, 2, len(k)), 0, 7), rng.integers(0, 8, pad)]).astype(np.int32)
    desc = np.concatenate([d, synth.random_descriptors(pad, seed)])
    valid = (rng.uniform(size=n) < 0.93).astype(np.uint8)
    return valid, u, v, level, desc


@pytest.mark.parametrize("gate,stereo,th", [(False, False, 3.0), (True, False, 3.0), (True, True, 4.0), (False, True, 7.5)])
def test_search_window_best_fuse_loops(orbx, oracle, synth, frames, gate, stereo, th):
    k, d = frames["a"]
    rng = np.random.default_rng(6)
    n = len(frames["a_jit"][0]) + 800
    valid, u, v, level, desc = _projected_points(frames, synth, "a_jit", n, 8)
    radius = (np.float32(th) * frames["scale"][level]).astype(np.float32)
    uR = None; ur = None
    if stereo:
        uR = np.where(rng.uniform(size=len(k)) < 0.6, k["x"] - rng.uniform(0, 30, len(k)), -1).astype(np.float32)
        ur = (u - rng.uniform(0, 30, n)).astype(np.float32)
    ref_i, ref_d = oracle.search_window_best(k, d, W, H, valid, u, v, radius, level, desc, int(gate), frames["inv_sigma2"], uR, ur)
    F = orbx.FrameView(k, d, frames["scale"], W, H, u_right=uR)
    m = orbx.ORBmatcher()
    bi, bd = m.SearchWindowBest(F, valid, u, v, radius, level, desc, gate, frames["inv_sigma2"], ur)
    assert np.array_equal(bi, ref_i) and np.array_equal(bd, ref_d)
    assert (ref_d <= 50).sum() > 200
    if gate:  # the chi-square gate must actually remove candidates that the ungated search accepts
        ui, ud = oracle.search_window_best(k, d, W, H, valid, u, v, radius, level, desc, 0, None, uR, ur)
        assert (ui != ref_i).any()
    m.close()


def test_search_by_sim3(orbx, oracle, synth, frames):
    (k1, d1), (k2, d2) = frames["a"], frames["a_jit"]
    n1, n2 = len(k1), len(k2)
    rng = np.random.default_rng(9)
    # pass 1: map point of keypoint i1 of KF1 projected into KF2 (near its true match when there is one); pass 2 likewise
    def proj(ksrc, seed):
        r = np.random.default_rng(seed)
        u = (ksrc["x"] + r.normal(0, 1.0, len(ksrc))).astype(np.float32)
        v = (ksrc["y"] + r.normal(0, 1.0, len(ksrc))).astype(np.float32)
        lvl = np.clip(ksrc["octave"] + r.integers(0, 2, len(ksrc)), 0, 7).astype(np.int32)
        valid = (r.uniform(size=len(ksrc)) < 0.8).astype(np.uint8)   # has a good map point, not already matched, in image
        return valid, u, v, lvl
    va1, u1, v1, l1 = proj(k1, 1); va2, u2, v2, l2 = proj(k2, 2)
    r1 = (np.float32(7.5) * frames["scale"][l1]).astype(np.float32); r2 = (np.float32(7.5) * frames["scale"][l2]).astype(np.float32)
    F1 = orbx.FrameView(k1, d1, frames["scale"], W, H); F2 = orbx.FrameView(k2, d2, frames["scale"], W, H)
    m = orbx.ORBmatcher()
    nfound, m12 = m.SearchBySim3Projected(F1, F2, va1, u1, v1, r1, l1, d1, va2, u2, v2, r2, l2, d2)
    # oracle: the reference's two loops + agreement (ORBmatcher.cc:1866-1999)
    b1, dd1 = oracle.search_window_best(k2, d2, W, H, va1, u1, v1, r1, l1, d1)
    b2, dd2 = oracle.search_window_best(k1, d1, W, H, va2, u2, v2, r2, l2, d2)
    vn1 = np.where(dd1 <= 100, b1, -1); vn2 = np.where(dd2 <= 100, b2, -1)
    ref = np.full(n1, -1, np.int32)
    for i1 in range(n1):
        if vn1[i1] >= 0 and vn2[vn1[i1]] == i1:
            ref[i1] = vn1[i1]
    assert np.array_equal(m12, ref) and nfound == int((ref >= 0).sum())
    assert nfound > 300
    m.close()


@pytest.mark.parametrize("ratio_hamming,th", [(1.0, 3), (0.8, 5)])
def test_search_by_projection_keyframe_scw(orbx, oracle, synth, frames, ratio_hamming, th):
    """SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th, ratioHamming), ORBmatcher.cc:489-602, after projection"""
    k, d = frames["a"]
    n = len(frames["a_jit"][0]) + 500
    valid, u, v, level, desc = _projected_points(frames, synth, "a_jit", n, 12)
    rng = np.random.default_rng(13)
    kp0 = np.full(len(k), -1, np.int32)
    kp0[rng.integers(0, len(k), 150)] = -2          # vpMatched[idx] already set
    dist_th = int(np.floor(np.float32(50) * np.float32(ratio_hamming)))
    ref_n, ref = oracle.search_by_projection_window(k, d, frames["scale"], W, H, valid, u, v, level, np.zeros(n, np.float32), desc,
                                                    float(th), 3, dist_th, True, False, kp0)
    F = orbx.FrameView(k, d, frames["scale"], W, H)
    m = orbx.ORBmatcher(0.6, False)
    got_n, got = m.SearchByProjectionWindow(F, valid, u, v, level, np.zeros(n, np.float32), desc, float(th), 3, dist_th, True, kp0)
    assert got_n == ref_n and np.array_equal(got, ref)
    assert ref_n > 200
    m.close()


def test_parallel_resolve_on_long_claim_chains(orbx, oracle, synth, frames):
    """The resolve kernels iterate a parallel fixed point instead of replaying the reference's loop (match_kernels.cuh,
    k_resolve_par / k_bow_resolve_par).  Worst case for that scheme: many queries that compete for the same keypoints, so that
    every claim depends on the one before (map points / source points / keyframe features repeated up to 12 times)."""
    k, d = frames["a"]
    kj, dj = frames["a_jit"]
    rng = np.random.default_rng(21)
    rep = rng.integers(1, 13, len(kj))
    idx = np.repeat(np.arange(len(kj)), rep)
    rng.shuffle(idx)
    n = len(idx)
    u = (kj["x"][idx] + rng.normal(0, 0.3, n)).astype(np.float32); v = (kj["y"][idx] + rng.normal(0, 0.3, n)).astype(np.float32)
    lvl = kj["octave"][idx].astype(np.int32); desc = dj[idx]
    observed = (rng.uniform(size=n) < 0.7).astype(np.uint8)
    F = orbx.FrameView(k, d, frames["scale"], W, H)
    # SearchByProjection(F, map points): ratio 1.0 so that most repeated points pass the ratio test and claim in turn
    m = orbx.ORBmatcher(1.0, True)
    ref = oracle.search_by_projection_mappoints(k, d, frames["scale"], W, H, np.ones(n, np.uint8), u, v, lvl, np.ones(n, np.float32), desc,
                                                3.0, 1.0, observed=observed)
    got = m.SearchByProjection(F, np.ones(n, np.uint8), u, v, lvl, np.ones(n, np.float32), desc, 3.0, observed=observed)
    assert got[0] == ref[0] and np.array_equal(got[1], ref[1]) and ref[0] > 800
    # window search (Frame / Frame rule and the any-map-point-blocks rule of Frame / KeyFrame)
    ang = kj["angle"][idx]
    for any_blocks in (False, True):
        ref = oracle.search_by_projection_window(k, d, frames["scale"], W, H, np.ones(n, np.uint8), u, v, lvl, ang, desc, 7.0, 0, 100,
                                                 any_blocks, True, observed=observed)
        got = m.SearchByProjectionWindow(F, np.ones(n, np.uint8), u, v, lvl, ang, desc, 7.0, 0, 100, any_blocks, observed=observed)
        assert got[0] == ref[0] and np.array_equal(got[1], ref[1]) and ref[0] > 300
    # SearchByBoW: the keyframe side holds every feature several times, all in the same few nodes
    k1, d1 = kj[idx], dj[idx]
    fv1, fv2 = feature_vector(d1, 6, 2), feature_vector(d, 6, 2)
    ref = oracle.search_by_bow(k1, d1, fv1, np.ones(n, np.uint8), k, d, fv2, None, 50, 1.0, True)
    got = m.SearchByBoW(k1, d1, fv1, np.ones(n, np.uint8), k, d, fv2, None, 50)
    assert got[0] == ref[0] and np.array_equal(got[1], ref[1]) and ref[0] > 100
    m.close()


@pytest.mark.parametrize("th,all_observed,pre_share", [(1.0, True, 0.0), (4.0, True, 0.2), (3.0, False, 0.2), (5.0, False, 0.0)])
def test_search_by_projection_two_cameras_vs_oracle(orbx, oracle, th, all_observed, pre_share):
    """orbx_search_by_projection_mappoints_stereo through the C-ABI against the CPU restatement (ORBmatcher.cc:44-211 on a fisheye-stereo
    frame; the restatement is pinned to the reference's compiled source in test_oracle_refmatch.py): every map point observed -> the cluster
    fixed point k_resolve_fisheye_par, otherwise the one-warp replay k_resolve_fisheye.  Tracking fields are synthetic here (no _ref needed):
    projections of the keypoints' own map points plus jitter, so windows overlap and claims collide."""
    import world_cases as wc
    W = wc.World(seed=21, n_pts=1800, cam=wc.KB8, n_dup=0)
    rng = np.random.default_rng(31)
    n_obs = np.maximum(W.n_obs, 1) if all_observed else W.n_obs
    R, t = W.pose(seed=4)
    obs = W.observe_two_cameras(R, t, seed=9)
    nl = obs["n_left"]
    kl, kr = obs["kps"], obs["kps_right"]
    w, h = W.cam["size"]
    sf = np.float32(1.2) ** np.arange(8, dtype=np.float32)

    def track(kps, mp):
        inv = np.zeros(W.n, np.uint8); px = np.zeros(W.n, np.float32); py = np.zeros(W.n, np.float32)
        lv = np.full(W.n, -1, np.int32); vc = np.full(W.n, 0.9, np.float32)
        sel = mp >= 0
        idx = mp[sel]
        inv[idx] = 1
        px[idx] = kps["x"][sel] + rng.normal(size=sel.sum()).astype(np.float32) * 2.0
        py[idx] = kps["y"][sel] + rng.normal(size=sel.sum()).astype(np.float32) * 2.0
        lv[idx] = np.clip(kps["octave"][sel] + rng.integers(0, 2, sel.sum()), 0, 7)
        vc[idx] = np.where(rng.random(sel.sum()) < 0.5, 0.999, 0.9).astype(np.float32)
        return dict(in_view=inv, proj_x=px, proj_y=py, level=lv, view_cos=vc)
    left, right = track(kl, obs["mp_idx"][:nl]), track(kr, obs["mp_idx"][nl:])
    right["level"][rng.random(W.n) < 0.05] = -1          # mnTrackScaleLevelR == -1: no right pass (:146)
    pre = np.where((rng.random(len(obs["assoc"])) < pre_share) & (obs["assoc"] >= 0), obs["assoc"], -1)
    code = np.where(pre < 0, -1, np.where(n_obs[np.maximum(pre, 0)] > 0, -2, -3)).astype(np.int32)
    observed = (n_obs > 0).astype(np.uint8)
    nm_o, kp_o = oracle.search_by_projection_mappoints_stereo(kl, obs["desc"][:nl], kr, obs["desc"][nl:], sf, w, h, left, right, W.desc,
                                                              obs["l2r"], obs["r2l"], th, 0.8, observed=observed, kp_mp=code)
    m = orbx.ORBmatcher(0.8)
    FL, FR = orbx.FrameView(kl, obs["desc"][:nl], sf, w, h), orbx.FrameView(kr, obs["desc"][nl:], sf, w, h)
    nm_g, kp_g = m.SearchByProjectionTwoCameras(FL, FR, left, right, W.desc, obs["l2r"], obs["r2l"], th, observed=observed, kp_mp=code)
    m.close()
    assert nm_g == nm_o and nm_o > 600
    assert np.array_equal(kp_g, kp_o)
    assert (kp_o[:nl] >= 0).sum() > 200 and (kp_o[nl:] >= 0).sum() > 200


@pytest.mark.parametrize("ori,ratio", [(True, 0.7), (False, 0.7), (True, 0.9)])
def test_search_by_bow_two_cameras_vs_oracle(orbx, oracle, ori, ratio):
    """orbx_search_by_bow_stereo through the C-ABI against the CPU restatement of ORBmatcher.cc:338-431 (pinned to the reference's compiled
    source in test_oracle_refmatch.py): left / right matching of a keyframe feature inside the cluster fixed point, shared rotation filter"""
    import world_cases as wc
    W = wc.World(seed=33, n_pts=1800, cam=wc.KB8, n_dup=0)
    R, t = W.pose(seed=5)
    obs = W.observe_two_cameras(R, t, seed=12)
    R2, t2 = W.pose(seed=6)
    ko = W.observe(R2, t2, seed=14)
    fvF = feature_vector(obs["desc"], 24, seed=2)
    fvK = feature_vector(ko["desc"], 24, seed=2)
    kps2 = np.concatenate([obs["kps"], obs["kps_right"]])
    elig = ((ko["assoc"] >= 0) & (W.bad[np.maximum(ko["assoc"], 0)] == 0)).astype(np.uint8)
    nm_o, mL_o, mR_o = oracle.search_by_bow_stereo(ko["kps"], ko["desc"], fvK, elig, kps2, obs["desc"], fvF, obs["n_left"], 50, ratio, ori)
    m = orbx.ORBmatcher(ratio, ori)
    nm_g, mL_g, mR_g = m.SearchByBoWTwoCameras(ko["kps"], ko["desc"], fvK, elig, kps2, obs["desc"], fvF, obs["n_left"])
    m.close()
    assert nm_g == nm_o and nm_o > 300
    assert np.array_equal(mL_g, mL_o) and np.array_equal(mR_g, mR_o)
    assert (mL_o >= 0).sum() > 100 and (mR_o >= 0).sum() > 100 and (mR_o[mR_o >= 0] >= obs["n_left"]).all()


# ---- KannalaBrandt8::epipolarConstrain on the device (the gate of SearchForTriangulation on fisheye keyframes) -----------------------
def _kb8_cam():
    import world_cases as wc
    return np.array(list(wc.KB8["K"]) + list(wc.KB8["kb"]), np.float32)


def _kb8_world(seed, dt=0.25, two=False):
    import world_cases as wc
    Wd = wc.World(seed=seed, n_pts=1600, cam=wc.KB8, n_dup=0)
    Ra, ta = Wd.pose(seed=30, dt=dt); Rb, tb = Wd.pose(seed=31, dt=dt)
    if two:
        return Wd, Wd.observe_two_cameras(Ra, ta, seed=30), Wd.observe_two_cameras(Rb, tb, seed=50)
    return Wd, Wd.observe(Ra, ta, seed=40), Wd.observe(Rb, tb, seed=41)


@pytest.mark.parametrize("seed,dt", [(16, 0.25), (17, 0.6), (18, 0.05)])
def test_kb8_triangulate_matches_vs_oracle(orbx, oracle, seed, dt):
    """orbx_kb8_triangulate_matches (k_kb8_triangulate_pairs: unproject, parallax, the 4x4 system through the Jacobi SVD, depths, both
    reprojections) against the CPU restatement in its portable libm mode -- itself equal, in host-libm mode, to the reference's compiled
    KannalaBrandt8::TriangulateMatches (tests/test_oracle_refmatch.py).  Depth and point bit for bit; CUDA's and glibc's double tan /
    atan2 / cos / sin / hypot may differ in the last place, which after the rounding to float is allowed to show in a handful of values."""
    from tri_cases import triangulation_geometry_kb8
    Wd, A, B = _kb8_world(seed, dt)
    pairs, _ = triangulation_geometry_kb8(oracle, A["Tcw"], B["Tcw"], _kb8_cam())
    ia = {int(mp): i for i, mp in enumerate(A["mp_idx"]) if mp >= 0}
    tr = [(ia[int(mp)], j) for j, mp in enumerate(B["mp_idx"]) if mp >= 0 and int(mp) in ia]
    rng = np.random.default_rng(seed)
    rd = list(zip(rng.integers(0, len(A["kps"]), 3000), rng.integers(0, len(B["kps"]), 3000)))
    p = np.array(tr + rd)
    xy1 = np.stack([A["kps"]["x"][p[:, 0]], A["kps"]["y"][p[:, 0]]], 1); xy2 = np.stack([B["kps"]["x"][p[:, 1]], B["kps"]["y"][p[:, 1]]], 1)
    s2 = (np.float32(1.2) ** np.arange(8, dtype=np.float32)) ** 2
    sg1, sg2 = s2[A["kps"]["octave"][p[:, 0]]], s2[B["kps"]["octave"][p[:, 1]]]
    z_o, p_o = oracle.kb8_triangulate_matches(pairs[0], xy1, xy2, sg1, sg2, oracle.LIBM_PORTABLE)
    m = orbx.ORBmatcher(0.6)
    z_g, p_g = m.TriangulateMatchesKB8(pairs[0], xy1, xy2, sg1, sg2)
    m.close()
    assert np.array_equal(z_g > 0.0001, z_o > 0.0001)
    assert (z_g.view(np.uint32) != z_o.view(np.uint32)).sum() <= 3 and (p_g.view(np.uint32) != p_o.view(np.uint32)).sum() <= 9
    acc = z_o > 0.0001
    assert acc[:len(tr)].sum() > (100 if dt >= 0.25 else 0) and (~acc[:len(tr)]).sum() > 5 and acc[len(tr):].mean() < 0.2


@pytest.mark.parametrize("only_stereo,coarse,ori", [(False, False, True), (False, True, True), (False, False, False), (True, False, True)])
def test_search_for_triangulation_kb8_vs_oracle(orbx, oracle, only_stereo, coarse, ori):
    """orbx_search_for_triangulation_kb8 (k_bow_triangulate_kb8) on monocular-fisheye keyframes against the CPU restatement, which
    tests/test_oracle_refmatch.py pins to the reference's compiled SearchForTriangulation: same pairs, same count"""
    from tri_cases import triangulation_geometry_kb8
    Wd, A, B = _kb8_world(16)
    fva, fvb = feature_vector(A["desc"], 24, 1), feature_vector(B["desc"], 24, 1)
    rng = np.random.default_rng(1)
    aa = np.where(rng.random(len(A["assoc"])) < 0.25, A["assoc"], -1); bb = np.where(rng.random(len(B["assoc"])) < 0.25, B["assoc"], -1)
    pairs, ep = triangulation_geometry_kb8(oracle, A["Tcw"], B["Tcw"], _kb8_cam())
    sf = np.float32(1.2) ** np.arange(8, dtype=np.float32)
    s1, s2 = (A["u_right"] >= 0).astype(np.uint8), (B["u_right"] >= 0).astype(np.uint8)
    if only_stereo:     # the reference's flag on a camera without stereo measurements: nothing is eligible
        e1 = ((aa < 0) & (s1 > 0)).astype(np.uint8); e2 = ((bb < 0) & (s2 > 0)).astype(np.uint8)
    else:
        e1 = (aa < 0).astype(np.uint8); e2 = (bb < 0).astype(np.uint8)
    ref_n, ref = oracle.search_for_triangulation_kb8(A["kps"], A["desc"], fva, e1, s1, -1, B["kps"], B["desc"], fvb, e2, s2, -1, pairs, ep,
                                                     sf, sf * sf, sf * sf, coarse, ori, oracle.LIBM_PORTABLE)
    m = orbx.ORBmatcher(0.6, ori)
    n, got = m.SearchForTriangulationKB8(A["kps"], A["desc"], fva, e1, s1, -1, B["kps"], B["desc"], fvb, e2, s2, -1, pairs, ep, sf, sf * sf,
                                         sf * sf, coarse)
    m.close()
    assert n == ref_n and np.array_equal(got, ref)
    assert ref_n > (100 if not only_stereo else -1)


@pytest.mark.parametrize("coarse,ori", [(False, True), (False, False), (True, True)])
def test_search_for_triangulation_two_cameras_vs_oracle(orbx, oracle, coarse, ori):
    """the same on fisheye-stereo keyframes: features of both cameras in one list, the pair (left|right, left|right) selects one of four
    relative poses (src/ORBmatcher.cc:1103-1140), no epipole gate"""
    from tri_cases import triangulation_geometry_kb8
    Wd, A, B = _kb8_world(8, two=True)
    fva, fvb = feature_vector(A["desc"], 30, 1), feature_vector(B["desc"], 30, 1)
    rng = np.random.default_rng(2)
    aa = np.where(rng.random(len(A["assoc"])) < 0.25, A["assoc"], -1); bb = np.where(rng.random(len(B["assoc"])) < 0.25, B["assoc"], -1)
    pairs, ep = triangulation_geometry_kb8(oracle, A["Tcw"], B["Tcw"], _kb8_cam(), Tlr=A["Tlr"], Tlr2=B["Tlr"])
    k1, k2 = np.concatenate([A["kps"], A["kps_right"]]), np.concatenate([B["kps"], B["kps_right"]])
    sf = np.float32(1.2) ** np.arange(8, dtype=np.float32)
    e1, e2 = (aa < 0).astype(np.uint8), (bb < 0).astype(np.uint8)
    ref_n, ref = oracle.search_for_triangulation_kb8(k1, A["desc"], fva, e1, None, A["n_left"], k2, B["desc"], fvb, e2, None, B["n_left"], pairs,
                                                     ep, sf, sf * sf, sf * sf, coarse, ori, oracle.LIBM_PORTABLE)
    m = orbx.ORBmatcher(0.6, ori)
    n, got = m.SearchForTriangulationKB8(k1, A["desc"], fva, e1, None, A["n_left"], k2, B["desc"], fvb, e2, None, B["n_left"], pairs, ep, sf,
                                         sf * sf, sf * sf, coarse)
    assert n == ref_n and np.array_equal(got, ref) and ref_n > 150
    r1, r2 = np.nonzero(got >= 0)[0] >= A["n_left"], got[got >= 0] >= B["n_left"]
    assert all(((r1 == a) & (r2 == b)).sum() > 5 for a in (False, True) for b in (False, True))
    # argument checks: a rig on one side only, a wrong record count
    import ctypes
    with pytest.raises(Exception):
        m.SearchForTriangulationKB8(k1, A["desc"], fva, e1, None, A["n_left"], k2, B["desc"], fvb, e2, None, -1, pairs, ep, sf, sf * sf, sf * sf)
    with pytest.raises(Exception):
        m.SearchForTriangulationKB8(k1, A["desc"], fva, e1, None, A["n_left"], k2, B["desc"], fvb, e2, None, B["n_left"], pairs[:1], ep, sf,
                                    sf * sf, sf * sf)
    m.close()
