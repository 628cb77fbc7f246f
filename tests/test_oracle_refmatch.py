"""CPU: the oracle restatement of the ORBmatcher / Frame searches against the reference's OWN compiled source
(oracle/_ref/libref_match.so = src/ORBmatcher.cc + the Frame / KeyFrame / MapPoint / camera member bodies, unmodified; see
oracle/Makefile) on synthetic worlds.  For the searches that start from poses and 3-D points, the per-point prologue is the numpy
restatement of tests/prologue_np.py (what k_project_points computes on the device) in front of the oracle's search loops."""
import numpy as np
import pytest

import bowsynth
import prologue_np as pro
import world_cases as wc
from oracle import refmatch as rm

pytestmark = pytest.mark.skipif(not rm.available(), reason="oracle/_ref/libref_match.so not built (needs /root/reference)")


@pytest.fixture(scope="module", params=["euroc", "kb8"])
def scene(request):
    cam = wc.EUROC if request.param == "euroc" else wc.KB8
    W = wc.World(seed=15 if request.param == "euroc" else 16, n_pts=1500, cam=cam, n_dup=150)
    Ra, ta = W.pose(seed=30)
    Rb, tb = W.pose(seed=31)
    A, B = W.observe(Ra, ta, seed=40), W.observe(Rb, tb, seed=41)
    return dict(W=W, A=A, B=B, pts=W.points(rm), Ra=Ra, ta=ta, Rb=Rb, tb=tb, name=request.param)


def _pose32(obs):
    T = obs["Tcw"]
    return T[:3, :3].copy(), T[:3, 3].copy()


def _finish(kp_mp, src_to_mp, pre):
    """oracle kp_mp (source index / -1 / untouched -2,-3) -> map-point index per keypoint as the reference leaves mvpMapPoints"""
    return np.where(kp_mp >= 0, src_to_mp[np.clip(kp_mp, 0, None)], np.where(kp_mp == -1, -1, pre)).astype(np.int32)


@pytest.mark.parametrize("mono,th,dz,ori", [(True, 7.0, 0.0, True), (False, 15.0, 0.0, True), (False, 7.0, 0.6, True), (False, 7.0, -0.6, False)])
def test_search_by_projection_frame_frame(oracle, scene, mono, th, dz, ori):
    W, A, B = scene["W"], scene["A"], scene["B"]
    if dz:
        B = W.observe(scene["Rb"], scene["tb"] - np.array([0, 0, dz]), seed=43)
    cam = W.cam
    pre = np.where(np.random.default_rng(4).random(len(B["kps"])) < 0.1, B["assoc"], -1).astype(np.int32)
    cur = W.view(rm, B, assoc=pre)
    last = W.view(rm, A, assoc=A["assoc"], outlier=A["outlier"])
    nm_ref, kp_ref = rm.search_by_projection_frame(cur, last, scene["pts"], th, mono, 0.9, ori)
    # restatement: src/ORBmatcher.cc:2038-2118 then the window search
    Rcw, tcw = _pose32(B); Rlw, tlw = _pose32(A)
    twc = pro.cam_center(Rcw, tcw)
    tlc = pro.gemv_add(Rlw, twc[None], tlw)
    mb = np.float32(cam.get("bf", 0.0) / cam["K"][0]) if cam.get("bf", 0.0) else np.float32(0)
    mb = np.float32(np.float32(cam.get("bf", 0.0)) / np.float32(cam["K"][0]))
    fwd = bool(tlc[2][0] > mb) and not mono
    bwd = bool(-tlc[2][0] > mb) and not mono
    mp = A["assoc"]
    valid = (mp >= 0) & (A["outlier"] == 0)
    P = W.P[np.clip(mp, 0, None)]
    pj = pro.project(P, valid, Rcw, tcw, twc, cam["K"], W.bounds(), kb=cam.get("kb"), depth_mode=0, invz_double=True, bounds_mode=0,
                     bf=cam.get("bf", 0.0), level_mode=0, octave=A["kps"]["octave"], scale_factors=cur.scale, th=th)
    kp0 = np.where(pre < 0, -1, np.where(W.n_obs[np.clip(pre, 0, None)] > 0, -2, -3)).astype(np.int32)
    nm, kp = oracle.search_by_projection_window(B["kps"], B["desc"], cur.scale, *cam["size"], pj["valid"], pj["u"], pj["v"], pj["level"],
                                                A["kps"]["angle"], W.desc[np.clip(mp, 0, None)], th, mode=1 if fwd else (2 if bwd else 0),
                                                dist_th=100, any_mp_blocks=False, check_orientation=ori, kp_mp=kp0,
                                                observed=W.n_obs[np.clip(mp, 0, None)] > 0,   # :2134-2136: a point without observations does not hold its keypoint
                                                uRight=B["u_right"] if cam.get("bf", 0.0) > 0 else None, ur=pj["ur"])
    assert nm == nm_ref and nm > 50
    assert np.array_equal(_finish(kp, mp, pre), kp_ref)


@pytest.mark.parametrize("th,orb_dist,ori", [(10.0, 100, True), (3.0, 64, False)])
def test_search_by_projection_frame_keyframe(oracle, scene, th, orb_dist, ori):
    W, A, B = scene["W"], scene["A"], scene["B"]
    cam = W.cam
    found = (np.random.default_rng(7).random(W.n) < 0.15).astype(np.uint8)
    pre = np.where(np.random.default_rng(8).random(len(B["kps"])) < 0.2, B["assoc"], -1).astype(np.int32)
    cur = W.view(rm, B, assoc=pre)
    kf = W.view(rm, A, assoc=A["assoc"])
    nm_ref, kp_ref = rm.search_by_projection_keyframe(cur, kf, scene["pts"], found, th, orb_dist, 0.9, ori)
    Rcw, tcw = _pose32(B)
    Ow = pro.cam_center(Rcw, tcw)
    mp = A["assoc"]
    mpc = np.clip(mp, 0, None)
    valid = (mp >= 0) & (W.bad[mpc] == 0) & (found[mpc] == 0)
    pj = pro.project(W.P[mpc], valid, Rcw, tcw, Ow, cam["K"], W.bounds(), kb=cam.get("kb"), depth_mode=2, bounds_mode=0,
                     max_d=W.max_dist[mpc], min_d=W.min_dist[mpc], level_mode=1, log_scale=np.log(np.float32(1.2)), scale_factors=cur.scale, th=th)
    kp0 = np.where(pre < 0, -1, -2).astype(np.int32)
    nm, kp = oracle.search_by_projection_window(B["kps"], B["desc"], cur.scale, *cam["size"], pj["valid"], pj["u"], pj["v"], pj["level"],
                                                A["kps"]["angle"], W.desc[mpc], th, mode=0, dist_th=orb_dist, any_mp_blocks=True,
                                                check_orientation=ori, kp_mp=kp0)
    assert nm == nm_ref and nm > 50
    assert np.array_equal(_finish(kp, mp, pre), kp_ref)


def _scw(B, s):
    Scw = B["Tcw"].copy()
    Scw[:3, :] *= np.float32(s)
    sR = Scw[:3, :3]
    scw = np.float32(np.sqrt(np.float64(sR[0].astype(np.float64) @ sR[0].astype(np.float64))))   # sqrt(row.dot(row)): dot in double, sqrt(double) -> float
    R = (sR / scw).astype(np.float32)
    t = (Scw[:3, 3] / scw).astype(np.float32)
    return Scw, R, t


@pytest.mark.parametrize("th,ratio,with_kfs,s", [(4, 1.0, False, 1.07), (8, 0.9, True, 0.93)])
def test_search_by_projection_scw(oracle, scene, th, ratio, with_kfs, s):
    W, B = scene["W"], scene["B"]
    cam = W.cam
    Scw, R, t = _scw(B, s)
    kf = W.view(rm, B, assoc=B["assoc"])
    matched = np.where(np.random.default_rng(1).random(len(B["kps"])) < 0.2, B["assoc"], -1).astype(np.int32)
    out = rm.search_by_projection_scw(kf, Scw, scene["pts"], th, ratio, matched, with_kfs)
    nm_ref, m_ref = out[0], out[1]
    Ow = pro.cam_center(R, t)
    already = np.zeros(W.n, bool)
    already[matched[matched >= 0]] = True
    valid = (W.bad == 0) & ~already
    w, h = cam["size"]
    pj = pro.project(W.P, valid, R, t, Ow, cam["K"], (0, int(w), 0, int(h)), kb=None if with_kfs else cam.get("kb"), depth_mode=1,
                     invz_double=False, bounds_mode=1, max_d=W.max_dist, min_d=W.min_dist, normal=W.normal, level_mode=1,
                     log_scale=np.log(np.float32(1.2)), scale_factors=kf.scale, th=float(th), pinhole_formula=with_kfs)
    kp0 = np.where(matched < 0, -1, -2).astype(np.int32)
    nm, kp = oracle.search_by_projection_window(B["kps"], B["desc"], kf.scale, w, h, pj["valid"], pj["u"], pj["v"], pj["level"],
                                                np.zeros(W.n, np.float32), W.desc, float(th), mode=3, dist_th=int(np.floor(50 * ratio)),
                                                any_mp_blocks=True, check_orientation=False, kp_mp=kp0)
    assert nm == nm_ref and nm > 20
    assert np.array_equal(_finish(kp, np.arange(W.n, dtype=np.int32), matched), m_ref)
    if with_kfs:
        assert np.array_equal(out[2] >= 0, (m_ref >= 0) & (matched < 0))


@pytest.mark.parametrize("th", [3.0, 6.5])
def test_fuse(oracle, scene, th):
    """src/ORBmatcher.cc:1425-1662: the gated best-match search + the Replace / AddObservation bookkeeping in candidate order"""
    W, B = scene["W"], scene["B"]
    cam = W.cam
    kf = W.view(rm, B, assoc=B["assoc"])
    cand = np.random.default_rng(3).permutation(W.n).astype(np.int32)
    nf_ref, ev_ref = rm.fuse(kf, scene["pts"], cand, th)
    R, t = _pose32(B)
    Ow = pro.cam_center(R, t)
    in_kf = np.zeros(W.n, bool)
    in_kf[B["assoc"][B["assoc"] >= 0]] = True
    valid = (W.bad[cand] == 0) & ~in_kf[cand]
    w, h = cam["size"]
    pj = pro.project(W.P[cand], valid, R, t, Ow, cam["K"], (0, int(w), 0, int(h)), kb=cam.get("kb"), depth_mode=1, invz_double=False,
                     bounds_mode=1, bf=cam.get("bf", 0.0), max_d=W.max_dist[cand], min_d=W.min_dist[cand], normal=W.normal[cand], level_mode=1,
                     log_scale=np.log(np.float32(1.2)), scale_factors=kf.scale, th=th)
    _, _, _, inv_s2 = rm.scale_tables()
    bi, bd = oracle.search_window_best(B["kps"], B["desc"], w, h, pj["valid"], pj["u"], pj["v"], pj["radius"], pj["level"], W.desc[cand],
                                       gate=1, inv_level_sigma2=inv_s2, uRight=B["u_right"], ur=pj["ur"])
    events, nfused = [], 0
    holder = B["assoc"].copy()
    for i, p in enumerate(cand):
        if not pj["valid"][i] or bi[i] < 0 or bd[i] > 50:
            continue
        other = holder[bi[i]]
        if other >= 0:
            if not W.bad[other]:
                events.append((0, p, other, -1) if W.n_obs[other] > W.n_obs[p] else (0, other, p, -1))
        else:
            events.append((1, p, -1, bi[i]))
            holder[bi[i]] = p
        nfused += 1
    assert nfused == nf_ref and nfused > 50
    assert np.array_equal(np.array(events, np.int32).reshape(-1, 4), ev_ref)


def test_search_for_initialization_and_bow(oracle, scene):
    """the searches without a projection prologue: SearchForInitialization (:722-837), both SearchByBoW (:273-475, :839-979)"""
    W, A, B = scene["W"], scene["A"], scene["B"]
    w, h = W.cam["size"]
    v1, v2 = W.view(rm, A, with_pose=False), W.view(rm, B, with_pose=False)
    lvl0 = A["kps"]["octave"] == 0
    prev = np.stack([A["kps"]["x"], A["kps"]["y"]], 1).astype(np.float32)
    nm_ref, m_ref, pm_ref = rm.search_for_initialization(v1, v2, prev, 60, 0.9, True)
    nm, m12, pm = oracle.search_for_initialization(A["kps"], A["desc"], B["kps"], B["desc"], w, h, prev, 60, 0.9, True)
    assert nm == nm_ref and np.array_equal(m12, m_ref) and np.array_equal(pm, pm_ref) and nm > 10 and lvl0.any()
    fva = bowsynth.feature_vector(A["desc"], 24, seed=1)
    fvb = bowsynth.feature_vector(B["desc"], 24, seed=1)
    kfa = W.view(rm, A, assoc=A["assoc"], feat_vec=fva)
    frb = W.view(rm, B, feat_vec=fvb)
    nm_ref, out_ref = rm.search_by_bow_frame(kfa, frb, scene["pts"], 0.7, True)
    elig = ((A["assoc"] >= 0) & (W.bad[np.clip(A["assoc"], 0, None)] == 0)).astype(np.uint8)
    nm, m12 = oracle.search_by_bow(A["kps"], A["desc"], fva, elig, B["kps"], B["desc"], fvb, None, 50, 0.7, True)
    got = np.full(len(B["kps"]), -1, np.int32)
    got[m12[m12 >= 0]] = A["assoc"][np.nonzero(m12 >= 0)[0]]
    assert nm == nm_ref and np.array_equal(got, out_ref) and nm > 20
    kfb = W.view(rm, B, assoc=B["assoc"], feat_vec=fvb)
    nm_ref, out_ref = rm.search_by_bow_keyframes(kfa, kfb, scene["pts"], 0.75, True)
    elig2 = ((B["assoc"] >= 0) & (W.bad[np.clip(B["assoc"], 0, None)] == 0)).astype(np.uint8)
    nm, m12 = oracle.search_by_bow(A["kps"], A["desc"], fva, elig, B["kps"], B["desc"], fvb, elig2, 50, 0.75, True)
    got = np.where(m12 >= 0, B["assoc"][np.clip(m12, 0, None)], -1).astype(np.int32)
    assert nm == nm_ref and np.array_equal(got, out_ref) and nm > 10


@pytest.mark.parametrize("th,all_observed", [(1.0, False), (3.0, False), (4.0, True)])
def test_search_by_projection_two_cameras_restatement(th, all_observed):
    """oracle.search_by_projection_mappoints_stereo (ORBmatcher.cc:44-211 on a fisheye-stereo frame) == the reference's compiled ORBmatcher.cc,
    with the tracking fields of both cameras written by the reference's own Frame::isInFrustum (src/Frame.cc:643-660, 1270-1343)"""
    import oracle
    W = wc.World(seed=8, n_pts=1600, cam=wc.KB8, n_dup=200)
    if all_observed:
        W.n_obs[:] = np.maximum(W.n_obs, 1)
    R, t = W.pose(seed=12)
    obs = W.observe_two_cameras(R, t, seed=30)
    n = len(obs["assoc"])
    pre = np.where(np.random.default_rng(5).random(n) < 0.2, obs["assoc"], -1).astype(np.int32)
    view = W.view_two_cameras(rm, obs, assoc=pre)
    pts = W.points(rm)
    tr = rm.is_in_frustum(view, pts, 0.5, right=True)
    left = dict(in_view=tr["in_view"], proj_x=tr["proj_x"], proj_y=tr["proj_y"], level=tr["level"], view_cos=tr["view_cos"])
    right = dict(in_view=tr["in_view_r"], proj_x=tr["proj_xr_r"], proj_y=tr["proj_y_r"], level=tr["level_r"], view_cos=tr["view_cos_r"])
    nm_r, kp_r = rm.search_by_projection_mappoints_stereo(view, pts, th, 0.8, left, right, pre)
    # the restatement takes mvpMapPoints coded -1 / -2 (holder with observations) / -3 (holder without) and the gates of :52-62 folded into in_view
    bad = W.bad != 0
    gate = lambda f: dict(f, in_view=(np.asarray(f["in_view"]) != 0) & ~bad)
    code = np.where(pre < 0, -1, np.where(W.n_obs[np.maximum(pre, 0)] > 0, -2, -3)).astype(np.int32)
    nl = obs["n_left"]
    w, h = W.cam["size"]
    nm_o, kp_o = oracle.search_by_projection_mappoints_stereo(obs["kps"], obs["desc"][:nl], obs["kps_right"], obs["desc"][nl:], view.scale, w, h,
                                                              gate(left), gate(right), W.desc, obs["l2r"], obs["r2l"], th, 0.8,
                                                              observed=(W.n_obs > 0).astype(np.uint8), kp_mp=code)
    assert nm_o == nm_r and nm_r > 500
    got = np.where(kp_o >= 0, kp_o, np.where(code == -1, -1, pre))   # untouched holders keep their map point
    assert np.array_equal(got, kp_r)


@pytest.mark.parametrize("ori", [True, False])
def test_search_by_bow_two_cameras_restatement(ori):
    """oracle.search_by_bow_stereo (ORBmatcher.cc:338-431) == the reference's compiled SearchByBoW(KeyFrame, Frame) on a fisheye-stereo frame"""
    import oracle
    W = wc.World(seed=8, n_pts=1600, cam=wc.KB8, n_dup=0)
    R, t = W.pose(seed=12)
    obs = W.observe_two_cameras(R, t, seed=30)
    fvF = bowsynth.feature_vector(obs["desc"], 40, seed=3)
    frame = W.view_two_cameras(rm, obs, feat_vec=fvF)
    R2, t2 = W.pose(seed=13)
    ko = W.observe(R2, t2, seed=41)
    fvK = bowsynth.feature_vector(ko["desc"], 40, seed=3)
    kf = W.view(rm, ko, assoc=ko["assoc"], feat_vec=fvK)
    nm_r, out_r = rm.search_by_bow_frame(kf, frame, W.points(rm), 0.7, ori)
    elig = (ko["assoc"] >= 0) & (W.bad[np.maximum(ko["assoc"], 0)] == 0)
    kps2 = np.concatenate([obs["kps"], obs["kps_right"]])
    nm_o, mL, mR = oracle.search_by_bow_stereo(ko["kps"], ko["desc"], fvK, elig.astype(np.uint8), kps2, obs["desc"], fvF, obs["n_left"], 50, 0.7, ori)
    got = np.full(len(kps2), -1, np.int32)
    for m in (mL, mR):
        sel = m >= 0
        got[m[sel]] = ko["assoc"][sel]
    assert nm_o == nm_r and nm_r > 300
    assert np.array_equal(got, out_r)
    assert (mL >= 0).sum() > 100 and (mR >= 0).sum() > 100


# ---- KannalaBrandt8::epipolarConstrain, the gate of SearchForTriangulation on fisheye keyframes -------------------------------------
def _kb8_cam():
    return np.array(list(wc.KB8["K"]) + list(wc.KB8["kb"]), np.float32)


def test_kb8_unproject_restatement():
    """oracle.kb8_unproject in the host-libm mode == the reference's compiled KannalaBrandt8::unproject (KannalaBrandt8.cpp:112-137), bit
    for bit; the portable mode ((float)tan((double)theta), the CUDA arithmetic) moves ~3 % of the rays, by at most 4 ulp of a coordinate"""
    import oracle
    rng = np.random.default_rng(2)
    xy = np.concatenate([rng.uniform(0, 512, (20000, 2)), [[wc.KB8["K"][2], wc.KB8["K"][3]]]]).astype(np.float32)   # incl. the principal point
    ref = rm.kb8_unproject(_kb8_cam(), xy)
    lit = oracle.kb8_unproject(_kb8_cam(), xy, oracle.LIBM_F32)
    assert np.array_equal(lit.view(np.uint32), ref.view(np.uint32))
    por = oracle.kb8_unproject(_kb8_cam(), xy, oracle.LIBM_PORTABLE)
    d = np.abs(por.view(np.int32).astype(np.int64) - ref.view(np.int32))
    assert d.max() <= 4 and (d > 0).any(1).mean() < 0.05    # glibc's tanf is within an ulp of tan, not always the nearest float


def _pair_candidates(W, A, B, n_rand, seed):
    """keypoint pairs for the triangulation gate: the true correspondences of two observations plus random pairs"""
    ia = {int(m): i for i, m in enumerate(A["mp_idx"]) if m >= 0}
    tr = [(ia[int(m)], j) for j, m in enumerate(B["mp_idx"]) if m >= 0 and int(m) in ia]
    rng = np.random.default_rng(seed)
    rd = list(zip(rng.integers(0, len(A["kps"]), n_rand), rng.integers(0, len(B["kps"]), n_rand)))
    p = np.array(tr + rd)
    return p[:, 0], p[:, 1], len(tr)


@pytest.mark.parametrize("seed,dt", [(16, 0.25), (17, 0.6), (18, 0.05)])
def test_kb8_triangulate_matches_restatement(seed, dt):
    """oracle.kb8_triangulate_matches == the reference's compiled KannalaBrandt8::TriangulateMatches / Triangulate (KannalaBrandt8.cpp:343-444)
    over the stand-in cv:: (gemm, dot, norm, hconcat, row assignment, cv::SVD -- each pinned to cv2 in test_oracle_cvmat.py): the returned
    depth and the accepted point, bit for bit, on true correspondences (accepted, rejected by parallax or by reprojection) and random pairs"""
    import oracle
    from tri_cases import triangulation_geometry_kb8
    W = wc.World(seed=seed, n_pts=1500, cam=wc.KB8, n_dup=0)
    Ra, ta = W.pose(seed=30, dt=dt)
    Rb, tb = W.pose(seed=31, dt=dt)
    A, B = W.observe(Ra, ta, seed=40), W.observe(Rb, tb, seed=41)
    pairs, _ = triangulation_geometry_kb8(oracle, A["Tcw"], B["Tcw"], _kb8_cam(), gemm=rm.cv_gemm, libm_mode=oracle.LIBM_F32)
    G = pairs[0]
    i1, i2, ntrue = _pair_candidates(W, A, B, 1500, seed)
    xy1 = np.stack([A["kps"]["x"][i1], A["kps"]["y"][i1]], 1); xy2 = np.stack([B["kps"]["x"][i2], B["kps"]["y"][i2]], 1)
    _, _, s2, _ = rm.scale_tables(1.2, 8)
    sg1, sg2 = s2[A["kps"]["octave"][i1]], s2[B["kps"]["octave"][i2]]
    z_r, p_r = rm.kb8_triangulate_matches(_kb8_cam(), _kb8_cam(), G[:9], G[9:12], xy1, xy2, sg1, sg2)
    z_o, p_o = oracle.kb8_triangulate_matches(G, xy1, xy2, sg1, sg2, oracle.LIBM_F32)
    assert np.array_equal(z_o.view(np.uint32), z_r.view(np.uint32))
    assert np.array_equal(p_o.view(np.uint32), p_r.view(np.uint32))
    acc = z_r > 0.0001
    assert acc[:ntrue].mean() > (0.5 if dt >= 0.25 else 0.0) and acc[ntrue:].mean() < 0.2 and (~acc[:ntrue]).sum() > 5
    # portable libm policy: the same decisions on this sample (a flip needs a value within an ulp of a threshold)
    z_p, _ = oracle.kb8_triangulate_matches(G, xy1, xy2, sg1, sg2, oracle.LIBM_PORTABLE)
    assert np.array_equal(z_p > 0.0001, acc)


def _free(assoc, keep, seed):
    return np.where(np.random.default_rng(seed).random(len(assoc)) < keep, assoc, -1).astype(np.int32)


@pytest.mark.parametrize("only_stereo,coarse,ori", [(False, False, True), (False, True, True), (False, False, False), (True, False, True)])
def test_search_for_triangulation_kb8_restatement(only_stereo, coarse, ori):
    """oracle.search_for_triangulation_kb8 == the reference's compiled SearchForTriangulation (ORBmatcher.cc:981-1222) on monocular-fisheye
    keyframes: the gate is KannalaBrandt8::epipolarConstrain, R12 / t12 / the epipole are the reference's own cv::Mat lines"""
    import oracle
    from tri_cases import triangulation_geometry_kb8
    W = wc.World(seed=16, n_pts=1500, cam=wc.KB8, n_dup=0)
    Ra, ta = W.pose(seed=30); Rb, tb = W.pose(seed=31)
    A, B = W.observe(Ra, ta, seed=40), W.observe(Rb, tb, seed=41)
    fva, fvb = bowsynth.feature_vector(A["desc"], 24, seed=1), bowsynth.feature_vector(B["desc"], 24, seed=1)
    aa, bb = _free(A["assoc"], 0.25, 1), _free(B["assoc"], 0.25, 2)
    kf1, kf2 = W.view(rm, A, assoc=aa, feat_vec=fva), W.view(rm, B, assoc=bb, feat_vec=fvb)
    nm_r, pr = rm.search_for_triangulation(kf1, kf2, W.points(rm), np.eye(3, dtype=np.float32), only_stereo, coarse, 0.6, ori)
    pairs, ep = triangulation_geometry_kb8(oracle, A["Tcw"], B["Tcw"], _kb8_cam(), gemm=rm.cv_gemm, libm_mode=oracle.LIBM_F32)
    s1, s2 = (A["u_right"] >= 0), (B["u_right"] >= 0)
    e1 = ((aa < 0) & (~np.bool_(only_stereo) | s1)).astype(np.uint8); e2 = ((bb < 0) & (~np.bool_(only_stereo) | s2)).astype(np.uint8)
    nm_o, m12 = oracle.search_for_triangulation_kb8(A["kps"], A["desc"], fva, e1, s1, -1, B["kps"], B["desc"], fvb, e2, s2, -1, pairs, ep,
                                                    kf2.scale, kf1.scale ** 2, kf2.scale ** 2, coarse, ori, oracle.LIBM_F32)
    got = np.stack([np.nonzero(m12 >= 0)[0], m12[m12 >= 0]], 1)
    assert nm_o == nm_r and np.array_equal(got, pr)
    assert nm_r > (100 if not only_stereo else -1)


@pytest.mark.parametrize("coarse,ori", [(False, True), (True, False)])
def test_search_for_triangulation_two_cameras_restatement(coarse, ori):
    """the same on fisheye-stereo keyframes (mpCamera2 set, ORBmatcher.cc:1005-1019, 1103-1140): the pair (left|right, left|right) selects
    Rll..Rrr / tll..trr and the two cameras; the epipole gate is off"""
    import oracle
    from tri_cases import triangulation_geometry_kb8
    W = wc.World(seed=8, n_pts=1600, cam=wc.KB8, n_dup=0)
    Ra, ta = W.pose(seed=12); Rb, tb = W.pose(seed=13)
    A, B = W.observe_two_cameras(Ra, ta, seed=30), W.observe_two_cameras(Rb, tb, seed=50)
    fva, fvb = bowsynth.feature_vector(A["desc"], 30, seed=1), bowsynth.feature_vector(B["desc"], 30, seed=1)
    aa, bb = _free(A["assoc"], 0.25, 1), _free(B["assoc"], 0.25, 2)
    kf1, kf2 = W.view_two_cameras(rm, A, assoc=aa, feat_vec=fva), W.view_two_cameras(rm, B, assoc=bb, feat_vec=fvb)
    nm_r, pr = rm.search_for_triangulation(kf1, kf2, W.points(rm), np.eye(3, dtype=np.float32), False, coarse, 0.6, ori)
    pairs, ep = triangulation_geometry_kb8(oracle, A["Tcw"], B["Tcw"], _kb8_cam(), Tlr=A["Tlr"], Tlr2=B["Tlr"], gemm=rm.cv_gemm,
                                           libm_mode=oracle.LIBM_F32)
    k1, k2 = np.concatenate([A["kps"], A["kps_right"]]), np.concatenate([B["kps"], B["kps_right"]])
    nm_o, m12 = oracle.search_for_triangulation_kb8(k1, A["desc"], fva, (aa < 0).astype(np.uint8), None, A["n_left"], k2, B["desc"], fvb,
                                                    (bb < 0).astype(np.uint8), None, B["n_left"], pairs, ep, kf2.scale, kf1.scale ** 2,
                                                    kf2.scale ** 2, coarse, ori, oracle.LIBM_F32)
    got = np.stack([np.nonzero(m12 >= 0)[0], m12[m12 >= 0]], 1)
    assert nm_o == nm_r and np.array_equal(got, pr) and nm_r > 150
    right1, right2 = got[:, 0] >= A["n_left"], got[:, 1] >= B["n_left"]
    assert all(((right1 == a) & (right2 == b)).sum() > 5 for a in (False, True) for b in (False, True))   # ll, lr, rl, rr all occur


def test_distinctive_descriptors_restatement(oracle):
    """MapPoint::ComputeDistinctiveDescriptors: the restatement (oracle.distinctive_descriptors, what k_distinctive is held against)
    equals the reference's own body (src/MapPoint.cc:350-436, cut out at build time) -- chosen descriptor per map point, incl. points
    without observations, observations in bad keyframes (skipped), keyframes that see the point with both cameras (two rows),
    clusters of identical descriptors (ties go to the first row in the reference's map order) and small descriptors (low distances)."""
    rng = np.random.default_rng(12)
    counts = np.concatenate([[0, 1, 2, 3, 8, 33, 64, 5, 1, 2], rng.integers(1, 40, 60)]).astype(np.int32)
    start = np.concatenate([[0], np.cumsum(counts)]).astype(np.int32)
    nkf = int(start[-1])
    kind = rng.choice(np.array([0, 0, 0, 1, 2], np.uint8), nkf)
    kind[start[8]:start[9]] = 1            # a point all of whose keyframes are bad: mDescriptor stays empty
    kind[start[9]:start[10]] = [1, 2]
    rows = np.where(kind == 2, 2, 1)
    row0 = np.concatenate([[0], np.cumsum(rows)])
    desc = rng.integers(0, 256, (int(row0[-1]), 32), dtype=np.uint8)
    for p in range(12, 40, 3):             # families of near-duplicates and exact duplicates: equal medians
        a, b = int(row0[start[p]]), int(row0[start[p + 1]])
        desc[a:b] = desc[a]
        if b - a > 3:
            desc[a + 1, 0] ^= 1
            desc[b - 1, 5] ^= 3
    ref_desc, has = rm.distinctive(desc, start, kind)
    # what a caller of the product gathers: the rows of the keyframes that are not bad, in map order
    keep_rows, pstart = [], [0]
    for p in range(len(counts)):
        for k in range(start[p], start[p + 1]):
            if kind[k] != 1:
                keep_rows += list(range(int(row0[k]), int(row0[k + 1])))
        pstart.append(len(keep_rows))
    d2 = desc[keep_rows] if keep_rows else np.zeros((0, 32), np.uint8)
    bi, bm = oracle.distinctive_descriptors(d2, np.asarray(pstart, np.int32))
    for p in range(len(counts)):
        n = pstart[p + 1] - pstart[p]
        assert bool(has[p]) == (n > 0), p
        if n:
            assert 0 <= bi[p] < n
            assert np.array_equal(ref_desc[p], d2[pstart[p] + bi[p]]), p


def _sorted_children(voc):
    """the same tree with every node's children stored in ascending id order (what a text vocabulary file can express)"""
    v = dict(voc)
    ci = voc["child_ids"].copy()
    cs = voc["child_start"]
    for p in range(len(cs) - 1):
        ci[cs[p]:cs[p + 1]] = np.sort(ci[cs[p]:cs[p + 1]])
    v["child_ids"] = ci
    return v


@pytest.mark.parametrize("k,L,levelsup", [(10, 3, 1), (10, 3, 4), (6, 4, 2), (9, 2, 0), (3, 5, 4)])
def test_vocabulary_transform_restatement(oracle, tmp_path, k, L, levelsup):
    """Frame::ComputeBoW (src/Frame.cc:828-835): the restatement (oracle.vocab_transform + vocab_fold, what k_vocab_transform and the
    host fold are held against) equals the reference's own vendored DBoW2 -- the vocabulary loaded by loadFromTextFile as ORB-SLAM3
    loads ORBvoc.txt, TemplatedVocabulary::transform with its strict-'<' child choice, TF-IDF accumulation in feature order, the L1
    normalisation over ascending word ids and the FeatureVector at `levelsup` levels above the leaves (nodes, feature lists, order);
    doubles compared bit for bit.  Queries include exact node descriptors (distance 0) and descriptors equidistant from two
    children (ties -> the first child); a few words carry weight 0 (stopped words: no BowVector / FeatureVector entry)."""
    from oracle import refbow
    if not refbow.available():
        pytest.skip("oracle/_ref/libref_bow.so not built (needs /root/reference)")
    voc = _sorted_children(bowsynth.vocabulary(k, L, 7 * k + L, ragged=False))
    path = str(tmp_path / "voc.txt")
    refbow.write_text_vocabulary(path, voc, k)
    rv = refbow.Vocabulary(path)
    try:
        leaves = np.flatnonzero(voc["node_word_id"] >= 0)
        assert rv.words() == len(leaves)
        rng = np.random.default_rng(100 + k + L)
        desc = np.concatenate([rng.integers(0, 256, (300, 32), dtype=np.uint8), voc["node_desc"][rng.choice(leaves, 40)],
                               voc["node_desc"][1:1 + min(20, len(voc["node_desc"]) - 1)]])
        # ties: a query exactly between two sibling descriptors (flip half of the differing bits of the first one)
        cs, ci = voc["child_start"], voc["child_ids"]
        a, b = voc["node_desc"][ci[cs[0]]], voc["node_desc"][ci[cs[0] + 1]]
        diff = np.flatnonzero(np.unpackbits(a ^ b))
        t = np.unpackbits(a).copy(); t[diff[: len(diff) // 2]] ^= 1
        desc = np.concatenate([desc, np.packbits(t)[None, :]])
        w, nid, wt = oracle.vocab_transform(voc, desc, levelsup)
        assert np.array_equal(rv.transform_words(desc), w.astype(np.uint32))
        ids, vals, (fn, fs, fx) = oracle.vocab_fold(w, nid, wt)
        rids, rvals, (rfn, rfs, rfx) = rv.transform(desc, levelsup)
        assert np.array_equal(ids, rids) and np.array_equal(vals.view(np.uint64), rvals.view(np.uint64))
        assert np.array_equal(fn, rfn) and np.array_equal(fs, rfs) and np.array_equal(fx, rfx)
        # empty input: empty containers
        e = rv.transform(np.zeros((0, 32), np.uint8), levelsup)
        assert len(e[0]) == 0 and len(e[2][0]) == 0
    finally:
        rv.close()


@pytest.mark.parametrize("w,h,nfeat,fx,bf", [(752, 480, 1200, 435.2046959714599, 47.90639384423901), (512, 512, 1500, 190.97847715128717, 21.0),
                                             (1241, 376, 2000, 718.856, 386.1448)])
def test_compute_stereo_matches_restatement(oracle, w, h, nfeat, fx, bf):
    """Frame::ComputeStereoMatches (src/Frame.cc:901-1079, cut out of the reference's Frame.cc) behind two reference extractors ==
    oracle.stereo_matches behind two oracle extractors on seeded rectified pairs (EuRoC, a square rig, KITTI): left keypoints, mvuRight
    and mvDepth byte for byte, on the CPU (the GPU test test_batch256_matches_oracle holds the kernels against both)."""
    import importlib
    synth = importlib.import_module("3_orb_slam3_selfnote_b200.synth")
    mbf = np.float32(bf); mb = np.float32(mbf / np.float32(fx))
    ref = rm.StereoFrames(nfeat)
    exL = oracle.Extractor(nfeat, sincos_mode=oracle.SINCOS_LIBM_F32); exR = oracle.Extractor(nfeat, sincos_mode=oracle.SINCOS_LIBM_F32)
    matched = 0
    for f in (0, 3, 7, 12):
        left, right = synth.frame(w, h, 2000 + w, f, False), synth.frame(w, h, 2000 + w, f, True)
        r = ref(left, right, float(mb), float(mbf))
        _, kl, dl = exL(left, (0, 0)); _, kr, dr = exR(right, (0, 0))
        uo, do, _, _ = oracle.stereo_matches(exL, exR, kl, dl, kr, dr, mb, mbf)
        assert r["kpsL"].tobytes() == kl.tobytes(), f
        assert r["uRight"].tobytes() == uo.tobytes() and r["depth"].tobytes() == do.tobytes(), f
        matched += int((uo >= 0).sum())
    assert matched > 200


@pytest.mark.parametrize("kb", [None, "tumvi"])
def test_is_in_frustum_and_grid_restatement(oracle, kb):
    """Frame::isInFrustum (src/Frame.cc:572-642 + MapPoint::PredictScale + Pinhole / KannalaBrandt8 project, all cut from the reference's
    files) == oracle.is_in_frustum in its literal-libm mode, field by field, on 3 000 map points in front of / behind / beside the
    camera, in and out of the scale-invariance range, facing towards and away; and Frame::GetFeaturesInArea (src/Frame.cc:744-813 over the
    grid of AssignFeaturesToGrid) == oracle.features_in_area, indices in the reference's order, on random windows and level ranges."""
    import frame_cases as fc
    kbv = None if kb is None else fc.KB8_TUMVI
    cam, P, normal, maxd, mind = fc.frustum_case(oracle, 3000, seed=5, log_mode=0, kb=kbv)
    cam.logScaleFactor = float(np.log(np.float32(1.2)))
    Tcw = np.eye(4, dtype=np.float32)
    Tcw[:3, :3] = np.array(cam.Rcw[:], np.float32).reshape(3, 3); Tcw[:3, 3] = np.array(cam.tcw[:], np.float32)
    # the camera centre as Frame::UpdatePoseMatrices forms it (mOw = -mRcw.t() * mtcw, a float gemm), not numpy's product
    cam.Ow[:] = [float(x) for x in np.asarray(pro.cam_center(Tcw[:3, :3].copy(), Tcw[:3, 3].copy())).reshape(-1)]
    rng = np.random.default_rng(9)
    n = 1500
    kps = np.zeros(n, oracle.KP_DTYPE)
    kps["x"] = rng.uniform(0, 752, n).astype(np.float32); kps["y"] = rng.uniform(0, 480, n).astype(np.float32)
    kps["octave"] = rng.integers(0, 8, n); kps["size"] = 31; kps["angle"] = rng.uniform(0, 360, n).astype(np.float32)
    desc = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    view = rm.View(kps, desc, (cam.minX, cam.maxX, cam.minY, cam.maxY), K=(cam.fx, cam.fy, cam.cx, cam.cy), bf=cam.bf, b=cam.bf / cam.fx,
                   u_right=np.full(n, -1, np.float32), Tcw=Tcw, kb=kbv)
    pts = rm.Points(P, rng.integers(0, 256, (len(P), 32), dtype=np.uint8), normal, mind, maxd, np.ones(len(P), np.int32), np.zeros(len(P), np.uint8))
    r = rm.is_in_frustum(view, pts, 0.5)
    o = oracle.is_in_frustum(cam, P, normal, maxd, mind)
    assert np.array_equal(r["in_view"], o["in_view"]) and 300 < int(o["in_view"].sum()) < 2700
    iv = o["in_view"] != 0
    for k in ("proj_x", "proj_y", "proj_xr", "depth", "view_cos"):
        assert r[k][iv].tobytes() == o[k][iv].tobytes(), k
    assert np.array_equal(r["level"][iv], o["level"][iv])
    if kb is None:
        # the grid over the image bounds (0, w, 0, h) that oracle.features_in_area assumes
        gview = rm.View(kps, desc, (0.0, 752.0, 0.0, 480.0), K=(cam.fx, cam.fy, cam.cx, cam.cy), u_right=np.full(n, -1, np.float32))
        hits = 0
        for _ in range(200):
            x, y, rad = float(rng.uniform(-20, 780)), float(rng.uniform(-20, 500)), float(rng.uniform(1, 90))
            lo = int(rng.integers(-1, 6)); hi = int(rng.integers(lo, 8)) if lo >= 0 and rng.random() < 0.7 else -1
            a = rm.features_in_area(gview, x, y, rad, lo, hi)
            b = oracle.features_in_area(kps, 752, 480, x, y, rad, lo, hi)
            assert np.array_equal(a, b), (x, y, rad, lo, hi)
            hits += len(a)
        assert hits > 500
