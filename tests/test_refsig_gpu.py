"""The reference-signature overloads of host/ORBmatcher.h (pose algebra on the host in the reference's own cv::Mat lines, per-point
projection prologue + search on the device) against the reference's OWN src/ORBmatcher.cc: oracle/_ref/libref_match.so and its twin
oracle/_ref/libshim_match.so are the same harness -- same caller code, same Frame / KeyFrame / MapPoint model, same stand-in cv::
-- linked once with the reference's ORBmatcher and once with the drop-in.  Every observable output must be identical."""
import numpy as np
import pytest

import world_cases as wc
from oracle import refmatch as rm

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not (rm.available() and rm.twin_available()), reason="oracle/_ref twin libraries not built")]


def both(fn, *a, **k):
    ref = fn(*a, **k)
    with rm.twin():
        got = fn(*a, **k)
    return ref, got


def same(ref, got, what):
    assert len(ref) == len(got)
    for i, (r, g) in enumerate(zip(ref, got)):
        if isinstance(r, np.ndarray):
            assert r.shape == g.shape and (r == g).all(), "%s: output %d differs at %s" % (what, i, np.nonzero(r != g)[0][:8])
        else:
            assert r == g, "%s: output %d: reference %r, drop-in %r" % (what, i, r, g)


@pytest.fixture(scope="module", params=["euroc", "kb8"])
def scene(request):
    cam = wc.EUROC if request.param == "euroc" else wc.KB8
    W = wc.World(seed=5 if request.param == "euroc" else 6, n_pts=1800, cam=cam, n_dup=200)
    Ra, ta = W.pose(seed=10)
    Rb, tb = W.pose(seed=11)
    A, B = W.observe(Ra, ta, seed=20), W.observe(Rb, tb, seed=21)
    return dict(W=W, A=A, B=B, pts=W.points(rm), Ra=Ra, ta=ta, Rb=Rb, tb=tb, name=request.param)


@pytest.mark.parametrize("mono,th,dt,ori", [(True, 7.0, None, True), (False, 15.0, None, True), (False, 7.0, (0, 0, 0.6), True),
                                            (False, 7.0, (0, 0, -0.6), False), (True, 2.0, None, False)])
def test_search_by_projection_frame_frame(scene, mono, th, dt, ori):
    """src/ORBmatcher.cc:2027-2289: plain / forward / backward windows (|tlc.z| > mb), stereo ur check, rotation histogram"""
    W, A, B = scene["W"], scene["A"], scene["B"]
    if dt is not None:   # move the current camera along its optical axis so that bForward / bBackward trigger
        B = W.observe(scene["Rb"], scene["tb"] - np.array(dt), seed=23)
    pre = np.where(np.random.default_rng(4).random(len(B["kps"])) < 0.1, B["assoc"], -1).astype(np.int32)
    for assoc in (np.full(len(B["kps"]), -1, np.int32), pre):
        cur = W.view(rm, B, assoc=assoc)
        last = W.view(rm, A, assoc=A["assoc"], outlier=A["outlier"])
        ref, got = both(rm.search_by_projection_frame, cur, last, scene["pts"], th, mono, 0.9, ori)
        same(ref, got, "SearchByProjection(Frame, Frame)")
        assert ref[0] > 50


def test_search_by_projection_frame_frame_shared_points(scene):
    """every map point is held by two keypoints of the last frame: the second query lands on the keypoint the first one took, and takes it
    again when the point has no observations (:2134-2136) -- the count differs from the blocking case even where the final holders agree"""
    W, A, B = scene["W"], scene["A"], scene["B"]
    A2 = {k: np.concatenate([A[k], A[k]]) for k in ("kps", "desc", "u_right", "assoc", "outlier")}
    A2["Tcw"] = A["Tcw"]
    cur = W.view(rm, B, assoc=np.full(len(B["kps"]), -1, np.int32))
    last = W.view(rm, A2, assoc=A2["assoc"], outlier=A2["outlier"])
    ref, got = both(rm.search_by_projection_frame, cur, last, scene["pts"], 7.0, True, 0.9, True)
    same(ref, got, "SearchByProjection(Frame, Frame), shared points")
    assert ref[0] > (ref[1] >= 0).sum() + 20   # re-taken keypoints were counted twice


@pytest.mark.parametrize("th,orb_dist,ori", [(10.0, 100, True), (3.0, 64, True), (10.0, 100, False)])
def test_search_by_projection_frame_keyframe(scene, th, orb_dist, ori):
    """src/ORBmatcher.cc:2291-2413 (relocalisation)"""
    W, A, B = scene["W"], scene["A"], scene["B"]
    found = (np.random.default_rng(7).random(W.n) < 0.15).astype(np.uint8)
    pre = np.where(np.random.default_rng(8).random(len(B["kps"])) < 0.2, B["assoc"], -1).astype(np.int32)
    cur = W.view(rm, B, assoc=pre)
    kf = W.view(rm, A, assoc=A["assoc"])
    ref, got = both(rm.search_by_projection_keyframe, cur, kf, scene["pts"], found, th, orb_dist, 0.9, ori)
    same(ref, got, "SearchByProjection(Frame, KeyFrame)")
    assert ref[0] > 50


@pytest.mark.parametrize("th,ratio,with_kfs,s", [(4, 1.0, False, 1.07), (8, 0.9, True, 0.93), (3, 0.5, False, 1.0)])
def test_search_by_projection_scw(scene, th, ratio, with_kfs, s):
    """src/ORBmatcher.cc:489-602 and :604-720 (loop detection / place recognition)"""
    W, B = scene["W"], scene["B"]
    Scw = B["Tcw"].copy()
    Scw[:3, :] *= np.float32(s)
    kf = W.view(rm, B, assoc=B["assoc"])
    matched = np.where(np.random.default_rng(1).random(len(B["kps"])) < 0.2, B["assoc"], -1).astype(np.int32)
    ref, got = both(rm.search_by_projection_scw, kf, Scw, scene["pts"], th, ratio, matched, with_kfs)
    same(ref, got, "SearchByProjection(KeyFrame, Scw)")
    assert ref[0] > 20


@pytest.mark.parametrize("th", [3.0, 6.5])
def test_fuse(scene, th):
    """src/ORBmatcher.cc:1425-1662 (bRight = false): AddObservation / AddMapPoint and both Replace directions"""
    if scene["name"] == "kb8":
        pytest.skip("Fuse's stereo branch needs bf > 0; the mono branch is covered by the euroc world's non-stereo keypoints")
    W, B = scene["W"], scene["B"]
    kf = W.view(rm, B, assoc=B["assoc"])   # the world's duplicated points arrive as candidates: Replace in both directions
    cand = np.random.default_rng(3).permutation(W.n).astype(np.int32)
    ref, got = both(rm.fuse, kf, scene["pts"], cand, th)
    same(ref, got, "Fuse")
    ev = ref[1]
    assert (ev[:, 0] == 0).sum() > 10 and (ev[:, 0] == 1).sum() > 50


def test_fuse_mono_keyframe(scene):
    """the same on a keyframe without stereo measurements (mvuRight all -1) and a fisheye camera model"""
    W, B = scene["W"], dict(scene["B"])
    B["u_right"] = np.full(len(B["kps"]), -1.0, np.float32)
    kf = W.view(rm, B, assoc=B["assoc"])
    cand = np.arange(W.n, dtype=np.int32)
    ref, got = both(rm.fuse, kf, scene["pts"], cand, 3.0)
    same(ref, got, "Fuse (mono)")
    assert ref[0] > 50


@pytest.mark.parametrize("th,s", [(4.0, 1.05), (7.5, 0.97)])
def test_fuse_scw(scene, th, s):
    """src/ORBmatcher.cc:1664-1786 (loop closing)"""
    W, B = scene["W"], scene["B"]
    Scw = B["Tcw"].copy()
    Scw[:3, :] *= np.float32(s)
    kf = W.view(rm, B, assoc=B["assoc"])
    cand = np.random.default_rng(9).permutation(W.n).astype(np.int32)
    ref, got = both(rm.fuse_scw, kf, Scw, scene["pts"], cand, th)
    same(ref, got, "Fuse(Scw)")
    assert ref[0] > 50 and (ref[1] >= 0).sum() > 10


@pytest.mark.parametrize("th,s12", [(7.5, 1.0), (10.0, 1.04)])
def test_search_by_sim3(scene, th, s12):
    """src/ORBmatcher.cc:1788-2018: both projection passes and the agreement check"""
    if scene["name"] == "kb8":
        pytest.skip("SearchBySim3 projects with fx, fy, cx, cy only (pinhole), src/ORBmatcher.cc:1792-1795")
    W, A, B = scene["W"], scene["A"], scene["B"]
    Ra, ta, Rb, tb = scene["Ra"], scene["ta"], scene["Rb"], scene["tb"]
    R12 = Ra @ Rb.T
    t12 = ta - R12 @ tb
    kf1 = W.view(rm, A, assoc=A["assoc"])
    kf2 = W.view(rm, B, assoc=B["assoc"])
    # some pairs already matched: keypoints of KF1 whose point is also held by KF2
    in2 = set(B["assoc"][B["assoc"] >= 0].tolist())
    rng = np.random.default_rng(2)
    m12 = np.array([a if (a >= 0 and a in in2 and rng.random() < 0.3) else -1 for a in A["assoc"]], np.int32)
    ref, got = both(rm.search_by_sim3, kf1, kf2, scene["pts"], m12, s12, R12, t12 * s12, th)
    same(ref, got, "SearchBySim3")
    assert ref[0] > 20


def test_search_for_triangulation_reference_signature(scene):
    """src/ORBmatcher.cc:981-1222 through the cv::Mat F12 parameter list (epipole projected by the shim).  Pinhole keyframes: the epipolar
    line test with the caller's F12; KannalaBrandt8 keyframes: the reference's own KannalaBrandt8::epipolarConstrain (two-view
    triangulation through cv::SVD, compiled in oracle/_ref) against k_bow_triangulate_kb8 -- the reference's binary calls the host's tanf /
    atan2f, the device rounds the double functions, and the matched pairs are the same."""
    import bowsynth
    W, A, B = scene["W"], scene["A"], scene["B"]
    Ra, ta, Rb, tb = scene["Ra"], scene["ta"], scene["Rb"], scene["tb"]
    fx, fy, cx, cy = W.cam["K"]
    K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]])
    R12 = Ra @ Rb.T
    t12 = ta - R12 @ tb
    tx = np.array([[0, -t12[2], t12[1]], [t12[2], 0, -t12[0]], [-t12[1], t12[0], 0]])
    F12 = (np.linalg.inv(K).T @ tx @ R12 @ np.linalg.inv(K)).astype(np.float32)
    fva = bowsynth.feature_vector(A["desc"], 24, seed=1)
    fvb = bowsynth.feature_vector(B["desc"], 24, seed=1)
    free = lambda a, seed: np.where(np.random.default_rng(seed).random(len(a)) < 0.3, a, -1).astype(np.int32)
    for a1, a2, floor in ((A["assoc"], B["assoc"], 10), (free(A["assoc"], 1), free(B["assoc"], 2), 100)):
        kf1 = W.view(rm, A, assoc=a1, feat_vec=fva)
        kf2 = W.view(rm, B, assoc=a2, feat_vec=fvb)
        for only_stereo, coarse in ((False, False), (True, False), (False, True)):
            ref, got = both(rm.search_for_triangulation, kf1, kf2, scene["pts"], F12, only_stereo, coarse)
            same(ref, got, "SearchForTriangulation")
        assert ref[0] > floor


@pytest.mark.parametrize("coarse,ori", [(False, True), (True, False)])
def test_search_for_triangulation_two_cameras(coarse, ori):
    """src/ORBmatcher.cc:1005-1019, 1103-1140 on fisheye-stereo keyframes: Rll..Rrr / tll..trr from the shim's own cv::Mat lines, features of
    both cameras in one list, the KannalaBrandt8 gate with the camera pair the two features belong to"""
    import bowsynth
    W = wc.World(seed=8, n_pts=1600, cam=wc.KB8, n_dup=0)
    Ra, ta = W.pose(seed=12); Rb, tb = W.pose(seed=13)
    A, B = W.observe_two_cameras(Ra, ta, seed=30), W.observe_two_cameras(Rb, tb, seed=50)
    fva, fvb = bowsynth.feature_vector(A["desc"], 30, seed=1), bowsynth.feature_vector(B["desc"], 30, seed=1)
    free = lambda a, seed: np.where(np.random.default_rng(seed).random(len(a)) < 0.25, a, -1).astype(np.int32)
    kf1 = W.view_two_cameras(rm, A, assoc=free(A["assoc"], 1), feat_vec=fva)
    kf2 = W.view_two_cameras(rm, B, assoc=free(B["assoc"], 2), feat_vec=fvb)
    ref, got = both(rm.search_for_triangulation, kf1, kf2, W.points(rm), np.eye(3, dtype=np.float32), False, coarse, 0.6, ori)
    same(ref, got, "SearchForTriangulation (two cameras)")
    assert ref[0] > 150
    r1, r2 = ref[1][:, 0] >= A["n_left"], ref[1][:, 1] >= B["n_left"]
    assert all(((r1 == a) & (r2 == b)).sum() > 5 for a in (False, True) for b in (False, True))


# ---- fisheye-stereo frames (F.Nleft != -1; Examples/Stereo/TUM_512.yaml-shaped rigs) ------------------------------------------
@pytest.fixture(scope="module")
def rig():
    W = wc.World(seed=8, n_pts=2200, cam=wc.KB8, n_dup=300)
    R, t = W.pose(seed=12)
    obs = W.observe_two_cameras(R, t, seed=30)
    return dict(W=W, obs=obs, pts=W.points(rm))


def _tracks(view, pts):
    """the MapPoint tracking fields of both cameras as the reference's Frame::isInFrustum writes them (src/Frame.cc:643-660, 1270-1343)"""
    t = rm.is_in_frustum(view, pts, 0.5, right=True)
    left = dict(in_view=t["in_view"], proj_x=t["proj_x"], proj_y=t["proj_y"], level=t["level"], view_cos=t["view_cos"])
    right = dict(in_view=t["in_view_r"], proj_x=t["proj_xr_r"], proj_y=t["proj_y_r"], level=t["level_r"], view_cos=t["view_cos_r"])
    return left, right, t["depth"]


@pytest.mark.parametrize("th,pre_share,far", [(1.0, 0.0, False), (3.0, 0.15, False), (5.0, 0.4, True)])
def test_search_by_projection_two_cameras(rig, th, pre_share, far):
    """src/ORBmatcher.cc:44-211 on a fisheye-stereo frame: left pass, right pass, stereo-partner hand-over, claims across both"""
    W, obs = rig["W"], rig["obs"]
    n = len(obs["assoc"])
    pre = np.where(np.random.default_rng(5).random(n) < pre_share, obs["assoc"], -1).astype(np.int32)
    view = W.view_two_cameras(rm, obs, assoc=pre)
    left, right, depth = _tracks(view, rig["pts"])
    assert left["in_view"].sum() > 300 and right["in_view"].sum() > 300
    ref, got = both(rm.search_by_projection_mappoints_stereo, view, rig["pts"], th, 0.8, left, right, pre, depth, far, 6.0)
    same(ref, got, "SearchByProjection(Frame, MapPoints) two cameras")
    nl = obs["n_left"]
    new = (ref[1] >= 0) & (ref[1] != pre)
    assert new[:nl].sum() > 100 and new[nl:].sum() > 100
    # the integrated path (isInFrustum inside the harness, then the search)
    ref2, got2 = both(rm.search_by_projection_mappoints, view, rig["pts"], th, 0.8, None, 0.5, pre)
    same(ref2, got2, "SearchLocalPoints core, two cameras")


@pytest.mark.parametrize("th,pre_share,rep", [(3.0, 0.0, 1), (5.0, 0.2, 1), (4.0, 0.1, 3)])
def test_search_by_projection_two_cameras_all_observed(rig, th, pre_share, rep):
    """every map point has observations (what Tracking::SearchLocalPoints hands over): the drop-in resolves both camera passes as a
    parallel fixed point (k_resolve_fisheye_par) instead of the one-warp replay; rep = 3 crowds every keypoint with three candidates"""
    W, obs = rig["W"], rig["obs"]
    rng = np.random.default_rng(23)
    n = len(obs["assoc"])
    view0 = W.view_two_cameras(rm, obs)
    left, right, depth = _tracks(view0, rig["pts"])
    order = rng.permutation(W.n * rep)
    t3 = lambda a: np.tile(a, (rep,) + (1,) * (a.ndim - 1))[order]
    n_obs = rng.integers(1, 6, W.n * rep).astype(np.int32)
    P = rm.Points(t3(W.P), t3(W.desc), t3(W.normal), t3(W.min_dist), t3(W.max_dist), n_obs, t3(W.bad))

    def widen(f):
        g = {k: t3(v) for k, v in f.items()}
        if rep > 1:
            g["proj_x"] = (g["proj_x"] + rng.normal(size=W.n * rep) * 1.5).astype(np.float32)
            g["proj_y"] = (g["proj_y"] + rng.normal(size=W.n * rep) * 1.5).astype(np.float32)
        return g
    L, R = widen(left), widen(right)
    # pre-existing associations must point at observed points of P: map an original index to one of its copies
    where = np.full(W.n, -1, np.int64)
    where[order % W.n] = np.arange(W.n * rep)
    pre = np.where((rng.random(n) < pre_share) & (obs["assoc"] >= 0), where[np.maximum(obs["assoc"], 0)], -1).astype(np.int32)
    view = W.view_two_cameras(rm, obs, assoc=pre)
    ref, got = both(rm.search_by_projection_mappoints_stereo, view, P, th, 0.8, L, R, pre)
    same(ref, got, "SearchByProjection two cameras, all observed")
    nl = obs["n_left"]
    new = (ref[1] >= 0) & (ref[1] != pre)
    assert new[:nl].sum() > 100 and new[nl:].sum() > 100


def test_search_by_projection_two_cameras_crowded(rig):
    """many map points per keypoint and map points without observations: claims get replaced through the partner hand-over"""
    W, obs = rig["W"], rig["obs"]
    rng = np.random.default_rng(17)
    view = W.view_two_cameras(rm, obs)
    left, right, depth = _tracks(view, rig["pts"])
    # three copies of every map point (jittered projections, shuffled), a third of them without observations
    rep = 3
    order = rng.permutation(W.n * rep)
    t3 = lambda a: np.tile(a, (rep,) + (1,) * (a.ndim - 1))[order]
    n_obs = (rng.random(W.n * rep) < 0.66).astype(np.int32)
    P = rm.Points(t3(W.P), t3(W.desc), t3(W.normal), t3(W.min_dist), t3(W.max_dist), n_obs, t3(W.bad))

    def widen(f):
        g = {k: t3(v) for k, v in f.items()}
        g["proj_x"] = (g["proj_x"] + rng.normal(size=W.n * rep) * 1.5).astype(np.float32)
        g["proj_y"] = (g["proj_y"] + rng.normal(size=W.n * rep) * 1.5).astype(np.float32)
        return g
    L, R = widen(left), widen(right)
    ref, got = both(rm.search_by_projection_mappoints_stereo, view, P, 4.0, 0.9, L, R)
    same(ref, got, "SearchByProjection two cameras, crowded")
    assert ref[0] > 400


@pytest.mark.parametrize("kf_two_cameras,ori", [(False, True), (True, True), (True, False)])
def test_search_by_bow_two_cameras(rig, kf_two_cameras, ori):
    """src/ORBmatcher.cc:338-431: the frame is a fisheye-stereo one; a keyframe feature may take a left and a right feature"""
    import bowsynth
    W, obs = rig["W"], rig["obs"]
    fvF = bowsynth.feature_vector(obs["desc"], 40, seed=3)
    frame = W.view_two_cameras(rm, obs, feat_vec=fvF)
    R2, t2 = W.pose(seed=13)
    if kf_two_cameras:
        ko = W.observe_two_cameras(R2, t2, seed=40)
        kf = W.view_two_cameras(rm, ko, assoc=ko["assoc"], feat_vec=bowsynth.feature_vector(ko["desc"], 40, seed=3))
    else:
        ko = W.observe(R2, t2, seed=41)
        kf = W.view(rm, ko, assoc=ko["assoc"], feat_vec=bowsynth.feature_vector(ko["desc"], 40, seed=3))
    ref, got = both(rm.search_by_bow_frame, kf, frame, rig["pts"], 0.7, ori)
    same(ref, got, "SearchByBoW(KeyFrame, Frame) two cameras")
    nl = obs["n_left"]
    assert (ref[1][:nl] >= 0).sum() > 50 and (ref[1][nl:] >= 0).sum() > 50


def test_is_in_frustum_checks_two_cameras(rig):
    """Frame::isInFrustum on a fisheye-stereo frame (src/Frame.cc:643-660 -> isInFrustumChecks :1270-1343, both cameras): the device
    kernel with the right camera's pose composed as :1274-1281 against the reference's compiled Frame.cc.  Fields are compared where the
    reference wrote them (points in view).  Tolerance: the compiled reference calls libm's float atan2f / logf, the
    kernel evaluates (float)atan2((double)..) / (float)log((double)..): decisions and predicted levels identical, projections within
    2e-4 px; depth and viewing cosine (no transcendental involved) bit-equal."""
    import importlib
    pkg = importlib.import_module("3_orb_slam3_selfnote_b200")
    pre_mod = importlib.import_module("3_orb_slam3_selfnote_b200.preproc")
    W, obs = rig["W"], rig["obs"]
    view = W.view_two_cameras(rm, obs)
    t = rm.is_in_frustum(view, rig["pts"], 0.5, right=True)
    cam = W.cam
    fx, fy, cx, cy = cam["K"]
    Tcw = obs["Tcw"]
    left = pre_mod.frustum(Tcw[:3, :3], Tcw[:3, 3], fx, fy, cx, cy, W.bounds(), 0.0, W.scale, W.nlevels, 0.5, kb=cam["kb"])
    right = pre_mod.frustum_right_camera(left, obs["Tlr"], fx, fy, cx, cy, kb=cam["kb"])
    pre = pre_mod.Preprocessor()
    try:
        for g, sfx, px_key in ((left, "", "proj_x"), (right, "_r", "proj_xr_r")):
            out = pre.isInFrustum(g, W.P, W.normal, W.max_dist, W.min_dist)
            inv = t["in_view" + sfx] != 0
            assert np.array_equal(out["in_view"] != 0, inv) and inv.sum() > 300
            assert np.array_equal(out["level"][inv], t["level" + sfx][inv])
            for ours, ref in ((out["proj_x"], t[px_key]), (out["proj_y"], t["proj_y" + sfx])):
                d = np.abs(ours[inv] - ref[inv])
                assert d.max() <= 2e-4
            assert np.array_equal(out["depth"][inv].view(np.uint32), t["depth" + sfx][inv].view(np.uint32)), "depth" + sfx
            assert np.array_equal(out["view_cos"][inv].view(np.uint32), t["view_cos" + sfx][inv].view(np.uint32)), "view_cos" + sfx
    finally:
        pre.close()


@pytest.mark.parametrize("mono,th,dz,ori", [(False, 7.0, 0.0, True), (False, 10.0, 0.6, True), (False, 7.0, -0.6, False), (True, 15.0, 0.0, True)])
def test_search_by_projection_frame_frame_two_cameras(rig, mono, th, dz, ori):
    """src/ORBmatcher.cc:2027-2289 with a fisheye-stereo current frame: the right-image pass (:2188-2255) is skipped where the left one hit
    a `continue`, claims stay inside each image, one rotation histogram over both; the last frame is a two-camera frame too (octaves
    and angles of its right keypoints, :2100-2101, :2169-2171)"""
    W = rig["W"]
    Ra, ta = W.pose(seed=14)
    Rb, tb = W.pose(seed=15)
    last = W.observe_two_cameras(Ra, ta, seed=50)
    cur = W.observe_two_cameras(Rb, tb - np.array([0, 0, dz]), seed=52)
    n = len(cur["assoc"])
    pre = np.where(np.random.default_rng(6).random(n) < 0.1, cur["assoc"], -1).astype(np.int32)
    for assoc in (np.full(n, -1, np.int32), pre):
        vc = W.view_two_cameras(rm, cur, assoc=assoc)
        vl = W.view_two_cameras(rm, last, assoc=last["assoc"])
        ref, got = both(rm.search_by_projection_frame, vc, vl, rig["pts"], th, mono, 0.9, ori)
        same(ref, got, "SearchByProjection(Frame, Frame) two cameras")
        nl = cur["n_left"]
        new = (ref[1] >= 0) & (ref[1] != assoc)
        assert new[:nl].sum() > 50 and new[nl:].sum() > 50


def test_search_by_projection_frame_keyframe_two_cameras(rig):
    """src/ORBmatcher.cc:2291-2413 on a fisheye-stereo frame: the reference searches its left image only"""
    W, obs = rig["W"], rig["obs"]
    R2, t2 = W.pose(seed=16)
    ko = W.observe(R2, t2, seed=60)
    n = len(obs["assoc"])
    pre = np.where(np.random.default_rng(8).random(n) < 0.2, obs["assoc"], -1).astype(np.int32)
    cur = W.view_two_cameras(rm, obs, assoc=pre)
    kf = W.view(rm, ko, assoc=ko["assoc"])
    found = (np.random.default_rng(7).random(W.n) < 0.15).astype(np.uint8)
    ref, got = both(rm.search_by_projection_keyframe, cur, kf, rig["pts"], found, 10.0, 100, 0.9, True)
    same(ref, got, "SearchByProjection(Frame, KeyFrame) two cameras")
    assert ref[0] > 50 and (ref[1][obs["n_left"]:] != pre[obs["n_left"]:]).sum() == 0


@pytest.mark.parametrize("right,th", [(False, 3.0), (True, 3.0), (True, 6.5)])
def test_fuse_two_cameras(rig, right, th):
    """src/ORBmatcher.cc:1425-1662 on a fisheye-stereo keyframe: bRight selects the right pose, mpCamera2, mvKeysRight / mGridRight and
    offsets descriptor rows and indices by NLeft"""
    W, obs = rig["W"], rig["obs"]
    kf = W.view_two_cameras(rm, obs, assoc=obs["assoc"])
    cand = np.random.default_rng(3).permutation(W.n).astype(np.int32)
    ref, got = both(rm.fuse, kf, rig["pts"], cand, th, right)
    same(ref, got, "Fuse two cameras")
    ev = ref[1]
    assert ref[0] > 50 and (ev[:, 0] == 1).sum() > 30
