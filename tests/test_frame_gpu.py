"""GPU parity tests of SURVEY.md §8(f) rank 2 through the C-ABI: orbx_undistort_keypoints and orbx_is_in_frustum against the
CPU oracle (bit-exact: every float compared by its bits), then isInFrustum's output driving SearchByProjection as in
Tracking::SearchLocalPoints (src/Tracking.cc:3490-3520)."""
import importlib

import numpy as np
import pytest

import frame_cases as fc

pytestmark = pytest.mark.gpu
KEYS = ("in_view", "proj_x", "proj_y", "proj_xr", "depth", "level", "view_cos")


@pytest.fixture(scope="module")
def pre():
    mod = importlib.import_module("3_orb_slam3_selfnote_b200.preproc")
    p = mod.Preprocessor()
    p.mod = mod
    yield p
    p.close()


def _bits(a):
    a = np.ascontiguousarray(a)
    return a.view(np.int32) if a.dtype == np.float32 else a


@pytest.mark.parametrize("name", list(fc.CAMERAS))
def test_undistort_keypoints(pre, orbx, oracle, name):
    cam = fc.CAMERAS[name]
    fx, fy, cx, cy = [np.float32(v) for v in cam["K"]]
    pts = fc.pixel_points(cam, 5000, seed=3)
    kps = np.zeros(len(pts), orbx.KP_DTYPE)
    kps["x"], kps["y"] = pts[:, 0], pts[:, 1]
    rng = np.random.default_rng(1)
    kps["size"], kps["angle"], kps["response"] = 31, rng.uniform(0, 360, len(pts)), rng.integers(7, 200, len(pts))
    kps["octave"], kps["class_id"] = rng.integers(0, 8, len(pts)), -1
    un = pre.UndistortKeyPoints(kps, fx, fy, cx, cy, cam["D"])
    ref = oracle.undistort_points(pts, fx, fy, cx, cy, np.array(cam["D"], np.float32))
    assert np.array_equal(_bits(un["x"]), _bits(ref[:, 0])) and np.array_equal(_bits(un["y"]), _bits(ref[:, 1]))
    for f in ("size", "angle", "response", "octave", "class_id"):
        assert np.array_equal(un[f], kps[f])
    assert np.abs(un["x"] - kps["x"]).max() > 1.0  # the distortion is not a no-op
    # mDistCoef(0) == 0: mvKeysUn = mvKeys (Frame.cc:839-843), whatever the other coefficients say
    same = pre.UndistortKeyPoints(kps, fx, fy, cx, cy, [0.0, 0.3, 0.01, 0.01])
    assert same.tobytes() == kps.tobytes()
    assert len(pre.UndistortKeyPoints(kps[:0], fx, fy, cx, cy, cam["D"])) == 0


@pytest.mark.parametrize("n,seed", [(5000, 5), (1, 6), (70000, 7)])
def test_is_in_frustum(pre, oracle, n, seed):
    cam, P, normal, maxd, mind = fc.frustum_case(oracle, n, seed=seed, log_mode=oracle.LOG_PORTABLE)
    ref = oracle.is_in_frustum(cam, P, normal, maxd, mind)
    g = pre.mod.frustum(np.array(cam.Rcw[:]).reshape(3, 3), cam.tcw[:], cam.fx, cam.fy, cam.cx, cam.cy,
                        (cam.minX, cam.maxX, cam.minY, cam.maxY), cam.bf, 1.2, cam.nScaleLevels, cam.viewingCosLimit)
    assert np.array_equal(np.array(g.Ow[:], np.float32), np.array(cam.Ow[:], np.float32))
    out = pre.isInFrustum(g, P, normal, maxd, mind)
    for k in KEYS:
        assert np.array_equal(_bits(out[k]), _bits(ref[k])), k
    if n >= 5000:
        assert 0.1 < out["in_view"].mean() < 0.8 and len(np.unique(out["level"][out["in_view"] == 1])) >= 6
    # reference-literal logf on the host: same decisions and levels on this case
    cam.logMode = oracle.LOG_LIBM_F32
    lit = oracle.is_in_frustum(cam, P, normal, maxd, mind)
    assert np.array_equal(lit["level"], out["level"])


def test_frustum_feeds_search_by_projection(pre, orbx, oracle, synth):
    """SearchLocalPoints: map points = keypoints of a jittered copy of the frame back-projected to 3-D at random depths; the
    frustum outputs (in_view, proj, level, viewCos) go straight into SearchByProjection and must reproduce the oracle chain."""
    W, H = 752, 480
    img = synth.frame(W, H, 21, 0)
    ex = oracle.Extractor(1000, blur_taps=oracle.TAPS_OCV346, sincos_mode=oracle.SINCOS_PORTABLE)
    _, k, d = ex(img, (0, 0))
    scale = np.array([np.float32(1.2) ** i for i in range(8)], np.float32)
    rng = np.random.default_rng(4)
    n = len(k)
    R, t = fc.pose(9)
    fx, fy, cx, cy = [np.float32(v) for v in fc.CAMERAS["euroc"]["K"]]
    z = rng.uniform(1.5, 9.0, n)
    Pc = np.stack([(k["x"] + rng.uniform(-1, 1, n) - cx) / fx * z, (k["y"] + rng.uniform(-1, 1, n) - cy) / fy * z, z], 1)
    P = ((Pc - t) @ R).astype(np.float32)
    Ow = -(R.T @ t).astype(np.float32)
    dist = np.linalg.norm(P - Ow, axis=1)
    maxd = (dist * scale[k["octave"]] * 1.1).astype(np.float32)   # so that PredictScale lands near the keypoint's octave
    mind = (maxd / scale[7] * 0.8).astype(np.float32)
    nrm = ((P - Ow) / dist[:, None]).astype(np.float32)
    cam = oracle.FrustumCam()
    cam.Rcw[:] = R.reshape(-1).tolist(); cam.tcw[:] = t.tolist(); cam.Ow[:] = Ow.tolist()
    cam.fx, cam.fy, cam.cx, cam.cy = fx, fy, cx, cy
    cam.minX, cam.maxX, cam.minY, cam.maxY = 0, W, 0, H
    cam.bf, cam.logScaleFactor, cam.viewingCosLimit, cam.nScaleLevels, cam.logMode = 47.9, float(np.log(np.float32(1.2))), 0.5, 8, 1
    ref = oracle.is_in_frustum(cam, P, nrm, maxd, mind)
    nm_o, kp_o = oracle.search_by_projection_mappoints(k, d, scale, W, H, ref["in_view"], ref["proj_x"], ref["proj_y"], ref["level"],
                                                       ref["view_cos"], d, 1.0, 0.8)
    g = pre.mod.frustum(R, t, fx, fy, cx, cy, (0, W, 0, H), 47.9, 1.2, 8)
    out = pre.isInFrustum(g, P, nrm, maxd, mind)
    F = orbx.FrameView(k, d, scale, W, H)
    m = orbx.ORBmatcher(0.8)
    nm_g, kp_g = m.SearchByProjection(F, out["in_view"], out["proj_x"], out["proj_y"], out["level"], out["view_cos"], d, 1.0)
    assert nm_g == nm_o and np.array_equal(kp_g, kp_o) and nm_o > 400
    m.close()
