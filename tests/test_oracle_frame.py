"""Pins oracle/frame_oracle.cc.  cv::undistortPoints against the real OpenCV (python cv2, live,
bit-exact on the float outputs); isInFrustum against an independent numpy float32 / float64 restatement of
src/Frame.cc:572-642 written from the reference's text."""
import numpy as np
import pytest

import frame_cases as fc


def test_undistort_points_matches_cv2(oracle):
    cv2 = pytest.importorskip("cv2")
    for cam in fc.CAMERAS.values():
        fx, fy, cx, cy = [np.float32(v) for v in cam["K"]]
        K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float32)
        D = np.array(cam["D"], np.float32)
        pts = fc.pixel_points(cam, 20000)
        ref = cv2.undistortPoints(pts.reshape(-1, 1, 2), K, D.reshape(1, -1), None, K).reshape(-1, 2)
        got = oracle.undistort_points(pts, fx, fy, cx, cy, D)
        assert np.array_equal(ref.view(np.int32), got.view(np.int32))


def test_undistort_points_random_cameras(oracle):
    """cv::undistortPoints restatement against live cv2 on 40 seeded random cameras (focal lengths, principal points, 4 or 5 distortion
    coefficients incl. strong barrel / pincushion) x 300 points, float bits; a 200-camera run (60 000 points) found no difference."""
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(101)
    for _ in range(40):
        fx, fy = float(rng.uniform(150, 1200)), float(rng.uniform(150, 1200))
        w, h = int(rng.integers(200, 2000)), int(rng.integers(200, 1500))
        cx, cy = float(rng.uniform(0.3, 0.7) * w), float(rng.uniform(0.3, 0.7) * h)
        nd = int(rng.choice([4, 5]))
        D = np.array([rng.uniform(-0.4, 0.4), rng.uniform(-0.3, 0.3), rng.uniform(-0.01, 0.01), rng.uniform(-0.01, 0.01), rng.uniform(-0.5, 0.5)][:nd], np.float32)
        K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float32)
        pts = (rng.random((300, 2)) * [w, h]).astype(np.float32)
        ref = cv2.undistortPoints(pts.reshape(-1, 1, 2), K, D.reshape(1, -1), None, K).reshape(-1, 2)
        got = oracle.undistort_points(pts, np.float32(fx), np.float32(fy), np.float32(cx), np.float32(cy), D)
        assert np.array_equal(ref.view(np.int32), got.view(np.int32))


def _frustum_numpy(cam, P, normal, maxd, mind, log_mode):
    f32 = np.float32
    R = np.array(cam.Rcw[:], f32).reshape(3, 3); t = np.array(cam.tcw[:], f32); Ow = np.array(cam.Ow[:], f32)
    out = []
    for p, nrm, mx, mn in zip(P, normal, maxd, mind):
        Pc = np.array([f32(f32(f32(R[r, 0] * p[0]) + f32(R[r, 1] * p[1])) + f32(R[r, 2] * p[2])) + t[r] for r in range(3)], f32)
        pcd = f32(np.sqrt(np.sum(Pc.astype(np.float64) ** 2)))
        if Pc[2] < 0:
            out.append((0, -1, -1, 0, 0, -1, 0)); continue
        with np.errstate(all="ignore"):
            invz = f32(1) / Pc[2]
            if cam.cameraModel == 1:   # KannalaBrandt8::project, portable atan2 policy (log_mode 1)
                th = f32(np.arctan2(np.float64(np.sqrt(f32(f32(Pc[0] * Pc[0]) + f32(Pc[1] * Pc[1])))), np.float64(Pc[2])))
                psi = f32(np.arctan2(np.float64(Pc[1]), np.float64(Pc[0])))
                t2 = f32(th * th); t3 = f32(th * t2); t5 = f32(t3 * t2); t7 = f32(t5 * t2); t9 = f32(t7 * t2)
                k = [f32(v) for v in cam.kb[:]]
                r = f32(f32(f32(f32(th + f32(k[0] * t3)) + f32(k[1] * t5)) + f32(k[2] * t7)) + f32(k[3] * t9))
                ux = f32(np.float64(f32(f32(cam.fx) * r)) * np.cos(np.float64(psi)) + np.float64(f32(cam.cx)))
                uy = f32(np.float64(f32(f32(cam.fy) * r)) * np.sin(np.float64(psi)) + np.float64(f32(cam.cy)))
            else:
                ux = f32(f32(f32(cam.fx) * Pc[0]) / Pc[2]) + f32(cam.cx); uy = f32(f32(f32(cam.fy) * Pc[1]) / Pc[2]) + f32(cam.cy)
        if not (ux >= f32(cam.minX) and ux <= f32(cam.maxX) and uy >= f32(cam.minY) and uy <= f32(cam.maxY)):
            if ux < f32(cam.minX) or ux > f32(cam.maxX) or uy < f32(cam.minY) or uy > f32(cam.maxY):
                out.append((0, -1, -1, 0, 0, -1, 0)); continue
        PO = (p - Ow).astype(f32)
        dist = f32(np.sqrt(np.sum(PO.astype(np.float64) ** 2)))
        if dist < f32(f32(0.8) * mn) or dist > f32(f32(1.2) * mx):   # Get{Min,Max}DistanceInvariance, MapPoint.cc:555-565
            out.append((0, ux, uy, 0, 0, -1, 0)); continue
        vc = f32(np.sum(PO.astype(np.float64) * nrm.astype(np.float64)) / np.float64(dist))
        if vc < f32(cam.viewingCosLimit):
            out.append((0, ux, uy, 0, 0, -1, 0)); continue
        ratio = f32(mx / dist)   # PredictScale divides the raw mfMaxDistance, MapPoint.cc:587-602
        lg = f32(np.log(np.float64(ratio))) if log_mode else np.log(ratio)
        lvl = int(np.ceil(f32(lg / f32(cam.logScaleFactor))))
        lvl = 0 if lvl < 0 else min(lvl, cam.nScaleLevels - 1)
        out.append((1, ux, uy, f32(ux - f32(f32(cam.bf) * invz)), pcd, lvl, vc))
    return out


@pytest.mark.parametrize("kb", [None, fc.KB8_TUMVI])
def test_is_in_frustum_matches_numpy_restatement(oracle, kb):
    cam, P, normal, maxd, mind = fc.frustum_case(oracle, 3000, seed=5, kb=kb)
    o = oracle.is_in_frustum(cam, P, normal, maxd, mind)
    ref = _frustum_numpy(cam, P, normal, maxd, mind, cam.logMode)
    keys = ("in_view", "proj_x", "proj_y", "proj_xr", "depth", "level", "view_cos")
    for i, r in enumerate(ref):
        for k, v in zip(keys, r):
            assert o[k][i] == v or (np.isnan(o[k][i]) and np.isnan(v)), (i, k, o[k][i], v)
    frac = o["in_view"].mean()
    assert 0.1 < frac < 0.8  # the case exercises every rejection branch and the accept path
    # the reference-literal logf mode may only differ in the predicted level, and in practice does not
    cam.logMode = oracle.LOG_LIBM_F32
    o2 = oracle.is_in_frustum(cam, P, normal, maxd, mind)
    assert np.array_equal(o2["in_view"], o["in_view"]) and (o2["level"] != o["level"]).sum() == 0
    # ... and, for the fisheye model, move a projection by at most a few float ulps (libm atan2f against the rounded double atan2)
    assert np.abs(o2["proj_x"] - o["proj_x"]).max() <= 2e-4 and np.abs(o2["proj_y"] - o["proj_y"]).max() <= 2e-4


def test_predicted_level_uses_raw_max_distance(oracle):
    """MapPoint::PredictScale (MapPoint.cc:587-602): level = ceil(log(mfMaxDistance / dist) / log(scaleFactor)) with the RAW
    mfMaxDistance, while the range gate of Frame::isInFrustum (Frame.cc:609-615) uses 1.2f * mfMaxDistance / 0.8f * mfMinDistance.
    Expected levels are built directly from that definition: a point seen at dist = mfMaxDistance / 1.2^(k - 0.5) is at level k."""
    cam = fc.frustum_case(oracle, 1, seed=1)[0]
    cam.Rcw[:] = [1, 0, 0, 0, 1, 0, 0, 0, 1]; cam.tcw[:] = [0, 0, 0]; cam.Ow[:] = [0, 0, 0]
    dist = np.float32(4.0)
    P = np.tile(np.array([[0, 0, dist]], np.float32), (12, 1))
    normal = np.tile(np.array([[0, 0, 1]], np.float32), (12, 1))
    k = np.arange(12)
    maxd = (dist * np.float64(1.2) ** (k - 0.5)).astype(np.float32)       # k = 0: dist just above mfMaxDistance but inside 1.2 x
    mind = np.full(12, 0.1, np.float32)
    o = oracle.is_in_frustum(cam, P, normal, maxd, mind)
    assert o["in_view"].all()                                            # dist <= 1.2 * mfMaxDistance even for k = 0
    assert o["level"].tolist() == [0, 1, 2, 3, 4, 5, 6, 7, 7, 7, 7, 7]     # clamped to nScaleLevels - 1 = 7
    # the gates: beyond 1.2 x mfMaxDistance / below 0.8 x mfMinDistance the point is rejected
    o = oracle.is_in_frustum(cam, P[:4], normal[:4], np.array([dist / 1.19, dist / 1.21, 50, 50], np.float32),
                             np.array([0.1, 0.1, dist / 0.81, dist / 0.79], np.float32))
    assert o["in_view"].tolist() == [1, 0, 1, 0]


def test_right_camera_pose_composition_matches_cv2_gemm():
    """frustum_right_camera (src/Frame.cc:1274-1281 + :1192-1195) composes the right camera's pose with the float op order of cv::gemm's
    small-matrix path: checked bit for bit against cv2.gemm, the arithmetic behind the reference's cv::Mat expressions."""
    import importlib
    cv2 = pytest.importorskip("cv2")
    pp = importlib.import_module("3_orb_slam3_selfnote_b200.preproc")
    rng = np.random.default_rng(3)
    for _ in range(200):
        R, _ = np.linalg.qr(rng.normal(size=(3, 3)))
        Rcw = R.astype(np.float32); tcw = rng.normal(size=3).astype(np.float32)
        Q, _ = np.linalg.qr(np.eye(3) + rng.normal(size=(3, 3)) * 0.01)
        Tlr = np.concatenate([Q.astype(np.float32), rng.normal(size=(3, 1)).astype(np.float32) * 0.1], 1)
        left = pp.frustum(Rcw, tcw, 190.0, 190.0, 255.0, 257.0, (0, 512, 0, 512), 0.0, 1.2, 8, 0.5, kb=(0.003, 0.0007, -0.002, 0.0002))
        g = pp.frustum_right_camera(left, Tlr, 191.0, 191.5, 254.0, 256.0, kb=(0.003, 0.0007, -0.002, 0.0002))
        Rrl = np.ascontiguousarray(Tlr[:, :3].T)
        trl = cv2.gemm(Rrl, (np.float32(-1) * Tlr[:, 3:4]).astype(np.float32), 1, None, 0)
        Ow = np.array(left.Ow[:], np.float32).reshape(3, 1)
        assert np.array_equal(np.array(g.Rcw[:], np.float32).reshape(3, 3), cv2.gemm(Rrl, Rcw, 1, None, 0))
        assert np.array_equal(np.array(g.tcw[:], np.float32).reshape(3, 1), cv2.gemm(Rrl, tcw.reshape(3, 1), 1, trl, 1))
        assert np.array_equal(np.array(g.Ow[:], np.float32).reshape(3, 1), cv2.gemm(np.ascontiguousarray(Rcw.T), Tlr[:, 3:4].copy(), 1, Ow, 1))
        assert (g.fx, g.camera_model) == (191.0, 1)
