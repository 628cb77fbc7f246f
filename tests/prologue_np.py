"""numpy restatement (float32, operation by operation) of the per-point prologue that precedes the window searches of
ORBmatcher.cc -- transform, depth test, projection, bounds, scale-invariance range, viewing angle, PredictScale, radius -- i.e.
what k_project_points (csrc/orbx_match.cu) computes on the device.  Test infrastructure: together with the oracle's search loops
it restates SearchByProjection(Frame, Frame) (src/ORBmatcher.cc:2027-2289), SearchByProjection(Frame, KeyFrame) (:2291-2413), the
KeyFrame-Scw searches (:489-720) and Fuse (:1425-1786), and tests/test_oracle_refmatch.py pins it to the reference's compiled source.
"""
import numpy as np

f32 = np.float32


def cam_center(R, t):
    """-R^T t as cv::Mat evaluates `-Rcw.t()*tcw` (ORBmatcher.cc:2038, Frame.cc:538): the unary minus MATERIALISES the transpose
    (MatOp::subtract(Scalar, expr) assigns the expression to a Mat and scales it by -1), so the product is a plain 3x3 * 3x1 gemm with
    alpha = -1 on the small-matrix path: float products summed left to right in float, then (float)(s * alpha).  (A gemm with the
    GEMM_1_T flag would accumulate in double; cv2.gemm shows both, and the reference's compiled Frame::UpdatePoseMatrices gives the float
    one -- tests/test_oracle_refmatch.py::test_is_in_frustum_and_grid_restatement depends on it to the last bit.)"""
    R = R.astype(np.float32); t = t.astype(np.float32).reshape(-1)
    out = np.empty(3, np.float32)
    for i in range(3):
        s = f32(f32(f32(R[0, i] * t[0]) + f32(R[1, i] * t[1])) + f32(R[2, i] * t[2]))
        out[i] = f32(np.float64(s) * -1.0)
    return out


def gemv_add(R, x, t):
    """R*x + t as one small-matrix gemm: float products accumulated left to right, then + t"""
    R = R.astype(f32); t = t.astype(f32)
    out = []
    for k in range(3):
        s = R[k, 0] * x[:, 0]
        s = s + R[k, 1] * x[:, 1]
        s = s + R[k, 2] * x[:, 2]
        out.append((s + t[k]).astype(f32))
    return out


def project(P, valid, R, t, Ow, K, bounds, *, kb=None, depth_mode=0, invz_double=True, bounds_mode=0, bf=0.0, max_d=None, min_d=None,
            normal=None, level_mode=0, octave=None, log_scale=None, nlevels=8, scale_factors=None, th=1.0, pinhole_formula=False):
    """-> dict(valid, u, v, ur, level, radius, invz).  depth_mode 0: invz < 0 rejects, 1: z < 0 rejects, 2: none.
    level_mode 0: level = octave[i]; 1: MapPoint::PredictScale on RAW max_d (MapPoint.cc:570-602)."""
    P = np.ascontiguousarray(P, f32)
    n = len(P)
    fx, fy, cx, cy = [f32(v) for v in K]
    x, y, z = gemv_add(R, P, t)
    with np.errstate(all="ignore"):
        invz = (1.0 / z.astype(np.float64)).astype(f32) if invz_double else (f32(1.0) / z).astype(f32)
        ok = np.asarray(valid, bool).copy()
        if depth_mode == 0:
            ok &= ~(invz < 0)
        elif depth_mode == 1:
            ok &= ~(z < 0)
        if pinhole_formula:
            u = (fx * (x * invz) + cx).astype(f32); v = (fy * (y * invz) + cy).astype(f32)
        elif kb is not None:   # KannalaBrandt8::project, KannalaBrandt8.cpp:29-45 (atan2f / cos / sin as the device evaluates them)
            x2y2 = (x * x + y * y).astype(f32)
            theta = np.arctan2(np.sqrt(x2y2).astype(np.float64), z.astype(np.float64)).astype(f32)
            psi = np.arctan2(y.astype(np.float64), x.astype(np.float64)).astype(f32)
            t2 = theta * theta; t3 = theta * t2; t5 = t3 * t2; t7 = t5 * t2; t9 = t7 * t2
            k = [f32(c) for c in kb]
            r = ((((theta + k[0] * t3) + k[1] * t5) + k[2] * t7) + k[3] * t9).astype(f32)
            u = ((fx * r).astype(np.float64) * np.cos(psi.astype(np.float64)) + np.float64(cx)).astype(f32)
            v = ((fy * r).astype(np.float64) * np.sin(psi.astype(np.float64)) + np.float64(cy)).astype(f32)
        else:
            u = ((fx * x) / z + cx).astype(f32); v = ((fy * y) / z + cy).astype(f32)
        minx, maxx, miny, maxy = [f32(b) for b in bounds]
        if bounds_mode == 0:
            ok &= ~((u < minx) | (u > maxx) | (v < miny) | (v > maxy))
        else:
            ok &= (u >= minx) & (u < maxx) & (v >= miny) & (v < maxy)
        ur = (u - f32(bf) * invz).astype(f32)
        level = np.zeros(n, np.int32)
        if max_d is not None or normal is not None or level_mode == 1:
            o = (P - Ow.astype(f32)[None]).astype(f32)
            dist = np.sqrt((o.astype(np.float64) ** 2).sum(1)).astype(f32)     # cv::norm: double accumulation
            if max_d is not None:
                ok &= ~((dist < f32(0.8) * min_d) | (dist > f32(1.2) * max_d))
            if normal is not None:
                dot = (o.astype(np.float64) * normal.astype(np.float64)).sum(1)
                ok &= ~(dot < 0.5 * dist.astype(np.float64))
            if level_mode == 1:
                ratio = (max_d / dist).astype(f32)
                lg = np.log(ratio.astype(np.float64)).astype(f32)
                lv = np.ceil(lg / f32(log_scale))
                lv = np.where(np.isfinite(lv), lv, 0)
                level = np.clip(lv, 0, nlevels - 1).astype(np.int32)
        if level_mode == 0:
            level = np.asarray(octave, np.int32)
        radius = (f32(th) * np.asarray(scale_factors, f32)[np.clip(level, 0, nlevels - 1)]).astype(f32)
    return dict(valid=ok.astype(np.uint8), u=u, v=v, ur=ur, level=level, radius=radius, invz=invz)
