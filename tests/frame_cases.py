"""Seeded inputs for the Frame-glue tests: cameras from the reference's own YAMLs and a synthetic
local map around a camera pose."""
import numpy as np

CAMERAS = {
    # Examples/Monocular/EuRoC.yaml (Camera.fx .. Camera.p2)
    "euroc": dict(K=(458.654, 457.296, 367.215, 248.375), D=(-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05), size=(752, 480)),
    # Examples/Monocular/TUM1.yaml (5 coefficients, k3 last)
    "tum1": dict(K=(517.306408, 516.469215, 318.643040, 255.313989), D=(0.262383, -0.953104, -0.005358, 0.002628, 1.163314), size=(640, 480)),
    "mild": dict(K=(600.0, 610.0, 320.5, 239.5), D=(0.1, -0.05, 0.001, -0.002, 0.003), size=(640, 480)),
}


def pixel_points(cam, n, seed=0):
    rng = np.random.default_rng(seed)
    w, h = cam["size"]
    pts = rng.random((n, 2)) * [w, h]
    corners = [[0, 0], [w, 0], [0, h], [w, h]]  # ComputeImageBounds, Frame.cc:876-879
    return np.concatenate([pts, corners]).astype(np.float32)


def pose(seed):
    rng = np.random.default_rng(seed)
    a = rng.normal(size=3) * 0.3
    th = np.linalg.norm(a)
    k = a / th
    K = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    R = np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * K @ K
    t = rng.normal(size=3)
    return R.astype(np.float32), t.astype(np.float32)


KB8_TUMVI = (0.003482389402, 0.000715034845, -0.002053236141, 0.000202936736)   # Examples/Monocular/TUM_512.yaml Camera.k1..k4


def frustum_case(oracle, n, seed=0, log_mode=1, kb=None):
    """camera + n map points: in front / behind, inside / outside the image, in / out of the scale-invariance range, facing
    towards / away"""
    rng = np.random.default_rng(seed)
    R, t = pose(seed)
    fx, fy, cx, cy = [np.float32(v) for v in CAMERAS["euroc"]["K"]]
    cam = oracle.FrustumCam()
    cam.Rcw[:] = R.reshape(-1).tolist(); cam.tcw[:] = t.tolist()
    Ow = -(R.T @ t).astype(np.float32)
    cam.Ow[:] = Ow.tolist()
    cam.fx, cam.fy, cam.cx, cam.cy = fx, fy, cx, cy
    cam.minX, cam.maxX, cam.minY, cam.maxY = -15.3, 770.2, -12.1, 495.7
    cam.bf = 47.90639384423901
    cam.logScaleFactor = float(np.log(np.float32(1.2)))
    cam.viewingCosLimit = 0.5
    cam.nScaleLevels = 8
    cam.logMode = log_mode
    cam.cameraModel = 0 if kb is None else 1
    if kb is not None:
        cam.kb[:] = [float(k) for k in kb]
    # points in camera coordinates, then to the world
    z = rng.uniform(-2, 12, n)
    x = (rng.uniform(-0.3, 1.3, n) * 752 - cx) / fx * z
    y = (rng.uniform(-0.3, 1.3, n) * 480 - cy) / fy * z
    Pc = np.stack([x, y, z], 1)
    P = ((Pc - t) @ R).astype(np.float32)  # R^T (Pc - t)
    d = np.linalg.norm(P - Ow, axis=1)
    maxd = (d * rng.uniform(0.7, 6.0, n)).astype(np.float32)
    mind = (maxd / np.float32(1.2 ** 7) * rng.uniform(0.5, 1.5, n)).astype(np.float32)
    nrm = (P - Ow) / np.maximum(d[:, None], 1e-6) + rng.normal(size=(n, 3)) * 0.6
    nrm = (nrm / np.linalg.norm(nrm, axis=1, keepdims=True)).astype(np.float32)
    return cam, P, nrm, maxd, mind
