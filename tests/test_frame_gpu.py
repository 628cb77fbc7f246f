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


@pytest.mark.parametrize("n,seed,kb", [(5000, 5, None), (1, 6, None), (70000, 7, None), (20000, 8, fc.KB8_TUMVI)])
def test_is_in_frustum(pre, oracle, n, seed, kb):
    cam, P, normal, maxd, mind = fc.frustum_case(oracle, n, seed=seed, log_mode=oracle.LOG_PORTABLE, kb=kb)
    ref = oracle.is_in_frustum(cam, P, normal, maxd, mind)
    g = pre.mod.frustum(np.array(cam.Rcw[:]).reshape(3, 3), cam.tcw[:], cam.fx, cam.fy, cam.cx, cam.cy,
                        (cam.minX, cam.maxX, cam.minY, cam.maxY), cam.bf, 1.2, cam.nScaleLevels, cam.viewingCosLimit, kb=kb, Ow=cam.Ow[:])
    assert np.array_equal(np.array(g.Ow[:], np.float32), np.array(cam.Ow[:], np.float32))
    out = pre.isInFrustum(g, P, normal, maxd, mind)
    if kb is None:
        for k in KEYS:
            assert np.array_equal(_bits(out[k]), _bits(ref[k])), k
    else:
        # KannalaBrandt8: atan2 / cos / sin come from two double-precision libraries (CUDA, glibc) that are each within 1-2 ulp
        # of the exact value, so after the rounding to float a projection may differ in its last bit once in ~10^8 values.
        # Tolerance: decisions and levels identical, at most 2 of the 2 x 20 000 projected coordinates off, by <= 1e-4 px.
        assert np.array_equal(out["in_view"], ref["in_view"]) and np.array_equal(out["level"], ref["level"])
        for k in ("proj_x", "proj_y", "proj_xr"):
            assert (_bits(out[k]) != _bits(ref[k])).sum() <= 2 and np.abs(out[k] - ref[k]).max() <= 1e-4, k
        for k in ("depth", "view_cos"):
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
    maxd = (dist * scale[k["octave"]] * 0.95).astype(np.float32)   # raw mfMaxDistance: PredictScale = ceil(octave - 0.28) = the octave
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
    g = pre.mod.frustum(R, t, fx, fy, cx, cy, (0, W, 0, H), 47.9, 1.2, 8, Ow=Ow)
    out = pre.isInFrustum(g, P, nrm, maxd, mind)
    F = orbx.FrameView(k, d, scale, W, H)
    m = orbx.ORBmatcher(0.8)
    nm_g, kp_g = m.SearchByProjection(F, out["in_view"], out["proj_x"], out["proj_y"], out["level"], out["view_cos"], d, 1.0)
    assert nm_g == nm_o and np.array_equal(kp_g, kp_o) and nm_o > 400
    m.close()


def _local_map_case(orbx, synth, pre, nframes, nmp, seed):
    """frames on the device + a shared local map: the keypoints of frame 0 back-projected through frame 0's pose, padded with
    random points; every frame sees the map from its own (slightly different) pose"""
    import torch
    W = H = 512
    frames = synth.frames(W, H, seed, range(nframes), False)
    dev = torch.from_numpy(frames).cuda()
    gx = orbx.ORBextractor(1500, max_batch=nframes)
    gx.extract_batch_device(dev.data_ptr(), nframes, W, H, W, W * H, (0, 0), 0)
    torch.cuda.synchronize()
    counts, _, kps, desc = gx.results_host(nframes)
    rng = np.random.default_rng(seed)
    fx, fy, cx, cy = np.float32(190.978), np.float32(190.973), np.float32(254.93), np.float32(256.9)  # TUM-VI cam0 (pinhole part)
    R0, t0 = fc.pose(seed)
    c0 = int(counts[0])
    k0 = kps[0, :c0]
    z = rng.uniform(1.0, 8.0, c0)
    Pc = np.stack([(k0["x"] + 1.0 - cx) / fx * z, (k0["y"] - 1.0 - cy) / fy * z, z], 1)
    P = ((Pc - t0) @ R0).astype(np.float32)
    Ow0 = -(R0.T @ t0).astype(np.float32)
    scale = gx.GetScaleFactors()
    d0 = np.linalg.norm(P - Ow0, axis=1)
    maxd = (d0 * scale[k0["octave"]] * 0.95).astype(np.float32)
    npad = nmp - c0
    Ppad = (((np.stack([rng.uniform(-3, 3, npad), rng.uniform(-3, 3, npad), rng.uniform(-1, 8, npad)], 1)) - t0) @ R0).astype(np.float32)
    P = np.concatenate([P, Ppad]); maxd = np.concatenate([maxd, rng.uniform(2, 20, npad).astype(np.float32)])
    mind = (maxd / scale[7] * 0.8).astype(np.float32)
    nrm = (P - Ow0) / np.maximum(np.linalg.norm(P - Ow0, axis=1, keepdims=True), 1e-6) + rng.normal(size=(nmp, 3)) * 0.3
    nrm = (nrm / np.linalg.norm(nrm, axis=1, keepdims=True)).astype(np.float32)
    mdesc = np.concatenate([desc[0, :c0], rng.integers(0, 256, (npad, 32), dtype=np.uint8)])
    observed = (rng.uniform(size=nmp) < 0.95).astype(np.uint8)
    poses = []
    for f in range(nframes):
        dR, dt = fc.pose(1000 + f)
        Rf = (np.eye(3) * (1 - 0.02 * (f > 0)) + 0.02 * (f > 0) * dR) @ R0   # not exactly orthonormal: irrelevant to the arithmetic
        poses.append((Rf.astype(np.float32), (t0 + 0.01 * f * dt).astype(np.float32)))
    cams = [pre.mod.frustum(Rf, tf, fx, fy, cx, cy, (0, W, 0, H), 0.0, 1.2, 8) for Rf, tf in poses]
    return gx, (counts, kps, desc), (P, nrm, maxd, mind, mdesc, observed), poses, cams, (fx, fy, cx, cy, W, H)


@pytest.mark.parametrize("nframes,nmp,th", [(6, 5000, 1.0), (3, 2500, 3.0)])
def test_search_local_points_batch(pre, orbx, oracle, synth, nframes, nmp, th):
    gx, (counts, kps, desc), (P, nrm, maxd, mind, mdesc, observed), poses, cams, (fx, fy, cx, cy, W, H) = \
        _local_map_case(orbx, synth, pre, nframes, nmp, 31)
    m = orbx.ORBmatcher(0.8)
    lm = orbx.LocalMap(m, P, nrm, maxd, mind, mdesc, observed)
    nm, kp_mp = m.SearchLocalPointsBatch(gx, nframes, cams, lm, th)
    scale = gx.GetScaleFactors()
    total = 0
    for f in range(nframes):
        Rf, tf = poses[f]
        cam = oracle.FrustumCam()
        cam.Rcw[:] = Rf.reshape(-1).tolist(); cam.tcw[:] = tf.tolist(); cam.Ow[:] = cams[f].Ow[:]   # the frame's mOw as frustum() forms it (Frame.cc:565)
        cam.fx, cam.fy, cam.cx, cam.cy = fx, fy, cx, cy
        cam.minX, cam.maxX, cam.minY, cam.maxY = 0, W, 0, H
        cam.bf, cam.logScaleFactor, cam.viewingCosLimit, cam.nScaleLevels, cam.logMode = 0.0, float(np.log(np.float32(1.2))), 0.5, 8, 1
        fr = oracle.is_in_frustum(cam, P, nrm, maxd, mind)
        c = int(counts[f])
        nm_o, kp_o = oracle.search_by_projection_mappoints(kps[f, :c], desc[f, :c], scale, W, H, fr["in_view"], fr["proj_x"], fr["proj_y"],
                                                           fr["level"], fr["view_cos"], mdesc, th, 0.8, observed=observed)
        assert nm[f] == nm_o and np.array_equal(kp_mp[f, :c], kp_o), f
        total += nm_o
    assert total > 300 * nframes // 2
    lm.close(), m.close(), gx.close()


def test_device_resident_forms_and_arguments(pre, orbx, oracle):
    """ORBX_MEM_DEVICE: device pointers in / out on the caller's stream (in place for the keypoints), plus argument checking."""
    import ctypes as C
    import torch
    lib = orbx.lib()
    cam = fc.CAMERAS["euroc"]
    fx, fy, cx, cy = [np.float32(v) for v in cam["K"]]
    pts = fc.pixel_points(cam, 3000, seed=8)
    kps = np.zeros(len(pts), orbx.KP_DTYPE)
    kps["x"], kps["y"], kps["octave"] = pts[:, 0], pts[:, 1], 3
    d_k = torch.from_numpy(kps.view(np.uint8).reshape(-1, 28).copy()).cuda()
    D = np.zeros(5, np.float32); D[:4] = cam["D"]
    st = torch.cuda.Stream()
    vp = lambda t: C.c_void_p(t.data_ptr())
    fl = lambda v: C.c_float(float(v))
    assert lib.orbx_undistort_keypoints(pre._h, vp(d_k), len(kps), fl(fx), fl(fy), fl(cx), fl(cy), D.ctypes.data_as(C.c_void_p), vp(d_k), 1,
                                        C.c_void_p(st.cuda_stream)) == 0
    st.synchronize()
    un = d_k.cpu().numpy().reshape(-1).view(orbx.KP_DTYPE)
    ref = oracle.undistort_points(pts, fx, fy, cx, cy, D)
    assert un["x"].tobytes() == ref[:, 0].tobytes() and un["y"].tobytes() == ref[:, 1].tobytes() and (un["octave"] == 3).all()
    # frustum on device buffers
    ocam, P, normal, maxd, mind = fc.frustum_case(oracle, 4000, seed=12)
    g = pre.mod.frustum(np.array(ocam.Rcw[:]).reshape(3, 3), ocam.tcw[:], ocam.fx, ocam.fy, ocam.cx, ocam.cy,
                        (ocam.minX, ocam.maxX, ocam.minY, ocam.maxY), ocam.bf, 1.2, 8, 0.5, Ow=ocam.Ow[:])
    n = len(P)
    dP, dN, dMx, dMn = [torch.from_numpy(a).cuda() for a in (P, normal, maxd, mind)]
    dIn = torch.empty(n, dtype=torch.uint8, device="cuda")
    dF = [torch.empty(n, dtype=torch.float32, device="cuda") for _ in range(5)]
    dL = torch.empty(n, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    assert lib.orbx_is_in_frustum(pre._h, C.byref(g), n, vp(dP), vp(dN), vp(dMx), vp(dMn), vp(dIn), vp(dF[0]), vp(dF[1]), vp(dF[2]), vp(dF[3]),
                                  vp(dL), vp(dF[4]), 1, C.c_void_p(st.cuda_stream)) == 0
    st.synchronize()
    ref = oracle.is_in_frustum(ocam, P, normal, maxd, mind)
    assert np.array_equal(dIn.cpu().numpy(), ref["in_view"]) and np.array_equal(dL.cpu().numpy(), ref["level"])
    for t, k in zip(dF, ("proj_x", "proj_y", "proj_xr", "depth", "view_cos")):
        assert t.cpu().numpy().tobytes() == ref[k].tobytes(), k
    # arguments
    assert lib.orbx_is_in_frustum(pre._h, C.byref(g), 0, None, None, None, None, None, None, None, None, None, None, None, 0, None) == 0
    assert lib.orbx_is_in_frustum(pre._h, C.byref(g), 5, None, None, None, None, None, None, None, None, None, None, None, 0, None) == orbx.ORBX_ERR_ARG
    assert lib.orbx_undistort_keypoints(pre._h, None, 3, fl(fx), fl(fy), fl(cx), fl(cy), D.ctypes.data_as(C.c_void_p), None, 0, None) == orbx.ORBX_ERR_ARG
    assert lib.orbx_undistort_keypoints(pre._h, vp(d_k), 3, fl(fx), fl(fy), fl(cx), fl(cy), D.ctypes.data_as(C.c_void_p), vp(d_k), 7, None) == orbx.ORBX_ERR_ARG
