"""ctypes binding of oracle/_ref/libref_match.so: the reference's OWN src/ORBmatcher.cc (+ ORBextractor.cc) compiled unmodified
against the stand-in cv:: header and data model (oracle/ref_shim/), plus the member functions of Frame / KeyFrame / MapPoint /
Pinhole / KannalaBrandt8 that oracle/Makefile cuts out of the reference's sources at build time.

TEST INFRASTRUCTURE ONLY (same rule as oracle/__init__.py).  Built by `make -C oracle ref` where /root/reference exists; the GPU
box uses the prebuilt file.  A View is the plain-array image of a Frame / KeyFrame, Points of a set of MapPoints.
"""
import ctypes as C
import os

import numpy as np

from . import KP_DTYPE
from . import ref as _ref

_LIB = os.path.join(_ref._DIR, "libref_match.so")
_TWIN = os.path.join(_ref._DIR, "libshim_match.so")
_lib = None
_twin = None
_use_twin = False


class twin:
    """`with refmatch.twin():` routes the calls of this module to oracle/_ref/libshim_match.so -- the same harness compiled
    against the product's host/ORBmatcher.h (liborbx.so, CUDA) instead of the reference's ORBmatcher.cc.  GPU tests only."""

    def __enter__(self):
        global _use_twin
        self.prev = _use_twin
        _use_twin = True
        return self

    def __exit__(self, *a):
        global _use_twin
        _use_twin = self.prev


def twin_available():
    return os.path.exists(_TWIN)


def available():
    if not os.path.exists(_LIB) and _ref.can_build():
        _ref.build()
    return os.path.exists(_LIB)


def lib():
    global _lib, _twin
    if _use_twin:
        if _twin is None:
            if not twin_available():
                raise RuntimeError("oracle/_ref/libshim_match.so missing: run `make -C oracle twin` in the build container")
            _twin = C.CDLL(_TWIN)
        return _twin
    if _lib is None:
        if not available():
            raise RuntimeError("oracle/_ref/libref_match.so missing: run `make -C oracle ref` in the build container")
        _lib = C.CDLL(_LIB)
        _lib.rm_stereo_create.restype = C.c_void_p
        _lib.rm_stereo_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        _lib.rm_stereo_destroy.argtypes = [C.c_void_p]
        _lib.rm_time_stereo.restype = C.c_double
    return _lib


class _View(C.Structure):
    _fields_ = [("N", C.c_int), ("Nleft", C.c_int), ("kps", C.c_void_p), ("kpsUn", C.c_void_p), ("kpsRight", C.c_void_p), ("nRight", C.c_int),
                ("desc", C.c_void_p), ("uRight", C.c_void_p), ("nlevels", C.c_int), ("scaleFactors", C.c_void_p),
                ("invScaleFactors", C.c_void_p), ("levelSigma2", C.c_void_p), ("invLevelSigma2", C.c_void_p), ("scaleFactor", C.c_float),
                ("logScaleFactor", C.c_float), ("minX", C.c_float), ("maxX", C.c_float), ("minY", C.c_float), ("maxY", C.c_float),
                ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float), ("bf", C.c_float), ("b", C.c_float),
                ("camModel", C.c_int), ("kb", C.c_float * 4), ("cam2Model", C.c_int), ("cam2", C.c_float * 8), ("Tcw", C.c_void_p),
                ("Tlr", C.c_void_p), ("mapPointIdx", C.c_void_p), ("outlier", C.c_void_p), ("fvNodes", C.c_int), ("fvIds", C.c_void_p),
                ("fvStart", C.c_void_p), ("fvIdx", C.c_void_p), ("leftToRight", C.c_void_p), ("rightToLeft", C.c_void_p)]


class _Points(C.Structure):
    _fields_ = [("n", C.c_int), ("pos", C.c_void_p), ("normal", C.c_void_p), ("desc", C.c_void_p), ("minDist", C.c_void_p),
                ("maxDist", C.c_void_p), ("nObs", C.c_void_p), ("bad", C.c_void_p)]


def _p(a):
    return a.ctypes.data_as(C.c_void_p).value if a is not None else None


def scale_tables(scale_factor=1.2, nlevels=8):
    """mvScaleFactors, mvInvScaleFactors, mvLevelSigma2, mvInvLevelSigma2 as ORBextractor::ORBextractor builds them (ORBextractor.cc:413-429)."""
    sf = np.empty(nlevels, np.float32); s2 = np.empty(nlevels, np.float32)
    sf[0] = 1.0; s2[0] = 1.0
    f = np.float64(np.float32(scale_factor))
    for i in range(1, nlevels):
        sf[i] = np.float32(np.float64(sf[i - 1]) * f)
        s2[i] = sf[i] * sf[i]
    return sf, (np.float32(1.0) / sf).astype(np.float32), s2, (np.float32(1.0) / s2).astype(np.float32)


class View:
    """Members of a Frame or KeyFrame.  kps = mvKeys (= mvKeysUn unless kps_un is given); bounds = (mnMinX, mnMaxX, mnMinY, mnMaxY)."""

    def __init__(self, kps, desc, bounds, K=(1.0, 1.0, 0.0, 0.0), bf=0.0, b=0.0, u_right=None, kps_un=None, scale_factor=1.2, nlevels=8,
                 Tcw=None, map_point_idx=None, outlier=None, feat_vec=None, kb=None, kps_right=None, n_left=-1, cam2=None, Tlr=None,
                 left_to_right=None, right_to_left=None):
        self.keep = []
        k = lambda a, dt: None if a is None else self._hold(np.ascontiguousarray(a, dt))
        v = _View()
        kps = k(kps, KP_DTYPE); desc = k(desc, np.uint8)
        nr = 0 if kps_right is None else len(kps_right)
        v.N = len(kps) + (nr if n_left != -1 else 0)
        v.Nleft = n_left
        assert desc is None or desc.reshape(-1, 32).shape[0] == v.N
        v.kps = _p(kps); v.kpsUn = _p(k(kps_un, KP_DTYPE)); v.kpsRight = _p(k(kps_right, KP_DTYPE)); v.nRight = nr
        v.desc = _p(desc); v.uRight = _p(k(u_right, np.float32))
        sf, isf, s2, is2 = scale_tables(scale_factor, nlevels)
        self.scale = sf
        v.nlevels = nlevels
        v.scaleFactors, v.invScaleFactors, v.levelSigma2, v.invLevelSigma2 = [_p(self._hold(a)) for a in (sf, isf, s2, is2)]
        v.scaleFactor = np.float32(scale_factor)
        v.logScaleFactor = np.log(np.float32(scale_factor))      # Frame.cc:110: log(float) -> logf
        v.minX, v.maxX, v.minY, v.maxY = [np.float32(x) for x in bounds]
        v.fx, v.fy, v.cx, v.cy = [np.float32(x) for x in K]
        v.bf, v.b = np.float32(bf), np.float32(b)
        v.camModel = 0 if kb is None else 1
        if kb is not None:
            v.kb[:] = [float(x) for x in kb]
        v.cam2Model = -1
        if cam2 is not None:                                      # (model, fx, fy, cx, cy, k1..k4)
            v.cam2Model = int(cam2[0])
            v.cam2[:] = [float(x) for x in cam2[1:9]]
        v.Tcw = _p(k(Tcw, np.float32)); v.Tlr = _p(k(Tlr, np.float32))
        v.mapPointIdx = _p(k(map_point_idx, np.int32)); v.outlier = _p(k(outlier, np.uint8))
        v.leftToRight = _p(k(left_to_right, np.int32)); v.rightToLeft = _p(k(right_to_left, np.int32))
        if feat_vec is not None:
            ids, start, idx = feat_vec
            v.fvNodes = len(ids)
            v.fvIds, v.fvStart, v.fvIdx = _p(k(ids, np.uint32)), _p(k(start, np.int32)), _p(k(idx, np.uint32))
        self.c = v
        self.N = v.N

    def _hold(self, a):
        self.keep.append(a)
        return a


class Points:
    """MapPoints as a structure of arrays: mWorldPos, mNormalVector, mDescriptor, RAW mfMinDistance / mfMaxDistance, Observations(), isBad()."""

    def __init__(self, pos, desc=None, normal=None, min_dist=None, max_dist=None, n_obs=None, bad=None):
        k = lambda a, dt: None if a is None else np.ascontiguousarray(a, dt)
        self.a = [k(pos, np.float32), k(normal, np.float32), k(desc, np.uint8), k(min_dist, np.float32), k(max_dist, np.float32),
                  k(n_obs, np.int32), k(bad, np.uint8)]
        p = _Points()
        p.n = len(self.a[0])
        p.pos, p.normal, p.desc, p.minDist, p.maxDist, p.nObs, p.bad = [_p(a) for a in self.a]
        self.c = p
        self.n = p.n


def _r(x):
    return C.byref(x.c) if x is not None else None


def features_in_area(view, x, y, r, min_level=-1, max_level=-1, right=False):
    out = np.empty(max(view.N, 1), np.int32)
    n = lib().rm_features_in_area(_r(view), C.c_float(x), C.c_float(y), C.c_float(r), min_level, max_level, int(right), C.c_void_p(_p(out)), len(out))
    return out[:n].copy()


def is_in_frustum(view, pts, viewing_cos_limit=0.5, right=False):
    n = pts.n
    o = {"in_view": np.empty(n, np.uint8), "proj_x": np.empty(n, np.float32), "proj_y": np.empty(n, np.float32), "proj_xr": np.empty(n, np.float32),
         "depth": np.empty(n, np.float32), "level": np.empty(n, np.int32), "view_cos": np.empty(n, np.float32)}
    r = {k + "_r": np.empty_like(v) for k, v in o.items()} if right else {}
    args = [C.c_void_p(_p(o[k])) for k in ("in_view", "proj_x", "proj_y", "proj_xr", "depth", "level", "view_cos")]
    args += [C.c_void_p(_p(r[k + "_r"])) if right else None for k in ("in_view", "proj_xr", "proj_y", "depth", "level", "view_cos")]
    lib().rm_is_in_frustum(_r(view), _r(pts), C.c_float(viewing_cos_limit), *args)
    o.update(r)
    return o


def search_by_projection_mappoints(view, pts, th, nnratio, fields=None, viewing_cos_limit=0.5, kp_mp=None, far_points=False, th_far=50.0):
    """fields = None: Frame::isInFrustum runs first (Tracking::SearchLocalPoints); else dict(in_view, proj_x, proj_y, level, view_cos[, proj_xr, depth])."""
    kp_mp = np.full(view.N, -1, np.int32) if kp_mp is None else np.ascontiguousarray(kp_mp, np.int32).copy()
    g = (lambda k, dt: None) if fields is None else (lambda k, dt: None if fields.get(k) is None else np.ascontiguousarray(fields[k], dt))
    a = [g("in_view", np.uint8), g("proj_x", np.float32), g("proj_y", np.float32), g("proj_xr", np.float32), g("level", np.int32),
         g("view_cos", np.float32), g("depth", np.float32)]
    nm = lib().rm_search_by_projection_mappoints(_r(view), _r(pts), int(fields is None), C.c_float(viewing_cos_limit), *[C.c_void_p(_p(x)) for x in a],
                                                 C.c_float(th), C.c_float(nnratio), int(far_points), C.c_float(th_far), C.c_void_p(_p(kp_mp)))
    return nm, kp_mp


def search_by_projection_mappoints_stereo(view, pts, th, nnratio, left, right, kp_mp=None, depth=None, far_points=False, th_far=50.0):
    """Fisheye-stereo frame (view.Nleft != -1): left / right = dict(in_view, proj_x, proj_y, level, view_cos) of the two cameras."""
    kp_mp = np.full(view.N, -1, np.int32) if kp_mp is None else np.ascontiguousarray(kp_mp, np.int32).copy()
    a = []
    for f in (left, right):
        a += [np.ascontiguousarray(f["in_view"], np.uint8), np.ascontiguousarray(f["proj_x"], np.float32), np.ascontiguousarray(f["proj_y"], np.float32),
              np.ascontiguousarray(f["level"], np.int32), np.ascontiguousarray(f["view_cos"], np.float32)]
    a.append(None if depth is None else np.ascontiguousarray(depth, np.float32))
    nm = lib().rm_search_by_projection_mappoints_stereo(_r(view), _r(pts), *[C.c_void_p(_p(x)) for x in a], C.c_float(th), C.c_float(nnratio),
                                                        int(far_points), C.c_float(th_far), C.c_void_p(_p(kp_mp)))
    return nm, kp_mp


def search_by_projection_frame(cur, last, pts, th, mono, nnratio=0.9, check_orientation=True):
    kp_mp = np.empty(cur.N, np.int32)
    nm = lib().rm_search_by_projection_frame(_r(cur), _r(last), _r(pts), C.c_float(th), int(mono), C.c_float(nnratio), int(check_orientation),
                                             C.c_void_p(_p(kp_mp)))
    return nm, kp_mp


def search_by_projection_keyframe(cur, kf, pts, already_found, th, orb_dist, nnratio=0.9, check_orientation=True):
    kp_mp = np.empty(cur.N, np.int32)
    af = None if already_found is None else np.ascontiguousarray(already_found, np.uint8)
    nm = lib().rm_search_by_projection_keyframe(_r(cur), _r(kf), _r(pts), C.c_void_p(_p(af)), C.c_float(th), int(orb_dist), C.c_float(nnratio),
                                                int(check_orientation), C.c_void_p(_p(kp_mp)))
    return nm, kp_mp


def search_by_projection_scw(kf, Scw, pts, th, ratio_hamming=1.0, matched=None, with_kfs=False):
    matched = np.full(kf.N, -1, np.int32) if matched is None else np.ascontiguousarray(matched, np.int32).copy()
    mkf = np.full(kf.N, -1, np.int32)
    Scw = np.ascontiguousarray(Scw, np.float32)
    nm = lib().rm_search_by_projection_scw(_r(kf), C.c_void_p(_p(Scw)), _r(pts), int(th), C.c_float(ratio_hamming), int(with_kfs),
                                           C.c_void_p(_p(matched)), C.c_void_p(_p(mkf)))
    return (nm, matched, mkf) if with_kfs else (nm, matched)


def search_for_initialization(v1, v2, prev_matched, window, nnratio=0.9, check_orientation=True):
    pm = np.ascontiguousarray(prev_matched, np.float32).copy()
    m12 = np.empty(v1.N, np.int32)
    nm = lib().rm_search_for_initialization(_r(v1), _r(v2), C.c_void_p(_p(pm)), int(window), C.c_float(nnratio), int(check_orientation),
                                            C.c_void_p(_p(m12)))
    return nm, m12, pm


def search_by_bow_frame(kf, frame, pts, nnratio=0.7, check_orientation=True):
    out = np.empty(frame.N, np.int32)
    nm = lib().rm_search_by_bow_frame(_r(kf), _r(frame), _r(pts), C.c_float(nnratio), int(check_orientation), C.c_void_p(_p(out)))
    return nm, out


def search_by_bow_keyframes(kf1, kf2, pts, nnratio=0.75, check_orientation=True):
    out = np.empty(kf1.N, np.int32)
    nm = lib().rm_search_by_bow_keyframes(_r(kf1), _r(kf2), _r(pts), C.c_float(nnratio), int(check_orientation), C.c_void_p(_p(out)))
    return nm, out


def search_for_triangulation(kf1, kf2, pts, F12, only_stereo=False, coarse=False, nnratio=0.6, check_orientation=True):
    cap = kf1.N
    pairs = np.empty((max(cap, 1), 2), np.int32)
    F12 = np.ascontiguousarray(F12, np.float32)
    nm = lib().rm_search_for_triangulation(_r(kf1), _r(kf2), _r(pts), C.c_void_p(_p(F12)), int(only_stereo), int(coarse), C.c_float(nnratio),
                                           int(check_orientation), C.c_void_p(_p(pairs)), cap)
    return nm, pairs[:nm].copy()


def kb8_unproject(cam, xy):
    """KannalaBrandt8::unproject, the reference's body (libm tanf)"""
    cam = np.ascontiguousarray(cam, np.float32).reshape(8); xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    rays = np.empty((xy.shape[0], 3), np.float32)
    lib().rm_kb8_unproject(C.c_void_p(_p(cam)), xy.shape[0], C.c_void_p(_p(xy)), C.c_void_p(_p(rays)))
    return rays


def kb8_triangulate_matches(cam1, cam2, R12, t12, xy1, xy2, sigma1, sigma2):
    """KannalaBrandt8::TriangulateMatches, the reference's body over the stand-in cv:: (z1 or -1, accepted point) per keypoint pair"""
    f = lambda a, shape: np.ascontiguousarray(a, np.float32).reshape(shape)
    cam1, cam2, R12, t12, xy1, xy2 = f(cam1, 8), f(cam2, 8), f(R12, 9), f(t12, 3), f(xy1, (-1, 2)), f(xy2, (-1, 2))
    n = xy1.shape[0]
    sigma1 = np.ascontiguousarray(np.broadcast_to(np.asarray(sigma1, np.float32), (n,)))
    sigma2 = np.ascontiguousarray(np.broadcast_to(np.asarray(sigma2, np.float32), (n,)))
    z1 = np.empty(n, np.float32); p3d = np.empty((n, 3), np.float32)
    lib().rm_kb8_triangulate_matches(*[C.c_void_p(_p(a)) for a in (cam1, cam2, R12, t12)], n, *[C.c_void_p(_p(a)) for a in (xy1, xy2, sigma1, sigma2, z1, p3d)])
    return z1, p3d


def cv_svd(A):
    A = np.ascontiguousarray(A, np.float32); n = A.shape[0]
    w = np.empty(n, np.float32); vt = np.empty((n, n), np.float32)
    lib().rm_cv_svd(C.c_void_p(_p(A)), n, C.c_void_p(_p(w)), C.c_void_p(_p(vt)))
    return w, vt


def cv_scaled_sub_row(a, s, b):
    a = np.ascontiguousarray(a, np.float32).ravel(); b = np.ascontiguousarray(b, np.float32).ravel()
    out = np.empty(a.size, np.float32)
    lib().rm_cv_scaled_sub_row(C.c_void_p(_p(a)), C.c_float(s), C.c_void_p(_p(b)), a.size, C.c_void_p(_p(out)))
    return out


def cv_div_scalar(a, s):
    a = np.ascontiguousarray(a, np.float32).ravel()
    out = np.empty(a.size, np.float32)
    lib().rm_cv_div_scalar(C.c_void_p(_p(a)), a.size, C.c_float(s), C.c_void_p(_p(out)))
    return out


def cv_gemm(a, b, alpha=1.0, c=None, beta=0.0, flags=0):
    a = np.ascontiguousarray(a, np.float32); b = np.ascontiguousarray(b, np.float32)
    M = a.shape[1] if flags & 1 else a.shape[0]; N = b.shape[0] if flags & 2 else b.shape[1]
    cc = None if c is None else np.ascontiguousarray(c, np.float32)
    out = np.empty((M, N), np.float32)
    lib().rm_cv_gemm(C.c_void_p(_p(a)), a.shape[0], a.shape[1], C.c_void_p(_p(b)), b.shape[0], b.shape[1], C.c_double(alpha),
                     C.c_void_p(_p(cc)) if cc is not None else None, C.c_double(beta), int(flags), C.c_void_p(_p(out)))
    return out


def cv_dot(a, b):
    a = np.ascontiguousarray(a, np.float32).ravel(); b = np.ascontiguousarray(b, np.float32).ravel()
    fn = lib().rm_cv_dot; fn.restype = C.c_double
    return fn(C.c_void_p(_p(a)), C.c_void_p(_p(b)), a.size)


def cv_norm(a):
    a = np.ascontiguousarray(a, np.float32).ravel()
    fn = lib().rm_cv_norm; fn.restype = C.c_double
    return fn(C.c_void_p(_p(a)), a.size)


def fuse(kf, pts, candidates, th=3.0, right=False):
    cand = np.ascontiguousarray(candidates, np.int32)
    ev = np.empty((2 * len(cand) + 4, 4), np.int32)
    ne = C.c_int(0)
    nf = lib().rm_fuse(_r(kf), _r(pts), C.c_void_p(_p(cand)), len(cand), C.c_float(th), int(right), C.c_void_p(_p(ev)), len(ev), C.byref(ne))
    return nf, ev[:ne.value].copy()


def fuse_scw(kf, Scw, pts, candidates, th):
    cand = np.ascontiguousarray(candidates, np.int32)
    ev = np.empty((2 * len(cand) + 4, 4), np.int32)
    rep = np.empty(len(cand), np.int32)
    ne = C.c_int(0)
    Scw = np.ascontiguousarray(Scw, np.float32)
    nf = lib().rm_fuse_scw(_r(kf), C.c_void_p(_p(Scw)), _r(pts), C.c_void_p(_p(cand)), len(cand), C.c_float(th), C.c_void_p(_p(rep)),
                           C.c_void_p(_p(ev)), len(ev), C.byref(ne))
    return nf, rep, ev[:ne.value].copy()


def search_by_sim3(kf1, kf2, pts, matches12, s12, R12, t12, th):
    m = np.ascontiguousarray(matches12, np.int32).copy()
    R12 = np.ascontiguousarray(R12, np.float32); t12 = np.ascontiguousarray(t12, np.float32)
    nm = lib().rm_search_by_sim3(_r(kf1), _r(kf2), _r(pts), C.c_float(s12), C.c_void_p(_p(R12)), C.c_void_p(_p(t12)),
                                 C.c_float(th), m.ctypes.data_as(C.c_void_p))
    return nm, m


class StereoFrames:
    """The data-parallel part of Frame::Frame(stereo) (Frame.cc:89-187): two reference ORBextractors + Frame::ComputeStereoMatches."""

    def __init__(self, nfeatures=1200, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7, blur_taps=0):
        self.nfeatures, self.nlevels, self.blur_taps = nfeatures, nlevels, blur_taps
        self._h = lib().rm_stereo_create(nfeatures, C.c_float(scaleFactor), nlevels, iniThFAST, minThFAST)

    def __del__(self):
        try:
            lib().rm_stereo_destroy(C.c_void_p(self._h))
        except Exception:
            pass

    def __call__(self, left, right, mb, mbf):
        left = np.ascontiguousarray(left, np.uint8); right = np.ascontiguousarray(right, np.uint8)
        h, w = left.shape
        cap = 4 * self.nfeatures + 64 * self.nlevels
        kl, kr = np.empty(cap, KP_DTYPE), np.empty(cap, KP_DTYPE)
        dl, dr = np.empty((cap, 32), np.uint8), np.empty((cap, 32), np.uint8)
        ur, dp = np.empty(cap, np.float32), np.empty(cap, np.float32)
        nl, nr = C.c_int(0), C.c_int(0)
        rc = lib().rm_stereo_frame(C.c_void_p(self._h), C.c_void_p(_p(left)), C.c_void_p(_p(right)), w, h, w, C.c_float(mb), C.c_float(mbf),
                                   self.blur_taps, C.c_void_p(_p(kl)), C.c_void_p(_p(dl)), C.byref(nl), C.c_void_p(_p(kr)), C.c_void_p(_p(dr)),
                                   C.byref(nr), C.c_void_p(_p(ur)), C.c_void_p(_p(dp)), cap)
        assert rc == 0, rc
        n, m = nl.value, nr.value
        return dict(kpsL=kl[:n].copy(), descL=dl[:n].copy(), kpsR=kr[:m].copy(), descR=dr[:m].copy(), uRight=ur[:n].copy(), depth=dp[:n].copy())


def time_stereo(left, right, nthreads, nfeatures, mb, mbf, blur_taps=0):
    """bench.py's reference arm: seconds for the reference code over frames[n,h,w] pairs, frame-parallel, 2 threads per frame."""
    left = np.ascontiguousarray(left, np.uint8); right = np.ascontiguousarray(right, np.uint8)
    n, h, w = left.shape
    matched = C.c_longlong(0)
    s = lib().rm_time_stereo(nfeatures, C.c_float(1.2), 8, 20, 7, blur_taps, C.c_void_p(_p(left)), C.c_void_p(_p(right)), n, w, h, C.c_float(mb),
                             C.c_float(mbf), nthreads, C.byref(matched))
    return float(s), int(matched.value)


def distinctive(desc, start, kind=None):
    """MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:350-436), the reference's own body, for a batch of map points:
    start = keyframe ranges per point, kind[k] = 0 left feature only / 1 bad keyframe / 2 left + right feature (two rows);
    desc rows in that order -> (mDescriptor[npoints, 32], has[npoints])"""
    desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    start = np.ascontiguousarray(start, np.int32)
    npoints = len(start) - 1
    kind = np.zeros(int(start[-1]), np.uint8) if kind is None else np.ascontiguousarray(kind, np.uint8)
    assert len(desc) == int(np.where(kind == 2, 2, 1).sum())
    out = np.zeros((npoints, 32), np.uint8)
    has = np.zeros(npoints, np.int32)
    lib().rm_distinctive(desc.ctypes.data_as(C.c_void_p), start.ctypes.data_as(C.c_void_p), kind.ctypes.data_as(C.c_void_p), npoints,
                         out.ctypes.data_as(C.c_void_p), has.ctypes.data_as(C.c_void_p))
    return out, has
