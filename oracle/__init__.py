"""ctypes binding of the CPU oracle (oracle/liborb_oracle.so).

TEST INFRASTRUCTURE ONLY: importable from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs.  The product package never imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "liborb_oracle.so")

TAPS_OCV346, TAPS_OCV4 = 0, 1
SINCOS_LIBM_F32, SINCOS_PORTABLE = 0, 1

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
assert KP_DTYPE.itemsize == 28


def _host_tag():
    """CPU model of this host: liborb_oracle.so is built -march=native, so a library built on another
    machine (the build container's .so travels to the GPU box) is rebuilt here instead of being reused."""
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("model name") or line.startswith("flags"):
                    if line.startswith("flags"):
                        return model + "|" + str(hash(line) & 0xFFFFFFFF)
                    model = line.split(":", 1)[1].strip()
    except Exception:
        pass
    return "unknown"


def build(force=False):
    srcs = [os.path.join(_HERE, f) for f in ("orb_oracle.cc", "match_oracle.cc", "oracle_capi.cc", "preproc_oracle.cc", "frame_oracle.cc", "orb_oracle.h",
                                             "match_oracle.h", "Makefile")]
    tagfile = _LIB + ".host"
    tag = _host_tag()
    try:
        built_for = open(tagfile).read()
    except OSError:
        built_for = ""
    if force or not os.path.exists(_LIB) or built_for != tag or any(os.path.getmtime(s) > os.path.getmtime(_LIB) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-B", "liborb_oracle.so", "ORACLE_MARCH=native"], stdout=subprocess.DEVNULL)
        with open(tagfile, "w") as f:
            f.write(tag)
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            build()
        _lib = C.CDLL(_LIB)
        _lib.oro_fast_atan2.restype = C.c_float
        _lib.oro_fast_atan2.argtypes = [C.c_float, C.c_float]
        _lib.oro_extractor_create.restype = C.c_void_p
        _lib.oro_extractor_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
        _lib.oro_extractor_destroy.argtypes = [C.c_void_p]
        _lib.oro_time_extract.restype = C.c_double
        _lib.oro_time_stereo.restype = C.c_double
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _u8(a):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    return a


def resize(src, dw, dh):
    src = _u8(src)
    dst = np.empty((dh, dw), np.uint8)
    lib().oro_resize(_p(src), src.shape[1], src.shape[0], src.strides[0], _p(dst), dw, dh, dw)
    return dst


def blur7(src, taps=TAPS_OCV4):
    src = _u8(src)
    dst = np.empty_like(src)
    lib().oro_blur7(_p(src), src.shape[1], src.shape[0], src.strides[0], _p(dst), dst.strides[0], taps)
    return dst


def fast(img, threshold, nonmax=True):
    """returns int32 array (n,3): x, y, response in cv::FAST order"""
    img = np.asarray(img, dtype=np.uint8)
    assert img.strides[1] == 1
    cap = img.shape[0] * img.shape[1]
    out = np.empty((max(cap, 1), 3), np.int32)
    n = lib().oro_fast(_p(img), img.shape[1], img.shape[0], img.strides[0], threshold, int(nonmax), _p(out), cap)
    return out[:n].copy()


def fast_atan2(y, x):
    return float(lib().oro_fast_atan2(C.c_float(y), C.c_float(x)))


def hamming(a, b):
    return lib().oro_hamming(_p(_u8(a)), _p(_u8(b)))


def knn2(q, t):
    q, t = _u8(q), _u8(t)
    idx = np.empty((q.shape[0], 2), np.int32)
    dist = np.empty((q.shape[0], 2), np.int32)
    lib().oro_knn2(_p(q), q.shape[0], _p(t), t.shape[0], _p(idx), _p(dist))
    return idx, dist


def knn2_ratio(q, t):
    q, t = _u8(q), _u8(t)
    idx = np.empty((q.shape[0], 2), np.int32)
    dist = np.empty((q.shape[0], 2), np.int32)
    good = np.empty(q.shape[0], np.uint8)
    lib().oro_knn2_ratio(_p(q), q.shape[0], _p(t), t.shape[0], _p(idx), _p(dist), _p(good))
    return idx, dist, good


def distribute(cands, minX, maxX, minY, maxY, N):
    """cands: (n,3) int32 x,y,resp relative to (minX,minY).  Returns kept candidate indices in the reference's order."""
    cands = np.ascontiguousarray(cands, dtype=np.int32)
    keep = np.empty(max(N + 16, 64) + 16, np.int32)
    n = lib().oro_distribute(_p(cands), cands.shape[0], minX, maxX, minY, maxY, N, _p(keep), keep.shape[0])
    assert n <= keep.shape[0]
    return keep[:n].copy()


def descriptor(blurred, x, y, angle, sincos_mode=SINCOS_PORTABLE):
    blurred = _u8(blurred)
    d = np.empty(32, np.uint8)
    lib().oro_descriptor(_p(blurred), blurred.strides[0], int(x), int(y), C.c_float(angle), sincos_mode, _p(d))
    return d


def three_maxima(hist):
    hist = np.ascontiguousarray(hist, np.int32)
    out = np.empty(3, np.int32)
    lib().oro_three_maxima(_p(hist), hist.shape[0], _p(out))
    return tuple(int(v) for v in out)


class Extractor:
    """Mirror of ORB_SLAM3::ORBextractor (include/ORBextractor.h:43-109) on numpy arrays."""

    def __init__(self, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7, blur_taps=TAPS_OCV346,
                 sincos_mode=SINCOS_PORTABLE):
        self.args = (nfeatures, float(scaleFactor), nlevels, iniThFAST, minThFAST, blur_taps, sincos_mode)
        self.nlevels = nlevels
        self.nfeatures = nfeatures
        self._h = lib().oro_extractor_create(nfeatures, C.c_float(scaleFactor), nlevels, iniThFAST, minThFAST, blur_taps,
                                             sincos_mode)
        self.scale = np.empty(nlevels, np.float32)
        self.inv_scale = np.empty(nlevels, np.float32)
        self.sigma2 = np.empty(nlevels, np.float32)
        self.inv_sigma2 = np.empty(nlevels, np.float32)
        self.quota = np.empty(nlevels, np.int32)
        self.umax = np.empty(16, np.int32)
        lib().oro_extractor_tables(C.c_void_p(self._h), _p(self.scale), _p(self.inv_scale), _p(self.sigma2),
                                   _p(self.inv_sigma2), _p(self.quota), _p(self.umax))

    def __del__(self):
        try:
            lib().oro_extractor_destroy(C.c_void_p(self._h))
        except Exception:
            pass

    def __call__(self, img, lapping=(0, 0)):
        """returns (monoIndex, keypoints[KP_DTYPE], descriptors[n,32])"""
        if img is None or img.size == 0:
            return -1, np.empty(0, KP_DTYPE), np.empty((0, 32), np.uint8)
        img = np.asarray(img, dtype=np.uint8)
        assert img.ndim == 2 and img.strides[1] == 1
        cap = 4 * self.nfeatures + 64 * self.nlevels
        kps = np.empty(cap, KP_DTYPE)
        desc = np.empty((cap, 32), np.uint8)
        n = C.c_int(0)
        mono = lib().oro_extract(C.c_void_p(self._h), _p(img), img.shape[1], img.shape[0], img.strides[0], int(lapping[0]),
                                 int(lapping[1]), _p(kps), _p(desc), cap, C.byref(n))
        assert n.value <= cap
        return mono, kps[:n.value].copy(), desc[:n.value].copy()

    def level(self, l):
        w, h = C.c_int(), C.c_int()
        lib().oro_level_size(C.c_void_p(self._h), l, C.byref(w), C.byref(h))
        out = np.empty((h.value, w.value), np.uint8)
        lib().oro_level_pixels(C.c_void_p(self._h), l, _p(out))
        return out

    def level_candidates(self, l):
        cap = 1 << 18
        out = np.empty((cap, 3), np.int32)
        n = lib().oro_level_candidates(C.c_void_p(self._h), l, _p(out), cap)
        return out[:n].copy()

    def level_keypoints(self, l):
        cap = 4 * self.nfeatures + 64
        out = np.empty(cap, KP_DTYPE)
        n = lib().oro_level_keypoints(C.c_void_p(self._h), l, _p(out), cap)
        return out[:n].copy()

    def fallback_cells(self, l):
        return lib().oro_level_fallback_cells(C.c_void_p(self._h), l)


def _bounds(w, h):
    return np.array([0.0, 0.0, float(w), float(h)], np.float32)


def features_in_area(kps, w, h, x, y, r, minLevel=-1, maxLevel=-1):
    kps = np.ascontiguousarray(kps, KP_DTYPE)
    out = np.empty(max(kps.shape[0], 1), np.int32)
    n = lib().oro_features_in_area(kps.shape[0], _p(kps), _p(_bounds(w, h)), C.c_float(x), C.c_float(y), C.c_float(r), minLevel,
                                   maxLevel, _p(out), out.shape[0])
    return out[:n].copy()


def search_by_projection_mappoints(kps, desc, scale_factors, w, h, in_view, projX, projY, level, viewCos, mpDesc, th, nnratio,
                                   kp_mp=None, observed=None, uRight=None, projXR=None):
    kps = np.ascontiguousarray(kps, KP_DTYPE)
    desc = _u8(desc)
    n = kps.shape[0]
    kp_mp = np.full(n, -1, np.int32) if kp_mp is None else np.ascontiguousarray(kp_mp, np.int32).copy()
    f32 = lambda a: None if a is None else np.ascontiguousarray(a, np.float32)
    u8 = lambda a: None if a is None else np.ascontiguousarray(a, np.uint8)
    sf = f32(scale_factors); in_view = u8(in_view); projX = f32(projX); projY = f32(projY)
    level = np.ascontiguousarray(level, np.int32); viewCos = f32(viewCos); mpDesc = _u8(mpDesc)
    observed = u8(observed); uRight = f32(uRight); projXR = f32(projXR)
    nm = lib().oro_search_by_projection_mappoints(n, _p(kps), _p(desc), _p(uRight), _p(sf), _p(_bounds(w, h)), in_view.shape[0],
                                                  _p(in_view), _p(projX), _p(projY), _p(level), _p(viewCos), _p(mpDesc),
                                                  _p(observed), _p(projXR), C.c_float(th), C.c_float(nnratio), _p(kp_mp))
    return nm, kp_mp


def search_by_projection_mappoints_stereo(kps_l, desc_l, kps_r, desc_r, scale_factors, w, h, left, right, mp_desc, l2r, r2l, th, nnratio,
                                          observed=None, kp_mp=None):
    """ORBmatcher::SearchByProjection(Frame, MapPoints) on a fisheye-stereo frame (ORBmatcher.cc:44-211); left / right =
    dict(in_view, proj_x, proj_y, level, view_cos) of the two cameras -> (nmatches, kp_mp[nL + nR])"""
    kps_l = np.ascontiguousarray(kps_l, KP_DTYPE); kps_r = np.ascontiguousarray(kps_r, KP_DTYPE)
    desc_l = _u8(desc_l); desc_r = _u8(desc_r)
    nl, nr = kps_l.shape[0], kps_r.shape[0]
    kp_mp = np.full(nl + nr, -1, np.int32) if kp_mp is None else np.ascontiguousarray(kp_mp, np.int32).copy()
    f32 = lambda a: np.ascontiguousarray(a, np.float32)
    u8 = lambda a: None if a is None else np.ascontiguousarray(a, np.uint8)
    i32 = lambda a: np.ascontiguousarray(a, np.int32)
    sf = f32(scale_factors)
    t = []
    for f in (left, right):
        t += [u8(f["in_view"]), f32(f["proj_x"]), f32(f["proj_y"]), i32(f["level"]), f32(f["view_cos"])]
    mp_desc = _u8(mp_desc); observed = u8(observed); l2r = i32(l2r); r2l = i32(r2l)
    nm = lib().oro_search_by_projection_mappoints_stereo(nl, _p(kps_l), _p(desc_l), nr, _p(kps_r), _p(desc_r), _p(sf), _p(_bounds(w, h)),
                                                         t[0].shape[0], *[_p(x) for x in t], _p(mp_desc), _p(observed), _p(l2r), _p(r2l),
                                                         C.c_float(th), C.c_float(nnratio), _p(kp_mp))
    return nm, kp_mp


def search_by_projection_window(kps, desc, scale_factors, w, h, valid, u, v, lvl, angle, srcDesc, th, mode=0, dist_th=100,
                                any_mp_blocks=False, check_orientation=True, kp_mp=None, observed=None, uRight=None, ur=None):
    kps = np.ascontiguousarray(kps, KP_DTYPE)
    desc = _u8(desc)
    n = kps.shape[0]
    kp_mp = np.full(n, -1, np.int32) if kp_mp is None else np.ascontiguousarray(kp_mp, np.int32).copy()
    f32 = lambda a: None if a is None else np.ascontiguousarray(a, np.float32)
    u8 = lambda a: None if a is None else np.ascontiguousarray(a, np.uint8)
    sf = f32(scale_factors); valid = u8(valid); u = f32(u); v = f32(v); lvl = np.ascontiguousarray(lvl, np.int32)
    angle = f32(angle); srcDesc = _u8(srcDesc); observed = u8(observed); uRight = f32(uRight); ur = f32(ur)
    nm = lib().oro_search_by_projection_window(n, _p(kps), _p(desc), _p(uRight), _p(sf), _p(_bounds(w, h)), valid.shape[0],
                                               _p(valid), _p(u), _p(v), _p(lvl), _p(ur), _p(angle), _p(srcDesc), _p(observed),
                                               C.c_float(th), mode, dist_th, int(any_mp_blocks), int(check_orientation), _p(kp_mp))
    return nm, kp_mp


def search_for_initialization(kps1, desc1, kps2, desc2, w, h, prev_matched, window, nnratio=0.9, check_orientation=True):
    kps1 = np.ascontiguousarray(kps1, KP_DTYPE); kps2 = np.ascontiguousarray(kps2, KP_DTYPE)
    desc1, desc2 = _u8(desc1), _u8(desc2)
    pm = np.ascontiguousarray(prev_matched, np.float32).copy()
    m12 = np.empty(kps1.shape[0], np.int32)
    nm = lib().oro_search_for_initialization(kps1.shape[0], _p(kps1), _p(desc1), kps2.shape[0], _p(kps2), _p(desc2),
                                             _p(_bounds(w, h)), _p(pm), int(window), C.c_float(nnratio), int(check_orientation),
                                             _p(m12))
    return nm, m12, pm


def _fv(fv):
    """fv = (node_ids uint32 ascending, node_start int32 [nnodes+1], indices uint32)"""
    ids = np.ascontiguousarray(fv[0], np.uint32); start = np.ascontiguousarray(fv[1], np.int32)
    ind = np.ascontiguousarray(fv[2], np.uint32)
    return ids, start, ind


def search_by_bow(kps1, desc1, fv1, elig1, kps2, desc2, fv2, elig2, max_dist, nnratio, check_orientation=True):
    kps1 = np.ascontiguousarray(kps1, KP_DTYPE); kps2 = np.ascontiguousarray(kps2, KP_DTYPE)
    desc1, desc2 = _u8(desc1), _u8(desc2)
    i1, s1, x1 = _fv(fv1); i2, s2, x2 = _fv(fv2)
    elig1 = _u8(elig1); elig2 = None if elig2 is None else _u8(elig2)
    m12 = np.empty(max(kps1.shape[0], 1), np.int32)
    nm = lib().oro_search_by_bow(kps1.shape[0], _p(kps1), _p(desc1), i1.shape[0], _p(i1), _p(s1), _p(x1), _p(elig1), kps2.shape[0],
                                 _p(kps2), _p(desc2), i2.shape[0], _p(i2), _p(s2), _p(x2), _p(elig2), int(max_dist),
                                 C.c_float(nnratio), int(check_orientation), _p(m12))
    return nm, m12[:kps1.shape[0]]


def search_by_bow_stereo(kps1, desc1, fv1, elig1, kps2, desc2, fv2, n_left2, max_dist, nnratio, check_orientation=True):
    """SearchByBoW(KeyFrame, Frame) with a fisheye-stereo frame on side 2 (ORBmatcher.cc:338-431) -> (nmatches, matches12_left, matches12_right)"""
    kps1 = np.ascontiguousarray(kps1, KP_DTYPE); kps2 = np.ascontiguousarray(kps2, KP_DTYPE)
    desc1, desc2 = _u8(desc1), _u8(desc2)
    i1, s1, x1 = _fv(fv1); i2, s2, x2 = _fv(fv2)
    elig1 = _u8(elig1)
    mL = np.empty(max(kps1.shape[0], 1), np.int32); mR = np.empty(max(kps1.shape[0], 1), np.int32)
    nm = lib().oro_search_by_bow_stereo(kps1.shape[0], _p(kps1), _p(desc1), i1.shape[0], _p(i1), _p(s1), _p(x1), _p(elig1), kps2.shape[0],
                                        int(n_left2), _p(kps2), _p(desc2), i2.shape[0], _p(i2), _p(s2), _p(x2), int(max_dist),
                                        C.c_float(nnratio), int(check_orientation), _p(mL), _p(mR))
    return nm, mL[:kps1.shape[0]], mR[:kps1.shape[0]]


def search_for_triangulation(kps1, desc1, fv1, elig1, stereo1, kps2, desc2, fv2, elig2, stereo2, F12, ep, scale_factors2,
                             level_sigma2_2, coarse=False, check_orientation=True):
    kps1 = np.ascontiguousarray(kps1, KP_DTYPE); kps2 = np.ascontiguousarray(kps2, KP_DTYPE)
    desc1, desc2 = _u8(desc1), _u8(desc2)
    i1, s1, x1 = _fv(fv1); i2, s2, x2 = _fv(fv2)
    u8 = lambda a: None if a is None else np.ascontiguousarray(a, np.uint8)
    elig1, elig2, stereo1, stereo2 = u8(elig1), u8(elig2), u8(stereo1), u8(stereo2)
    F12 = np.ascontiguousarray(F12, np.float32).reshape(9)
    sf = np.ascontiguousarray(scale_factors2, np.float32); ls = np.ascontiguousarray(level_sigma2_2, np.float32)
    m12 = np.empty(max(kps1.shape[0], 1), np.int32)
    nm = lib().oro_search_for_triangulation(kps1.shape[0], _p(kps1), _p(desc1), i1.shape[0], _p(i1), _p(s1), _p(x1), _p(elig1),
                                            _p(stereo1), kps2.shape[0], _p(kps2), _p(desc2), i2.shape[0], _p(i2), _p(s2), _p(x2),
                                            _p(elig2), _p(stereo2), _p(F12), C.c_float(ep[0]), C.c_float(ep[1]), _p(sf), _p(ls),
                                            int(coarse), int(check_orientation), _p(m12))
    return nm, m12[:kps1.shape[0]]


LIBM_F32, LIBM_PORTABLE = 0, 1


def kb8_pair(R12, t12, cam1, cam2):
    """One (camera of KF1, camera of KF2) geometry record of the KannalaBrandt8 gate: R12[9] t12[3] cam1[8] cam2[8] (cam = fx fy cx cy k1..k4)"""
    return np.concatenate([np.asarray(R12, np.float32).reshape(9), np.asarray(t12, np.float32).reshape(3),
                           np.asarray(cam1, np.float32).reshape(8), np.asarray(cam2, np.float32).reshape(8)])


def search_for_triangulation_kb8(kps1, desc1, fv1, elig1, stereo1, n_left1, kps2, desc2, fv2, elig2, stereo2, n_left2, pairs, ep,
                                 scale_factors2, level_sigma2_1, level_sigma2_2, coarse=False, check_orientation=True,
                                 libm_mode=LIBM_PORTABLE):
    """ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:981-1222) on KannalaBrandt8 keyframes: pairs = [kb8_pair(..)] (n_left = -1) or
    the four records ll, lr, rl, rr of a fisheye-stereo rig (kps = mvKeys followed by mvKeysRight)."""
    kps1 = np.ascontiguousarray(kps1, KP_DTYPE); kps2 = np.ascontiguousarray(kps2, KP_DTYPE)
    desc1, desc2 = _u8(desc1), _u8(desc2)
    i1, s1, x1 = _fv(fv1); i2, s2, x2 = _fv(fv2)
    u8 = lambda a: None if a is None else np.ascontiguousarray(a, np.uint8)
    elig1, elig2, stereo1, stereo2 = u8(elig1), u8(elig2), u8(stereo1), u8(stereo2)
    G = np.ascontiguousarray(np.stack(pairs), np.float32)
    assert G.shape == ((4 if n_left1 != -1 else 1), 28) and (n_left1 == -1) == (n_left2 == -1)
    sf = np.ascontiguousarray(scale_factors2, np.float32)
    l1 = np.ascontiguousarray(level_sigma2_1, np.float32); l2 = np.ascontiguousarray(level_sigma2_2, np.float32)
    m12 = np.empty(max(kps1.shape[0], 1), np.int32)
    nm = lib().oro_search_for_triangulation_kb8(kps1.shape[0], _p(kps1), _p(desc1), i1.shape[0], _p(i1), _p(s1), _p(x1), _p(elig1),
                                                _p(stereo1), int(n_left1), kps2.shape[0], _p(kps2), _p(desc2), i2.shape[0], _p(i2),
                                                _p(s2), _p(x2), _p(elig2), _p(stereo2), int(n_left2), _p(G), C.c_float(ep[0]),
                                                C.c_float(ep[1]), _p(sf), _p(l1), _p(l2), int(coarse), int(check_orientation),
                                                int(libm_mode), _p(m12))
    return nm, m12[:kps1.shape[0]]


def kb8_triangulate_matches(pair, xy1, xy2, sigma1, sigma2, libm_mode=LIBM_PORTABLE):
    """KannalaBrandt8::TriangulateMatches (KannalaBrandt8.cpp:343-418) for n independent keypoint pairs: (z1 or -1, point) per pair"""
    pair = np.ascontiguousarray(pair, np.float32).reshape(28)
    xy1 = np.ascontiguousarray(xy1, np.float32).reshape(-1, 2); xy2 = np.ascontiguousarray(xy2, np.float32).reshape(-1, 2)
    n = xy1.shape[0]
    sigma1 = np.ascontiguousarray(np.broadcast_to(np.asarray(sigma1, np.float32), (n,)))
    sigma2 = np.ascontiguousarray(np.broadcast_to(np.asarray(sigma2, np.float32), (n,)))
    z1 = np.empty(n, np.float32); p3d = np.empty((n, 3), np.float32)
    lib().oro_kb8_triangulate_matches(_p(pair), n, _p(xy1), _p(xy2), _p(sigma1), _p(sigma2), int(libm_mode), _p(z1), _p(p3d))
    return z1, p3d


def kb8_unproject(cam, xy, libm_mode=LIBM_PORTABLE):
    cam = np.ascontiguousarray(cam, np.float32).reshape(8)
    xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    rays = np.empty((xy.shape[0], 3), np.float32)
    lib().oro_kb8_unproject(_p(cam), xy.shape[0], _p(xy), int(libm_mode), _p(rays))
    return rays


def kb8_project(cam, X, libm_mode=LIBM_PORTABLE):
    cam = np.ascontiguousarray(cam, np.float32).reshape(8)
    X = np.ascontiguousarray(X, np.float32).reshape(-1, 3)
    uv = np.empty((X.shape[0], 2), np.float32)
    lib().oro_kb8_project(_p(cam), X.shape[0], _p(X), int(libm_mode), _p(uv))
    return uv


def jacobi_svd4(A):
    """cv::SVD::compute on 4x4 float32 matrices (n x 4 x 4): (w n x 4, vt n x 4 x 4)"""
    A = np.ascontiguousarray(A, np.float32).reshape(-1, 4, 4)
    w = np.empty((A.shape[0], 4), np.float32); vt = np.empty((A.shape[0], 4, 4), np.float32)
    lib().oro_jacobi_svd4(_p(A), A.shape[0], _p(w), _p(vt))
    return w, vt


def search_window_best(kps, desc, w, h, valid, u, v, radius, level, qdesc, gate=0, inv_level_sigma2=None, uRight=None, ur=None):
    kps = np.ascontiguousarray(kps, KP_DTYPE); desc = _u8(desc)
    f32 = lambda a: None if a is None else np.ascontiguousarray(a, np.float32)
    valid = _u8(valid); u = f32(u); v = f32(v); radius = f32(radius); level = np.ascontiguousarray(level, np.int32)
    qdesc = _u8(qdesc); ils = f32(inv_level_sigma2); uRight = f32(uRight); ur = f32(ur)
    n = valid.shape[0]
    bi = np.empty(max(n, 1), np.int32); bd = np.empty(max(n, 1), np.int32)
    lib().oro_search_window_best(kps.shape[0], _p(kps), _p(desc), _p(uRight), _p(_bounds(w, h)), n, _p(valid), _p(u), _p(v),
                                 _p(radius), _p(level), _p(ur), _p(qdesc), int(gate), _p(ils), _p(bi), _p(bd))
    return bi[:n], bd[:n]


def distinctive_descriptors(desc, start):
    """MapPoint::ComputeDistinctiveDescriptors for a batch: -> (best_idx[npoints], best_median[npoints])"""
    desc = _u8(desc); start = np.ascontiguousarray(start, np.int32)
    npts = start.shape[0] - 1
    bi = np.empty(max(npts, 1), np.int32); bm = np.empty(max(npts, 1), np.int32)
    lib().oro_distinctive_descriptors(_p(desc), _p(start), npts, _p(bi), _p(bm))
    return bi[:npts], bm[:npts]


def vocab_transform(vocab, desc, levelsup):
    """vocab = dict(L, child_start, child_ids, node_desc, node_weight, node_word_id) -> (word_id, node_id, weight) per feature"""
    desc = _u8(desc).reshape(-1, 32)
    cs = np.ascontiguousarray(vocab["child_start"], np.int32); ci = np.ascontiguousarray(vocab["child_ids"], np.int32)
    nd = _u8(vocab["node_desc"]); nw = np.ascontiguousarray(vocab["node_weight"], np.float64)
    wi = np.ascontiguousarray(vocab["node_word_id"], np.int32)
    n = desc.shape[0]
    w = np.empty(max(n, 1), np.int32); nid = np.empty(max(n, 1), np.int32); wt = np.empty(max(n, 1), np.float64)
    lib().oro_vocab_transform(cs.shape[0] - 1, int(vocab["L"]), _p(cs), _p(ci), _p(nd), _p(nw), _p(wi), _p(desc), n, int(levelsup),
                              _p(w), _p(nid), _p(wt))
    return w[:n], nid[:n], wt[:n]


def vocab_fold(word_id, node_id, weight):
    """per-feature (word, node, weight) -> (bow_ids, bow_values, (fv_node_ids, fv_start, fv_indices)); TF_IDF + L1"""
    word_id = np.ascontiguousarray(word_id, np.int32); node_id = np.ascontiguousarray(node_id, np.int32)
    weight = np.ascontiguousarray(weight, np.float64)
    n = word_id.shape[0]
    bi = np.empty(max(n, 1), np.uint32); bv = np.empty(max(n, 1), np.float64)
    fi = np.empty(max(n, 1), np.uint32); fs = np.empty(n + 1, np.int32); fx = np.empty(max(n, 1), np.uint32)
    nn = C.c_int(0)
    k = lib().oro_vocab_fold(_p(word_id), _p(node_id), _p(weight), n, _p(bi), _p(bv), _p(fi), _p(fs), _p(fx), C.byref(nn))
    return bi[:k].copy(), bv[:k].copy(), (fi[:nn.value].copy(), fs[:nn.value + 1].copy(), fx[:fs[nn.value]].copy())


def stereo_matches(exL, exR, kpL, descL, kpR, descR, mb, mbf):
    """exL / exR: oracle Extractor objects that have just extracted the left / right image."""
    kpL = np.ascontiguousarray(kpL, KP_DTYPE); kpR = np.ascontiguousarray(kpR, KP_DTYPE)
    descL, descR = _u8(descL), _u8(descR)
    n = kpL.shape[0]
    uR = np.empty(n, np.float32); depth = np.empty(n, np.float32)
    bd = np.empty(n, np.int32); bi = np.empty(n, np.int32)
    lib().oro_stereo_matches(C.c_void_p(exL._h), C.c_void_p(exR._h), n, _p(kpL), _p(descL), kpR.shape[0], _p(kpR), _p(descR),
                             C.c_float(mb), C.c_float(mbf), _p(uR), _p(depth), _p(bd), _p(bi))
    return uR, depth, bd, bi


def time_stereo(left, right, nthreads, nfeatures, mb, mbf, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7,
                blur_taps=TAPS_OCV346, sincos_mode=SINCOS_LIBM_F32):
    """whole stereo frames (extract L+R on two threads per frame + ComputeStereoMatches).  Returns (seconds, matched keypoints)."""
    left, right = _u8(left), _u8(right)
    n, h, w = left.shape
    tot = C.c_longlong(0)
    s = lib().oro_time_stereo(nfeatures, C.c_float(scaleFactor), nlevels, iniThFAST, minThFAST, blur_taps, sincos_mode, _p(left),
                              _p(right), n, w, h, C.c_float(mb), C.c_float(mbf), int(nthreads), C.byref(tot))
    return float(s), int(tot.value)


def time_extract(frames, nthreads, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7,
                 blur_taps=TAPS_OCV346, sincos_mode=SINCOS_LIBM_F32, lapping=(0, 0)):
    """frames: (n,h,w) uint8.  Returns (seconds, total keypoints)."""
    frames = _u8(frames)
    n, h, w = frames.shape
    tot = C.c_longlong(0)
    s = lib().oro_time_extract(nfeatures, C.c_float(scaleFactor), nlevels, iniThFAST, minThFAST, blur_taps, sincos_mode,
                               _p(frames), n, w, h, int(lapping[0]), int(lapping[1]), int(nthreads), C.byref(tot))
    return float(s), int(tot.value)


# ---- pre-processing before the extractor (oracle/preproc_oracle.cc) -------------------------------
GRAY_OCV346, GRAY_OCV4 = 0, 1


def cvt_gray(src, rgb_order=False, coeffs=GRAY_OCV4):
    """cvtColor(src, CV_{RGB,BGR,RGBA,BGRA}2GRAY) -- src/Tracking.cc:989-1014."""
    src = _u8(src)
    h, w, c = src.shape
    dst = np.empty((h, w), np.uint8)
    lib().oro_cvt_gray(_p(src), w, h, w * c, c, int(bool(rgb_order)), int(coeffs), _p(dst), w)
    return dst


def remap_linear(src, mapx, mapy):
    """cv::remap(src, dst, M1, M2, INTER_LINEAR) -- Examples/Stereo/stereo_euroc.cc:166-167."""
    src = _u8(src)
    mapx = np.ascontiguousarray(mapx, np.float32)
    mapy = np.ascontiguousarray(mapy, np.float32)
    dh, dw = mapx.shape
    dst = np.empty((dh, dw), np.uint8)
    lib().oro_remap_linear(_p(src), src.shape[1], src.shape[0], src.shape[1], _p(mapx), _p(mapy), dw, _p(dst), dw, dh, dw)
    return dst


def clahe(src, clip_limit=3.0, tiles=(8, 8)):
    """cv::createCLAHE(clip_limit, tiles)->apply(src, dst) -- Examples/Monocular/mono_tum_vi.cc:101,109."""
    src = _u8(src)
    h, w = src.shape
    dst = np.empty((h, w), np.uint8)
    lib().oro_clahe(_p(src), w, h, w, C.c_double(clip_limit), int(tiles[0]), int(tiles[1]), _p(dst), w)
    return dst


# ---- Frame glue (oracle/frame_oracle.cc) -----------------------------------------------------------
LOG_LIBM_F32, LOG_PORTABLE = 0, 1


class FrustumCam(C.Structure):
    _fields_ = [("Rcw", C.c_float * 9), ("tcw", C.c_float * 3), ("Ow", C.c_float * 3), ("fx", C.c_float), ("fy", C.c_float),
                ("cx", C.c_float), ("cy", C.c_float), ("minX", C.c_float), ("maxX", C.c_float), ("minY", C.c_float),
                ("maxY", C.c_float), ("bf", C.c_float), ("logScaleFactor", C.c_float), ("viewingCosLimit", C.c_float),
                ("nScaleLevels", C.c_int), ("logMode", C.c_int), ("cameraModel", C.c_int), ("kb", C.c_float * 4)]


def undistort_points(pts, fx, fy, cx, cy, dist):
    """cv::undistortPoints(mat, mat, K, D, Mat(), K) -- src/Frame.cc:856, 884."""
    pts = np.ascontiguousarray(pts, np.float32).reshape(-1, 2)
    d = np.zeros(5, np.float32)
    d[:len(dist)] = dist
    out = np.empty_like(pts)
    lib().oro_undistort_points(_p(pts), len(pts), C.c_float(fx), C.c_float(fy), C.c_float(cx), C.c_float(cy), _p(d), _p(out))
    return out


def is_in_frustum(cam, P, normal, max_dist, min_dist):
    """Frame::isInFrustum (src/Frame.cc:572-642) over n map points; cam is a FrustumCam."""
    P = np.ascontiguousarray(P, np.float32).reshape(-1, 3)
    normal = np.ascontiguousarray(normal, np.float32).reshape(-1, 3)
    max_dist = np.ascontiguousarray(max_dist, np.float32)
    min_dist = np.ascontiguousarray(min_dist, np.float32)
    n = len(P)
    out = dict(in_view=np.empty(n, np.uint8), proj_x=np.empty(n, np.float32), proj_y=np.empty(n, np.float32),
               proj_xr=np.empty(n, np.float32), depth=np.empty(n, np.float32), level=np.empty(n, np.int32),
               view_cos=np.empty(n, np.float32))
    lib().oro_is_in_frustum(C.byref(cam), n, _p(P), _p(normal), _p(max_dist), _p(min_dist), _p(out["in_view"]), _p(out["proj_x"]),
                            _p(out["proj_y"]), _p(out["proj_xr"]), _p(out["depth"]), _p(out["level"]), _p(out["view_cos"]))
    return out
