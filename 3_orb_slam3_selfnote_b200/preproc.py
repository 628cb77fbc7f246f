"""ctypes mirror of the pre-processing entry points of include/orbx.h, named after the OpenCV calls
the reference makes before ORBextractor::operator():

    cvtColor(im, CV_{RGB,BGR,RGBA,BGRA}2GRAY)         src/Tracking.cc:989-1014, 1059-1073, 1121-1135
    remap(im, M1, M2, INTER_LINEAR)                   Examples/Stereo/stereo_euroc.cc:117-118, 166-167
    createCLAHE(3.0, (8, 8)).apply(im)                Examples/Monocular/mono_tum_vi.cc:101, 109

Host arrays (numpy) in / out through ORBX_MEM_HOST; the *_device methods take raw device pointers and enqueue on a stream.
No CPU fallback: everything runs in liborbx.so's kernels.
"""
import ctypes as C

import numpy as np

from . import _check, _p, lib

MEM_HOST, MEM_DEVICE = 0, 1
GRAY_OCV346, GRAY_OCV4 = 0, 1
COLOR_BGR2GRAY, COLOR_RGB2GRAY, COLOR_BGRA2GRAY, COLOR_RGBA2GRAY = "bgr", "rgb", "bgra", "rgba"
_LL = C.c_longlong


def _frames(a, ndim):
    """(frames?, h, w[, c]) uint8 array -> contiguous batch + whether a batch axis was given."""
    a = np.ascontiguousarray(a, dtype=np.uint8)
    if a.ndim == ndim:
        return a[None], False
    if a.ndim != ndim + 1:
        raise ValueError("expected %d or %d dimensions" % (ndim, ndim + 1))
    return a, True


class RemapTable:
    def __init__(self, owner, mapx, mapy, src_size):
        mapx = np.ascontiguousarray(mapx, np.float32)
        mapy = np.ascontiguousarray(mapy, np.float32)
        assert mapx.shape == mapy.shape and mapx.ndim == 2
        self.dh, self.dw = mapx.shape
        self.sw, self.sh = int(src_size[0]), int(src_size[1])
        self._t = C.c_void_p()
        _check(lib().orbx_remap_create(owner._h, _p(mapx), _p(mapy), self.dw, self.dh, self.dw, self.sw, self.sh, C.byref(self._t)))

    def close(self):
        if getattr(self, "_t", None) and self._t.value:
            lib().orbx_remap_destroy(self._t)
            self._t = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Preprocessor:
    def __init__(self, device=0):
        self._h = C.c_void_p()
        _check(lib().orbx_preproc_create(int(device), C.byref(self._h)))
        self.device = device

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            lib().orbx_preproc_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- cvtColor ---------------------------------------------------------------------------------------------------------
    def cvtColor(self, src, code, coeffs=GRAY_OCV4):
        src, batched = _frames(src, 3)
        n, h, w, c = src.shape
        if {"bgr": 3, "rgb": 3, "bgra": 4, "rgba": 4}[code] != c:
            raise ValueError("channel count does not match the conversion code")
        dst = np.empty((n, h, w), np.uint8)
        _check(lib().orbx_cvt_gray(self._h, _p(src), n, w, h, w * c, _LL(h * w * c), c, int(code[0] == "r"), int(coeffs), _p(dst), w,
                                   _LL(h * w), MEM_HOST, None))
        return dst if batched else dst[0]

    def cvt_gray_device(self, d_src, n, w, h, src_stride, src_frame_stride, channels, rgb_order, d_dst, dst_stride, dst_frame_stride,
                        stream=None, coeffs=GRAY_OCV4):
        _check(lib().orbx_cvt_gray(self._h, C.c_void_p(d_src), n, w, h, src_stride, _LL(src_frame_stride), channels, int(rgb_order),
                                   int(coeffs), C.c_void_p(d_dst), dst_stride, _LL(dst_frame_stride), MEM_DEVICE,
                                   C.c_void_p(stream or 0)))

    # -- remap ------------------------------------------------------------------------------------------------------------
    def remap_table(self, mapx, mapy, src_size):
        """src_size = (width, height) of the images that will be remapped."""
        return RemapTable(self, mapx, mapy, src_size)

    def remap(self, src, table):
        src, batched = _frames(src, 2)
        n, h, w = src.shape
        if (w, h) != (table.sw, table.sh):
            raise ValueError("image size differs from the table's source size")
        dst = np.empty((n, table.dh, table.dw), np.uint8)
        _check(lib().orbx_remap(self._h, table._t, _p(src), n, w, _LL(h * w), _p(dst), table.dw, _LL(table.dh * table.dw), MEM_HOST, None))
        return dst if batched else dst[0]

    def remap_device(self, table, d_src, n, src_stride, src_frame_stride, d_dst, dst_stride, dst_frame_stride, stream=None):
        _check(lib().orbx_remap(self._h, table._t, C.c_void_p(d_src), n, src_stride, _LL(src_frame_stride), C.c_void_p(d_dst), dst_stride,
                                _LL(dst_frame_stride), MEM_DEVICE, C.c_void_p(stream or 0)))

    # -- CLAHE ------------------------------------------------------------------------------------------------------------
    def clahe(self, src, clip_limit=3.0, tiles=(8, 8)):
        src, batched = _frames(src, 2)
        n, h, w = src.shape
        dst = np.empty((n, h, w), np.uint8)
        _check(lib().orbx_clahe(self._h, _p(src), n, w, h, w, _LL(h * w), C.c_double(clip_limit), int(tiles[0]), int(tiles[1]), _p(dst), w,
                                _LL(h * w), MEM_HOST, None))
        return dst if batched else dst[0]

    def clahe_device(self, d_src, n, w, h, src_stride, src_frame_stride, d_dst, dst_stride, dst_frame_stride, clip_limit=3.0, tiles=(8, 8),
                     stream=None):
        _check(lib().orbx_clahe(self._h, C.c_void_p(d_src), n, w, h, src_stride, _LL(src_frame_stride), C.c_double(clip_limit),
                                int(tiles[0]), int(tiles[1]), C.c_void_p(d_dst), dst_stride, _LL(dst_frame_stride), MEM_DEVICE,
                                C.c_void_p(stream or 0)))


# ---- Frame glue ---------------------------------------------------------------------------------
class _Frustum(C.Structure):
    _fields_ = [("Rcw", C.c_float * 9), ("tcw", C.c_float * 3), ("Ow", C.c_float * 3), ("fx", C.c_float), ("fy", C.c_float),
                ("cx", C.c_float), ("cy", C.c_float), ("min_x", C.c_float), ("max_x", C.c_float), ("min_y", C.c_float),
                ("max_y", C.c_float), ("bf", C.c_float), ("log_scale_factor", C.c_float), ("viewing_cos_limit", C.c_float),
                ("n_scale_levels", C.c_int), ("camera_model", C.c_int), ("kb", C.c_float * 4)]


def _mm3(A, B):
    """3x3 (or 3x3 * 3x1) float32 product in the order of OpenCV's small-matrix gemm: ((a0*b0 + a1*b1) + a2*b2), every step rounded to float"""
    A = np.asarray(A, np.float32); B = np.asarray(B, np.float32).reshape(3, -1)
    out = np.empty((3, B.shape[1]), np.float32)
    for i in range(3):
        for j in range(B.shape[1]):
            out[i, j] = np.float32(np.float32(np.float32(A[i, 0] * B[0, j]) + np.float32(A[i, 1] * B[1, j])) + np.float32(A[i, 2] * B[2, j]))
    return out


def frustum(Rcw, tcw, fx, fy, cx, cy, bounds, bf, scale_factor, nlevels, viewing_cos_limit=0.5, kb=None, Ow=None):
    """The members of Frame that isInFrustum reads.  mfLogScaleFactor = log(mfScaleFactor) in float32 (Frame.cc:110).  Ow: the
    frame's mOw; when omitted it is formed as Frame::UpdatePoseMatrices does (src/Frame.cc:561-566, mOw = -mRcw.t()*mtcw: OpenCV
    materialises the transpose for the unary minus, then runs one small-matrix gemm with alpha = -1)."""
    Rcw = np.asarray(Rcw, np.float32).reshape(3, 3)
    tcw = np.asarray(tcw, np.float32).reshape(3)
    Ow = (-_mm3(np.ascontiguousarray(Rcw.T), tcw).reshape(3)).astype(np.float32) if Ow is None else np.asarray(Ow, np.float32).reshape(3)
    f = _Frustum()
    f.Rcw[:] = Rcw.reshape(-1).tolist(); f.tcw[:] = tcw.tolist(); f.Ow[:] = Ow.tolist()
    f.fx, f.fy, f.cx, f.cy = float(fx), float(fy), float(cx), float(cy)
    f.min_x, f.max_x, f.min_y, f.max_y = [float(b) for b in bounds]
    f.bf = float(bf)
    f.log_scale_factor = float(np.log(np.float32(scale_factor)))
    f.viewing_cos_limit = float(viewing_cos_limit)
    f.n_scale_levels = int(nlevels)
    f.camera_model = 0 if kb is None else 1   # ORBX_CAMERA_PINHOLE / ORBX_CAMERA_KB8 (k1..k4 of KannalaBrandt8)
    if kb is not None:
        f.kb[:] = [float(k) for k in kb]
    return f


def frustum_right_camera(left, Tlr, fx, fy, cx, cy, kb=None):
    """The right camera of a fisheye-stereo frame as Frame::isInFrustumChecks(pMP, limit, bRight = true) sets it up
    (src/Frame.cc:1274-1281): mR = Rrl * mRcw, mt = Rrl * mtcw + trl, twc = mRwc * tlr + mOw, projected through mpCamera2
    (fx .. kb).  Tlr = mTlr (3x4, left <- right); Trl is derived as the constructor does (Frame.cc:1192-1195:
    Rrl = Rlr^T, trl = Rrl * (-1 * tlr)).  `left` is the frustum() of the left camera; same bounds, scale and limits."""
    Tlr = np.asarray(Tlr, np.float32).reshape(3, 4)
    Rlr, tlr = Tlr[:, :3], Tlr[:, 3]
    Rrl = np.ascontiguousarray(Rlr.T)
    trl = _mm3(Rrl, (np.float32(-1) * tlr).astype(np.float32)).reshape(3)
    Rcw = np.array(left.Rcw[:], np.float32).reshape(3, 3); tcw = np.array(left.tcw[:], np.float32); Ow = np.array(left.Ow[:], np.float32)
    g = _Frustum.from_buffer_copy(left)
    g.Rcw[:] = _mm3(Rrl, Rcw).reshape(-1).tolist()
    g.tcw[:] = (_mm3(Rrl, tcw).reshape(3) + trl).astype(np.float32).tolist()
    g.Ow[:] = (_mm3(np.ascontiguousarray(Rcw.T), tlr).reshape(3) + Ow).astype(np.float32).tolist()
    g.fx, g.fy, g.cx, g.cy = float(fx), float(fy), float(cx), float(cy)
    g.camera_model = 0 if kb is None else 1
    if kb is not None:
        g.kb[:] = [float(k) for k in kb]
    return g


def _undistort(self, kps, fx, fy, cx, cy, dist):
    """Frame::UndistortKeyPoints (src/Frame.cc:837-870): mvKeys -> mvKeysUn."""
    from . import KP_DTYPE
    kps = np.ascontiguousarray(kps, dtype=KP_DTYPE)
    d = np.zeros(5, np.float32)
    d[:len(dist)] = dist
    out = np.empty_like(kps)
    _check(lib().orbx_undistort_keypoints(self._h, _p(kps), len(kps), C.c_float(fx), C.c_float(fy), C.c_float(cx), C.c_float(cy), _p(d),
                                          _p(out), MEM_HOST, None))
    return out


def _is_in_frustum(self, cam, P, normal, max_dist, min_dist):
    """Frame::isInFrustum (src/Frame.cc:572-642) over n map points -> dict of the MapPoint tracking fields."""
    P = np.ascontiguousarray(P, np.float32).reshape(-1, 3)
    normal = np.ascontiguousarray(normal, np.float32).reshape(-1, 3)
    max_dist = np.ascontiguousarray(max_dist, np.float32)
    min_dist = np.ascontiguousarray(min_dist, np.float32)
    n = len(P)
    o = dict(in_view=np.empty(n, np.uint8), proj_x=np.empty(n, np.float32), proj_y=np.empty(n, np.float32),
             proj_xr=np.empty(n, np.float32), depth=np.empty(n, np.float32), level=np.empty(n, np.int32),
             view_cos=np.empty(n, np.float32))
    _check(lib().orbx_is_in_frustum(self._h, C.byref(cam), n, _p(P), _p(normal), _p(max_dist), _p(min_dist), _p(o["in_view"]),
                                    _p(o["proj_x"]), _p(o["proj_y"]), _p(o["proj_xr"]), _p(o["depth"]), _p(o["level"]), _p(o["view_cos"]),
                                    MEM_HOST, None))
    return o


Preprocessor.UndistortKeyPoints = _undistort
Preprocessor.isInFrustum = _is_in_frustum
