"""B200-native ORB feature frontend (extract + Hamming searches) behind ORB-SLAM3's own interface.

This Python module is a thin ctypes mirror of the reference's C++ classes for tests and benchmarks:

    ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)      include/ORBextractor.h:49-50
    ORBextractor.__call__(image, mask, vLappingArea) -> (monoIndex, keypoints, descriptors)   src/ORBextractor.cc:1071
    GetLevels / GetScaleFactor / GetScaleFactors / GetInverseScaleFactors / GetScaleSigmaSquares /
    GetInverseScaleSigmaSquares / mvImagePyramid                             include/ORBextractor.h:61-84
    ORBmatcher(nnratio, checkOri), DescriptorDistance, SearchByProjection, ...  include/ORBmatcher.h:39-94

All arithmetic runs in the hand-written sm_100a kernels of csrc/ through the C-ABI of include/orbx.h (liborbx.so).
There is no CPU fallback: if the library is missing, or no CUDA device is usable, calls raise.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liborbx.so")

ORBX_OK, ORBX_EMPTY_IMAGE, ORBX_ERR_ARG, ORBX_ERR_CUDA, ORBX_ERR_CAPACITY, ORBX_ERR_UNSUPPORTED = 0, -1, -2, -3, -4, -5
TAPS_OCV346, TAPS_OCV4 = 0, 1

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])  # == cv::KeyPoint, 28 bytes


class OrbxError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("orbx error %d: %s" % (code, msg))
        self.code = code


class _Params(C.Structure):
    _fields_ = [("nfeatures", C.c_int), ("scale_factor", C.c_float), ("nlevels", C.c_int), ("ini_th_fast", C.c_int),
                ("min_th_fast", C.c_int), ("blur_taps", C.c_int), ("max_batch", C.c_int), ("device", C.c_int)]


_lib = None


def lib():
    """Loads liborbx.so; raises if it has not been built (python 3_orb_slam3_selfnote_b200/build.py)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("liborbx.so is missing: build it with `python 3_orb_slam3_selfnote_b200/build.py` "
                              "(there is no CPU fallback)")
        _lib = C.CDLL(LIB_PATH)
        _lib.orbx_last_error.restype = C.c_char_p
        _lib.orbx_version.restype = C.c_char_p
        _lib.orbx_get_scale_factor.restype = C.c_float
    return _lib


class HostBuffer:
    """Page-locked host array from orbx_host_alloc (full-rate uploads / downloads); `.array` is a numpy view, freed on close()."""

    def __init__(self, shape, dtype=np.uint8):
        self.shape, self.dtype = tuple(int(x) for x in np.atleast_1d(shape)), np.dtype(dtype)
        self.nbytes = int(np.prod(self.shape)) * self.dtype.itemsize
        self._p = C.c_void_p()
        _check(lib().orbx_host_alloc(C.c_size_t(max(self.nbytes, 1)), C.byref(self._p)))
        buf = (C.c_uint8 * max(self.nbytes, 1)).from_address(self._p.value)
        self.array = np.frombuffer(buf, dtype=self.dtype, count=int(np.prod(self.shape))).reshape(self.shape)

    @property
    def ptr(self):
        return C.c_void_p(self._p.value)

    def close(self):
        if self._p is not None and self._p.value:
            self.array = None
            lib().orbx_host_free(self._p)
            self._p = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _check(rc):
    if rc < 0:
        raise OrbxError(rc, lib().orbx_last_error().decode())
    return rc


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class ORBextractor:
    def __init__(self, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7, blur_taps=TAPS_OCV346,
                 max_batch=1, device=0):
        self._h = C.c_void_p()
        p = _Params(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, blur_taps, max_batch, device)
        _check(lib().orbx_create(C.byref(p), C.byref(self._h)))
        self.nfeatures, self.nlevels, self.max_batch, self.device = nfeatures, nlevels, max_batch, device
        self._last_shape = None

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            lib().orbx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- getters (include/ORBextractor.h:61-82) ------------------------------------------------------------------
    def GetLevels(self):
        return lib().orbx_get_levels(self._h)

    def GetScaleFactor(self):
        return float(lib().orbx_get_scale_factor(self._h))

    def _tables(self):
        t = [np.empty(self.nlevels, np.float32) for _ in range(4)]
        _check(lib().orbx_get_scale_tables(self._h, *[_p(a) for a in t]))
        return t

    def GetScaleFactors(self):
        return self._tables()[0]

    def GetInverseScaleFactors(self):
        return self._tables()[1]

    def GetScaleSigmaSquares(self):
        return self._tables()[2]

    def GetInverseScaleSigmaSquares(self):
        return self._tables()[3]

    def features_per_level(self):
        q = np.empty(self.nlevels, np.int32)
        _check(lib().orbx_get_features_per_level(self._h, _p(q)))
        return q

    def max_keypoints(self, w, h):
        return _check(lib().orbx_max_keypoints(self._h, w, h))

    # -- operator() ----------------------------------------------------------------------------------------------
    def __call__(self, image, mask=None, vLappingArea=(0, 0)):
        """-> (monoIndex, keypoints[KP_DTYPE], descriptors[n,32]); monoIndex == -1 for an empty image (ORBextractor.cc:1075)"""
        if image is None or image.size == 0:
            return -1, np.empty(0, KP_DTYPE), np.empty((0, 32), np.uint8)
        image = np.asarray(image)
        assert image.dtype == np.uint8 and image.ndim == 2, "CV_8UC1 expected (ORBextractor.cc:1080)"
        if image.strides[1] != 1:
            image = np.ascontiguousarray(image)
        h, w = image.shape
        cap = self.max_keypoints(w, h)
        kps = np.empty(cap, KP_DTYPE)
        desc = np.empty((cap, 32), np.uint8)
        n, mono = C.c_int(0), C.c_int(0)
        rc = lib().orbx_extract(self._h, _p(image), w, h, image.strides[0], int(vLappingArea[0]), int(vLappingArea[1]), _p(kps),
                                _p(desc), cap, C.byref(n), C.byref(mono))
        _check(rc)
        self._last_shape = (h, w)
        return mono.value, kps[:n.value].copy(), desc[:n.value].copy()

    def extract_batch(self, images, vLappingArea=(0, 0)):
        """images: (n,h,w) uint8 host array.  -> (counts[n], mono[n], keypoints[n,cap], descriptors[n,cap,32])"""
        images = np.ascontiguousarray(images, dtype=np.uint8)
        n, h, w = images.shape
        cap = self.max_keypoints(w, h)
        kps = np.empty((n, cap), KP_DTYPE)
        desc = np.empty((n, cap, 32), np.uint8)
        counts = np.empty(n, np.int32)
        mono = np.empty(n, np.int32)
        _check(lib().orbx_extract_batch_host(self._h, _p(images), n, w, h, w, C.c_size_t(w * h), int(vLappingArea[0]),
                                             int(vLappingArea[1]), _p(kps), _p(desc), cap, _p(counts), _p(mono)))
        self._last_shape = (h, w)
        return counts, mono, kps, desc

    def extract_batch_device(self, d_ptr, n, w, h, pitch, frame_stride, vLappingArea=(0, 0), stream=0, slot=0):
        """frames resident in HBM (raw device pointer); results stay on the device (see results_device).  slot: which of the handle's
        two workspace sets the batch uses (alternate them to overlap the readers of batch i with the extraction of batch i+1)."""
        _check(lib().orbx_extract_batch_device_slot(self._h, int(slot), C.c_void_p(d_ptr), n, w, h, pitch, C.c_size_t(frame_stride),
                                                    int(vLappingArea[0]), int(vLappingArea[1]), C.c_void_p(stream)))
        self._last_shape = (h, w)

    def results_device(self):
        """-> dict of raw device pointers of the last batch + kp_stride"""
        k, d, c, m, s = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p()
        stride = C.c_int(0)
        _check(lib().orbx_batch_results_device(self._h, C.byref(k), C.byref(d), C.byref(c), C.byref(m), C.byref(s), C.byref(stride)))
        return dict(keypoints=k.value, descriptors=d.value, counts=c.value, mono=m.value, status=s.value, kp_stride=stride.value)

    def results_host(self, n):
        """host copy of the last device batch -> (counts[n], mono[n], keypoints[n,cap], descriptors[n,cap,32])"""
        h, w = self._last_shape
        cap = self.max_keypoints(w, h)
        kps = np.empty((n, cap), KP_DTYPE)
        desc = np.empty((n, cap, 32), np.uint8)
        counts = np.empty(n, np.int32)
        mono = np.empty(n, np.int32)
        _check(lib().orbx_batch_results_host(self._h, _p(kps), _p(desc), cap, _p(counts), _p(mono)))
        return counts, mono, kps, desc

    def set_profiling(self, enable=True):
        _check(lib().orbx_set_profiling(self._h, int(enable)))

    def stage_times(self):
        """-> list of (stage name, ms) of the last profiled batch"""
        ms = (C.c_float * 8)()
        names = (C.c_char_p * 8)()
        n = _check(lib().orbx_stage_times(self._h, ms, names))
        return [(names[i].decode(), float(ms[i])) for i in range(n)]

    def launches_per_batch(self):
        return _check(lib().orbx_launches_per_batch(self._h))

    # -- mvImagePyramid (include/ORBextractor.h:84) ----------------------------------------------------------------
    def pyramid_level(self, level, frame=0):
        w, h = C.c_int(0), C.c_int(0)
        _check(lib().orbx_pyramid_level(self._h, frame, level, None, 0, C.byref(w), C.byref(h)))
        out = np.empty((h.value, w.value), np.uint8)
        _check(lib().orbx_pyramid_level(self._h, frame, level, _p(out), w.value, C.byref(w), C.byref(h)))
        return out

    @property
    def mvImagePyramid(self):
        return [self.pyramid_level(l) for l in range(self.nlevels)]

    # -- stage taps for parity tests ---------------------------------------------------------------------------------
    def debug_blurred_level(self, level, frame=0):
        ref = self.pyramid_level(level, frame)
        out = np.empty_like(ref)
        _check(lib().orbx_debug_blurred_level(self._h, frame, level, _p(out), out.shape[1]))
        return out

    def debug_candidates(self, level, frame=0):
        n = C.c_int(0)
        _check(lib().orbx_debug_candidates(self._h, frame, level, None, 0, C.byref(n)))
        out = np.empty((max(n.value, 1), 3), np.int32)
        _check(lib().orbx_debug_candidates(self._h, frame, level, _p(out), n.value, C.byref(n)))
        return out[:n.value]

    def debug_level_keypoints(self, level, frame=0):
        n = C.c_int(0)
        _check(lib().orbx_debug_level_keypoints(self._h, frame, level, None, 0, C.byref(n)))
        out = np.empty(max(n.value, 1), KP_DTYPE)
        _check(lib().orbx_debug_level_keypoints(self._h, frame, level, _p(out), n.value, C.byref(n)))
        return out[:n.value]


def DescriptorDistance(a, b):
    """ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:2463-2483) on two 32-byte rows"""
    a = np.ascontiguousarray(a, np.uint8)
    b = np.ascontiguousarray(b, np.uint8)
    return lib().orbx_hamming(_p(a), _p(b))


from .matcher import FrameView, LocalMap, ORBmatcher, ORBVocabulary  # noqa: E402,F401  (mirror of include/ORBmatcher.h)
