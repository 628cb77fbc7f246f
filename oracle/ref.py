"""ctypes binding of oracle/_ref/libref_orb.so: the reference's OWN ORBextractor.cc, compiled unmodified.

TEST INFRASTRUCTURE ONLY (same rule as oracle/__init__.py): tests/, __graft_entry__.smoke() and bench.py's CPU legs.

The library is built by `make -C oracle ref` from /root/reference/src/ORBextractor.cc (by path, nothing copied) against
the stand-in cv:: headers of oracle/ref_shim/, whose six image primitives are the cv2-pinned functions of
oracle/orb_oracle.cc.  /root/reference exists only in the build container, so the GPU box uses the prebuilt files
(oracle/_ref/ is git-ignored but travels with the snapshot).  `available()` says whether the library can be loaded.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from . import KP_DTYPE, TAPS_OCV346, TAPS_OCV4  # noqa: F401

_HERE = os.path.dirname(os.path.abspath(__file__))
_DIR = os.path.join(_HERE, "_ref")
REFERENCE_ROOT = os.environ.get("ORBX_REFERENCE_ROOT", "/root/reference")
_LIBS = {"strict": os.path.join(_DIR, "libref_orb.so"),        # -ffp-contract=off: the declared arithmetic
         "literal": os.path.join(_DIR, "libref_orb_literal.so")}  # gcc default contraction (FMA), as CMakeLists.txt builds it
_loaded = {}


def can_build():
    return os.path.exists(os.path.join(REFERENCE_ROOT, "src", "ORBextractor.cc"))


def build(force=False):
    """Compiles the reference sources where they lie.  No-op (returns False) when /root/reference is absent."""
    if not can_build():
        return False
    cmd = ["make", "-C", _HERE, "REF=" + REFERENCE_ROOT, "ref"] + (["-B"] if force else [])
    subprocess.check_call(cmd, stdout=subprocess.DEVNULL)
    return True


def build_twin():
    """libshim_match.so: the reference-side harness linked against the product's host/ORBmatcher.h + liborbx.so (tests/test_refsig_gpu.py).
    Needs the product library built first; returns False when /root/reference is absent."""
    if not can_build():
        return False
    subprocess.check_call(["make", "-C", _HERE, "REF=" + REFERENCE_ROOT, "twin"], stdout=subprocess.DEVNULL)
    return True


def available(variant="strict"):
    if not os.path.exists(_LIBS[variant]) and can_build():
        build()
    return os.path.exists(_LIBS[variant])


def lib(variant="strict"):
    if variant not in _loaded:
        if not available(variant):
            raise RuntimeError("oracle/_ref/%s missing and %s absent: run `make -C oracle ref` in the build container"
                               % (os.path.basename(_LIBS[variant]), REFERENCE_ROOT))
        L = C.CDLL(_LIBS[variant])
        L.ref_extractor_create.restype = C.c_void_p
        L.ref_extractor_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.ref_extractor_destroy.argtypes = [C.c_void_p]
        L.ref_scale_factor.restype = C.c_float
        L.ref_time_extract.restype = C.c_double
        _loaded[variant] = L
    return _loaded[variant]


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class RefExtractor:
    """ORB_SLAM3::ORBextractor of the reference (include/ORBextractor.h:43-109), same call shape as oracle.Extractor."""

    def __init__(self, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7, blur_taps=TAPS_OCV346,
                 variant="strict"):
        self._L = lib(variant)
        self.nfeatures, self.nlevels, self.blur_taps = nfeatures, nlevels, blur_taps
        self._h = self._L.ref_extractor_create(nfeatures, C.c_float(scaleFactor), nlevels, iniThFAST, minThFAST)
        self.scale = np.empty(nlevels, np.float32)
        self.inv_scale = np.empty(nlevels, np.float32)
        self.sigma2 = np.empty(nlevels, np.float32)
        self.inv_sigma2 = np.empty(nlevels, np.float32)
        self.quota = np.empty(nlevels, np.int32)
        self.umax = np.empty(16, np.int32)
        self._L.ref_tables(C.c_void_p(self._h), _p(self.scale), _p(self.inv_scale), _p(self.sigma2), _p(self.inv_sigma2),
                           _p(self.quota), _p(self.umax))

    def __del__(self):
        try:
            self._L.ref_extractor_destroy(C.c_void_p(self._h))
        except Exception:
            pass

    def __call__(self, img, lapping=(0, 0)):
        """returns (monoIndex, keypoints[KP_DTYPE], descriptors[n,32]) -- operator() of the reference"""
        cap = 4 * self.nfeatures + 64 * self.nlevels
        kps = np.empty(cap, KP_DTYPE)
        desc = np.empty((cap, 32), np.uint8)
        n = C.c_int(0)
        if img is None or img.size == 0:
            w = h = stride = 0
            ptr = None
        else:
            img = np.asarray(img, dtype=np.uint8)
            assert img.ndim == 2 and img.strides[1] == 1
            h, w = img.shape
            stride, ptr = img.strides[0], _p(img)
        mono = self._L.ref_extract(C.c_void_p(self._h), ptr, w, h, stride, int(lapping[0]), int(lapping[1]), self.blur_taps,
                                   _p(kps), _p(desc), cap, C.byref(n))
        if mono == -1:
            return -1, np.empty(0, KP_DTYPE), np.empty((0, 32), np.uint8)
        assert mono >= 0, "ref_extract failed: %d" % mono
        return mono, kps[:n.value].copy(), desc[:n.value].copy()

    def level(self, l):
        """mvImagePyramid[l] after the last call."""
        ptr, w, h, s = C.c_void_p(), C.c_int(), C.c_int(), C.c_int()
        rc = self._L.ref_pyramid_level(C.c_void_p(self._h), l, C.byref(ptr), C.byref(w), C.byref(h), C.byref(s))
        assert rc == 0
        buf = (C.c_uint8 * (s.value * h.value)).from_address(ptr.value)
        a = np.frombuffer(buf, np.uint8).reshape(h.value, s.value)[:, :w.value]
        return a.copy()

    def bordered_level(self, l, border=19):
        """The level with its EDGE_THRESHOLD frame (copyMakeBorder, ORBextractor.cc:1203-1214)."""
        ptr, w, h, s = C.c_void_p(), C.c_int(), C.c_int(), C.c_int()
        self._L.ref_pyramid_level(C.c_void_p(self._h), l, C.byref(ptr), C.byref(w), C.byref(h), C.byref(s))
        start = ptr.value - border * s.value - border
        buf = (C.c_uint8 * (s.value * (h.value + 2 * border))).from_address(start)
        return np.frombuffer(buf, np.uint8).reshape(h.value + 2 * border, s.value)[:, :w.value + 2 * border].copy()


def descriptor_distance(a, b, variant="strict"):
    """ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:2463-2483) on n pairs of rows."""
    a = np.ascontiguousarray(a, np.uint8).reshape(-1, 32)
    b = np.ascontiguousarray(b, np.uint8).reshape(-1, 32)
    out = np.empty(a.shape[0], np.int32)
    lib(variant).ref_descriptor_distance(_p(a), _p(b), a.shape[0], _p(out))
    return out


def three_maxima(hist, variant="strict"):
    """ORBmatcher::ComputeThreeMaxima (src/ORBmatcher.cc:2416-2458) on bin sizes; indices the call leaves untouched stay -1."""
    hist = np.ascontiguousarray(hist, np.int32)
    out = np.empty(3, np.int32)
    lib(variant).ref_three_maxima(_p(hist), hist.shape[0], _p(out))
    return tuple(int(v) for v in out)


def time_extract(frames, nfeatures, nthreads, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7, blur_taps=TAPS_OCV346,
                 variant="strict"):
    """Seconds to run the reference operator() over frames[n,h,w], frame-parallel on nthreads; returns (seconds, keypoints)."""
    frames = np.ascontiguousarray(frames, np.uint8)
    n, h, w = frames.shape
    tot = C.c_longlong(0)
    s = lib(variant).ref_time_extract(nfeatures, C.c_float(scaleFactor), nlevels, iniThFAST, minThFAST, blur_taps, _p(frames), n, w,
                                      h, 0, 0, nthreads, C.byref(tot))
    return float(s), int(tot.value)
