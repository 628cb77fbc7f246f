"""The restated oracle against the reference's OWN code (oracle/_ref: /root/reference/src/ORBextractor.cc compiled unmodified
against the stand-in cv:: headers of oracle/ref_shim/, node allocator = monotone bump, see oracle/ref_shim/ref_harness.cc).

This is the pin VERDICT r01 asked for: quadtree, cell loop + threshold fallback, lapping order, IC_Angle, rBRIEF, pattern
table and constructor tables are compared with the reference's compiled source, byte for byte, on every shape the GPU
parity tests use plus the edge cases.  Also ORBmatcher::DescriptorDistance / ComputeThreeMaxima (cut out of
src/ORBmatcher.cc at build time).  CPU-only; skipped when neither /root/reference nor a prebuilt oracle/_ref exists."""
import numpy as np
import pytest

from oracle import ref as oref

pytestmark = pytest.mark.skipif(not oref.available(), reason="oracle/_ref not built and /root/reference absent")

SHAPES = [  # the parametrisation of tests/test_extract_gpu.py::test_extract_matches_oracle
    (752, 480, 1000, (0, 1000), 0), (752, 480, 1200, (0, 0), 0), (752, 480, 1000, (0, 0), 1), (512, 512, 1500, (0, 0), 0),
    (512, 512, 1000, (0, 255), 0), (640, 480, 1000, (0, 0), 0), (1241, 376, 2000, (0, 0), 0), (1280, 720, 2000, (0, 0), 0),
    (333, 251, 500, (100, 200), 0), (752, 480, 5000, (0, 1000), 0),
]
GEOMETRIES = [  # ::test_extract_other_geometries
    (1920, 1080, 2000, 1.2, 8, 20, 7), (752, 480, 1000, 1.5, 6, 20, 7), (640, 480, 800, 2.4, 4, 20, 7),
    (752, 480, 1000, 1.2, 8, 40, 12), (752, 480, 300, 1.1, 12, 20, 7),
]


def _same(r, o, what):
    (mr, kr, dr), (mo, ko, do) = r, o
    assert mr == mo, what
    assert len(kr) == len(ko), "%s: %d vs %d keypoints" % (what, len(kr), len(ko))
    assert kr.tobytes() == ko.tobytes(), "%s: keypoint records differ" % what
    assert np.array_equal(dr, do), "%s: %d descriptor rows differ" % (what, int((dr != do).any(axis=1).sum()))


@pytest.mark.parametrize("w,h,nfeat,lap,taps", SHAPES)
def test_restatement_equals_reference_source(oracle, synth, w, h, nfeat, lap, taps):
    img = synth.frame(w, h, 100 + nfeat, 3)
    rx = oref.RefExtractor(nfeat, blur_taps=taps)
    ex = oracle.Extractor(nfeat, blur_taps=taps, sincos_mode=oracle.SINCOS_LIBM_F32)  # the reference's cos(float) is cosf
    _same(rx(img, lap), ex(img, lap), "%dx%d" % (w, h))
    for l in range(8):
        assert np.array_equal(rx.level(l), ex.level(l)), "mvImagePyramid[%d]" % l
    # the portable sin/cos definition the CUDA path implements gives the same descriptors here (F6)
    ep = oracle.Extractor(nfeat, blur_taps=taps, sincos_mode=oracle.SINCOS_PORTABLE)
    _same(rx(img, lap), ep(img, lap), "%dx%d portable sincos" % (w, h))


@pytest.mark.parametrize("w,h,nfeat,scale,nlevels,ini,mn", GEOMETRIES)
def test_other_geometries(oracle, synth, w, h, nfeat, scale, nlevels, ini, mn):
    img = synth.frame(w, h, 300 + nfeat, 1)
    rx = oref.RefExtractor(nfeat, scale, nlevels, ini, mn)
    ex = oracle.Extractor(nfeat, scale, nlevels, ini, mn, sincos_mode=oracle.SINCOS_LIBM_F32)
    _same(rx(img, (0, 0)), ex(img, (0, 0)), "%dx%d scale %.1f" % (w, h, scale))
    for l in range(nlevels):
        assert np.array_equal(rx.level(l), ex.level(l))
    for a, b in ((rx.scale, ex.scale), (rx.inv_scale, ex.inv_scale), (rx.sigma2, ex.sigma2), (rx.inv_sigma2, ex.inv_sigma2),
                 (rx.quota, ex.quota), (rx.umax, ex.umax)):
        assert np.array_equal(a, b)


def test_constructor_tables(oracle):
    for nf, sf, nl in ((1000, 1.2, 8), (1200, 1.2, 8), (1500, 1.2, 8), (2000, 1.2, 8), (5000, 1.2, 8), (1000, 1.5, 6), (300, 1.1, 12),
                       (800, 2.4, 4), (1, 1.2, 8), (7, 2.0, 3)):
        rx, ex = oref.RefExtractor(nf, sf, nl), oracle.Extractor(nf, sf, nl)
        assert np.array_equal(rx.quota, ex.quota), (nf, sf, nl)
        assert rx.scale.tobytes() == ex.scale.tobytes() and rx.inv_scale.tobytes() == ex.inv_scale.tobytes()
        assert rx.sigma2.tobytes() == ex.sigma2.tobytes() and rx.inv_sigma2.tobytes() == ex.inv_sigma2.tobytes()
        assert np.array_equal(rx.umax, ex.umax)
    assert list(oref.RefExtractor(1000).umax) == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]


def test_edge_cases(oracle):
    rx = oref.RefExtractor(1000)
    ex = oracle.Extractor(1000, sincos_mode=oracle.SINCOS_LIBM_F32)
    assert rx(np.empty((0, 0), np.uint8))[0] == -1                       # ORBextractor.cc:1075-1076
    m, k, d = rx(np.full((480, 752), 128, np.uint8))                      # flat: no keypoints, descriptors released
    assert m == 0 and len(k) == 0
    rng = np.random.default_rng(3)
    noise = rng.integers(0, 256, (480, 752), dtype=np.uint8)
    sparse = np.full((480, 752), 100, np.uint8)
    for _ in range(40):
        y, x = int(rng.integers(30, 440)), int(rng.integers(30, 700))
        sparse[y:y + 9, x:x + 9] = 112
    small = (rng.integers(0, 256, (120, 160), dtype=np.uint8) // 4 * 4).astype(np.uint8)
    # periodic texture: many nodes with equal point counts -> exercises the (size, address) tie of the quadtree sort
    yy, xx = np.mgrid[0:480, 0:752]
    checker = (((xx // 12 + yy // 12) % 2) * 90 + 60).astype(np.uint8)
    dots = np.full((480, 752), 90, np.uint8)
    dots[8::16, 8::16] = 200
    dots[9::16, 8::16] = 200
    dots[8::16, 9::16] = 200
    for name, img in (("noise", noise), ("sparse", sparse), ("small", small), ("checker", checker), ("dots", dots)):
        _same(rx(img, (0, 0)), ex(img, (0, 0)), name)
        _same(rx(img, (100, 300)), ex(img, (100, 300)), name + " lapping")
    big = np.random.default_rng(5).integers(0, 256, (500, 800), dtype=np.uint8)
    view = big[10:490, 24:776]                                           # strided input Mat
    _same(rx(view, (0, 0)), ex(np.ascontiguousarray(view), (0, 0)), "strided")


def test_same_object_reused_across_sizes_and_frames(oracle, synth):
    """mvImagePyramid and the extractor object are reused call after call (Frame.cc:480-482); results must not depend on history."""
    rx = oref.RefExtractor(1200)
    ex = oracle.Extractor(1200, sincos_mode=oracle.SINCOS_LIBM_F32)
    for (w, h, fid) in ((752, 480, 0), (512, 512, 1), (752, 480, 2), (752, 480, 0)):
        img = synth.frame(w, h, 21, fid)
        _same(rx(img, (0, 0)), ex(img, (0, 0)), "frame %d" % fid)


def test_bordered_pyramid_is_reflect101(synth):
    """copyMakeBorder(.., BORDER_REFLECT_101) frame around every level (ORBextractor.cc:1203-1214); never read on the path."""
    rx = oref.RefExtractor(1000)
    rx(synth.frame(320, 240, 2, 0), (0, 0))
    for l in (0, 3, 7):
        lvl, b = rx.level(l), rx.bordered_level(l)
        assert np.array_equal(b, np.pad(lvl, 19, mode="reflect"))


def test_descriptor_distance_and_three_maxima(oracle):
    rng = np.random.default_rng(11)
    a = rng.integers(0, 256, (4000, 32), dtype=np.uint8)
    b = rng.integers(0, 256, (4000, 32), dtype=np.uint8)
    b[:100] = a[:100]
    b[100:200] = ~a[100:200]
    d = oref.descriptor_distance(a, b)
    want = np.unpackbits(a ^ b, axis=1).sum(axis=1)
    assert np.array_equal(d, want) and d[:100].max() == 0 and d[100:200].min() == 256
    assert all(oracle.hamming(a[i], b[i]) == d[i] for i in range(0, 4000, 37))
    for _ in range(300):
        hist = rng.integers(0, 40, 30) * (rng.random(30) < rng.random())
        assert oref.three_maxima(hist) == oracle.three_maxima(hist), hist
    for hist in ([0] * 30, [5] + [0] * 29, [50, 4] + [0] * 28, [50, 5, 4] + [0] * 27, [3, 3, 3] + [0] * 27):
        assert oref.three_maxima(hist) == oracle.three_maxima(hist), hist


def test_literal_fma_build_difference_is_reported(synth):
    """gcc contracts x*b + y*a into FMA by default (-std=c++11 does not switch that off for C++), so the reference's own
    build on an FMA host rounds the rBRIEF rotation differently from the declared no-FMA arithmetic.  Measured,
    not hidden: keypoints are identical, descriptor rows that change are counted."""
    if not oref.available("literal"):
        pytest.skip("literal build missing")
    rs, rl = oref.RefExtractor(1200), oref.RefExtractor(1200, variant="literal")
    rows = diff = 0
    for fid in range(6):
        img = synth.frame(752, 480, 77, fid)
        (ms, ks, ds), (ml, kl, dl) = rs(img, (0, 0)), rl(img, (0, 0))
        assert ms == ml and ks.tobytes() == kl.tobytes()
        rows += len(ks)
        diff += int((ds != dl).any(axis=1).sum())
    print("descriptor rows differing between the FMA-contracted and the strict build: %d of %d" % (diff, rows))
    assert diff <= max(2, rows // 500)


def test_random_geometries_equal_reference_source(oracle, synth):
    """Seeded random image sizes, feature counts, scale factors, level counts, FAST thresholds and lapping areas (a 120-case run of the
    same generator found no difference): restatement == the reference's compiled ORBextractor.cc, keypoints, descriptors, all levels.
    Cases where a level's bordered region is more than twice as tall as wide are skipped: the reference divides by zero there
    (nIni = round(width / height) = 0, ORBextractor.cc:541-543; the product answers ORBX_ERR_UNSUPPORTED)."""
    rng = np.random.default_rng(2024)
    ran = 0
    for t in range(16):
        w, h = int(rng.integers(80, 900)), int(rng.integers(80, 700))
        if w < 0.55 * h:
            w = int(0.6 * h)
        nfeat = int(rng.integers(50, 2500)); scale = float(np.float32(rng.uniform(1.05, 1.6)))
        nlev = int(rng.integers(2, 11)); ini = int(rng.integers(8, 40)); mn = int(rng.integers(2, ini + 1))
        lap = (0, 0) if t % 3 else (int(rng.integers(0, w // 2)), int(rng.integers(w // 2, w)))
        if min(w, h) / scale ** (nlev - 1) < 60:
            nlev = max(2, int(np.log(min(w, h) / 60) / np.log(scale)))
        if any(round((round(w / scale ** l) - 32) / max(1, round(h / scale ** l) - 32)) < 1 for l in range(nlev)):
            continue
        img = synth.frame(w, h, 500 + t, t)
        rx = oref.RefExtractor(nfeat, scale, nlev, ini, mn)
        ex = oracle.Extractor(nfeat, scale, nlev, ini, mn, sincos_mode=oracle.SINCOS_LIBM_F32)
        _same(rx(img, lap), ex(img, lap), "case %d: %dx%d nfeat %d scale %.3f levels %d fast %d/%d lap %s" % (t, w, h, nfeat, scale, nlev, ini, mn, lap))
        for l in range(nlev):
            assert np.array_equal(rx.level(l), ex.level(l))
        ran += 1
    assert ran >= 12
