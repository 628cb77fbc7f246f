"""The oracle against the committed golden vectors (tests/golden/, made by tests/golden/gen_golden.py from the REFERENCE'S OWN
ORBextractor.cc compiled unmodified -- `source` == "reference") plus structural checks of the restated DistributeOctTree /
operator() ordering.  The reference has no tests of its own for this path: these fixtures carry the
reference's output to hosts where /root/reference does not exist; the cv2 cross-checks (test_oracle_cv2.py) pin the primitives."""
import glob
import os
import zlib

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_oracle_reproduces_golden(oracle, synth):
    files = sorted(glob.glob(os.path.join(GOLDEN, "*.npz")))
    assert len(files) >= 5
    for fn in files:
        g = np.load(fn)
        w, h, nfeat, seed, fid, right, lap0, lap1 = [int(v) for v in g["spec"]]
        assert str(g["source"]) == "reference", "fixture was not generated from the reference source"
        img = synth.frame(w, h, seed, fid, bool(right))
        assert zlib.crc32(img.tobytes()) == int(g["image_crc"]), "synthetic generator changed"
        ex = oracle.Extractor(nfeat, blur_taps=oracle.TAPS_OCV346, sincos_mode=oracle.SINCOS_PORTABLE)
        mono, k, d = ex(img, (lap0, lap1))
        assert mono == int(g["mono"])
        assert k.tobytes() == g["keypoints"].tobytes()
        assert np.array_equal(d, g["descriptors"])
        assert [zlib.crc32(ex.level(l).tobytes()) for l in range(8)] == [int(c) for c in g["pyramid_crc"]]


def test_lapping_area_ordering(oracle, synth):
    """operator() writes lapping keypoints from the back and the rest from the front (ORBextractor.cc:1169-1178)."""
    img = synth.frame(752, 480, 4, 0)
    ex = oracle.Extractor(1000)
    mono_f, kf, df = ex(img, (0, 0))          # nothing laps: forward level-major order
    mono_b, kb, db = ex(img, (0, 1000))       # everything laps: exact reverse
    assert mono_f == len(kf) and mono_b == 0
    assert kf[::-1].tobytes() == kb.tobytes() and np.array_equal(df[::-1], db)
    mono_m, km, dm = ex(img, (200, 400))
    lap = (kf["x"] >= 200) & (kf["x"] <= 400)
    assert mono_m == int((~lap).sum())
    assert km[:mono_m].tobytes() == kf[~lap].tobytes()
    assert km[mono_m:].tobytes() == kf[lap][::-1].tobytes()
    assert np.array_equal(dm[:mono_m], df[~lap]) and np.array_equal(dm[mono_m:], df[lap][::-1])


def test_distribute_octtree_properties(oracle):
    rng = np.random.default_rng(0)
    for trial in range(30):
        W, H = int(rng.integers(100, 900)), int(rng.integers(100, 500))
        if round(W / H) < 1:
            continue
        n = int(rng.integers(0, 1500))
        pts = np.unique(np.stack([rng.integers(0, W, n), rng.integers(0, H, n)], 1), axis=0)
        rng.shuffle(pts)
        cands = np.concatenate([pts, rng.integers(7, 60, (len(pts), 1))], 1).astype(np.int32)
        N = int(rng.integers(1, 400))
        keep = oracle.distribute(cands, 16, 16 + W, 16, 16 + H, N)
        assert len(set(keep.tolist())) == len(keep)
        nIni = round(W / H)
        assert len(keep) <= max(N + 3, 4 * nIni)
        if len(pts) >= N + 3:
            assert len(keep) >= min(N, 1)
        if len(pts) <= 1:
            assert len(keep) == len(pts)
    # all candidates kept when the quota exceeds them and every point can be isolated
    cands = np.array([[10, 10, 9], [300, 10, 9], [10, 200, 30], [300, 200, 8], [150, 100, 50]], np.int32)
    keep = oracle.distribute(cands, 16, 16 + 400, 16, 16 + 300, 100)
    assert sorted(keep.tolist()) == [0, 1, 2, 3, 4]
    # one node, two points: the stronger wins; ties keep the first in candidate order
    assert oracle.distribute(np.array([[5, 5, 9], [6, 5, 12]], np.int32), 16, 416, 16, 316, 1).tolist() in ([1], [0, 1], [1, 0])


def test_three_maxima(oracle):
    h = np.zeros(30, np.int32); h[3] = 50; h[7] = 20; h[9] = 4
    assert oracle.three_maxima(h) == (3, 7, -1)       # third below 10 % of the first
    h[9] = 6
    assert oracle.three_maxima(h) == (3, 7, 9)
    h[7] = 4
    assert oracle.three_maxima(h) == (3, 9, -1)
    h[9] = 4
    assert oracle.three_maxima(h) == (3, -1, -1)
    assert oracle.three_maxima(np.zeros(30, np.int32)) == (-1, -1, -1)
