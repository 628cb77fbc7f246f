"""Pins oracle/preproc_oracle.cc (cvtColor -> gray, remap, CLAHE) against the real OpenCV:
live against python cv2 when it is importable, and against tests/golden/preproc/preproc_cv2.npz (cv2 outputs committed by
tools/gen_golden_preproc.py) always.  Bit-exact in every case."""
import os
import zlib

import numpy as np
import pytest

import preproc_cases as pc

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "preproc", "preproc_cv2.npz"))
CODES = ("bgr", "rgb", "bgra", "rgba")


def _same(gold_key, img):
    assert np.array_equal(GOLD[gold_key + "_corner"], img[:64, :64])
    assert int(GOLD[gold_key + "_crc"]) == zlib.crc32(img.tobytes())


@pytest.mark.parametrize("name", list(pc.SHAPES))
def test_oracle_gray_matches_cv2_golden(oracle, name):
    col = pc.colour_image(name, 4)
    for code in CODES:
        g = oracle.cvt_gray(np.ascontiguousarray(col[:, :, :len(code)]), rgb_order=code[0] == "r", coeffs=oracle.GRAY_OCV4)
        _same("gray_%s_%s" % (name, code), g)


def test_oracle_gray_coefficient_sets_differ_by_at_most_one(oracle):
    col = pc.colour_image("euroc", 3)
    a = oracle.cvt_gray(col, False, oracle.GRAY_OCV346).astype(int)
    b = oracle.cvt_gray(col, False, oracle.GRAY_OCV4).astype(int)
    assert np.abs(a - b).max() <= 1 and (a != b).any()


@pytest.mark.parametrize("name", list(pc.SHAPES))
def test_oracle_remap_matches_cv2_golden(oracle, name):
    mx, my = pc.rectify_maps(name)
    _same("remap_%s" % name, oracle.remap_linear(pc.gray_image(name), mx, my))


@pytest.mark.parametrize("name", list(pc.SHAPES))
def test_oracle_clahe_matches_cv2_golden(oracle, name):
    img = pc.gray_image(name)
    for clip, tiles in pc.CLAHE_PARAMS:
        _same("clahe_%s_%g_%dx%d" % (name, clip, tiles[0], tiles[1]), oracle.clahe(img, clip, tiles))


def test_oracle_matches_live_cv2_on_random_inputs(oracle):
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(7)
    for h, w in [(64, 64), (75, 203), (240, 320)]:
        col = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
        assert np.array_equal(cv2.cvtColor(col, cv2.COLOR_RGBA2GRAY), oracle.cvt_gray(col, True))
        assert np.array_equal(cv2.cvtColor(col[:, :, :3].copy(), cv2.COLOR_BGR2GRAY), oracle.cvt_gray(col[:, :, :3], False))
        g = rng.integers(0, 256, (h, w), dtype=np.uint8)
        mx = (rng.random((h, w)) * (w + 8) - 4).astype(np.float32)
        my = (rng.random((h, w)) * (h + 8) - 4).astype(np.float32)
        assert np.array_equal(cv2.remap(g, mx, my, cv2.INTER_LINEAR), oracle.remap_linear(g, mx, my))
        for clip, tiles in [(3.0, (8, 8)), (1.0, (2, 7)), (0.0, (8, 8))]:
            assert np.array_equal(cv2.createCLAHE(clip, tiles).apply(g), oracle.clahe(g, clip, tiles))


def test_preproc_on_random_sizes(oracle):
    """cvtColor (BGR / RGBA), remap (INTER_LINEAR, a random distortion field incl. samples outside the image) and CLAHE (random clip
    limits and tile grids, incl. sizes that need padding) against live cv2 on seeded random image sizes; an 80-case run found no
    difference."""
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(55)
    for t in range(12):
        w, h = int(rng.integers(16, 1000)), int(rng.integers(16, 700))
        col = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
        if t % 3 == 0:
            col = cv2.GaussianBlur(col, (9, 9), 3)
        assert np.array_equal(cv2.cvtColor(col, cv2.COLOR_RGBA2GRAY), oracle.cvt_gray(col, True)), (w, h)
        c3 = np.ascontiguousarray(col[:, :, :3])
        g = cv2.cvtColor(c3, cv2.COLOR_BGR2GRAY)
        assert np.array_equal(g, oracle.cvt_gray(c3, False)), (w, h)
        xx, yy = np.meshgrid(np.arange(w, dtype=np.float32), np.arange(h, dtype=np.float32))
        mx = (xx * np.float32(rng.uniform(0.9, 1.1)) + np.float32(rng.uniform(-8, 8)) + np.float32(3) * np.sin(yy / 17).astype(np.float32)).astype(np.float32)
        my = (yy * np.float32(rng.uniform(0.9, 1.1)) + np.float32(rng.uniform(-8, 8)) + np.float32(2) * np.cos(xx / 23).astype(np.float32)).astype(np.float32)
        assert np.array_equal(cv2.remap(g, mx, my, cv2.INTER_LINEAR), oracle.remap_linear(g, mx, my)), (w, h)
        clip = float(rng.choice([1.0, 2.0, 3.0, 4.5, 40.0])); tiles = (int(rng.integers(1, 12)), int(rng.integers(1, 12)))
        if w >= tiles[0] and h >= tiles[1]:
            assert np.array_equal(cv2.createCLAHE(clip, tiles).apply(g), oracle.clahe(g, clip, tiles)), (w, h, clip, tiles)
