"""Seeded inputs shared by the pre-processing tests and tools/gen_golden_preproc.py (numpy only, no cv2)."""
import zlib

import numpy as np

SHAPES = {"euroc": (480, 752), "tumvi": (512, 512), "ragged": (101, 131), "tiny": (37, 53)}
CLAHE_PARAMS = [(3.0, (8, 8)), (2.0, (4, 4)), (40.0, (8, 8)), (0.5, (3, 5))]


def _rng(name, salt):
    return np.random.default_rng(zlib.crc32(name.encode()) + salt)


def _smooth(a):
    """3x3 box filter with edge replication (integer), keeps images feature-like rather than white noise"""
    p = np.pad(a.astype(np.int32), ((1, 1), (1, 1)) + ((0, 0),) * (a.ndim - 2), mode="edge")
    s = sum(p[dy:dy + a.shape[0], dx:dx + a.shape[1]] for dy in range(3) for dx in range(3))
    return ((s + 4) // 9).astype(np.uint8)


def colour_image(name, channels=4):
    h, w = SHAPES[name]
    return _smooth(_rng(name, 1).integers(0, 256, (h, w, channels), dtype=np.uint8))


def gray_image(name):
    h, w = SHAPES[name]
    rng = _rng(name, 2)
    img = _smooth(rng.integers(0, 256, (h, w), dtype=np.uint8)).astype(np.int32)
    yy, xx = np.mgrid[0:h, 0:w]
    img = img // 2 + (xx * 90 // w) + (yy * 40 // h)  # low-contrast ramp so CLAHE's clipping is exercised
    return img.clip(0, 255).astype(np.uint8)


def rectify_maps(name):
    """float32 maps shaped like initUndistortRectifyMap output (radial distortion + small rotation), leaving the image on
    all four sides so the constant border is exercised, plus a few NaN / huge entries"""
    h, w = SHAPES[name]
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float64)
    cx, cy, f = w * 0.49, h * 0.52, 0.6 * w
    x, y = (xx - cx) / f, (yy - cy) / f
    c, s = np.cos(0.01), np.sin(0.01)
    x, y = c * x - s * y, s * x + c * y
    r2 = x * x + y * y
    k = 1 - 0.28 * r2 + 0.07 * r2 * r2
    mx = (x * k * f * 1.08 + cx + 0.3).astype(np.float32)
    my = (y * k * f * 1.08 + cy - 0.7).astype(np.float32)
    rng = _rng(name, 3)
    for _ in range(6):
        mx[rng.integers(h), rng.integers(w)] = rng.choice([np.nan, 1e12, -1e12, -1.0, w - 1.0, w - 0.5])
        my[rng.integers(h), rng.integers(w)] = rng.choice([np.nan, 1e12, -1e12, -1.0, h - 1.0, h - 0.5])
    return mx, my
