"""Deterministic synthetic frames for parity tests and benchmarks.

All-integer generator (splitmix64), reproducible bit-for-bit from (w, h, seed, id, right): 400 painted rectangles on a
grey background, two 3x3 box blurs, +-3 uniform noise.  `right=True` shifts every rectangle left by its own disparity
(1..40 px), which gives EuRoC-shaped rectified stereo pairs.  `jitter=True` moves every rectangle centre by -3..+3 px
(used to fabricate the "previous frame" for the projection-search workload).
"""
import numpy as np

_G = np.uint64(0x9E3779B97F4A7C15)
_M1 = np.uint64(0xBF58476D1CE4E5B9)
_M2 = np.uint64(0x94D049BB133111EB)


def splitmix64(seed, n):
    """the first n outputs of splitmix64 started from state `seed` (uint64 array)"""
    with np.errstate(over="ignore"):
        x = np.uint64(seed) + _G * np.arange(1, n + 1, dtype=np.uint64)
        z = (x ^ (x >> np.uint64(30))) * _M1
        z = (z ^ (z >> np.uint64(27))) * _M2
        return z ^ (z >> np.uint64(31))


def _box3(img):
    p = np.pad(img, 1, mode="edge")
    s = (p[:-2, :-2] + p[:-2, 1:-1] + p[:-2, 2:] + p[1:-1, :-2] + p[1:-1, 1:-1] + p[1:-1, 2:] + p[2:, :-2] + p[2:, 1:-1]
         + p[2:, 2:])
    return (s + 4) // 9


def frame(w, h, seed, fid, right=False, jitter=False, nrect=400):
    with np.errstate(over="ignore"):
        base = np.uint64(seed) * _G + np.uint64(fid)
    r = splitmix64(base, nrect * 6).reshape(nrect, 6)
    with np.errstate(over="ignore"):
        jr = splitmix64(base + np.uint64(0xAAAA), nrect * 2).reshape(nrect, 2) if jitter else None
    img = np.full((h, w), 128, np.int32)
    for k in range(nrect):
        cx = int(r[k, 0] % np.uint64(w)); cy = int(r[k, 1] % np.uint64(h))
        hx = 2 + int(r[k, 2] % np.uint64(29)); hy = 2 + int(r[k, 3] % np.uint64(29))
        g = int(r[k, 4] % np.uint64(256)); d = 1 + int(r[k, 5] % np.uint64(40))
        if jitter:
            cx += int(jr[k, 0] % np.uint64(7)) - 3
            cy += int(jr[k, 1] % np.uint64(7)) - 3
        if right:
            cx -= d
        y0, y1, x0, x1 = max(0, cy - hy), max(0, cy + hy), max(0, cx - hx), max(0, cx + hx)
        img[y0:y1, x0:x1] = g
    img = _box3(_box3(img))
    with np.errstate(over="ignore"):
        nseed = base + np.uint64(1 if right else 0) + np.uint64(0x5555)
    noise = (splitmix64(nseed, w * h) % np.uint64(7)).astype(np.int32).reshape(h, w) - 3
    return np.clip(img + noise, 0, 255).astype(np.uint8)


def frames(w, h, seed, ids, right=False):
    return np.stack([frame(w, h, seed, i, right) for i in ids])


def random_descriptors(n, seed):
    return splitmix64(np.uint64(seed) ^ np.uint64(0xD15C), n * 4).view(np.uint8).reshape(n, 32).copy()
