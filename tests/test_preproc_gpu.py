"""GPU parity tests of SURVEY.md §8(f) rank 3 through the C-ABI: orbx_cvt_gray / orbx_remap / orbx_clahe against the CPU
oracle (itself pinned to cv2 in test_oracle_preproc.py) and directly against the committed cv2 outputs.  Bit-exact."""
import importlib
import os
import zlib

import numpy as np
import pytest

import preproc_cases as pc

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "preproc", "preproc_cv2.npz"))


@pytest.fixture(scope="module")
def pre():
    mod = importlib.import_module("3_orb_slam3_selfnote_b200.preproc")
    p = mod.Preprocessor()
    yield p
    p.close()


def _gold(key, img):
    assert int(GOLD[key + "_crc"]) == zlib.crc32(img.tobytes()), key


@pytest.mark.parametrize("name", list(pc.SHAPES))
def test_cvt_gray(pre, oracle, name):
    col = pc.colour_image(name, 4)
    for code in ("bgr", "rgb", "bgra", "rgba"):
        src = np.ascontiguousarray(col[:, :, :len(code)])
        g = pre.cvtColor(src, code)
        _gold("gray_%s_%s" % (name, code), g)
        for coeffs in (oracle.GRAY_OCV346, oracle.GRAY_OCV4):
            assert np.array_equal(pre.cvtColor(src, code, coeffs), oracle.cvt_gray(src, code[0] == "r", coeffs))


@pytest.mark.parametrize("name", list(pc.SHAPES))
def test_remap(pre, oracle, name):
    img = pc.gray_image(name)
    mx, my = pc.rectify_maps(name)
    t = pre.remap_table(mx, my, (img.shape[1], img.shape[0]))
    r = pre.remap(img, t)
    _gold("remap_%s" % name, r)
    assert np.array_equal(r, oracle.remap_linear(img, mx, my))
    # a different destination size than the source, random maps reaching well outside the image
    rng = np.random.default_rng(3)
    dh, dw = img.shape[0] // 2 + 3, img.shape[1] + 5
    mx2 = (rng.random((dh, dw)) * (img.shape[1] + 40) - 20).astype(np.float32)
    my2 = (rng.random((dh, dw)) * (img.shape[0] + 40) - 20).astype(np.float32)
    t2 = pre.remap_table(mx2, my2, (img.shape[1], img.shape[0]))
    assert np.array_equal(pre.remap(img, t2), oracle.remap_linear(img, mx2, my2))
    t.close(), t2.close()


@pytest.mark.parametrize("name", list(pc.SHAPES))
def test_clahe(pre, oracle, name):
    img = pc.gray_image(name)
    for clip, tiles in pc.CLAHE_PARAMS:
        c = pre.clahe(img, clip, tiles)
        _gold("clahe_%s_%g_%dx%d" % (name, clip, tiles[0], tiles[1]), c)
        assert np.array_equal(c, oracle.clahe(img, clip, tiles))
    for clip, tiles in [(0.0, (8, 8)), (3.0, (1, 1)), (3.0, (16, 16)), (1000.0, (2, 2))]:
        assert np.array_equal(pre.clahe(img, clip, tiles), oracle.clahe(img, clip, tiles))


def test_batches_and_flat_images(pre, oracle):
    rng = np.random.default_rng(11)
    col = rng.integers(0, 256, (5, 60, 77, 3), dtype=np.uint8)
    g = pre.cvtColor(col, "bgr")
    for f in range(5):
        assert np.array_equal(g[f], oracle.cvt_gray(col[f], False))
    c = pre.clahe(g, 3.0, (8, 8))
    for f in range(5):
        assert np.array_equal(c[f], oracle.clahe(g[f], 3.0, (8, 8)))
    flat = np.full((48, 64), 200, np.uint8)  # one histogram bin holds everything
    assert np.array_equal(pre.clahe(flat), oracle.clahe(flat))


def test_error_codes(pre, orbx):
    with pytest.raises(orbx.OrbxError) as e:
        pre.clahe(np.zeros((32, 32), np.uint8), 3.0, (64, 64))
    assert e.value.code == orbx.ORBX_ERR_ARG
    with pytest.raises(orbx.OrbxError) as e:
        pre.remap_table(np.zeros((4, 4), np.float32), np.zeros((4, 4), np.float32), (4000, 100))
    assert e.value.code == orbx.ORBX_ERR_UNSUPPORTED


def test_device_pipeline_rectify_then_extract(pre, orbx, oracle):
    """stereo_euroc.cc:166-167 followed by Frame's extraction, without leaving HBM: colour -> gray -> remap (odd pitches, so the
    byte paths run) -> orbx_extract_batch_device, against the oracle run on the same chain."""
    import torch
    name = "euroc"
    h, w = pc.SHAPES[name]
    col = np.stack([pc.colour_image(name, 3), pc.colour_image(name, 3)[::-1].copy()])
    mx, my = pc.rectify_maps(name)
    t = pre.remap_table(mx, my, (w, h))
    st = torch.cuda.Stream()
    d_col = torch.from_numpy(col).cuda()
    gp = w + 3  # unaligned gray pitch
    d_gray = torch.zeros((2, h, gp), dtype=torch.uint8, device="cuda")
    d_rect = torch.zeros((2, h, w), dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    pre.cvt_gray_device(d_col.data_ptr(), 2, w, h, w * 3, h * w * 3, 3, False, d_gray.data_ptr(), gp, h * gp, st.cuda_stream)
    pre.remap_device(t, d_gray.data_ptr(), 2, gp, h * gp, d_rect.data_ptr(), w, h * w, st.cuda_stream)
    gx = orbx.ORBextractor(1000, max_batch=2)
    gx.extract_batch_device(d_rect.data_ptr(), 2, w, h, w, w * h, (0, 0), st.cuda_stream)
    st.synchronize()
    counts, mono, kps, desc = gx.results_host(2)
    rect = d_rect.cpu().numpy()
    ex = oracle.Extractor(1000, blur_taps=oracle.TAPS_OCV346, sincos_mode=oracle.SINCOS_PORTABLE)
    for f in range(2):
        ref = oracle.remap_linear(oracle.cvt_gray(col[f], False), mx, my)
        assert np.array_equal(rect[f], ref)
        _, kps_o, desc_o = ex(ref, (0, 0))
        n = int(counts[f])
        assert n == len(kps_o) > 300
        assert kps[f][:n].tobytes() == kps_o.tobytes() and np.array_equal(desc[f][:n], desc_o)
    gx.close(), t.close()
