"""Builds tests/cpp/pre_test.cc (host/FramePreproc.h, the C++ wrappers of SURVEY.md §8-f ranks 2-3) with g++, runs it on the GPU
and checks every output against the CPU oracle."""
import os
import shutil
import subprocess

import numpy as np
import pytest

import preproc_cases as pc

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cpp_preproc_wrappers(orbx, oracle, tmp_path):
    cxx = shutil.which("g++") or "/opt/gcc/bin/g++"
    exe = str(tmp_path / "pre_test")
    libdir = os.path.join(ROOT, "3_orb_slam3_selfnote_b200")
    subprocess.check_call([cxx, "-std=c++17", "-O1", os.path.join(ROOT, "tests", "cpp", "pre_test.cc"), "-o", exe, "-L" + libdir,
                           "-lorbx", "-Wl,-rpath," + libdir])
    H, W = pc.SHAPES["euroc"]
    bgr = pc.colour_image("euroc", 3)
    mx, my = pc.rectify_maps("euroc")
    bgr.tofile(tmp_path / "bgr.raw"); mx.tofile(tmp_path / "mx.f32"); my.tofile(tmp_path / "my.f32")
    out = str(tmp_path / "out.bin")
    subprocess.check_call([exe, str(tmp_path / "bgr.raw"), str(tmp_path / "mx.f32"), str(tmp_path / "my.f32"), str(W), str(H), out])
    buf = open(out, "rb").read()
    pos = 0

    def take(dtype, count):
        nonlocal pos
        a = np.frombuffer(buf, dtype, count, pos)
        pos += a.nbytes
        return a

    gray = take(np.uint8, W * H).reshape(H, W); rect = take(np.uint8, W * H).reshape(H, W); eq = take(np.uint8, W * H).reshape(H, W)
    g_o = oracle.cvt_gray(bgr, False)
    r_o = oracle.remap_linear(g_o, mx, my)
    assert np.array_equal(gray, g_o) and np.array_equal(rect, r_o) and np.array_equal(eq, oracle.clahe(r_o, 3.0, (8, 8)))
    nk = int(take(np.int32, 1)[0])
    keys = take(orbx.KP_DTYPE, nk); un = take(orbx.KP_DTYPE, nk)
    ref = oracle.undistort_points(np.stack([keys["x"], keys["y"]], 1), np.float32(458.654), np.float32(457.296), np.float32(367.215),
                                  np.float32(248.375), np.array([-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05], np.float32))
    assert nk > 100 and un["x"].tobytes() == ref[:, 0].tobytes() and un["y"].tobytes() == ref[:, 1].tobytes()
    assert np.array_equal(un["octave"], keys["octave"]) and np.array_equal(un["size"], keys["size"])
    n, visible = [int(v) for v in take(np.int32, 2)]
    P = take(np.float32, 3 * n); N = take(np.float32, 3 * n); maxd = take(np.float32, n); mind = take(np.float32, n)
    got = dict(in_view=take(np.uint8, n), proj_x=take(np.float32, n), proj_y=take(np.float32, n), proj_xr=take(np.float32, n),
               depth=take(np.float32, n), level=take(np.int32, n), view_cos=take(np.float32, n))
    cam = oracle.FrustumCam()
    cam.Rcw[:] = [1, 0, 0, 0, 1, 0, 0, 0, 1]
    cam.fx, cam.fy, cam.cx, cam.cy = 458.654, 457.296, 367.215, 248.375
    cam.minX, cam.maxX, cam.minY, cam.maxY = 0, W, 0, H
    cam.bf, cam.logScaleFactor, cam.viewingCosLimit, cam.nScaleLevels, cam.logMode = 47.9, 0.18232156, 0.5, 8, oracle.LOG_PORTABLE
    ref = oracle.is_in_frustum(cam, P, N, maxd, mind)
    for k in got:
        assert got[k].tobytes() == ref[k].tobytes(), k
    assert visible == int(ref["in_view"].sum()) and 20 < visible < n
