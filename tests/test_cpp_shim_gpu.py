"""Builds tests/cpp/shim_test.cc (the reference-signature C++ shims over the C-ABI) with g++, runs it on the GPU and checks
its outputs against the CPU oracle: ORBextractor::operator(), Frame::ComputeStereoMatches and
ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th), both ORBmatcher::SearchByBoW overloads through the same class / member names the reference uses."""
import os
import shutil
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cpp_shim_end_to_end(orbx, oracle, synth, tmp_path):
    cxx = shutil.which("g++") or "/opt/gcc/bin/g++"
    exe = str(tmp_path / "shim_test")
    libdir = os.path.join(ROOT, "3_orb_slam3_selfnote_b200")
    subprocess.check_call([cxx, "-std=c++17", "-O1", os.path.join(ROOT, "tests", "cpp", "shim_test.cc"), "-o", exe, "-L" + libdir,
                           "-lorbx", "-Wl,-rpath," + libdir])
    W, H = 752, 480
    left, right = synth.frame(W, H, 71, 2, False), synth.frame(W, H, 71, 2, True)
    left.tofile(tmp_path / "l.raw"); right.tofile(tmp_path / "r.raw")
    out = str(tmp_path / "out.bin")
    subprocess.check_call([exe, str(tmp_path / "l.raw"), str(tmp_path / "r.raw"), str(W), str(H), out])
    buf = open(out, "rb").read()
    pos = 0

    def take(dtype, count):
        nonlocal pos
        a = np.frombuffer(buf, dtype, count, pos)
        pos += a.nbytes
        return a

    # mono extraction with vLappingArea {0,1000}
    n, mono = take(np.int32, 2)
    kps = take(orbx.KP_DTYPE, n); desc = take(np.uint8, 32 * n).reshape(n, 32)
    ex = oracle.Extractor(1000, sincos_mode=oracle.SINCOS_PORTABLE)
    mono_o, ko, do = ex(left, (0, 1000))
    assert mono == mono_o and kps.tobytes() == ko.tobytes() and np.array_equal(desc, do)
    # stereo
    exL = oracle.Extractor(1200, sincos_mode=oracle.SINCOS_PORTABLE); exR = oracle.Extractor(1200, sincos_mode=oracle.SINCOS_PORTABLE)
    _, kl, dl = exL(left, (0, 0)); _, kr, dr = exR(right, (0, 0))
    N = int(take(np.int32, 1)[0])
    assert N == len(kl)
    uR = take(np.float32, N); depth = take(np.float32, N)
    mbf = np.float32(47.90639384423901); mb = np.float32(mbf / np.float32(435.2046959714599))
    uo, dpo, _, _ = oracle.stereo_matches(exL, exR, kl, dl, kr, dr, mb, mbf)
    assert uR.tobytes() == uo.tobytes() and depth.tobytes() == dpo.tobytes() and (uo >= 0).sum() > 100
    # SearchByProjection against map points built from the right frame; the frame is a rectified-stereo one (mvuRight gate active)
    nm = int(take(np.int32, 1)[0])
    kp_mp = take(np.int32, N)
    nmp = len(kr)
    nm_o, kp_o = oracle.search_by_projection_mappoints(kl, dl, exL.scale, W, H, np.ones(nmp, np.uint8), kr["x"] + np.float32(5), kr["y"],
                                                       kr["octave"], np.where(np.arange(nmp) % 2 == 1, 0.9, 1.0).astype(np.float32), dr,
                                                       3.0, 0.8, uRight=uo, projXR=kr["x"])
    assert nm == nm_o and np.array_equal(kp_mp, kp_o) and nm_o > 50
    assert int(take(np.int32, 1)[0]) == oracle.hamming(dl[0], dr[0])
    # SearchByBoW through the reference signatures: keyframe = left image (map point on every keypoint with i % 5 != 0)
    def fv(d):
        node = d[:, 0] & 31
        ids = np.unique(node)
        start = [0]; ind = []
        for w in ids:
            ind.extend(np.nonzero(node == w)[0].tolist()); start.append(len(ind))
        return ids.astype(np.uint32), np.array(start, np.int32), np.array(ind, np.uint32)
    e1 = (np.arange(len(kl)) % 5 != 0).astype(np.uint8)
    nb = int(take(np.int32, 1)[0]); got = take(np.int32, len(kr))
    nb_o, m12 = oracle.search_by_bow(kl, dl, fv(dl), e1, kr, dr, fv(dr), None, 50, 0.75, True)
    ref = np.full(len(kr), -1, np.int32)
    ref[m12[m12 >= 0]] = np.nonzero(m12 >= 0)[0]
    assert nb == nb_o and np.array_equal(got, ref) and nb_o > 10
    e2 = (np.arange(len(kr)) % 7 != 0).astype(np.uint8)
    nb2 = int(take(np.int32, 1)[0]); got2 = take(np.int32, len(kl))
    nb2_o, m12 = oracle.search_by_bow(kl, dl, fv(dl), e1, kr, dr, fv(dr), e2, 49, 0.75, True)
    assert nb2 == nb2_o and np.array_equal(got2, m12) and nb2_o > 10
