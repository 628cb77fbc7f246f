"""GPU parity tests of the Hamming searches (C-ABI matcher section) against the CPU oracle's restatement of
ORBmatcher / Frame (oracle/match_oracle.cc): bit-exact match indices, distances, match counts, uRight / depth floats."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

W, H = 752, 480


@pytest.fixture(scope="module")
def frames(orbx, oracle, synth):
    """keypoints / descriptors of a few synthetic frames (oracle extraction == GPU extraction, tested elsewhere)"""
    out = {}
    ex = oracle.Extractor(1200, sincos_mode=oracle.SINCOS_PORTABLE)
    for name, fid, right, jitter in (("a", 0, False, False), ("a_right", 0, True, False), ("a_jit", 0, False, True), ("b", 1, False, False)):
        img = synth.frame(W, H, 31, fid, right, jitter)
        mono, k, d = ex(img, (0, 0))
        out[name] = (img, k, d)
    out["scale"] = ex.scale.copy()
    return out


def _mappoints(frames, synth, nmp, seed):
    """config-3 style map points: keypoints of the jittered frame + random padding"""
    _, kj, dj = frames["a_jit"]
    rng = np.random.default_rng(seed)
    n0 = len(kj)
    pad = nmp - n0
    projX = np.concatenate([kj["x"], rng.uniform(0, W, pad)]).astype(np.float32)
    projY = np.concatenate([kj["y"], rng.uniform(0, H, pad)]).astype(np.float32)
    level = np.concatenate([kj["octave"], rng.integers(0, 8, pad)]).astype(np.int32)
    desc = np.concatenate([dj, synth.random_descriptors(pad, seed)])
    viewCos = np.where(np.arange(nmp) % 2 == 0, 1.0, 0.9).astype(np.float32)
    in_view = np.ones(nmp, np.uint8)
    in_view[rng.integers(0, nmp, nmp // 20)] = 0
    return in_view, projX, projY, level, viewCos, desc


def test_hamming_batch(orbx, oracle):
    rng = np.random.default_rng(0)
    a = rng.integers(0, 256, (5000, 32), dtype=np.uint8)
    b = rng.integers(0, 256, (5000, 32), dtype=np.uint8)
    b[:100] = a[:100]
    m = orbx.ORBmatcher()
    d = m.hamming_batch(a, b)
    ref = np.unpackbits(a ^ b, axis=1).sum(axis=1)
    assert np.array_equal(d, ref)
    assert all(orbx.ORBmatcher.DescriptorDistance(a[i], b[i]) == oracle.hamming(a[i], b[i]) for i in range(200))
    m.close()


def test_features_in_area_matches_frame_grid(orbx, oracle, frames):
    _, k, d = frames["a"]
    F = orbx.FrameView(k, d, frames["scale"], W, H)
    m = orbx.ORBmatcher()
    rng = np.random.default_rng(1)
    nonempty = 0
    for t in range(120):
        x, y = float(rng.uniform(-30, W + 30)), float(rng.uniform(-30, H + 30))
        r = float(rng.choice([2.5, 7.0, 15.0, 40.0, 100.0]))
        lo, hi = [(-1, -1), (0, 0), (2, 3), (1, -1), (0, 5)][t % 5]
        ref = oracle.features_in_area(k, W, H, x, y, r, lo, hi)
        got = m.GetFeaturesInArea(F, x, y, r, lo, hi)
        assert np.array_equal(ref, got), (x, y, r, lo, hi)
        nonempty += len(ref) > 0
    assert nonempty > 40
    m.close()


@pytest.mark.parametrize("th,ratio,nmp", [(1.0, 0.8, 5000), (3.0, 0.8, 2000), (1.0, 0.6, 700)])
def test_search_by_projection_mappoints(orbx, oracle, synth, frames, th, ratio, nmp):
    _, k, d = frames["a"]
    nmp = max(nmp, len(frames["a_jit"][1]) + 10)
    in_view, projX, projY, level, viewCos, desc = _mappoints(frames, synth, nmp, 5)
    rng = np.random.default_rng(2)
    observed = (rng.uniform(size=nmp) < 0.9).astype(np.uint8)
    kp0 = np.full(len(k), -1, np.int32)
    kp0[rng.integers(0, len(k), 60)] = -2   # keypoints that already hold an observed map point
    kp0[rng.integers(0, len(k), 30)] = -3   # ... and some holding a map point without observations
    nm_o, kp_o = oracle.search_by_projection_mappoints(k, d, frames["scale"], W, H, in_view, projX, projY, level, viewCos, desc, th, ratio,
                                                       kp_mp=kp0, observed=observed)
    F = orbx.FrameView(k, d, frames["scale"], W, H)
    m = orbx.ORBmatcher(ratio)
    nm_g, kp_g = m.SearchByProjection(F, in_view, projX, projY, level, viewCos, desc, th, kp_mp=kp0, observed=observed)
    assert nm_g == nm_o and np.array_equal(kp_g, kp_o)
    assert nm_o > 300
    # stereo frame: the mvuRight gate (ORBmatcher.cc:93-98)
    uR = np.where(rng.uniform(size=len(k)) < 0.5, k["x"] - rng.uniform(0, 40, len(k)), -1).astype(np.float32)
    projXR = (projX - rng.uniform(0, 40, nmp)).astype(np.float32)
    nm_o2, kp_o2 = oracle.search_by_projection_mappoints(k, d, frames["scale"], W, H, in_view, projX, projY, level, viewCos, desc, th, ratio,
                                                         uRight=uR, projXR=projXR)
    F2 = orbx.FrameView(k, d, frames["scale"], W, H, u_right=uR)
    nm_g2, kp_g2 = m.SearchByProjection(F2, in_view, projX, projY, level, viewCos, desc, th, projXR=projXR)
    assert nm_g2 == nm_o2 and np.array_equal(kp_g2, kp_o2) and nm_o2 < nm_o + 100
    m.close()


@pytest.mark.parametrize("mode,any_blocks,dist_th,check", [(0, False, 100, True), (1, False, 100, True), (2, False, 100, False),
                                                          (0, True, 64, True)])
def test_search_by_projection_window(orbx, oracle, frames, mode, any_blocks, dist_th, check):
    _, k, d = frames["a"]
    _, kj, dj = frames["a_jit"]
    rng = np.random.default_rng(3)
    n = len(kj)
    valid = (rng.uniform(size=n) < 0.93).astype(np.uint8)
    u = (kj["x"] + rng.uniform(-2, 2, n)).astype(np.float32)
    v = (kj["y"] + rng.uniform(-2, 2, n)).astype(np.float32)
    lvl = kj["octave"].astype(np.int32)
    ang = kj["angle"].astype(np.float32)
    kp0 = np.full(len(k), -1, np.int32)
    kp0[rng.integers(0, len(k), 40)] = -2
    for th in (7.0, 15.0):
        nm_o, kp_o = oracle.search_by_projection_window(k, d, frames["scale"], W, H, valid, u, v, lvl, ang, dj, th, mode, dist_th,
                                                        any_blocks, check, kp_mp=kp0)
        F = orbx.FrameView(k, d, frames["scale"], W, H)
        m = orbx.ORBmatcher(0.9, check)
        nm_g, kp_g = m.SearchByProjectionWindow(F, valid, u, v, lvl, ang, dj, th, mode, dist_th, any_blocks, kp_mp=kp0)
        assert nm_g == nm_o and np.array_equal(kp_g, kp_o), (th, nm_g, nm_o)
        assert nm_o > 200
        m.close()


def test_search_for_initialization(orbx, oracle, synth):
    ex = oracle.Extractor(5000, sincos_mode=oracle.SINCOS_PORTABLE)   # the mono-init extractor (Tracking.cc:844)
    _, k1, d1 = ex(synth.frame(W, H, 41, 0), (0, 1000))
    _, k2, d2 = ex(synth.frame(W, H, 41, 0, False, True), (0, 1000))
    prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)        # Tracking.cc:2427-2429
    for win, ratio, check in ((100, 0.9, True), (30, 0.9, False)):
        nm_o, m12_o, pm_o = oracle.search_for_initialization(k1, d1, k2, d2, W, H, prev, win, ratio, check)
        m = orbx.ORBmatcher(ratio, check)
        F1 = orbx.FrameView(k1, d1, ex.scale, W, H)
        F2 = orbx.FrameView(k2, d2, ex.scale, W, H)
        nm_g, m12_g, pm_g = m.SearchForInitialization(F1, F2, prev, win)
        assert nm_g == nm_o and np.array_equal(m12_g, m12_o) and np.array_equal(pm_g, pm_o)
        assert nm_o > 100
        m.close()


def test_knn2_and_ratio(orbx, oracle, frames):
    _, k, d = frames["a"]
    _, kr, dr = frames["a_right"]
    m = orbx.ORBmatcher()
    idx, dist, good = m.knn2(d, dr)
    idx_o, dist_o, good_o = oracle.knn2_ratio(d, dr)
    assert np.array_equal(idx, idx_o) and np.array_equal(dist, dist_o) and np.array_equal(good, good_o)
    # exact ties and degenerate train sets
    t = dr[:300].copy(); t[50] = t[10]; t[120] = t[10]
    q = d[:64].copy(); q[3] = t[10]; q[:, 8:] = 0; t[:, 8:] = 0
    a, b, _ = m.knn2(q, t)
    ao, bo = oracle.knn2(q, t)
    assert np.array_equal(a, ao) and np.array_equal(b, bo)
    a, b, g = m.knn2(q, t[:1])
    assert (a[:, 1] == -1).all() and (b[:, 1] == -1).all() and not g.any()
    a, b, g = m.knn2(q, t[:0])
    assert (a == -1).all()
    m.close()


def test_stereo_matches_batch(orbx, oracle, synth):
    """Frame::ComputeStereoMatches on EuRoC-shaped rectified pairs (bf 47.906, fx 435.2, Examples/Stereo/EuRoC.yaml:9,28)"""
    n = 5
    left = synth.frames(W, H, 51, range(n), False)
    right = synth.frames(W, H, 51, range(n), True)
    gl = orbx.ORBextractor(1200, max_batch=n)
    gr = orbx.ORBextractor(1200, max_batch=n)
    cl, _, kl, dl = gl.extract_batch(left, (0, 0))
    cr, _, kr, dr = gr.extract_batch(right, (0, 0))
    mbf, fx = np.float32(47.90639384423901), np.float32(435.2046959714599)
    mb = np.float32(mbf / fx)
    m = orbx.ORBmatcher()
    m.stereo_match_batch(gl, gr, mb, mbf)
    uR, depth, bd, bi = m.stereo_results(n)
    exL = oracle.Extractor(1200, sincos_mode=oracle.SINCOS_PORTABLE)
    exR = oracle.Extractor(1200, sincos_mode=oracle.SINCOS_PORTABLE)
    matched = 0
    for f in range(n):
        _, kol, dol = exL(left[f], (0, 0))
        _, kor, dor = exR(right[f], (0, 0))
        assert kol.tobytes() == kl[f, :cl[f]].tobytes() and kor.tobytes() == kr[f, :cr[f]].tobytes()
        uo, do, bdo, bio = oracle.stereo_matches(exL, exR, kol, dol, kor, dor, mb, mbf)
        nL = len(kol)
        assert np.array_equal(bi[f, :nL], bio), "Hamming-stage best index, frame %d" % f
        has = bio >= 0
        assert np.array_equal(bd[f, :nL][has], bdo[has])
        assert uR[f, :nL].tobytes() == uo.tobytes(), "mvuRight, frame %d" % f
        assert depth[f, :nL].tobytes() == do.tobytes(), "mvDepth, frame %d" % f
        matched += int((uo >= 0).sum())
    assert matched > 150 * n
    gl.close(); gr.close(); m.close()


def test_search_by_projection_batch_on_device_results(orbx, oracle, synth):
    """config 3 shape: 512x512, 1500 features, 5000 projected map points per frame, th = 1, ratio 0.8 (Tracking.cc:3508)"""
    n, w, h, nmp = 3, 512, 512, 5000
    imgs = synth.frames(w, h, 61, range(n))
    gx = orbx.ORBextractor(1500, max_batch=n)
    counts, _, kps, desc = gx.extract_batch(imgs, (0, 0))
    ex = oracle.Extractor(1500, sincos_mode=oracle.SINCOS_PORTABLE)
    rng = np.random.default_rng(7)
    arrs = []
    for f in range(n):
        _, kj, dj = ex(synth.frame(w, h, 61, f, False, True), (0, 0))
        pad = nmp - len(kj)
        arrs.append((np.ones(nmp, np.uint8), np.concatenate([kj["x"], rng.uniform(0, w, pad)]).astype(np.float32),
                     np.concatenate([kj["y"], rng.uniform(0, h, pad)]).astype(np.float32),
                     np.concatenate([kj["octave"], rng.integers(0, 8, pad)]).astype(np.int32),
                     np.where(np.arange(nmp) % 2 == 0, 1.0, 0.9).astype(np.float32),
                     np.concatenate([dj, synth.random_descriptors(pad, f)])))
    stack = [np.stack([a[i] for a in arrs]) for i in range(6)]
    m = orbx.ORBmatcher(0.8)
    nm, kp_mp = m.SearchByProjectionBatch(gx, n, *stack, th=1.0)
    for f in range(n):
        c = counts[f]
        nm_o, kp_o = oracle.search_by_projection_mappoints(kps[f, :c], desc[f, :c], ex.scale, w, h, *arrs[f], 1.0, 0.8)
        assert nm[f] == nm_o and np.array_equal(kp_mp[f, :c], kp_o)
        assert nm_o > 300
    gx.close(); m.close()


def test_stereo_frames_host_pipeline(orbx, oracle, synth):
    """orbx_stereo_frames_host: chunked (4,4,3), double-buffered; every frame must equal the oracle's stereo frame"""
    n = 11
    left = synth.frames(W, H, 52, range(n), False)
    right = synth.frames(W, H, 52, range(n), True)
    gl = orbx.ORBextractor(1200, max_batch=4); gr = orbx.ORBextractor(1200, max_batch=4)
    m = orbx.ORBmatcher()
    mbf = np.float32(47.90639384423901); mb = np.float32(mbf / np.float32(435.2046959714599))
    o = m.stereo_frames_host(gl, gr, left, right, mb, mbf)
    exL = oracle.Extractor(1200, sincos_mode=oracle.SINCOS_PORTABLE); exR = oracle.Extractor(1200, sincos_mode=oracle.SINCOS_PORTABLE)
    for f in range(n):
        _, kl, dl = exL(left[f], (0, 0)); _, kr, dr = exR(right[f], (0, 0))
        uo, do, _, _ = oracle.stereo_matches(exL, exR, kl, dl, kr, dr, mb, mbf)
        cl, cr = int(o["countsL"][f]), int(o["countsR"][f])
        assert cl == len(kl) and cr == len(kr)
        assert o["kpsL"][f, :cl].tobytes() == kl.tobytes() and o["kpsR"][f, :cr].tobytes() == kr.tobytes()
        assert np.array_equal(o["descL"][f, :cl], dl) and np.array_equal(o["descR"][f, :cr], dr)
        assert o["uRight"][f, :cl].tobytes() == uo.tobytes() and o["depth"][f, :cl].tobytes() == do.tobytes()
    # the right image's records are optional (rectified stereo reads only mvuRight / mvDepth afterwards): same left outputs
    o2 = m.stereo_frames_host(gl, gr, left, right, mb, mbf, right_records=False)
    assert o2["kpsR"] is None and np.array_equal(o2["countsR"], o["countsR"]) and np.array_equal(o2["countsL"], o["countsL"])
    for f in range(n):
        cl = int(o["countsL"][f])
        assert o2["kpsL"][f, :cl].tobytes() == o["kpsL"][f, :cl].tobytes() and o2["uRight"][f, :cl].tobytes() == o["uRight"][f, :cl].tobytes()
        assert o2["depth"][f, :cl].tobytes() == o["depth"][f, :cl].tobytes()
    gl.close(); gr.close(); m.close()


def test_batch256_matches_oracle(orbx, oracle, synth):
    """parity at the bench's own scale: ONE 256-frame extract_batch_device per image side + one stereo_match_batch launch; 8 sampled
    frames against the oracle and, where built, against the reference's own source (oracle/_ref)"""
    import torch
    from oracle import refmatch
    B = 256
    distinct = 32
    left = np.concatenate([synth.frames(W, H, 1000, range(distinct), False)] * (B // distinct))
    right = np.concatenate([synth.frames(W, H, 1000, range(distinct), True)] * (B // distinct))
    dl_, dr_ = torch.from_numpy(left).cuda(), torch.from_numpy(right).cuda()
    gl = orbx.ORBextractor(1200, max_batch=B); gr = orbx.ORBextractor(1200, max_batch=B)
    m = orbx.ORBmatcher()
    mbf = np.float32(47.90639384423901); mb = np.float32(mbf / np.float32(435.2046959714599))
    st = torch.cuda.current_stream().cuda_stream
    gl.extract_batch_device(dl_.data_ptr(), B, W, H, W, W * H, (0, 0), st)
    gr.extract_batch_device(dr_.data_ptr(), B, W, H, W, W * H, (0, 0), st)
    m.stereo_match_batch(gl, gr, mb, mbf, st)
    torch.cuda.synchronize()
    cL, _, kL, dL = gl.results_host(B)
    cR, _, kR, dR = gr.results_host(B)
    uR, depth, _, _ = m.stereo_results(B)
    exL = oracle.Extractor(1200, sincos_mode=oracle.SINCOS_PORTABLE); exR = oracle.Extractor(1200, sincos_mode=oracle.SINCOS_PORTABLE)
    ref = refmatch.StereoFrames(1200) if refmatch.available() else None
    for f in (0, 1, 31, 32, 100, 177, 254, 255):
        _, kl, dl = exL(left[f], (0, 0)); _, kr, dr = exR(right[f], (0, 0))
        uo, do, _, _ = oracle.stereo_matches(exL, exR, kl, dl, kr, dr, mb, mbf)
        cl, cr = int(cL[f]), int(cR[f])
        assert cl == len(kl) and cr == len(kr)
        assert kL[f, :cl].tobytes() == kl.tobytes() and kR[f, :cr].tobytes() == kr.tobytes()
        assert np.array_equal(dL[f, :cl], dl) and np.array_equal(dR[f, :cr], dr)
        assert uR[f, :cl].tobytes() == uo.tobytes() and depth[f, :cl].tobytes() == do.tobytes()
        if ref is not None and f in (0, 100, 255):
            r = ref(left[f], right[f], float(mb), float(mbf))
            assert r["kpsL"].tobytes() == kl.tobytes() and r["uRight"].tobytes() == uo.tobytes() and r["depth"].tobytes() == do.tobytes()
    # frames that repeat an image must repeat its result (32 distinct pairs tiled 8 times)
    for f in range(distinct, B):
        assert cL[f] == cL[f % distinct] and uR[f, :cL[f]].tobytes() == uR[f % distinct, :cL[f]].tobytes()
    gl.close(); gr.close(); m.close()


def test_device_batches_pipelined_over_two_slots(orbx, synth):
    """orbx_extract_batch_device_slot: launches alternate the handles' two workspace slots, the stereo matcher runs on its own stream while
    the next launch is extracted (bench.py's device-resident loop).  Every launch's results must equal the serial single-slot run."""
    import torch
    B, NL = 24, 5
    left = [np.stack([synth.frame(W, H, 300 + k, i % 5, False) for i in range(B)]) for k in range(NL)]
    right = [np.stack([synth.frame(W, H, 300 + k, i % 5, True) for i in range(B)]) for k in range(NL)]
    dL = [torch.from_numpy(a).cuda() for a in left]; dR = [torch.from_numpy(a).cuda() for a in right]
    mbf = np.float32(47.90639384423901); mb = np.float32(mbf / np.float32(435.2046959714599))
    gl = orbx.ORBextractor(1200, max_batch=B); gr = orbx.ORBextractor(1200, max_batch=B); m = orbx.ORBmatcher()
    st = torch.cuda.Stream().cuda_stream
    serial = []
    for k in range(NL):
        gl.extract_batch_device(dL[k].data_ptr(), B, W, H, W, W * H, (0, 0), st)
        gr.extract_batch_device(dR[k].data_ptr(), B, W, H, W, W * H, (0, 0), st)
        m.stereo_match_batch(gl, gr, mb, mbf, st)
        torch.cuda.synchronize()
        serial.append((gl.results_host(B), gr.results_host(B), m.stereo_results(B)))
    sA, sB, sS = torch.cuda.Stream(), torch.cuda.Stream(), torch.cuda.Stream()
    evL, evR, evDone = [torch.cuda.Event() for _ in range(2)], [torch.cuda.Event() for _ in range(2)], [torch.cuda.Event() for _ in range(2)]
    got = []
    for k in range(NL):
        slot = k & 1
        sA.wait_event(evDone[slot]); sB.wait_event(evDone[slot])
        gl.extract_batch_device(dL[k].data_ptr(), B, W, H, W, W * H, (0, 0), sA.cuda_stream, slot=slot)
        evL[slot].record(sA)
        gr.extract_batch_device(dR[k].data_ptr(), B, W, H, W, W * H, (0, 0), sB.cuda_stream, slot=slot)
        evR[slot].record(sB)
        sS.wait_event(evL[slot]); sS.wait_event(evR[slot])
        m.stereo_match_batch(gl, gr, mb, mbf, sS.cuda_stream)
        evDone[slot].record(sS)
    torch.cuda.synchronize()
    # results are read from the slot used last: run the pipeline again stopping after launch `stop`, for every stop
    for stop in range(NL):
        for k in range(stop + 1):
            slot = k & 1
            sA.wait_event(evDone[slot]); sB.wait_event(evDone[slot])
            gl.extract_batch_device(dL[k].data_ptr(), B, W, H, W, W * H, (0, 0), sA.cuda_stream, slot=slot)
            evL[slot].record(sA)
            gr.extract_batch_device(dR[k].data_ptr(), B, W, H, W, W * H, (0, 0), sB.cuda_stream, slot=slot)
            evR[slot].record(sB)
            sS.wait_event(evL[slot]); sS.wait_event(evR[slot])
            m.stereo_match_batch(gl, gr, mb, mbf, sS.cuda_stream)
            evDone[slot].record(sS)
        torch.cuda.synchronize()
        got.append((gl.results_host(B), gr.results_host(B), m.stereo_results(B)))
    for k in range(NL):
        (cl, ml, kl, dl), (cr, mr, kr, dr), (ur, dp, _, _) = serial[k]
        (cl2, ml2, kl2, dl2), (cr2, mr2, kr2, dr2), (ur2, dp2, _, _) = got[k]
        assert np.array_equal(cl, cl2) and np.array_equal(cr, cr2) and cl.min() > 300
        for f in range(B):
            n, nr = int(cl[f]), int(cr[f])
            assert kl[f, :n].tobytes() == kl2[f, :n].tobytes() and np.array_equal(dl[f, :n], dl2[f, :n]), (k, f)
            assert kr[f, :nr].tobytes() == kr2[f, :nr].tobytes() and np.array_equal(dr[f, :nr], dr2[f, :nr]), (k, f)
            assert ur[f, :n].tobytes() == ur2[f, :n].tobytes() and dp[f, :n].tobytes() == dp2[f, :n].tobytes(), (k, f)
    gl.close(); gr.close(); m.close()


def test_host_buffers_and_chunked_pipeline(orbx, oracle, synth):
    """orbx_host_alloc staging buffers + orbx_stereo_frames_host over several ramped / tail chunks (graph cache hit on the
    second call) against the per-frame oracle."""
    W, H, n = 752, 480, 11
    left = np.stack([synth.frame(W, H, 55, i % 3, False) for i in range(n)])
    right = np.stack([synth.frame(W, H, 55, i % 3, True) for i in range(n)])
    hl, hr = orbx.HostBuffer((n, H, W)), orbx.HostBuffer((n, H, W))
    hl.array[...] = left
    hr.array[...] = right
    mbf = np.float32(47.90639384423901); mb = np.float32(mbf / np.float32(435.2046959714599))
    exL, exR, mt = orbx.ORBextractor(1200, max_batch=8), orbx.ORBextractor(1200, max_batch=8), orbx.ORBmatcher()   # chunks 2, 8, 1
    ref = []
    for i in range(3):
        oL = oracle.Extractor(1200, sincos_mode=oracle.SINCOS_PORTABLE); oR = oracle.Extractor(1200, sincos_mode=oracle.SINCOS_PORTABLE)
        _, kl, dl = oL(left[i], (0, 0)); _, kr, dr = oR(right[i], (0, 0))
        uo, dpo, _, _ = oracle.stereo_matches(oL, oR, kl, dl, kr, dr, mb, mbf)
        ref.append((kl, dl, kr, dr, uo, dpo))
    for rep in range(2):
        o = mt.stereo_frames_host(exL, exR, hl.array, hr.array, mb, mbf)
        for i in range(n):
            kl, dl, kr, dr, uo, dpo = ref[i % 3]
            c = int(o["countsL"][i])
            assert c == len(kl) and int(o["countsR"][i]) == len(kr)
            assert o["kpsL"][i, :c].tobytes() == kl.tobytes() and np.array_equal(o["descL"][i, :c], dl)
            assert o["kpsR"][i, :len(kr)].tobytes() == kr.tobytes() and np.array_equal(o["descR"][i, :len(kr)], dr)
            assert o["uRight"][i, :c].tobytes() == uo.tobytes() and o["depth"][i, :c].tobytes() == dpo.tobytes()
    for b in (hl, hr):
        b.close()
    exL.close(); exR.close(); mt.close()


def test_concurrent_host_threads_on_independent_handles(orbx, synth):
    """Handles are independent (include/orbx.h): several host threads, each with its own extractor / matcher handles, run the
    host-buffer pipeline at once -- including their first calls (geometry set-up, allocation, graph capture) -- and every call
    returns exactly what a single thread gets."""
    import threading
    W, H, n = 752, 480, 12
    left = np.stack([synth.frame(W, H, 77, i % 4, False) for i in range(n)])
    right = np.stack([synth.frame(W, H, 77, i % 4, True) for i in range(n)])
    mbf = np.float32(47.90639384423901); mb = np.float32(mbf / np.float32(435.2046959714599))
    exL, exR, mt = orbx.ORBextractor(1200, max_batch=4), orbx.ORBextractor(1200, max_batch=4), orbx.ORBmatcher()
    ref = mt.stereo_frames_host(exL, exR, left, right, mb, mbf)
    exL.close(); exR.close(); mt.close()
    errs = []

    def worker(k):
        try:
            a, b, m = orbx.ORBextractor(1200, max_batch=4), orbx.ORBextractor(1200, max_batch=4), orbx.ORBmatcher()
            for rep in range(4):
                o = m.stereo_frames_host(a, b, left, right, mb, mbf)
                for key in ("countsL", "countsR"):
                    assert np.array_equal(o[key], ref[key]), (k, rep, key)
                for i in range(n):
                    c = int(ref["countsL"][i])
                    assert o["kpsL"][i, :c].tobytes() == ref["kpsL"][i, :c].tobytes()
                    assert np.array_equal(o["descL"][i, :c], ref["descL"][i, :c])
                    assert o["uRight"][i, :c].tobytes() == ref["uRight"][i, :c].tobytes()
            a.close(); b.close(); m.close()
        except Exception as e:   # noqa: BLE001
            errs.append(repr(e))

    ths = [threading.Thread(target=worker, args=(k,)) for k in range(4)]
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    assert not errs, errs
