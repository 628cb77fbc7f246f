"""CPU checks of the oracle's restatement of the §8(f) rank-1 searches against straight Python loops written from the
reference text (src/ORBmatcher.cc:273-475, 839-979, 981-1222, 1524-1590) on small inputs."""
import numpy as np

from bowsynth import feature_vector, triangulation_geometry

W, H = 320, 240


def _small(oracle, synth, fid, jitter=False, right=False):
    ex = oracle.Extractor(300, sincos_mode=oracle.SINCOS_PORTABLE)
    img = synth.frame(W, H, 5, fid, right, jitter)
    _, k, d = ex(img, (0, 0))
    return k, d, ex.scale.copy()


def _ham(a, b):
    return int(np.unpackbits(a ^ b).sum())


def _three_maxima(oracle, hist):
    return set(int(x) for x in oracle.three_maxima(hist))


def _rot_bin(a1, a2):
    rot = np.float32(a1) - np.float32(a2)
    if rot < 0:
        rot = np.float32(rot + np.float32(360.0))
    b = int(np.round(np.float32(rot * np.float32(1.0 / 30))))   # np.round is half-even; ties at .5 cannot occur for bin edges used here
    return 0 if b == 30 else b


def _py_bow(oracle, k1, d1, fv1, e1, k2, d2, fv2, e2, max_dist, ratio, ori):
    m12 = np.full(len(k1), -1, np.int32)
    claimed = np.zeros(len(k2), bool)
    nodes2 = {int(i): n for n, i in enumerate(fv2[0])}
    hist = [[] for _ in range(30)]
    nm = 0
    for a, nid in enumerate(fv1[0]):
        if int(nid) not in nodes2:
            continue
        b = nodes2[int(nid)]
        for i1 in fv1[2][fv1[1][a]:fv1[1][a + 1]]:
            if not e1[i1]:
                continue
            bd1, bi, bd2 = 256, -1, 256
            for i2 in fv2[2][fv2[1][b]:fv2[1][b + 1]]:
                if claimed[i2] or (e2 is not None and not e2[i2]):
                    continue
                dist = _ham(d1[i1], d2[i2])
                if dist < bd1:
                    bd2, bd1, bi = bd1, dist, i2
                elif dist < bd2:
                    bd2 = dist
            if bd1 <= max_dist and np.float32(bd1) < np.float32(ratio) * np.float32(bd2):
                m12[i1] = bi
                claimed[bi] = True
                nm += 1
                if ori:
                    hist[_rot_bin(k1["angle"][i1], k2["angle"][bi])].append(i1)
    if ori:
        keep = _three_maxima(oracle, np.array([len(h) for h in hist], np.int32))
        for b, lst in enumerate(hist):
            if b not in keep:
                for i1 in lst:
                    m12[i1] = -1
                    nm -= 1
    return nm, m12


def test_search_by_bow_matches_python_loops(oracle, synth):
    k1, d1, _ = _small(oracle, synth, 0)
    k2, d2, _ = _small(oracle, synth, 0, jitter=True)
    rng = np.random.default_rng(0)
    for nwords, kfkf, ratio, ori in ((12, False, 0.8, True), (40, True, 0.7, False), (3, True, 0.9, True)):
        fv1, fv2 = feature_vector(d1, nwords, 2), feature_vector(d2, nwords, 2)
        e1 = (rng.uniform(size=len(k1)) < 0.8).astype(np.uint8)
        e2 = (rng.uniform(size=len(k2)) < 0.8).astype(np.uint8) if kfkf else None
        md = 49 if kfkf else 50
        ref = _py_bow(oracle, k1, d1, fv1, e1, k2, d2, fv2, e2, md, ratio, ori)
        got = oracle.search_by_bow(k1, d1, fv1, e1, k2, d2, fv2, e2, md, ratio, ori)
        assert got[0] == ref[0] and np.array_equal(got[1], ref[1])
        assert got[0] > 20


def test_search_for_triangulation_matches_python_loops(oracle, synth):
    k1, d1, scale = _small(oracle, synth, 0)
    k2, d2, _ = _small(oracle, synth, 0, right=True)
    sigma2 = (scale * scale).astype(np.float32)
    F12, ep = triangulation_geometry(W, H, 1)
    fv1, fv2 = feature_vector(d1, 10, 3), feature_vector(d2, 10, 3)
    rng = np.random.default_rng(1)
    e1 = (rng.uniform(size=len(k1)) < 0.8).astype(np.uint8); e2 = (rng.uniform(size=len(k2)) < 0.8).astype(np.uint8)
    s1 = (rng.uniform(size=len(k1)) < 0.5).astype(np.uint8); s2 = (rng.uniform(size=len(k2)) < 0.5).astype(np.uint8)
    f = F12.reshape(9)
    f32 = np.float32

    def epi(kp1, kp2, unc):
        a = f32(f32(kp1["x"] * f[0]) + f32(kp1["y"] * f[3])) + f[6]
        b = f32(f32(kp1["x"] * f[1]) + f32(kp1["y"] * f[4])) + f[7]
        c = f32(f32(kp1["x"] * f[2]) + f32(kp1["y"] * f[5])) + f[8]
        num = f32(f32(f32(a * kp2["x"]) + f32(b * kp2["y"])) + c)
        den = f32(f32(a * a) + f32(b * b))
        if den == 0:
            return False
        return float(f32(f32(num * num) / den)) < 3.84 * float(unc)

    m12 = np.full(len(k1), -1, np.int32)
    nodes2 = {int(i): n for n, i in enumerate(fv2[0])}
    for a, nid in enumerate(fv1[0]):
        if int(nid) not in nodes2:
            continue
        b = nodes2[int(nid)]
        for i1 in fv1[2][fv1[1][a]:fv1[1][a + 1]]:
            if not e1[i1]:
                continue
            best, bi = 50, -1
            for i2 in fv2[2][fv2[1][b]:fv2[1][b + 1]]:
                if not e2[i2]:
                    continue
                dist = _ham(d1[i1], d2[i2])
                if dist > 50 or dist > best:
                    continue
                if not s1[i1] and not s2[i2]:
                    ex, ey = f32(f32(ep[0]) - k2["x"][i2]), f32(f32(ep[1]) - k2["y"][i2])
                    if f32(f32(ex * ex) + f32(ey * ey)) < f32(100 * scale[k2["octave"][i2]]):
                        continue
                if epi(k1[i1], k2[i2], sigma2[k2["octave"][i2]]):
                    bi, best = i2, dist
            m12[i1] = bi
    got_n, got = oracle.search_for_triangulation(k1, d1, fv1, e1, s1, k2, d2, fv2, e2, s2, F12, ep, scale, sigma2, False, False)
    assert np.array_equal(got, m12) and got_n == int((m12 >= 0).sum())
    assert got_n > 3
    # coarse mode accepts every candidate within TH_LOW: at least as many matches
    assert oracle.search_for_triangulation(k1, d1, fv1, e1, s1, k2, d2, fv2, e2, s2, F12, ep, scale, sigma2, True, False)[0] >= got_n


def test_search_window_best_is_gated_argmin(oracle, synth):
    k, d, scale = _small(oracle, synth, 0)
    kj, dj, _ = _small(oracle, synth, 0, jitter=True)
    inv = (np.float32(1) / (scale * scale)).astype(np.float32)
    n = len(kj)
    valid = np.ones(n, np.uint8)
    u, v, lvl = kj["x"].copy(), kj["y"].copy(), kj["octave"].astype(np.int32)
    radius = (np.float32(3) * scale[lvl]).astype(np.float32)
    for gate in (0, 1):
        bi, bd = oracle.search_window_best(k, d, W, H, valid, u, v, radius, lvl, dj, gate, inv)
        for i in range(0, n, 7):
            cand = oracle.features_in_area(k, W, H, float(u[i]), float(v[i]), float(radius[i]))
            best, bidx = 256, -1
            for j in cand:
                if k["octave"][j] < lvl[i] - 1 or k["octave"][j] > lvl[i]:
                    continue
                if gate:
                    ex, ey = np.float32(u[i] - k["x"][j]), np.float32(v[i] - k["y"][j])
                    e2 = np.float32(np.float32(ex * ex) + np.float32(ey * ey))
                    if float(np.float32(e2 * inv[k["octave"][j]])) > 5.99:
                        continue
                dist = _ham(dj[i], d[j])
                if dist < best:
                    best, bidx = dist, j
            assert (bidx, best) == (int(bi[i]), int(bd[i]))


def test_distinctive_descriptors_matches_numpy(oracle):
    """src/MapPoint.cc:385-428 restated with numpy sorting"""
    rng = np.random.default_rng(3)
    counts = np.array([0, 1, 2, 3, 8, 33, 64, 5], np.int32)
    start = np.concatenate([[0], np.cumsum(counts)]).astype(np.int32)
    desc = rng.integers(0, 256, (start[-1], 32), dtype=np.uint8)
    desc[20:30] = desc[20]   # ties
    bi, bm = oracle.distinctive_descriptors(desc, start)
    for p, n in enumerate(counts):
        if n == 0:
            assert bi[p] == -1
            continue
        d = desc[start[p]:start[p + 1]]
        D = np.unpackbits(d[:, None, :] ^ d[None, :, :], axis=2).sum(axis=2)
        med = np.sort(D, axis=1)[:, int(0.5 * (n - 1))]
        assert bi[p] == int(np.argmin(med)) and bm[p] == int(med.min())


def test_vocab_transform_matches_python_descent(oracle):
    """Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1218-1259 restated as a Python loop"""
    from bowsynth import vocabulary
    voc = vocabulary(5, 3, 2, True)
    rng = np.random.default_rng(4)
    desc = np.concatenate([rng.integers(0, 256, (60, 32), dtype=np.uint8), voc["node_desc"][5:25]])
    for levelsup in (0, 1, 2, 5):
        w, nid, wt = oracle.vocab_transform(voc, desc, levelsup)
        for f in range(len(desc)):
            node, lvl, n_at = 0, 0, 0
            while voc["child_start"][node + 1] > voc["child_start"][node]:
                lvl += 1
                ch = voc["child_ids"][voc["child_start"][node]:voc["child_start"][node + 1]]
                dist = [_ham(desc[f], voc["node_desc"][c]) for c in ch]
                node = int(ch[int(np.argmin(dist))])      # argmin = first minimum, like the strict '<' of the reference
                if lvl == voc["L"] - levelsup:
                    n_at = node
            assert (w[f], nid[f], wt[f]) == (voc["node_word_id"][node], n_at, voc["node_weight"][node])
    # fold: TF-IDF sums in feature order, L1 over ascending word ids
    w, nid, wt = oracle.vocab_transform(voc, desc, 1)
    ids, vals, fv = oracle.vocab_fold(w, nid, wt)
    acc = {}
    for f in range(len(desc)):
        if wt[f] > 0:
            acc[int(w[f])] = acc.get(int(w[f]), 0.0) + float(wt[f])
    norm = 0.0
    for k in sorted(acc):
        norm += abs(acc[k])
    assert list(ids) == sorted(acc) and [acc[k] / norm for k in sorted(acc)] == list(vals)
    assert sorted(fv[2].tolist()) == [f for f in range(len(desc)) if wt[f] > 0]
