"""GPU parity tests of SURVEY.md §8(f) rank 4: MapPoint::ComputeDistinctiveDescriptors and the DBoW2 tree descent
(Frame::ComputeBoW) against the CPU oracle; the BoW output then drives SearchByBoW end to end."""
import numpy as np
import pytest

from bowsynth import vocabulary

pytestmark = pytest.mark.gpu

W, H = 752, 480


def _observations(rng, npoints, max_obs):
    """descriptors of map points: noisy copies of a per-point prototype, 0..max_obs observations each"""
    counts = rng.integers(0, max_obs + 1, npoints)
    counts[:4] = [0, 1, 2, max_obs]
    start = np.concatenate([[0], np.cumsum(counts)]).astype(np.int32)
    desc = np.empty((start[-1], 32), np.uint8)
    for p in range(npoints):
        proto = np.unpackbits(rng.integers(0, 256, 32, dtype=np.uint8))
        for r in range(start[p], start[p + 1]):
            d = proto.copy()
            d[rng.choice(256, size=int(rng.integers(0, 40)), replace=False)] ^= 1
            desc[r] = np.packbits(d)
    return desc, start


@pytest.mark.parametrize("npoints,max_obs", [(3000, 12), (200, 150), (1, 700)])
def test_distinctive_descriptors(orbx, oracle, npoints, max_obs):
    rng = np.random.default_rng(npoints)
    desc, start = _observations(rng, max(npoints, 4), max_obs)
    ref_i, ref_m = oracle.distinctive_descriptors(desc, start)
    m = orbx.ORBmatcher()
    bi, bm = m.DistinctiveDescriptors(desc, start)
    assert np.array_equal(bi, ref_i) and np.array_equal(bm, ref_m)
    # duplicated observations: every median ties, the first row must win
    d2 = np.repeat(desc[:1], 9, axis=0)
    bi, bm = m.DistinctiveDescriptors(d2, np.array([0, 9], np.int32))
    assert bi[0] == 0 and bm[0] == 0
    m.close()


@pytest.mark.parametrize("k,L,levelsup,ragged", [(10, 4, 2, False), (10, 3, 4, False), (7, 5, 3, True), (33, 2, 1, True)])
def test_vocabulary_transform(orbx, oracle, synth, k, L, levelsup, ragged):
    voc = vocabulary(k, L, 5, ragged)
    ex = oracle.Extractor(1500, sincos_mode=oracle.SINCOS_PORTABLE)
    _, kps, desc = ex(synth.frame(W, H, 31, 0, False), (0, 0))
    rng = np.random.default_rng(1)
    # mix of real descriptors, exact node descriptors (distance-0 hits, ties between equal children) and random rows
    desc = np.concatenate([desc, voc["node_desc"][rng.integers(0, len(voc["node_desc"]), 200)], synth.random_descriptors(300, 2)])
    rw, rn, rwt = oracle.vocab_transform(voc, desc, levelsup)
    m = orbx.ORBmatcher()
    V = orbx.ORBVocabulary(voc)
    w, nid, wt = V.transform_features(m, desc, levelsup)
    assert np.array_equal(w, rw) and np.array_equal(nid, rn) and wt.tobytes() == rwt.tobytes()
    if L - levelsup > 0:
        assert (voc["level"][nid] <= L - levelsup).all() and len(np.unique(nid)) > 1
    else:
        assert (nid == 0).all()
    # BowVector / FeatureVector fold (TF_IDF + L1): ids, CSR and the doubles bit for bit
    bi, bv, fv = V.transform(m, desc, levelsup)
    obi, obv, ofv = oracle.vocab_fold(rw, rn, rwt)
    assert np.array_equal(bi, obi) and bv.tobytes() == obv.tobytes()
    assert all(np.array_equal(a, b) for a, b in zip(fv, ofv))
    assert abs(bv.sum() - 1.0) < 1e-9
    V.close(); m.close()


def test_bow_pipeline_feeds_search_by_bow(orbx, oracle, synth):
    """extract two frames -> ComputeBoW on the GPU -> SearchByBoW on the GPU == the same chain in the oracle"""
    voc = vocabulary(10, 4, 9, False)
    ex = oracle.Extractor(1200, sincos_mode=oracle.SINCOS_PORTABLE)
    _, k1, d1 = ex(synth.frame(W, H, 31, 0, False), (0, 0))
    _, k2, d2 = ex(synth.frame(W, H, 31, 0, False, True), (0, 0))
    m = orbx.ORBmatcher(0.75, True)
    V = orbx.ORBVocabulary(voc)
    fv1 = V.transform(m, d1, 2)[2]; fv2 = V.transform(m, d2, 2)[2]
    o1 = oracle.vocab_fold(*oracle.vocab_transform(voc, d1, 2))[2]; o2 = oracle.vocab_fold(*oracle.vocab_transform(voc, d2, 2))[2]
    e1 = np.ones(len(k1), np.uint8)
    ref = oracle.search_by_bow(k1, d1, o1, e1, k2, d2, o2, None, 50, 0.75, True)
    got = m.SearchByBoW(k1, d1, fv1, e1, k2, d2, fv2, None, 50)
    assert got[0] == ref[0] and np.array_equal(got[1], ref[1]) and ref[0] > 30
    V.close(); m.close()


def test_vocabulary_rejects_malformed_trees(orbx):
    voc = vocabulary(4, 2, 1, False)
    bad = dict(voc); bad["child_ids"] = voc["child_ids"].copy(); bad["child_ids"][3] = bad["child_ids"][2]   # shared child
    with pytest.raises(Exception):
        orbx.ORBVocabulary(bad)
    bad = dict(voc); bad["L"] = 1                                                                            # deeper than L
    with pytest.raises(Exception):
        orbx.ORBVocabulary(bad)


@pytest.mark.parametrize("k,L,levelsup", [(10, 4, 4), (10, 3, 2), (5, 5, 1)])
def test_vocabulary_transform_vs_reference_dbow2(orbx, oracle, synth, tmp_path, k, L, levelsup):
    """Frame::ComputeBoW on the GPU against the reference's OWN vendored DBoW2 (oracle/_ref/libref_bow.so: Thirdparty/DBoW2 compiled
    unmodified; the vocabulary goes through loadFromTextFile as ORBvoc.txt does): BowVector ids and doubles, FeatureVector nodes and
    feature lists, bit for bit.  levelsup = 4 is what Frame.cc:833 passes."""
    from oracle import refbow
    if not refbow.available():
        pytest.skip("oracle/_ref/libref_bow.so not prebuilt (needs /root/reference at build time)")
    voc = vocabulary(k, L, 11, False)
    ci, cs = voc["child_ids"].copy(), voc["child_start"]
    for p in range(len(cs) - 1):          # a text vocabulary lists children in id order
        ci[cs[p]:cs[p + 1]] = np.sort(ci[cs[p]:cs[p + 1]])
    voc["child_ids"] = ci
    path = str(tmp_path / "voc.txt")
    refbow.write_text_vocabulary(path, voc, k)
    rv = refbow.Vocabulary(path)
    ex = oracle.Extractor(1500, sincos_mode=oracle.SINCOS_PORTABLE)
    _, kps, desc = ex(synth.frame(W, H, 31, 3, False), (0, 0))
    rng = np.random.default_rng(2)
    desc = np.concatenate([desc, voc["node_desc"][rng.integers(1, len(voc["node_desc"]), 200)], synth.random_descriptors(300, 4)])
    m = orbx.ORBmatcher()
    V = orbx.ORBVocabulary(voc)
    try:
        bi, bv, fv = V.transform(m, desc, levelsup)
        rbi, rbv, rfv = rv.transform(desc, levelsup)
        assert np.array_equal(bi, rbi) and bv.tobytes() == rbv.tobytes() and len(bi) > 50
        assert all(np.array_equal(a, b) for a, b in zip(fv, rfv))
        w, _, _ = V.transform_features(m, desc, levelsup)
        assert np.array_equal(w.astype(np.uint32), rv.transform_words(desc))
    finally:
        V.close(); m.close(); rv.close()


def test_distinctive_descriptors_vs_reference_body(orbx):
    """k_distinctive against the reference's OWN MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:350-436, cut out of the
    reference's file into oracle/_ref/libref_match.so): the descriptor the reference stores in mDescriptor is the row the GPU picks,
    for clustered observations, tie families (duplicated rows) and single observations."""
    from oracle import refmatch
    if not refmatch.available():
        pytest.skip("oracle/_ref/libref_match.so not prebuilt (needs /root/reference at build time)")
    rng = np.random.default_rng(77)
    desc, start = _observations(rng, 400, 40)
    for p in range(0, 400, 7):            # tie families
        a, b = int(start[p]), int(start[p + 1])
        desc[a:b] = desc[a]
        if b - a > 2:
            desc[b - 1, 3] ^= 5
    ref_desc, has = refmatch.distinctive(desc, start)
    m = orbx.ORBmatcher()
    try:
        bi, bm = m.DistinctiveDescriptors(desc, start)
        for p in range(400):
            n = int(start[p + 1] - start[p])
            assert bool(has[p]) == (n > 0)
            if n:
                assert np.array_equal(desc[start[p] + bi[p]], ref_desc[p]), p
    finally:
        m.close()
