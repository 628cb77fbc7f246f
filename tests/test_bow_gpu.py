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
