"""Synthetic stand-ins for the DBoW2 side inputs of the bag-of-words searches (test helper, no reference data involved).

feature_vector(): a one-level "vocabulary" of K seeded random 256-bit words; a descriptor's node is its nearest word
(lowest Hamming distance, lowest word index on ties), which is what DBoW2's tree descent does at one level
(Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1127-1160).  Node ids are made sparse (id*7+3) and every node lists its
features in ascending feature order, like DBoW2::FeatureVector::addFeature (FeatureVector.cpp:28-42).
"""
import numpy as np


def feature_vector(desc, nwords, seed):
    desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    rng = np.random.default_rng(1000 + seed)
    words = rng.integers(0, 256, (nwords, 32), dtype=np.uint8)
    if len(desc) == 0:
        return np.zeros(0, np.uint32), np.zeros(1, np.int32), np.zeros(0, np.uint32)
    bits_d = np.unpackbits(desc, axis=1).astype(np.int16)
    bits_w = np.unpackbits(words, axis=1).astype(np.int16)
    dist = bits_d.sum(1)[:, None] + bits_w.sum(1)[None, :] - 2 * bits_d @ bits_w.T
    node = dist.argmin(axis=1)
    ids, start, indices = [], [0], []
    for w in range(nwords):
        members = np.nonzero(node == w)[0]
        if len(members) == 0:
            continue
        ids.append(w * 7 + 3)
        indices.extend(members.tolist())
        start.append(len(indices))
    return np.array(ids, np.uint32), np.array(start, np.int32), np.array(indices, np.uint32)


def triangulation_geometry(w, h, seed):
    """F12 (row-major) of a mostly horizontal baseline with a small rotation, and the epipole in image 2"""
    rng = np.random.default_rng(2000 + seed)
    F = np.array([[0, 0, 0], [0, 0, -1], [0, 1, 0]], np.float64)      # pure x translation: epipolar lines are image rows
    F += rng.normal(0, 2e-4, (3, 3)) * np.array([[1e-3, 1e-3, 1], [1e-3, 1e-3, 1], [1, 1, 1]])
    ep = (float(w) * 0.3 + float(rng.uniform(-5, 5)), float(h) * 0.5 + float(rng.uniform(-5, 5)))
    return F.astype(np.float32), ep


def vocabulary(k, L, seed, ragged=True):
    """A seeded random k-ary vocabulary tree of depth L, flattened like DBoW2's m_nodes (TemplatedVocabulary.h): node 0 = root;
    children get their ids in breadth-first creation order but are *stored* in a shuffled order per node (ids need not be
    contiguous or sorted in DBoW2 files); with ragged=True some inner nodes have fewer children and some branches stop early
    (leaves above level L), as k-means on few descriptors produces.  Child descriptors are bit-flipped copies of the parent's so
    that the descent is discriminative; a few words get weight 0 (stopped words)."""
    rng = np.random.default_rng(3000 + seed)
    desc = [rng.integers(0, 256, 32, dtype=np.uint8)]
    children = [[]]
    level = [0]
    frontier = [0]
    for l in range(1, L + 1):
        nxt = []
        for p in frontier:
            if ragged and l > 1 and rng.uniform() < 0.05:
                continue                      # early leaf
            kk = k if not ragged else int(rng.integers(max(2, k - 3), k + 1))
            for _ in range(kk):
                d = np.unpackbits(desc[p]).copy()
                flip = rng.choice(256, size=max(4, 96 >> l), replace=False)
                d[flip] ^= 1
                desc.append(np.packbits(d)); children.append([]); level.append(l)
                children[p].append(len(desc) - 1)
                nxt.append(len(desc) - 1)
            rng.shuffle(children[p])
        frontier = nxt
    n = len(desc)
    child_start = np.zeros(n + 1, np.int32)
    child_ids = []
    for i in range(n):
        child_ids.extend(children[i]); child_start[i + 1] = len(child_ids)
    word_id = np.full(n, -1, np.int32)
    leaves = [i for i in range(n) if not children[i]]
    word_id[leaves] = np.arange(len(leaves))
    weight = np.zeros(n, np.float64)
    weight[leaves] = rng.uniform(0.5, 9.0, len(leaves))
    weight[rng.choice(leaves, size=max(1, len(leaves) // 50), replace=False)] = 0.0
    return dict(L=L, child_start=child_start, child_ids=np.array(child_ids, np.int32), node_desc=np.stack(desc),
                node_weight=weight, node_word_id=word_id, level=np.array(level, np.int32))
