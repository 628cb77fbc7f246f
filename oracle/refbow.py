"""ctypes binding of oracle/_ref/libref_bow.so: the reference's vendored DBoW2 (Thirdparty/DBoW2: TemplatedVocabulary.h, FORB.cpp,
BowVector.cpp, FeatureVector.cpp, ScoringObject.cpp) compiled unmodified against the stand-in cv:: header -- the vocabulary as
ORB-SLAM3 loads it (loadFromTextFile) and the transform of Frame::ComputeBoW (src/Frame.cc:828-835).

TEST INFRASTRUCTURE ONLY (same rule as oracle/__init__.py).  Built by `make -C oracle ref` where /root/reference exists.
"""
import ctypes as C
import os

import numpy as np

from . import ref as _ref

_LIB = os.path.join(_ref._DIR, "libref_bow.so")
_lib = None


def available():
    if not os.path.exists(_LIB) and _ref.can_build():
        _ref.build()
    return os.path.exists(_LIB)


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise RuntimeError("oracle/_ref/libref_bow.so missing: run `make -C oracle ref` in the build container")
        _lib = C.CDLL(_LIB)
        _lib.rb_vocab_load.restype = C.c_void_p
        _lib.rb_vocab_load.argtypes = [C.c_char_p]
        _lib.rb_vocab_destroy.argtypes = [C.c_void_p]
        _lib.rb_vocab_words.argtypes = [C.c_void_p]
    return _lib


def write_text_vocabulary(path, vocab, k):
    """The ORBvoc.txt format loadFromTextFile reads (TemplatedVocabulary.h:1338-1423): header `k L scoring weighting` (0 0 = L1_NORM,
    TF_IDF, what ORBvoc.txt carries), then one line per node in id order: parent, is-leaf, 32 descriptor bytes, weight.  Node ids are the
    line numbers, children attach in file order and word ids count the leaves in file order, so `vocab` (tests/bowsynth.py layout) must
    list every node's children in ascending id order.  No trailing newline: the loader's `while(!f.eof())` would parse an empty last
    line into a node with an uninitialised parent."""
    cs, ci = vocab["child_start"], vocab["child_ids"]
    n = len(cs) - 1
    parent = np.zeros(n, np.int64)
    for p in range(n):
        ch = ci[cs[p]:cs[p + 1]]
        assert list(ch) == sorted(ch), "children must be stored in ascending id order"
        parent[ch] = p
    lines = ["%d %d 0 0" % (k, vocab["L"])]
    for i in range(1, n):
        leaf = cs[i + 1] == cs[i]
        lines.append("%d %d %s %s" % (parent[i], 1 if leaf else 0, " ".join(str(int(b)) for b in vocab["node_desc"][i]),
                                      repr(float(vocab["node_weight"][i])) if leaf else "0"))
    with open(path, "w") as f:
        f.write("\n".join(lines))


class Vocabulary:
    def __init__(self, path):
        self._h = lib().rb_vocab_load(os.fsencode(path))
        if not self._h:
            raise RuntimeError("loadFromTextFile failed: " + path)

    def words(self):
        return lib().rb_vocab_words(C.c_void_p(self._h))

    def transform(self, desc, levelsup):
        """-> (bow_ids, bow_values, (fv_node_ids, fv_start, fv_indices)) in the containers' ascending-key order"""
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = len(desc)
        bi = np.empty(max(n, 1), np.uint32); bv = np.empty(max(n, 1), np.float64)
        fi = np.empty(max(n, 1), np.uint32); fs = np.empty(n + 1, np.int32); fx = np.empty(max(n, 1), np.uint32)
        nb, nf = C.c_int(0), C.c_int(0)
        p = lambda a: a.ctypes.data_as(C.c_void_p)
        lib().rb_transform(C.c_void_p(self._h), p(desc), n, int(levelsup), p(bi), p(bv), C.byref(nb), p(fi), p(fs), p(fx), C.byref(nf))
        return bi[:nb.value].copy(), bv[:nb.value].copy(), (fi[:nf.value].copy(), fs[:nf.value + 1].copy(), fx[:fs[nf.value]].copy())

    def transform_words(self, desc):
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        w = np.empty(max(len(desc), 1), np.uint32)
        lib().rb_transform_words(C.c_void_p(self._h), desc.ctypes.data_as(C.c_void_p), len(desc), w.ctypes.data_as(C.c_void_p))
        return w[:len(desc)]

    def close(self):
        if self._h:
            lib().rb_vocab_destroy(C.c_void_p(self._h))
            self._h = None
