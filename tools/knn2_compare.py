#!/usr/bin/env python3
"""POPC fold vs tensor-core fold of the all-pairs 2-NN through the C-ABI (orbx_knn2_chunk_device), same inputs:

    python tools/knn2_compare.py [nq=16384] [nt=262144] [reps=5] [kernel=both|popc|tensor]

prints one JSON line: pairs/s of k_knn2 (xor + popc) and k_knn2_tc (tcgen05.mma kind::i8), and whether the (best, second) keys are equal.
"""
import importlib
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
orbx = importlib.import_module("3_orb_slam3_selfnote_b200")

nq = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
nt = int(sys.argv[2]) if len(sys.argv) > 2 else 262144
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
which = sys.argv[4] if len(sys.argv) > 4 else "both"
g = torch.Generator(device="cuda"); g.manual_seed(3)
q = torch.randint(0, 256, (nq, 32), dtype=torch.uint8, device="cuda", generator=g)
t = torch.randint(0, 256, (nt, 32), dtype=torch.uint8, device="cuda", generator=g)
m = orbx.ORBmatcher()
st = torch.cuda.Stream()
out, keys = {"nq": nq, "nt": nt, "pairs": float(nq) * nt}, {}
for name, kind in (("popc", m.KNN2_POPC), ("tensor", m.KNN2_TENSOR)):
    if which not in ("both", name):
        continue
    m.set_knn2_kernel(kind)
    k0 = torch.empty(nq, dtype=torch.int64, device="cuda"); k1 = torch.empty_like(k0)
    best = 1e30
    for r in range(reps + 1):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(st):
            e0.record(st)
            m.knn2_chunk_device(q.data_ptr(), nq, t.data_ptr(), nt, 0, 0, k0.data_ptr(), k1.data_ptr(), 1, st.cuda_stream)
            e1.record(st)
        st.synchronize()
        if r > 0:
            best = min(best, e0.elapsed_time(e1))
    keys[name] = (k0.clone(), k1.clone())
    out[name + "_ms"] = best
    out[name + "_pairs_per_s"] = float(nq) * nt / (best * 1e-3)
if len(keys) == 2:
    out["identical_keys"] = bool(torch.equal(keys["popc"][0], keys["tensor"][0]) and torch.equal(keys["popc"][1], keys["tensor"][1]))
    out["speedup"] = out["popc_ms"] / out["tensor_ms"]
print(json.dumps(out))
