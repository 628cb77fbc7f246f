"""GPU diagnostic: all-pairs 2-NN vs brute force, prints the first mismatches."""
import importlib, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_allpairs import brute_force, make_desc
orbx = importlib.import_module("3_orb_slam3_selfnote_b200")
ap = importlib.import_module("3_orb_slam3_selfnote_b200.allpairs")
per = 100
desc = make_desc(30 * per, 3)
m = orbx.ORBmatcher()
for nch in (1, 7):
    idx, dst = ap.allpairs_knn2(torch.from_numpy(desc).cuda(), per, matcher=m, nchunks=nch)
    ridx, rdst = brute_force(desc, per)
    bad = np.nonzero((idx != ridx).any(1) | (dst != rdst).any(1))[0]
    print("nchunks", nch, "mismatching rows", len(bad), "of", len(idx))
    for i in bad[:8]:
        print(i, idx[i], dst[i], ridx[i], rdst[i])
idx2, dst2, good = m.knn2(desc[:500], desc[500:])
d = np.unpackbits(desc[:500, None, :] ^ desc[None, 500:, :], axis=2).sum(2).astype(np.int64)
key = (d << 32) | np.arange(d.shape[1])[None]
key.sort(1)
print("plain knn2 ok:", np.array_equal(idx2, key[:, :2] & 0xffffffff), np.array_equal(dst2, key[:, :2] >> 32))
