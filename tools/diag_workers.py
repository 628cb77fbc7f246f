#!/usr/bin/env python3
"""Concurrency diagnosis: N host threads, each with its own handles, run orbx_stereo_frames_host repeatedly on shared input
buffers; every call's outputs are compared with a single-threaded run.   python tools/diag_workers.py [B=128] [workers=2] [chunks=4] [calls=13]"""
import importlib
import os
import sys
import threading

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
orbx = importlib.import_module("3_orb_slam3_selfnote_b200")
synth = importlib.import_module("3_orb_slam3_selfnote_b200.synth")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
NW = int(sys.argv[2]) if len(sys.argv) > 2 else 2
CH = int(sys.argv[3]) if len(sys.argv) > 3 else 4
CALLS = int(sys.argv[4]) if len(sys.argv) > 4 else 13
W, H = 752, 480
mbf = np.float32(47.90639384423901); mb = np.float32(mbf / np.float32(435.2046959714599))
tile = lambda a: np.concatenate([a] * ((B + 31) // 32))[:B]
hl, hr = orbx.HostBuffer((B, H, W)), orbx.HostBuffer((B, H, W))
hl.array[...] = tile(synth.frames(W, H, 1000, range(32), False))
hr.array[...] = tile(synth.frames(W, H, 1000, range(32), True))
nb = max(1, B // CH)
a, b, m = orbx.ORBextractor(1200, max_batch=nb), orbx.ORBextractor(1200, max_batch=nb), orbx.ORBmatcher()
ref = m.stereo_frames_host(a, b, hl.array, hr.array, mb, mbf)
ref2 = m.stereo_frames_host(a, b, hl.array, hr.array, mb, mbf)
for k in ref:
    assert ref[k].tobytes() == ref2[k].tobytes() or k in ("kpsL", "kpsR", "descL", "descR", "uRight", "depth"), k
lock = threading.Lock()
bad = []


def cmp(o, tag):
    for key in ("countsL", "countsR"):
        if not np.array_equal(o[key], ref[key]):
            d = np.nonzero(o[key] != ref[key])[0]
            with lock:
                bad.append((tag, key, d[:8].tolist(), o[key][d[:4]].tolist(), ref[key][d[:4]].tolist()))
            return
    for i in range(B):
        c = int(ref["countsL"][i]); cr = int(ref["countsR"][i])
        for key, n in (("kpsL", c), ("descL", c), ("kpsR", cr), ("descR", cr), ("uRight", c), ("depth", c)):
            if o[key][i, :n].tobytes() != ref[key][i, :n].tobytes():
                with lock:
                    bad.append((tag, key, i))
                return


import ctypes as C
PINNED = os.environ.get("DIAG_PINNED", "1") == "1"
cap = a.max_keypoints(W, H)


class PinnedOut:
    def __init__(self):
        hb = orbx.HostBuffer
        self.b = dict(kpsL=hb((B, cap, 28)), kpsR=hb((B, cap, 28)), descL=hb((B, cap, 32)), descR=hb((B, cap, 32)),
                      countsL=hb((B,), np.int32), countsR=hb((B,), np.int32), uRight=hb((B, cap), np.float32), depth=hb((B, cap), np.float32))

    def call(self, z, x, y):
        b = self.b
        rc = orbx.lib().orbx_stereo_frames_host(z._m, x._h, y._h, hl.ptr, hr.ptr, B, W, H, W, C.c_size_t(W * H), C.c_float(mb), C.c_float(mbf),
                                                b["kpsL"].ptr, b["descL"].ptr, b["countsL"].ptr, b["kpsR"].ptr, b["descR"].ptr, b["countsR"].ptr,
                                                b["uRight"].ptr, b["depth"].ptr, cap)
        assert rc == 0, (rc, orbx.lib().orbx_last_error())
        o = {k: v.array for k, v in b.items()}
        o["kpsL"] = o["kpsL"].view(orbx.KP_DTYPE).reshape(B, cap); o["kpsR"] = o["kpsR"].view(orbx.KP_DTYPE).reshape(B, cap)
        return o


def worker(k):
    x, y, z = orbx.ORBextractor(1200, max_batch=nb), orbx.ORBextractor(1200, max_batch=nb), orbx.ORBmatcher()
    po = PinnedOut() if PINNED else None
    for c in range(CALLS):
        o = po.call(z, x, y) if PINNED else z.stereo_frames_host(x, y, hl.array, hr.array, mb, mbf)
        cmp(o, (k, c))


ths = [threading.Thread(target=worker, args=(k,)) for k in range(NW)]
[t.start() for t in ths]
[t.join() for t in ths]
print("B", B, "workers", NW, "chunks", CH, "calls", CALLS, "mismatches", len(bad))
for x in bad[:12]:
    print("  ", x)
