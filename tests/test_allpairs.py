"""All-pairs keyframe 2-NN sweep (config 5): chunk schedule + index bookkeeping on CPU with gloo (world_size 2, injected numpy
fold), and the real k_knn2 path on the GPU (single rank; two NCCL ranks when the box has two GPUs)."""
import importlib
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = "3_orb_slam3_selfnote_b200"


def brute_force(desc_all, desc_per_kf):
    """top-2 (dist, global row) over all rows of other keyframes; ties -> lower global row"""
    n = desc_all.shape[0]
    d = np.unpackbits(desc_all[:, None, :] ^ desc_all[None, :, :], axis=2).sum(axis=2).astype(np.int64)
    kf = np.arange(n) // desc_per_kf
    key = (d << 32) | np.arange(n)[None, :]
    key[kf[:, None] == kf[None, :]] = np.iinfo(np.int64).max
    key.sort(axis=1)
    k = key[:, :2]
    none = k == np.iinfo(np.int64).max
    return np.where(none, -1, k & 0xFFFFFFFF), np.where(none, -1, k >> 32).astype(np.int32)


def make_desc(n, seed):
    rng = np.random.default_rng(seed)
    d = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    d[:, 6:] = d[0, 6:]          # few distinct distances => many ties across chunks and ranks
    d[n // 3] = d[n // 2]        # exact duplicates in different keyframes
    return d


WORKER = r"""
import importlib, os, sys, numpy as np, torch, torch.distributed as dist
sys.path.insert(0, {root!r})
ap = importlib.import_module("{pkg}.allpairs")
backend = sys.argv[1]
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
if backend == "nccl":
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
else:
    dist.init_process_group("gloo")
allrows = np.load(sys.argv[2])
per = allrows.shape[0] // world
mine = torch.from_numpy(allrows[rank * per:(rank + 1) * per].copy())
if backend == "nccl":
    orbx = importlib.import_module("{pkg}")
    m = orbx.ORBmatcher(device=rank)
    idx, dst = ap.allpairs_knn2(mine.cuda(), int(sys.argv[3]), matcher=m, nchunks=3)
else:
    idx, dst = ap.allpairs_knn2(mine, int(sys.argv[3]), nchunks=3, fold=ap.numpy_fold)
np.savez(sys.argv[4] + ".%d.npz" % rank, idx=idx, dst=dst)
dist.barrier()
dist.destroy_process_group()
"""


def _run_ranks(tmp_path, backend, world, desc_all, desc_per_kf):
    script = tmp_path / "worker.py"
    script.write_text(WORKER.format(root=ROOT, pkg=PKG))
    np.save(tmp_path / "rows.npy", desc_all)
    port = 29500 + (os.getpid() % 2000)
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(world), LOCAL_RANK=str(r), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, str(script), backend, str(tmp_path / "rows.npy"), str(desc_per_kf), str(tmp_path / "out")],
                                      env=env))
    for p in procs:
        assert p.wait(timeout=300) == 0
    idx = np.concatenate([np.load(str(tmp_path / "out") + ".%d.npz" % r)["idx"] for r in range(world)])
    dst = np.concatenate([np.load(str(tmp_path / "out") + ".%d.npz" % r)["dst"] for r in range(world)])
    return idx, dst


def test_allpairs_gloo_two_ranks_cpu(tmp_path):
    desc_per_kf, kf_per_rank, world = 25, 6, 2
    desc_all = make_desc(world * kf_per_rank * desc_per_kf, 1)
    idx, dst = _run_ranks(tmp_path, "gloo", world, desc_all, desc_per_kf)
    ridx, rdst = brute_force(desc_all, desc_per_kf)
    assert np.array_equal(idx, ridx) and np.array_equal(dst, rdst)


def test_allpairs_single_rank_schedule_cpu():
    import torch
    ap = importlib.import_module(PKG + ".allpairs")
    desc = make_desc(7 * 20, 2)
    idx, dst = ap.allpairs_knn2(torch.from_numpy(desc), 20, nchunks=5, fold=ap.numpy_fold)
    ridx, rdst = brute_force(desc, 20)
    assert np.array_equal(idx, ridx) and np.array_equal(dst, rdst)
    # a single keyframe has no neighbours at all
    idx, dst = ap.allpairs_knn2(torch.from_numpy(desc[:20]), 20, nchunks=2, fold=ap.numpy_fold)
    assert (idx == -1).all() and (dst == -1).all()
    with pytest.raises(RuntimeError):
        ap.allpairs_knn2(torch.from_numpy(desc), 20, matcher=None)   # default fold: CUDA only, no CPU fallback


@pytest.mark.gpu
def test_allpairs_single_gpu(orbx):
    import torch
    ap = importlib.import_module(PKG + ".allpairs")
    desc_per_kf = 100
    desc = make_desc(30 * desc_per_kf, 3)
    m = orbx.ORBmatcher()
    idx, dst = ap.allpairs_knn2(torch.from_numpy(desc).cuda(), desc_per_kf, matcher=m, nchunks=7)
    ridx, rdst = brute_force(desc, desc_per_kf)
    assert np.array_equal(idx, ridx) and np.array_equal(dst, rdst)
    m.close()


@pytest.mark.gpu
def test_allpairs_nccl_two_gpus(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (run with gpurun --gpus 2)")
    desc_per_kf, kf_per_rank, world = 100, 12, 2
    desc_all = make_desc(world * kf_per_rank * desc_per_kf, 4)
    idx, dst = _run_ranks(tmp_path, "nccl", world, desc_all, desc_per_kf)
    ridx, rdst = brute_force(desc_all, desc_per_kf)
    assert np.array_equal(idx, ridx) and np.array_equal(dst, rdst)


def _fold(m, q, chunks, ex, base, kind):
    """fold the train chunks one call each (init on the first) with the given kernel -> (key0, key1) as numpy uint64"""
    import torch
    m.set_knn2_kernel(kind)
    nq = q.shape[0]
    k0 = torch.empty(nq, dtype=torch.int64, device="cuda"); k1 = torch.empty_like(k0)
    off = 0
    for i, t in enumerate(chunks):
        m.knn2_chunk_device(q.data_ptr(), nq, t.data_ptr(), t.shape[0], base + off, ex.data_ptr() if ex is not None else 0, k0.data_ptr(),
                            k1.data_ptr(), init=(i == 0), stream=torch.cuda.current_stream().cuda_stream)
        off += t.shape[0]
    torch.cuda.synchronize()
    m.set_knn2_kernel(m.KNN2_AUTO)
    return k0.cpu().numpy().view(np.uint64), k1.cpu().numpy().view(np.uint64)


@pytest.mark.gpu
@pytest.mark.parametrize("nq,sizes,ties", [(1, (1,), True), (77, (5, 1, 300), True), (256, (128, 128), False), (1000, (129, 2047, 640), True),
                                            (5000, (20000, 12345), False)])
def test_knn2_tc_matches_popc_kernel(orbx, nq, sizes, ties):
    """k_knn2_tc (tcgen05 kind::i8) writes the keys k_knn2 writes: ragged tiles, chunks smaller than a tile, ties across chunks,
    exact duplicates, own-keyframe exclusion ranges"""
    import torch
    rng = np.random.default_rng(nq)
    nt = sum(sizes)
    q = rng.integers(0, 256, (nq, 32), dtype=np.uint8); t = rng.integers(0, 256, (nt, 32), dtype=np.uint8)
    if ties:
        t[:, 5:] = t[0, 5:]; q[:, 5:] = t[0, 5:]
    t[nt - 1] = q[0]; t[nt // 2] = q[nq // 2]
    base = 7_000_000
    ex = np.stack([base + (np.arange(nq) // 50) * 50 % max(nt, 1), base + (np.arange(nq) // 50) * 50 % max(nt, 1) + 50], 1).astype(np.int32)
    dq = torch.from_numpy(q).cuda(); dex = torch.from_numpy(ex).cuda()
    chunks = [torch.from_numpy(c.copy()).cuda() for c in np.split(t, np.cumsum(sizes)[:-1])]
    m = orbx.ORBmatcher()
    for e in (dex, None):
        p0, p1 = _fold(m, dq, chunks, e, base, m.KNN2_POPC)
        g0, g1 = _fold(m, dq, chunks, e, base, m.KNN2_TENSOR)
        assert np.array_equal(p0, g0) and np.array_equal(p1, g1)
    if nq * nt <= 4_000_000:   # and both equal the brute force
        d = np.unpackbits(q[:, None, :] ^ t[None, :, :], axis=2).sum(axis=2).astype(np.uint64)
        key = (d << np.uint64(32)) | (np.uint64(base) + np.arange(nt, dtype=np.uint64))[None, :]
        key.sort(axis=1)
        assert np.array_equal(g0, key[:, 0])
        if nt > 1:
            assert np.array_equal(g1, key[:, 1])
    m.close()


@pytest.mark.gpu
def test_allpairs_single_gpu_tensor_path(orbx):
    """the sweep at a size where the tensor-core fold is picked by itself (>= 2^26 pairs per chunk), against sampled brute force"""
    import torch
    ap = importlib.import_module(PKG + ".allpairs")
    desc_per_kf, nkf = 500, 80
    rng = np.random.default_rng(9)
    desc = rng.integers(0, 256, (nkf * desc_per_kf, 32), dtype=np.uint8)
    desc[:, 8:] = desc[0, 8:]
    desc[123] = desc[30123]
    m = orbx.ORBmatcher()
    idx, dst = ap.allpairs_knn2(torch.from_numpy(desc).cuda(), desc_per_kf, matcher=m, nchunks=2)
    m.set_knn2_kernel(m.KNN2_POPC)
    idx_p, dst_p = ap.allpairs_knn2(torch.from_numpy(desc).cuda(), desc_per_kf, matcher=m, nchunks=2)
    assert np.array_equal(idx, idx_p) and np.array_equal(dst, dst_p)
    rows = rng.choice(len(desc), 64, replace=False)
    d = np.unpackbits(desc[rows][:, None, :] ^ desc[None, :, :], axis=2).sum(axis=2).astype(np.int64)
    key = (d << 32) | np.arange(len(desc))[None, :]
    key[(rows // desc_per_kf)[:, None] == (np.arange(len(desc)) // desc_per_kf)[None, :]] = np.iinfo(np.int64).max
    key.sort(axis=1)
    assert np.array_equal(idx[rows], key[:, :2] & 0xFFFFFFFF) and np.array_equal(dst[rows], (key[:, :2] >> 32).astype(np.int32))
    m.close()
