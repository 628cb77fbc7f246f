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
