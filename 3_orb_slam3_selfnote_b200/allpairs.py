"""All-pairs keyframe-descriptor 2-NN sweep (BASELINE.json config 5) -- the one step of the path with a real exchange.

Every rank owns a contiguous shard of keyframes (`kf_per_rank` keyframes x `desc_per_kf` descriptors of 32 bytes, global row
index = rank * rows_per_rank + local row).  For each of its own descriptors a rank wants the two nearest descriptors among
ALL descriptors of all OTHER keyframes.

The product path is the C-ABI call `orbx_allpairs_knn2` (csrc/orbx_allpairs.cu): raw ncclAllGather of the shards in `nchunks`
pieces on a side stream while the 2-NN tile kernel folds the previous piece; CUDA tensors take that path and nothing else (no
CPU fallback).  The same chunk schedule is written out below in torch.distributed terms with an injectable `fold`, so that
the index bookkeeping (global rows, own-keyframe exclusion, ties across chunks and ranks) is testable with the gloo backend on
CPU (tests/test_allpairs.py); that form never runs on a GPU.
"""
import numpy as np

KEY_NONE = np.uint64(0xFFFFFFFFFFFFFFFF)


def allpairs_knn2(desc, desc_per_kf, matcher=None, group=None, nchunks=4, fold=None, comm=None, return_stats=False):
    """desc: torch.uint8 [rows_per_rank, 32] (this rank's shard; same shape on every rank).
    Returns (idx[rows,2] int64 global row indices, dist[rows,2] int32), -1 where fewer than two neighbours exist.
    CUDA tensors: orbx_allpairs_knn2 through `matcher`; `comm` = a matcher.NcclComm to reuse (else one is made from `group`)."""
    import torch
    import torch.distributed as dist
    if desc.is_cuda:
        if matcher is None:
            raise RuntimeError("allpairs: CUDA tensors need an ORBmatcher (the sweep runs in liborbx.so)")
        assert desc.dtype == torch.uint8 and desc.dim() == 2 and desc.shape[1] == 32 and desc.is_contiguous()
        world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        own = None
        if world > 1 and comm is None:
            from .matcher import nccl_comm_from_torch
            comm = own = nccl_comm_from_torch(desc.device.index, group)
        torch.cuda.synchronize(desc.device)      # the library runs on its own streams
        try:
            idx, dst, stats = matcher.allpairs_knn2(desc.data_ptr(), desc.shape[0], desc_per_kf, nchunks, comm if world > 1 else None)
        finally:
            if own is not None:
                own.close()
        return (idx, dst, stats) if return_stats else (idx, dst)
    if fold is None:
        raise RuntimeError("allpairs: host tensors are accepted only with an injected fold (tests); there is no CPU fallback")
    return _allpairs_impl(desc, desc_per_kf, matcher, group, nchunks, fold)


def _allpairs_impl(desc, desc_per_kf, matcher, group, nchunks, fold):
    import torch
    import torch.distributed as dist
    rows = desc.shape[0]
    assert rows % desc_per_kf == 0 and desc.dtype == torch.uint8 and desc.shape[1] == 32
    world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
    rank = dist.get_rank(group) if world > 1 else 0
    dev = desc.device
    # own-keyframe exclusion range per query, in global rows
    kf = torch.arange(rows, device=dev, dtype=torch.int64) // desc_per_kf
    start = rank * rows + kf * desc_per_kf
    exclude = torch.stack([start, start + desc_per_kf], 1).to(torch.int32).contiguous()
    key0 = torch.empty(rows, dtype=torch.int64, device=dev)
    key1 = torch.empty(rows, dtype=torch.int64, device=dev)
    nchunks = max(1, min(nchunks, rows))
    bounds = [(c * rows) // nchunks for c in range(nchunks + 1)]
    init = 1
    if world == 1:
        for c in range(nchunks):
            a, b = bounds[c], bounds[c + 1]
            fold(desc, desc[a:b], a, exclude, key0, key1, init)
            init = 0
    else:
        maxr = max(bounds[c + 1] - bounds[c] for c in range(nchunks))
        bufs = [torch.empty((world, maxr, 32), dtype=torch.uint8, device=dev) for _ in range(2)]
        send = [torch.zeros((maxr, 32), dtype=torch.uint8, device=dev) for _ in range(2)]

        def launch(c):
            a, b = bounds[c], bounds[c + 1]
            send[c & 1][: b - a].copy_(desc[a:b])
            return dist.all_gather_into_tensor(bufs[c & 1].view(world * maxr, 32), send[c & 1], group=group, async_op=True)

        work = launch(0)
        for c in range(nchunks):
            work.wait()                       # the current stream waits for chunk c
            nxt = launch(c + 1) if c + 1 < nchunks else None   # chunk c+1 travels while chunk c is folded
            a, b = bounds[c], bounds[c + 1]
            for r in range(world):
                fold(desc, bufs[c & 1][r, : b - a], r * rows + a, exclude, key0, key1, init)
                init = 0
            # buffers of chunk c are reused by chunk c+2, which is launched from this stream after these folds were enqueued:
            # torch's NCCL collectives wait for the launching stream, so the gather cannot overwrite them early
            work = nxt
    k0 = key0.cpu().numpy().view(np.uint64)
    k1 = key1.cpu().numpy().view(np.uint64)
    idx = np.stack([np.where(k0 == KEY_NONE, -1, (k0 & np.uint64(0xFFFFFFFF)).astype(np.int64)),
                    np.where(k1 == KEY_NONE, -1, (k1 & np.uint64(0xFFFFFFFF)).astype(np.int64))], 1)
    dst = np.stack([np.where(k0 == KEY_NONE, -1, (k0 >> np.uint64(32)).astype(np.int64)),
                    np.where(k1 == KEY_NONE, -1, (k1 >> np.uint64(32)).astype(np.int64))], 1).astype(np.int32)
    return idx, dst


def numpy_fold(q, train, train_base, exclude, key0, key1, init):
    """TESTS ONLY (gloo / CPU): the same fold as k_knn2 on numpy, so the chunk schedule can be checked without a GPU."""
    qn, tn = q.numpy(), train.numpy()
    ex = exclude.numpy()
    k0 = key0.numpy().view(np.uint64)
    k1 = key1.numpy().view(np.uint64)
    if init:
        k0[:] = KEY_NONE
        k1[:] = KEY_NONE
    if tn.shape[0] == 0:
        return
    d = np.unpackbits(qn[:, None, :] ^ tn[None, :, :], axis=2).sum(axis=2).astype(np.uint64)
    g = (np.arange(tn.shape[0], dtype=np.int64) + int(train_base))
    keys = (d << np.uint64(32)) | g.astype(np.uint64)[None, :]
    masked = (g[None, :] >= ex[:, 0:1]) & (g[None, :] < ex[:, 1:2])
    keys = np.where(masked, KEY_NONE, keys)
    allk = np.concatenate([k0[:, None], k1[:, None], keys], 1)
    allk.sort(axis=1)
    k0[:] = allk[:, 0]
    k1[:] = allk[:, 1]
