"""ctypes mirror of the reference's ORBmatcher / Frame Hamming searches over the C-ABI (include/orbx.h, matcher section).

    ORBmatcher(nnratio=0.6, checkOri=True)                    include/ORBmatcher.h:44
    ORBmatcher.DescriptorDistance(a, b)                       src/ORBmatcher.cc:2463
    ORBmatcher.SearchByProjection(F, mappoints, th)           src/ORBmatcher.cc:44     (map-point arrays instead of MapPoint*)
    ORBmatcher.SearchByProjectionWindow(...)                  src/ORBmatcher.cc:2027 / :2291 (after the caller's projection)
    ORBmatcher.SearchForInitialization(F1, F2, prev, win)     src/ORBmatcher.cc:722
    FrameView.GetFeaturesInArea(x, y, r, minLevel, maxLevel)  src/Frame.cc:744
    ORBmatcher.SearchByBoW(...)                               src/ORBmatcher.cc:273 / :839 (FeatureVectors as CSR triples)
    ORBmatcher.SearchForTriangulation(...)                    src/ORBmatcher.cc:981    (pinhole; F12 / epipole from the caller)
    ORBmatcher.SearchWindowBest(...)                          src/ORBmatcher.cc:1524 / :1728 / :1866 (Fuse x2, SearchBySim3 loops)
    ORBmatcher.SearchBySim3Projected(...)                     src/ORBmatcher.cc:1788   (both passes + the agreement check)
    knn2 / stereo_match_batch                                 src/Frame.cc:1246 / :901

A FrameView holds the members of Frame the searches read.  Everything runs in the CUDA kernels of csrc/match_kernels.cuh.
"""
import ctypes as C

import numpy as np

from . import KP_DTYPE, _check, _p, lib

TH_HIGH, TH_LOW, HISTO_LENGTH = 100, 50, 30  # src/ORBmatcher.cc:36-38


class AllpairsStats(C.Structure):   # orbx_allpairs_stats
    _fields_ = [("world", C.c_int), ("rank", C.c_int), ("nchunks", C.c_int), ("pairs", C.c_double), ("total_ms", C.c_double),
                ("fold_ms", C.c_double), ("gather_ms", C.c_double), ("gather_bytes", C.c_double), ("overlap_frac", C.c_double)]


class NcclComm:
    """An ncclComm_t created through the library's run-time NCCL binding (orbx_nccl_unique_id / orbx_nccl_comm_init).
    `exchange(id_bytes_or_None) -> id_bytes` ships rank 0's 128-byte id to every rank (any channel: torch.distributed, MPI, a file)."""

    def __init__(self, rank, world, device, exchange):
        uid = (C.c_uint8 * 128)()
        if rank == 0:
            _check(lib().orbx_nccl_unique_id(uid))
        got = exchange(bytes(uid) if rank == 0 else None)
        uid = (C.c_uint8 * 128).from_buffer_copy(got)
        self._c = C.c_void_p()
        _check(lib().orbx_nccl_comm_init(uid, world, rank, device, C.byref(self._c)))
        self.rank, self.world = rank, world

    def close(self):
        if self._c:
            lib().orbx_nccl_comm_destroy(self._c)
            self._c = C.c_void_p()


def nccl_comm_from_torch(device, group=None):
    """NcclComm over the ranks of an initialised torch.distributed group (torch only carries the 128-byte id)."""
    import torch.distributed as dist
    rank, world = dist.get_rank(group), dist.get_world_size(group)

    def exchange(b):
        box = [b]
        dist.broadcast_object_list(box, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        return box[0]
    return NcclComm(rank, world, device, exchange)


class _FrameViewC(C.Structure):
    _fields_ = [("n", C.c_int), ("keypoints", C.c_void_p), ("descriptors", C.c_void_p), ("u_right", C.c_void_p),
                ("scale_factors", C.c_void_p), ("nlevels", C.c_int), ("min_x", C.c_float), ("min_y", C.c_float),
                ("max_x", C.c_float), ("max_y", C.c_float)]


class _FeatVecC(C.Structure):
    _fields_ = [("nnodes", C.c_int), ("node_ids", C.c_void_p), ("node_start", C.c_void_p), ("indices", C.c_void_p)]


class _TrackC(C.Structure):
    _fields_ = [("in_view", C.c_void_p), ("proj_x", C.c_void_p), ("proj_y", C.c_void_p), ("level", C.c_void_p), ("view_cos", C.c_void_p)]


class FrameView:
    """N, mvKeysUn, mDescriptors, mvuRight, mvScaleFactors, mnMinX.. of a Frame (include/Frame.h)."""

    def __init__(self, keypoints, descriptors, scale_factors, width, height, u_right=None, bounds=None):
        self.kps = np.ascontiguousarray(keypoints, KP_DTYPE)
        self.desc = np.ascontiguousarray(descriptors, np.uint8).reshape(-1, 32)
        self.sf = np.ascontiguousarray(scale_factors, np.float32)
        self.u_right = None if u_right is None else np.ascontiguousarray(u_right, np.float32)
        b = bounds if bounds is not None else (0.0, 0.0, float(width), float(height))
        self.c = _FrameViewC(len(self.kps), self.kps.ctypes.data, self.desc.ctypes.data,
                             None if self.u_right is None else self.u_right.ctypes.data, self.sf.ctypes.data, len(self.sf),
                             b[0], b[1], b[2], b[3])

    @property
    def N(self):
        return len(self.kps)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, np.float32)


def _u8(a):
    return None if a is None else np.ascontiguousarray(a, np.uint8)


def _i32(a):
    return None if a is None else np.ascontiguousarray(a, np.int32)


class ORBmatcher:
    def __init__(self, nnratio=0.6, checkOri=True, device=0):
        self.mfNNratio = float(nnratio)
        self.mbCheckOrientation = bool(checkOri)
        self._m = C.c_void_p()
        _check(lib().orbx_matcher_create(device, C.byref(self._m)))

    def close(self):
        if getattr(self, "_m", None) and self._m.value:
            lib().orbx_matcher_destroy(self._m)
            self._m = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @staticmethod
    def DescriptorDistance(a, b):
        return lib().orbx_hamming(_p(_u8(a)), _p(_u8(b)))

    def hamming_batch(self, a, b):
        a, b = _u8(a).reshape(-1, 32), _u8(b).reshape(-1, 32)
        out = np.empty(len(a), np.int32)
        _check(lib().orbx_hamming_batch(self._m, _p(a), _p(b), len(a), _p(out)))
        return out

    def GetFeaturesInArea(self, F, x, y, r, minLevel=-1, maxLevel=-1):
        out = np.empty(max(F.N, 1), np.int32)
        n = C.c_int(0)
        _check(lib().orbx_features_in_area(self._m, C.byref(F.c), C.c_float(x), C.c_float(y), C.c_float(r), minLevel, maxLevel,
                                           _p(out), len(out), C.byref(n)))
        return out[:n.value].copy()

    def SearchByProjection(self, F, in_view, projX, projY, level, viewCos, mpDesc, th=1.0, kp_mp=None, observed=None, projXR=None):
        """local-map search; returns (nmatches, kp_mp) with kp_mp[k] = index of the map point matched to keypoint k (or -1)"""
        in_view, projX, projY, level, viewCos = _u8(in_view), _f32(projX), _f32(projY), _i32(level), _f32(viewCos)
        mpDesc, observed, projXR = _u8(mpDesc), _u8(observed), _f32(projXR)
        kp_mp = np.full(F.N, -1, np.int32) if kp_mp is None else _i32(kp_mp).copy()
        nm = C.c_int(0)
        _check(lib().orbx_search_by_projection_mappoints(self._m, C.byref(F.c), len(in_view), _p(in_view), _p(projX), _p(projY),
                                                         _p(level), _p(viewCos), _p(mpDesc), _p(observed), _p(projXR),
                                                         C.c_float(th), C.c_float(self.mfNNratio), _p(kp_mp), C.byref(nm)))
        return nm.value, kp_mp

    def SearchByProjectionBatch(self, extractor, nframes, in_view, projX, projY, level, viewCos, mpDesc, th=1.0, observed=None):
        """map-point arrays are [nframes, nmp]; keypoints are the device-resident results of `extractor`'s last batch"""
        in_view, projX, projY, level, viewCos = _u8(in_view), _f32(projX), _f32(projY), _i32(level), _f32(viewCos)
        mpDesc, observed = _u8(mpDesc), _u8(observed)
        nmp = in_view.shape[1]
        h, w = extractor._last_shape
        cap = extractor.max_keypoints(w, h)
        kp_mp = np.full((nframes, cap), -1, np.int32)
        nm = np.zeros(nframes, np.int32)
        _check(lib().orbx_search_by_projection_mappoints_batch(self._m, extractor._h, nframes, nmp, _p(in_view), _p(projX), _p(projY),
                                                               _p(level), _p(viewCos), _p(mpDesc), _p(observed), C.c_float(th),
                                                               C.c_float(self.mfNNratio), _p(kp_mp), cap, _p(nm)))
        return nm, kp_mp

    def SearchByProjectionWindow(self, F, valid, u, v, level, angle, srcDesc, th, mode=0, dist_th=TH_HIGH, any_mp_blocks=False,
                                 kp_mp=None, observed=None, ur=None):
        valid, u, v, level, angle = _u8(valid), _f32(u), _f32(v), _i32(level), _f32(angle)
        srcDesc, observed, ur = _u8(srcDesc), _u8(observed), _f32(ur)
        kp_mp = np.full(F.N, -1, np.int32) if kp_mp is None else _i32(kp_mp).copy()
        nm = C.c_int(0)
        _check(lib().orbx_search_by_projection_window(self._m, C.byref(F.c), len(valid), _p(valid), _p(u), _p(v), _p(level), _p(ur),
                                                      _p(angle), _p(srcDesc), _p(observed), C.c_float(th), int(mode), int(dist_th),
                                                      int(any_mp_blocks), int(self.mbCheckOrientation), _p(kp_mp), C.byref(nm)))
        return nm.value, kp_mp

    def SearchWindowBest(self, F, valid, u, v, radius, level, desc, gate=False, inv_level_sigma2=None, ur=None):
        """independent window argmin of Fuse / SearchBySim3: -> (best_idx[n], best_dist[n])"""
        valid, u, v, radius, level, desc = _u8(valid), _f32(u), _f32(v), _f32(radius), _i32(level), _u8(desc)
        ils, ur = _f32(inv_level_sigma2), _f32(ur)
        n = len(valid)
        bi = np.full(max(n, 1), -1, np.int32); bd = np.full(max(n, 1), 256, np.int32)
        _check(lib().orbx_search_window_best(self._m, C.byref(F.c), n, _p(valid), _p(u), _p(v), _p(radius), _p(level), _p(ur),
                                             _p(desc), int(gate), _p(ils), _p(bi), _p(bd)))
        return bi[:n], bd[:n]

    def SearchBySim3Projected(self, F1, F2, valid1, u1, v1, radius1, level1, desc1, valid2, u2, v2, radius2, level2, desc2):
        """SearchBySim3 after the caller's Sim3 projections: pass 1 = map points of KF1 (one per keypoint of F1) projected
        into KF2 (u1.. in F2's image), pass 2 the other way; returns match12[i1] = i2 where both passes agree (-1 otherwise),
        src/ORBmatcher.cc:1896-1900, 1976-1980, 1983-1999 (bestDist <= TH_HIGH)."""
        b1, d1 = self.SearchWindowBest(F2, valid1, u1, v1, radius1, level1, desc1)
        b2, d2 = self.SearchWindowBest(F1, valid2, u2, v2, radius2, level2, desc2)
        m1 = np.where(d1 <= TH_HIGH, b1, -1)
        m2 = np.where(d2 <= TH_HIGH, b2, -1)
        out = np.full(len(m1), -1, np.int32)
        ok = m1 >= 0
        i1 = np.nonzero(ok)[0]
        agree = m2[m1[i1]] == i1
        out[i1[agree]] = m1[i1[agree]]
        return int(agree.sum()), out

    @staticmethod
    def _fv(fv):
        ids = np.ascontiguousarray(fv[0], np.uint32); start = np.ascontiguousarray(fv[1], np.int32)
        ind = np.ascontiguousarray(fv[2], np.uint32)
        return _FeatVecC(len(ids), ids.ctypes.data, start.ctypes.data, ind.ctypes.data), (ids, start, ind)

    def SearchByBoW(self, kps1, desc1, fv1, elig1, kps2, desc2, fv2, elig2=None, max_dist=TH_LOW):
        """fv = (node_ids ascending, node_start[nnodes+1], indices).  KF-Frame: max_dist=TH_LOW, elig2=None; KF-KF: TH_LOW-1.
        -> (nmatches, matches12)"""
        kps1 = np.ascontiguousarray(kps1, KP_DTYPE); kps2 = np.ascontiguousarray(kps2, KP_DTYPE)
        desc1, desc2, elig1, elig2 = _u8(desc1), _u8(desc2), _u8(elig1), _u8(elig2)
        c1, keep1 = self._fv(fv1); c2, keep2 = self._fv(fv2)
        m12 = np.full(max(len(kps1), 1), -1, np.int32)
        nm = C.c_int(0)
        _check(lib().orbx_search_by_bow(self._m, len(kps1), _p(kps1), _p(desc1), C.byref(c1), _p(elig1), len(kps2), _p(kps2),
                                        _p(desc2), C.byref(c2), _p(elig2), int(max_dist), C.c_float(self.mfNNratio),
                                        int(self.mbCheckOrientation), _p(m12), C.byref(nm)))
        return nm.value, m12[:len(kps1)]

    def SearchByBoWTwoCameras(self, kps1, desc1, fv1, elig1, kps2, desc2, fv2, n_left2, max_dist=TH_LOW):
        """orbx_search_by_bow_stereo: SearchByBoW(KeyFrame, Frame) with a fisheye-stereo frame on side 2 (ORBmatcher.cc:338-431)
        -> (nmatches, matches12_left, matches12_right)"""
        kps1 = np.ascontiguousarray(kps1, KP_DTYPE); kps2 = np.ascontiguousarray(kps2, KP_DTYPE)
        desc1, desc2, elig1 = _u8(desc1), _u8(desc2), _u8(elig1)
        c1, keep1 = self._fv(fv1); c2, keep2 = self._fv(fv2)
        mL = np.full(max(len(kps1), 1), -1, np.int32); mR = np.full(max(len(kps1), 1), -1, np.int32)
        nm = C.c_int(0)
        _check(lib().orbx_search_by_bow_stereo(self._m, len(kps1), _p(kps1), _p(desc1), C.byref(c1), _p(elig1), len(kps2), int(n_left2), _p(kps2),
                                               _p(desc2), C.byref(c2), int(max_dist), C.c_float(self.mfNNratio), int(self.mbCheckOrientation),
                                               _p(mL), _p(mR), C.byref(nm)))
        return nm.value, mL[:len(kps1)], mR[:len(kps1)]

    def SearchForTriangulation(self, kps1, desc1, fv1, elig1, stereo1, kps2, desc2, fv2, elig2, stereo2, F12, ep, scale_factors2,
                               level_sigma2_2, coarse=False):
        kps1 = np.ascontiguousarray(kps1, KP_DTYPE); kps2 = np.ascontiguousarray(kps2, KP_DTYPE)
        desc1, desc2, elig1, elig2, stereo1, stereo2 = _u8(desc1), _u8(desc2), _u8(elig1), _u8(elig2), _u8(stereo1), _u8(stereo2)
        c1, keep1 = self._fv(fv1); c2, keep2 = self._fv(fv2)
        F12 = _f32(F12).reshape(9); sf = _f32(scale_factors2); ls = _f32(level_sigma2_2)
        m12 = np.full(max(len(kps1), 1), -1, np.int32)
        nm = C.c_int(0)
        _check(lib().orbx_search_for_triangulation(self._m, len(kps1), _p(kps1), _p(desc1), C.byref(c1), _p(elig1), _p(stereo1),
                                                   len(kps2), _p(kps2), _p(desc2), C.byref(c2), _p(elig2), _p(stereo2), _p(F12),
                                                   C.c_float(ep[0]), C.c_float(ep[1]), _p(sf), _p(ls), len(sf), int(coarse),
                                                   int(self.mbCheckOrientation), _p(m12), C.byref(nm)))
        return nm.value, m12[:len(kps1)]

    def SearchForTriangulationKB8(self, kps1, desc1, fv1, elig1, stereo1, n_left1, kps2, desc2, fv2, elig2, stereo2, n_left2, pairs, ep,
                                  scale_factors2, level_sigma2_1, level_sigma2_2, coarse=False):
        """SearchForTriangulation on KannalaBrandt8 keyframes (src/ORBmatcher.cc:981-1222 with the gate of KannalaBrandt8.cpp:244-247):
        pairs = one 28-float record R12[9] t12[3] cam1[8] cam2[8] (n_left = -1) or the four records ll, lr, rl, rr of a fisheye-stereo
        rig (kps / desc = left features followed by right ones)."""
        kps1 = np.ascontiguousarray(kps1, KP_DTYPE); kps2 = np.ascontiguousarray(kps2, KP_DTYPE)
        desc1, desc2, elig1, elig2 = _u8(desc1), _u8(desc2), _u8(elig1), _u8(elig2)
        stereo1 = None if stereo1 is None else _u8(stereo1); stereo2 = None if stereo2 is None else _u8(stereo2)
        c1, keep1 = self._fv(fv1); c2, keep2 = self._fv(fv2)
        G = np.ascontiguousarray(np.stack([_f32(p).reshape(28) for p in pairs]))
        sf = _f32(scale_factors2); l1 = _f32(level_sigma2_1); l2 = _f32(level_sigma2_2)
        m12 = np.full(max(len(kps1), 1), -1, np.int32)
        nm = C.c_int(0)
        _check(lib().orbx_search_for_triangulation_kb8(self._m, len(kps1), _p(kps1), _p(desc1), C.byref(c1), _p(elig1), _p(stereo1),
                                                       int(n_left1), len(kps2), _p(kps2), _p(desc2), C.byref(c2), _p(elig2), _p(stereo2),
                                                       int(n_left2), _p(G), len(G), C.c_float(ep[0]), C.c_float(ep[1]), _p(sf), _p(l1),
                                                       _p(l2), len(sf), int(coarse), int(self.mbCheckOrientation), _p(m12), C.byref(nm)))
        return nm.value, m12[:len(kps1)]

    def TriangulateMatchesKB8(self, pair, xy1, xy2, sigma1, sigma2):
        """KannalaBrandt8::TriangulateMatches (KannalaBrandt8.cpp:343-418) for n keypoint pairs -> (z1 or -1, accepted point) per pair"""
        G = _f32(pair).reshape(28)
        xy1 = _f32(xy1).reshape(-1, 2); xy2 = _f32(xy2).reshape(-1, 2)
        n = len(xy1)
        s1 = np.ascontiguousarray(np.broadcast_to(np.asarray(sigma1, np.float32), (n,)))
        s2 = np.ascontiguousarray(np.broadcast_to(np.asarray(sigma2, np.float32), (n,)))
        z1 = np.empty(max(n, 1), np.float32); p3d = np.empty((max(n, 1), 3), np.float32)
        _check(lib().orbx_kb8_triangulate_matches(self._m, _p(G), n, _p(xy1), _p(xy2), _p(s1), _p(s2), _p(z1), _p(p3d)))
        return z1[:n], p3d[:n]

    def DistinctiveDescriptors(self, desc, start):
        """MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:350) for a batch of map points: descriptors of point p
        are rows start[p]:start[p+1].  -> (best_idx[npoints] relative to start[p], best_median[npoints])"""
        desc = _u8(desc).reshape(-1, 32); start = _i32(start)
        npts = len(start) - 1
        bi = np.empty(max(npts, 1), np.int32); bm = np.empty(max(npts, 1), np.int32)
        _check(lib().orbx_distinctive_descriptors(self._m, _p(desc), _p(start), npts, _p(bi), _p(bm)))
        return bi[:npts], bm[:npts]

    def SearchForInitialization(self, F1, F2, vbPrevMatched, windowSize=10):
        pm = _f32(vbPrevMatched).reshape(-1, 2).copy()
        m12 = np.empty(F1.N, np.int32)
        nm = C.c_int(0)
        _check(lib().orbx_search_for_initialization(self._m, C.byref(F1.c), C.byref(F2.c), _p(pm), int(windowSize),
                                                    C.c_float(self.mfNNratio), int(self.mbCheckOrientation), _p(m12), C.byref(nm)))
        return nm.value, m12, pm

    def knn2(self, q, t):
        """BFMatcher(NORM_HAMMING).knnMatch(k=2) + the 0.7 ratio flag of Frame.cc:1253 -> (idx[nq,2], dist[nq,2], good[nq])"""
        q, t = _u8(q).reshape(-1, 32), _u8(t).reshape(-1, 32)
        idx = np.empty((len(q), 2), np.int32)
        dist = np.empty((len(q), 2), np.int32)
        good = np.empty(len(q), np.uint8)
        _check(lib().orbx_knn2(self._m, _p(q), len(q), _p(t), len(t), _p(idx), _p(dist), _p(good)))
        return idx, dist, good

    def SearchByProjectionTwoCameras(self, left, right, track_left, track_right, mp_desc, left_to_right, right_to_left, th, observed=None, kp_mp=None):
        """orbx_search_by_projection_mappoints_stereo: SearchByProjection(Frame, MapPoints) on a fisheye-stereo frame (ORBmatcher.cc:44-211).
        left / right: FrameView of mvKeys / mvKeysRight; track_*: dict(in_view, proj_x, proj_y, level, view_cos) -> (nmatches, kp_mp[nl+nr])"""
        keep = []

        def track(t):
            a = [_u8(t["in_view"]), _f32(t["proj_x"]), _f32(t["proj_y"]), _i32(t["level"]), _f32(t["view_cos"])]
            keep.append(a)
            return _TrackC(*[x.ctypes.data for x in a])
        tl, tr = track(track_left), track(track_right)
        nmp = len(keep[0][0])
        d = _u8(mp_desc).reshape(-1, 32)
        ob = _u8(observed)
        l2r, r2l = _i32(left_to_right), _i32(right_to_left)
        kp = np.full(left.N + right.N, -1, np.int32) if kp_mp is None else _i32(kp_mp).copy()
        nm = C.c_int(0)
        _check(lib().orbx_search_by_projection_mappoints_stereo(self._m, C.byref(left.c), C.byref(right.c), nmp, C.byref(tl), C.byref(tr), _p(d),
                                                                _p(ob), _p(l2r), _p(r2l), C.c_float(th), C.c_float(self.mfNNratio), _p(kp),
                                                                C.byref(nm)))
        return nm.value, kp

    KNN2_AUTO, KNN2_POPC, KNN2_TENSOR = 0, 1, 2

    def set_knn2_kernel(self, kind):
        """orbx_matcher_set_knn2_kernel: which kernel folds the 2-NN chunks (0 auto by size, 1 POPC, 2 tcgen05); same results"""
        _check(lib().orbx_matcher_set_knn2_kernel(self._m, int(kind)))

    def knn2_chunk_device(self, d_q, nq, d_t, nt, train_base, d_exclude, d_key0, d_key1, init, stream=0):
        _check(lib().orbx_knn2_chunk_device(self._m, C.c_void_p(d_q), nq, C.c_void_p(d_t), nt, C.c_longlong(train_base),
                                            C.c_void_p(d_exclude), C.c_void_p(d_key0), C.c_void_p(d_key1), int(init),
                                            C.c_void_p(stream)))

    def allpairs_knn2(self, d_desc, rows, desc_per_kf, nchunks=4, comm=None):
        """orbx_allpairs_knn2: the all-pairs keyframe sweep over this rank's `rows` device-resident descriptors (and, through the
        NcclComm `comm`, every other rank's).  -> (idx[rows,2] int64 global rows, dist[rows,2] int32, stats dict)"""
        idx = np.empty((rows, 2), np.int64)
        dist = np.empty((rows, 2), np.int32)
        st = AllpairsStats()
        _check(lib().orbx_allpairs_knn2(self._m, C.c_void_p(comm._c.value if comm is not None else None), C.c_void_p(d_desc),
                                        C.c_longlong(rows), int(desc_per_kf), int(nchunks), _p(idx), _p(dist), C.byref(st)))
        return idx, dist, {f: getattr(st, f) for f, _ in AllpairsStats._fields_}

    def stereo_match_batch(self, exL, exR, mb, mbf, stream=0):
        """Frame::ComputeStereoMatches for every pair of the two extractors' last batches (asynchronous)"""
        _check(lib().orbx_stereo_match_batch(self._m, exL._h, exR._h, C.c_float(mb), C.c_float(mbf), C.c_void_p(stream)))
        self._stereo_shape = exL._last_shape
        self._stereo_cap = exL.max_keypoints(exL._last_shape[1], exL._last_shape[0])

    def stereo_frames_host(self, exL, exR, left, right, mb, mbf, right_records=True):
        """whole stereo frames from host arrays (n,h,w): -> dict(countsL, kpsL, descL, countsR, kpsR, descR, uRight, depth);
        right_records=False leaves the right image's keypoints / descriptors on the device (kpsR = descR = None)"""
        left = np.ascontiguousarray(left, np.uint8); right = np.ascontiguousarray(right, np.uint8)
        n, h, w = left.shape
        cap = exL.max_keypoints(w, h)
        o = dict(countsL=np.empty(n, np.int32), kpsL=np.empty((n, cap), KP_DTYPE), descL=np.empty((n, cap, 32), np.uint8),
                 countsR=np.empty(n, np.int32), kpsR=np.empty((n, cap), KP_DTYPE) if right_records else None,
                 descR=np.empty((n, cap, 32), np.uint8) if right_records else None,
                 uRight=np.empty((n, cap), np.float32), depth=np.empty((n, cap), np.float32))
        pr = (lambda a: _p(a)) if right_records else (lambda a: None)
        _check(lib().orbx_stereo_frames_host(self._m, exL._h, exR._h, _p(left), _p(right), n, w, h, w, C.c_size_t(w * h), C.c_float(mb),
                                             C.c_float(mbf), _p(o["kpsL"]), _p(o["descL"]), _p(o["countsL"]), pr(o["kpsR"]),
                                             pr(o["descR"]), _p(o["countsR"]), _p(o["uRight"]), _p(o["depth"]), cap))
        exL._last_shape = exR._last_shape = (h, w)
        return o

    def stereo_results(self, nframes):
        cap = self._stereo_cap
        uR = np.empty((nframes, cap), np.float32); depth = np.empty((nframes, cap), np.float32)
        bd = np.empty((nframes, cap), np.int32); bi = np.empty((nframes, cap), np.int32)
        _check(lib().orbx_stereo_results_host(self._m, _p(uR), _p(depth), _p(bd), _p(bi), cap))
        return uR, depth, bd, bi


class ORBVocabulary:
    """The tree descent of DBoW2's TemplatedVocabulary<FORB> (the reference's ORBVocabulary, include/ORBVocabulary.h) on the GPU.

    vocab = dict(L, child_start, child_ids, node_desc, node_weight, node_word_id): the flattened m_nodes."""

    def __init__(self, vocab, device=0):
        self._keep = (_i32(vocab["child_start"]), _i32(vocab["child_ids"]), _u8(vocab["node_desc"]),
                      np.ascontiguousarray(vocab["node_weight"], np.float64), _i32(vocab["node_word_id"]))
        cs, ci, nd, nw, wi = self._keep
        self._v = C.c_void_p()
        _check(lib().orbx_vocabulary_create(device, len(cs) - 1, int(vocab["L"]), _p(cs), _p(ci), _p(nd), _p(nw), _p(wi),
                                            C.byref(self._v)))

    def close(self):
        if getattr(self, "_v", None) and self._v.value:
            lib().orbx_vocabulary_destroy(self._v)
            self._v = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def transform_features(self, matcher, desc, levelsup=4):
        """-> (word_id, node_id, weight) per feature, TemplatedVocabulary.h:1218-1259"""
        desc = _u8(desc).reshape(-1, 32)
        n = len(desc)
        w = np.empty(max(n, 1), np.int32); nid = np.empty(max(n, 1), np.int32); wt = np.empty(max(n, 1), np.float64)
        _check(lib().orbx_vocabulary_transform(matcher._m, self._v, _p(desc), n, int(levelsup), _p(w), _p(nid), _p(wt)))
        return w[:n], nid[:n], wt[:n]

    def transform(self, matcher, desc, levelsup=4):
        """Frame::ComputeBoW (src/Frame.cc:828-835): -> (bow_ids, bow_values, (fv_node_ids, fv_start, fv_indices)) for the
        TF_IDF / L1 settings of ORBvoc.  The fold is the reference's std::map insertion order restated on arrays: weights of
        one word are added in feature order, the L1 norm runs over ascending word ids (TemplatedVocabulary.h:1145-1195)."""
        w, nid, wt = self.transform_features(matcher, desc, levelsup)
        keep = wt > 0
        idx = np.nonzero(keep)[0]
        ids, inv = np.unique(w[idx], return_inverse=True)
        vals = np.zeros(len(ids), np.float64)
        np.add.at(vals, inv, wt[idx])                 # unbuffered: sequential in feature order, like addWeight
        norm = float(np.cumsum(np.abs(vals))[-1]) if len(vals) else 0.0   # sequential sum in ascending word order
        if norm > 0.0:
            vals = vals / norm
        nodes = np.unique(nid[idx])
        start = [0]; ind = []
        for nd in nodes:
            ind.extend(idx[nid[idx] == nd].tolist()); start.append(len(ind))
        return ids.astype(np.uint32), vals, (nodes.astype(np.uint32), np.array(start, np.int32), np.array(ind, np.uint32))


class LocalMap:
    """Device-resident local map for ORBmatcher.SearchLocalPointsBatch (Tracking::mvpLocalMapPoints as structure-of-arrays)."""

    def __init__(self, matcher, world_pos, normal, max_dist, min_dist, desc, observed=None):
        world_pos = np.ascontiguousarray(world_pos, np.float32).reshape(-1, 3)
        normal = np.ascontiguousarray(normal, np.float32).reshape(-1, 3)
        max_dist, min_dist, desc, observed = _f32(max_dist), _f32(min_dist), _u8(desc), _u8(observed)
        self.n = len(world_pos)
        self._l = C.c_void_p()
        _check(lib().orbx_local_map_create(matcher._m, self.n, _p(world_pos), _p(normal), _p(max_dist), _p(min_dist), _p(desc),
                                           _p(observed), C.byref(self._l)))

    def close(self):
        if getattr(self, "_l", None) and self._l.value:
            lib().orbx_local_map_destroy(self._l)
            self._l = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _search_local_points_batch(self, extractor, nframes, frustums, local_map, th=1.0, stream=None):
    """Tracking::SearchLocalPoints for the frames of `extractor`'s last device batch; frustums = list of preproc.frustum(...)
    -> (nmatches[nframes], kp_mp[nframes, cap])"""
    from .preproc import _Frustum
    arr = (_Frustum * nframes)(*frustums)
    h, w = extractor._last_shape
    cap = extractor.max_keypoints(w, h)
    kp_mp = np.empty((nframes, cap), np.int32)
    nm = np.zeros(nframes, np.int32)
    _check(lib().orbx_search_local_points_batch(self._m, extractor._h, nframes, arr, local_map._l, C.c_float(th), C.c_float(self.mfNNratio),
                                                _p(kp_mp), cap, _p(nm), C.c_void_p(stream or 0)))
    return nm, kp_mp


ORBmatcher.SearchLocalPointsBatch = _search_local_points_batch
