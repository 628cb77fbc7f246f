"""Geometry in front of ORBmatcher::SearchForTriangulation on KannalaBrandt8 keyframes (src/ORBmatcher.cc:988-1019), evaluated with
OpenCV's rounding: cv_gemm_np restates the float32 cv::gemm rules of oracle/ref_shim/cv_prims.cc (tests/test_oracle_cvmat.py holds both
against cv2.gemm), so that a test can hand the CPU restatement exactly the R12 / t12 / epipole the reference's own code forms."""
import numpy as np

f32 = np.float32


def cv_gemm_np(a, b, alpha=1.0, c=None, beta=0.0, flags=0):
    """cv::gemm on CV_32F: flags == 0 and inner dimension 2..4 matching an output side -> float products summed left to right, then
    (float)(t*alpha + c*beta) in double; otherwise products and sums in double."""
    a = np.asarray(a, f32); b = np.asarray(b, f32)
    A = a.T if flags & 1 else a
    B = b.T if flags & 2 else b
    M, K = A.shape; N = B.shape[1]
    out = np.empty((M, N), f32)
    small = flags == 0 and 2 <= K <= 4 and (K == N or K == M)
    for i in range(M):
        for j in range(N):
            cv = 0.0 if c is None or beta == 0 else float(np.asarray(c, f32).reshape(M, N)[i, j]) * beta
            if small:
                t = f32(A[i, 0] * B[0, j])
                for k in range(1, K):
                    t = f32(t + f32(A[i, k] * B[k, j]))
                out[i, j] = f32(float(t) * alpha + cv)
            else:
                s = 0.0
                for k in range(K):
                    s += float(A[i, k]) * float(B[k, j])
                out[i, j] = f32(s * alpha + cv)
    return out


def keyframe_poses(Tcw, Tlr=None, gemm=cv_gemm_np):
    """(R, t, Ow) of the left camera and, with Tlr (3x4, left <- right), of the right one, as oracle/ref_shim/ref_match_harness.cc builds a
    KeyFrame (KeyFrame::SetPose, Frame.cc:1186-1195 for Trl)."""
    Tcw = np.asarray(Tcw, f32)
    R, t = Tcw[:3, :3].copy(), Tcw[:3, 3:4].copy()
    Ow = gemm(np.ascontiguousarray(R.T), t, alpha=-1.0)
    out = dict(R=R, t=t, Ow=Ow)
    if Tlr is not None:
        Tlr = np.asarray(Tlr, f32)
        Rrl = np.ascontiguousarray(Tlr[:, :3].T)
        trl = gemm(Rrl, Tlr[:, 3:4].copy(), alpha=-1.0)
        Trl4 = np.eye(4, dtype=f32)
        Trl4[:3, :3] = Rrl; Trl4[:3, 3:4] = trl
        TcwR = gemm(Trl4, Tcw)
        out.update(Rr=TcwR[:3, :3].copy(), tr=TcwR[:3, 3:4].copy())
    return out


def triangulation_geometry_kb8(oracle, T1, T2, cam8, Tlr=None, gemm=cv_gemm_np, libm_mode=None, Tlr2=None):
    """(pairs, ep) for oracle.search_for_triangulation_kb8 / ORBmatcher.SearchForTriangulationKB8: one record without a second camera
    (ORBmatcher.cc:1000-1003), the four records ll, lr, rl, rr with one (:1005-1019); ep = pKF2->mpCamera->project(R2w*Cw + t2w) (:988-994)."""
    libm_mode = oracle.LIBM_PORTABLE if libm_mode is None else libm_mode
    k1, k2 = keyframe_poses(T1, Tlr, gemm), keyframe_poses(T2, Tlr if Tlr2 is None else Tlr2, gemm)
    C2 = gemm(k2["R"], k1["Ow"], c=k2["t"], beta=1.0)
    ep = oracle.kb8_project(cam8, C2.reshape(1, 3), libm_mode)[0]
    pairs = []
    if Tlr is None:
        R12 = gemm(k1["R"], k2["R"], flags=2)
        tmp = gemm(k1["R"], k2["R"], alpha=-1.0, flags=2)          # -R1w * R2w.t()  (one gemm, alpha = -1)
        t12 = gemm(tmp, k2["t"], c=k1["t"], beta=1.0)              # ... * t2w + t1w
        pairs.append(oracle.kb8_pair(R12, t12, cam8, cam8))
    else:
        for a in ("", "r"):
            for b in ("", "r"):
                R1, t1, R2, t2 = k1["R" + a], k1["t" + a], k2["R" + b], k2["t" + b]
                Rab = gemm(R1, R2, flags=2)
                tmp = gemm(np.ascontiguousarray(R2.T), t2, alpha=-1.0)   # -R2.t() * t2: the transpose is materialised, then one gemm
                tab = gemm(R1, tmp, c=t1, beta=1.0)
                pairs.append(oracle.kb8_pair(Rab, tab, cam8, cam8))
    return pairs, (float(ep[0]), float(ep[1]))
