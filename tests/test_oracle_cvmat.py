"""The float32 matrix arithmetic behind the reference's cv::Mat expressions, as oracle/ref_shim/cv_prims.cc (the stand-in cv:: the
reference's own sources are compiled against in oracle/_ref) and oracle/match_oracle.cc (the restatement) evaluate it, held bit for bit
against the live python cv2 of this image: cv::SVD::compute of a 4x4 (OpenCV's own Jacobi -- LAPACK is never reached for a matrix this
small), s*row - row written into a row view (addWeighted, one rounding), Mat / s (convertTo, float scale), cv::gemm small and general
paths, cv::norm."""
import numpy as np
import pytest

import oracle
from oracle import refmatch as rm
from tri_cases import cv_gemm_np

cv2 = pytest.importorskip("cv2")
f32 = np.float32
needs_ref = pytest.mark.skipif(not rm.available(), reason="oracle/_ref not built")


def _mats(n, seed):
    rng = np.random.default_rng(seed)
    for t in range(n):
        A = rng.standard_normal((4, 4)).astype(f32)
        if t % 3 == 0:      # nearly rank 3, like the triangulation system of a consistent pair
            A[3] = (A[0] * f32(0.4) - A[1] * f32(1.3) + A[2] * f32(0.2) + rng.standard_normal(4).astype(f32) * f32(1e-4)).astype(f32)
        if t % 7 == 0:      # the shape Triangulate builds: rows (-1, 0, x, 0), (0, -1, y, 0)
            A[0] = [-1, 0, rng.standard_normal(), 0]; A[1] = [0, -1, rng.standard_normal(), 0]
        yield A


def test_jacobi_svd4_equals_cv2_svdecomp():
    As = list(_mats(600, 1))
    w_o, vt_o = oracle.jacobi_svd4(np.stack(As))
    for i, A in enumerate(As):
        w, _, vt = cv2.SVDecomp(A.copy(), flags=cv2.SVD_MODIFY_A | cv2.SVD_FULL_UV)
        assert np.array_equal(w.ravel().view(np.uint32), w_o[i].view(np.uint32)), i
        assert np.array_equal(vt.view(np.uint32), vt_o[i].view(np.uint32)), i


@needs_ref
def test_standin_svd_equals_cv2_svdecomp():
    for i, A in enumerate(_mats(300, 2)):
        w, _, vt = cv2.SVDecomp(A.copy(), flags=cv2.SVD_MODIFY_A | cv2.SVD_FULL_UV)
        w_s, vt_s = rm.cv_svd(A)
        assert np.array_equal(w.ravel().view(np.uint32), w_s.view(np.uint32)) and np.array_equal(vt.view(np.uint32), vt_s.view(np.uint32)), i


@needs_ref
def test_standin_scaled_row_difference_equals_cv2_addweighted():
    rng = np.random.default_rng(3)
    two_roundings = 0
    for _ in range(1500):
        a = rng.standard_normal((1, 4)).astype(f32); b = rng.standard_normal((1, 4)).astype(f32)
        s = float(f32(rng.standard_normal()))
        ref = cv2.addWeighted(a, s, b, -1.0, 0.0)          # what MatOp_AddEx::assign calls for alpha != 1, beta == -1
        got = rm.cv_scaled_sub_row(a, s, b)
        assert np.array_equal(ref.ravel().view(np.uint32), got.view(np.uint32))
        two_roundings += np.array_equal(ref.ravel(), (a * f32(s)).astype(f32).ravel() - b.ravel())
    assert two_roundings < 1400      # the float-product-then-subtract reading is NOT what cv2 computes


@needs_ref
def test_standin_division_by_scalar_is_a_float_scale():
    if not hasattr(cv2, "dnn"):
        pytest.skip("cv2.dnn absent (blobFromImage is the python-visible user of Mat *= double -> convertTo)")
    rng = np.random.default_rng(4)
    for _ in range(800):
        a = rng.standard_normal(3).astype(f32)
        w = float(f32(rng.standard_normal()))
        ref = cv2.dnn.blobFromImage(a.reshape(3, 1, 1), scalefactor=1.0 / w, size=(1, 3), mean=(0, 0, 0), swapRB=False, crop=False).ravel()
        assert np.array_equal(ref.view(np.uint32), rm.cv_div_scalar(a, w).view(np.uint32))


@needs_ref
def test_standin_gemm_equals_cv2_gemm():
    rng = np.random.default_rng(5)
    for _ in range(300):
        A = rng.standard_normal((3, 3)).astype(f32); B = rng.standard_normal((3, 3)).astype(f32)
        v = rng.standard_normal((3, 1)).astype(f32); c = rng.standard_normal((3, 1)).astype(f32)
        T = rng.standard_normal((4, 4)).astype(f32); U = rng.standard_normal((4, 4)).astype(f32)
        cases = [(A, v, 1.0, None, 0.0, 0), (A, v, -1.0, None, 0.0, 0), (A, v, 1.0, c, 1.0, 0), (A, B, 1.0, None, 0.0, 2),
                 (A, B, -1.0, None, 0.0, 2), (A, B, 1.0, None, 0.0, 0), (T, U, 1.0, None, 0.0, 0), (A, v, 1.0, None, 0.0, 1)]
        for a, b, alpha, cc, beta, flags in cases:
            ref = cv2.gemm(a, b, alpha, cc, beta, flags=flags)
            assert np.array_equal(ref.view(np.uint32), rm.cv_gemm(a, b, alpha, cc, beta, flags).view(np.uint32)), (alpha, beta, flags)
            assert np.array_equal(ref.view(np.uint32), cv_gemm_np(a, b, alpha, cc, beta, flags).view(np.uint32)), (alpha, beta, flags)


@needs_ref
def test_standin_norm_equals_cv2_norm():
    rng = np.random.default_rng(6)
    for _ in range(500):
        a = rng.standard_normal(3).astype(f32)
        assert cv2.norm(a.reshape(3, 1)) == rm.cv_norm(a)
        b = rng.standard_normal(3).astype(f32)
        assert rm.cv_dot(a, b) == float(a[0]) * float(b[0]) + float(a[1]) * float(b[1]) + float(a[2]) * float(b[2])
