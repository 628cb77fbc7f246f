// CPU ORACLE support (test infrastructure, not product code).
// The image primitives behind the stand-in cv:: header (opencv2/core/core.hpp in this directory): thin adapters from
// cv::Mat views to the oracle's restatements in oracle/orb_oracle.cc, which tests/test_oracle_cv2.py pins bit-exact
// against python cv2 4.13 (resize on every level, FAST on full frames and ROI cells incl. order and response,
// GaussianBlur with the OCV4 taps, fastAtan2).
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <vector>

#include "../orb_oracle.h"
#include "opencv2/core/core.hpp"

namespace cvshim {
// Blur tap set used by cv::GaussianBlur below: 0 = OpenCV 3.4.1-3.4.8 (per-tap cvRound), 1 = OpenCV >= 3.4.9 / 4.x.
thread_local int g_blur_taps = 0;
}  // namespace cvshim

namespace cv {

float fastAtan2(float y, float x) { return oro::fast_atan2(y, x); }

// cv::FAST(image, keypoints, threshold, nonmax) -- TYPE_9_16; KeyPoint(x, y, 7.f, -1, score) in row-major order.
void FAST(InputArray image, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression) {
  Mat img = image.getMat();
  std::vector<oro::Candidate> c;
  oro::fast9_16(img.data, img.cols, img.rows, (int)img.step, threshold, nonmaxSuppression, c);
  keypoints.clear();
  for (const auto& k : c) keypoints.push_back(KeyPoint((float)k.x, (float)k.y, 7.f, -1, (float)k.resp));
}

// cv::resize(src, dst, dsize, 0, 0, INTER_LINEAR) on 8UC1.  dst.create() is a no-op when dst already has dsize, so a
// ROI destination is written in place (ORBextractor.cc:1201 relies on it).
void resize(InputArray src_, OutputArray dst_, Size dsize, double fx, double fy, int interpolation) {
  Mat src = src_.getMat();
  assert(interpolation == INTER_LINEAR && dsize.width > 0 && dsize.height > 0);
  (void)fx; (void)fy; (void)interpolation;
  dst_.create(dsize, CV_8UC1);
  Mat dst = dst_.getMat();
  oro::resize_linear_u8(src.data, src.cols, src.rows, (int)src.step, dst.data, dst.cols, dst.rows, (int)dst.step);
}

static inline int reflect101(int i, int n) {
  if (n == 1) return 0;
  while (i < 0 || i >= n) i = i < 0 ? -i : 2 * (n - 1) - i;
  return i;
}

// cv::copyMakeBorder with BORDER_REFLECT_101 (optionally | BORDER_ISOLATED).  The reference calls it on whole images
// or with BORDER_ISOLATED (ORBextractor.cc:1203, 1214), so pixels outside the source view are never consulted.
void copyMakeBorder(InputArray src_, OutputArray dst_, int top, int bottom, int left, int right, int borderType, const Scalar&) {
  Mat src = src_.getMat();
  assert((borderType & ~BORDER_ISOLATED) == BORDER_REFLECT_101);
  (void)borderType;
  dst_.create(src.rows + top + bottom, src.cols + left + right, CV_8UC1);
  Mat dst = dst_.getMat();
  const int w = src.cols, h = src.rows;
  // the source may be the interior view of dst itself (ORBextractor.cc:1203): go through a dense copy
  std::vector<uchar> tmp((size_t)w * h);
  for (int y = 0; y < h; y++) memcpy(&tmp[(size_t)y * w], src.ptr(y), (size_t)w);
  for (int y = 0; y < dst.rows; y++) {
    const uchar* s = &tmp[(size_t)reflect101(y - top, h) * w];
    uchar* d = dst.ptr(y);
    for (int x = 0; x < dst.cols; x++) d[x] = s[reflect101(x - left, w)];
  }
}

// cv::GaussianBlur(src, dst, Size(7,7), 2, 2, BORDER_REFLECT_101) on 8UC1 (in place allowed).
void GaussianBlur(InputArray src_, OutputArray dst_, Size ksize, double sigmaX, double sigmaY, int borderType) {
  Mat src = src_.getMat();
  assert(ksize.width == 7 && ksize.height == 7 && sigmaX == 2 && sigmaY == 2 && borderType == BORDER_REFLECT_101);
  (void)ksize; (void)sigmaX; (void)sigmaY; (void)borderType;
  std::vector<uchar> tmp((size_t)src.cols * src.rows);
  oro::gaussian_blur7_u8(src.data, src.cols, src.rows, (int)src.step, tmp.data(), src.cols, cvshim::g_blur_taps);
  dst_.create(src.rows, src.cols, CV_8UC1);
  Mat dst = dst_.getMat();
  for (int y = 0; y < dst.rows; y++) memcpy(dst.ptr(y), &tmp[(size_t)y * dst.cols], (size_t)dst.cols);
}

// cv::KeyPointsFilter::retainBest: keep the npoints strongest responses (and every tie with the weakest kept one).
void KeyPointsFilter::retainBest(std::vector<KeyPoint>& kp, int n) {
  if (n < 0 || (int)kp.size() <= n) return;
  if (n == 0) { kp.clear(); return; }
  std::nth_element(kp.begin(), kp.begin() + n - 1, kp.end(), [](const KeyPoint& a, const KeyPoint& b) { return a.response > b.response; });
  const float cut = kp[n - 1].response;
  auto end = std::partition(kp.begin() + n, kp.end(), [cut](const KeyPoint& k) { return k.response >= cut; });
  kp.resize(end - kp.begin());
}

}  // namespace cv

// ---- matrix arithmetic behind cv::MatExpr (CV_32F / CV_64F single channel) ------------------------------------------------
// Restated from the published OpenCV sources (modules/core/src/matmul.cpp, convert_scale.cpp, arithm.cpp, lapack.cpp, stat.cpp)
// and pinned against python cv2 4.13 (cv2.gemm / invert / norm / transpose on float32) in tests/test_oracle_cvmat.py.
namespace {
template <typename T> inline T el(const cv::Mat& m, int y, int x) { return m.at<T>(y, x); }
}  // namespace

namespace cv {

Mat Mat::ones(int r, int c, int type) {
  Mat m(r, c, type);
  for (int y = 0; y < r; y++)
    for (int x = 0; x < c; x++) {
      if (type == CV_32F) m.at<float>(y, x) = 1.f; else if (type == CV_64F) m.at<double>(y, x) = 1.0; else if (type == CV_32S) m.at<int>(y, x) = 1; else m.at<uchar>(y, x) = 1;
    }
  return m;
}
Mat Mat::eye(int r, int c, int type) {
  Mat m = zeros(r, c, type);
  for (int i = 0; i < std::min(r, c); i++) {
    if (type == CV_32F) m.at<float>(i, i) = 1.f; else if (type == CV_64F) m.at<double>(i, i) = 1.0; else if (type == CV_32S) m.at<int>(i, i) = 1; else m.at<uchar>(i, i) = 1;
  }
  return m;
}

// Mat::convertTo: 8U -> 32F with unit scale (Frame.cc:1001-1018) and 32F -> 32F with a scale (cvtScale_<float,float,float>:
// the scale and shift are rounded to float, the product is a float product).
void Mat::convertTo(const _OutputArray& dst_, int rtype, double alpha, double beta) const {
  if (rtype < 0) rtype = type_;
  Mat src = *this;   // in-place conversion allowed
  Mat dst(rows, cols, rtype);
  const float fa = (float)alpha, fb = (float)beta;
  for (int y = 0; y < rows; y++)
    for (int x = 0; x < cols; x++) {
      if (type_ == CV_8U && rtype == CV_32F) dst.at<float>(y, x) = (alpha == 1 && beta == 0) ? (float)src.at<uchar>(y, x) : (float)src.at<uchar>(y, x) * fa + fb;
      else if (type_ == CV_32F && rtype == CV_32F) dst.at<float>(y, x) = (alpha == 1 && beta == 0) ? src.at<float>(y, x) : src.at<float>(y, x) * fa + fb;
      else if (type_ == CV_64F && rtype == CV_64F) dst.at<double>(y, x) = src.at<double>(y, x) * alpha + beta;
      else if (type_ == CV_32F && rtype == CV_64F) dst.at<double>(y, x) = src.at<float>(y, x) * alpha + beta;
      else if (type_ == CV_64F && rtype == CV_32F) dst.at<float>(y, x) = (float)(src.at<double>(y, x) * alpha + beta);
      else if (type_ == CV_8U && rtype == CV_16S) dst.at<short>(y, x) = (short)src.at<uchar>(y, x);
      else if (type_ == CV_16S && rtype == CV_16S) dst.at<short>(y, x) = src.at<short>(y, x);
      else if (type_ == CV_16S && rtype == CV_32F) dst.at<float>(y, x) = (float)src.at<short>(y, x);
      else { assert(!"convertTo: unsupported type pair"); }
    }
  const_cast<_OutputArray&>(dst_).create(rows, cols, rtype);
  Mat out = dst_.getMat();
  if (out.data == dst.data) return;
  if (out.rows == rows && out.cols == cols && out.type() == rtype) dst.copyTo(out);
  // an output of another shape/type was re-created by create(); getMat() returned the new header
}

// Mat - scalar / Mat + scalar (MatOp_AddEx with a scalar term -> convertTo(type, alpha, s)): float work type, result saturated
// to the matrix type (cvRound for CV_16S).
void cvshim_add_scalar(const Mat& a, double alpha, double s, Mat& d) {
  Mat r(a.rows, a.cols, a.type());
  const float fa = (float)alpha, fs = (float)s;
  for (int y = 0; y < a.rows; y++)
    for (int x = 0; x < a.cols; x++) {
      if (a.type() == CV_16S) {
        const int v = cvRound((float)a.at<short>(y, x) * fa + fs);
        r.at<short>(y, x) = (short)std::min(32767, std::max(-32768, v));
      } else if (a.type() == CV_32F) r.at<float>(y, x) = a.at<float>(y, x) * fa + fs;
      else { assert(!"Mat +- scalar: unsupported type"); }
    }
  d = r;
}

void cvshim_transpose(const Mat& a, Mat& d) {
  Mat r(a.cols, a.rows, a.type());
  for (int y = 0; y < a.rows; y++)
    for (int x = 0; x < a.cols; x++) memcpy(r.ptr(x) + (size_t)y * r.elemSize(), a.ptr(y) + (size_t)x * a.elemSize(), a.elemSize());
  d = r;
}

// MatOp_AddEx::assign: alpha*a (+ beta*b).  Float matrices: cv::add / subtract are single float operations, scaleAdd / addWeighted /
// convertTo multiply by the scale rounded to float.
void cvshim_add_weighted(const Mat& a, double alpha, const Mat& b, double beta, Mat& d) {
  assert(a.type() == CV_32F || a.type() == CV_64F);
  Mat r(a.rows, a.cols, a.type());
  const bool hasB = !b.empty();
  if (hasB) assert(b.rows == a.rows && b.cols == a.cols && b.type() == a.type());
  for (int y = 0; y < a.rows; y++)
    for (int x = 0; x < a.cols; x++) {
      if (a.type() == CV_32F) {
        const float av = a.at<float>(y, x), fa = (float)alpha, fb = (float)beta;
        float o;
        if (!hasB) o = (alpha == 1) ? av : (alpha == -1 ? -av : av * fa);
        else {
          const float bv = b.at<float>(y, x);
          // cv::scaleAdd / cv::addWeighted on CV_32F as cv2 4.13 evaluates them on an FMA-capable host: the scaled sum is rounded
          // to float ONCE (tests/test_oracle_cvmat.py: equal to the double expression on every sample, to the two-rounding float
          // expression on a third of them)
          if (alpha == 1) o = beta == 1 ? av + bv : (beta == -1 ? av - bv : (float)((double)bv * (double)fb + (double)av));
          else if (beta == 1) o = alpha == -1 ? bv - av : (float)((double)av * (double)fa + (double)bv);
          else o = (float)((double)av * alpha + (double)bv * beta);
        }
        r.at<float>(y, x) = o;
      } else {
        const double av = a.at<double>(y, x);
        r.at<double>(y, x) = hasB ? av * alpha + b.at<double>(y, x) * beta : av * alpha;
      }
    }
  d = r;
}

// cv::gemm on CV_32F.  flags == 0 and inner dimension 2..4 with a matching output side: the small-matrix path -- every output
// element is t = a0*b0 + a1*b1 (+ a2*b2 (+ a3*b3)) in FLOAT, left to right, then (float)(t*alpha + c*beta) in double.
// Every other case (a transposed operand, other sizes): GEMMSingleMul<float,double> -- products and sums in double, then
// (float)(s*alpha + c*beta).
void cvshim_gemm(const Mat& a, const Mat& b, double alpha, const Mat& c, double beta, Mat& d, int flags) {
  assert(a.type() == CV_32F && b.type() == CV_32F);
  const bool ta = flags & GEMM_1_T, tb = flags & GEMM_2_T, tc = flags & GEMM_3_T;
  const int M = ta ? a.cols : a.rows, K = ta ? a.rows : a.cols, N = tb ? b.rows : b.cols;
  assert((tb ? b.cols : b.rows) == K);
  const bool hasC = !c.empty() && beta != 0;
  if (hasC) assert((tc ? c.cols : c.rows) == M && (tc ? c.rows : c.cols) == N);
  Mat r(M, N, CV_32F);
  const bool small = flags == 0 && K >= 2 && K <= 4 && (K == N || K == M);
  for (int i = 0; i < M; i++)
    for (int j = 0; j < N; j++) {
      const double cv_ = hasC ? (double)(tc ? c.at<float>(j, i) : c.at<float>(i, j)) : 0.0;
      if (small) {
        float t = a.at<float>(i, 0) * b.at<float>(0, j);
        for (int k = 1; k < K; k++) t = t + a.at<float>(i, k) * b.at<float>(k, j);
        r.at<float>(i, j) = (float)((double)t * alpha + cv_ * (hasC ? beta : 0.0));
      } else {
        double s = 0;
        for (int k = 0; k < K; k++) s += (double)(ta ? a.at<float>(k, i) : a.at<float>(i, k)) * (double)(tb ? b.at<float>(j, k) : b.at<float>(k, j));
        r.at<float>(i, j) = (float)(s * alpha + cv_ * (hasC ? beta : 0.0));
      }
    }
  d = r;
}

// cv::invert(DECOMP_LU) of a 3x3 (and 2x2) CV_32F matrix: closed form evaluated in double, rounded to float once per element.
void cvshim_invert(const Mat& a, Mat& d) {
  assert(a.type() == CV_32F && a.rows == a.cols && (a.rows == 2 || a.rows == 3));
  Mat r(a.rows, a.cols, CV_32F);
  auto S = [&](int y, int x) { return (double)a.at<float>(y, x); };
  if (a.rows == 2) {
    double det = S(0, 0) * S(1, 1) - S(0, 1) * S(1, 0);
    if (det != 0) {
      det = 1. / det;
      const double t0 = S(0, 0) * det, t1 = S(1, 1) * det;
      r.at<float>(1, 1) = (float)t0; r.at<float>(0, 0) = (float)t1;
      r.at<float>(0, 1) = (float)(-S(0, 1) * det); r.at<float>(1, 0) = (float)(-S(1, 0) * det);
    } else r = Mat::zeros(2, 2, CV_32F);
  } else {
    double det = S(0, 0) * (S(1, 1) * S(2, 2) - S(1, 2) * S(2, 1)) - S(0, 1) * (S(1, 0) * S(2, 2) - S(1, 2) * S(2, 0)) +
                 S(0, 2) * (S(1, 0) * S(2, 1) - S(1, 1) * S(2, 0));
    if (det != 0) {
      det = 1. / det;
      r.at<float>(0, 0) = (float)((S(1, 1) * S(2, 2) - S(1, 2) * S(2, 1)) * det);
      r.at<float>(0, 1) = (float)((S(0, 2) * S(2, 1) - S(0, 1) * S(2, 2)) * det);
      r.at<float>(0, 2) = (float)((S(0, 1) * S(1, 2) - S(0, 2) * S(1, 1)) * det);
      r.at<float>(1, 0) = (float)((S(1, 2) * S(2, 0) - S(1, 0) * S(2, 2)) * det);
      r.at<float>(1, 1) = (float)((S(0, 0) * S(2, 2) - S(0, 2) * S(2, 0)) * det);
      r.at<float>(1, 2) = (float)((S(0, 2) * S(1, 0) - S(0, 0) * S(1, 2)) * det);
      r.at<float>(2, 0) = (float)((S(1, 0) * S(2, 1) - S(1, 1) * S(2, 0)) * det);
      r.at<float>(2, 1) = (float)((S(0, 1) * S(2, 0) - S(0, 0) * S(2, 1)) * det);
      r.at<float>(2, 2) = (float)((S(0, 0) * S(1, 1) - S(0, 1) * S(1, 0)) * det);
    } else r = Mat::zeros(3, 3, CV_32F);
  }
  d = r;
}

// Mat::dot on CV_32F: products and the running sum in double (dotProd_32f's scalar path; vectors here have 3 elements).
double Mat::dot(const Mat& m) const {
  assert(type_ == CV_32F && m.type() == CV_32F && total() == m.total());
  double s = 0;
  const int n = (int)total();
  for (int i = 0; i < n; i++) s += (double)at<float>(i) * (double)m.at<float>(i);
  return s;
}

// cv::norm: NORM_L2 of a CV_32F matrix = sqrt of the double sum of squares; NORM_L1 of the difference of two CV_32F matrices =
// double sum of |a - b| with the difference taken in float (normDiffL1_<float,double>).
double norm(InputArray a_, int normType) {
  Mat a = a_.getMat();
  assert(a.type() == CV_32F && (normType == NORM_L2 || normType == NORM_L1));
  double s = 0;
  for (int y = 0; y < a.rows; y++)
    for (int x = 0; x < a.cols; x++) {
      const double v = a.at<float>(y, x);
      s += normType == NORM_L2 ? v * v : std::fabs(v);
    }
  return normType == NORM_L2 ? std::sqrt(s) : s;
}
double norm(InputArray a_, InputArray b_, int normType) {
  Mat a = a_.getMat(), b = b_.getMat();
  assert(a.type() == b.type() && a.rows == b.rows && a.cols == b.cols && (normType == NORM_L1 || normType == NORM_L2));
  if (a.type() == CV_16S) {   // normDiffL1_<short,int>: |a - b| summed in int (Frame.cc:1018)
    assert(normType == NORM_L1);
    long long si = 0;
    for (int y = 0; y < a.rows; y++)
      for (int x = 0; x < a.cols; x++) si += std::abs((int)a.at<short>(y, x) - (int)b.at<short>(y, x));
    return (double)si;
  }
  assert(a.type() == CV_32F);
  double s = 0;
  for (int y = 0; y < a.rows; y++)
    for (int x = 0; x < a.cols; x++) {
      const double v = (double)(a.at<float>(y, x) - b.at<float>(y, x));
      s += normType == NORM_L1 ? std::fabs(v) : v * v;
    }
  return normType == NORM_L1 ? s : std::sqrt(s);
}


void hconcat(InputArray a_, InputArray b_, OutputArray dst) {
  Mat a = a_.getMat(), b = b_.getMat();
  assert(a.rows == b.rows && a.type() == b.type());
  Mat r(a.rows, a.cols + b.cols, a.type());
  for (int y = 0; y < a.rows; y++) {
    memcpy(r.ptr(y), a.ptr(y), (size_t)a.cols * a.elemSize());
    memcpy(r.ptr(y) + (size_t)a.cols * a.elemSize(), b.ptr(y), (size_t)b.cols * b.elemSize());
  }
  dst.create(r.rows, r.cols, r.type());
  Mat d = dst.getMat();
  r.copyTo(d);
}

// JacobiSVDImpl_<float> (modules/core/src/lapack.cpp) for a square n x n CV_32F matrix, n <= 8: At = src^T (rows = columns of src),
// Hestenes rotations until no pair is rotated (at most max(n, 30) sweeps), W accumulated in double, c / s rounded to float, rotated
// elements float(c*x) + float(s*y); singular values sorted descending with the rows of Vt following.  u is not produced (the
// reference reads vt only, KannalaBrandt8.cpp:441-442).  Pinned to cv2.SVDecomp in tests/test_oracle_cvmat.py.
void SVD::compute(InputArray src_, OutputArray w_, OutputArray u_, OutputArray vt_, int) {
  Mat src = src_.getMat();
  assert(src.type() == CV_32F && src.rows == src.cols && src.rows <= 8);
  const int n = src.rows, m = n;
  float At[8][8], Vt[8][8];
  double W[8];
  for (int i = 0; i < n; i++)
    for (int k = 0; k < m; k++) { At[i][k] = src.at<float>(k, i); Vt[i][k] = i == k ? 1.f : 0.f; }
  for (int i = 0; i < n; i++) {
    double sd = 0;
    for (int k = 0; k < m; k++) { const float t = At[i][k]; sd += (double)t * t; }
    W[i] = sd;
  }
  const float eps = FLT_EPSILON * 2;
  const int max_iter = std::max(m, 30);
  for (int iter = 0; iter < max_iter; iter++) {
    bool changed = false;
    for (int i = 0; i < n - 1; i++)
      for (int j = i + 1; j < n; j++) {
        float *Ai = At[i], *Aj = At[j];
        double a = W[i], p = 0, b = W[j];
        for (int k = 0; k < m; k++) p += (double)Ai[k] * Aj[k];
        if (std::abs(p) <= eps * std::sqrt((double)a * b)) continue;
        p *= 2;
        const double beta = a - b, gamma = hypot((double)p, beta);
        float c, s;
        if (beta < 0) {
          const double delta = (gamma - beta) * 0.5;
          s = (float)std::sqrt(delta / gamma);
          c = (float)(p / (gamma * s * 2));
        } else {
          c = (float)std::sqrt((gamma + beta) / (gamma * 2));
          s = (float)(p / (gamma * c * 2));
        }
        a = b = 0;
        for (int k = 0; k < m; k++) {
          const float t0 = c * Ai[k] + s * Aj[k];
          const float t1 = -s * Ai[k] + c * Aj[k];
          Ai[k] = t0; Aj[k] = t1;
          a += (double)t0 * t0; b += (double)t1 * t1;
        }
        W[i] = a; W[j] = b;
        changed = true;
        float *Vi = Vt[i], *Vj = Vt[j];
        for (int k = 0; k < n; k++) {
          const float t0 = c * Vi[k] + s * Vj[k];
          const float t1 = -s * Vi[k] + c * Vj[k];
          Vi[k] = t0; Vj[k] = t1;
        }
      }
    if (!changed) break;
  }
  for (int i = 0; i < n; i++) {
    double sd = 0;
    for (int k = 0; k < m; k++) { const float t = At[i][k]; sd += (double)t * t; }
    W[i] = std::sqrt(sd);
  }
  for (int i = 0; i < n - 1; i++) {
    int j = i;
    for (int k = i + 1; k < n; k++)
      if (W[j] < W[k]) j = k;
    if (i != j) {
      std::swap(W[i], W[j]);
      for (int k = 0; k < m; k++) std::swap(At[i][k], At[j][k]);
      for (int k = 0; k < n; k++) std::swap(Vt[i][k], Vt[j][k]);
    }
  }
  w_.create(n, 1, CV_32F);
  vt_.create(n, n, CV_32F);
  Mat w = w_.getMat(), vt = vt_.getMat();
  for (int i = 0; i < n; i++) {
    w.at<float>(i, 0) = (float)W[i];
    for (int k = 0; k < n; k++) vt.at<float>(i, k) = Vt[i][k];
  }
  (void)u_;
}

}  // namespace cv
