// CPU ORACLE support (test infrastructure, not product code).
//
// Stand-in for the slice of OpenCV's C++ API that the reference's src/ORBextractor.cc (and
// ORBmatcher::DescriptorDistance, src/ORBmatcher.cc:2463-2483) touches, so that the UNMODIFIED reference source
// compiles here with g++ (the image has no C++ OpenCV).  Written from the public OpenCV API documentation; only the
// members the reference calls exist.  The six image primitives (FAST, resize, copyMakeBorder, GaussianBlur,
// fastAtan2, cvRound) are declared here and implemented in oracle/ref_shim/cv_prims.cc on top of the oracle's
// restatements, each of which is pinned bit-exact against python cv2 4.13 (tests/test_oracle_cv2.py).
//
// What this buys: the quadtree, the cell loop and its threshold fallback, the lapping order, IC_Angle, rBRIEF and the
// 256-pair pattern table are then the reference's OWN code, not a reading of it (oracle/_ref/libref_orb.so, built by
// oracle/Makefile from /root/reference/src/ORBextractor.cc by path).
#pragma once
#include <algorithm>
#include <cassert>
#include <cfloat>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>

typedef unsigned char uchar;
typedef unsigned short ushort;

#define CV_PI 3.1415926535897932384626433832795
#define CV_8U 0
#define CV_8UC1 0
#define CV_16S 3
#define CV_32S 4
#define CV_32F 5
#define CV_64F 6
#define CV_32FC1 5

// cvRound: SSE2 cvtsd2si / lrint under the default rounding mode (round-half-even), as OpenCV's fast_math.hpp.
static inline int cvRound(double v) { return (int)lrint(v); }
static inline int cvRound(float v) { return (int)lrintf(v); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v) { int i = (int)v; return i - (i > v); }
static inline int cvFloor(float v) { int i = (int)v; return i - (i > v); }
static inline int cvCeil(double v) { int i = (int)v; return i + (i < v); }
static inline int cvCeil(float v) { int i = (int)v; return i + (i < v); }

namespace cv {

template <typename T>
struct Point_ {
  T x, y;
  Point_() : x(0), y(0) {}
  Point_(T x_, T y_) : x(x_), y(y_) {}
  template <typename U>
  Point_(const Point_<U>& o) : x((T)o.x), y((T)o.y) {}
  Point_& operator*=(float s) { x = (T)(x * s); y = (T)(y * s); return *this; }   // saturate_cast<float> of a float product
  Point_& operator*=(double s) { x = (T)(x * s); y = (T)(y * s); return *this; }
  Point_& operator*=(int s) { x = (T)(x * s); y = (T)(y * s); return *this; }
};
typedef Point_<int> Point2i;
typedef Point2i Point;
typedef Point_<float> Point2f;
typedef Point_<double> Point2d;

template <typename T>
struct Point3_ {
  T x, y, z;
  Point3_() : x(0), y(0), z(0) {}
  Point3_(T x_, T y_, T z_) : x(x_), y(y_), z(z_) {}
};
typedef Point3_<float> Point3f;

template <typename T>
struct Size_ {
  T width, height;
  Size_() : width(0), height(0) {}
  Size_(T w, T h) : width(w), height(h) {}
};
typedef Size_<int> Size;

template <typename T>
struct Rect_ {
  T x, y, width, height;
  Rect_() : x(0), y(0), width(0), height(0) {}
  Rect_(T x_, T y_, T w, T h) : x(x_), y(y_), width(w), height(h) {}
};
typedef Rect_<int> Rect;

struct Scalar {
  double val[4];
  Scalar() { val[0] = val[1] = val[2] = val[3] = 0; }
  Scalar(double v0) { val[0] = v0; val[1] = val[2] = val[3] = 0; }
};

// Same seven 4-byte fields, in the same order, as cv::KeyPoint.
class KeyPoint {
 public:
  KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
  KeyPoint(float x, float y, float size_, float angle_ = -1, float response_ = 0, int octave_ = 0, int class_id_ = -1)
      : pt(x, y), size(size_), angle(angle_), response(response_), octave(octave_), class_id(class_id_) {}
  Point2f pt;
  float size, angle, response;
  int octave, class_id;
};

class _InputArray;
class _OutputArray;
class MatExpr;

static inline int cvshim_elem_size(int type) { return type == CV_8U ? 1 : (type == CV_16S ? 2 : (type == CV_64F ? 8 : 4)); }

// Reference-counted single-channel matrix header: CV_8U (images, descriptors), CV_32F (poses, points), CV_32S, CV_64F.
class Mat {
 public:
  struct Step {  // converts like cv::MatStep (bytes per row)
    size_t v;
    operator size_t() const { return v; }
  };
  Mat() : rows(0), cols(0), data(nullptr), step{0}, type_(CV_8U), base_(nullptr), refs_(nullptr) {}
  Mat(int r, int c, int type) : Mat() { create(r, c, type); }
  Mat(Size s, int type) : Mat() { create(s.height, s.width, type); }
  // user-allocated data, not owned (cv::Mat(rows, cols, type, void*, step))
  Mat(int r, int c, int type, void* d, size_t st = 0)
      : rows(r), cols(c), data((uchar*)d), step{st ? st : (size_t)c * cvshim_elem_size(type)}, type_(type), base_(nullptr), refs_(nullptr) {}
  Mat(const Mat& o) : rows(o.rows), cols(o.cols), data(o.data), step(o.step), type_(o.type_), base_(o.base_), refs_(o.refs_) { if (refs_) ++*refs_; }
  Mat(const MatExpr& e);
  Mat& operator=(const Mat& o) {
    if (this != &o) {
      if (o.refs_) ++*o.refs_;
      release();
      rows = o.rows; cols = o.cols; data = o.data; step = o.step; type_ = o.type_; base_ = o.base_; refs_ = o.refs_;
    }
    return *this;
  }
  Mat& operator=(const MatExpr& e);
  ~Mat() { release(); }
  void release() {
    if (refs_ && --*refs_ == 0) { free(base_); free(refs_); }
    rows = cols = 0; data = nullptr; base_ = nullptr; refs_ = nullptr; step.v = 0;
  }
  // cv::Mat::create: no-op when shape and type already match (that is what lets resize() write into a ROI in place)
  void create(int r, int c, int type) {
    if (data && rows == r && cols == c && type_ == type) return;
    release();
    rows = r; cols = c; type_ = type; step.v = (size_t)c * cvshim_elem_size(type);
    base_ = (uchar*)malloc((size_t)r * step.v + 64);   // plain malloc: matrix storage stays outside the node allocator
    refs_ = (int*)malloc(sizeof(int));
    *refs_ = 1;
    data = base_;
  }
  void create(Size s, int type) { create(s.height, s.width, type); }
  static Mat zeros(int r, int c, int type) { Mat m(r, c, type); for (int y = 0; y < r; y++) memset(m.ptr(y), 0, m.step.v); return m; }
  static Mat ones(int r, int c, int type);
  static Mat eye(int r, int c, int type);
  Mat operator()(const Rect& r) const {
    Mat m(*this);
    m.data = data + (size_t)r.y * step.v + (size_t)r.x * elemSize(); m.rows = r.height; m.cols = r.width;
    return m;
  }
  Mat rowRange(int a, int b) const { return (*this)(Rect(0, a, cols, b - a)); }
  Mat colRange(int a, int b) const { return (*this)(Rect(a, 0, b - a, rows)); }
  Mat row(int y) const { return (*this)(Rect(0, y, cols, 1)); }
  Mat col(int x) const { return (*this)(Rect(x, 0, 1, rows)); }
  Mat clone() const { Mat m(rows, cols, type_); for (int y = 0; y < rows; y++) memcpy(m.ptr(y), ptr(y), (size_t)cols * elemSize()); return m; }
  void copyTo(const _OutputArray& dst) const;
  void convertTo(const _OutputArray& dst, int rtype, double alpha = 1, double beta = 0) const;
  MatExpr t() const;
  MatExpr inv(int method = 0) const;
  double dot(const Mat& m) const;
  bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
  int type() const { return type_; }
  int depth() const { return type_; }
  int channels() const { return 1; }
  size_t elemSize() const { return (size_t)cvshim_elem_size(type_); }
  size_t step1() const { return step.v / elemSize(); }
  size_t total() const { return (size_t)rows * cols; }
  Size size() const { return Size(cols, rows); }
  bool isContinuous() const { return rows <= 1 || step.v == (size_t)cols * elemSize(); }
  uchar* ptr(int y = 0) { return data + (size_t)y * step.v; }
  const uchar* ptr(int y = 0) const { return data + (size_t)y * step.v; }
  template <typename T> T* ptr(int y = 0) { return (T*)(data + (size_t)y * step.v); }
  template <typename T> const T* ptr(int y = 0) const { return (const T*)(data + (size_t)y * step.v); }
  template <typename T> T& at(int y, int x) { return ((T*)(data + (size_t)y * step.v))[x]; }
  template <typename T> const T& at(int y, int x) const { return ((const T*)(data + (size_t)y * step.v))[x]; }
  // single-index access on a row or column vector (continuous or not)
  template <typename T> T& at(int i) { return rows == 1 ? ((T*)data)[i] : (cols == 1 ? *(T*)(data + (size_t)i * step.v) : ((T*)(data + (size_t)(i / cols) * step.v))[i % cols]); }
  template <typename T> const T& at(int i) const { return const_cast<Mat*>(this)->at<T>(i); }

  int rows, cols;
  uchar* data;
  Step step;

 protected:
  int type_;
  uchar* base_;
  int* refs_;
};

// cv::Mat_<float> with the comma initialiser (cv::Mat_<float>(3,3) << a, b, c, ...), Pinhole.cpp:195-200.
template <typename T> struct cvshim_type;
template <> struct cvshim_type<float> { enum { value = CV_32F }; };
template <> struct cvshim_type<double> { enum { value = CV_64F }; };
template <> struct cvshim_type<int> { enum { value = CV_32S }; };
template <> struct cvshim_type<short> { enum { value = CV_16S }; };
template <> struct cvshim_type<uchar> { enum { value = CV_8U }; };
template <typename T>
class Mat_ : public Mat {
 public:
  Mat_() {}
  Mat_(int r, int c) : Mat(r, c, cvshim_type<T>::value) {}
  T& operator()(int y, int x) { return this->template at<T>(y, x); }
  const T& operator()(int y, int x) const { return this->template at<T>(y, x); }
};
template <typename T>
class MatCommaInitializer_ {
 public:
  MatCommaInitializer_(Mat_<T>* m) : m_(m), i_(0) {}
  template <typename U> MatCommaInitializer_& operator,(U v) { put((T)v); return *this; }
  void put(T v) { m_->template at<T>(i_ / m_->cols, i_ % m_->cols) = v; ++i_; }
  operator Mat() const { return *m_; }
  operator Mat_<T>() const { return *m_; }
 private:
  Mat_<T>* m_;
  int i_;
};
template <typename T, typename U>
inline MatCommaInitializer_<T> operator<<(const Mat_<T>& m, U v) {
  MatCommaInitializer_<T> ci(const_cast<Mat_<T>*>(&m));
  ci.put((T)v);
  return ci;
}

// cv::InputArray / cv::OutputArray are typedefs of const references to proxy objects, as in OpenCV.
class _InputArray {
 public:
  _InputArray() : m_(nullptr) {}
  _InputArray(const Mat& m) : m_(const_cast<Mat*>(&m)) {}
  Mat getMat() const { return m_ ? *m_ : Mat(); }
  bool empty() const { return !m_ || m_->empty(); }
 protected:
  Mat* m_;
};
class _OutputArray : public _InputArray {
 public:
  _OutputArray() {}
  _OutputArray(Mat& m) : _InputArray(m) {}
  _OutputArray(const Mat& m) : _InputArray(m) {}   // OpenCV allows a temporary header (e.g. m.row(i)) as an output
  void create(int r, int c, int type) const { m_->create(r, c, type); }
  void create(Size s, int type) const { m_->create(s, type); }
  void release() const { if (m_) m_->release(); }
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;
typedef const _OutputArray& InputOutputArray;
inline InputArray noArray() { static _InputArray none; return none; }

inline void Mat::copyTo(const _OutputArray& dst) const {
  dst.create(rows, cols, type_);
  Mat d = dst.getMat();
  for (int y = 0; y < rows; y++) memmove(d.ptr(y), ptr(y), (size_t)cols * elemSize());
}

// ---- matrix expressions --------------------------------------------------------------------------------------------------
// OpenCV evaluates A*B+C, -A.t()*b, s*A ... lazily: operators build a cv::MatExpr that is folded into ONE gemm / scaled copy when
// assigned (modules/core/src/matop.cpp).  Which products get folded decides where intermediate results are rounded to float,
// so the folding rules are restated here (from the published OpenCV 3.4 sources; cvshim_gemm / cvshim_invert in cv_prims.cc
// hold the arithmetic and are pinned against python cv2's gemm / invert / norm in tests/test_oracle_cvmat.py):
//   Mat * Mat, with either side a transposed (A.t()) or scaled (s*A, -A, A/s) expression: one gemm(A, B, alpha, flags);
//   gemm-expression + Mat / - Mat: the addend becomes the gemm's C (beta = +-1);  s * expr, expr * s, -expr: alpha (and beta) scale;
//   -A.t(): the transpose is materialised, then scaled (MatOp::subtract(Scalar, expr) evaluates its operand);
//   any other operand that is itself an expression is evaluated to a temporary Mat first.
class MatExpr {
 public:
  enum Op { IDENT, T, ADDEX, GEMM, INV };
  MatExpr() : op(IDENT), alpha(1), beta(0), s(0), flags(0) {}
  MatExpr(const Mat& m) : op(IDENT), a(m), alpha(1), beta(0), s(0), flags(0) {}
  Mat eval() const;
  operator Mat() const { return eval(); }
  MatExpr t() const;
  MatExpr inv(int method = 0) const;
  double dot(const Mat& m) const { return eval().dot(m); }
  Mat row(int y) const { return eval().row(y); }
  Mat col(int x) const { return eval().col(x); }
  Mat rowRange(int r0, int r1) const { return eval().rowRange(r0, r1); }
  Mat colRange(int c0, int c1) const { return eval().colRange(c0, c1); }
  template <typename TT> TT at(int i) const { return eval().at<TT>(i); }
  template <typename TT> TT at(int y, int x) const { return eval().at<TT>(y, x); }
  Op op;
  Mat a, b, c;          // IDENT/T/INV: a.  ADDEX: alpha*a + beta*b.  GEMM: alpha*op(a)*op(b) + beta*c
  double alpha, beta;
  double s;             // ADDEX only: scalar addend (Mat - scalar, Frame.cc:1002)
  int flags;            // GEMM_1_T = 1, GEMM_2_T = 2, GEMM_3_T = 4
};
inline Mat::Mat(const MatExpr& e) : Mat() { *this = e.eval(); }
// Mat = MatExpr: OpenCV lets the expression write into the destination (MatOp::assign -> dst.create(), a no-op when shape and type
// already match), which is what makes `A.row(0) = s*B.row(2) - B.row(0)` (KannalaBrandt8.cpp:434-437) fill the row of A.
inline Mat& Mat::operator=(const MatExpr& e) {
  Mat m = e.eval();
  if (data && m.data != data && rows == m.rows && cols == m.cols && type_ == m.type()) {
    for (int y = 0; y < rows; y++) memmove(ptr(y), m.ptr(y), (size_t)cols * elemSize());
    return *this;
  }
  return *this = m;
}

enum { GEMM_1_T = 1, GEMM_2_T = 2, GEMM_3_T = 4, DECOMP_LU = 0 };
enum { NORM_INF = 1, NORM_L1 = 2, NORM_L2 = 4, NORM_HAMMING = 6 };

// arithmetic back ends (cv_prims.cc)
void cvshim_gemm(const Mat& a, const Mat& b, double alpha, const Mat& c, double beta, Mat& d, int flags);   // cv::gemm
void cvshim_add_weighted(const Mat& a, double alpha, const Mat& b, double beta, Mat& d);                     // MatOp_AddEx::assign
void cvshim_add_scalar(const Mat& a, double alpha, double s, Mat& d);                                         // a*alpha + s, saturated to a's type
void cvshim_transpose(const Mat& a, Mat& d);
void cvshim_invert(const Mat& a, Mat& d);                                                                    // cv::invert, DECOMP_LU
double norm(InputArray a, int normType = NORM_L2);
void hconcat(InputArray a, InputArray b, OutputArray dst);   // cv::hconcat of two matrices with equal rows and type
// cv::SVD::compute on a square CV_32F matrix: OpenCV's own one-sided Jacobi (modules/core/src/lapack.cpp JacobiSVDImpl_<float>);
// matrices this small never reach LAPACK even in a LAPACK build (hal_internal.cpp: HAL_SVD_SMALL_MATRIX_THRESH).  w and vt only.
struct SVD {
  enum { MODIFY_A = 1, NO_UV = 2, FULL_UV = 4 };
  static void compute(InputArray src, OutputArray w, OutputArray u, OutputArray vt, int flags = 0);
};
double norm(InputArray a, InputArray b, int normType = NORM_L2);

inline MatExpr cvshim_scaled(const Mat& m, double alpha) { MatExpr e; e.op = MatExpr::ADDEX; e.a = m; e.alpha = alpha; e.beta = 0; return e; }
inline bool cvshim_is_scaled(const MatExpr& e) { return e.op == MatExpr::ADDEX && (e.b.empty() || e.beta == 0) && e.s == 0; }
inline MatExpr cvshim_matmul(const MatExpr& e1, const MatExpr& e2) {   // MatOp::matmul
  MatExpr r; r.op = MatExpr::GEMM; r.alpha = 1; r.beta = 0; r.flags = 0;
  if (e1.op == MatExpr::T) { r.flags = GEMM_1_T; r.alpha = e1.alpha; r.a = e1.a; }
  else if (cvshim_is_scaled(e1)) { r.alpha = e1.alpha; r.a = e1.a; }
  else r.a = e1.eval();
  if (e2.op == MatExpr::T) { r.flags |= GEMM_2_T; r.alpha *= e2.alpha; r.b = e2.a; }
  else if (cvshim_is_scaled(e2)) { r.alpha *= e2.alpha; r.b = e2.a; }
  else r.b = e2.eval();
  return r;
}
inline MatExpr cvshim_add(const MatExpr& e1, const MatExpr& e2, double sign) {   // MatOp_GEMM::add/subtract, MatOp::add/subtract
  if (e1.op == MatExpr::GEMM && (e1.c.empty() || e1.beta == 0) && (e2.op == MatExpr::IDENT || cvshim_is_scaled(e2) || e2.op == MatExpr::T)) {
    MatExpr r = e1;
    r.c = e2.a; r.beta = sign * (e2.op == MatExpr::IDENT ? 1.0 : e2.alpha);
    if (e2.op == MatExpr::T) r.flags |= GEMM_3_T;
    return r;
  }
  if (e2.op == MatExpr::GEMM && (e2.c.empty() || e2.beta == 0) && (e1.op == MatExpr::IDENT || cvshim_is_scaled(e1) || e1.op == MatExpr::T)) {
    MatExpr r = e2;
    r.alpha = sign * e2.alpha; r.c = e1.a; r.beta = (e1.op == MatExpr::IDENT ? 1.0 : e1.alpha);
    if (e1.op == MatExpr::T) r.flags |= GEMM_3_T;
    return r;
  }
  MatExpr r; r.op = MatExpr::ADDEX;
  if (e1.op == MatExpr::IDENT) { r.a = e1.a; r.alpha = 1; } else if (cvshim_is_scaled(e1)) { r.a = e1.a; r.alpha = e1.alpha; } else { r.a = e1.eval(); r.alpha = 1; }
  if (e2.op == MatExpr::IDENT) { r.b = e2.a; r.beta = sign; } else if (cvshim_is_scaled(e2)) { r.b = e2.a; r.beta = sign * e2.alpha; } else { r.b = e2.eval(); r.beta = sign; }
  return r;
}
inline MatExpr cvshim_mul_scalar(const MatExpr& e, double s) {   // MatOp_*::multiply(e, s)
  if (e.op == MatExpr::IDENT) return cvshim_scaled(e.a, s);
  if (e.op == MatExpr::T || e.op == MatExpr::ADDEX || e.op == MatExpr::GEMM) { MatExpr r = e; r.alpha *= s; r.beta *= s; return r; }
  return cvshim_scaled(e.eval(), s);
}
inline MatExpr cvshim_neg(const MatExpr& e) {   // operator-(expr): MatOp::subtract(Scalar(0), expr) evaluates, AddEx / GEMM fold the sign
  if (e.op == MatExpr::IDENT) return cvshim_scaled(e.a, -1);
  if (e.op == MatExpr::ADDEX || e.op == MatExpr::GEMM) { MatExpr r = e; r.alpha = -r.alpha; r.beta = -r.beta; return r; }
  return cvshim_scaled(e.eval(), -1);
}

inline MatExpr operator*(const Mat& a, const Mat& b) { return cvshim_matmul(MatExpr(a), MatExpr(b)); }
inline MatExpr operator*(const MatExpr& a, const Mat& b) { return cvshim_matmul(a, MatExpr(b)); }
inline MatExpr operator*(const Mat& a, const MatExpr& b) { return cvshim_matmul(MatExpr(a), b); }
inline MatExpr operator*(const MatExpr& a, const MatExpr& b) { return cvshim_matmul(a, b); }
inline MatExpr operator*(const Mat& a, double s) { return cvshim_scaled(a, s); }
inline MatExpr operator*(double s, const Mat& a) { return cvshim_scaled(a, s); }
inline MatExpr operator*(const MatExpr& e, double s) { return cvshim_mul_scalar(e, s); }
inline MatExpr operator*(double s, const MatExpr& e) { return cvshim_mul_scalar(e, s); }
inline MatExpr operator/(const Mat& a, double s) { return cvshim_scaled(a, 1. / s); }
inline MatExpr operator/(const MatExpr& e, double s) { return cvshim_mul_scalar(e, 1. / s); }
inline MatExpr operator+(const Mat& a, const Mat& b) { return cvshim_add(MatExpr(a), MatExpr(b), 1); }
inline MatExpr operator+(const MatExpr& a, const Mat& b) { return cvshim_add(a, MatExpr(b), 1); }
inline MatExpr operator+(const Mat& a, const MatExpr& b) { return cvshim_add(MatExpr(a), b, 1); }
inline MatExpr operator+(const MatExpr& a, const MatExpr& b) { return cvshim_add(a, b, 1); }
inline MatExpr operator-(const Mat& a, const Mat& b) { return cvshim_add(MatExpr(a), MatExpr(b), -1); }
inline MatExpr operator-(const MatExpr& a, const Mat& b) { return cvshim_add(a, MatExpr(b), -1); }
inline MatExpr operator-(const Mat& a, const MatExpr& b) { return cvshim_add(MatExpr(a), b, -1); }
inline MatExpr operator-(const MatExpr& a, const MatExpr& b) { return cvshim_add(a, b, -1); }
inline MatExpr operator-(const Mat& a, double sc) { MatExpr e = cvshim_scaled(a, 1); e.s = -sc; return e; }
inline MatExpr operator+(const Mat& a, double sc) { MatExpr e = cvshim_scaled(a, 1); e.s = sc; return e; }
inline MatExpr operator-(const Mat& a) { return cvshim_scaled(a, -1); }
inline MatExpr operator-(const MatExpr& e) { return cvshim_neg(e); }

inline MatExpr Mat::t() const { MatExpr e; e.op = MatExpr::T; e.a = *this; e.alpha = 1; return e; }
inline MatExpr Mat::inv(int) const { MatExpr e; e.op = MatExpr::INV; e.a = *this; return e; }
inline MatExpr MatExpr::t() const {   // MatOp_T::transpose: (s * A.t()).t() = s * A; otherwise evaluate and transpose
  if (op == T) return alpha == 1 ? MatExpr(a) : cvshim_scaled(a, alpha);
  return eval().t();
}
inline MatExpr MatExpr::inv(int) const { return eval().inv(); }
inline Mat MatExpr::eval() const {
  Mat d;
  switch (op) {
    case IDENT: return a;
    case T:
      cvshim_transpose(a, d);
      if (alpha != 1) { Mat s; cvshim_add_weighted(d, alpha, Mat(), 0, s); return s; }
      return d;
    case ADDEX:
      if (s != 0 || ((b.empty() || beta == 0) && a.type() != CV_32F && a.type() != CV_64F)) { assert(b.empty() || beta == 0); cvshim_add_scalar(a, alpha, s, d); return d; }
      cvshim_add_weighted(a, alpha, b, beta, d);
      return d;
    case GEMM: cvshim_gemm(a, b, alpha, c, beta, d, flags); return d;
    case INV: cvshim_invert(a, d); return d;
  }
  return d;
}

enum { INTER_NEAREST = 0, INTER_LINEAR = 1, INTER_CUBIC = 2, INTER_AREA = 3 };
enum { BORDER_CONSTANT = 0, BORDER_REPLICATE = 1, BORDER_REFLECT = 2, BORDER_WRAP = 3, BORDER_REFLECT_101 = 4,
       BORDER_REFLECT101 = 4, BORDER_DEFAULT = 4, BORDER_ISOLATED = 16 };

// cv::fastAtan2 (degrees), cv::FAST (TYPE_9_16), cv::resize (INTER_LINEAR, 8UC1), cv::copyMakeBorder (REFLECT_101),
// cv::GaussianBlur (7x7, sigma 2, 8UC1 fixed point) -- cv_prims.cc.
float fastAtan2(float y, float x);
void FAST(InputArray image, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression = true);
void resize(InputArray src, OutputArray dst, Size dsize, double fx = 0, double fy = 0, int interpolation = INTER_LINEAR);
void copyMakeBorder(InputArray src, OutputArray dst, int top, int bottom, int left, int right, int borderType,
                    const Scalar& value = Scalar());
void GaussianBlur(InputArray src, OutputArray dst, Size ksize, double sigmaX, double sigmaY = 0, int borderType = BORDER_DEFAULT);

// Only ComputeKeyPointsOld (dead code in the reference, ORBextractor.cc:880-1057) names it.
struct KeyPointsFilter {
  static void retainBest(std::vector<KeyPoint>& keypoints, int npoints);
};

}  // namespace cv
