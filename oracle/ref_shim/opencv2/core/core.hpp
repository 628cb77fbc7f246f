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
#define CV_32S 4
#define CV_32F 5

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

// Reference-counted 8-bit single-channel matrix header (only 8UC1 is ever created on this path).
class Mat {
 public:
  struct Step {  // converts like cv::MatStep
    size_t v;
    operator size_t() const { return v; }
  };
  Mat() : rows(0), cols(0), data(nullptr), step{0}, base_(nullptr), refs_(nullptr) {}
  Mat(int r, int c, int type) : Mat() { create(r, c, type); }
  Mat(Size s, int type) : Mat() { create(s.height, s.width, type); }
  // user-allocated data, not owned (cv::Mat(rows, cols, type, void*, step))
  Mat(int r, int c, int type, void* d, size_t st = 0) : rows(r), cols(c), data((uchar*)d), step{st ? st : (size_t)c}, base_(nullptr), refs_(nullptr) {
    (void)type;
  }
  Mat(const Mat& o) : rows(o.rows), cols(o.cols), data(o.data), step(o.step), base_(o.base_), refs_(o.refs_) { if (refs_) ++*refs_; }
  Mat& operator=(const Mat& o) {
    if (this != &o) {
      if (o.refs_) ++*o.refs_;
      release();
      rows = o.rows; cols = o.cols; data = o.data; step = o.step; base_ = o.base_; refs_ = o.refs_;
    }
    return *this;
  }
  ~Mat() { release(); }
  void release() {
    if (refs_ && --*refs_ == 0) { free(base_); free(refs_); }
    rows = cols = 0; data = nullptr; base_ = nullptr; refs_ = nullptr; step.v = 0;
  }
  // cv::Mat::create: no-op when the shape already matches (that is what lets resize() write into a ROI in place)
  void create(int r, int c, int type) {
    (void)type;
    if (data && rows == r && cols == c) return;
    release();
    rows = r; cols = c; step.v = (size_t)c;
    base_ = (uchar*)malloc((size_t)r * c + 64);   // plain malloc: pixel storage stays outside the node allocator
    refs_ = (int*)malloc(sizeof(int));
    *refs_ = 1;
    data = base_;
  }
  void create(Size s, int type) { create(s.height, s.width, type); }
  static Mat zeros(int r, int c, int type) { Mat m(r, c, type); for (int y = 0; y < r; y++) memset(m.ptr(y), 0, (size_t)c); return m; }
  Mat operator()(const Rect& r) const { Mat m(*this); m.data = data + (size_t)r.y * step.v + r.x; m.rows = r.height; m.cols = r.width; return m; }
  Mat rowRange(int a, int b) const { return (*this)(Rect(0, a, cols, b - a)); }
  Mat colRange(int a, int b) const { return (*this)(Rect(a, 0, b - a, rows)); }
  Mat row(int y) const { return (*this)(Rect(0, y, cols, 1)); }
  Mat clone() const { Mat m(rows, cols, CV_8UC1); for (int y = 0; y < rows; y++) memcpy(m.ptr(y), ptr(y), (size_t)cols); return m; }
  void copyTo(const _OutputArray& dst) const;
  bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
  int type() const { return CV_8UC1; }
  size_t step1() const { return step.v; }
  size_t total() const { return (size_t)rows * cols; }
  Size size() const { return Size(cols, rows); }
  bool isContinuous() const { return step.v == (size_t)cols; }
  uchar* ptr(int y = 0) { return data + (size_t)y * step.v; }
  const uchar* ptr(int y = 0) const { return data + (size_t)y * step.v; }
  template <typename T> T* ptr(int y = 0) { return (T*)(data + (size_t)y * step.v); }
  template <typename T> const T* ptr(int y = 0) const { return (const T*)(data + (size_t)y * step.v); }
  template <typename T> T& at(int y, int x) { return ((T*)(data + (size_t)y * step.v))[x]; }
  template <typename T> const T& at(int y, int x) const { return ((const T*)(data + (size_t)y * step.v))[x]; }

  int rows, cols;
  uchar* data;
  Step step;

 private:
  uchar* base_;
  int* refs_;
};

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
  dst.create(rows, cols, CV_8UC1);
  Mat d = dst.getMat();
  for (int y = 0; y < rows; y++) memmove(d.ptr(y), ptr(y), (size_t)cols);
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
