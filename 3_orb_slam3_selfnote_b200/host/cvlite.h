// Minimal stand-ins for the handful of OpenCV types that appear in the signatures of ORB_SLAM3::ORBextractor / ORBmatcher.
// When the real OpenCV is available define ORBX_WITH_OPENCV before including the shim headers: `cvx` then aliases `cv` and the
// shim classes are source-compatible with the reference's (include/ORBextractor.h, include/ORBmatcher.h).  Without OpenCV
// (this image has no C++ OpenCV) the stand-ins below keep the same member names and memory layout so the same shim code
// compiles and the in-repo C++ test can exercise it.
#pragma once
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

// element type code of an 8-bit single-channel matrix (OpenCV's CV_8U is a macro, so it cannot be spelled cvx::CV_8U)
#define ORBX_MAT_8U 0

#ifdef ORBX_WITH_OPENCV
#include <opencv2/core/core.hpp>
namespace cvx = cv;
#else
namespace cvx {

template <class T> struct Point_ { T x, y; Point_() : x(0), y(0) {} Point_(T x_, T y_) : x(x_), y(y_) {} };
typedef Point_<float> Point2f;
typedef Point_<int> Point;

// cv::KeyPoint: pt, size, angle, response, octave, class_id (28 bytes)
struct KeyPoint {
  Point2f pt;
  float size = 0, angle = -1, response = 0;
  int octave = 0, class_id = -1;
};
static_assert(sizeof(KeyPoint) == 28, "cv::KeyPoint layout");


// 8-bit single channel matrix with shared ownership (the subset of cv::Mat the frontend touches)
class Mat {
 public:
  int rows = 0, cols = 0;
  size_t step = 0;
  uint8_t* data = nullptr;
  Mat() {}
  Mat(int r, int c, int /*type*/) { create(r, c, ORBX_MAT_8U); }
  Mat(int r, int c, int /*type*/, void* ext, size_t step_) : rows(r), cols(c), step(step_), data((uint8_t*)ext) {}
  void create(int r, int c, int /*type*/) {
    if (r == rows && c == cols && buf_) return;
    rows = r; cols = c; step = (size_t)c;
    buf_.reset(new uint8_t[(size_t)r * c + 1], std::default_delete<uint8_t[]>());
    data = buf_.get();
  }
  void release() { buf_.reset(); data = nullptr; rows = cols = 0; step = 0; }
  bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
  int type() const { return ORBX_MAT_8U; }
  uint8_t* ptr(int r = 0) { return data + (size_t)r * step; }
  const uint8_t* ptr(int r = 0) const { return data + (size_t)r * step; }
  template <class T> T* ptr(int r = 0) { return reinterpret_cast<T*>(data + (size_t)r * step); }
  template <class T> const T* ptr(int r = 0) const { return reinterpret_cast<const T*>(data + (size_t)r * step); }
  Mat row(int r) const { Mat m; m.rows = 1; m.cols = cols; m.step = step; m.data = data + (size_t)r * step; m.buf_ = buf_; return m; }
 private:
  std::shared_ptr<uint8_t> buf_;
};
typedef const Mat& InputArray;
typedef Mat& OutputArray;

}  // namespace cvx
#endif
