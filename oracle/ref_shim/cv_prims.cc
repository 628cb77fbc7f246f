// CPU ORACLE support (test infrastructure, not product code).
// The image primitives behind the stand-in cv:: header (opencv2/core/core.hpp in this directory): thin adapters from
// cv::Mat views to the oracle's restatements in oracle/orb_oracle.cc, which tests/test_oracle_cv2.py pins bit-exact
// against python cv2 4.13 (resize on every level, FAST on full frames and ROI cells incl. order and response,
// GaussianBlur with the OCV4 taps, fastAtan2).
#include <algorithm>
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
