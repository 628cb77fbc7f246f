// Drop-in replacement of the reference's include/ORBextractor.h (class ORB_SLAM3::ORBextractor, lines 43-109 of
// wangyunpeng139/3_ORB_SLAM3_SelfNote): same constructor, operator(), getters and the public mvImagePyramid member, so
// Frame::ExtractORB (src/Frame.cc:469-483), the Frame constructors (src/Frame.cc:108-114) and ComputeStereoMatches
// (src/Frame.cc:908,998,1012) compile unchanged against it.  All computation goes through the C-ABI of include/orbx.h
// (liborbx.so, CUDA sm_100a); nothing is computed on the host.
//
// Define ORBX_WITH_OPENCV to compile against the real OpenCV (cv::Mat / cv::KeyPoint / cv::InputArray); otherwise the
// layout-compatible stand-ins of cvlite.h are used (this image has no C++ OpenCV).
//
// Differences from the reference, all outside the observable contract of the path:
//   * mvImagePyramid holds tight copies of the levels (no 19-px replicated frame around them; nothing reads it).  It is a live
//     member: the levels are downloaded on first access after a call (operator[] / begin() / the vector conversion sync by
//     themselves), so Frame::ComputeStereoMatches' `mpORBextractorLeft->mvImagePyramid[kpL.octave].rowRange(..)` (src/Frame.cc:
//     908, 998, 1012) reads current pixels without any change at the call site; a caller that never touches the member never pays
//     the copy -- the GPU stereo matcher (host/ORBmatcher.h) reads the levels in HBM instead.
//   * the blur taps default to OpenCV 3.4.6's (ORBX_TAPS_OCV346); pass ORBX_TAPS_OCV4 to match an OpenCV >= 3.4.9 build.
#pragma once
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/orbx.h"
#include "cvlite.h"

namespace ORB_SLAM3 {

namespace orbx_shim {
#ifdef ORBX_WITH_OPENCV
inline cv::Mat as_mat(cv::InputArray a) { return a.getMat(); }
inline cv::Mat create(cv::OutputArray a, int rows, int cols) { a.create(rows, cols, ORBX_MAT_8U); return a.getMat(); }
inline void release(cv::OutputArray a) { a.release(); }
#else
inline cvx::Mat as_mat(cvx::InputArray a) { return a; }
inline cvx::Mat& create(cvx::OutputArray a, int rows, int cols) { a.create(rows, cols, ORBX_MAT_8U); return a; }
inline void release(cvx::OutputArray a) { a.release(); }
#endif
inline void check(int rc, const char* what) {
  if (rc < 0 && rc != ORBX_EMPTY_IMAGE) throw std::runtime_error(std::string(what) + ": " + orbx_last_error());
}
}  // namespace orbx_shim

class ORBextractor {
 public:
  enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

  ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST, int blurTaps = ORBX_TAPS_OCV346,
               int device = 0)
      : nfeatures(nfeatures), scaleFactor(scaleFactor), nlevels(nlevels), iniThFAST(iniThFAST), minThFAST(minThFAST) {
    orbx_params p;
    p.nfeatures = nfeatures; p.scale_factor = scaleFactor; p.nlevels = nlevels; p.ini_th_fast = iniThFAST; p.min_th_fast = minThFAST;
    p.blur_taps = blurTaps; p.max_batch = 1; p.device = device;
    orbx_shim::check(orbx_create(&p, &h_), "orbx_create");
    mvScaleFactor.resize(nlevels); mvInvScaleFactor.resize(nlevels); mvLevelSigma2.resize(nlevels); mvInvLevelSigma2.resize(nlevels);
    orbx_get_scale_tables(h_, mvScaleFactor.data(), mvInvScaleFactor.data(), mvLevelSigma2.data(), mvInvLevelSigma2.data());
    mnFeaturesPerLevel.resize(nlevels);
    orbx_get_features_per_level(h_, mnFeaturesPerLevel.data());
    mvImagePyramid.attach(this, nlevels);
  }
  ~ORBextractor() { orbx_destroy(h_); }
  ORBextractor(const ORBextractor&) = delete;
  ORBextractor& operator=(const ORBextractor&) = delete;

  // Compute the ORB features and descriptors on an image.  Mask is ignored (as in the reference, ORBextractor.h:56).
  int operator()(cvx::InputArray _image, cvx::InputArray /*_mask*/, std::vector<cvx::KeyPoint>& _keypoints, cvx::OutputArray _descriptors,
                 std::vector<int>& vLappingArea) {
    auto image = orbx_shim::as_mat(_image);
    if (image.empty()) return -1;  // src/ORBextractor.cc:1075-1076
    const int cap = orbx_max_keypoints(h_, image.cols, image.rows);
    orbx_shim::check(cap, "orbx_max_keypoints");
    kpBuf_.resize(cap);
    descBuf_.resize((size_t)cap * 32);
    int n = 0, mono = 0;
    const int rc = orbx_extract(h_, image.ptr(0), image.cols, image.rows, (int)image.step, vLappingArea[0], vLappingArea[1],
                                kpBuf_.data(), descBuf_.data(), cap, &n, &mono);
    if (rc == ORBX_EMPTY_IMAGE) return -1;
    orbx_shim::check(rc, "orbx_extract");
    static_assert(sizeof(cvx::KeyPoint) == sizeof(orbx_keypoint), "cv::KeyPoint must be 28 bytes");
    _keypoints.resize(n);
    if (n > 0) memcpy(static_cast<void*>(_keypoints.data()), kpBuf_.data(), (size_t)n * sizeof(orbx_keypoint));
    if (n == 0) orbx_shim::release(_descriptors);  // src/ORBextractor.cc:1100-1102
    else {
      auto d = orbx_shim::create(_descriptors, n, 32);
      for (int i = 0; i < n; i++) memcpy(d.ptr(i), descBuf_.data() + (size_t)i * 32, 32);
    }
    pyramidStale_ = true;
    return mono;
  }

  int inline GetLevels() { return nlevels; }
  float inline GetScaleFactor() { return (float)scaleFactor; }
  std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }
  std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }
  std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }
  std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

  // Host copies of the un-blurred levels of the last call (the reference's public member std::vector<cv::Mat> mvImagePyramid,
  // ORBextractor.h:84), fetched from the device on first access after a call.
  class Pyramid {
   public:
    typedef std::vector<cvx::Mat>::iterator iterator;
    cvx::Mat& operator[](size_t l) { sync(); return levels_[l]; }
    cvx::Mat& at(size_t l) { sync(); return levels_.at(l); }
    size_t size() const { return levels_.size(); }
    bool empty() const { return levels_.empty(); }
    iterator begin() { sync(); return levels_.begin(); }
    iterator end() { sync(); return levels_.end(); }
    operator std::vector<cvx::Mat>&() { sync(); return levels_; }
   private:
    friend class ORBextractor;
    void attach(ORBextractor* o, int n) { owner_ = o; levels_.resize(n); }
    void sync() { owner_->SyncPyramid(); }
    ORBextractor* owner_ = nullptr;
    std::vector<cvx::Mat> levels_;
  };
  Pyramid mvImagePyramid;
  void SyncPyramid() {   // explicit form of what the accessors above do
    if (!pyramidStale_) return;
    pyramidStale_ = false;
    for (int l = 0; l < nlevels; l++) {
      int w = 0, hgt = 0;
      orbx_shim::check(orbx_pyramid_level(h_, 0, l, nullptr, 0, &w, &hgt), "orbx_pyramid_level");
      cvx::Mat& m = mvImagePyramid.levels_[l];
      m.create(hgt, w, ORBX_MAT_8U);
      orbx_shim::check(orbx_pyramid_level(h_, 0, l, m.ptr(0), (int)m.step, &w, &hgt), "orbx_pyramid_level");
    }
  }

  orbx_extractor* handle() { return h_; }  // for host/ORBmatcher.h (device-resident stereo matching)

 protected:
  int nfeatures;
  double scaleFactor;
  int nlevels;
  int iniThFAST;
  int minThFAST;
  std::vector<int> mnFeaturesPerLevel;
  std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;

 private:
  orbx_extractor* h_ = nullptr;
  std::vector<orbx_keypoint> kpBuf_;
  std::vector<uint8_t> descBuf_;
  bool pyramidStale_ = false;
};

}  // namespace ORB_SLAM3
