// Host-side C++ wrappers (header-only, over the C-ABI of include/orbx.h) for the steps on either side of the extractor that
// SURVEY.md §8-f ranks 2 and 3 name.  Each call replaces one OpenCV / Frame call of the reference:
//
//   cvtColor(mImGray, mImGray, CV_RGB2GRAY ...)                 src/Tracking.cc:989-1014, 1059-1073, 1121-1135
//   cv::remap(imLeft, imLeftRect, M1l, M2l, cv::INTER_LINEAR)   Examples/Stereo/stereo_euroc.cc:166-167
//   clahe->apply(im, im)                                        Examples/Monocular/mono_tum_vi.cc:109
//   Frame::UndistortKeyPoints()                                 src/Frame.cc:837-870
//   Frame::isInFrustum(pMP, 0.5) over mvpLocalMapPoints         src/Frame.cc:572-642, src/Tracking.cc:3478-3500
//
// With ORBX_WITH_OPENCV the image arguments are cv::Mat (any channel count); without it (this image has no C++ OpenCV) the
// 8-bit stand-in of cvlite.h is used for gray images and colour input is passed as pointer + step.  Nothing is computed on
// the host; a missing GPU makes the constructor throw.
#pragma once
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/orbx.h"
#include "cvlite.h"

namespace ORB_SLAM3 {

class FramePreproc {
 public:
  explicit FramePreproc(int device = 0, int grayCoeffs = ORBX_GRAY_OCV4) : coeffs_(grayCoeffs) { ck(orbx_preproc_create(device, &h_), "orbx_preproc_create"); }
  ~FramePreproc() {
    for (orbx_remap_table* t : tables_) orbx_remap_destroy(t);
    orbx_preproc_destroy(h_);
  }
  FramePreproc(const FramePreproc&) = delete;
  FramePreproc& operator=(const FramePreproc&) = delete;

  // cvtColor(src, gray, CV_{RGB|BGR}[A]2GRAY); bRGB = Tracking::mbRGB.  src and gray must not alias.
  void cvtColorToGray(const uint8_t* src, int cols, int rows, size_t step, int channels, bool bRGB, cvx::Mat& gray) {
    gray.create(rows, cols, ORBX_MAT_8U);
    ck(orbx_cvt_gray(h_, src, 1, cols, rows, (int)step, 0, channels, bRGB ? 1 : 0, coeffs_, gray.data, (int)gray.step, 0, ORBX_MEM_HOST, nullptr),
       "orbx_cvt_gray");
  }
#ifdef ORBX_WITH_OPENCV
  void cvtColorToGray(const cv::Mat& src, bool bRGB, cv::Mat& gray) {
    cv::Mat out;
    cvtColorToGray(src.data, src.cols, src.rows, src.step, src.channels(), bRGB, out);
    gray = out;  // the reference converts in place (cvtColor(mImGray, mImGray, ...)); cv::Mat reallocates there too
  }
#endif

  // The pair of CV_32F maps from cv::initUndistortRectifyMap (stereo_euroc.cc:117-118), converted once.  Returns a map id.
  int addRectifyMap(const float* mapx, const float* mapy, int cols, int rows, size_t stepFloats, int srcCols, int srcRows) {
    orbx_remap_table* t = nullptr;
    ck(orbx_remap_create(h_, mapx, mapy, cols, rows, (int)stepFloats, srcCols, srcRows, &t), "orbx_remap_create");
    tables_.push_back(t);
    dims_.push_back(cols); dims_.push_back(rows);
    return (int)tables_.size() - 1;
  }
  // cv::remap(src, dst, M1, M2, cv::INTER_LINEAR)
  void remap(const cvx::Mat& src, cvx::Mat& dst, int mapId) {
    dst.create(dims_[2 * mapId + 1], dims_[2 * mapId], ORBX_MAT_8U);
    ck(orbx_remap(h_, tables_.at(mapId), src.data, 1, (int)src.step, 0, dst.data, (int)dst.step, 0, ORBX_MEM_HOST, nullptr), "orbx_remap");
  }
  // cv::createCLAHE(clipLimit, Size(tilesX, tilesY))->apply(im, im)
  void clahe(cvx::Mat& im, double clipLimit = 3.0, int tilesX = 8, int tilesY = 8) {
    ck(orbx_clahe(h_, im.data, 1, im.cols, im.rows, (int)im.step, 0, clipLimit, tilesX, tilesY, im.data, (int)im.step, 0, ORBX_MEM_HOST, nullptr),
       "orbx_clahe");
  }

  // Frame::UndistortKeyPoints: K = {fx, fy, cx, cy}; distCoef = mDistCoef (4 or 5 floats: k1 k2 p1 p2 [k3]).
  void UndistortKeyPoints(const std::vector<cvx::KeyPoint>& mvKeys, const float K[4], const float* distCoef, int nCoef,
                          std::vector<cvx::KeyPoint>& mvKeysUn) {
    static_assert(sizeof(cvx::KeyPoint) == sizeof(orbx_keypoint), "KeyPoint layout");
    float d[5] = {0, 0, 0, 0, 0};
    for (int i = 0; i < nCoef && i < 5; ++i) d[i] = distCoef[i];
    mvKeysUn.resize(mvKeys.size());
    ck(orbx_undistort_keypoints(h_, reinterpret_cast<const orbx_keypoint*>(mvKeys.data()), (int)mvKeys.size(), K[0], K[1], K[2], K[3], d,
                                reinterpret_cast<orbx_keypoint*>(mvKeysUn.data()), ORBX_MEM_HOST, nullptr),
       "orbx_undistort_keypoints");
  }

  // Frame::isInFrustum over a batch of map points (structure-of-arrays of the MapPoint getters the reference calls:
  // GetWorldPos, GetNormal) plus the RAW MapPoint::mfMaxDistance / mfMinDistance members -- not the *Invariance getters: the
  // gate multiplies by 1.2f / 0.8f itself and PredictScale divides the raw mfMaxDistance (MapPoint.cc:555-602); the members
  // are protected in the reference, INTEGRATION.md shows the two-line accessor.  Fills the MapPoint tracking fields.
  struct FrustumResult {
    std::vector<uint8_t> mbTrackInView;
    std::vector<float> mTrackProjX, mTrackProjY, mTrackProjXR, mTrackDepth, mTrackViewCos;
    std::vector<int> mnTrackScaleLevel;
  };
  int isInFrustum(const orbx_frustum& frame, const std::vector<float>& worldPos3, const std::vector<float>& normal3,
                  const std::vector<float>& maxDistance, const std::vector<float>& minDistance, FrustumResult& r) {
    const int n = (int)maxDistance.size();
    r.mbTrackInView.resize(n); r.mTrackProjX.resize(n); r.mTrackProjY.resize(n); r.mTrackProjXR.resize(n); r.mTrackDepth.resize(n);
    r.mTrackViewCos.resize(n); r.mnTrackScaleLevel.resize(n);
    ck(orbx_is_in_frustum(h_, &frame, n, worldPos3.data(), normal3.data(), maxDistance.data(), minDistance.data(), r.mbTrackInView.data(),
                          r.mTrackProjX.data(), r.mTrackProjY.data(), r.mTrackProjXR.data(), r.mTrackDepth.data(), r.mnTrackScaleLevel.data(),
                          r.mTrackViewCos.data(), ORBX_MEM_HOST, nullptr),
       "orbx_is_in_frustum");
    int visible = 0;
    for (uint8_t v : r.mbTrackInView) visible += v;
    return visible;
  }

  // The right camera of a fisheye-stereo frame as Frame::isInFrustumChecks(pMP, limit, bRight = true) sets it up
  // (src/Frame.cc:1274-1281): mR = Rrl * mRcw, mt = Rrl * mtcw + trl, twc = mRwc * tlr + mOw, projection through mpCamera2.
  // Float products in the order of cv::gemm's small-matrix path (((a0*b0 + a1*b1) + a2*b2), then + c), so the composed pose has the
  // bits the reference's cv::Mat lines produce.  Trl / Tlr are the frame's 3x4 row-major mTrl / mTlr; cam2 = fx fy cx cy of
  // mpCamera2 (kb2 = its k1..k4 for KannalaBrandt8, null for a pinhole).  Frame::isInFrustum on such a frame (:643-660) is
  // isInFrustum(left) for mbTrackInView + mTrackProjX/Y, mnTrackScaleLevel, mTrackViewCos, mTrackDepth and isInFrustum(right)
  // for mbTrackInViewR + mTrackProjXR (= its mTrackProjX), mTrackProjYR, mnTrackScaleLevelR, mTrackViewCosR, mTrackDepthR;
  // fields of points that fail keep their previous values in the reference (the result arrays hold -1 there).
  static orbx_frustum RightCameraFrustum(const orbx_frustum& left, const float* Trl, const float* Tlr, const float* cam2, const float* kb2) {
    orbx_frustum g = left;
    auto dot3 = [](float a0, float b0, float a1, float b1, float a2, float b2) {
      volatile float s = a0 * b0; volatile float p = a1 * b1; s = s + p; p = a2 * b2; s = s + p; return (float)s; };
    for (int i = 0; i < 3; i++) {
      for (int j = 0; j < 3; j++)
        g.Rcw[3 * i + j] = dot3(Trl[4 * i], left.Rcw[j], Trl[4 * i + 1], left.Rcw[3 + j], Trl[4 * i + 2], left.Rcw[6 + j]);
      volatile float t = dot3(Trl[4 * i], left.tcw[0], Trl[4 * i + 1], left.tcw[1], Trl[4 * i + 2], left.tcw[2]);
      t = t + Trl[4 * i + 3];
      g.tcw[i] = t;
      volatile float o = dot3(left.Rcw[i], Tlr[3], left.Rcw[3 + i], Tlr[7], left.Rcw[6 + i], Tlr[11]);   // mRwc = mRcw^T
      o = o + left.Ow[i];
      g.Ow[i] = o;
    }
    g.fx = cam2[0]; g.fy = cam2[1]; g.cx = cam2[2]; g.cy = cam2[3];
    g.camera_model = kb2 ? ORBX_CAMERA_KB8 : ORBX_CAMERA_PINHOLE;
    for (int k = 0; k < 4; k++) g.kb[k] = kb2 ? kb2[k] : 0.f;
    return g;
  }

  orbx_preproc* handle() { return h_; }

 private:
  static void ck(int rc, const char* what) {
    if (rc < 0) throw std::runtime_error(std::string(what) + ": " + orbx_last_error());
  }
  orbx_preproc* h_ = nullptr;
  int coeffs_;
  std::vector<orbx_remap_table*> tables_;
  std::vector<int> dims_;
};

}  // namespace ORB_SLAM3
