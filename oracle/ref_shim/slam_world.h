// CPU ORACLE support (test infrastructure, not product code).
//
// Stand-in data model for compiling the reference's src/ORBmatcher.cc UNMODIFIED (oracle/Makefile, target `ref`): this header is
// force-included (-include) ahead of the reference's own include/ORBmatcher.h and defines the include guards of Frame.h,
// KeyFrame.h, MapPoint.h and GeometricCamera.h, so that those heavyweight headers (DBoW2, Eigen, boost, IMU types) expand to
// nothing and the classes below are the ones ORBmatcher.cc sees.  Each class holds exactly the members ORBmatcher.cc touches, with
// the reference's names and types (include/Frame.h, include/KeyFrame.h, include/MapPoint.h, include/CameraModels/*.h).
// Member functions with arithmetic on the path -- Frame::GetFeaturesInArea / PosInGrid / AssignFeaturesToGrid / isInFrustum /
// isInFrustumChecks / ComputeStereoMatches, KeyFrame::GetFeaturesInArea / IsInImage, MapPoint::PredictScale /
// Get{Min,Max}DistanceInvariance, Pinhole::project, KannalaBrandt8::project -- are NOT restated here: oracle/Makefile cuts their
// bodies out of the reference's .cc files at build time (oracle/_ref/*_slice.inc) and ref_match_harness.cc includes them.
// What IS written here: plain accessors (GetRotation ...), MapPoint / KeyFrame bookkeeping reduced to an event log, and the
// two camera entry points that would need cv::solve / cv::SVD (Pinhole::epipolarConstrain takes the caller's F12 instead of
// rebuilding it from R12 / t12; matchAndtriangulate, which nothing calls, is not available).
#pragma once
#define FRAME_H
#define KEYFRAME_H
#define MAPPOINT_H
#define CAMERAMODELS_GEOMETRICCAMERA_H
#define __D_T_FEATURE_VECTOR__

#include <climits>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <list>
#include <map>
#include <mutex>
#include <set>
#include <tuple>
#include <vector>

#include <opencv2/core/core.hpp>

#ifndef SLAM_WORLD_NO_USING_STD
using namespace std;   // the reference's headers rely on it (include/ORBmatcher.h:75 names `pair` / `vector` unqualified)
#endif

namespace DBoW2 {
typedef unsigned int NodeId;
class FeatureVector : public std::map<NodeId, std::vector<unsigned int> > {};   // Thirdparty/DBoW2/DBoW2/FeatureVector.h:22-24
}  // namespace DBoW2

namespace ORB_SLAM3 {

#define FRAME_GRID_ROWS 48
#define FRAME_GRID_COLS 64

class MapPoint;
class KeyFrame;
class Frame;
class ORBextractor;

class GeometricCamera {
 public:
  GeometricCamera() {}
  explicit GeometricCamera(const std::vector<float>& p) : mvParameters(p) {}
  virtual ~GeometricCamera() {}
  virtual cv::Point2f project(const cv::Point3f& p3D) = 0;
  virtual cv::Point2f project(const cv::Mat& m3D) = 0;
  virtual cv::Mat unprojectMat(const cv::Point2f&) {
    fprintf(stderr, "oracle/_ref: unprojectMat is only needed for KannalaBrandt8 cameras\n");
    abort();
  }
  virtual bool epipolarConstrain(GeometricCamera* otherCamera, const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, const cv::Mat& R12,
                                 const cv::Mat& t12, const float sigmaLevel, const float unc) = 0;
  virtual bool matchAndtriangulate(const cv::KeyPoint&, const cv::KeyPoint&, GeometricCamera*, cv::Mat&, cv::Mat&, const float, const float, cv::Mat&) {
    fprintf(stderr, "oracle/_ref: matchAndtriangulate needs cv::SVD (KannalaBrandt8.cpp:249-341), not available in the stand-in\n");
    abort();
  }
  // include/CameraModels/GeometricCamera.h:81-88
  float getParameter(const int i) { return mvParameters[i]; }
  size_t size() { return mvParameters.size(); }
  unsigned int GetType() { return mnType; }
  const unsigned int CAM_PINHOLE = 0;
  const unsigned int CAM_FISHEYE = 1;
  std::vector<float> mvParameters;
  unsigned int mnType = 0;
};

class Pinhole : public GeometricCamera {
 public:
  explicit Pinhole(const std::vector<float>& p) : GeometricCamera(p) { mnType = CAM_PINHOLE; }
  cv::Point2f project(const cv::Point3f& p3D);   // body: src/CameraModels/Pinhole.cpp:46-49 (slice)
  cv::Point2f project(const cv::Mat& m3D);       // body: src/CameraModels/Pinhole.cpp:52-56 (slice)
  // Pinhole.cpp:143-165 with the fundamental matrix supplied by the harness (see the header comment): the line test itself.
  bool epipolarConstrain(GeometricCamera*, const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, const cv::Mat&, const cv::Mat&, const float,
                         const float unc) {
    const cv::Mat& F12 = F12given;
    const float a = kp1.pt.x * F12.at<float>(0, 0) + kp1.pt.y * F12.at<float>(1, 0) + F12.at<float>(2, 0);
    const float b = kp1.pt.x * F12.at<float>(0, 1) + kp1.pt.y * F12.at<float>(1, 1) + F12.at<float>(2, 1);
    const float c = kp1.pt.x * F12.at<float>(0, 2) + kp1.pt.y * F12.at<float>(1, 2) + F12.at<float>(2, 2);
    const float num = a * kp2.pt.x + b * kp2.pt.y + c;
    const float den = a * a + b * b;
    if (den == 0) return false;
    const float dsqr = num * num / den;
    return dsqr < 3.84 * unc;
  }
  cv::Mat F12given;
};

class KannalaBrandt8 : public GeometricCamera {
 public:
  explicit KannalaBrandt8(const std::vector<float>& p) : GeometricCamera(p) { mnType = CAM_FISHEYE; }
  cv::Point2f project(const cv::Point3f& p3D);   // body: src/CameraModels/KannalaBrandt8.cpp:29-45 (slice)
  cv::Point2f project(const cv::Mat& m3D);       // body: src/CameraModels/KannalaBrandt8.cpp:47-51 (slice)
  // bodies: src/CameraModels/KannalaBrandt8.cpp:103-137 (unprojectMat / unproject), :244-247 (epipolarConstrain), :343-418
  // (TriangulateMatches), :430-444 (Triangulate), cut by oracle/Makefile into camera_slice.inc; cv::SVD::compute behind Triangulate is
  // OpenCV's one-sided Jacobi restated in cv_prims.cc and pinned to cv2.SVDecomp (tests/test_oracle_cvmat.py)
  cv::Mat unprojectMat(const cv::Point2f& p2D);
  cv::Point3f unproject(const cv::Point2f& p2D);
  bool epipolarConstrain(GeometricCamera* pCamera2, const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, const cv::Mat& R12, const cv::Mat& t12,
                         const float sigmaLevel, const float unc);
  float TriangulateMatches(GeometricCamera* pCamera2, const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, const cv::Mat& R12, const cv::Mat& t12,
                           const float sigmaLevel, const float unc, cv::Mat& p3D);
 private:
  const float precision = 1e-6;   // include/CameraModels/KannalaBrandt8.h:47-52
  void Triangulate(const cv::Point2f& p1, const cv::Point2f& p2, const cv::Mat& Tcw1, const cv::Mat& Tcw2, cv::Mat& x3D);
};

// One entry per MapPoint / KeyFrame mutation ORBmatcher.cc performs (Fuse): kind 0 = pMP->Replace(pOther), 1 = AddObservation +
// AddMapPoint(pMP -> pKF, idx).
struct WorldEvent { int kind; MapPoint* a; MapPoint* b; KeyFrame* kf; int idx; };

class MapPoint {
 public:
  MapPoint() : mnId(0), mbBad(false), nObs(0), mfMinDistance(0), mfMaxDistance(0), mbTrackInView(false), mbTrackInViewR(false), mTrackProjX(0),
               mTrackProjY(0), mTrackDepth(0), mTrackDepthR(0), mTrackProjXR(0), mTrackProjYR(0), mnTrackScaleLevel(0), mnTrackScaleLevelR(0),
               mTrackViewCos(0), mTrackViewCosR(0), events(nullptr) {}
  cv::Mat GetWorldPos() { return mWorldPos.clone(); }     // MapPoint.cc: returns a clone
  cv::Mat GetNormal() { return mNormalVector.clone(); }
  cv::Mat GetDescriptor() { return mDescriptor.clone(); }
  bool isBad() { return mbBad; }
  int Observations() { return nObs; }
  float GetMinDistanceInvariance();   // body: src/MapPoint.cc:555-559 (slice)
  float GetMaxDistanceInvariance();   // body: src/MapPoint.cc:561-565 (slice)
  int PredictScale(const float& currentDist, KeyFrame* pKF);   // body: src/MapPoint.cc:570-585 (slice)
  int PredictScale(const float& currentDist, Frame* pF);       // body: src/MapPoint.cc:587-602 (slice)
  void ComputeDistinctiveDescriptors();                        // body: src/MapPoint.cc:350-436 (slice)
  bool IsInKeyFrame(KeyFrame* pKF) { return mObservations.count(pKF) != 0; }
  std::tuple<int, int> GetIndexInKeyFrame(KeyFrame* pKF) { return mObservations.count(pKF) ? mObservations[pKF] : std::tuple<int, int>(-1, -1); }
  void Replace(MapPoint* pMP) { if (events) events->push_back(WorldEvent{0, this, pMP, nullptr, -1}); }
  void AddObservation(KeyFrame* pKF, int idx) { if (events) events->push_back(WorldEvent{1, this, nullptr, pKF, idx}); }

  long unsigned int mnId;
  bool mbBad;
  int nObs;
  cv::Mat mWorldPos, mNormalVector, mDescriptor;
  float mfMinDistance, mfMaxDistance;
  std::mutex mMutexPos, mMutexFeatures;   // the sliced bodies lock them
  std::map<KeyFrame*, std::tuple<int, int> > mObservations;
  // tracking fields written by Frame::isInFrustum
  bool mbTrackInView, mbTrackInViewR;
  float mTrackProjX, mTrackProjY, mTrackDepth, mTrackDepthR, mTrackProjXR, mTrackProjYR;
  int mnTrackScaleLevel, mnTrackScaleLevelR;
  float mTrackViewCos, mTrackViewCosR;
  std::vector<WorldEvent>* events;   // (not named `log`: the sliced PredictScale bodies call log())
};

class KeyFrame {
 public:
  KeyFrame() : mnId(0), N(0), NLeft(-1), NRight(-1), fx(0), fy(0), cx(0), cy(0), invfx(0), invfy(0), mbf(0), mb(0), mnScaleLevels(0), mfScaleFactor(0),
               mfLogScaleFactor(0), mnMinX(0), mnMinY(0), mnMaxX(0), mnMaxY(0), mnGridCols(FRAME_GRID_COLS), mnGridRows(FRAME_GRID_ROWS),
               mfGridElementWidthInv(0), mfGridElementHeightInv(0), mpCamera(nullptr), mpCamera2(nullptr) {}
  cv::Mat GetPose() { return Tcw.clone(); }
  cv::Mat GetRotation() { return Tcw.rowRange(0, 3).colRange(0, 3).clone(); }
  cv::Mat GetTranslation() { return Tcw.rowRange(0, 3).col(3).clone(); }
  cv::Mat GetCameraCenter() { return Ow.clone(); }
  cv::Mat GetRightPose() { return TcwRight.clone(); }
  cv::Mat GetRightRotation() { return TcwRight.rowRange(0, 3).colRange(0, 3).clone(); }
  cv::Mat GetRightTranslation() { return TcwRight.rowRange(0, 3).col(3).clone(); }
  cv::Mat GetRightCameraCenter() { return OwRight.clone(); }
  std::vector<MapPoint*> GetMapPointMatches() { return mvpMapPoints; }
  std::set<MapPoint*> GetMapPoints() { std::set<MapPoint*> s; for (MapPoint* p : mvpMapPoints) if (p && !p->isBad()) s.insert(p); return s; }
  MapPoint* GetMapPoint(const size_t& idx) { return mvpMapPoints[idx]; }
  void AddMapPoint(MapPoint* pMP, const size_t& idx) { mvpMapPoints[idx] = pMP; }
  std::vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r, const bool bRight = false) const;   // src/KeyFrame.cc:798-842 (slice)
  bool IsInImage(const float& x, const float& y) const;                                                                      // src/KeyFrame.cc:844-847 (slice)
  bool isBad() { return mbBad; }

  long unsigned int mnId;
  bool mbBad = false;
  int N, NLeft, NRight;
  float fx, fy, cx, cy, invfx, invfy, mbf, mb;
  std::vector<cv::KeyPoint> mvKeys, mvKeysUn, mvKeysRight;
  std::vector<float> mvuRight, mvDepth;
  cv::Mat mDescriptors;
  DBoW2::FeatureVector mFeatVec;
  int mnScaleLevels;
  float mfScaleFactor, mfLogScaleFactor;
  std::vector<float> mvScaleFactors, mvLevelSigma2, mvInvLevelSigma2;
  int mnMinX, mnMinY, mnMaxX, mnMaxY;   // include/KeyFrame.h: const int
  int mnGridCols, mnGridRows;
  float mfGridElementWidthInv, mfGridElementHeightInv;
  std::vector<std::vector<std::vector<size_t> > > mGrid, mGridRight;
  GeometricCamera *mpCamera, *mpCamera2;
  cv::Mat Tcw, Ow, TcwRight, OwRight;
  std::vector<MapPoint*> mvpMapPoints;
};

class Frame {
 public:
  Frame() : N(0), Nleft(-1), Nright(-1), fx(0), fy(0), cx(0), cy(0), invfx(0), invfy(0), mbf(0), mb(0), mThDepth(0), mnScaleLevels(0), mfScaleFactor(0),
            mfLogScaleFactor(0), mnMinX(0), mnMaxX(0), mnMinY(0), mnMaxY(0), mfGridElementWidthInv(0), mfGridElementHeightInv(0), mpCamera(nullptr),
            mpCamera2(nullptr), mpORBextractorLeft(nullptr), mpORBextractorRight(nullptr) {}
  void UpdatePoseMatrices();                                                         // src/Frame.cc:533-539 (slice)
  void AssignFeaturesToGrid();                                                       // src/Frame.cc:434-465 (slice)
  bool PosInGrid(const cv::KeyPoint& kp, int& posX, int& posY);                      // src/Frame.cc:815-825 (slice)
  std::vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r, const int minLevel = -1, const int maxLevel = -1,
                                   const bool bRight = false) const;                 // src/Frame.cc:744-813 (slice)
  bool isInFrustum(MapPoint* pMP, float viewingCosLimit);                            // src/Frame.cc:572-661 (slice)
  bool isInFrustumChecks(MapPoint* pMP, float viewingCosLimit, bool bRight = false); // src/Frame.cc:1270-1343 (slice)
  void ComputeStereoMatches();                                                       // src/Frame.cc:901-1079 (slice)

  int N, Nleft, Nright;
  float fx, fy, cx, cy, invfx, invfy, mbf, mb, mThDepth;
  std::vector<cv::KeyPoint> mvKeys, mvKeysRight, mvKeysUn;
  std::vector<float> mvuRight, mvDepth;
  cv::Mat mDescriptors, mDescriptorsRight;
  DBoW2::FeatureVector mFeatVec;
  std::vector<MapPoint*> mvpMapPoints;
  std::vector<bool> mvbOutlier;
  std::vector<int> mvLeftToRightMatch, mvRightToLeftMatch;
  int mnScaleLevels;
  float mfScaleFactor, mfLogScaleFactor;
  std::vector<float> mvScaleFactors, mvInvScaleFactors, mvLevelSigma2, mvInvLevelSigma2;
  float mnMinX, mnMaxX, mnMinY, mnMaxY;   // static in the reference (one camera per process); per object here
  float mfGridElementWidthInv, mfGridElementHeightInv;
  std::vector<std::size_t> mGrid[FRAME_GRID_COLS][FRAME_GRID_ROWS];
  std::vector<std::size_t> mGridRight[FRAME_GRID_COLS][FRAME_GRID_ROWS];
  GeometricCamera *mpCamera, *mpCamera2;
  cv::Mat mTcw, mRcw, mtcw, mRwc, mOw;   // pose and its pieces (Frame::UpdatePoseMatrices)
  cv::Mat mTlr, mRlr, mtlr, mTrl;        // left-right rig transforms (fisheye stereo)
  ORBextractor *mpORBextractorLeft, *mpORBextractorRight;
};

}  // namespace ORB_SLAM3
