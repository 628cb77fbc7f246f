// Drop-in counterpart of the reference's include/ORBmatcher.h (class ORB_SLAM3::ORBmatcher, lines 39-107): constructor,
// the three constants, static DescriptorDistance and the data-parallel search entry points, computed on the GPU through
// the C-ABI of include/orbx.h.  The search methods are templates over the frame / map-point types and read exactly the
// members the reference's bodies read (F.N, F.mvKeysUn, F.mDescriptors, F.mvuRight, F.mvScaleFactors, F.mnMinX.., F.Nleft,
// F.mvpMapPoints, pMP->mbTrackInView, ...), so they bind to the reference's own Frame / MapPoint classes unchanged and to
// the small mocks of tests/cpp/shim_test.cc alike.
// With ORBX_WITH_OPENCV (real OpenCV, i.e. the drop-in build) the overloads that start from poses and 3-D map points carry the
// reference's exact parameter lists -- SearchByProjection(Frame&, const Frame&, th, bMono), (Frame&, KeyFrame*, set, th, ORBdist),
// both (KeyFrame*, cv::Mat Scw, ...), Fuse x2, SearchBySim3, SearchForTriangulation(.., cv::Mat F12, ..): the 3x3 pose algebra in
// front of each loop is the reference's own cv::Mat lines (a few per function), everything per point -- transform, projection,
// bounds / range / viewing-angle tests, PredictScale, radius -- runs on the device (orbx_search_projected_window / _best).
// The *Projected forms below them take already-projected points and work without OpenCV.
#pragma once
#include <climits>
#include <cmath>
#include <cstring>
#include <set>
#include <stdexcept>
#include <tuple>
#include <utility>
#include <vector>

#include "../../include/orbx.h"
#include "ORBextractor.h"
#include "cvlite.h"

namespace ORB_SLAM3 {

class ORBmatcher {
 public:
  static const int TH_LOW = 50;        // src/ORBmatcher.cc:37
  static const int TH_HIGH = 100;      // src/ORBmatcher.cc:36
  static const int HISTO_LENGTH = 30;  // src/ORBmatcher.cc:38

  ORBmatcher(float nnratio = 0.6, bool checkOri = true, int device = 0) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {
    orbx_shim::check(orbx_matcher_create(device, &m_), "orbx_matcher_create");
  }
  ~ORBmatcher() { orbx_matcher_destroy(m_); }
  ORBmatcher(const ORBmatcher&) = delete;
  ORBmatcher& operator=(const ORBmatcher&) = delete;

  // Computes the Hamming distance between two ORB descriptors (src/ORBmatcher.cc:2463-2483)
  static int DescriptorDistance(const cvx::Mat& a, const cvx::Mat& b) { return orbx_hamming(a.ptr(0), b.ptr(0)); }

  // Search matches between Frame keypoints and projected MapPoints (local-map tracking), src/ORBmatcher.cc:44-211: monocular /
  // rectified-stereo frames (F.Nleft == -1, :63-141) and fisheye-stereo frames (both camera passes, :63-211).
  template <class FrameT, class MapPointT>
  int SearchByProjection(FrameT& F, const std::vector<MapPointT*>& vpMapPoints, const float th = 3, const bool bFarPoints = false,
                         const float thFarPoints = 50.0f) {
    if (F.Nleft != -1) return SearchByProjectionTwoCameras(F, vpMapPoints, th, bFarPoints, thFarPoints);
    const int nmp = (int)vpMapPoints.size();
    std::vector<uint8_t> inView(nmp), observed(nmp), desc((size_t)nmp * 32);
    std::vector<float> px(nmp), py(nmp), vc(nmp), pxr(nmp);
    std::vector<int> lvl(nmp);
    for (int i = 0; i < nmp; i++) {  // the gating of src/ORBmatcher.cc:52-62
      MapPointT* pMP = vpMapPoints[i];
      bool ok = pMP->mbTrackInView;
      if (ok && bFarPoints && pMP->mTrackDepth > thFarPoints) ok = false;
      if (ok && pMP->isBad()) ok = false;
      inView[i] = ok;
      if (!ok) continue;
      px[i] = pMP->mTrackProjX; py[i] = pMP->mTrackProjY; pxr[i] = pMP->mTrackProjXR;
      lvl[i] = pMP->mnTrackScaleLevel; vc[i] = pMP->mTrackViewCos;
      observed[i] = pMP->Observations() > 0;
      const cvx::Mat d = pMP->GetDescriptor();
      memcpy(&desc[(size_t)i * 32], d.ptr(0), 32);
    }
    FrameViewHolder fv(F);
    std::vector<int> kpMp(F.N);
    std::vector<MapPointT*> pre(F.mvpMapPoints.begin(), F.mvpMapPoints.end());
    for (int k = 0; k < F.N; k++) kpMp[k] = !pre[k] ? -1 : (pre[k]->Observations() > 0 ? -2 : -3);
    int nmatches = 0;
    orbx_shim::check(orbx_search_by_projection_mappoints(m_, &fv.v, nmp, inView.data(), px.data(), py.data(), lvl.data(), vc.data(), desc.data(),
                                                         observed.data(), pxr.data(), th, mfNNratio, kpMp.data(), &nmatches),
                     "orbx_search_by_projection_mappoints");
    for (int k = 0; k < F.N; k++)
      if (kpMp[k] >= 0) F.mvpMapPoints[k] = vpMapPoints[kpMp[k]];
    return nmatches;
  }

  // The fisheye-stereo form of the search above (F.Nleft != -1): left pass over mvKeys / the left grid, right pass over mvKeysRight /
  // mGridRight, stereo partners through mvLeftToRightMatch / mvRightToLeftMatch (orbx_search_by_projection_mappoints_stereo).
  template <class FrameT, class MapPointT>
  int SearchByProjectionTwoCameras(FrameT& F, const std::vector<MapPointT*>& vpMapPoints, const float th, const bool bFarPoints,
                                   const float thFarPoints) {
    const int nmp = (int)vpMapPoints.size(), nl = F.Nleft, nr = F.N - F.Nleft;
    std::vector<uint8_t> inL(nmp), inR(nmp), observed(nmp), desc((size_t)nmp * 32);
    std::vector<float> pxL(nmp), pyL(nmp), vcL(nmp), pxR(nmp), pyR(nmp), vcR(nmp);
    std::vector<int> lvL(nmp), lvR(nmp);
    for (int i = 0; i < nmp; i++) {  // the gating of src/ORBmatcher.cc:52-62
      MapPointT* pMP = vpMapPoints[i];
      bool ok = pMP->mbTrackInView || pMP->mbTrackInViewR;
      if (ok && bFarPoints && pMP->mTrackDepth > thFarPoints) ok = false;
      if (ok && pMP->isBad()) ok = false;
      if (!ok) continue;
      inL[i] = pMP->mbTrackInView; inR[i] = pMP->mbTrackInViewR;
      pxL[i] = pMP->mTrackProjX; pyL[i] = pMP->mTrackProjY; lvL[i] = pMP->mnTrackScaleLevel; vcL[i] = pMP->mTrackViewCos;
      pxR[i] = pMP->mTrackProjXR; pyR[i] = pMP->mTrackProjYR; lvR[i] = pMP->mnTrackScaleLevelR; vcR[i] = pMP->mTrackViewCosR;
      observed[i] = pMP->Observations() > 0;
      const cvx::Mat d = pMP->GetDescriptor();
      memcpy(&desc[(size_t)i * 32], d.ptr(0), 32);
    }
    DescHolder dd(F.mDescriptors, F.N);
    orbx_frame_view L, R;
    L.n = nl; L.keypoints = reinterpret_cast<const orbx_keypoint*>(F.mvKeys.data()); L.descriptors = dd.p; L.u_right = nullptr;
    L.scale_factors = F.mvScaleFactors.data(); L.nlevels = (int)F.mvScaleFactors.size();
    L.min_x = F.mnMinX; L.min_y = F.mnMinY; L.max_x = F.mnMaxX; L.max_y = F.mnMaxY;
    R = L;
    R.n = nr; R.keypoints = reinterpret_cast<const orbx_keypoint*>(F.mvKeysRight.data()); R.descriptors = dd.p + (size_t)nl * 32;
    const orbx_mappoint_track tL = {inL.data(), pxL.data(), pyL.data(), lvL.data(), vcL.data()};
    const orbx_mappoint_track tR = {inR.data(), pxR.data(), pyR.data(), lvR.data(), vcR.data()};
    std::vector<int> kpMp(F.N);
    for (int k = 0; k < F.N; k++) kpMp[k] = !F.mvpMapPoints[k] ? -1 : (F.mvpMapPoints[k]->Observations() > 0 ? -2 : -3);
    int nmatches = 0;
    orbx_shim::check(orbx_search_by_projection_mappoints_stereo(m_, &L, &R, nmp, &tL, &tR, desc.data(), observed.data(),
                                                                F.mvLeftToRightMatch.data(), F.mvRightToLeftMatch.data(), th, mfNNratio,
                                                                kpMp.data(), &nmatches),
                     "orbx_search_by_projection_mappoints_stereo");
    for (int k = 0; k < F.N; k++)
      if (kpMp[k] >= 0) F.mvpMapPoints[k] = vpMapPoints[kpMp[k]];
    return nmatches;
  }

  // Matching for the Map Initialization (only used in the monocular case), src/ORBmatcher.cc:722-837.
  template <class FrameT>
  int SearchForInitialization(FrameT& F1, FrameT& F2, std::vector<cvx::Point2f>& vbPrevMatched, std::vector<int>& vnMatches12,
                              int windowSize = 10) {
    FrameViewHolder a(F1), b(F2);
    vnMatches12.assign(F1.mvKeysUn.size(), -1);
    int nmatches = 0;
    static_assert(sizeof(cvx::Point2f) == 8, "Point2f layout");
    orbx_shim::check(orbx_search_for_initialization(m_, &a.v, &b.v, reinterpret_cast<float*>(vbPrevMatched.data()), windowSize, mfNNratio,
                                                    mbCheckOrientation ? 1 : 0, vnMatches12.data(), &nmatches),
                     "orbx_search_for_initialization");
    return nmatches;
  }

  // The search loop shared by SearchByProjection(Frame& Cur, const Frame& Last, th, bMono) (src/ORBmatcher.cc:2027-2289) and
  // SearchByProjection(Frame&, KeyFrame*, sAlreadyFound, th, ORBdist) (:2291-2413) once the caller has projected the points
  // (valid / u / v / level / ur / angle / descriptors per source point; see include/orbx.h).  kpMp: -1 none, -2 blocking, -3 not.
  template <class FrameT>
  int SearchByProjectionProjected(FrameT& F, const std::vector<uint8_t>& valid, const std::vector<float>& u, const std::vector<float>& v,
                                  const std::vector<int>& level, const std::vector<float>* ur, const std::vector<float>& angle,
                                  const std::vector<uint8_t>& srcDesc, const std::vector<uint8_t>* observed, float th, int mode, int distTh,
                                  bool anyMpBlocks, std::vector<int>& kpMp) {
    FrameViewHolder fv(F);
    int nmatches = 0;
    orbx_shim::check(orbx_search_by_projection_window(m_, &fv.v, (int)valid.size(), valid.data(), u.data(), v.data(), level.data(),
                                                      ur ? ur->data() : nullptr, angle.data(), srcDesc.data(), observed ? observed->data() : nullptr,
                                                      th, mode, distTh, anyMpBlocks ? 1 : 0, mbCheckOrientation ? 1 : 0, kpMp.data(), &nmatches),
                     "orbx_search_by_projection_window");
    return nmatches;
  }

#ifdef ORBX_WITH_OPENCV
  // ---- reference signatures with the projection prologue on the device ------------------------------------------------------
  // Project MapPoints tracked in last frame into the current frame and search matches (Tracking::TrackWithMotionModel),
  // src/ORBmatcher.cc:2027-2289, incl. fisheye-stereo frames (Nleft != -1: the right-image pass of :2188-2255).
  template <class FrameT>
  int SearchByProjection(FrameT& CurrentFrame, const FrameT& LastFrame, const float th, const bool bMono) {
    const cv::Mat Rcw = CurrentFrame.mTcw.rowRange(0, 3).colRange(0, 3);   // :2038-2047, the reference's lines
    const cv::Mat tcw = CurrentFrame.mTcw.rowRange(0, 3).col(3);
    const cv::Mat twc = -Rcw.t() * tcw;
    const cv::Mat Rlw = LastFrame.mTcw.rowRange(0, 3).colRange(0, 3);
    const cv::Mat tlw = LastFrame.mTcw.rowRange(0, 3).col(3);
    const cv::Mat tlc = Rlw * twc + tlw;
    const bool bForward = tlc.at<float>(2) > CurrentFrame.mb && !bMono;
    const bool bBackward = -tlc.at<float>(2) > CurrentFrame.mb && !bMono;
    orbx_projection P = projection_of(Rcw, tcw, twc, CurrentFrame.mpCamera);
    frame_bounds(CurrentFrame, P);
    P.depth_mode = 0; P.invz_mode = 0; P.bf = CurrentFrame.mbf; P.want_xr = 1;
    P.level_mode = 0; P.window_mode = bForward ? 1 : (bBackward ? 2 : 0); P.th = th;
    const int n = LastFrame.N;
    PointArrays A(n);
    A.useOctave(); A.useAngle(); A.useObserved();
    for (int i = 0; i < n; i++) {
      auto* pMP = LastFrame.mvpMapPoints[i];
      if (!pMP || LastFrame.mvbOutlier[i]) continue;
      A.valid[i] = 1;
      A.setPos(i, pMP->GetWorldPos()); A.setDesc(i, pMP->GetDescriptor());
      A.observed[i] = pMP->Observations() > 0;   // a keypoint given to a point without observations can be taken again (:2134-2136)
      const bool lastLeft = LastFrame.Nleft == -1 || i < LastFrame.Nleft;                          // :2100-2101, :2169-2171
      A.octave[i] = lastLeft ? LastFrame.mvKeys[i].octave : LastFrame.mvKeysRight[i - LastFrame.Nleft].octave;
      A.angle[i] = LastFrame.Nleft == -1 ? LastFrame.mvKeysUn[i].angle : (lastLeft ? LastFrame.mvKeys[i].angle : LastFrame.mvKeysRight[i - LastFrame.Nleft].angle);
    }
    if (CurrentFrame.Nleft == -1)
      return run_window(CurrentFrame, P, A, TH_HIGH, /*anyMpBlocks=*/false, mbCheckOrientation, [&](int k, int i) { CurrentFrame.mvpMapPoints[k] = LastFrame.mvpMapPoints[i]; });
    // fisheye-stereo current frame: left image = mvKeys, right image = mvKeysRight, searched with x3Dr = Rrl * x3Dc + trl (:2189-2192)
    const int nl = CurrentFrame.Nleft, nr = CurrentFrame.N - nl;
    DescHolder dd(CurrentFrame.mDescriptors, CurrentFrame.N);
    orbx_frame_view L, R;
    L.n = nl; L.keypoints = reinterpret_cast<const orbx_keypoint*>(CurrentFrame.mvKeys.data()); L.descriptors = dd.p; L.u_right = nullptr;
    L.scale_factors = CurrentFrame.mvScaleFactors.data(); L.nlevels = (int)CurrentFrame.mvScaleFactors.size();
    L.min_x = CurrentFrame.mnMinX; L.min_y = CurrentFrame.mnMinY; L.max_x = CurrentFrame.mnMaxX; L.max_y = CurrentFrame.mnMaxY;
    R = L;
    R.n = nr; R.keypoints = reinterpret_cast<const orbx_keypoint*>(CurrentFrame.mvKeysRight.data()); R.descriptors = dd.p + (size_t)nl * 32;
    P.want_xr = 0;
    orbx_projection PR = P;
    PR.chain = 1; PR.depth_mode = 2; PR.bounds_mode = 2;
    put3x3(CurrentFrame.mTrl.colRange(0, 3).rowRange(0, 3), PR.R2); put3(CurrentFrame.mTrl.col(3), PR.t2);
    std::vector<int> kpMp(std::max(CurrentFrame.N, 1), -1);
    for (int k = 0; k < CurrentFrame.N; k++) {
      auto* p = CurrentFrame.mvpMapPoints[k];
      kpMp[k] = !p ? -1 : (p->Observations() > 0 ? -2 : -3);
    }
    const orbx_projection_points pts = A.c();
    int nmatches = 0;
    orbx_shim::check(orbx_search_projected_window_stereo(m_, &L, &R, &P, &PR, &pts, A.angle.data(), A.observed.data(), TH_HIGH, mbCheckOrientation ? 1 : 0, kpMp.data(),
                                                         &nmatches),
                     "orbx_search_projected_window_stereo");
    for (int k = 0; k < CurrentFrame.N; k++) {
      if (kpMp[k] >= 0) CurrentFrame.mvpMapPoints[k] = LastFrame.mvpMapPoints[kpMp[k]];
      else if (kpMp[k] == -1 && CurrentFrame.mvpMapPoints[k]) CurrentFrame.mvpMapPoints[k] = nullptr;
    }
    return nmatches;
  }

  // Project MapPoints seen in KeyFrame into the Frame and search matches (relocalisation), src/ORBmatcher.cc:2291-2413.
  template <class FrameT, class KeyFrameT, class MapPointT>
  int SearchByProjection(FrameT& CurrentFrame, KeyFrameT* pKF, const std::set<MapPointT*>& sAlreadyFound, const float th, const int ORBdist) {
    // (on a fisheye-stereo frame the reference searches the left image only here: GetFeaturesInArea without bRight, :2342; mvKeysUn = mvKeys)
    const cv::Mat Rcw = CurrentFrame.mTcw.rowRange(0, 3).colRange(0, 3);   // :2295-2297
    const cv::Mat tcw = CurrentFrame.mTcw.rowRange(0, 3).col(3);
    const cv::Mat Ow = -Rcw.t() * tcw;
    orbx_projection P = projection_of(Rcw, tcw, Ow, CurrentFrame.mpCamera);
    frame_bounds(CurrentFrame, P);
    P.depth_mode = 2; P.range_mode = 1; P.level_mode = 1; P.window_mode = 0; P.th = th;
    P.log_scale_factor = CurrentFrame.mfLogScaleFactor; P.n_scale_levels = CurrentFrame.mnScaleLevels;
    const std::vector<MapPointT*> vpMPs = pKF->GetMapPointMatches();
    const int n = (int)vpMPs.size();
    PointArrays A(n);
    A.useRange(); A.useAngle();
    for (int i = 0; i < n; i++) {
      MapPointT* pMP = vpMPs[i];
      if (!pMP || pMP->isBad() || sAlreadyFound.count(pMP)) continue;
      A.valid[i] = 1;
      A.setPos(i, pMP->GetWorldPos()); A.setDesc(i, pMP->GetDescriptor());
      A.setRange(i, pMP);
      A.angle[i] = pKF->mvKeysUn[i].angle;
    }
    return run_window(CurrentFrame, P, A, ORBdist, /*anyMpBlocks=*/true, mbCheckOrientation, [&](int k, int i) { CurrentFrame.mvpMapPoints[k] = vpMPs[i]; });
  }

  // Project MapPoints using a Similarity Transformation and search matches (loop detection), src/ORBmatcher.cc:489-602.
  template <class KeyFrameT, class MapPointT>
  int SearchByProjection(KeyFrameT* pKF, cv::Mat Scw, const std::vector<MapPointT*>& vpPoints, std::vector<MapPointT*>& vpMatched, int th,
                         float ratioHamming = 1.0) {
    return search_scw(pKF, Scw, vpPoints, vpMatched, th, ratioHamming, /*cameraProject=*/true, [](int, int) {});
  }
  // ... with the keyframe of every point (place recognition), src/ORBmatcher.cc:604-720.
  template <class KeyFrameT, class MapPointT>
  int SearchByProjection(KeyFrameT* pKF, cv::Mat Scw, const std::vector<MapPointT*>& vpPoints, const std::vector<KeyFrameT*>& vpPointsKFs,
                         std::vector<MapPointT*>& vpMatched, std::vector<KeyFrameT*>& vpMatchedKF, int th, float ratioHamming = 1.0) {
    // this overload projects with fx*(x/z)+cx whatever the camera model is (:646-652)
    return search_scw(pKF, Scw, vpPoints, vpMatched, th, ratioHamming, /*cameraProject=*/false, [&](int k, int i) { vpMatchedKF[k] = vpPointsKFs[i]; });
  }

  // Project MapPoints into KeyFrame and search for duplicated MapPoints, src/ORBmatcher.cc:1425-1662, incl. the right camera of a
  // fisheye-stereo keyframe (bRight: right pose, mpCamera2, mvKeysRight / mGridRight, descriptor rows and indices offset by NLeft).
  template <class KeyFrameT, class MapPointT>
  int Fuse(KeyFrameT* pKF, const std::vector<MapPointT*>& vpMapPoints, const float th = 3.0, const bool bRight = false) {
    const cv::Mat Rcw = bRight ? pKF->GetRightRotation() : pKF->GetRotation();   // :1431-1443
    const cv::Mat tcw = bRight ? pKF->GetRightTranslation() : pKF->GetTranslation();
    const cv::Mat Ow = bRight ? pKF->GetRightCameraCenter() : pKF->GetCameraCenter();
    orbx_projection P = projection_of(Rcw, tcw, Ow, bRight ? pKF->mpCamera2 : pKF->mpCamera);
    keyframe_bounds(pKF, P);
    P.depth_mode = 1; P.invz_mode = 1; P.bf = pKF->mbf; P.want_xr = 1; P.range_mode = 1; P.normal_check = 1; P.level_mode = 1; P.window_mode = 3; P.th = th;
    P.log_scale_factor = pKF->mfLogScaleFactor; P.n_scale_levels = pKF->mnScaleLevels;
    const int n = (int)vpMapPoints.size();
    PointArrays A(n);
    A.useRange(); A.useNormal();
    for (int i = 0; i < n; i++) {
      MapPointT* pMP = vpMapPoints[i];
      if (!pMP || pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;
      A.valid[i] = 1;
      A.setPos(i, pMP->GetWorldPos()); A.setDesc(i, pMP->GetDescriptor()); A.setRange(i, pMP); A.setNormal(i, pMP->GetNormal());
    }
    std::vector<int> bestIdx, bestDist;
    if (pKF->NLeft == -1) run_best(*pKF, P, A, /*gate=*/true, pKF->mvInvLevelSigma2.data(), bestIdx, bestDist);
    else {   // :1563-1565: mvKeys or mvKeysRight; mvuRight is read with the image-local index (:1574), rows / results carry + NLeft (:1602, :1610)
      const int off = bRight ? pKF->NLeft : 0;
      const std::vector<cv::KeyPoint>& keys = bRight ? pKF->mvKeysRight : pKF->mvKeys;
      DescHolder dd(pKF->mDescriptors, pKF->N);
      orbx_frame_view v;
      v.n = (int)keys.size(); v.keypoints = reinterpret_cast<const orbx_keypoint*>(keys.data()); v.descriptors = dd.p + (size_t)off * 32;
      v.u_right = pKF->mvuRight.size() >= keys.size() && v.n > 0 ? pKF->mvuRight.data() : nullptr;
      v.scale_factors = pKF->mvScaleFactors.data(); v.nlevels = (int)pKF->mvScaleFactors.size();
      v.min_x = pKF->mnMinX; v.min_y = pKF->mnMinY; v.max_x = pKF->mnMaxX; v.max_y = pKF->mnMaxY;
      bestIdx.assign(std::max(A.n, 1), -1); bestDist.assign(std::max(A.n, 1), 256);
      const orbx_projection_points pts = A.c();
      orbx_shim::check(orbx_search_projected_best(m_, &v, &P, &pts, 1, pKF->mvInvLevelSigma2.data(), bestIdx.data(), bestDist.data(), nullptr),
                       "orbx_search_projected_best");
      for (int i = 0; i < n; i++) if (bestIdx[i] >= 0) bestIdx[i] += off;
    }
    int nFused = 0;
    for (int i = 0; i < n; i++) {   // the bookkeeping of :1626-1650 in the reference's order (it can change what later points see)
      MapPointT* pMP = vpMapPoints[i];
      if (!A.valid[i] || pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;
      if (bestIdx[i] < 0 || bestDist[i] > TH_LOW) continue;
      MapPointT* pMPinKF = pKF->GetMapPoint(bestIdx[i]);
      if (pMPinKF) {
        if (!pMPinKF->isBad()) {
          if (pMPinKF->Observations() > pMP->Observations()) pMP->Replace(pMPinKF);
          else pMPinKF->Replace(pMP);
        }
      } else {
        pMP->AddObservation(pKF, bestIdx[i]);
        pKF->AddMapPoint(pMP, bestIdx[i]);
      }
      nFused++;
    }
    return nFused;
  }

  // Project MapPoints into KeyFrame using a given Sim3 and search for duplicated MapPoints, src/ORBmatcher.cc:1664-1786.
  template <class KeyFrameT, class MapPointT>
  int Fuse(KeyFrameT* pKF, cv::Mat Scw, const std::vector<MapPointT*>& vpPoints, float th, std::vector<MapPointT*>& vpReplacePoint) {
    cv::Mat sRcw = Scw.rowRange(0, 3).colRange(0, 3);   // :1673-1677
    const float scw = sqrt(sRcw.row(0).dot(sRcw.row(0)));
    cv::Mat Rcw = sRcw / scw;
    cv::Mat tcw = Scw.rowRange(0, 3).col(3) / scw;
    cv::Mat Ow = -Rcw.t() * tcw;
    orbx_projection P = projection_of(Rcw, tcw, Ow, pKF->mpCamera);
    keyframe_bounds(pKF, P);
    P.depth_mode = 1; P.range_mode = 1; P.normal_check = 1; P.level_mode = 1; P.window_mode = 3; P.th = th;
    P.log_scale_factor = pKF->mfLogScaleFactor; P.n_scale_levels = pKF->mnScaleLevels;
    const std::set<MapPointT*> spAlreadyFound = pKF->GetMapPoints();
    const int n = (int)vpPoints.size();
    PointArrays A(n);
    A.useRange(); A.useNormal();
    for (int i = 0; i < n; i++) {
      MapPointT* pMP = vpPoints[i];
      if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;
      A.valid[i] = 1;
      A.setPos(i, pMP->GetWorldPos()); A.setDesc(i, pMP->GetDescriptor()); A.setRange(i, pMP); A.setNormal(i, pMP->GetNormal());
    }
    std::vector<int> bestIdx, bestDist;
    run_best(*pKF, P, A, /*gate=*/false, nullptr, bestIdx, bestDist);
    int nFused = 0;
    for (int i = 0; i < n; i++) {   // :1765-1782
      MapPointT* pMP = vpPoints[i];
      if (!A.valid[i] || bestIdx[i] < 0 || bestDist[i] > TH_LOW) continue;
      MapPointT* pMPinKF = pKF->GetMapPoint(bestIdx[i]);
      if (pMPinKF) {
        if (!pMPinKF->isBad()) vpReplacePoint[i] = pMPinKF;
      } else {
        pMP->AddObservation(pKF, bestIdx[i]);
        pKF->AddMapPoint(pMP, bestIdx[i]);
      }
      nFused++;
    }
    return nFused;
  }

  // Search matches between MapPoints seen in KF1 and KF2 transforming by a Sim3 [s12*R12|t12], src/ORBmatcher.cc:1788-2018.
  template <class KeyFrameT, class MapPointT>
  int SearchBySim3(KeyFrameT* pKF1, KeyFrameT* pKF2, std::vector<MapPointT*>& vpMatches12, const float& s12, const cv::Mat& R12, const cv::Mat& t12,
                   const float th) {
    cv::Mat R1w = pKF1->GetRotation(), t1w = pKF1->GetTranslation();   // :1797-1807
    cv::Mat R2w = pKF2->GetRotation(), t2w = pKF2->GetTranslation();
    cv::Mat sR12 = s12 * R12;
    cv::Mat sR21 = (1.0 / s12) * R12.t();
    cv::Mat t21 = -sR21 * t12;
    const std::vector<MapPointT*> vpMapPoints1 = pKF1->GetMapPointMatches(), vpMapPoints2 = pKF2->GetMapPointMatches();
    const int N1 = (int)vpMapPoints1.size(), N2 = (int)vpMapPoints2.size();
    std::vector<bool> vbAlreadyMatched1(N1, false), vbAlreadyMatched2(N2, false);
    for (int i = 0; i < N1; i++) {
      MapPointT* pMP = vpMatches12[i];
      if (pMP) {
        vbAlreadyMatched1[i] = true;
        const int idx2 = std::get<0>(pMP->GetIndexInKeyFrame(pKF2));
        if (idx2 >= 0 && idx2 < N2) vbAlreadyMatched2[idx2] = true;
      }
    }
    auto pass = [&](KeyFrameT* from, KeyFrameT* into, const std::vector<MapPointT*>& pts, const std::vector<bool>& done, const cv::Mat& Ra, const cv::Mat& ta,
                    const cv::Mat& Rb, const cv::Mat& tb, std::vector<int>& match) {
      (void)from;
      orbx_projection P = projection_of(Ra, ta, ta, into->mpCamera);
      put3x3(Rb, P.R2); put3(tb, P.t2); P.chain = 1;
      P.fx = into->fx; P.fy = into->fy; P.cx = into->cx; P.cy = into->cy; P.proj_mode = 1;   // :1792-1795 (pKF1's intrinsics; same camera)
      keyframe_bounds(into, P);
      P.depth_mode = 1; P.invz_mode = 0; P.range_mode = 1; P.dist_mode = 1; P.level_mode = 1; P.window_mode = 3; P.th = th;
      P.log_scale_factor = into->mfLogScaleFactor; P.n_scale_levels = into->mnScaleLevels;
      const int n = (int)pts.size();
      PointArrays A(n);
      A.useRange();
      for (int i = 0; i < n; i++) {
        MapPointT* pMP = pts[i];
        if (!pMP || done[i] || pMP->isBad()) continue;
        A.valid[i] = 1;
        A.setPos(i, pMP->GetWorldPos()); A.setDesc(i, pMP->GetDescriptor()); A.setRange(i, pMP);
      }
      std::vector<int> bestIdx, bestDist;
      run_best(*into, P, A, false, nullptr, bestIdx, bestDist);
      match.assign(n, -1);
      for (int i = 0; i < n; i++)
        if (A.valid[i] && bestIdx[i] >= 0 && bestDist[i] <= TH_HIGH) match[i] = bestIdx[i];
    };
    std::vector<int> vnMatch1, vnMatch2;
    pass(pKF1, pKF2, vpMapPoints1, vbAlreadyMatched1, R1w, t1w, sR21, t21, vnMatch1);   // :1838-1915
    pass(pKF2, pKF1, vpMapPoints2, vbAlreadyMatched2, R2w, t2w, sR12, t12, vnMatch2);   // :1917-1996
    int nFound = 0;   // :1998-2015
    for (int i1 = 0; i1 < N1; i1++) {
      const int idx2 = vnMatch1[i1];
      if (idx2 >= 0) {
        const int idx1 = vnMatch2[idx2];
        if (idx1 == i1) { vpMatches12[i1] = vpMapPoints2[idx2]; nFound++; }
      }
    }
    return nFound;
  }

  // Matching to triangulate new MapPoints with the reference's parameter list (cv::Mat F12), src/ORBmatcher.cc:981-1222: pinhole
  // keyframes without a second camera, KannalaBrandt8 keyframes with one camera (TUM-VI monocular) or two (fisheye stereo).  For
  // pinhole keyframes F12 is the matrix LocalMapping::ComputeF12 formed (K1^-T [t12]x R12 K2^-1, the very
  // expression Pinhole::epipolarConstrain re-evaluates per pair, Pinhole.cpp:145-148); the epipole is projected here (:988-994).
  template <class KeyFrameT>
  int SearchForTriangulation(KeyFrameT* pKF1, KeyFrameT* pKF2, cv::Mat F12, std::vector<std::pair<size_t, size_t> >& vMatchedPairs,
                             const bool bOnlyStereo, const bool bCoarse = false) {
    // the gate is pCamera1->epipolarConstrain (:1146): Pinhole's epipolar-line distance and KannalaBrandt8's two-view triangulation
    // (KannalaBrandt8.cpp:244-247, 343-444; the cv::SVD of the 4x4 system is OpenCV's Jacobi, restated in csrc/kb8_triangulate.cuh)
    // both run on the device.  The 3x3 pose algebra in front of the loop is the reference's own lines (:988-1019).
    cv::Mat Cw = pKF1->GetCameraCenter();
    cv::Mat R2w = pKF2->GetRotation();
    cv::Mat t2w = pKF2->GetTranslation();
    cv::Mat C2 = R2w * Cw + t2w;
    cv::Point2f ep = pKF2->mpCamera->project(C2);
    const bool fisheye1 = pKF1->mpCamera->GetType() == 1, fisheye2 = pKF2->mpCamera->GetType() == 1;   // GeometricCamera::CAM_FISHEYE
    if (!fisheye1 && !fisheye2 && !pKF1->mpCamera2 && !pKF2->mpCamera2) {
      float f[9];
      put3x3(F12, f);
      return SearchForTriangulation(pKF1, pKF2, f, ep, vMatchedPairs, bOnlyStereo, bCoarse);
    }
    if (!fisheye1 || !fisheye2 || (pKF1->mpCamera2 != nullptr) != (pKF2->mpCamera2 != nullptr) ||
        (pKF1->mpCamera2 && (pKF1->mpCamera2->GetType() != 1 || pKF2->mpCamera2->GetType() != 1)))
      throw std::runtime_error("ORBmatcher::SearchForTriangulation: keyframes must share a camera model (Pinhole without a second camera, or "
                               "KannalaBrandt8 with or without one on both sides)");
    cv::Mat R1w = pKF1->GetRotation();
    cv::Mat t1w = pKF1->GetTranslation();
    orbx_kb8_pair G[4];
    memset(G, 0, sizeof G);
    auto cam8 = [](decltype(pKF1->mpCamera) c, float* o) { for (int i = 0; i < 8; i++) o[i] = c->getParameter(i); };
    const bool two = pKF1->mpCamera2 != nullptr;
    if (!two) {
      cv::Mat R12 = R1w * R2w.t();
      cv::Mat t12 = -R1w * R2w.t() * t2w + t1w;
      put3x3(R12, G[0].R12); put3(t12, G[0].t12); cam8(pKF1->mpCamera, G[0].cam1); cam8(pKF2->mpCamera, G[0].cam2);
    } else {
      cv::Mat Rll = pKF1->GetRotation() * pKF2->GetRotation().t();
      cv::Mat Rlr = pKF1->GetRotation() * pKF2->GetRightRotation().t();
      cv::Mat Rrl = pKF1->GetRightRotation() * pKF2->GetRotation().t();
      cv::Mat Rrr = pKF1->GetRightRotation() * pKF2->GetRightRotation().t();
      cv::Mat tll = pKF1->GetRotation() * (-pKF2->GetRotation().t() * pKF2->GetTranslation()) + pKF1->GetTranslation();
      cv::Mat tlr = pKF1->GetRotation() * (-pKF2->GetRightRotation().t() * pKF2->GetRightTranslation()) + pKF1->GetTranslation();
      cv::Mat trl = pKF1->GetRightRotation() * (-pKF2->GetRotation().t() * pKF2->GetTranslation()) + pKF1->GetRightTranslation();
      cv::Mat trr = pKF1->GetRightRotation() * (-pKF2->GetRightRotation().t() * pKF2->GetRightTranslation()) + pKF1->GetRightTranslation();
      put3x3(Rll, G[0].R12); put3(tll, G[0].t12); cam8(pKF1->mpCamera, G[0].cam1); cam8(pKF2->mpCamera, G[0].cam2);
      put3x3(Rlr, G[1].R12); put3(tlr, G[1].t12); cam8(pKF1->mpCamera, G[1].cam1); cam8(pKF2->mpCamera2, G[1].cam2);
      put3x3(Rrl, G[2].R12); put3(trl, G[2].t12); cam8(pKF1->mpCamera2, G[2].cam1); cam8(pKF2->mpCamera, G[2].cam2);
      put3x3(Rrr, G[3].R12); put3(trr, G[3].t12); cam8(pKF1->mpCamera2, G[3].cam1); cam8(pKF2->mpCamera2, G[3].cam2);
    }
    const int n1 = pKF1->N, n2 = pKF2->N;
    // kp = NLeft == -1 ? mvKeysUn[i] : i < NLeft ? mvKeys[i] : mvKeysRight[i - NLeft] (:1062-1064, 1102-1104): one list per keyframe
    std::vector<cv::KeyPoint> k1, k2;
    auto keys = [](KeyFrameT* K, std::vector<cv::KeyPoint>& out) {
      if (K->NLeft == -1) { out = K->mvKeysUn; return; }
      out.assign(K->mvKeys.begin(), K->mvKeys.begin() + K->NLeft);
      out.insert(out.end(), K->mvKeysRight.begin(), K->mvKeysRight.end());
    };
    keys(pKF1, k1); keys(pKF2, k2);
    if ((int)k1.size() != n1 || (int)k2.size() != n2) throw std::runtime_error("ORBmatcher::SearchForTriangulation: N != NLeft + NRight");
    std::vector<uint8_t> e1(std::max(n1, 1)), e2(std::max(n2, 1)), s1(std::max(n1, 1)), s2(std::max(n2, 1));
    for (int i = 0; i < n1; i++) { s1[i] = !two && pKF1->mvuRight[i] >= 0; e1[i] = !pKF1->GetMapPoint(i) && (!bOnlyStereo || s1[i]); }
    for (int i = 0; i < n2; i++) { s2[i] = !two && pKF2->mvuRight[i] >= 0; e2[i] = !pKF2->GetMapPoint(i) && (!bOnlyStereo || s2[i]); }
    FeatVecHolder fv1(pKF1->mFeatVec), fv2(pKF2->mFeatVec);
    DescHolder d1(pKF1->mDescriptors, n1), d2(pKF2->mDescriptors, n2);
    std::vector<int> m12(std::max(n1, 1), -1);
    int nmatches = 0;
    orbx_shim::check(orbx_search_for_triangulation_kb8(m_, n1, reinterpret_cast<const orbx_keypoint*>(k1.data()), d1.p, &fv1.v, e1.data(), s1.data(),
                                                       two ? pKF1->NLeft : -1, n2, reinterpret_cast<const orbx_keypoint*>(k2.data()), d2.p, &fv2.v,
                                                       e2.data(), s2.data(), two ? pKF2->NLeft : -1, G, two ? 4 : 1, ep.x, ep.y,
                                                       pKF2->mvScaleFactors.data(), pKF1->mvLevelSigma2.data(), pKF2->mvLevelSigma2.data(),
                                                       (int)pKF2->mvScaleFactors.size(), bCoarse ? 1 : 0, mbCheckOrientation ? 1 : 0, m12.data(),
                                                       &nmatches),
                     "orbx_search_for_triangulation_kb8");
    vMatchedPairs.clear();
    vMatchedPairs.reserve(nmatches);
    for (int i = 0; i < n1; i++)
      if (m12[i] >= 0) vMatchedPairs.push_back(std::make_pair((size_t)i, (size_t)m12[i]));
    return nmatches;
  }
#endif  // ORBX_WITH_OPENCV

  // ---- SURVEY.md section 8(f) rank 1 ---------------------------------------------------------------------------------------
  // Search matches between MapPoints in a KeyFrame and ORB in a Frame, brute force constrained to ORB that belong to the same
  // vocabulary node (relocalisation / reference-keyframe tracking), src/ORBmatcher.cc:273-475, incl. fisheye-stereo frames
  // (F.Nleft != -1, :338-431: a keyframe feature may take one left-image and one right-image feature of the frame).
  template <class KeyFrameT, class FrameT, class MapPointT>
  int SearchByBoW(KeyFrameT* pKF, FrameT& F, std::vector<MapPointT*>& vpMapPointMatches) {
    const std::vector<MapPointT*> vpMapPointsKF = pKF->GetMapPointMatches();
    vpMapPointMatches.assign(F.N, static_cast<MapPointT*>(nullptr));
    const int n1 = (int)vpMapPointsKF.size();
    std::vector<uint8_t> elig1(std::max(n1, 1));
    for (int i = 0; i < n1; i++) elig1[i] = vpMapPointsKF[i] && !vpMapPointsKF[i]->isBad();
    FeatVecHolder fv1(pKF->mFeatVec), fv2(F.mFeatVec);
    DescHolder d1(pKF->mDescriptors, n1), d2(F.mDescriptors, F.N);
    std::vector<int> m12(std::max(n1, 1), -1);
    int nmatches = 0;
    if (F.Nleft != -1) {
      // keypoints whose angles feed the rotation histogram (:381-389, :409-416): mvKeys then mvKeysRight on both sides when the
      // keyframe has a second camera, else the keyframe's mvKeysUn
      std::vector<orbx_keypoint> k1, k2;
      const orbx_keypoint* kp1 = reinterpret_cast<const orbx_keypoint*>(pKF->mvKeysUn.data());
      if (pKF->mpCamera2) {
        k1.resize(std::max(n1, 1));
        for (int i = 0; i < n1; i++)
          memcpy(&k1[i], i >= pKF->NLeft ? (const void*)&pKF->mvKeysRight[i - pKF->NLeft] : (const void*)&pKF->mvKeys[i], sizeof(orbx_keypoint));
        kp1 = k1.data();
      }
      k2.resize(std::max(F.N, 1));
      for (int i = 0; i < F.N; i++)
        memcpy(&k2[i], i >= F.Nleft ? (const void*)&F.mvKeysRight[i - F.Nleft] : (const void*)&F.mvKeys[i], sizeof(orbx_keypoint));
      std::vector<int> m12R(std::max(n1, 1), -1);
      orbx_shim::check(orbx_search_by_bow_stereo(m_, n1, kp1, d1.p, &fv1.v, elig1.data(), F.N, F.Nleft, k2.data(), d2.p, &fv2.v, TH_LOW, mfNNratio,
                                                 mbCheckOrientation ? 1 : 0, m12.data(), m12R.data(), &nmatches),
                       "orbx_search_by_bow_stereo");
      for (int i = 0; i < n1; i++) {
        if (m12[i] >= 0) vpMapPointMatches[m12[i]] = vpMapPointsKF[i];
        if (m12R[i] >= 0) vpMapPointMatches[m12R[i]] = vpMapPointsKF[i];
      }
      return nmatches;
    }
    orbx_shim::check(orbx_search_by_bow(m_, n1, reinterpret_cast<const orbx_keypoint*>(pKF->mvKeysUn.data()), d1.p, &fv1.v, elig1.data(), F.N,
                                        reinterpret_cast<const orbx_keypoint*>(F.mvKeys.data()), d2.p, &fv2.v, nullptr, TH_LOW, mfNNratio,
                                        mbCheckOrientation ? 1 : 0, m12.data(), &nmatches),
                     "orbx_search_by_bow");
    for (int i = 0; i < n1; i++)
      if (m12[i] >= 0) vpMapPointMatches[m12[i]] = vpMapPointsKF[i];
    return nmatches;
  }

  // Matches between the map points of two keyframes (loop detection / merging), src/ORBmatcher.cc:839-979.
  template <class KeyFrameT, class MapPointT>
  int SearchByBoW(KeyFrameT* pKF1, KeyFrameT* pKF2, std::vector<MapPointT*>& vpMatches12) {
    const std::vector<MapPointT*> vpMapPoints1 = pKF1->GetMapPointMatches(), vpMapPoints2 = pKF2->GetMapPointMatches();
    const int n1 = (int)std::min(vpMapPoints1.size(), pKF1->mvKeysUn.size()), n2 = (int)std::min(vpMapPoints2.size(), pKF2->mvKeysUn.size());
    vpMatches12.assign(vpMapPoints1.size(), static_cast<MapPointT*>(nullptr));
    std::vector<uint8_t> elig1(std::max(n1, 1)), elig2(std::max(n2, 1));
    for (int i = 0; i < n1; i++) elig1[i] = vpMapPoints1[i] && !vpMapPoints1[i]->isBad();
    for (int i = 0; i < n2; i++) elig2[i] = vpMapPoints2[i] && !vpMapPoints2[i]->isBad();
    FeatVecHolder fv1(pKF1->mFeatVec), fv2(pKF2->mFeatVec);
    DescHolder d1(pKF1->mDescriptors, n1), d2(pKF2->mDescriptors, n2);
    std::vector<int> m12(std::max(n1, 1), -1);
    int nmatches = 0;
    orbx_shim::check(orbx_search_by_bow(m_, n1, reinterpret_cast<const orbx_keypoint*>(pKF1->mvKeysUn.data()), d1.p, &fv1.v, elig1.data(), n2,
                                        reinterpret_cast<const orbx_keypoint*>(pKF2->mvKeysUn.data()), d2.p, &fv2.v, elig2.data(), TH_LOW - 1,
                                        mfNNratio, mbCheckOrientation ? 1 : 0, m12.data(), &nmatches),
                     "orbx_search_by_bow");
    for (int i = 0; i < n1; i++)
      if (m12[i] >= 0) vpMatches12[i] = vpMapPoints2[m12[i]];
    return nmatches;
  }

  // Matching to triangulate new MapPoints, src/ORBmatcher.cc:981-1222, for pinhole keyframes without a second camera.  The
  // reference ignores its F12 argument and lets Pinhole::epipolarConstrain rebuild the matrix per pair
  // (src/CameraModels/Pinhole.cpp:145-148); here the caller forms it once (row-major 3x3) together with the epipole
  // ep = pKF2->mpCamera->project(R2w*Cw + t2w) (src/ORBmatcher.cc:988-994).
  template <class KeyFrameT>
  int SearchForTriangulation(KeyFrameT* pKF1, KeyFrameT* pKF2, const float* F12, const cvx::Point2f& ep,
                             std::vector<std::pair<size_t, size_t> >& vMatchedPairs, const bool bOnlyStereo, const bool bCoarse = false) {
    const int n1 = pKF1->N, n2 = pKF2->N;
    std::vector<uint8_t> e1(std::max(n1, 1)), e2(std::max(n2, 1)), s1(std::max(n1, 1)), s2(std::max(n2, 1));
    for (int i = 0; i < n1; i++) { s1[i] = pKF1->mvuRight[i] >= 0; e1[i] = !pKF1->GetMapPoint(i) && (!bOnlyStereo || s1[i]); }
    for (int i = 0; i < n2; i++) { s2[i] = pKF2->mvuRight[i] >= 0; e2[i] = !pKF2->GetMapPoint(i) && (!bOnlyStereo || s2[i]); }
    FeatVecHolder fv1(pKF1->mFeatVec), fv2(pKF2->mFeatVec);
    DescHolder d1(pKF1->mDescriptors, n1), d2(pKF2->mDescriptors, n2);
    std::vector<int> m12(std::max(n1, 1), -1);
    int nmatches = 0;
    orbx_shim::check(orbx_search_for_triangulation(m_, n1, reinterpret_cast<const orbx_keypoint*>(pKF1->mvKeysUn.data()), d1.p, &fv1.v, e1.data(),
                                                   s1.data(), n2, reinterpret_cast<const orbx_keypoint*>(pKF2->mvKeysUn.data()), d2.p, &fv2.v,
                                                   e2.data(), s2.data(), F12, ep.x, ep.y, pKF2->mvScaleFactors.data(), pKF2->mvLevelSigma2.data(),
                                                   (int)pKF2->mvScaleFactors.size(), bCoarse ? 1 : 0, mbCheckOrientation ? 1 : 0, m12.data(),
                                                   &nmatches),
                     "orbx_search_for_triangulation");
    vMatchedPairs.clear();
    vMatchedPairs.reserve(nmatches);
    for (int i = 0; i < n1; i++)
      if (m12[i] >= 0) vMatchedPairs.push_back(std::make_pair((size_t)i, (size_t)m12[i]));
    return nmatches;
  }

  // The search loop of Fuse (src/ORBmatcher.cc:1524-1590, :1728-1756; gate = true for the first overload) and of both passes of
  // SearchBySim3 (:1866-1895, :1946-1975) once the caller has projected the points: best keypoint index / distance per point.
  // Replace / AddObservation / the Sim3 agreement check stay with the caller, in the reference's order.
  template <class KeyFrameT>
  void SearchWindowBest(KeyFrameT* pKF, const std::vector<uint8_t>& valid, const std::vector<float>& u, const std::vector<float>& v,
                        const std::vector<float>& radius, const std::vector<int>& level, const std::vector<float>* ur,
                        const std::vector<uint8_t>& desc, bool gate, std::vector<int>& bestIdx, std::vector<int>& bestDist) {
    FrameViewHolder fv(*pKF);
    bestIdx.assign(valid.size(), -1); bestDist.assign(valid.size(), 256);
    orbx_shim::check(orbx_search_window_best(m_, &fv.v, (int)valid.size(), valid.data(), u.data(), v.data(), radius.data(), level.data(),
                                             ur ? ur->data() : nullptr, desc.data(), gate ? 1 : 0, gate ? pKF->mvInvLevelSigma2.data() : nullptr,
                                             bestIdx.data(), bestDist.data()),
                     "orbx_search_window_best");
  }

  // Frame::ComputeStereoMatches (src/Frame.cc:901-1079) for a frame whose two extractors have just run: fills F.mvuRight / F.mvDepth.
  template <class FrameT>
  void ComputeStereoMatches(FrameT& F, ORBextractor& left, ORBextractor& right) {
    orbx_shim::check(orbx_stereo_match_batch(m_, left.handle(), right.handle(), F.mb, F.mbf, nullptr), "orbx_stereo_match_batch");
    const int cap = std::max(F.N, 1);
    F.mvuRight.assign(F.N, -1.0f);
    F.mvDepth.assign(F.N, -1.0f);
    orbx_shim::check(orbx_stereo_results_host(m_, F.mvuRight.data(), F.mvDepth.data(), nullptr, nullptr, cap), "orbx_stereo_results_host");
  }

  // The brute-force 2-NN + Lowe ratio of Frame::ComputeStereoFishEyeMatches (src/Frame.cc:1246-1253); triangulation stays on the host.
  void KnnMatch2(const cvx::Mat& query, const cvx::Mat& train, std::vector<int>& idx, std::vector<int>& dist, std::vector<uint8_t>& good) {
    idx.assign((size_t)query.rows * 2, -1); dist.assign((size_t)query.rows * 2, -1); good.assign(query.rows, 0);
    std::vector<uint8_t> q((size_t)query.rows * 32), t((size_t)train.rows * 32);
    for (int i = 0; i < query.rows; i++) memcpy(&q[(size_t)i * 32], query.ptr(i), 32);
    for (int i = 0; i < train.rows; i++) memcpy(&t[(size_t)i * 32], train.ptr(i), 32);
    orbx_shim::check(orbx_knn2(m_, q.data(), query.rows, t.data(), train.rows, idx.data(), dist.data(), good.data()), "orbx_knn2");
  }

  // MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:385-428) for a batch of map points: vDescriptors[p] is the list
  // the reference gathers per point (:366-382); returns the index of the descriptor to keep per point (-1 for an empty list).
  std::vector<int> DistinctiveDescriptors(const std::vector<std::vector<cvx::Mat> >& vDescriptors) {
    std::vector<int> start(1, 0), best(vDescriptors.size(), -1), median(vDescriptors.size(), 0);
    std::vector<uint8_t> rows;
    for (size_t p = 0; p < vDescriptors.size(); p++) {
      for (size_t k = 0; k < vDescriptors[p].size(); k++) rows.insert(rows.end(), vDescriptors[p][k].ptr(0), vDescriptors[p][k].ptr(0) + 32);
      start.push_back((int)(rows.size() / 32));
    }
    if (rows.empty()) rows.resize(32);
    orbx_shim::check(orbx_distinctive_descriptors(m_, rows.data(), start.data(), (int)vDescriptors.size(), best.data(), median.data()),
                     "orbx_distinctive_descriptors");
    return best;
  }

  orbx_matcher* handle() const { return m_; }

 protected:
  float mfNNratio;
  bool mbCheckOrientation;

 private:
#ifdef ORBX_WITH_OPENCV
  static void put3x3(const cv::Mat& R, float* o) { for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) o[3 * r + c] = R.at<float>(r, c); }
  static void put3(const cv::Mat& t, float* o) { for (int i = 0; i < 3; i++) o[i] = t.at<float>(i); }
  template <class CameraT>
  static orbx_projection projection_of(const cv::Mat& R, const cv::Mat& t, const cv::Mat& Ow, CameraT* cam) {
    orbx_projection P;
    memset(&P, 0, sizeof P);
    put3x3(R, P.R); put3(t, P.t); put3(Ow, P.Ow);
    P.fx = cam->getParameter(0); P.fy = cam->getParameter(1); P.cx = cam->getParameter(2); P.cy = cam->getParameter(3);
    P.camera_model = cam->GetType() == 1 ? ORBX_CAMERA_KB8 : ORBX_CAMERA_PINHOLE;   // GeometricCamera::CAM_FISHEYE
    if (P.camera_model == ORBX_CAMERA_KB8) for (int i = 0; i < 4; i++) P.kb[i] = cam->getParameter(4 + i);
    return P;
  }
  template <class FrameT> static void frame_bounds(const FrameT& F, orbx_projection& P) {
    P.bounds_mode = 0; P.min_x = F.mnMinX; P.max_x = F.mnMaxX; P.min_y = F.mnMinY; P.max_y = F.mnMaxY;
  }
  template <class KeyFrameT> static void keyframe_bounds(const KeyFrameT* K, orbx_projection& P) {
    P.bounds_mode = 1; P.min_x = (float)K->mnMinX; P.max_x = (float)K->mnMaxX; P.min_y = (float)K->mnMinY; P.max_y = (float)K->mnMaxY;
  }
  // structure-of-arrays image of the map points a search walks over
  struct PointArrays {
    int n;
    std::vector<uint8_t> valid, desc, observed;   // observed: Observations() > 0 of the point (it blocks the keypoint it takes); empty = all
    std::vector<float> pos, normal, maxD, minD, angle;
    std::vector<int> octave;
    explicit PointArrays(int n_) : n(n_), valid(std::max(n_, 1), 0), desc((size_t)std::max(n_, 1) * 32, 0), pos((size_t)std::max(n_, 1) * 3, 0.f) {}
    void useOctave() { octave.assign(std::max(n, 1), 0); }
    void useAngle() { angle.assign(std::max(n, 1), 0.f); }
    void useObserved() { observed.assign(std::max(n, 1), 1); }
    void useRange() { maxD.assign(std::max(n, 1), 0.f); minD.assign(std::max(n, 1), 0.f); }
    void useNormal() { normal.assign((size_t)std::max(n, 1) * 3, 0.f); }
    void setPos(int i, const cv::Mat& p) { for (int k = 0; k < 3; k++) pos[3 * (size_t)i + k] = p.at<float>(k); }
    void setNormal(int i, const cv::Mat& p) { for (int k = 0; k < 3; k++) normal[3 * (size_t)i + k] = p.at<float>(k); }
    void setDesc(int i, const cv::Mat& d) { memcpy(&desc[(size_t)i * 32], d.ptr(0), 32); }
    // RAW mfMaxDistance / mfMinDistance recovered from the public getters (1.2f * max, 0.8f * min, MapPoint.cc:555-565): the
    // device multiplies by the same constants again, so what must round-trip is x -> 1.2f*x -> /1.2f -> *1.2f.  A MapPoint class
    // that exposes GetMaxDistance() / GetMinDistance() avoids the division.
    template <class MapPointT> void setRange(int i, MapPointT* pMP) { raw_range(pMP, maxD[i], minD[i], 0); }
    template <class MapPointT> static auto raw_range(MapPointT* p, float& mx, float& mn, int) -> decltype(p->GetMaxDistance(), void()) {
      mx = p->GetMaxDistance(); mn = p->GetMinDistance();
    }
    template <class MapPointT> static void raw_range(MapPointT* p, float& mx, float& mn, long) {
      mx = unscale(p->GetMaxDistanceInvariance(), 1.2f); mn = unscale(p->GetMinDistanceInvariance(), 0.8f);
    }
    static float unscale(float scaled, float k) {   // the float x with k * x == scaled (nearest candidate first)
      float x = scaled / k;
      if (k * x == scaled) return x;
      const float lo = std::nextafter(x, -INFINITY), hi = std::nextafter(x, INFINITY);
      if (k * lo == scaled) return lo;
      if (k * hi == scaled) return hi;
      return x;
    }
    orbx_projection_points c() const {
      orbx_projection_points p;
      p.n = n; p.world_pos = pos.data(); p.normal = normal.empty() ? nullptr : normal.data();
      p.max_distance = maxD.empty() ? nullptr : maxD.data(); p.min_distance = minD.empty() ? nullptr : minD.data();
      p.octave = octave.empty() ? nullptr : octave.data(); p.valid = valid.data(); p.desc = desc.data();
      return p;
    }
  };
  // window search with claiming: kpMp = the frame's current map points as -1 / -2 (blocking) / -3; assign(k, i) stores match i -> keypoint k
  template <class FrameT, class Assign>
  int run_window(FrameT& F, const orbx_projection& P, const PointArrays& A, int distTh, bool anyMpBlocks, bool checkOri, Assign assign) {
    FrameViewHolder fv(F);
    std::vector<int> kpMp(std::max(F.N, 1), -1);
    for (int k = 0; k < F.N; k++) {
      auto* p = F.mvpMapPoints[k];
      kpMp[k] = !p ? -1 : ((anyMpBlocks || p->Observations() > 0) ? -2 : -3);
    }
    const orbx_projection_points pts = A.c();
    int nmatches = 0;
    orbx_shim::check(orbx_search_projected_window(m_, &fv.v, &P, &pts, A.angle.empty() ? nullptr : A.angle.data(),
                                                  A.observed.empty() ? nullptr : A.observed.data(), distTh, anyMpBlocks ? 1 : 0,
                                                  checkOri ? 1 : 0, kpMp.data(), &nmatches, nullptr),
                     "orbx_search_projected_window");
    for (int k = 0; k < F.N; k++) {
      if (kpMp[k] >= 0) assign(k, kpMp[k]);
      else if (kpMp[k] == -1 && F.mvpMapPoints[k]) F.mvpMapPoints[k] = nullptr;   // a non-blocking point was re-matched, then dropped by the
    }                                                                             // rotation histogram (:2270-2281 writes NULL)
    return nmatches;
  }
  template <class KeyFrameT>
  void run_best(KeyFrameT& K, const orbx_projection& P, const PointArrays& A, bool gate, const float* invSigma2, std::vector<int>& bestIdx,
                std::vector<int>& bestDist) {
    FrameViewHolder fv(K);
    bestIdx.assign(std::max(A.n, 1), -1); bestDist.assign(std::max(A.n, 1), 256);
    const orbx_projection_points pts = A.c();
    orbx_shim::check(orbx_search_projected_best(m_, &fv.v, &P, &pts, gate ? 1 : 0, invSigma2, bestIdx.data(), bestDist.data(), nullptr),
                     "orbx_search_projected_best");
  }
  template <class KeyFrameT, class MapPointT, class Extra>
  int search_scw(KeyFrameT* pKF, cv::Mat Scw, const std::vector<MapPointT*>& vpPoints, std::vector<MapPointT*>& vpMatched, int th, float ratioHamming,
                 bool cameraProject, Extra extra) {
    cv::Mat sRcw = Scw.rowRange(0, 3).colRange(0, 3);   // :499-503 / :614-618, the reference's lines
    const float scw = sqrt(sRcw.row(0).dot(sRcw.row(0)));
    cv::Mat Rcw = sRcw / scw;
    cv::Mat tcw = Scw.rowRange(0, 3).col(3) / scw;
    cv::Mat Ow = -Rcw.t() * tcw;
    orbx_projection P = projection_of(Rcw, tcw, Ow, pKF->mpCamera);
    keyframe_bounds(pKF, P);
    P.depth_mode = 1; P.range_mode = 1; P.normal_check = 1; P.level_mode = 1; P.window_mode = 3; P.th = (float)th;
    P.log_scale_factor = pKF->mfLogScaleFactor; P.n_scale_levels = pKF->mnScaleLevels;
    if (!cameraProject) { P.proj_mode = 1; P.invz_mode = 1; P.fx = pKF->fx; P.fy = pKF->fy; P.cx = pKF->cx; P.cy = pKF->cy; }
    std::set<MapPointT*> spAlreadyFound(vpMatched.begin(), vpMatched.end());
    spAlreadyFound.erase(static_cast<MapPointT*>(nullptr));
    const int n = (int)vpPoints.size();
    PointArrays A(n);
    A.useRange(); A.useNormal();
    for (int i = 0; i < n; i++) {
      MapPointT* pMP = vpPoints[i];
      if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;
      A.valid[i] = 1;
      A.setPos(i, pMP->GetWorldPos()); A.setDesc(i, pMP->GetDescriptor()); A.setRange(i, pMP); A.setNormal(i, pMP->GetNormal());
    }
    FrameViewHolder fv(*pKF);
    const int N = (int)vpMatched.size();
    std::vector<int> kpMp(std::max(N, 1), -1);
    for (int k = 0; k < N; k++) kpMp[k] = vpMatched[k] ? -2 : -1;   // `if(vpMatched[idx]) continue` (:578): every match blocks
    const orbx_projection_points pts = A.c();
    int nmatches = 0;
    orbx_shim::check(orbx_search_projected_window(m_, &fv.v, &P, &pts, nullptr, nullptr, (int)std::floor(TH_LOW * ratioHamming), 1, 0, kpMp.data(),
                                                  &nmatches, nullptr),
                     "orbx_search_projected_window");
    for (int k = 0; k < N; k++)
      if (kpMp[k] >= 0) { vpMatched[k] = vpPoints[kpMp[k]]; extra(k, kpMp[k]); }
    return nmatches;
  }
#endif
  // gathers the Frame members of include/orbx.h's orbx_frame_view (descriptors are copied only if the Mat is not dense)
  struct FrameViewHolder {
    orbx_frame_view v;
    std::vector<uint8_t> dense;
    template <class FrameT>
    explicit FrameViewHolder(FrameT& F) {
      v.n = (int)F.mvKeysUn.size();
      v.keypoints = reinterpret_cast<const orbx_keypoint*>(F.mvKeysUn.data());
      if (F.mDescriptors.step == 32 || v.n == 0) v.descriptors = F.mDescriptors.ptr(0);
      else {
        dense.resize((size_t)v.n * 32);
        for (int i = 0; i < v.n; i++) memcpy(&dense[(size_t)i * 32], F.mDescriptors.ptr(i), 32);
        v.descriptors = dense.data();
      }
      v.u_right = (F.mvuRight.size() == (size_t)v.n && v.n > 0) ? F.mvuRight.data() : nullptr;
      v.scale_factors = F.mvScaleFactors.data();
      v.nlevels = (int)F.mvScaleFactors.size();
      v.min_x = F.mnMinX; v.min_y = F.mnMinY; v.max_x = F.mnMaxX; v.max_y = F.mnMaxY;
    }
  };
  // DBoW2::FeatureVector (an ordered map NodeId -> vector<unsigned>) flattened for the C-ABI
  struct FeatVecHolder {
    orbx_featvec v;
    std::vector<uint32_t> ids, indices;
    std::vector<int> start;
    template <class FeatVecT>
    explicit FeatVecHolder(const FeatVecT& fv) {
      start.push_back(0);
      for (typename FeatVecT::const_iterator it = fv.begin(); it != fv.end(); ++it) {
        ids.push_back((uint32_t)it->first);
        for (size_t k = 0; k < it->second.size(); k++) indices.push_back((uint32_t)it->second[k]);
        start.push_back((int)indices.size());
      }
      v.nnodes = (int)ids.size(); v.node_ids = ids.data(); v.node_start = start.data(); v.indices = indices.data();
    }
  };
  struct DescHolder {  // dense n x 32 view of a descriptor matrix
    const uint8_t* p = nullptr;
    std::vector<uint8_t> dense;
    DescHolder(const cvx::Mat& m, int n) {
      if (n == 0) { dense.resize(32); p = dense.data(); }
      else if (m.step == 32) p = m.ptr(0);
      else {
        dense.resize((size_t)n * 32);
        for (int i = 0; i < n; i++) memcpy(&dense[(size_t)i * 32], m.ptr(i), 32);
        p = dense.data();
      }
    }
  };
  orbx_matcher* m_ = nullptr;
};

// The tree descent of ORBVocabulary::transform (DBoW2 TemplatedVocabulary<FORB>, Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1130-1195,
// 1218-1259) on the GPU, for Frame::ComputeBoW / KeyFrame::ComputeBoW (src/Frame.cc:828-835).  Built from the flattened m_nodes
// (a subclass of the reference's vocabulary can export them: m_nodes is protected); BowVectorT / FeatureVectorT are the
// reference's DBoW2::BowVector / DBoW2::FeatureVector (addWeight / normalize / addFeature are called in the reference's order).
class ORBVocabularyDevice {
 public:
  ORBVocabularyDevice(int L, const std::vector<int>& childStart, const std::vector<int>& childIds, const std::vector<uint8_t>& nodeDesc,
                      const std::vector<double>& nodeWeight, const std::vector<int>& nodeWordId, int device = 0) {
    orbx_shim::check(orbx_vocabulary_create(device, (int)nodeWordId.size(), L, childStart.data(), childIds.data(), nodeDesc.data(),
                                            nodeWeight.data(), nodeWordId.data(), &v_),
                     "orbx_vocabulary_create");
  }
  ~ORBVocabularyDevice() { orbx_vocabulary_destroy(v_); }
  ORBVocabularyDevice(const ORBVocabularyDevice&) = delete;
  ORBVocabularyDevice& operator=(const ORBVocabularyDevice&) = delete;

  template <class BowVectorT, class FeatureVectorT, class NormT>
  void transform(ORBmatcher& matcher, const std::vector<cvx::Mat>& features, BowVectorT& v, FeatureVectorT& fv, int levelsup, NormT l1) const {
    v.clear(); fv.clear();
    const int n = (int)features.size();
    if (n == 0) return;
    std::vector<uint8_t> rows((size_t)n * 32);
    for (int i = 0; i < n; i++) memcpy(&rows[(size_t)i * 32], features[i].ptr(0), 32);
    std::vector<int> word(n), node(n);
    std::vector<double> weight(n);
    orbx_shim::check(orbx_vocabulary_transform(matcher.handle(), v_, rows.data(), n, levelsup, word.data(), node.data(), weight.data()),
                     "orbx_vocabulary_transform");
    for (int i = 0; i < n; i++)
      if (weight[i] > 0) { v.addWeight(word[i], weight[i]); fv.addFeature(node[i], (unsigned)i); }   // TemplatedVocabulary.h:1160-1168
    v.normalize(l1);                                                                                   // :1194 (L1 for ORBvoc)
  }

 private:
  orbx_vocabulary* v_ = nullptr;
};

}  // namespace ORB_SLAM3
