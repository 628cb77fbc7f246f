// CPU ORACLE support (test infrastructure, not product code).
//
// C entry points around the reference's own, UNMODIFIED src/ORBmatcher.cc (compiled by oracle/Makefile by path, against the
// stand-in cv:: header and the stand-in data model of slam_world.h) and around the member functions of Frame / KeyFrame /
// MapPoint / Pinhole / KannalaBrandt8 that sit on the path, whose bodies oracle/Makefile cuts out of the reference's .cc files at
// build time (oracle/_ref/*_slice.inc -- nothing is copied into the repository).  Loaded by oracle/refmatch.py for tests/ and
// bench.py's CPU legs only.
//
// A "view" (RefView) carries the members of a Frame or KeyFrame as plain arrays; map points are a structure of arrays.  Every
// call builds the reference's objects from them, runs the reference function, and copies the observable result out.
#include <chrono>
#include <cstdint>
#include <cstring>
#include <memory>
#include <thread>

#ifdef ORBX_SHIM_TWIN
// The twin build (oracle/_ref/libshim_match.so): the SAME caller code, data model and stand-in cv:: header, but the ORBmatcher is
// the product's drop-in (host/ORBmatcher.h -> liborbx.so, CUDA) instead of the reference's class.  tests/ run both libraries on
// the same worlds and compare every observable output, which exercises the reference-signature overloads end to end.
#define ORBX_WITH_OPENCV 1
#include "host/ORBmatcher.h"
#else
#include "ORBextractor.h"
#include "ORBmatcher.h"
#endif
#include "ref_alloc.h"

namespace cvshim { extern thread_local int g_blur_taps; }

namespace ORB_SLAM3 {
// ---- bodies cut from the reference (oracle/Makefile: frame_slice.inc <- src/Frame.cc, keyframe_slice.inc <- src/KeyFrame.cc,
// mappoint_slice.inc <- src/MapPoint.cc; the camera bodies are compiled in ref_camera.cc)
#include "frame_slice.inc"
#include "keyframe_slice.inc"
#include "mappoint_slice.inc"
}  // namespace ORB_SLAM3

using namespace ORB_SLAM3;

extern "C" {

struct RefView {
  int N, Nleft;                 // Nleft = -1: monocular / rectified stereo; else N = Nleft + Nright (fisheye stereo)
  const cv::KeyPoint* kps;      // mvKeys      (Nleft == -1: N entries; else Nleft entries)
  const cv::KeyPoint* kpsUn;    // mvKeysUn    (Nleft == -1: N entries; else may be null)
  const cv::KeyPoint* kpsRight; // mvKeysRight (nRight entries) or null
  int nRight;
  const uint8_t* desc;          // mDescriptors, N x 32
  const float* uRight;          // mvuRight (N) or null
  int nlevels;
  const float* scaleFactors; const float* invScaleFactors; const float* levelSigma2; const float* invLevelSigma2;
  float scaleFactor, logScaleFactor;
  float minX, maxX, minY, maxY;
  float fx, fy, cx, cy, bf, b;
  int camModel;                 // 0 Pinhole, 1 KannalaBrandt8 (kb = k1..k4)
  float kb[4];
  int cam2Model;                // -1: no second camera; else as camModel with cam2 = fx fy cx cy k1..k4
  float cam2[8];
  const float* Tcw;             // 4x4 row-major or null
  const float* Tlr;             // 3x4 row-major (left <- right) or null; Trl is derived as Frame.cc:1192-1195 does
  const int* mapPointIdx;       // mvpMapPoints as indices into the map-point arrays (-1 = none), N entries, or null
  const uint8_t* outlier;       // mvbOutlier or null
  int fvNodes; const uint32_t* fvIds; const int* fvStart; const uint32_t* fvIdx;   // mFeatVec as CSR (ids ascending) or fvNodes = 0
  const int* leftToRight;       // mvLeftToRightMatch (Nleft entries) or null = all -1   (fisheye stereo, Frame.cc:1228-1262)
  const int* rightToLeft;       // mvRightToLeftMatch (Nright entries) or null = all -1
};

struct RefPoints {
  int n;
  const float* pos; const float* normal;   // n x 3
  const uint8_t* desc;                     // n x 32
  const float* minDist; const float* maxDist;   // mfMinDistance / mfMaxDistance (raw)
  const int* nObs;                         // Observations(); null = 1
  const uint8_t* bad;                      // isBad(); null = false
};

}  // extern "C"

namespace {

cv::Mat mat_from(const float* p, int r, int c) {
  cv::Mat m(r, c, CV_32F);
  for (int y = 0; y < r; y++) for (int x = 0; x < c; x++) m.at<float>(y, x) = p[y * c + x];
  return m;
}
cv::Mat desc_rows(const uint8_t* d, int n) {
  cv::Mat m(std::max(n, 1), 32, CV_8U);
  if (n > 0) memcpy(m.data, d, (size_t)n * 32);
  return n > 0 ? m : m.rowRange(0, 0);
}

struct World {
  std::unique_ptr<MapPoint[]> mps;   // MapPoint holds a mutex: not movable
  int nmps = 0;
  std::vector<WorldEvent> log;
  std::vector<GeometricCamera*> cams;
  ~World() { for (auto* c : cams) delete c; }
  GeometricCamera* camera(int model, float fx, float fy, float cx, float cy, const float* k) {
    std::vector<float> p{fx, fy, cx, cy};
    GeometricCamera* c;
    if (model == 1) { for (int i = 0; i < 4; i++) p.push_back(k[i]); c = new KannalaBrandt8(p); }
    else c = new Pinhole(p);
    cams.push_back(c);
    return c;
  }
  void points(const RefPoints* P) {
    mps.reset(); nmps = 0;
    if (!P) return;
    mps.reset(new MapPoint[P->n]); nmps = P->n;
    for (int i = 0; i < P->n; i++) {
      MapPoint& m = mps[i];
      m.mnId = i;
      m.mWorldPos = mat_from(P->pos + 3 * i, 3, 1);
      if (P->normal) m.mNormalVector = mat_from(P->normal + 3 * i, 3, 1);
      if (P->desc) { m.mDescriptor = cv::Mat(1, 32, CV_8U); memcpy(m.mDescriptor.data, P->desc + 32 * (size_t)i, 32); }
      m.mfMinDistance = P->minDist ? P->minDist[i] : 0.f;
      m.mfMaxDistance = P->maxDist ? P->maxDist[i] : 0.f;
      m.nObs = P->nObs ? P->nObs[i] : 1;
      m.mbBad = P->bad ? P->bad[i] != 0 : false;
      m.events = &log;
    }
  }
  MapPoint* mp(int i) { return (i >= 0 && i < nmps) ? &mps[i] : nullptr; }

  void fill_common(const RefView& V, Frame& F) {
    F.N = V.N; F.Nleft = V.Nleft; F.Nright = V.Nleft == -1 ? -1 : V.N - V.Nleft;
    const int nl = V.Nleft == -1 ? V.N : V.Nleft;
    if (V.kps) F.mvKeys.assign(V.kps, V.kps + nl);
    if (V.kpsUn) F.mvKeysUn.assign(V.kpsUn, V.kpsUn + nl);
    else if (V.kps) F.mvKeysUn = F.mvKeys;
    if (V.kpsRight) F.mvKeysRight.assign(V.kpsRight, V.kpsRight + V.nRight);
    F.mDescriptors = desc_rows(V.desc, V.N);
    if (V.uRight) F.mvuRight.assign(V.uRight, V.uRight + V.N); else F.mvuRight.assign(V.N, -1.f);
    F.mvDepth.assign(V.N, -1.f);
    F.mnScaleLevels = V.nlevels; F.mfScaleFactor = V.scaleFactor; F.mfLogScaleFactor = V.logScaleFactor;
    if (V.scaleFactors) F.mvScaleFactors.assign(V.scaleFactors, V.scaleFactors + V.nlevels);
    if (V.invScaleFactors) F.mvInvScaleFactors.assign(V.invScaleFactors, V.invScaleFactors + V.nlevels);
    if (V.levelSigma2) F.mvLevelSigma2.assign(V.levelSigma2, V.levelSigma2 + V.nlevels);
    if (V.invLevelSigma2) F.mvInvLevelSigma2.assign(V.invLevelSigma2, V.invLevelSigma2 + V.nlevels);
    F.mnMinX = V.minX; F.mnMaxX = V.maxX; F.mnMinY = V.minY; F.mnMaxY = V.maxY;
    F.mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(F.mnMaxX - F.mnMinX);    // Frame.cc:268-269
    F.mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(F.mnMaxY - F.mnMinY);
    F.fx = V.fx; F.fy = V.fy; F.cx = V.cx; F.cy = V.cy; F.invfx = 1.0f / V.fx; F.invfy = 1.0f / V.fy; F.mbf = V.bf; F.mb = V.b;
    F.mpCamera = camera(V.camModel, V.fx, V.fy, V.cx, V.cy, V.kb);
    F.mpCamera2 = V.cam2Model >= 0 ? camera(V.cam2Model, V.cam2[0], V.cam2[1], V.cam2[2], V.cam2[3], V.cam2 + 4) : nullptr;
    if (V.Tcw) { F.mTcw = mat_from(V.Tcw, 4, 4); F.UpdatePoseMatrices(); }
    if (V.Tlr) {   // Frame.cc:1186-1195
      F.mTlr = mat_from(V.Tlr, 3, 4);
      F.mRlr = F.mTlr.rowRange(0, 3).colRange(0, 3);
      F.mtlr = F.mTlr.col(3);
      cv::Mat Rrl = F.mTlr.rowRange(0, 3).colRange(0, 3).t();
      cv::Mat trl = Rrl * (-1 * F.mTlr.col(3));
      F.mTrl = cv::Mat(3, 4, CV_32F);
      for (int y = 0; y < 3; y++) { for (int x = 0; x < 3; x++) F.mTrl.at<float>(y, x) = Rrl.at<float>(y, x); F.mTrl.at<float>(y, 3) = trl.at<float>(y); }
    }
    if (V.Nleft != -1) {
      F.mvLeftToRightMatch.assign(V.Nleft, -1); F.mvRightToLeftMatch.assign(V.N - V.Nleft, -1);
      if (V.leftToRight) F.mvLeftToRightMatch.assign(V.leftToRight, V.leftToRight + V.Nleft);
      if (V.rightToLeft) F.mvRightToLeftMatch.assign(V.rightToLeft, V.rightToLeft + (V.N - V.Nleft));
    }
    F.mvpMapPoints.assign(V.N, nullptr);
    if (V.mapPointIdx) for (int i = 0; i < V.N; i++) F.mvpMapPoints[i] = mp(V.mapPointIdx[i]);
    F.mvbOutlier.assign(V.N, false);
    if (V.outlier) for (int i = 0; i < V.N; i++) F.mvbOutlier[i] = V.outlier[i] != 0;
    for (int n = 0; n < V.fvNodes; n++) {
      std::vector<unsigned int>& v = F.mFeatVec[V.fvIds[n]];
      for (int k = V.fvStart[n]; k < V.fvStart[n + 1]; k++) v.push_back(V.fvIdx[k]);
    }
    F.AssignFeaturesToGrid();
  }

  // KeyFrame::KeyFrame(Frame&, ...) copies these members from the frame (src/KeyFrame.cc:41-100)
  KeyFrame* keyframe(const RefView& V, std::vector<Frame*>& framesOwned, std::vector<KeyFrame*>& kfsOwned) {
    Frame* F = new Frame();
    framesOwned.push_back(F);
    fill_common(V, *F);
    KeyFrame* K = new KeyFrame();
    kfsOwned.push_back(K);
    K->mnId = kfsOwned.size();
    K->N = F->N; K->NLeft = F->Nleft; K->NRight = F->Nright;
    K->fx = F->fx; K->fy = F->fy; K->cx = F->cx; K->cy = F->cy; K->invfx = F->invfx; K->invfy = F->invfy; K->mbf = F->mbf; K->mb = F->mb;
    K->mvKeys = F->mvKeys; K->mvKeysUn = F->mvKeysUn; K->mvKeysRight = F->mvKeysRight; K->mvuRight = F->mvuRight; K->mvDepth = F->mvDepth;
    K->mDescriptors = F->mDescriptors.clone(); K->mFeatVec = F->mFeatVec;
    K->mnScaleLevels = F->mnScaleLevels; K->mfScaleFactor = F->mfScaleFactor; K->mfLogScaleFactor = F->mfLogScaleFactor;
    K->mvScaleFactors = F->mvScaleFactors; K->mvLevelSigma2 = F->mvLevelSigma2; K->mvInvLevelSigma2 = F->mvInvLevelSigma2;
    K->mnMinX = F->mnMinX; K->mnMinY = F->mnMinY; K->mnMaxX = F->mnMaxX; K->mnMaxY = F->mnMaxY;   // const int members: truncation as in the reference
    K->mfGridElementWidthInv = F->mfGridElementWidthInv; K->mfGridElementHeightInv = F->mfGridElementHeightInv;
    K->mGrid.resize(K->mnGridCols); K->mGridRight.resize(K->mnGridCols);
    for (int i = 0; i < K->mnGridCols; i++) {
      K->mGrid[i].resize(K->mnGridRows); K->mGridRight[i].resize(K->mnGridRows);
      for (int j = 0; j < K->mnGridRows; j++) { K->mGrid[i][j] = F->mGrid[i][j]; K->mGridRight[i][j] = F->mGridRight[i][j]; }
    }
    K->mpCamera = F->mpCamera; K->mpCamera2 = F->mpCamera2;
    K->mvpMapPoints = F->mvpMapPoints;
    if (V.Tcw) {   // KeyFrame::SetPose, src/KeyFrame.cc:118-140
      K->Tcw = F->mTcw.clone();
      cv::Mat Rcw = K->Tcw.rowRange(0, 3).colRange(0, 3);
      cv::Mat tcw = K->Tcw.rowRange(0, 3).col(3);
      cv::Mat Rwc = Rcw.t();
      K->Ow = -Rwc * tcw;
      if (V.Tlr) {   // KeyFrame::GetRightPose / GetRightCameraCenter, src/KeyFrame.cc (mTrl * Tcw, Rwc * tlr + Ow)
        cv::Mat Trl4 = cv::Mat::eye(4, 4, CV_32F);
        for (int y = 0; y < 3; y++) for (int x = 0; x < 4; x++) Trl4.at<float>(y, x) = F->mTrl.at<float>(y, x);
        K->TcwRight = Trl4 * K->Tcw;
        cv::Mat Rwl = K->Tcw.rowRange(0, 3).colRange(0, 3).t();
        cv::Mat tlr = F->mTlr.col(3);
        K->OwRight = Rwl * tlr + K->Ow;
      }
    }
    for (size_t i = 0; i < K->mvpMapPoints.size(); i++)
      if (K->mvpMapPoints[i]) K->mvpMapPoints[i]->mObservations[K] = tuple<int, int>((int)i, -1);
    return K;
  }
};

struct Owned {
  std::vector<Frame*> frames;
  std::vector<KeyFrame*> kfs;
  ~Owned() { for (auto* f : frames) delete f; for (auto* k : kfs) delete k; }
};

void export_matches(const std::vector<MapPoint*>& v, const World& W, int* out) {
  for (size_t i = 0; i < v.size(); i++) out[i] = v[i] ? (int)(v[i] - &W.mps[0]) : -1;
}

}  // namespace

extern "C" {

// Frame::GetFeaturesInArea (src/Frame.cc:744-813) after AssignFeaturesToGrid; returns the count, indices in the reference's order.
int rm_features_in_area(const RefView* V, float x, float y, float r, int minLevel, int maxLevel, int bRight, int* out, int cap) {
  World W; Frame F;
  W.fill_common(*V, F);
  std::vector<size_t> v = F.GetFeaturesInArea(x, y, r, minLevel, maxLevel, bRight != 0);
  for (size_t i = 0; i < v.size() && (int)i < cap; i++) out[i] = (int)v[i];
  return (int)v.size();
}

// Frame::isInFrustum (src/Frame.cc:572-661) for every map point: the MapPoint tracking fields it writes.  The *R outputs (may be
// null) are the right-camera fields of a fisheye-stereo frame (Nleft != -1).
void rm_is_in_frustum(const RefView* V, const RefPoints* P, float viewingCosLimit, uint8_t* inView, float* projX, float* projY, float* projXR,
                      float* depth, int* level, float* viewCos, uint8_t* inViewR, float* projXRr, float* projYR, float* depthR, int* levelR,
                      float* viewCosR) {
  World W; Frame F;
  W.points(P);
  W.fill_common(*V, F);
  for (int i = 0; i < P->n; i++) {
    MapPoint& m = W.mps[i];
    m.mnTrackScaleLevel = -1; m.mnTrackScaleLevelR = -1;
    F.isInFrustum(&m, viewingCosLimit);
    inView[i] = m.mbTrackInView; projX[i] = m.mTrackProjX; projY[i] = m.mTrackProjY; projXR[i] = m.mTrackProjXR; depth[i] = m.mTrackDepth;
    level[i] = m.mnTrackScaleLevel; viewCos[i] = m.mTrackViewCos;
    if (inViewR) { inViewR[i] = m.mbTrackInViewR; projXRr[i] = m.mTrackProjXR; projYR[i] = m.mTrackProjYR; depthR[i] = m.mTrackDepthR; levelR[i] = m.mnTrackScaleLevelR; viewCosR[i] = m.mTrackViewCosR; }
  }
}

// Tracking::SearchLocalPoints' core: isInFrustum for every map point (optional: run = 1) or the given tracking fields (run = 0),
// then ORBmatcher::SearchByProjection(F, vpMapPoints, th, bFarPoints, thFarPoints) (src/ORBmatcher.cc:44-222).
// kpMp (in/out, N): F.mvpMapPoints as indices.  Returns nmatches.
int rm_search_by_projection_mappoints(const RefView* V, const RefPoints* P, int runFrustum, float viewingCosLimit, const uint8_t* inView,
                                      const float* projX, const float* projY, const float* projXR, const int* level, const float* viewCos,
                                      const float* trackDepth, float th, float nnratio, int bFarPoints, float thFarPoints, int* kpMp) {
  World W; Frame F;
  W.points(P);
  RefView v2 = *V; v2.mapPointIdx = kpMp;
  W.fill_common(v2, F);
  std::vector<MapPoint*> vp;
  for (int i = 0; i < P->n; i++) {
    MapPoint& m = W.mps[i];
    if (runFrustum) F.isInFrustum(&m, viewingCosLimit);
    else {
      m.mbTrackInView = inView[i] != 0; m.mTrackProjX = projX[i]; m.mTrackProjY = projY[i]; m.mTrackProjXR = projXR ? projXR[i] : 0.f;
      m.mnTrackScaleLevel = level[i]; m.mTrackViewCos = viewCos[i]; m.mTrackDepth = trackDepth ? trackDepth[i] : 0.f;
    }
    vp.push_back(&m);
  }
  ORBmatcher matcher(nnratio, true);
  const int nm = matcher.SearchByProjection(F, vp, th, bFarPoints != 0, thFarPoints);
  export_matches(F.mvpMapPoints, W, kpMp);
  return nm;
}

// The same call on a fisheye-stereo frame (V->Nleft != -1) with the tracking fields of BOTH cameras given (the fields
// Frame::isInFrustumChecks writes, src/Frame.cc:1324-1339): L / R = in_view, proj_x, proj_y, level, view_cos per map point.
int rm_search_by_projection_mappoints_stereo(const RefView* V, const RefPoints* P, const uint8_t* inL, const float* pxL, const float* pyL,
                                             const int* lvL, const float* vcL, const uint8_t* inR, const float* pxR, const float* pyR,
                                             const int* lvR, const float* vcR, const float* trackDepth, float th, float nnratio, int bFarPoints,
                                             float thFarPoints, int* kpMp) {
  World W; Frame F;
  W.points(P);
  RefView v2 = *V; v2.mapPointIdx = kpMp;
  W.fill_common(v2, F);
  std::vector<MapPoint*> vp;
  for (int i = 0; i < P->n; i++) {
    MapPoint& m = W.mps[i];
    m.mbTrackInView = inL[i] != 0; m.mTrackProjX = pxL[i]; m.mTrackProjY = pyL[i]; m.mnTrackScaleLevel = lvL[i]; m.mTrackViewCos = vcL[i];
    m.mbTrackInViewR = inR[i] != 0; m.mTrackProjXR = pxR[i]; m.mTrackProjYR = pyR[i]; m.mnTrackScaleLevelR = lvR[i]; m.mTrackViewCosR = vcR[i];
    m.mTrackDepth = trackDepth ? trackDepth[i] : 0.f;
    vp.push_back(&m);
  }
  ORBmatcher matcher(nnratio, true);
  const int nm = matcher.SearchByProjection(F, vp, th, bFarPoints != 0, thFarPoints);
  export_matches(F.mvpMapPoints, W, kpMp);
  return nm;
}

// ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, th, bMono), src/ORBmatcher.cc:2027-2289.
// Both views carry Tcw; Last carries mapPointIdx / outlier.  kpMp (out, Cur->N).
int rm_search_by_projection_frame(const RefView* Cur, const RefView* Last, const RefPoints* P, float th, int bMono, float nnratio, int checkOri, int* kpMp) {
  World W; Frame FC, FL;
  W.points(P);
  W.fill_common(*Cur, FC);
  W.fill_common(*Last, FL);
  ORBmatcher matcher(nnratio, checkOri != 0);
  const int nm = matcher.SearchByProjection(FC, FL, th, bMono != 0);
  export_matches(FC.mvpMapPoints, W, kpMp);
  return nm;
}

// ORBmatcher::SearchByProjection(Frame&, KeyFrame*, const set<MapPoint*>& sAlreadyFound, th, ORBdist), src/ORBmatcher.cc:2291-2413.
int rm_search_by_projection_keyframe(const RefView* Cur, const RefView* KF, const RefPoints* P, const uint8_t* alreadyFound, float th, int ORBdist,
                                     float nnratio, int checkOri, int* kpMp) {
  World W; Frame FC; Owned O;
  W.points(P);
  W.fill_common(*Cur, FC);
  KeyFrame* K = W.keyframe(*KF, O.frames, O.kfs);
  std::set<MapPoint*> found;
  if (alreadyFound) for (int i = 0; i < P->n; i++) if (alreadyFound[i]) found.insert(&W.mps[i]);
  ORBmatcher matcher(nnratio, checkOri != 0);
  const int nm = matcher.SearchByProjection(FC, K, found, th, ORBdist);
  export_matches(FC.mvpMapPoints, W, kpMp);
  return nm;
}

// ORBmatcher::SearchByProjection(KeyFrame*, cv::Mat Scw, vpPoints, vpMatched, th, ratioHamming), src/ORBmatcher.cc:489-602, and the
// overload with vpPointsKFs / vpMatchedKF (:604-720; withKFs != 0: every point's keyframe is reported as its index).
// matched (in/out, KF->N): indices into P or -1.
int rm_search_by_projection_scw(const RefView* KF, const float* Scw, const RefPoints* P, int th, float ratioHamming, int withKFs, int* matched,
                                int* matchedKF) {
  World W; Owned O;
  W.points(P);
  KeyFrame* K = W.keyframe(*KF, O.frames, O.kfs);
  std::vector<MapPoint*> vp, vm(KF->N, nullptr);
  for (int i = 0; i < P->n; i++) vp.push_back(&W.mps[i]);
  for (int i = 0; i < KF->N; i++) vm[i] = W.mp(matched[i]);
  ORBmatcher matcher(0.75f, true);
  int nm;
  if (withKFs) {
    std::vector<KeyFrame*> pkfs(P->n), mkfs(KF->N, nullptr);
    std::vector<KeyFrame> dummy(P->n);
    for (int i = 0; i < P->n; i++) pkfs[i] = &dummy[i];
    nm = matcher.SearchByProjection(K, mat_from(Scw, 4, 4), vp, pkfs, vm, mkfs, th, ratioHamming);
    if (matchedKF) for (int i = 0; i < KF->N; i++) matchedKF[i] = mkfs[i] ? (int)(mkfs[i] - &dummy[0]) : -1;
  } else nm = matcher.SearchByProjection(K, mat_from(Scw, 4, 4), vp, vm, th, ratioHamming);
  export_matches(vm, W, matched);
  return nm;
}

// ORBmatcher::SearchForInitialization, src/ORBmatcher.cc:722-837.
int rm_search_for_initialization(const RefView* V1, const RefView* V2, float* prevMatched, int windowSize, float nnratio, int checkOri, int* matches12) {
  World W; Frame F1, F2;
  W.fill_common(*V1, F1);
  W.fill_common(*V2, F2);
  std::vector<cv::Point2f> pm(V1->N);
  for (int i = 0; i < V1->N; i++) pm[i] = cv::Point2f(prevMatched[2 * i], prevMatched[2 * i + 1]);
  std::vector<int> m12;
  ORBmatcher matcher(nnratio, checkOri != 0);
  const int nm = matcher.SearchForInitialization(F1, F2, pm, m12, windowSize);
  for (int i = 0; i < V1->N; i++) { matches12[i] = m12[i]; prevMatched[2 * i] = pm[i].x; prevMatched[2 * i + 1] = pm[i].y; }
  return nm;
}

// ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches), src/ORBmatcher.cc:273-475.  out (F->N): map-point indices.
int rm_search_by_bow_frame(const RefView* KF, const RefView* Fv, const RefPoints* P, float nnratio, int checkOri, int* out) {
  World W; Frame F; Owned O;
  W.points(P);
  KeyFrame* K = W.keyframe(*KF, O.frames, O.kfs);
  W.fill_common(*Fv, F);
  std::vector<MapPoint*> vm;
  ORBmatcher matcher(nnratio, checkOri != 0);
  const int nm = matcher.SearchByBoW(K, F, vm);
  export_matches(vm, W, out);
  return nm;
}

// ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12), src/ORBmatcher.cc:839-979.  out (KF1->N): map-point indices of pKF2's points.
int rm_search_by_bow_keyframes(const RefView* KF1, const RefView* KF2, const RefPoints* P, float nnratio, int checkOri, int* out) {
  World W; Owned O;
  W.points(P);
  KeyFrame* K1 = W.keyframe(*KF1, O.frames, O.kfs);
  KeyFrame* K2 = W.keyframe(*KF2, O.frames, O.kfs);
  std::vector<MapPoint*> vm;
  ORBmatcher matcher(nnratio, checkOri != 0);
  const int nm = matcher.SearchByBoW(K1, K2, vm);
  export_matches(vm, W, out);
  return nm;
}

// ---- the stand-in cv:: arithmetic, exposed so that tests/test_oracle_cvmat.py can hold it against python cv2 -------------------
// MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:350-436) for npoints map points.  The observations of point p are the
// keyframes start[p] .. start[p+1]-1; kind[k] = 0: keyframe k sees the point with its left feature only (one descriptor row), 1: the
// keyframe is bad (one row, skipped by the reference), 2: left and right feature (two consecutive rows).  Rows are consumed in that
// order from desc.  The keyframes of a point live in one array, so the reference's map<KeyFrame*, ..> (ordered by ADDRESS) iterates
// them in index order.  out: the chosen mDescriptor (32 bytes per point), has[p] = 0 where the reference leaves mDescriptor empty.
void rm_distinctive(const uint8_t* desc, const int* start, const uint8_t* kind, int npoints, uint8_t* out, int* has) {
  size_t row = 0;
  for (int p = 0; p < npoints; p++) {
    const int n = start[p + 1] - start[p];
    std::vector<KeyFrame> kfs(n);
    MapPoint mp;
    for (int k = 0; k < n; k++) {
      const int kd = kind[start[p] + k];
      const int rows = kd == 2 ? 2 : 1;
      kfs[k].mbBad = kd == 1;
      kfs[k].mDescriptors = cv::Mat(rows, 32, CV_8U);
      for (int r = 0; r < rows; r++) memcpy(kfs[k].mDescriptors.ptr(r), desc + (row + r) * 32, 32);
      row += rows;
      mp.mObservations[&kfs[k]] = std::tuple<int, int>(0, kd == 2 ? 1 : -1);
    }
    mp.ComputeDistinctiveDescriptors();
    has[p] = mp.mDescriptor.empty() ? 0 : 1;
    if (has[p]) memcpy(out + (size_t)p * 32, mp.mDescriptor.ptr(0), 32);
  }
}

void rm_cv_svd(const float* A, int n, float* w, float* vt) {   // cv::SVD::compute(A, w, u, vt, MODIFY_A | FULL_UV) of an n x n matrix
  cv::Mat a = mat_from(A, n, n), W, U, Vt;
  cv::SVD::compute(a, W, U, Vt, cv::SVD::MODIFY_A | cv::SVD::FULL_UV);
  for (int i = 0; i < n; i++) { w[i] = W.at<float>(i); for (int k = 0; k < n; k++) vt[i * n + k] = Vt.at<float>(i, k); }
}
void rm_cv_scaled_sub_row(const float* a, float s, const float* b, int cols, float* out) {   // M.row(0) = s*A - B written into a row view
  cv::Mat A = mat_from(a, 1, cols), B = mat_from(b, 1, cols), M = cv::Mat::zeros(2, cols, CV_32F);
  M.row(0) = s * A - B;
  for (int k = 0; k < cols; k++) out[k] = M.at<float>(0, k);
}
void rm_cv_div_scalar(const float* a, int n, float s, float* out) {   // Mat / s
  cv::Mat A = mat_from(a, n, 1);
  cv::Mat r = A / s;
  for (int k = 0; k < n; k++) out[k] = r.at<float>(k);
}
void rm_cv_gemm(const float* a, int ar, int ac, const float* b, int br, int bc, double alpha, const float* c, double beta, int flags, float* out) {
  cv::Mat A = mat_from(a, ar, ac), B = mat_from(b, br, bc), Cm, D;
  const int M = (flags & cv::GEMM_1_T) ? ac : ar, N = (flags & cv::GEMM_2_T) ? br : bc;
  if (c) Cm = mat_from(c, (flags & cv::GEMM_3_T) ? N : M, (flags & cv::GEMM_3_T) ? M : N);
  cv::cvshim_gemm(A, B, alpha, Cm, beta, D, flags);
  for (int i = 0; i < M; i++) for (int k = 0; k < N; k++) out[i * N + k] = D.at<float>(i, k);
}
double rm_cv_dot(const float* a, const float* b, int n) { return mat_from(a, n, 1).dot(mat_from(b, n, 1)); }
double rm_cv_norm(const float* a, int n) { return cv::norm(mat_from(a, n, 1)); }

// KannalaBrandt8::unproject (KannalaBrandt8.cpp:112-137) and ::TriangulateMatches (:343-418), the reference's bodies: cam = fx fy cx cy
// k1..k4, n independent keypoint pairs; z1 (or -1) and the accepted point per pair.
void rm_kb8_unproject(const float* cam, int n, const float* xy, float* rays) {
  KannalaBrandt8 K(std::vector<float>(cam, cam + 8));
  for (int i = 0; i < n; i++) {
    const cv::Point3f r = K.unproject(cv::Point2f(xy[2 * i], xy[2 * i + 1]));
    rays[3 * i] = r.x; rays[3 * i + 1] = r.y; rays[3 * i + 2] = r.z;
  }
}
void rm_kb8_triangulate_matches(const float* cam1, const float* cam2, const float* R12, const float* t12, int n, const float* xy1,
                                const float* xy2, const float* sigma1, const float* sigma2, float* z1, float* p3D) {
  KannalaBrandt8 K1(std::vector<float>(cam1, cam1 + 8)), K2(std::vector<float>(cam2, cam2 + 8));
  const cv::Mat R = mat_from(R12, 3, 3), t = mat_from(t12, 3, 1);
  for (int i = 0; i < n; i++) {
    cv::KeyPoint kp1, kp2;
    kp1.pt = cv::Point2f(xy1[2 * i], xy1[2 * i + 1]); kp2.pt = cv::Point2f(xy2[2 * i], xy2[2 * i + 1]);
    cv::Mat X;
    z1[i] = K1.TriangulateMatches(&K2, kp1, kp2, R, t, sigma1[i], sigma2[i], X);
    for (int k = 0; k < 3; k++) p3D[3 * i + k] = X.empty() ? 0.f : X.at<float>(k);
  }
}

// ORBmatcher::SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo, bCoarse), src/ORBmatcher.cc:981-1222.  Pinhole
// cameras: the epipolar test uses the F12 given here (see slam_world.h); KannalaBrandt8 cameras (one per keyframe or a fisheye-stereo
// rig) run the reference's own triangulation gate and ignore F12.  pairs: up to cap (i1, i2) pairs; returns nmatches.
int rm_search_for_triangulation(const RefView* KF1, const RefView* KF2, const RefPoints* P, const float* F12, int bOnlyStereo, int bCoarse, float nnratio,
                                int checkOri, int* pairs, int cap) {
  World W; Owned O;
  W.points(P);
  KeyFrame* K1 = W.keyframe(*KF1, O.frames, O.kfs);
  KeyFrame* K2 = W.keyframe(*KF2, O.frames, O.kfs);
  cv::Mat F = mat_from(F12, 3, 3);
  if (Pinhole* ph = dynamic_cast<Pinhole*>(K1->mpCamera)) ph->F12given = F;
  std::vector<pair<size_t, size_t> > vp;
  ORBmatcher matcher(nnratio, checkOri != 0);
  const int nm = matcher.SearchForTriangulation(K1, K2, F, vp, bOnlyStereo != 0, bCoarse != 0);
  for (size_t i = 0; i < vp.size() && (int)i < cap; i++) { pairs[2 * i] = (int)vp[i].first; pairs[2 * i + 1] = (int)vp[i].second; }
  return nm;
}

// ORBmatcher::Fuse(pKF, vpMapPoints, th, bRight), src/ORBmatcher.cc:1425-1662.  events: up to cap records (kind, point, other point or
// -1, keypoint index or -1): kind 0 = point->Replace(other), 1 = AddObservation / AddMapPoint(point, idx).  Returns nFused.
static int export_events(const World& W, int* events, int cap) {
  int n = 0;
  for (const WorldEvent& e : W.log) {
    if (n >= cap) break;
    events[4 * n] = e.kind; events[4 * n + 1] = (int)(e.a - &W.mps[0]); events[4 * n + 2] = e.b ? (int)(e.b - &W.mps[0]) : -1; events[4 * n + 3] = e.idx;
    n++;
  }
  return n;
}
int rm_fuse(const RefView* KF, const RefPoints* P, const int* candidates, int ncand, float th, int bRight, int* events, int cap, int* nevents) {
  World W; Owned O;
  W.points(P);
  KeyFrame* K = W.keyframe(*KF, O.frames, O.kfs);
  std::vector<MapPoint*> vp;
  for (int i = 0; i < ncand; i++) vp.push_back(W.mp(candidates[i]));
  ORBmatcher matcher(0.6f, true);
  const int nf = matcher.Fuse(K, vp, th, bRight != 0);
  *nevents = export_events(W, events, cap);
  return nf;
}
// ORBmatcher::Fuse(pKF, Scw, vpPoints, th, vpReplacePoint), src/ORBmatcher.cc:1664-1786.  replace (ncand): index of the KeyFrame's
// map point to be replaced by candidate i, or -1.
int rm_fuse_scw(const RefView* KF, const float* Scw, const RefPoints* P, const int* candidates, int ncand, float th, int* replace, int* events, int cap,
                int* nevents) {
  World W; Owned O;
  W.points(P);
  KeyFrame* K = W.keyframe(*KF, O.frames, O.kfs);
  std::vector<MapPoint*> vp, vr(ncand, nullptr);
  for (int i = 0; i < ncand; i++) vp.push_back(W.mp(candidates[i]));
  ORBmatcher matcher(0.6f, true);
  const int nf = matcher.Fuse(K, mat_from(Scw, 4, 4), vp, th, vr);
  export_matches(vr, W, replace);
  *nevents = export_events(W, events, cap);
  return nf;
}

// ORBmatcher::SearchBySim3, src/ORBmatcher.cc:1788-2018.  matches12 (in/out, KF1->N): map-point indices (points of pKF2).
int rm_search_by_sim3(const RefView* KF1, const RefView* KF2, const RefPoints* P, float s12, const float* R12, const float* t12, float th, int* matches12) {
  World W; Owned O;
  W.points(P);
  KeyFrame* K1 = W.keyframe(*KF1, O.frames, O.kfs);
  KeyFrame* K2 = W.keyframe(*KF2, O.frames, O.kfs);
  std::vector<MapPoint*> vm(KF1->N, nullptr);
  for (int i = 0; i < KF1->N; i++) vm[i] = W.mp(matches12[i]);
  ORBmatcher matcher(0.75f, true);
  const int nm = matcher.SearchBySim3(K1, K2, vm, s12, mat_from(R12, 3, 3), mat_from(t12, 3, 1), th);
  export_matches(vm, W, matches12);
  return nm;
}

#ifndef ORBX_SHIM_TWIN   // the stereo frame entry points time / pin the reference only
// Whole stereo frame as Frame::Frame(stereo) does the data-parallel part (src/Frame.cc:89-187): both extractions with the
// reference's ORBextractor (vLappingArea {0,0}) and Frame::ComputeStereoMatches (src/Frame.cc:901-1079).
// Outputs: keypoints / descriptors of both images (cap records each), uRight / depth (cap floats).
struct RefStereo { ORBextractor* L; ORBextractor* R; int nfeatures; };
void* rm_stereo_create(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh) {
  RefStereo* s = new RefStereo();
  s->L = new ORBextractor(nfeatures, scaleFactor, nlevels, iniTh, minTh);
  s->R = new ORBextractor(nfeatures, scaleFactor, nlevels, iniTh, minTh);
  s->nfeatures = nfeatures;
  return s;
}
void rm_stereo_destroy(void* h) { RefStereo* s = (RefStereo*)h; delete s->L; delete s->R; delete s; }
// The two extractions allocate from one bump arena each (ref_alloc.h); the arenas stay bound to the frame until it is consumed.
struct FrameArenas {
  refalloc::Arena *aL, *aR;
  FrameArenas() : aL(refalloc::acquire()), aR(refalloc::acquire()) {}
  ~FrameArenas() { refalloc::unbind(); refalloc::release(aL); refalloc::release(aR); }
  bool ok() const { return aL && aR; }
};
static void stereo_frame(RefStereo* s, FrameArenas& A, const uint8_t* left, const uint8_t* right, int w, int hgt, int stride, float mb, float mbf, int blurTaps,
                         bool twoThreads, Frame& F) {
  std::vector<int> lap{0, 0};
  cv::Mat imL(hgt, w, CV_8UC1, (void*)left, (size_t)stride), imR(hgt, w, CV_8UC1, (void*)right, (size_t)stride);
  cv::Mat descR;
  auto runR = [&]() {
    cvshim::g_blur_taps = blurTaps;
    refalloc::bind(A.aR);
    std::vector<int> lapR{0, 0};
    (*s->R)(imR, cv::Mat(), F.mvKeysRight, descR, lapR);
    refalloc::unbind();
  };
  cvshim::g_blur_taps = blurTaps;
  if (twoThreads) {   // Frame.cc:120-123
    std::thread tr(runR);
    refalloc::bind(A.aL);
    (*s->L)(imL, cv::Mat(), F.mvKeys, F.mDescriptors, lap);
    tr.join();
  } else {
    runR();
    refalloc::bind(A.aL);
    (*s->L)(imL, cv::Mat(), F.mvKeys, F.mDescriptors, lap);
  }
  F.mDescriptorsRight = descR;
  F.N = (int)F.mvKeys.size();
  F.mpORBextractorLeft = s->L; F.mpORBextractorRight = s->R;
  F.mvScaleFactors = s->L->GetScaleFactors(); F.mvInvScaleFactors = s->L->GetInverseScaleFactors();
  F.mb = mb; F.mbf = mbf;
  F.ComputeStereoMatches();
}
int rm_stereo_frame(void* h, const uint8_t* left, const uint8_t* right, int w, int hgt, int stride, float mb, float mbf, int blurTaps, void* kpsL,
                    uint8_t* descL, int* nL, void* kpsR, uint8_t* descR, int* nR, float* uRight, float* depth, int cap) {
  RefStereo* s = (RefStereo*)h;
  FrameArenas A;
  if (!A.ok()) return -3;
  Frame F;
  stereo_frame(s, A, left, right, w, hgt, stride, mb, mbf, blurTaps, false, F);
  *nL = F.N; *nR = (int)F.mvKeysRight.size();
  if (F.N > cap || *nR > cap) return -2;
  if (F.N) memcpy(kpsL, F.mvKeys.data(), (size_t)F.N * sizeof(cv::KeyPoint));
  for (int i = 0; i < F.N; i++) { memcpy(descL + 32 * (size_t)i, F.mDescriptors.ptr(i), 32); uRight[i] = F.mvuRight[i]; depth[i] = F.mvDepth[i]; }
  if (*nR) memcpy(kpsR, F.mvKeysRight.data(), (size_t)*nR * sizeof(cv::KeyPoint));
  for (int i = 0; i < *nR; i++) memcpy(descR + 32 * (size_t)i, F.mDescriptorsRight.ptr(i), 32);
  return 0;
}
// ComputeStereoMatches alone on given keypoints / descriptors, reading the pyramids the two extractors hold from their last call.
void rm_compute_stereo_matches(void* h, int N, const cv::KeyPoint* kpL, const uint8_t* descL, int Nr, const cv::KeyPoint* kpR, const uint8_t* descR,
                               float mb, float mbf, float* uRight, float* depth) {
  RefStereo* s = (RefStereo*)h;
  Frame F;
  F.N = N; F.mvKeys.assign(kpL, kpL + N); F.mvKeysRight.assign(kpR, kpR + Nr);
  F.mDescriptors = desc_rows(descL, N); F.mDescriptorsRight = desc_rows(descR, Nr);
  F.mpORBextractorLeft = s->L; F.mpORBextractorRight = s->R;
  F.mvScaleFactors = s->L->GetScaleFactors(); F.mvInvScaleFactors = s->L->GetInverseScaleFactors();
  F.mb = mb; F.mbf = mbf;
  F.ComputeStereoMatches();
  for (int i = 0; i < N; i++) { uRight[i] = F.mvuRight[i]; depth[i] = F.mvDepth[i]; }
}
// Timed reference arm of bench.py: frame-parallel workers, each builds stereo frames with two threads per frame exactly as
// Frame::Frame(stereo) does.  nthreads = total threads.  Returns seconds; *matched = number of keypoints with a stereo match.
double rm_time_stereo(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh, int blurTaps, const uint8_t* left, const uint8_t* right,
                      int nframes, int w, int hgt, float mb, float mbf, int nthreads, long long* matched) {
  const int workers = std::max(1, nthreads / 2);
  std::vector<std::thread> th;
  std::vector<long long> counts(workers, 0);
  auto t0 = std::chrono::steady_clock::now();
  for (int t = 0; t < workers; t++)
    th.emplace_back([&, t]() {
      RefStereo* s = (RefStereo*)rm_stereo_create(nfeatures, scaleFactor, nlevels, iniTh, minTh);
      for (int f = t; f < nframes; f += workers) {
        FrameArenas A;
        Frame F;
        stereo_frame(s, A, left + (size_t)f * w * hgt, right + (size_t)f * w * hgt, w, hgt, w, mb, mbf, blurTaps, true, F);
        for (int i = 0; i < F.N; i++) counts[t] += F.mvuRight[i] >= 0;
      }
      rm_stereo_destroy(s);
    });
  for (auto& x : th) x.join();
  auto t1 = std::chrono::steady_clock::now();
  long long tot = 0;
  for (auto c : counts) tot += c;
  if (matched) *matched = tot;
  return std::chrono::duration<double>(t1 - t0).count();
}

#endif  // ORBX_SHIM_TWIN

}  // extern "C"
