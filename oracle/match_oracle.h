// CPU ORACLE (test infrastructure, not product code) -- matcher side.
// Restates, on plain arrays, the Hamming searches of the reference:
//   Frame::AssignFeaturesToGrid / PosInGrid / GetFeaturesInArea   src/Frame.cc:434-465, 815-825, 744-813
//   Frame::ComputeStereoMatches                                   src/Frame.cc:901-1079
//   Frame::ComputeStereoFishEyeMatches (kNN + ratio part)         src/Frame.cc:1228-1268
//   ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&..)  src/ORBmatcher.cc:44-214   (mono / Nleft==-1 branch)
//   ORBmatcher::SearchByProjection(Frame&, const Frame&, ..)      src/ORBmatcher.cc:2027-2289 (after projection)
//   ORBmatcher::SearchByProjection(Frame&, KeyFrame*, set, ..)    src/ORBmatcher.cc:2291-2413 (after projection)
//   ORBmatcher::SearchForInitialization                           src/ORBmatcher.cc:722-837
//   ORBmatcher::ComputeThreeMaxima                                src/ORBmatcher.cc:2416-2458
//   ORBmatcher::SearchByBoW(KeyFrame*, Frame&, ..)                src/ORBmatcher.cc:273-475   (Nleft == -1 branch)
//   ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*, ..)             src/ORBmatcher.cc:839-979
//   ORBmatcher::SearchForTriangulation(.., bOnlyStereo, bCoarse)  src/ORBmatcher.cc:981-1222  (pinhole, no second camera)
//   Pinhole::epipolarConstrain                                    src/CameraModels/Pinhole.cpp:143-165 (F12 given)
//   ORBmatcher::SearchByProjection(KeyFrame*, Scw, ..) x2         src/ORBmatcher.cc:489-602, 604-720 (after projection)
//   ORBmatcher::Fuse(KeyFrame*, vector<MapPoint*>, th, bRight)    src/ORBmatcher.cc:1425-1662 (search loop, after projection)
//   ORBmatcher::Fuse(KeyFrame*, Scw, ..)                          src/ORBmatcher.cc:1664-1786 (search loop, after projection)
//   ORBmatcher::SearchBySim3                                      src/ORBmatcher.cc:1788-2018 (search loops + agreement)
// Pose algebra / camera projection / MapPoint bookkeeping stay with the caller (out of scope); the
// functions here take the already-projected (u,v), predicted level etc. exactly as the reference computes them
// before entering the search loop.  Parity status: unpinned by the reference (it has no tests), see orb_oracle.h.
#pragma once
#include <cstdint>
#include <map>
#include <vector>

#include "orb_oracle.h"

namespace oro {

static const int FRAME_GRID_ROWS = 48;  // include/Frame.h:38
static const int FRAME_GRID_COLS = 64;  // include/Frame.h:39
static const int TH_HIGH = 100, TH_LOW = 50, HISTO_LENGTH = 30;  // src/ORBmatcher.cc:36-38

struct FrameView {
  int N = 0;
  const KeyPoint* kps = nullptr;   // mvKeysUn (pt, octave, angle)
  const uint8_t* desc = nullptr;   // mDescriptors, N x 32
  const float* uRight = nullptr;   // mvuRight or null (monocular)
  const float* scaleFactors = nullptr;
  float minX = 0, minY = 0, maxX = 0, maxY = 0;  // mnMinX.. (image bounds)
  float invW = 0, invH = 0;                      // mfGridElementWidthInv / HeightInv
  std::vector<int> grid[FRAME_GRID_COLS][FRAME_GRID_ROWS];
  void setBounds(float mnX, float mnY, float mxX, float mxY);
  void assignFeaturesToGrid();
  std::vector<int> getFeaturesInArea(float x, float y, float r, int minLevel, int maxLevel) const;
};

void compute_three_maxima(const int* hist_sizes, int L, int& ind1, int& ind2, int& ind3);

// SearchByProjection(Frame&, vector<MapPoint*>&, th, ...) -- local-map tracking.  Per map point i: in_view[i]
// (mbTrackInView && !isBad && !far), projX/projY/level/viewCos, 32-byte descriptor, observed[i] (Observations()>0),
// projXR (mTrackProjXR; used only when F.uRight != null).  kp_mp[] is F.mvpMapPoints as indices: -1 empty,
// -2 = pre-existing map point with observations (blocks), -3 = pre-existing map point without observations (does not block).
// Returns nmatches.
int search_by_projection_mappoints_stereo(FrameView& FL, FrameView& FR, int nMP, const uint8_t* inL, const float* pxL, const float* pyL,
                                          const int* lvL, const float* vcL, const uint8_t* inR, const float* pxR, const float* pyR,
                                          const int* lvR, const float* vcR, const uint8_t* mpDesc, const uint8_t* observed,
                                          const int* l2r, const int* r2l, float th, float nnratio, int* kp_mp);
int search_by_projection_mappoints(FrameView& F, int nMP, const uint8_t* in_view, const float* projX, const float* projY,
                                   const int* level, const float* viewCos, const uint8_t* mpDesc, const uint8_t* observed,
                                   const float* projXR, float th, float nnratio, int* kp_mp);

// Window argmin used by SearchByProjection(Frame&,Frame&) (mode: 0 = [lvl-1,lvl+1], 1 = forward [lvl,inf), 2 = backward [0,lvl])
// and SearchByProjection(Frame&,KeyFrame*) (mode 0, any_mp_blocks = true, dist_th = ORBdist, no right check).
// valid[i]: the point exists, projects inside the image (and invzc >= 0).  ur[i] = uv.x - mbf*invzc (used when F.uRight != null).
// angle[i]: source keypoint angle for the rotation histogram.  Returns nmatches after the rotation-consistency filter.
int search_by_projection_window(FrameView& F, int n, const uint8_t* valid, const float* u, const float* v, const int* lvl,
                                const float* ur, const float* angle, const uint8_t* srcDesc, const uint8_t* observed, float th,
                                int mode, int dist_th, bool any_mp_blocks, bool check_orientation, int* kp_mp);

// SearchForInitialization.  prevMatched: n1 x 2 floats (in/out).  matches12: n1 ints (out).  Returns nmatches.
int search_for_initialization(const FrameView& F1, FrameView& F2, float* prevMatched, int windowSize, float nnratio,
                              bool check_orientation, int* matches12);

struct StereoParams { float mb, mbf; };
// ComputeStereoMatches.  Pyramids are the un-blurred levels of the left/right extractor.  Outputs uRight/depth (N floats, -1 = none),
// and for inspection bestDistOut/bestIdxOut (Hamming stage, -1 = skipped).
void compute_stereo_matches(int N, const KeyPoint* kpL, const uint8_t* descL, int Nr, const KeyPoint* kpR, const uint8_t* descR,
                            const std::vector<Image>& pyrL, const std::vector<Image>& pyrR, const float* scaleFactors,
                            const float* invScaleFactors, StereoParams sp, float* uRight, float* depth, int* bestDistOut,
                            int* bestIdxOut);

// kNN(2) + Lowe ratio of ComputeStereoFishEyeMatches: good[i] = (#neighbours >= 2 && d0 < d1 * 0.7)
void knn2_ratio(const uint8_t* q, int nq, const uint8_t* t, int nt, int* idx, int* dist, uint8_t* good);

// ---- the remaining ORBmatcher entry points ------------------------------------------------------
// DBoW2::FeatureVector (std::map<NodeId, std::vector<unsigned>>) flattened: node ids ascending, CSR offsets, feature indices.
struct FeatVec {
  int nnodes = 0;
  const uint32_t* nodeIds = nullptr;
  const int* nodeStart = nullptr;   // [nnodes + 1]
  const uint32_t* indices = nullptr;
};

// The bucketed best/second search shared by both SearchByBoW overloads.  Side 1 is walked (node order, then position in the
// node's list), side 2 is claimed.  elig1[i]: the feature takes part (has a good map point; idx < mvKeysUn.size());
// elig2[j]: static eligibility of a side-2 feature (KF-KF: has a good map point; KF-Frame: always).  A claimed side-2
// feature is skipped by later side-1 features.  Accept when best <= maxDist and (float)best < nnratio * (float)second.
// maxDist = TH_LOW for KF-Frame (":<=TH_LOW"), TH_LOW - 1 for KF-KF ("<TH_LOW").  matches12[i] = j or -1; returns nmatches
// after the rotation filter (angle1 - angle2 histogram).
int search_by_bow_stereo(int n1, const KeyPoint* kps1, const uint8_t* desc1, const FeatVec& fv1, const uint8_t* elig1, int n2, int nLeft2,
                         const KeyPoint* kps2, const uint8_t* desc2, const FeatVec& fv2, int maxDist, float nnratio, bool check_orientation,
                         int* matches12L, int* matches12R);
int search_by_bow(int n1, const KeyPoint* kps1, const uint8_t* desc1, const FeatVec& fv1, const uint8_t* elig1, int n2,
                  const KeyPoint* kps2, const uint8_t* desc2, const FeatVec& fv2, const uint8_t* elig2, int maxDist, float nnratio,
                  bool check_orientation, int* matches12);

// Pinhole::epipolarConstrain with the fundamental matrix already formed (F12 row-major 3x3), unc = mvLevelSigma2[kp2.octave].
bool epipolar_constrain_pinhole(const KeyPoint& kp1, const KeyPoint& kp2, const float* F12, float unc);

// SearchForTriangulation (pinhole, mpCamera2 == null).  elig1[i] = !GetMapPoint(i) && (!bOnlyStereo || stereo1[i]);
// elig2[j] likewise for pKF2.  stereo1/2 = mvuRight[.] >= 0.  ep = projection of KF1's centre in KF2; levelSigma2 /
// scaleFactors are pKF2's tables.  matches12[i] = j or -1.
int search_for_triangulation(int n1, const KeyPoint* kps1, const uint8_t* desc1, const FeatVec& fv1, const uint8_t* elig1,
                             const uint8_t* stereo1, int n2, const KeyPoint* kps2, const uint8_t* desc2, const FeatVec& fv2,
                             const uint8_t* elig2, const uint8_t* stereo2, const float* F12, float epx, float epy,
                             const float* scaleFactors2, const float* levelSigma2_2, bool coarse, bool check_orientation,
                             int* matches12);

// ---- KannalaBrandt8 cameras: the gate of SearchForTriangulation is a two-view triangulation (KannalaBrandt8.cpp:244-247) ----------
// Relative geometry of one (camera of KF1, camera of KF2) combination: R12 / t12 as ORBmatcher.cc:1000-1019 forms them,
// cam = fx fy cx cy k1 k2 k3 k4.
struct Kb8Pair { float R12[9]; float t12[3]; float cam1[8]; float cam2[8]; };
// libm policy (same switch as frame_oracle.cc): 0 = tanf / atan2f of the host libm (what the reference's binary calls);
// 1 = (float)tan((double)x), (float)atan2((double)..) -- portable, the arithmetic of the CUDA path.
enum { LIBM_F32 = 0, LIBM_PORTABLE = 1 };
// KannalaBrandt8::unproject (KannalaBrandt8.cpp:112-137): Newton iterations on the distortion polynomial, ray = (x, y, 1).
void kb8_unproject(const float* cam, float u, float v, int libmMode, float* ray);
// KannalaBrandt8::project (KannalaBrandt8.cpp:29-45).
void kb8_project(const float* cam, const float* X, int libmMode, float* u, float* v);
// cv::SVD::compute(A, w, u, vt, MODIFY_A | FULL_UV) of a 4x4 CV_32F matrix (row-major), w and vt only: OpenCV's one-sided Jacobi
// (modules/core/src/lapack.cpp, JacobiSVDImpl_<float>), which is what cv::SVD runs for a matrix this small in every build.
void jacobi_svd4(const float* A, float* w, float* vt);
// KannalaBrandt8::TriangulateMatches (KannalaBrandt8.cpp:343-418) + Triangulate (:430-444): z1 of the triangulated point in camera 1, or
// -1.  p3D (3 floats, may be null) receives the point when it is accepted.  epipolarConstrain is `> 0.0001f` on the result.
float kb8_triangulate_matches(const Kb8Pair& G, float x1, float y1, float x2, float y2, float sigmaLevel, float unc, int libmMode,
                              float* p3D);
// SearchForTriangulation for KannalaBrandt8 keyframes (ORBmatcher.cc:981-1222).  One camera per keyframe: nLeft1 = nLeft2 = -1, G[0]
// is the pair, the epipole gate (:1093-1101) applies.  Fisheye-stereo keyframes (mpCamera2 set, NLeft != -1): kps = mvKeys followed
// by mvKeysRight, G[2 * bRight1 + bRight2] = (ll, lr, rl, rr), no feature counts as stereo and the epipole gate is off (:1093).
// levelSigma2_1 / _2 = mvLevelSigma2 of pKF1 / pKF2 (sigmaLevel and unc of the gate).
int search_for_triangulation_kb8(int n1, const KeyPoint* kps1, const uint8_t* desc1, const FeatVec& fv1, const uint8_t* elig1,
                                 const uint8_t* stereo1, int nLeft1, int n2, const KeyPoint* kps2, const uint8_t* desc2,
                                 const FeatVec& fv2, const uint8_t* elig2, const uint8_t* stereo2, int nLeft2, const Kb8Pair* G,
                                 float epx, float epy, const float* scaleFactors2, const float* levelSigma2_1,
                                 const float* levelSigma2_2, bool coarse, bool check_orientation, int libmMode, int* matches12);

// The per-point search loop shared by Fuse (both overloads), SearchBySim3 and (without blocking) SearchByProjection(KF,Scw):
// KeyFrame::GetFeaturesInArea(u, v, radius) -> octave in [level-1, level] -> optional Fuse chi-square gate
// (gate != 0: uses F.uRight / ur / invLevelSigma2; 7.8 with a right coordinate, 5.99 without) -> strict argmin Hamming.
// bestIdx[i] = -1 / bestDist[i] = 256 when nothing qualifies.  The queries are independent of each other.
void search_window_best(const FrameView& F, int n, const uint8_t* valid, const float* u, const float* v, const float* radius,
                        const int* level, const float* ur, const uint8_t* desc, int gate, const float* invLevelSigma2,
                        int* bestIdx, int* bestDist);

// ---- pure Hamming workloads next to the path ------------------------------------------------------------
// MapPoint::ComputeDistinctiveDescriptors, src/MapPoint.cc:350-436, for a batch of map points: descriptors of point p are
// rows start[p] .. start[p+1]-1 of desc.  bestIdx[p] = row (relative to start[p]) with the least median distance to the others
// (median = sorted row [0.5*(N-1)], first minimum wins), bestMedian[p] that median; -1 / INT_MAX for a point without rows.
void distinctive_descriptors(const uint8_t* desc, const int* start, int npoints, int* bestIdx, int* bestMedian);

// DBoW2 vocabulary tree flattened (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h: m_nodes[id].children / descriptor / weight /
// word_id).  Node 0 is the root; children of node i are childIds[childStart[i] .. childStart[i+1]) in their stored order.
struct VocabTree {
  int nnodes = 0, L = 0;
  const int* childStart = nullptr;
  const int* childIds = nullptr;
  const uint8_t* nodeDesc = nullptr;   // nnodes x 32
  const double* nodeWeight = nullptr;
  const int* nodeWordId = nullptr;     // -1 for inner nodes
};
// TemplatedVocabulary::transform(feature, word_id, weight, nid, levelsup), TemplatedVocabulary.h:1218-1259, for n features.
// The folding of per-feature (word, weight, node) into BowVector / FeatureVector for TF_IDF weighting + L1 norm (ORBvoc's
// settings), TemplatedVocabulary.h:1145-1195, BowVector.cpp:34-84, FeatureVector.cpp:28-42.  Outputs: sorted word ids + values,
// and the FeatureVector as CSR.  Returns the number of words; nnodesOut = number of FeatureVector nodes.
int vocab_fold(const int* wordId, const int* nodeId, const double* weight, int n, uint32_t* bowIds, double* bowValues,
               uint32_t* fvNodeIds, int* fvStart, uint32_t* fvIndices, int* nnodesOut);
void vocab_transform(const VocabTree& V, const uint8_t* desc, int n, int levelsup, int* wordId, int* nodeId, double* weight);

}  // namespace oro
