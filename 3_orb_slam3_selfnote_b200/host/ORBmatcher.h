// Drop-in counterpart of the reference's include/ORBmatcher.h (class ORB_SLAM3::ORBmatcher, lines 39-107): constructor,
// the three constants, static DescriptorDistance and the data-parallel search entry points, computed on the GPU through
// the C-ABI of include/orbx.h.  The search methods are templates over the frame / map-point types and read exactly the
// members the reference's bodies read (F.N, F.mvKeysUn, F.mDescriptors, F.mvuRight, F.mvScaleFactors, F.mnMinX.., F.Nleft,
// F.mvpMapPoints, pMP->mbTrackInView, ...), so they bind to the reference's own Frame / MapPoint classes unchanged and to
// the small mocks of tests/cpp/shim_test.cc alike.  Pose algebra / camera projection (the loop prologues of the
// Frame-Frame and Frame-KeyFrame overloads) stay on the host: see SearchByProjectionProjected and INTEGRATION.md.
// SearchByProjectionProjected with mode 3 is the search loop of both SearchByProjection(KeyFrame*, Scw, ...) overloads
// (src/ORBmatcher.cc:489-720: anyMpBlocks = true, distTh = floor(TH_LOW*ratioHamming), kpMp = vpMatched as -1 / -2).
#pragma once
#include <climits>
#include <cstring>
#include <stdexcept>
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

  // Search matches between Frame keypoints and projected MapPoints (local-map tracking), src/ORBmatcher.cc:44-142.
  // Monocular / rectified-stereo frames (F.Nleft == -1); the fisheye-stereo right-camera pass (:144-211) is not offloaded.
  template <class FrameT, class MapPointT>
  int SearchByProjection(FrameT& F, const std::vector<MapPointT*>& vpMapPoints, const float th = 3, const bool bFarPoints = false,
                         const float thFarPoints = 50.0f) {
    if (F.Nleft != -1) throw std::runtime_error("ORBmatcher::SearchByProjection: fisheye stereo (Nleft != -1) is not offloaded");
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

  // ---- SURVEY.md section 8(f) rank 1 ---------------------------------------------------------------------------------------
  // Search matches between MapPoints in a KeyFrame and ORB in a Frame, brute force constrained to ORB that belong to the same
  // vocabulary node (relocalisation / reference-keyframe tracking), src/ORBmatcher.cc:273-475.  Nleft == -1 frames.
  template <class KeyFrameT, class FrameT, class MapPointT>
  int SearchByBoW(KeyFrameT* pKF, FrameT& F, std::vector<MapPointT*>& vpMapPointMatches) {
    if (F.Nleft != -1) throw std::runtime_error("ORBmatcher::SearchByBoW: fisheye stereo (Nleft != -1) is not offloaded");
    const std::vector<MapPointT*> vpMapPointsKF = pKF->GetMapPointMatches();
    vpMapPointMatches.assign(F.N, static_cast<MapPointT*>(nullptr));
    const int n1 = (int)vpMapPointsKF.size();
    std::vector<uint8_t> elig1(std::max(n1, 1));
    for (int i = 0; i < n1; i++) elig1[i] = vpMapPointsKF[i] && !vpMapPointsKF[i]->isBad();
    FeatVecHolder fv1(pKF->mFeatVec), fv2(F.mFeatVec);
    DescHolder d1(pKF->mDescriptors, n1), d2(F.mDescriptors, F.N);
    std::vector<int> m12(std::max(n1, 1), -1);
    int nmatches = 0;
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
