// CPU ORACLE (test infrastructure, not product code): flat C entry points for ctypes / bench cpu_baseline.
#include <chrono>
#include <cstring>
#include <thread>

#include "match_oracle.h"
#include "orb_oracle.h"

using namespace oro;

extern "C" {

void oro_resize(const uint8_t* src, int sw, int sh, int sstride, uint8_t* dst, int dw, int dh, int dstride) {
  resize_linear_u8(src, sw, sh, sstride, dst, dw, dh, dstride);
}
void oro_blur7(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride, int taps) {
  gaussian_blur7_u8(src, w, h, sstride, dst, dstride, taps);
}
// returns the number of keypoints (may exceed cap; only the first cap are written)
int oro_fast(const uint8_t* img, int w, int h, int stride, int threshold, int nonmax, int* xyr, int cap) {
  std::vector<Candidate> out;
  fast9_16(img, w, h, stride, threshold, nonmax != 0, out);
  for (size_t i = 0; i < out.size() && (int)i < cap; i++) { xyr[3 * i] = out[i].x; xyr[3 * i + 1] = out[i].y; xyr[3 * i + 2] = out[i].resp; }
  return (int)out.size();
}
float oro_fast_atan2(float y, float x) { return fast_atan2(y, x); }
int oro_hamming(const uint8_t* a, const uint8_t* b) { return hamming256(a, b); }
void oro_knn2(const uint8_t* q, int nq, const uint8_t* t, int nt, int* idx, int* dist) { knn2_hamming(q, nq, t, nt, idx, dist); }

int oro_distribute(const int* xyr, int n, int minX, int maxX, int minY, int maxY, int N, int* keep, int cap) {
  std::vector<Candidate> c(n);
  for (int i = 0; i < n; i++) c[i] = {xyr[3 * i], xyr[3 * i + 1], xyr[3 * i + 2]};
  std::vector<int> k = distribute_octtree(c, minX, maxX, minY, maxY, N);
  for (size_t i = 0; i < k.size() && (int)i < cap; i++) keep[i] = k[i];
  return (int)k.size();
}

void* oro_extractor_create(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh, int blur_taps, int sincos_mode) {
  return new Extractor(nfeatures, scaleFactor, nlevels, iniTh, minTh, blur_taps, sincos_mode);
}
void oro_extractor_destroy(void* e) { delete (Extractor*)e; }
void oro_extractor_tables(void* e, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2, int* quota, int* umax16) {
  Extractor* x = (Extractor*)e;
  for (int i = 0; i < x->nlevels; i++) {
    scale[i] = x->mvScaleFactor[i]; inv_scale[i] = x->mvInvScaleFactor[i];
    sigma2[i] = x->mvLevelSigma2[i]; inv_sigma2[i] = x->mvInvLevelSigma2[i];
    quota[i] = x->mnFeaturesPerLevel[i];
  }
  for (int i = 0; i < 16; i++) umax16[i] = x->umax[i];
}
// returns monoIndex (or -1); *n = number of keypoints; kps (28 B each) / desc (32 B each) hold min(*n, cap) rows.
int oro_extract(void* e, const uint8_t* img, int w, int h, int stride, int lap0, int lap1, KeyPoint* kps, uint8_t* desc, int cap, int* n) {
  Extractor* x = (Extractor*)e;
  std::vector<KeyPoint> k;
  std::vector<uint8_t> d;
  int mono = x->extract(img, w, h, stride, lap0, lap1, k, d);
  *n = (int)k.size();
  int m = std::min((int)k.size(), cap);
  if (m > 0) { memcpy(kps, k.data(), sizeof(KeyPoint) * m); memcpy(desc, d.data(), 32 * (size_t)m); }
  return mono;
}
int oro_level_size(void* e, int level, int* w, int* h) {
  Extractor* x = (Extractor*)e;
  *w = x->pyramid[level].w; *h = x->pyramid[level].h;
  return 0;
}
void oro_level_pixels(void* e, int level, uint8_t* dst) {
  Extractor* x = (Extractor*)e;
  memcpy(dst, x->pyramid[level].px.data(), x->pyramid[level].px.size());
}
int oro_level_candidates(void* e, int level, int* xyr, int cap) {
  Extractor* x = (Extractor*)e;
  auto& c = x->candidates[level];
  for (size_t i = 0; i < c.size() && (int)i < cap; i++) { xyr[3 * i] = c[i].x; xyr[3 * i + 1] = c[i].y; xyr[3 * i + 2] = c[i].resp; }
  return (int)c.size();
}
int oro_level_keypoints(void* e, int level, KeyPoint* kps, int cap) {
  Extractor* x = (Extractor*)e;
  auto& c = x->levelKeys[level];
  for (size_t i = 0; i < c.size() && (int)i < cap; i++) kps[i] = c[i];
  return (int)c.size();
}
int oro_level_fallback_cells(void* e, int level) { return ((Extractor*)e)->fallbackCells[level]; }
void oro_descriptor(const uint8_t* blurred, int stride, int x, int y, float angle, int sincos_mode, uint8_t* desc) {
  orb_descriptor(blurred, stride, x, y, angle, sincos_mode, desc);
}

// Timed CPU baseline: extract `nframes` frames (dense w*h each) with `nthreads` threads, one frame per thread at a time
// exactly as the reference runs it (one extractor object per thread).  Returns wall seconds.
double oro_time_extract(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh, int blur_taps, int sincos_mode,
                        const uint8_t* frames, int nframes, int w, int h, int lap0, int lap1, int nthreads, long long* total_kps) {
  std::vector<std::thread> th;
  std::vector<long long> counts(nthreads, 0);
  auto t0 = std::chrono::steady_clock::now();
  for (int t = 0; t < nthreads; t++)
    th.emplace_back([&, t]() {
      Extractor ex(nfeatures, scaleFactor, nlevels, iniTh, minTh, blur_taps, sincos_mode);
      std::vector<KeyPoint> k;
      std::vector<uint8_t> d;
      for (int f = t; f < nframes; f += nthreads) {
        ex.extract(frames + (size_t)f * w * h, w, h, w, lap0, lap1, k, d);
        counts[t] += (long long)k.size();
      }
    });
  for (auto& x : th) x.join();
  auto t1 = std::chrono::steady_clock::now();
  long long tot = 0;
  for (auto c : counts) tot += c;
  if (total_kps) *total_kps = tot;
  return std::chrono::duration<double>(t1 - t0).count();
}

// ---- matcher side -----------------------------------------------------------------------------------------------
static void fill_view(FrameView& F, int N, const KeyPoint* kps, const uint8_t* desc, const float* uRight, const float* scaleFactors,
                      const float* bounds4) {
  F.N = N; F.kps = kps; F.desc = desc; F.uRight = uRight; F.scaleFactors = scaleFactors;
  F.setBounds(bounds4[0], bounds4[1], bounds4[2], bounds4[3]);
  F.assignFeaturesToGrid();
}
int oro_features_in_area(int N, const KeyPoint* kps, const float* bounds4, float x, float y, float r, int minLevel, int maxLevel,
                         int* out, int cap) {
  FrameView F;
  fill_view(F, N, kps, nullptr, nullptr, nullptr, bounds4);
  std::vector<int> v = F.getFeaturesInArea(x, y, r, minLevel, maxLevel);
  for (size_t i = 0; i < v.size() && (int)i < cap; i++) out[i] = v[i];
  return (int)v.size();
}
int oro_search_by_projection_mappoints(int N, const KeyPoint* kps, const uint8_t* desc, const float* uRight, const float* scaleFactors,
                                       const float* bounds4, int nMP, const uint8_t* in_view, const float* projX, const float* projY,
                                       const int* level, const float* viewCos, const uint8_t* mpDesc, const uint8_t* observed,
                                       const float* projXR, float th, float nnratio, int* kp_mp) {
  FrameView F;
  fill_view(F, N, kps, desc, uRight, scaleFactors, bounds4);
  return search_by_projection_mappoints(F, nMP, in_view, projX, projY, level, viewCos, mpDesc, observed, projXR, th, nnratio, kp_mp);
}
int oro_search_by_projection_mappoints_stereo(int nL, const KeyPoint* kpsL, const uint8_t* descL, int nR, const KeyPoint* kpsR, const uint8_t* descR,
                                              const float* scaleFactors, const float* bounds4, int nMP, const uint8_t* inL, const float* pxL,
                                              const float* pyL, const int* lvL, const float* vcL, const uint8_t* inR, const float* pxR,
                                              const float* pyR, const int* lvR, const float* vcR, const uint8_t* mpDesc, const uint8_t* observed,
                                              const int* l2r, const int* r2l, float th, float nnratio, int* kp_mp) {
  FrameView FL, FR;
  fill_view(FL, nL, kpsL, descL, nullptr, scaleFactors, bounds4);
  fill_view(FR, nR, kpsR, descR, nullptr, scaleFactors, bounds4);
  return search_by_projection_mappoints_stereo(FL, FR, nMP, inL, pxL, pyL, lvL, vcL, inR, pxR, pyR, lvR, vcR, mpDesc, observed, l2r, r2l, th, nnratio, kp_mp);
}
int oro_search_by_projection_window(int N, const KeyPoint* kps, const uint8_t* desc, const float* uRight, const float* scaleFactors,
                                    const float* bounds4, int n, const uint8_t* valid, const float* u, const float* v, const int* lvl,
                                    const float* ur, const float* angle, const uint8_t* srcDesc, const uint8_t* observed, float th,
                                    int mode, int dist_th, int any_mp_blocks, int check_orientation, int* kp_mp) {
  FrameView F;
  fill_view(F, N, kps, desc, uRight, scaleFactors, bounds4);
  return search_by_projection_window(F, n, valid, u, v, lvl, ur, angle, srcDesc, observed, th, mode, dist_th, any_mp_blocks != 0,
                                     check_orientation != 0, kp_mp);
}
int oro_search_for_initialization(int N1, const KeyPoint* kps1, const uint8_t* desc1, int N2, const KeyPoint* kps2, const uint8_t* desc2,
                                  const float* bounds4, float* prevMatched, int windowSize, float nnratio, int check_orientation,
                                  int* matches12) {
  FrameView F1, F2;
  F1.N = N1; F1.kps = kps1; F1.desc = desc1;
  fill_view(F2, N2, kps2, desc2, nullptr, nullptr, bounds4);
  return search_for_initialization(F1, F2, prevMatched, windowSize, nnratio, check_orientation != 0, matches12);
}
// Uses the un-blurred pyramids held by the two extractor objects after their last extract() (Frame.cc:908,998,1012).
void oro_stereo_matches(void* exL, void* exR, int N, const KeyPoint* kpL, const uint8_t* descL, int Nr, const KeyPoint* kpR,
                        const uint8_t* descR, float mb, float mbf, float* uRight, float* depth, int* bestDist, int* bestIdx) {
  Extractor* L = (Extractor*)exL; Extractor* R = (Extractor*)exR;
  compute_stereo_matches(N, kpL, descL, Nr, kpR, descR, L->pyramid, R->pyramid, L->mvScaleFactor.data(), L->mvInvScaleFactor.data(),
                         StereoParams{mb, mbf}, uRight, depth, bestDist, bestIdx);
}
void oro_knn2_ratio(const uint8_t* q, int nq, const uint8_t* t, int nt, int* idx, int* dist, uint8_t* good) {
  knn2_ratio(q, nq, t, nt, idx, dist, good);
}
// ---- SURVEY.md §8(f) rank 1 ----
static FeatVec fv_of(int nnodes, const uint32_t* ids, const int* start, const uint32_t* indices) {
  FeatVec f; f.nnodes = nnodes; f.nodeIds = ids; f.nodeStart = start; f.indices = indices; return f;
}
int oro_search_by_bow(int n1, const KeyPoint* kps1, const uint8_t* desc1, int nn1, const uint32_t* ids1, const int* start1,
                      const uint32_t* ind1, const uint8_t* elig1, int n2, const KeyPoint* kps2, const uint8_t* desc2, int nn2,
                      const uint32_t* ids2, const int* start2, const uint32_t* ind2, const uint8_t* elig2, int maxDist, float nnratio,
                      int check_orientation, int* matches12) {
  return search_by_bow(n1, kps1, desc1, fv_of(nn1, ids1, start1, ind1), elig1, n2, kps2, desc2, fv_of(nn2, ids2, start2, ind2), elig2,
                       maxDist, nnratio, check_orientation != 0, matches12);
}
int oro_search_by_bow_stereo(int n1, const KeyPoint* kps1, const uint8_t* desc1, int nn1, const uint32_t* ids1, const int* start1,
                             const uint32_t* ind1, const uint8_t* elig1, int n2, int nLeft2, const KeyPoint* kps2, const uint8_t* desc2, int nn2,
                             const uint32_t* ids2, const int* start2, const uint32_t* ind2, int maxDist, float nnratio, int check_orientation,
                             int* matches12L, int* matches12R) {
  return search_by_bow_stereo(n1, kps1, desc1, fv_of(nn1, ids1, start1, ind1), elig1, n2, nLeft2, kps2, desc2, fv_of(nn2, ids2, start2, ind2),
                              maxDist, nnratio, check_orientation != 0, matches12L, matches12R);
}
int oro_search_for_triangulation(int n1, const KeyPoint* kps1, const uint8_t* desc1, int nn1, const uint32_t* ids1, const int* start1,
                                 const uint32_t* ind1, const uint8_t* elig1, const uint8_t* stereo1, int n2, const KeyPoint* kps2,
                                 const uint8_t* desc2, int nn2, const uint32_t* ids2, const int* start2, const uint32_t* ind2,
                                 const uint8_t* elig2, const uint8_t* stereo2, const float* F12, float epx, float epy,
                                 const float* scaleFactors2, const float* levelSigma2_2, int coarse, int check_orientation,
                                 int* matches12) {
  return search_for_triangulation(n1, kps1, desc1, fv_of(nn1, ids1, start1, ind1), elig1, stereo1, n2, kps2, desc2,
                                  fv_of(nn2, ids2, start2, ind2), elig2, stereo2, F12, epx, epy, scaleFactors2, levelSigma2_2,
                                  coarse != 0, check_orientation != 0, matches12);
}
// KannalaBrandt8 gate of SearchForTriangulation.  pairs: npairs x 28 floats (R12[9] t12[3] cam1[8] cam2[8]), npairs = 1 (one camera per
// keyframe, nLeft = -1) or 4 (ll, lr, rl, rr of a fisheye-stereo rig).
int oro_search_for_triangulation_kb8(int n1, const KeyPoint* kps1, const uint8_t* desc1, int nn1, const uint32_t* ids1, const int* start1,
                                     const uint32_t* ind1, const uint8_t* elig1, const uint8_t* stereo1, int nLeft1, int n2,
                                     const KeyPoint* kps2, const uint8_t* desc2, int nn2, const uint32_t* ids2, const int* start2,
                                     const uint32_t* ind2, const uint8_t* elig2, const uint8_t* stereo2, int nLeft2, const float* pairs,
                                     float epx, float epy, const float* scaleFactors2, const float* levelSigma2_1,
                                     const float* levelSigma2_2, int coarse, int check_orientation, int libmMode, int* matches12) {
  static_assert(sizeof(Kb8Pair) == 28 * sizeof(float), "Kb8Pair layout");
  return search_for_triangulation_kb8(n1, kps1, desc1, fv_of(nn1, ids1, start1, ind1), elig1, stereo1, nLeft1, n2, kps2, desc2,
                                      fv_of(nn2, ids2, start2, ind2), elig2, stereo2, nLeft2, (const Kb8Pair*)pairs, epx, epy,
                                      scaleFactors2, levelSigma2_1, levelSigma2_2, coarse != 0, check_orientation != 0, libmMode, matches12);
}
// n independent pairs through KannalaBrandt8::TriangulateMatches: xy1 / xy2 n x 2, sigma1 / sigma2 n; z1 (or -1) and the point per pair.
void oro_kb8_triangulate_matches(const float* pair, int n, const float* xy1, const float* xy2, const float* sigma1, const float* sigma2,
                                 int libmMode, float* z1, float* p3D) {
  for (int i = 0; i < n; i++) {
    p3D[3 * i] = p3D[3 * i + 1] = p3D[3 * i + 2] = 0.f;
    z1[i] = kb8_triangulate_matches(*(const Kb8Pair*)pair, xy1[2 * i], xy1[2 * i + 1], xy2[2 * i], xy2[2 * i + 1], sigma1[i], sigma2[i], libmMode,
                                    p3D + 3 * i);
  }
}
void oro_kb8_unproject(const float* cam, int n, const float* xy, int libmMode, float* rays) {
  for (int i = 0; i < n; i++) kb8_unproject(cam, xy[2 * i], xy[2 * i + 1], libmMode, rays + 3 * i);
}
void oro_kb8_project(const float* cam, int n, const float* X, int libmMode, float* uv) {
  for (int i = 0; i < n; i++) kb8_project(cam, X + 3 * i, libmMode, uv + 2 * i, uv + 2 * i + 1);
}
void oro_jacobi_svd4(const float* A, int n, float* w, float* vt) {
  for (int i = 0; i < n; i++) jacobi_svd4(A + 16 * i, w + 4 * i, vt + 16 * i);
}
void oro_search_window_best(int N, const KeyPoint* kps, const uint8_t* desc, const float* uRight, const float* bounds4, int n,
                            const uint8_t* valid, const float* u, const float* v, const float* radius, const int* level,
                            const float* ur, const uint8_t* qdesc, int gate, const float* invLevelSigma2, int* bestIdx, int* bestDist) {
  FrameView F;
  fill_view(F, N, kps, desc, uRight, nullptr, bounds4);
  search_window_best(F, n, valid, u, v, radius, level, ur, qdesc, gate, invLevelSigma2, bestIdx, bestDist);
}
// ---- SURVEY.md §8(f) rank 4 ----
void oro_distinctive_descriptors(const uint8_t* desc, const int* start, int npoints, int* bestIdx, int* bestMedian) {
  distinctive_descriptors(desc, start, npoints, bestIdx, bestMedian);
}
void oro_vocab_transform(int nnodes, int L, const int* childStart, const int* childIds, const uint8_t* nodeDesc, const double* nodeWeight,
                         const int* nodeWordId, const uint8_t* desc, int n, int levelsup, int* wordId, int* nodeId, double* weight) {
  VocabTree V; V.nnodes = nnodes; V.L = L; V.childStart = childStart; V.childIds = childIds; V.nodeDesc = nodeDesc;
  V.nodeWeight = nodeWeight; V.nodeWordId = nodeWordId;
  vocab_transform(V, desc, n, levelsup, wordId, nodeId, weight);
}
int oro_vocab_fold(const int* wordId, const int* nodeId, const double* weight, int n, uint32_t* bowIds, double* bowValues,
                   uint32_t* fvNodeIds, int* fvStart, uint32_t* fvIndices, int* nnodesOut) {
  return vocab_fold(wordId, nodeId, weight, n, bowIds, bowValues, fvNodeIds, fvStart, fvIndices, nnodesOut);
}
void oro_three_maxima(const int* hist, int L, int* ind3) {
  int a = -1, b = -1, c = -1;
  compute_three_maxima(hist, L, a, b, c);
  ind3[0] = a; ind3[1] = b; ind3[2] = c;
}

// Timed CPU baseline for whole stereo frames: frame-parallel workers; each worker extracts left and right on two threads
// exactly as Frame::Frame(stereo) does (src/Frame.cc:120-123), then runs ComputeStereoMatches.  nthreads = total threads used.
double oro_time_stereo(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh, int blur_taps, int sincos_mode,
                       const uint8_t* left, const uint8_t* right, int nframes, int w, int h, float mb, float mbf, int nthreads,
                       long long* total_matched) {
  const int workers = std::max(1, nthreads / 2);
  std::vector<std::thread> th;
  std::vector<long long> counts(workers, 0);
  auto t0 = std::chrono::steady_clock::now();
  for (int t = 0; t < workers; t++)
    th.emplace_back([&, t]() {
      Extractor exL(nfeatures, scaleFactor, nlevels, iniTh, minTh, blur_taps, sincos_mode), exR(nfeatures, scaleFactor, nlevels, iniTh, minTh, blur_taps, sincos_mode);
      std::vector<KeyPoint> kl, kr;
      std::vector<uint8_t> dl, dr;
      std::vector<float> uR, depth;
      for (int f = t; f < nframes; f += workers) {
        std::thread tr([&]() { exR.extract(right + (size_t)f * w * h, w, h, w, 0, 0, kr, dr); });
        exL.extract(left + (size_t)f * w * h, w, h, w, 0, 0, kl, dl);
        tr.join();
        uR.resize(kl.size() + 1); depth.resize(kl.size() + 1);
        compute_stereo_matches((int)kl.size(), kl.data(), dl.data(), (int)kr.size(), kr.data(), dr.data(), exL.pyramid, exR.pyramid,
                               exL.mvScaleFactor.data(), exL.mvInvScaleFactor.data(), StereoParams{mb, mbf}, uR.data(), depth.data(), nullptr, nullptr);
        for (size_t i = 0; i < kl.size(); i++) counts[t] += uR[i] >= 0;
      }
    });
  for (auto& x : th) x.join();
  auto t1 = std::chrono::steady_clock::now();
  long long tot = 0;
  for (auto c : counts) tot += c;
  if (total_matched) *total_matched = tot;
  return std::chrono::duration<double>(t1 - t0).count();
}

}  // extern "C"
