// CPU ORACLE support (test infrastructure, not product code).
//
// C entry points around the reference's own, unmodified ORB_SLAM3::ORBextractor (compiled by oracle/Makefile from
// /root/reference/src/ORBextractor.cc by path) and the two dependency-free members of ORBmatcher
// (DescriptorDistance, ComputeThreeMaxima; oracle/Makefile cuts them out of /root/reference/src/ORBmatcher.cc at
// build time into oracle/_ref/).  Loaded by oracle/ref.py for tests/ and bench.py's CPU legs only.
//
// Node allocator: every operator new issued while a reference call runs is served from a monotone bump arena (ref_alloc.cc),
// so the (size, ExtractorNode*) tie of DistributeOctTree's sort is "later created = higher address" (SURVEY F5).
#include <atomic>
#include <chrono>
#include <thread>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

#include "ORBextractor.h"
#include "ref_alloc.h"

namespace cvshim { extern thread_local int g_blur_taps; }

namespace {
// Read access to the protected tables (mnFeaturesPerLevel, umax) for the constructor parity test.
struct Probe : public ORB_SLAM3::ORBextractor {
  using ORB_SLAM3::ORBextractor::ORBextractor;
  const std::vector<int>& quotas() const { return mnFeaturesPerLevel; }
  const std::vector<int>& discRows() const { return umax; }
};
}  // namespace

namespace ORB_SLAM3 {
// The two members of ORBmatcher that need nothing but integers (include/ORBmatcher.h:47, :104); their bodies are the
// reference's own lines, cut out of src/ORBmatcher.cc by oracle/Makefile.
class ORBmatcher {
 public:
  static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b);
  void ComputeThreeMaxima(std::vector<int>* histo, const int L, int& ind1, int& ind2, int& ind3);
};
using std::vector;
#include "orbmatcher_slice.inc"
}  // namespace ORB_SLAM3

extern "C" {

void* ref_extractor_create(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST) {
  return new Probe(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST);
}
void ref_extractor_destroy(void* h) { delete (Probe*)h; }

// ORBextractor::operator() on a host image.  Returns monoIndex (or -1 for an empty image, ORBextractor.cc:1075),
// -2 if cap is too small, -3 if the node arena overflowed.  kps_out: 28-byte cv::KeyPoint records.
int ref_extract(void* h, const uint8_t* img, int w, int hgt, int stride, int lap0, int lap1, int blur_taps, void* kps_out,
                uint8_t* desc_out, int cap, int* n_out) {
  Probe* ex = (Probe*)h;
  cvshim::g_blur_taps = blur_taps;
  int mono, n;
  {
    refalloc::Scope scope;
    if (!scope.ok()) return -3;
    {
      std::vector<cv::KeyPoint> kps;
      cv::Mat desc;
      std::vector<int> lap{lap0, lap1};
      cv::Mat image = (img && w > 0 && hgt > 0) ? cv::Mat(hgt, w, CV_8UC1, (void*)img, (size_t)stride) : cv::Mat();
      mono = (*ex)(image, cv::Mat(), kps, desc, lap);
      n = (int)kps.size();
      if (n_out) *n_out = mono < 0 ? 0 : n;
      if (mono >= 0 && n <= cap) {
        static_assert(sizeof(cv::KeyPoint) == 28, "cv::KeyPoint layout");
        if (n) memcpy(kps_out, kps.data(), (size_t)n * sizeof(cv::KeyPoint));
        for (int i = 0; i < n; i++) memcpy(desc_out + (size_t)i * 32, desc.ptr(i), 32);
      }
    }
    if (refalloc::overflowed()) return -3;
  }
  if (mono < 0) return -1;
  if (n > cap) return -2;
  return mono;
}

// mvImagePyramid[level] after the last ref_extract on this object (a view into the bordered level image).
int ref_pyramid_level(void* h, int level, const uint8_t** ptr, int* w, int* hgt, int* stride) {
  Probe* ex = (Probe*)h;
  if (level < 0 || level >= (int)ex->mvImagePyramid.size()) return -1;
  const cv::Mat& m = ex->mvImagePyramid[level];
  *ptr = m.data; *w = m.cols; *hgt = m.rows; *stride = (int)m.step;
  return 0;
}

// Constructor tables (ORBextractor.cc:408-468): four float vectors of nlevels, the per-level quotas and umax[16].
void ref_tables(void* h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2, int* quotas, int* umax16) {
  Probe* ex = (Probe*)h;
  std::vector<float> a = ex->GetScaleFactors(), b = ex->GetInverseScaleFactors(), c = ex->GetScaleSigmaSquares(),
                     d = ex->GetInverseScaleSigmaSquares();
  const int L = ex->GetLevels();
  for (int i = 0; i < L; i++) { scale[i] = a[i]; inv_scale[i] = b[i]; sigma2[i] = c[i]; inv_sigma2[i] = d[i]; quotas[i] = ex->quotas()[i]; }
  for (int i = 0; i < 16; i++) umax16[i] = ex->discRows()[i];
}
int ref_levels(void* h) { return ((Probe*)h)->GetLevels(); }
float ref_scale_factor(void* h) { return ((Probe*)h)->GetScaleFactor(); }

// ORBmatcher::DescriptorDistance (ORBmatcher.cc:2463-2483) on n pairs of 32-byte rows.
void ref_descriptor_distance(const uint8_t* a, const uint8_t* b, int n, int* out) {
  for (int i = 0; i < n; i++) {
    cv::Mat ma(1, 32, CV_8U, (void*)(a + (size_t)i * 32)), mb(1, 32, CV_8U, (void*)(b + (size_t)i * 32));
    out[i] = ORB_SLAM3::ORBmatcher::DescriptorDistance(ma, mb);
  }
}

// ORBmatcher::ComputeThreeMaxima (ORBmatcher.cc:2416-2458) on bin sizes (the function only reads histo[i].size()).
void ref_three_maxima(const int* bin_sizes, int L, int* ind) {
  std::vector<std::vector<int>> histo(L);
  for (int i = 0; i < L; i++) histo[i].assign(bin_sizes[i], 0);
  int i1 = -1, i2 = -1, i3 = -1;
  ORB_SLAM3::ORBmatcher m;
  m.ComputeThreeMaxima(histo.data(), L, i1, i2, i3);
  ind[0] = i1; ind[1] = i2; ind[2] = i3;
}

// Frame-parallel timing of the reference operator(): one extractor object per thread, as one ORBextractor per
// camera thread in the reference (Frame.cc:120-123).  Returns seconds; *total_kps = keypoints over all frames.
double ref_time_extract(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh, int blur_taps, const uint8_t* frames,
                        int nframes, int w, int hgt, int lap0, int lap1, int nthreads, long long* total_kps) {
  std::vector<std::thread> th;
  std::vector<long long> counts(nthreads, 0);
  auto t0 = std::chrono::steady_clock::now();
  for (int t = 0; t < nthreads; t++)
    th.emplace_back([&, t]() {
      cvshim::g_blur_taps = blur_taps;
      Probe ex(nfeatures, scaleFactor, nlevels, iniTh, minTh);
      for (int f = t; f < nframes; f += nthreads) {
        refalloc::Scope scope;
        std::vector<cv::KeyPoint> kps;
        cv::Mat desc;
        std::vector<int> lap{lap0, lap1};
        cv::Mat image(hgt, w, CV_8UC1, (void*)(frames + (size_t)f * w * hgt), (size_t)w);
        ex(image, cv::Mat(), kps, desc, lap);
        counts[t] += (long long)kps.size();
      }
    });
  for (auto& x : th) x.join();
  auto t1 = std::chrono::steady_clock::now();
  long long tot = 0;
  for (auto c : counts) tot += c;
  if (total_kps) *total_kps = tot;
  return std::chrono::duration<double>(t1 - t0).count();
}

}  // extern "C"
