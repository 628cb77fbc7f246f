// CPU ORACLE (test infrastructure, not product code).
//
// A dependency-free C++ restatement of the reference's ORB frontend
// (src/ORBextractor.cc, src/ORBmatcher.cc, src/Frame.cc of
// wangyunpeng139/3_ORB_SLAM3_SelfNote) together with the OpenCV-3.4 8-bit primitives it
// calls (resize / FAST / GaussianBlur / fastAtan2 / BFMatcher).  OpenCV is an external,
// un-vendored dependency of the reference (CMakeLists.txt:43-48, "opencv346"), so those
// primitives are restated from their published algorithm and pinned against
// the only executable OpenCV available (python cv2 4.13) in tests/test_oracle_cv2.py.
//
// PARITY STATUS: pinned.  The reference ships no tests or golden vectors for this path, so the pin is the reference's own code:
// oracle/Makefile (`make ref`) compiles the UNMODIFIED /root/reference/src/ORBextractor.cc and src/ORBmatcher.cc (plus the member
// functions of Frame / KeyFrame / MapPoint / Pinhole / KannalaBrandt8 on the path, cut out of the reference's sources at build
// time) against the stand-in cv:: header of oracle/ref_shim/ into oracle/_ref/*.so, and tests/test_oracle_ref.py /
// tests/test_oracle_refmatch.py assert that this restatement equals it byte for byte (keypoints, descriptors, pyramid levels,
// match indices, uRight / depth, frustum fields).  The OpenCV primitives underneath both are the restatements below, pinned
// against python cv2 4.13 in tests/test_oracle_cv2.py; the committed golden vectors in tests/golden/ were generated from oracle/_ref.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
// may load this library.  The product (liborbx.so) never links or calls it.
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>

namespace oro {

// Layout-compatible with cv::KeyPoint (7 x 4 bytes).
struct KeyPoint {
  float x, y, size, angle, response;
  int octave, class_id;
};

struct Image {
  int w = 0, h = 0;
  std::vector<uint8_t> px;  // dense, pitch == w
  uint8_t* row(int y) { return px.data() + (size_t)y * w; }
  const uint8_t* row(int y) const { return px.data() + (size_t)y * w; }
  void alloc(int w_, int h_) { w = w_; h = h_; px.assign((size_t)w_ * h_, 0); }
};

enum BlurTaps { TAPS_OCV346 = 0, TAPS_OCV4 = 1 };
enum SinCosMode { SINCOS_LIBM_F32 = 0 /* cosf/sinf, reference-literal */, SINCOS_PORTABLE = 1 /* (float)cos((double)x) */ };

struct Candidate { int x, y, resp; };  // level coordinates relative to (minBorderX,minBorderY), FAST response

// ---- OpenCV primitives ---------------------------------------------
void resize_linear_u8(const uint8_t* src, int sw, int sh, int sstride, uint8_t* dst, int dw, int dh, int dstride);
void gaussian_blur7_u8(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride, int taps);
// FAST-9/16 with optional 3x3 non-max suppression on a w x h sub-image; emits row-major.
void fast9_16(const uint8_t* img, int w, int h, int stride, int threshold, bool nonmax, std::vector<Candidate>& out);
float fast_atan2(float y, float x);
void knn2_hamming(const uint8_t* q, int nq, const uint8_t* t, int nt, int* idx /*nq*2*/, int* dist /*nq*2*/);

// ---- reference functions -------------------------------------------------------------
float ic_angle(const uint8_t* img, int stride, int x, int y, const int* umax);
void orb_descriptor(const uint8_t* blurred, int stride, int x, int y, float angle_deg, int sincos_mode, uint8_t* desc32);
// DistributeOctTree.  xs/ys/resp in candidate order; returns indices (into the candidate list) of the kept
// points in the reference's output (list) order.
std::vector<int> distribute_octtree(const std::vector<Candidate>& cands, int minX, int maxX, int minY, int maxY, int N);
int hamming256(const uint8_t* a, const uint8_t* b);

class Extractor {
 public:
  Extractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST, int blur_taps, int sincos_mode);
  // returns monoIndex, or -1 on empty image (ORBextractor.cc:1075)
  int extract(const uint8_t* img, int w, int h, int stride, int lap0, int lap1, std::vector<KeyPoint>& kps,
              std::vector<uint8_t>& desc);

  int nfeatures, nlevels, iniTh, minTh, blur_taps, sincos_mode;
  double scaleFactor;
  std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;
  std::vector<int> mnFeaturesPerLevel, umax;
  std::vector<Image> pyramid;                      // mvImagePyramid (ROI only; the 19-px frame is not materialised)
  std::vector<std::vector<Candidate>> candidates;  // per level, in ComputeKeyPointsOctTree order (debug/inspection)
  std::vector<std::vector<KeyPoint>> levelKeys;    // per level, after DistributeOctTree + orientation (level coords)
  std::vector<int> fallbackCells;                  // per level: number of cells that needed minThFAST
 private:
  void computePyramid(const uint8_t* img, int w, int h, int stride);
  void computeKeyPointsOctTree();
};

}  // namespace oro
