// CPU ORACLE (test infrastructure, not product code) -- see orb_oracle.h for scope and parity status.
// Build: g++ -O2 -std=c++17 -ffp-contract=off -fPIC -shared (see oracle/Makefile).
// -ffp-contract=off is REQUIRED: the float expressions below must not be fused.
#include "orb_oracle.h"

#include <algorithm>
#include <cassert>
#include <cfloat>
#include <climits>
#include <cmath>
#include <cstring>
#include <list>
#include <utility>

namespace oro {

static const int PATCH_SIZE = 31;       // ORBextractor.cc:70
static const int HALF_PATCH_SIZE = 15;  // ORBextractor.cc:71
static const int EDGE_THRESHOLD = 19;   // ORBextractor.cc:72

static const signed char kPattern[256 * 4] = {
#include "../3_orb_slam3_selfnote_b200/csrc/orb_pattern.inc"
};

// cvRound(float/double): SSE cvtss2si / lrint under the default rounding mode = round-half-even.
static inline int cv_round(float v) { return (int)lrintf(v); }
static inline int cv_round(double v) { return (int)lrint(v); }
static inline int cv_floor(double v) { int i = (int)v; return i - (i > v); }
static inline int cv_ceil(double v) { int i = (int)v; return i + (i < v); }

// ------------------------------------------------------------------------------------------------
// cv::resize(INTER_LINEAR) for 8UC1 -- OpenCV imgproc/resize.cpp fixed-point path, SURVEY.md §9.1.
// Coefficients are Q11 shorts (INTER_RESIZE_COEF_BITS = 11); the horizontal pass keeps int32 rows,
// the vertical pass is ((b0*(S0>>4))>>16) + ((b1*(S1>>4))>>16) + 2) >> 2.
// Called by the reference at src/ORBextractor.cc:1201.
// ------------------------------------------------------------------------------------------------
static inline short sat_short_round(float v) {
  int i = cv_round(v);
  return (short)std::min(32767, std::max(-32768, i));
}

void resize_linear_u8(const uint8_t* src, int sw, int sh, int sstride, uint8_t* dst, int dw, int dh, int dstride) {
  const double inv_scale_x = (double)dw / sw, inv_scale_y = (double)dh / sh;
  const double scale_x = 1. / inv_scale_x, scale_y = 1. / inv_scale_y;
  std::vector<int> xofs(dw), yofs(dh);
  std::vector<short> ialpha(dw * 2), ibeta(dh * 2);
  for (int dx = 0; dx < dw; dx++) {
    float fx = (float)((dx + 0.5) * scale_x - 0.5);
    int sx = cv_floor(fx);
    fx -= sx;
    if (sx < 0) { fx = 0; sx = 0; }
    if (sx >= sw - 1) { fx = 0; sx = sw - 1; }
    xofs[dx] = sx;
    ialpha[dx * 2] = sat_short_round((1.f - fx) * 2048.f);
    ialpha[dx * 2 + 1] = sat_short_round(fx * 2048.f);
  }
  for (int dy = 0; dy < dh; dy++) {
    float fy = (float)((dy + 0.5) * scale_y - 0.5);
    int sy = cv_floor(fy);
    fy -= sy;
    yofs[dy] = sy;
    ibeta[dy * 2] = sat_short_round((1.f - fy) * 2048.f);
    ibeta[dy * 2 + 1] = sat_short_round(fy * 2048.f);
  }
  std::vector<int> r0(dw), r1(dw);
  auto hrow = [&](int sy, std::vector<int>& out) {
    sy = std::min(std::max(sy, 0), sh - 1);  // rows are clipped, beta is not zeroed (resizeGeneric_Invoker)
    const uint8_t* S = src + (size_t)sy * sstride;
    for (int dx = 0; dx < dw; dx++) {
      int sx = xofs[dx];
      int sx1 = std::min(sx + 1, sw - 1);
      out[dx] = S[sx] * ialpha[dx * 2] + S[sx1] * ialpha[dx * 2 + 1];
    }
  };
  for (int dy = 0; dy < dh; dy++) {
    hrow(yofs[dy], r0);
    hrow(yofs[dy] + 1, r1);
    const int b0 = ibeta[dy * 2], b1 = ibeta[dy * 2 + 1];
    uint8_t* D = dst + (size_t)dy * dstride;
    for (int dx = 0; dx < dw; dx++) {
      int v = (((b0 * (r0[dx] >> 4)) >> 16) + ((b1 * (r1[dx] >> 4)) >> 16) + 2) >> 2;
      D[dx] = (uint8_t)std::min(255, std::max(0, v));
    }
  }
}

// ------------------------------------------------------------------------------------------------
// cv::GaussianBlur(7x7, sigma 2, BORDER_REFLECT_101) on 8UC1 -- fixed-point branch, SURVEY.md §9.2.
// Called by the reference at src/ORBextractor.cc:1136 on a clone of each level.
// ------------------------------------------------------------------------------------------------
static const int kTaps[2][7] = {{18, 34, 49, 55, 49, 34, 18},   // OpenCV 3.4.1-3.4.8: per-tap cvRound(256*g)  (unverified, F7)
                                {18, 34, 48, 56, 48, 34, 18}};  // OpenCV >= 3.4.9/4.2: bit-exact kernel (probed on cv2 4.13)
static inline int refl101(int i, int n) {
  if (n == 1) return 0;
  while (i < 0 || i >= n) i = i < 0 ? -i : 2 * (n - 1) - i;
  return i;
}

void gaussian_blur7_u8(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride, int taps) {
  const int* k = kTaps[taps];
  std::vector<uint16_t> hbuf((size_t)w * h);
  for (int y = 0; y < h; y++) {
    const uint8_t* S = src + (size_t)y * sstride;
    for (int x = 0; x < w; x++) {
      uint32_t acc = 0;
      for (int i = 0; i < 7; i++) {
        acc += (uint32_t)k[i] * S[refl101(x + i - 3, w)];
        if (acc > 65535u) acc = 65535u;  // ufixedpoint16 saturating add
      }
      hbuf[(size_t)y * w + x] = (uint16_t)acc;
    }
  }
  for (int y = 0; y < h; y++) {
    uint8_t* D = dst + (size_t)y * dstride;
    for (int x = 0; x < w; x++) {
      uint32_t acc = 0;
      for (int i = 0; i < 7; i++) acc += (uint32_t)k[i] * hbuf[(size_t)refl101(y + i - 3, h) * w + x];
      uint32_t v = (acc + 32768u) >> 16;
      D[x] = (uint8_t)std::min(255u, v);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// cv::FAST(img, kps, t, nonmax, TYPE_9_16) -- SURVEY.md §9.3.  Called at src/ORBextractor.cc:808,827.
// ------------------------------------------------------------------------------------------------
static const int kRingDx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
static const int kRingDy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};

// S(p): max over the 16 arcs of 9 contiguous ring pixels of max(min(v - ring), min(ring - v)).
static inline int fast_S(const uint8_t* p, int stride) {
  const int v = p[0];
  int d[16];
  for (int k = 0; k < 16; k++) d[k] = v - p[kRingDy[k] * stride + kRingDx[k]];
  int best = -256;
  for (int k = 0; k < 16; k++) {
    int mn = INT_MAX, mx = INT_MIN;
    for (int j = 0; j < 9; j++) {
      int dv = d[(k + j) & 15];
      mn = std::min(mn, dv);
      mx = std::max(mx, dv);
    }
    best = std::max(best, std::max(mn, -mx));
  }
  return best;
}

void fast9_16(const uint8_t* img, int w, int h, int stride, int threshold, bool nonmax, std::vector<Candidate>& out) {
  out.clear();
  if (w < 7 || h < 7) return;
  std::vector<int> score((size_t)w * h, 0);      // cornerScore (= S-1) where corner at this threshold, else 0
  std::vector<uint8_t> corner((size_t)w * h, 0);
  for (int y = 3; y < h - 3; y++)
    for (int x = 3; x < w - 3; x++) {
      const uint8_t* p = img + (size_t)y * stride + x;
      // exact necessary condition (every 9-arc holds one pixel of each opposite pair k, k+8), as in OpenCV's fast.cpp
      const int v = p[0], lo = v - threshold, hi = v + threshold;
      int d = 3;  // bit0: darker arc possible, bit1: brighter arc possible
      for (int k = 0; k < 8 && d; k++) {
        const int a = p[kRingDy[k] * stride + kRingDx[k]], b = p[kRingDy[k + 8] * stride + kRingDx[k + 8]];
        d &= ((a < lo || b < lo) ? 1 : 0) | ((a > hi || b > hi) ? 2 : 0);
      }
      if (!d) continue;
      int S = fast_S(p, stride);
      if (S > threshold) { score[(size_t)y * w + x] = S - 1; corner[(size_t)y * w + x] = 1; }
    }
  for (int y = 3; y < h - 3; y++)
    for (int x = 3; x < w - 3; x++) {
      if (!corner[(size_t)y * w + x]) continue;
      const int s = score[(size_t)y * w + x];
      bool keep = true;
      if (nonmax) {  // strictly greater than all 8 neighbours; non-corners and untested pixels count as 0
        for (int dy = -1; dy <= 1 && keep; dy++)
          for (int dx = -1; dx <= 1; dx++) {
            if (!dx && !dy) continue;
            if (score[(size_t)(y + dy) * w + x + dx] >= s) { keep = false; break; }
          }
      }
      if (keep) out.push_back({x, y, s});
    }
}

// cv::fastAtan2 scalar (core/mathfuncs_core): 7th-order odd polynomial in fp32, Horner as written, no FMA.
float fast_atan2(float y, float x) {
  static const float scale = (float)(180.0 / 3.1415926535897932384626433832795);
  static const float p1 = 0.9997878412794807f * scale, p3 = -0.3258083974640975f * scale, p5 = 0.1555786518463281f * scale,
                     p7 = -0.04432655554792128f * scale;
  float ax = std::fabs(x), ay = std::fabs(y), a, c, c2;
  if (ax >= ay) {
    c = ay / (ax + (float)DBL_EPSILON);
    c2 = c * c;
    a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
  } else {
    c = ax / (ay + (float)DBL_EPSILON);
    c2 = c * c;
    a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
  }
  if (x < 0) a = 180.f - a;
  if (y < 0) a = 360.f - a;
  return a;
}

// cv::BFMatcher(NORM_HAMMING).knnMatch(q, t, k=2): ascending distance, ties keep the lower train index first.
// Used by the reference at src/Frame.cc:1246.  Missing neighbours are reported as idx -1 / dist -1.
void knn2_hamming(const uint8_t* q, int nq, const uint8_t* t, int nt, int* idx, int* dist) {
  for (int i = 0; i < nq; i++) {
    int d0 = INT_MAX, d1 = INT_MAX, i0 = -1, i1 = -1;
    for (int j = 0; j < nt; j++) {
      int d = hamming256(q + 32 * (size_t)i, t + 32 * (size_t)j);
      if (d < d0) { d1 = d0; i1 = i0; d0 = d; i0 = j; }
      else if (d < d1) { d1 = d; i1 = j; }
    }
    idx[2 * i] = i0; idx[2 * i + 1] = i1;
    dist[2 * i] = i0 < 0 ? -1 : d0; dist[2 * i + 1] = i1 < 0 ? -1 : d1;
  }
}

// ------------------------------------------------------------------------------------------------
// Reference functions.
// ------------------------------------------------------------------------------------------------

// ORBmatcher::DescriptorDistance, src/ORBmatcher.cc:2463-2483 (SWAR popcount over 8 x u32).
int hamming256(const uint8_t* a, const uint8_t* b) {
  int dist = 0;
  for (int i = 0; i < 8; i++) {
    uint32_t pa, pb;
    memcpy(&pa, a + 4 * i, 4);
    memcpy(&pb, b + 4 * i, 4);
    uint32_t v = pa ^ pb;
    v = v - ((v >> 1) & 0x55555555);
    v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
    dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
  }
  return dist;
}

// IC_Angle, src/ORBextractor.cc:75-102.
float ic_angle(const uint8_t* img, int stride, int x, int y, const int* umax) {
  int m_01 = 0, m_10 = 0;
  const uint8_t* center = img + (size_t)y * stride + x;
  for (int u = -HALF_PATCH_SIZE; u <= HALF_PATCH_SIZE; ++u) m_10 += u * center[u];
  for (int v = 1; v <= HALF_PATCH_SIZE; ++v) {
    int v_sum = 0;
    int d = umax[v];
    for (int u = -d; u <= d; ++u) {
      int val_plus = center[u + v * stride], val_minus = center[u - v * stride];
      v_sum += (val_plus - val_minus);
      m_10 += u * (val_plus + val_minus);
    }
    m_01 += v * v_sum;
  }
  return fast_atan2((float)m_01, (float)m_10);
}

// computeOrbDescriptor, src/ORBextractor.cc:105-145.
void orb_descriptor(const uint8_t* img, int stride, int x, int y, float angle_deg, int sincos_mode, uint8_t* desc) {
  const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);  // ORBextractor.cc:105
  float angle = angle_deg * factorPI;
  float a, b;
  if (sincos_mode == SINCOS_LIBM_F32) { a = cosf(angle); b = sinf(angle); }       // what `cos(float)` resolves to (F6)
  else { a = (float)cos((double)angle); b = (float)sin((double)angle); }          // portable definition implemented on the GPU
  const uint8_t* center = img + (size_t)y * stride + x;
  const signed char* pat = kPattern;
  for (int i = 0; i < 32; ++i) {
    int val = 0;
    for (int j = 0; j < 8; ++j, pat += 4) {
      float x0 = (float)pat[0], y0 = (float)pat[1], x1 = (float)pat[2], y1 = (float)pat[3];
      int t0 = center[cv_round(x0 * b + y0 * a) * stride + cv_round(x0 * a - y0 * b)];
      int t1 = center[cv_round(x1 * b + y1 * a) * stride + cv_round(x1 * a - y1 * b)];
      val |= (t0 < t1) << j;
    }
    desc[i] = (uint8_t)val;
  }
}

// ExtractorNode + DistributeOctTree, src/ORBextractor.cc:479-761 -- literal restatement on std::list.
// Declared tie rule: the reference sorts (size, node address); here the address is replaced by the
// node's creation sequence number, i.e. what the reference does under a monotone allocator.
namespace {
struct Node {
  std::vector<int> keys;  // indices into the candidate list, order preserved
  int ULx, ULy, URx, URy, BLx, BLy, BRx, BRy;
  std::list<Node>::iterator lit;
  bool noMore = false;
  int seq = 0;
};

void divide(const Node& p, const std::vector<Candidate>& c, Node& n1, Node& n2, Node& n3, Node& n4) {
  const int halfX = (int)std::ceil((float)(p.URx - p.ULx) / 2);
  const int halfY = (int)std::ceil((float)(p.BRy - p.ULy) / 2);
  n1.ULx = p.ULx; n1.ULy = p.ULy; n1.URx = p.ULx + halfX; n1.URy = p.ULy;
  n1.BLx = p.ULx; n1.BLy = p.ULy + halfY; n1.BRx = p.ULx + halfX; n1.BRy = p.ULy + halfY;
  n2.ULx = n1.URx; n2.ULy = n1.URy; n2.URx = p.URx; n2.URy = p.URy;
  n2.BLx = n1.BRx; n2.BLy = n1.BRy; n2.BRx = p.URx; n2.BRy = p.ULy + halfY;
  n3.ULx = n1.BLx; n3.ULy = n1.BLy; n3.URx = n1.BRx; n3.URy = n1.BRy;
  n3.BLx = p.BLx; n3.BLy = p.BLy; n3.BRx = n1.BRx; n3.BRy = p.BLy;
  n4.ULx = n3.URx; n4.ULy = n3.URy; n4.URx = n2.BRx; n4.URy = n2.BRy;
  n4.BLx = n3.BRx; n4.BLy = n3.BRy; n4.BRx = p.BRx; n4.BRy = p.BRy;
  for (int k : p.keys) {
    const float kx = (float)c[k].x, ky = (float)c[k].y;
    if (kx < n1.URx) {
      if (ky < n1.BRy) n1.keys.push_back(k); else n3.keys.push_back(k);
    } else if (ky < n1.BRy) n2.keys.push_back(k);
    else n4.keys.push_back(k);
  }
  if (n1.keys.size() == 1) n1.noMore = true;
  if (n2.keys.size() == 1) n2.noMore = true;
  if (n3.keys.size() == 1) n3.noMore = true;
  if (n4.keys.size() == 1) n4.noMore = true;
}
}  // namespace

std::vector<int> distribute_octtree(const std::vector<Candidate>& cands, int minX, int maxX, int minY, int maxY, int N) {
  std::vector<int> result;
  const int nIni = (int)std::round((float)(maxX - minX) / (maxY - minY));
  if (nIni <= 0) return result;  // the reference divides by zero / indexes an empty vector here (portrait aspect < 0.5)
  const float hX = (float)(maxX - minX) / nIni;
  std::list<Node> lNodes;
  std::vector<Node*> ini(nIni);
  int seq = 0;
  for (int i = 0; i < nIni; i++) {
    Node ni;
    ni.ULx = (int)(hX * (float)i); ni.ULy = 0;
    ni.URx = (int)(hX * (float)(i + 1)); ni.URy = 0;
    ni.BLx = ni.ULx; ni.BLy = maxY - minY;
    ni.BRx = ni.URx; ni.BRy = maxY - minY;
    ni.seq = seq++;
    lNodes.push_back(ni);
    ini[i] = &lNodes.back();
  }
  for (size_t i = 0; i < cands.size(); i++) {
    size_t r = (size_t)((float)cands[i].x / hX);
    if (r >= (size_t)nIni) r = nIni - 1;  // out of range = UB in the reference; cannot happen for x < maxX-minX
    ini[r]->keys.push_back((int)i);
  }
  for (auto lit = lNodes.begin(); lit != lNodes.end();) {
    if (lit->keys.size() == 1) { lit->noMore = true; ++lit; }
    else if (lit->keys.empty()) lit = lNodes.erase(lit);
    else ++lit;
  }
  bool bFinish = false;
  typedef std::pair<int, int> SizeSeq;  // (size, creation seq) -- stands in for (size, node address)
  std::vector<std::pair<SizeSeq, Node*>> vSizeAndPointerToNode;
  auto push_child = [&](Node& n, bool count, int& nToExpand) {
    if (n.keys.empty()) return;
    n.seq = seq++;
    lNodes.push_front(n);
    if (n.keys.size() > 1) {
      if (count) nToExpand++;
      vSizeAndPointerToNode.push_back({{(int)n.keys.size(), lNodes.front().seq}, &lNodes.front()});
      lNodes.front().lit = lNodes.begin();
    }
  };
  while (!bFinish) {
    int prevSize = (int)lNodes.size();
    auto lit = lNodes.begin();
    int nToExpand = 0;
    vSizeAndPointerToNode.clear();
    while (lit != lNodes.end()) {
      if (lit->noMore) { ++lit; continue; }
      Node n1, n2, n3, n4;
      divide(*lit, cands, n1, n2, n3, n4);
      push_child(n1, true, nToExpand);
      push_child(n2, true, nToExpand);
      push_child(n3, true, nToExpand);
      push_child(n4, true, nToExpand);
      lit = lNodes.erase(lit);
    }
    if ((int)lNodes.size() >= N || (int)lNodes.size() == prevSize) {
      bFinish = true;
    } else if (((int)lNodes.size() + nToExpand * 3) > N) {
      while (!bFinish) {
        prevSize = (int)lNodes.size();
        auto vPrev = vSizeAndPointerToNode;
        vSizeAndPointerToNode.clear();
        std::sort(vPrev.begin(), vPrev.end(),
                  [](const std::pair<SizeSeq, Node*>& a, const std::pair<SizeSeq, Node*>& b) { return a.first < b.first; });
        for (int j = (int)vPrev.size() - 1; j >= 0; j--) {
          Node n1, n2, n3, n4;
          divide(*vPrev[j].second, cands, n1, n2, n3, n4);
          int dummy = 0;
          push_child(n1, false, dummy);
          push_child(n2, false, dummy);
          push_child(n3, false, dummy);
          push_child(n4, false, dummy);
          lNodes.erase(vPrev[j].second->lit);
          if ((int)lNodes.size() >= N) break;
        }
        if ((int)lNodes.size() >= N || (int)lNodes.size() == prevSize) bFinish = true;
      }
    }
  }
  result.reserve(lNodes.size());
  for (auto& n : lNodes) {
    int best = n.keys[0];
    float maxResponse = (float)cands[best].resp;
    for (size_t k = 1; k < n.keys.size(); k++)
      if ((float)cands[n.keys[k]].resp > maxResponse) { best = n.keys[k]; maxResponse = (float)cands[best].resp; }
    result.push_back(best);
  }
  return result;
}

// ORBextractor::ORBextractor, src/ORBextractor.cc:408-468.
Extractor::Extractor(int nfeatures_, float scaleFactor_, int nlevels_, int iniThFAST, int minThFAST, int blur_taps_, int sincos_mode_)
    : nfeatures(nfeatures_), nlevels(nlevels_), iniTh(iniThFAST), minTh(minThFAST), blur_taps(blur_taps_), sincos_mode(sincos_mode_),
      scaleFactor(scaleFactor_) {
  mvScaleFactor.resize(nlevels); mvLevelSigma2.resize(nlevels);
  mvScaleFactor[0] = 1.0f; mvLevelSigma2[0] = 1.0f;
  for (int i = 1; i < nlevels; i++) {
    mvScaleFactor[i] = (float)(mvScaleFactor[i - 1] * scaleFactor);  // float * double member (ORBextractor.h:91)
    mvLevelSigma2[i] = mvScaleFactor[i] * mvScaleFactor[i];
  }
  mvInvScaleFactor.resize(nlevels); mvInvLevelSigma2.resize(nlevels);
  for (int i = 0; i < nlevels; i++) {
    mvInvScaleFactor[i] = 1.0f / mvScaleFactor[i];
    mvInvLevelSigma2[i] = 1.0f / mvLevelSigma2[i];
  }
  pyramid.resize(nlevels);
  mnFeaturesPerLevel.resize(nlevels);
  float factor = (float)(1.0f / scaleFactor);
  float nDesiredFeaturesPerScale = nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nlevels));
  int sumFeatures = 0;
  for (int level = 0; level < nlevels - 1; level++) {
    mnFeaturesPerLevel[level] = cv_round(nDesiredFeaturesPerScale);
    sumFeatures += mnFeaturesPerLevel[level];
    nDesiredFeaturesPerScale *= factor;
  }
  mnFeaturesPerLevel[nlevels - 1] = std::max(nfeatures - sumFeatures, 0);
  umax.resize(HALF_PATCH_SIZE + 1);
  int v, v0, vmax = cv_floor(HALF_PATCH_SIZE * sqrt(2.f) / 2 + 1);
  int vmin = cv_ceil(HALF_PATCH_SIZE * sqrt(2.f) / 2);
  const double hp2 = HALF_PATCH_SIZE * HALF_PATCH_SIZE;
  for (v = 0; v <= vmax; ++v) umax[v] = cv_round(sqrt(hp2 - v * v));
  for (v = HALF_PATCH_SIZE, v0 = 0; v >= vmin; --v) {
    while (umax[v0] == umax[v0 + 1]) ++v0;
    umax[v] = v0;
    ++v0;
  }
}

// ComputePyramid, src/ORBextractor.cc:1186-1219.  The EDGE_THRESHOLD frame (copyMakeBorder) is not materialised:
// no consumer on the path reads it.
void Extractor::computePyramid(const uint8_t* img, int w, int h, int stride) {
  for (int level = 0; level < nlevels; ++level) {
    float scale = mvInvScaleFactor[level];
    int sw = cv_round((float)w * scale), sh = cv_round((float)h * scale);
    pyramid[level].alloc(sw, sh);
    if (level != 0) {
      resize_linear_u8(pyramid[level - 1].px.data(), pyramid[level - 1].w, pyramid[level - 1].h, pyramid[level - 1].w,
                       pyramid[level].px.data(), sw, sh, sw);
    } else {
      for (int y = 0; y < h; y++) memcpy(pyramid[0].row(y), img + (size_t)y * stride, w);
    }
  }
}

// ComputeKeyPointsOctTree, src/ORBextractor.cc:763-878.
void Extractor::computeKeyPointsOctTree() {
  candidates.assign(nlevels, {});
  levelKeys.assign(nlevels, {});
  fallbackCells.assign(nlevels, 0);
  const float W = 30;
  for (int level = 0; level < nlevels; ++level) {
    const Image& im = pyramid[level];
    const int minBorderX = EDGE_THRESHOLD - 3;
    const int minBorderY = minBorderX;
    const int maxBorderX = im.w - EDGE_THRESHOLD + 3;
    const int maxBorderY = im.h - EDGE_THRESHOLD + 3;
    std::vector<Candidate>& vToDistributeKeys = candidates[level];
    const float width = (float)(maxBorderX - minBorderX);
    const float height = (float)(maxBorderY - minBorderY);
    const int nCols = (int)(width / W);
    const int nRows = (int)(height / W);
    if (nCols > 0 && nRows > 0) {
      const int wCell = (int)std::ceil(width / nCols);
      const int hCell = (int)std::ceil(height / nRows);
      std::vector<Candidate> cell;
      for (int i = 0; i < nRows; i++) {
        const float iniY = (float)(minBorderY + i * hCell);
        float maxY = iniY + hCell + 6;
        if (iniY >= maxBorderY - 3) continue;
        if (maxY > maxBorderY) maxY = (float)maxBorderY;
        for (int j = 0; j < nCols; j++) {
          const float iniX = (float)(minBorderX + j * wCell);
          float maxX = iniX + wCell + 6;
          if (iniX >= maxBorderX - 6) continue;
          if (maxX > maxBorderX) maxX = (float)maxBorderX;
          const int x0 = (int)iniX, y0 = (int)iniY, cw = (int)maxX - x0, ch = (int)maxY - y0;
          fast9_16(im.row(y0) + x0, cw, ch, im.w, iniTh, true, cell);
          if (cell.empty()) {
            fast9_16(im.row(y0) + x0, cw, ch, im.w, minTh, true, cell);
            fallbackCells[level]++;
          }
          for (auto& c : cell) vToDistributeKeys.push_back({c.x + j * wCell, c.y + i * hCell, c.resp});
        }
      }
    }
    std::vector<int> keep = distribute_octtree(vToDistributeKeys, minBorderX, maxBorderX, minBorderY, maxBorderY,
                                               mnFeaturesPerLevel[level]);
    const int scaledPatchSize = (int)(PATCH_SIZE * mvScaleFactor[level]);
    std::vector<KeyPoint>& kps = levelKeys[level];
    kps.reserve(keep.size());
    for (int k : keep) {
      KeyPoint kp;
      kp.x = (float)(vToDistributeKeys[k].x + minBorderX);
      kp.y = (float)(vToDistributeKeys[k].y + minBorderY);
      kp.size = (float)scaledPatchSize;
      kp.angle = -1.f;
      kp.response = (float)vToDistributeKeys[k].resp;
      kp.octave = level;
      kp.class_id = -1;
      kps.push_back(kp);
    }
  }
  for (int level = 0; level < nlevels; ++level)
    for (auto& kp : levelKeys[level])
      kp.angle = ic_angle(pyramid[level].px.data(), pyramid[level].w, cv_round(kp.x), cv_round(kp.y), umax.data());
}

// ORBextractor::operator(), src/ORBextractor.cc:1071-1184.
int Extractor::extract(const uint8_t* img, int w, int h, int stride, int lap0, int lap1, std::vector<KeyPoint>& out,
                       std::vector<uint8_t>& desc) {
  if (!img || w <= 0 || h <= 0) return -1;
  computePyramid(img, w, h, stride);
  computeKeyPointsOctTree();
  int nkeypoints = 0;
  for (int level = 0; level < nlevels; ++level) nkeypoints += (int)levelKeys[level].size();
  out.assign(nkeypoints, KeyPoint{});
  desc.assign((size_t)nkeypoints * 32, 0);
  int monoIndex = 0, stereoIndex = nkeypoints - 1;
  Image working;
  uint8_t d[32];
  for (int level = 0; level < nlevels; ++level) {
    std::vector<KeyPoint>& keypoints = levelKeys[level];
    if (keypoints.empty()) continue;
    const Image& im = pyramid[level];
    working.alloc(im.w, im.h);
    gaussian_blur7_u8(im.px.data(), im.w, im.h, im.w, working.px.data(), im.w, blur_taps);
    float scale = mvScaleFactor[level];
    for (auto kp : keypoints) {  // by value: levelKeys keeps level coordinates for inspection
      orb_descriptor(working.px.data(), working.w, cv_round(kp.x), cv_round(kp.y), kp.angle, sincos_mode, d);
      if (level != 0) { kp.x *= scale; kp.y *= scale; }
      if (kp.x >= lap0 && kp.x <= lap1) {
        out[stereoIndex] = kp;
        memcpy(&desc[(size_t)stereoIndex * 32], d, 32);
        stereoIndex--;
      } else {
        out[monoIndex] = kp;
        memcpy(&desc[(size_t)monoIndex * 32], d, 32);
        monoIndex++;
      }
    }
  }
  return monoIndex;
}

}  // namespace oro
