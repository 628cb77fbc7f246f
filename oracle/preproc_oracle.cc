// CPU ORACLE (test infrastructure, not product code) -- SURVEY.md §8(f) rank 3: the 8-bit image operations the reference
// applies to a frame BEFORE ORBextractor::operator() sees it.
//
//   cvtColor RGB/BGR/RGBA/BGRA -> GRAY     src/Tracking.cc:989-1014 (stereo), and the mono / RGB-D twins below it
//   remap(INTER_LINEAR, float maps)        Examples/Stereo/stereo_euroc.cc:166-167 (maps from :117-118)
//   CLAHE(3.0, 8x8)->apply                 Examples/Monocular/mono_tum_vi.cc:101,109
//
// All three are OpenCV calls (external, un-vendored: CMakeLists.txt:43-48), restated here from the published algorithm
// and PINNED against python cv2 4.13 in tests/test_oracle_preproc.py (bit-exact).  Only tests/, smoke() and bench.py's
// CPU legs may load this library.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

// cvRound = cvtss2si: round-half-even, and the "integer indefinite" INT_MIN for NaN / out-of-range inputs
static inline int cv_round(float v) { return (v >= -2147483648.f && v < 2147483648.f) ? (int)lrintf(v) : INT32_MIN; }
static inline int cv_floor(float v) { return (int)floorf(v); }
static inline uint8_t sat_u8(int v) { return (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v); }
static inline int reflect101(int p, int len) {
  if (len == 1) return 0;
  while (p < 0 || p >= len) p = p < 0 ? -p : 2 * len - 2 - p;
  return p;
}

extern "C" {

// cvtColor(..., CV_{RGB,BGR,RGBA,BGRA}2GRAY) on 8U.  coeffs 1 (OpenCV >= 4 / late 3.4): 15-bit weights R 9798, G 19235,
// B 3735, (sum + 2^14) >> 15.  coeffs 0 (OpenCV 3.4.0-3.4.x table version): 14-bit weights R 4899, G 9617, B 1868,
// (sum + 2^13) >> 14.  rgb_order = 1 when channel 0 is red (mbRGB, Tracking.cc:991).
void oro_cvt_gray(const uint8_t* src, int w, int h, int stride, int channels, int rgb_order, int coeffs, uint8_t* dst, int dstride) {
  const int shift = coeffs ? 15 : 14;
  const unsigned cr = coeffs ? 9798 : 4899, cg = coeffs ? 19235 : 9617, cb = coeffs ? 3735 : 1868;
  const unsigned c0 = rgb_order ? cr : cb, c2 = rgb_order ? cb : cr;
  for (int y = 0; y < h; ++y) {
    const uint8_t* s = src + (size_t)y * stride;
    uint8_t* d = dst + (size_t)y * dstride;
    for (int x = 0; x < w; ++x, s += channels) d[x] = (uint8_t)((s[0] * c0 + s[1] * cg + s[2] * c2 + (1u << (shift - 1))) >> shift);
  }
}

// cv::remap(src, dst, mapx, mapy, INTER_LINEAR) with CV_32FC1 maps, BORDER_CONSTANT, value 0 (the defaults the reference
// relies on).  OpenCV's fixed-point path: s = cvRound(map * 32) (INTER_BITS = 5), integer part s >> 5, the 2x2 weights
// come from a table of the float products (1-fy)(1-fx).. scaled by 2^15 -- exact integers (32-fx)(32-fy)*32 for the
// bilinear case, so the table's sum correction never fires --, out-of-image taps read the border value 0, result
// (sum + 2^14) >> 15.
void oro_remap_linear(const uint8_t* src, int sw, int sh, int sstride, const float* mapx, const float* mapy, int mstride /*floats*/,
                      uint8_t* dst, int dw, int dh, int dstride) {
  for (int y = 0; y < dh; ++y)
    for (int x = 0; x < dw; ++x) {
      int sx = cv_round(mapx[(size_t)y * mstride + x] * 32.f), sy = cv_round(mapy[(size_t)y * mstride + x] * 32.f);
      int fx = sx & 31, fy = sy & 31;
      int ix = std::max(-32768, std::min(32767, sx >> 5)), iy = std::max(-32768, std::min(32767, sy >> 5));
      auto px = [&](int xx, int yy) -> int { return (xx >= 0 && xx < sw && yy >= 0 && yy < sh) ? src[(size_t)yy * sstride + xx] : 0; };
      int w00 = (32 - fx) * (32 - fy) * 32, w01 = fx * (32 - fy) * 32, w10 = (32 - fx) * fy * 32, w11 = fx * fy * 32;
      int v = px(ix, iy) * w00 + px(ix + 1, iy) * w01 + px(ix, iy + 1) * w10 + px(ix + 1, iy + 1) * w11;
      dst[(size_t)y * dstride + x] = sat_u8((v + (1 << 14)) >> 15);
    }
}

// cv::CLAHE::apply on 8U (modules/imgproc/src/clahe.cpp): per-tile clipped histogram -> LUT, then bilinear blend of the four
// neighbouring tiles' LUTs in float.  Sizes that are not a multiple of the tile grid are padded bottom / right with
// BORDER_REFLECT_101 by tiles - (size % tiles) pixels in BOTH directions as soon as one of them needs it.
void oro_clahe(const uint8_t* src, int w, int h, int stride, double clip_limit, int tilesX, int tilesY, uint8_t* dst, int dstride) {
  int ew = w, eh = h;
  if (w % tilesX != 0 || h % tilesY != 0) { ew = w + tilesX - (w % tilesX); eh = h + tilesY - (h % tilesY); }
  const int tw = ew / tilesX, th = eh / tilesY, total = tw * th;
  const float lutScale = (float)255 / total;
  int clip = 0;
  if (clip_limit > 0.0) clip = std::max((int)(clip_limit * total / 256), 1);
  std::vector<uint8_t> lut((size_t)tilesX * tilesY * 256);
  for (int ty = 0; ty < tilesY; ++ty)
    for (int tx = 0; tx < tilesX; ++tx) {
      int hist[256] = {0};
      for (int y = ty * th; y < (ty + 1) * th; ++y)
        for (int x = tx * tw; x < (tx + 1) * tw; ++x) hist[src[(size_t)reflect101(y, h) * stride + reflect101(x, w)]]++;
      if (clip > 0) {
        int clipped = 0;
        for (int i = 0; i < 256; ++i)
          if (hist[i] > clip) { clipped += hist[i] - clip; hist[i] = clip; }
        int batch = clipped / 256, residual = clipped - batch * 256;
        for (int i = 0; i < 256; ++i) hist[i] += batch;
        if (residual != 0) {
          int step = std::max(256 / residual, 1);
          for (int i = 0; i < 256 && residual > 0; i += step, residual--) hist[i]++;
        }
      }
      uint8_t* L = lut.data() + ((size_t)ty * tilesX + tx) * 256;
      int sum = 0;
      for (int i = 0; i < 256; ++i) { sum += hist[i]; L[i] = sat_u8(cv_round(sum * lutScale)); }
    }
  const float inv_tw = 1.0f / tw, inv_th = 1.0f / th;
  for (int y = 0; y < h; ++y) {
    float tyf = y * inv_th - 0.5f;
    int ty1 = cv_floor(tyf), ty2 = ty1 + 1;
    float ya = tyf - ty1, ya1 = 1.0f - ya;
    ty1 = std::max(ty1, 0); ty2 = std::min(ty2, tilesY - 1);
    for (int x = 0; x < w; ++x) {
      float txf = x * inv_tw - 0.5f;
      int tx1 = cv_floor(txf), tx2 = tx1 + 1;
      float xa = txf - tx1, xa1 = 1.0f - xa;
      tx1 = std::max(tx1, 0); tx2 = std::min(tx2, tilesX - 1);
      int v = src[(size_t)y * stride + x];
      const uint8_t* P1 = lut.data() + (size_t)ty1 * tilesX * 256;
      const uint8_t* P2 = lut.data() + (size_t)ty2 * tilesX * 256;
      float res = (P1[tx1 * 256 + v] * xa1 + P1[tx2 * 256 + v] * xa) * ya1 + (P2[tx1 * 256 + v] * xa1 + P2[tx2 * 256 + v] * xa) * ya;
      dst[(size_t)y * dstride + x] = sat_u8(cv_round(res));
    }
  }
}

}  // extern "C"
