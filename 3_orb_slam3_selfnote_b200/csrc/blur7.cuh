// k_blur7v -- cv::GaussianBlur(7x7, sigma 2, BORDER_REFLECT_101) on 8U in OpenCV's fixed point (reference
// call site src/ORBextractor.cc:1136).  Tuned replacement of k_blur7 (extract_kernels.cuh: issue-bound at 88 %, ~30
// instructions per pixel, 39 M shared-memory bank conflicts per launch in profiles/r01).
//
// OpenCV's 8-bit path is exact integer arithmetic: row pass sum(tap*px) in Q8.8 without rounding (<= 65280 because the
// taps sum to 256), column pass sum(tap*row) in Q16.16, one rounding (+0x8000 >> 16) at the end.  The result is therefore
// (sum_ij t_i t_j p_ij + 0x8000) >> 16 whatever the order of the two passes, and this kernel runs the COLUMN pass first:
//   A. warp = strip of R output rows, lane = one 32-bit word column (4 pixels) read straight from HBM/L2 (coalesced 128-B
//      rows, no staging); bytes are spread to u16 pairs (PRMT) and each tap is ONE 32-bit IMAD for two pixels at once
//      (no carry between the halves since every partial sum <= 65280); the u16 sums go to shared memory (STS.64).
//   B. row pass on the u16 sums with IDP.2A (two 16-bit x 8-bit products per instruction) on aligned word pairs -- the tap
//      bytes are pre-arranged per output parity so no funnel shifts are needed; rounding rides in the accumulator init.
// CTA = 120 x 8R output pixels: the 32 word columns of a strip cover x0-4 .. x0+123 (3-pixel halo each side).
// Algorithmic bytes per launch: read + write every level.
#pragma once
#include "dev_common.cuh"
#include "geom.h"

namespace orbx {

constexpr int kBvW = 120;          // output pixels per tile row
constexpr int kBvR = 4;            // output rows per warp
constexpr int kBvH = 8 * kBvR;     // output rows per tile

__device__ __forceinline__ int bv_refl(int i, int n) {
  if (i < 0) i = -i;
  if (i >= n) i = 2 * (n - 1) - i;
  return max(0, min(n - 1, i));  // (twice-reflected only on degenerate tiny levels)
}

__global__ void __launch_bounds__(256) k_blur7v(const __grid_constant__ Geom g, const __grid_constant__ Batch b) {
  __shared__ __align__(16) uint32_t sV[kBvH * 64];  // column-pass sums, u16 pairs: [row][64 words = 128 columns]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, f = blockIdx.y;
  int level = 0;
  while (level + 1 < g.nlevels && (int)blockIdx.x >= g.lv[level + 1].blurTileBase) level++;
  const LevelGeom& L = g.lv[level];
  const int t = blockIdx.x - L.blurTileBase;
  const int ty = t / L.blurTilesX, tx = t - ty * L.blurTilesX;
  const int x0 = tx * kBvW, y0 = ty * kBvH;
  int sp;
  const uint8_t* src = level_ptr(g, b, level, f, sp);

  // ---- A. column pass
  {
    const int ys = y0 + warp * kBvR;      // first output row of the strip
    if (ys < L.h) {
      const int gx = x0 - 4 + 4 * lane;   // first pixel column of this lane's word
      const bool edge = (gx < 0) || (gx + 3 >= L.w);
      int c0 = 0, c1 = 0, c2 = 0, c3 = 0;
      if (edge) { c0 = bv_refl(gx, L.w); c1 = bv_refl(gx + 1, L.w); c2 = bv_refl(gx + 2, L.w); c3 = bv_refl(gx + 3, L.w); }
      uint32_t lo[kBvR + 6], hi[kBvR + 6];
      const bool inner = (ys >= 3) && (ys + kBvR + 2 < L.h);   // warp-uniform: no row reflection needed
      const uint8_t* rowIn = src + (long long)(ys - 3) * sp;
#pragma unroll
      for (int i = 0; i < kBvR + 6; i++) {
        const uint8_t* row = inner ? rowIn + i * sp : src + (long long)bv_refl(ys - 3 + i, L.h) * sp;
        uint32_t w;
        if (!edge) w = __ldg(reinterpret_cast<const uint32_t*>(row + gx));
        else w = (uint32_t)__ldg(row + c0) | ((uint32_t)__ldg(row + c1) << 8) | ((uint32_t)__ldg(row + c2) << 16) | ((uint32_t)__ldg(row + c3) << 24);
        lo[i] = __byte_perm(w, 0u, 0x4140);
        hi[i] = __byte_perm(w, 0u, 0x4342);
      }
      const uint32_t t0 = g.taps[0], t1 = g.taps[1], t2 = g.taps[2], t3 = g.taps[3], t4 = g.taps[4], t5 = g.taps[5], t6 = g.taps[6];
#pragma unroll
      for (int j = 0; j < kBvR; j++) {
        uint2 v;
        v.x = t0 * lo[j] + t1 * lo[j + 1] + t2 * lo[j + 2] + t3 * lo[j + 3] + t4 * lo[j + 4] + t5 * lo[j + 5] + t6 * lo[j + 6];
        v.y = t0 * hi[j] + t1 * hi[j + 1] + t2 * hi[j + 2] + t3 * hi[j + 3] + t4 * hi[j + 4] + t5 * hi[j + 5] + t6 * hi[j + 6];
        *reinterpret_cast<uint2*>(&sV[(warp * kBvR + j) * 64 + 2 * lane]) = v;
      }
    }
  }
  __syncthreads();

  // ---- B. row pass: lane = 4 output pixels; output pixel p (tile-local) reads sum columns p+1 .. p+7
  if (lane < kBvW / 4) {
    const int x = x0 + 4 * lane;
    if (x < L.w) {
      const uint32_t t0 = g.taps[0], t1 = g.taps[1], t2 = g.taps[2], t3 = g.taps[3], t4 = g.taps[4], t5 = g.taps[5], t6 = g.taps[6];
      const uint32_t Ta = (t0 << 8) | (t1 << 16) | (t2 << 24);               // ( 0, t0 | t1, t2)
      const uint32_t Tb = t3 | (t4 << 8) | (t5 << 16) | (t6 << 24);          // (t3, t4 | t5, t6)
      const uint32_t Tc = t0 | (t1 << 8) | (t2 << 16) | (t3 << 24);          // (t0, t1 | t2, t3)
      const uint32_t Td = t4 | (t5 << 8) | (t6 << 16);                       // (t4, t5 | t6,  0)
      uint8_t* dst = b.blur + (long long)f * b.blurFrameStride + L.blurOff + x;
#pragma unroll
      for (int i = 0; i < kBvR; i++) {
        const int r = warp + 8 * i, y = y0 + r;
        if (y >= L.h) break;
        const uint2* p = reinterpret_cast<const uint2*>(&sV[r * 64 + 2 * lane]);
        const uint2 A = p[0], B = p[1], C = p[2];
        const uint32_t w0 = A.x, w1 = A.y, w2 = B.x, w3 = B.y, w4 = C.x, w5 = C.y;
        uint32_t a0 = 32768u, a1 = 32768u, a2 = 32768u, a3 = 32768u;
        a0 = __dp2a_lo(w0, Ta, a0); a0 = __dp2a_hi(w1, Ta, a0); a0 = __dp2a_lo(w2, Tb, a0); a0 = __dp2a_hi(w3, Tb, a0);
        a1 = __dp2a_lo(w1, Tc, a1); a1 = __dp2a_hi(w2, Tc, a1); a1 = __dp2a_lo(w3, Td, a1); a1 = __dp2a_hi(w4, Td, a1);
        a2 = __dp2a_lo(w1, Ta, a2); a2 = __dp2a_hi(w2, Ta, a2); a2 = __dp2a_lo(w3, Tb, a2); a2 = __dp2a_hi(w4, Tb, a2);
        a3 = __dp2a_lo(w2, Tc, a3); a3 = __dp2a_hi(w3, Tc, a3); a3 = __dp2a_lo(w4, Td, a3); a3 = __dp2a_hi(w5, Td, a3);
        // saturate_cast<uchar>: the OCV346 taps sum to 257, so a bright patch reaches 257.5 in Q16.16 (a >= 2^24)
        a0 = min(a0, 0x00FFFFFFu); a1 = min(a1, 0x00FFFFFFu); a2 = min(a2, 0x00FFFFFFu); a3 = min(a3, 0x00FFFFFFu);
        const uint32_t o01 = __byte_perm(a0, a1, 0x0062), o23 = __byte_perm(a2, a3, 0x0062);   // bits 16..23 of each sum
        *reinterpret_cast<uint32_t*>(dst + (long long)y * L.blurPitch) = __byte_perm(o01, o23, 0x5410);
      }
    }
  }
}

}  // namespace orbx
