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
// Second cut (profiles/r01c: 71 % of the instructions were in pass A -- 4-row strips re-read 10 rows for 4, the byte-wise
// reflect path of edge lanes ran in 41 % of the warps at 200 instructions, per-row reflect selects in every warp): strips of
// R = 8 rows, edge lanes load two aligned words + one PRMT, the row-reflect form only runs in the top / bottom strips.
// Algorithmic bytes per launch: read + write every level.
#pragma once
#include "dev_common.cuh"
#include "geom.h"

namespace orbx {

constexpr int kBvW = 120;          // output pixels per tile row
constexpr int kBvR = 8;            // output rows per warp (14 source rows: 1.75 loads per output row)
constexpr int kBvH = 8 * kBvR;     // output rows per tile

__device__ __forceinline__ int bv_refl(int i, int n) {
  if (i < 0) i = -i;
  if (i >= n) i = 2 * (n - 1) - i;
  return max(0, min(n - 1, i));  // (twice-reflected only on degenerate tiny levels)
}

// one source row of the column pass: the lane's 4 pixels (columns gx .. gx+3, reflected at the level's left / right edge).
// pa = src + gx (inner lanes) or src + wb (edge lanes), pb = src + wb2; off = byte offset of the row
template <bool EDGE>
__device__ __forceinline__ uint32_t bv_load(const uint8_t* pa, const uint8_t* pb, unsigned long long off, uint32_t sel) {
  if (!EDGE) return __ldg(reinterpret_cast<const uint32_t*>(pa + off));
  // reflected columns are 4 consecutive ones in reverse order: they lie in two adjacent aligned words, picked apart by PRMT
  return __byte_perm(__ldg(reinterpret_cast<const uint32_t*>(pa + off)), __ldg(reinterpret_cast<const uint32_t*>(pb + off)), sel);
}

template <bool EDGE>
__device__ __forceinline__ void bv_column_pass(const Geom& g, const uint8_t* src, int sp, int h, int ys, int gx, int wb, int wb2, uint32_t sel,
                                               uint32_t* dst /* &sV[(warp * kBvR) * 64 + 2 * lane] */) {
  uint32_t w[kBvR + 6];
  const uint8_t* pa = src + (EDGE ? wb : gx);
  const uint8_t* pb = src + wb2;
  asm volatile("" : "+l"(pa));   // whole per-thread pointers in registers: one IMAD.WIDE per row instead of re-adding the uniform base
  if (EDGE) asm volatile("" : "+l"(pb));
  const unsigned usp = (unsigned)sp;
  if (ys >= 3 && ys + kBvR + 2 < h) {   // warp-uniform: no row reflection needed
#pragma unroll
    for (int i = 0; i < kBvR + 6; i++) w[i] = bv_load<EDGE>(pa, pb, (unsigned long long)(unsigned)(ys - 3 + i) * usp, sel);
  } else {
#pragma unroll
    for (int i = 0; i < kBvR + 6; i++) w[i] = bv_load<EDGE>(pa, pb, (unsigned long long)(unsigned)bv_refl(ys - 3 + i, h) * usp, sel);
  }
  const uint32_t t[7] = {(uint32_t)g.taps[0], (uint32_t)g.taps[1], (uint32_t)g.taps[2], (uint32_t)g.taps[3], (uint32_t)g.taps[4],
                         (uint32_t)g.taps[5], (uint32_t)g.taps[6]};
  uint32_t alo[kBvR], ahi[kBvR];
#pragma unroll
  for (int j = 0; j < kBvR; j++) { alo[j] = 0; ahi[j] = 0; }
#pragma unroll
  for (int i = 0; i < kBvR + 6; i++) {
    const uint32_t lo = __byte_perm(w[i], 0u, 0x4140), hi = __byte_perm(w[i], 0u, 0x4342);   // bytes -> u16 pairs
#pragma unroll
    for (int j = 0; j < kBvR; j++)
      if (i - j >= 0 && i - j < 7) { alo[j] += t[i - j] * lo; ahi[j] += t[i - j] * hi; }   // one IMAD = two pixels (sums <= 65535)
  }
#pragma unroll
  for (int j = 0; j < kBvR; j++) *reinterpret_cast<uint2*>(dst + j * 64) = make_uint2(alo[j], ahi[j]);
}

__global__ void __launch_bounds__(256) k_blur7v(const __grid_constant__ Geom g, const __grid_constant__ Batch b) {
  __shared__ __align__(16) uint32_t sV[kBvH * 64];  // column-pass sums, u16 pairs: [row][64 words = 128 columns]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, f = blockIdx.y;
  const uint32_t tinfo = __ldg(b.blurTiles + blockIdx.x);
  const int level = tinfo & 15, tx = (tinfo >> 4) & 0x3fff, ty = tinfo >> 18;
  const LevelGeom& L = g.lv[level];
  const int x0 = tx * kBvW, y0 = ty * kBvH;
  int sp;
  const uint8_t* src = level_ptr(g, b, level, f, sp);

  // ---- A. column pass: warp = strip of kBvR output rows, lane = one word column
  {
    const int ys = y0 + warp * kBvR;      // first output row of the strip
    if (ys < L.h) {
      const int gx = x0 - 4 + 4 * lane;   // first pixel column of this lane's word
      const bool edge = (gx < 0) || (gx + 3 >= L.w);
      int wb = 0, wb2 = 0;
      uint32_t sel = 0x3210u;
      if (edge) {
        const int c0 = bv_refl(gx, L.w), c1 = bv_refl(gx + 1, L.w), c2 = bv_refl(gx + 2, L.w), c3 = bv_refl(gx + 3, L.w);
        wb = min(min(c0, c1), min(c2, c3)) & ~3;
        wb2 = min(wb + 4, sp - 4);          // never selected when it had to be clamped: all four columns are then inside word wb
        sel = (uint32_t)min(c0 - wb, 7) | ((uint32_t)min(c1 - wb, 7) << 4) | ((uint32_t)min(c2 - wb, 7) << 8) | ((uint32_t)min(c3 - wb, 7) << 12);
      } else {
        wb = gx; wb2 = gx;
      }
      uint32_t* dst = &sV[(warp * kBvR) * 64 + 2 * lane];
      if (__any_sync(0xffffffffu, edge)) bv_column_pass<true>(g, src, sp, L.h, ys, gx, wb, wb2, sel, dst);
      else bv_column_pass<false>(g, src, sp, L.h, ys, gx, wb, wb2, sel, dst);
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
      asm volatile("" : "+l"(dst));   // keep the per-thread base whole in registers (else its uniform part is re-added per row)
      const unsigned ubp = (unsigned)L.blurPitch;
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
        // saturate_cast<uchar>: the OCV346 taps sum to 257, so a bright patch reaches 257.5 in Q16.16 (a >> 16 <= 257): the
        // upper halves of two sums are clamped together as a u16 pair
        const uint32_t o01 = __vminu2(__byte_perm(a0, a1, 0x7632), 0x00ff00ffu), o23 = __vminu2(__byte_perm(a2, a3, 0x7632), 0x00ff00ffu);
        *reinterpret_cast<uint32_t*>(dst + (unsigned long long)(unsigned)y * ubp) = __byte_perm(o01, o23, 0x6420);
      }
    }
  }
}

}  // namespace orbx
