// k_pyr_resize2 -- one pyramid level from the previous one, cv::resize(INTER_LINEAR) on 8U in OpenCV's fixed point
// (reference call site src/ORBextractor.cc:1201).  Tuned replacement of k_pyr_resize (extract_kernels.cuh),
// which issued 16 byte-wide LDGs per thread and was LSU/latency bound (profiles/r01: issue 30 %, DRAM 5 %).
//   CTA = 128x16 destination pixels of one frame.
//   1. the source window (<= ~156x21 px at scale 1.2) is staged with aligned 32-bit loads, coalesced along rows;
//   2. horizontal pass ONCE per source row (OpenCV's order: rows are interpolated horizontally first):
//      h = S[sx]*a0 + S[sx+1]*a1, kept as (h >> 4) in shared memory; a thread owns one destination column, so its
//      table entry (source columns + Q11 weights) lives in registers for all window rows;
//   3. vertical pass: thread = 4 destination pixels of one row, two LDS.128 (rows sy, sy+1), one 32-bit store:
//      v = ((b0*(h0>>4))>>16) + ((b1*(h1>>4))>>16) + 2 >> 2.
// Algorithmic bytes per launch: read level l-1 + write level l.
#pragma once
#include "dev_common.cuh"
#include "geom.h"

namespace orbx {

constexpr int kRsTileW = 128, kRsTileH = 16;

__global__ void __launch_bounds__(256) k_pyr_resize2(const __grid_constant__ Geom g, const __grid_constant__ Batch b, int level) {
  extern __shared__ __align__(16) uint32_t rs_smem[];
  const LevelGeom& L = g.lv[level];
  const int winWords = g.rsWinWords, winRows = g.rsWinRows;
  uint32_t* sRaw = rs_smem;                                               // [winRows][winWords]  source bytes
  int* sH = reinterpret_cast<int*>(rs_smem + winRows * winWords);          // [winRows][128]       (h >> 4)
  const int tid = threadIdx.x, f = blockIdx.z;
  const int x0 = blockIdx.x * kRsTileW, y0 = blockIdx.y * kRsTileH;
  int sp;
  const uint8_t* src = level_ptr(g, b, level - 1, f, sp);
  const RsX* tx = b.rsx + L.rsxOff;
  const RsY* ty = b.rsy + L.rsyOff;
  const int wpad = (L.w + 3) & ~3;                                         // the column table is padded to a multiple of 4
  const int xl = min(x0 + kRsTileW, wpad) - 1, yl = min(y0 + kRsTileH, L.h) - 1;
  const int gx0 = (int)tx[x0].s0 & ~3;                                     // first staged source column (word aligned)
  const int nW = (((int)tx[xl].s1 - gx0) >> 2) + 1;                        // words per window row   (<= winWords)
  const int wy0 = ty[y0].s0;
  const int nR = (int)ty[yl].s1 - wy0 + 1;                                 // window rows            (<= winRows)

  // ---- 1. stage the window
  for (int r = tid >> 5; r < nR; r += 8) {
    const uint32_t* row = reinterpret_cast<const uint32_t*>(src + (long long)(wy0 + r) * sp + gx0);
    for (int k = tid & 31; k < nW; k += 32) sRaw[r * winWords + k] = __ldg(row + k);
  }
  __syncthreads();

  // ---- 2. horizontal pass: thread = destination column c, row group rg
  {
    const int c = tid & (kRsTileW - 1), rg = tid >> 7;
    if (x0 + c < wpad) {
      const RsX e = tx[x0 + c];
      const int o0 = e.s0 - gx0, o1 = e.s1 - gx0, a0 = e.a0, a1 = e.a1;
      const uint8_t* rawb = reinterpret_cast<const uint8_t*>(sRaw);
#pragma unroll 4
      for (int r = rg; r < nR; r += 2) {
        const uint8_t* rb = rawb + r * (winWords * 4);
        const int h = (int)rb[o0] * a0 + (int)rb[o1] * a1;
        sH[r * kRsTileW + c] = h >> 4;
      }
    }
  }
  __syncthreads();

  // ---- 3. vertical pass: task = 4 destination pixels
  uint8_t* dstBase = b.pyr + (long long)f * b.pyrFrameStride + L.pyrOff;
#pragma unroll
  for (int it = 0; it < (kRsTileW / 4) * kRsTileH / 256; it++) {
    const int task = tid + it * 256;
    const int q = task & 31, r = task >> 5;
    const int x = x0 + 4 * q, y = y0 + r;
    if (x >= L.w || y >= L.h) continue;
    const RsY e = ty[y];
    const int b0 = e.b0, b1 = e.b1;
    const int4 h0 = *reinterpret_cast<const int4*>(sH + (e.s0 - wy0) * kRsTileW + 4 * q);
    const int4 h1 = *reinterpret_cast<const int4*>(sH + (e.s1 - wy0) * kRsTileW + 4 * q);
    const int v0 = min(255, max(0, (((b0 * h0.x) >> 16) + ((b1 * h1.x) >> 16) + 2) >> 2));
    const int v1 = min(255, max(0, (((b0 * h0.y) >> 16) + ((b1 * h1.y) >> 16) + 2) >> 2));
    const int v2 = min(255, max(0, (((b0 * h0.z) >> 16) + ((b1 * h1.z) >> 16) + 2) >> 2));
    const int v3 = min(255, max(0, (((b0 * h0.w) >> 16) + ((b1 * h1.w) >> 16) + 2) >> 2));
    *reinterpret_cast<uint32_t*>(dstBase + (long long)y * L.pitch + x) = (uint32_t)v0 | ((uint32_t)v1 << 8) | ((uint32_t)v2 << 16) | ((uint32_t)v3 << 24);
  }
}



// ---------------------------------------------------------------------------------------------------------------------------
// k_pyr_resize3 -- register-resident version (profiles/r01c: k_pyr_resize2 executed ~70 thread-instructions per destination
// pixel, issue-bound at 82 %: partially filled staging loops, one shared-memory round trip for the horizontal pass, scalar
// vertical arithmetic with clamps).
//   thread = 4 destination columns (one 32-bit store) walking down a strip of kRs3Rows destination rows;
//   per source row: three aligned 32-bit loads cover the <= 8 source bytes the 4 columns touch (scale <= 2), two funnel shifts
//   line them up at the first source column, and per column one PRMT picks its two taps and one IDP.2A computes
//   S[sx]*a0 + S[sx+1]*a1 (the Q11 weights as a u16 pair); source rows are interpolated horizontally once and carried in
//   registers from one destination row to the next (the row tables are warp-uniform, so the reuse test does not diverge);
//   vertical: ((b0*(h0>>4))>>16) + ((b1*(h1>>4))>>16) as two IMAD.HI with the weights pre-shifted by 16, then (+2)>>2.
//   No clamp: h>>4 <= 32655 and b0+b1 <= 2049 bound the result by 255.
// Levels whose 4-column groups span more than 8 source bytes (scale factor > 2) use k_pyr_resize2.
// Measured alternatives: strips of 8 / 16 rows (0.197 / 0.219 ms against 0.190 for 4), and loading every source row of a
// strip before the first use (80 registers, 0.207 ms).
// ---------------------------------------------------------------------------------------------------------------------------
constexpr int kRs3Rows = 4;                    // destination rows per warp
constexpr int kRs3TileW = 128, kRs3TileH = 8 * kRs3Rows;

struct Rs3Col { uint32_t sel[4]; uint32_t wt[4]; int base; uint32_t sh; };

__device__ __forceinline__ void rs3_hrow(const uint8_t* row, int sp, const Rs3Col& c, uint32_t (&h)[4]) {
  const int o1 = min(c.base + 4, sp - 4), o2 = min(c.base + 8, sp - 4);   // words past the row end hold no tap (s1 <= w-1 < sp)
  const uint32_t w0 = __ldg(reinterpret_cast<const uint32_t*>(row + c.base));
  const uint32_t w1 = __ldg(reinterpret_cast<const uint32_t*>(row + o1));
  const uint32_t w2 = __ldg(reinterpret_cast<const uint32_t*>(row + o2));
  const uint32_t u0 = __funnelshift_r(w0, w1, c.sh), u1 = __funnelshift_r(w1, w2, c.sh);   // 8 bytes from the first source column on
#pragma unroll
  for (int k = 0; k < 4; k++) h[k] = __dp2a_lo(c.wt[k], __byte_perm(u0, u1, c.sel[k]), 0u) >> 4;
}

__global__ void __launch_bounds__(256) k_pyr_resize3(const __grid_constant__ Geom g, const __grid_constant__ Batch b, int level) {
  const LevelGeom& L = g.lv[level];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, f = blockIdx.z;
  const int x = blockIdx.x * kRs3TileW + 4 * lane;
  const int ys = blockIdx.y * kRs3TileH + warp * kRs3Rows;
  const int wpad = (L.w + 3) & ~3;   // the column table is padded to a multiple of 4
  if (x >= wpad || ys >= L.h) return;
  int sp;
  const uint8_t* src = level_ptr(g, b, level - 1, f, sp);
  const RsY* ty = b.rsy + L.rsyOff;
  Rs3Col c;
  {
    const uint4* tx = reinterpret_cast<const uint4*>(b.rsx + L.rsxOff + x);   // 4 entries of 8 bytes
    const uint4 p0 = __ldg(tx), p1 = __ldg(tx + 1);
    const uint32_t e[8] = {p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w};   // (s0 | s1 << 16), (a0 | a1 << 16) per column
    const int sx0 = (int)(e[0] & 0xffffu);
    c.base = sx0 & ~3; c.sh = (uint32_t)(sx0 & 3) * 8;
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int s0 = (int)(e[2 * k] & 0xffffu), s1 = (int)(e[2 * k] >> 16);
      c.sel[k] = (uint32_t)min(s0 - sx0, 7) | ((uint32_t)min(s1 - sx0, 7) << 4);   // upper two bytes: don't care (byte 0)
      c.wt[k] = e[2 * k + 1];
    }
  }
  uint8_t* dst = b.pyr + (long long)f * b.pyrFrameStride + L.pyrOff + x;
  int r0 = -1, r1 = -1;
  uint32_t hA[4] = {0, 0, 0, 0}, hB[4] = {0, 0, 0, 0};
#pragma unroll
  for (int j = 0; j < kRs3Rows; j++) {
    const int y = ys + j;
    if (y >= L.h) break;
    const RsY e = ty[y];
    const int s0 = e.s0, s1 = e.s1;
    if (s0 != r0) {
      if (s0 == r1) {
#pragma unroll
        for (int k = 0; k < 4; k++) hA[k] = hB[k];
      } else rs3_hrow(src + (long long)s0 * sp, sp, c, hA);
      r0 = s0;
    }
    if (s1 != r1) {
      if (s1 == r0) {
#pragma unroll
        for (int k = 0; k < 4; k++) hB[k] = hA[k];
      } else rs3_hrow(src + (long long)s1 * sp, sp, c, hB);
      r1 = s1;
    }
    const uint32_t B0 = (uint32_t)(uint16_t)e.b0 << 16, B1 = (uint32_t)(uint16_t)e.b1 << 16;
    uint32_t v[4];
#pragma unroll
    for (int k = 0; k < 4; k++) v[k] = (__umulhi(B0, hA[k]) + __umulhi(B1, hB[k]) + 2u) >> 2;
    *reinterpret_cast<uint32_t*>(dst + (long long)y * L.pitch) = v[0] | (v[1] << 8) | (v[2] << 16) | (v[3] << 24);
  }
}
}  // namespace orbx
