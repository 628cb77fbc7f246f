// k_fast_nms -- FAST-9/16 corner score + per-cell 3x3 NMS, tuned version (see extract_kernels.cuh for the scheme).
//   S(p) = max over the 16 arcs of 9 ring pixels of max(min(ring) - v, v - max(ring))      (cv::FAST)
//   keypoint of its cell at threshold t  <=>  S > t  and  S strictly greater than S of the 8 neighbours inside the same cell
// Differences from v1 (profiles/r01: 38 % of the instructions were the scalar NMS, 12 % staging index arithmetic):
//   * staging: one aligned 32-bit LDG per 4 window bytes, funnel-shift to the window alignment, PRMT to the two u16-pair
//     copies, 64-bit STS; rows are walked warp-per-row so no integer division is needed
//   * S is kept as u16 pairs in shared memory; NMS runs on pairs too: neighbours of both pixels of a lane come from the
//     lane's own / adjacent words (SHFL + funnel shift), 3x3 maximum as VIMNMX3.U16x2, cell borders as AND masks
//   * validity / cell flags are per-lane or per-row constants hoisted out of the loops
#pragma once
#include "dev_common.cuh"
#include "geom.h"

namespace orbx {

constexpr int kFWinH = kTileH + 8;   // 38 window rows
constexpr int kFPairWords = 36;      // 72 window columns as u16 pairs

template <int DX, int DY>
__device__ __forceinline__ uint32_t fring(const uint32_t* sA, const uint32_t* sB, int rowBase, int p) {
  // rowBase = (sr + 3) * kFPairWords; the pair of lane p starts at window column 2p+3
  constexpr bool odd = (DX & 1) != 0;
  if (odd) return sA[rowBase + DY * kFPairWords + p + (3 + DX) / 2];
  return sB[rowBase + DY * kFPairWords + p + 1 + DX / 2];
}

__global__ void __launch_bounds__(256) k_fast_nms(const __grid_constant__ Geom g, const __grid_constant__ Batch b) {
  __shared__ __align__(16) uint32_t sA[kFWinH * kFPairWords];  // (col 2j, 2j+1)
  __shared__ __align__(16) uint32_t sB[kFWinH * kFPairWords];  // (col 2j+1, 2j+2)
  __shared__ uint32_t sS[32 * 32];                             // 64x32 scores as u16 pairs

  const int tid = threadIdx.x, warp = tid >> 5, p = tid & 31;
  const int f = blockIdx.y;
  int level = 0;
  while (level + 1 < g.nlevels && (int)blockIdx.x >= g.lv[level + 1].tileBase) level++;
  const LevelGeom& L = g.lv[level];
  const int t = blockIdx.x - L.tileBase;
  const int ty = t / L.tilesX, tx = t - ty * L.tilesX;
  const int ox = kEdge + kTileW * tx, oy = kEdge + kTileH * ty;  // first output pixel of the tile
  const int wx0 = ox - 4, wy0 = oy - 4;                          // window origin (>= 15)
  const int a8 = (wx0 & 3) * 8, gx0 = wx0 & ~3;
  int sp;
  const uint8_t* src = level_ptr(g, b, level, f, sp);

  // ---- stage the 70x38 window as two u16-pair copies
  {
    const int gx = min(gx0 + 4 * p, sp - 4);
    const uint8_t* colp = src + gx;
    for (int r = warp; r < kFWinH; r += 8) {
      const int gy = min(wy0 + r, L.h - 1);
      uint32_t gw = 0;
      if (p < 20) gw = __ldg(reinterpret_cast<const uint32_t*>(colp + (long long)gy * sp));
      const uint32_t gn = __shfl_down_sync(ORBX_FULL, gw, 1);
      const uint32_t u = __funnelshift_r(gw, gn, a8);              // window bytes 4p .. 4p+3
      const uint32_t un = __shfl_down_sync(ORBX_FULL, u, 1) & 0xffu;  // window byte 4p+4
      if (p < 18) {
        uint2 va, vb;
        va.x = __byte_perm(u, 0u, 0x4140); va.y = __byte_perm(u, 0u, 0x4342);
        vb.x = __byte_perm(u, 0u, 0x4241); vb.y = __byte_perm(u, un, 0x5453);
        *reinterpret_cast<uint2*>(&sA[r * kFPairWords + 2 * p]) = va;
        *reinterpret_cast<uint2*>(&sB[r * kFPairWords + 2 * p]) = vb;
      }
    }
  }
  __syncthreads();

  // ---- corner score of the pixel pair (columns 2p, 2p+1 of the 64-wide score tile)
  const uint32_t kBias = 0x01000100u;
  const uint32_t thrB = (uint32_t)g.minTh * 0x00010001u + kBias;
#pragma unroll 1
  for (int sr = warp; sr < 32; sr += 8) {
    const int rb = (sr + 3) * kFPairWords;
    const uint32_t v = sB[rb + p + 1];
    uint32_t S2 = 0;
    // compass bound: every 9-arc holds one of (r0|r8) and one of (r4|r12)  =>  S <= Q
    const uint32_t r0 = fring<0, 3>(sA, sB, rb, p), r8 = fring<0, -3>(sA, sB, rb, p);
    const uint32_t r4 = fring<3, 0>(sA, sB, rb, p), r12 = fring<-3, 0>(sA, sB, rb, p);
    const uint32_t hi = __vminu2(__vmaxu2(r0, r8), __vmaxu2(r4, r12));
    const uint32_t lo = __vmaxu2(__vminu2(r0, r8), __vminu2(r4, r12));
    const uint32_t q = __vmaxu2((hi + kBias) - v, (v + kBias) - lo);  // Q + 256 per lane half
    if (__any_sync(ORBX_FULL, __vmaxu2(q, thrB) != thrB)) {           // some pixel of the row segment may exceed minTh
      const uint32_t r1 = fring<1, 3>(sA, sB, rb, p), r2 = fring<2, 2>(sA, sB, rb, p), r3 = fring<3, 1>(sA, sB, rb, p);
      const uint32_t r5 = fring<3, -1>(sA, sB, rb, p), r6 = fring<2, -2>(sA, sB, rb, p), r7 = fring<1, -3>(sA, sB, rb, p);
      const uint32_t r9 = fring<-1, -3>(sA, sB, rb, p), r10 = fring<-2, -2>(sA, sB, rb, p), r11 = fring<-3, -1>(sA, sB, rb, p);
      const uint32_t r13 = fring<-3, 1>(sA, sB, rb, p), r14 = fring<-2, 2>(sA, sB, rb, p), r15 = fring<-1, 3>(sA, sB, rb, p);
      const uint32_t r[16] = {r0, r1, r2, r3, r4, r5, r6, r7, r8, r9, r10, r11, r12, r13, r14, r15};
      uint32_t n3[16], x3[16];
#pragma unroll
      for (int k = 0; k < 16; k++) {
        n3[k] = __vimin3_u16x2(r[k], r[(k + 1) & 15], r[(k + 2) & 15]);
        x3[k] = __vimax3_u16x2(r[k], r[(k + 1) & 15], r[(k + 2) & 15]);
      }
      uint32_t mxn = 0, mnx = 0xffffffffu;
#pragma unroll
      for (int k = 0; k < 16; k += 2) {  // two 9-arcs per step: max3 / min3 absorb both
        const uint32_t na = __vimin3_u16x2(n3[k], n3[(k + 3) & 15], n3[(k + 6) & 15]);
        const uint32_t nb = __vimin3_u16x2(n3[k + 1], n3[(k + 4) & 15], n3[(k + 7) & 15]);
        const uint32_t xa = __vimax3_u16x2(x3[k], x3[(k + 3) & 15], x3[(k + 6) & 15]);
        const uint32_t xb = __vimax3_u16x2(x3[k + 1], x3[(k + 4) & 15], x3[(k + 7) & 15]);
        mxn = __vimax3_u16x2(mxn, na, nb);
        mnx = __vimin3_u16x2(mnx, xa, xb);
      }
      S2 = __vmaxu2(__vmaxu2((mxn + kBias) - v, (v + kBias) - mnx), kBias) - kBias;  // max(S, 0) per half
    }
    sS[sr * 32 + p] = S2;
  }
  __syncthreads();

  // ---- NMS inside the owning cell + emission of pre-candidates (S > minTh)
  const int xa = ox - 1 + 2 * p;  // level x of the lane's low half; high half = xa + 1
  uint32_t ciA = 0, ciB = 0;
  const bool okA = (p >= 1) && (xa < L.w - kEdge), okB = (p <= 30) && (xa + 1 < L.w - kEdge);
  if (okA) ciA = b.info[L.colInfoOff + xa];
  if (okB) ciB = b.info[L.colInfoOff + xa + 1];
  // neighbour masks: the left neighbour of a first-column pixel and the right neighbour of a last-column pixel belong to another cell
  const uint32_t maskL = ((ciA & 0x4000u) ? 0u : 0x0000ffffu) | ((ciB & 0x4000u) ? 0u : 0xffff0000u);
  const uint32_t maskR = ((ciA & 0x8000u) ? 0u : 0x0000ffffu) | ((ciB & 0x8000u) ? 0u : 0xffff0000u);
  const uint32_t valid = (okA ? 0x0000ffffu : 0u) | (okB ? 0xffff0000u : 0u);
  const uint32_t thr2 = (uint32_t)g.minTh * 0x00010001u;
  PreCand* cand = b.cand + (long long)f * g.candSlab + L.candOff;
  int* candCount = b.candCount + f * ORBX_MAX_LEVELS + level;
  int* cellFlags = b.cellFlags + (long long)f * g.nCellsTotal + L.cellBase;
#pragma unroll 1
  for (int sr = 1 + warp; sr <= kTileH; sr += 8) {
    const int y = oy - 1 + sr;
    if (y >= L.h - kEdge) break;  // rows below the region (warp-uniform)
    const uint32_t ri = b.info[L.rowInfoOff + y];
    const uint32_t cu = sS[(sr - 1) * 32 + p], cc = sS[sr * 32 + p], cd = sS[(sr + 1) * 32 + p];
    if (!__any_sync(ORBX_FULL, (cc & valid) != 0)) continue;
    const uint32_t pu = __shfl_up_sync(ORBX_FULL, cu, 1), pc = __shfl_up_sync(ORBX_FULL, cc, 1), pd = __shfl_up_sync(ORBX_FULL, cd, 1);
    const uint32_t nu = __shfl_down_sync(ORBX_FULL, cu, 1), nc = __shfl_down_sync(ORBX_FULL, cc, 1), nd = __shfl_down_sync(ORBX_FULL, cd, 1);
    // left / right neighbours of both halves: (col-1 | col) and (col+1 | col+2) of the pair
    const uint32_t lC = __funnelshift_r(pc, cc, 16) & maskL, rC = __funnelshift_r(cc, nc, 16) & maskR;
    uint32_t nb = __vmaxu2(lC, rC);
    if (!(ri & 0x4000u)) nb = __vimax3_u16x2(nb, cu, __vmaxu2(__funnelshift_r(pu, cu, 16) & maskL, __funnelshift_r(cu, nu, 16) & maskR));
    if (!(ri & 0x8000u)) nb = __vimax3_u16x2(nb, cd, __vmaxu2(__funnelshift_r(pd, cd, 16) & maskL, __funnelshift_r(cd, nd, 16) & maskR));
    nb = __vmaxu2(nb, thr2);
    const uint32_t tt = ((cc & valid) + kBias) - nb;      // per half: > 0x100  <=>  S > max(neighbours, minTh)
    const bool keepA = (tt & 0xffffu) > 0x100u, keepB = (tt >> 16) > 0x100u;
    const uint32_t mA = __ballot_sync(ORBX_FULL, keepA), mB = __ballot_sync(ORBX_FULL, keepB);
    if ((mA | mB) == 0) continue;
    const int nA = __popc(mA), total = nA + __popc(mB);
    int base = 0;
    if (p == 0) base = atomicAdd(candCount, total);
    base = __shfl_sync(ORBX_FULL, base, 0);
    const int rowCell = (int)(ri & 0x3fffu) * L.nCols;
    if (keepA) {
      const int idx = base + __popc(mA & ((1u << p) - 1)), s = cc & 0xffffu, cell = rowCell + (int)(ciA & 0x3fffu);
      if (idx < L.candCap) {
        PreCand pc2; pc2.x = (uint16_t)xa; pc2.y = (uint16_t)y; pc2.cell = (uint16_t)cell; pc2.S = (uint8_t)s; pc2.pad = 0;
        cand[idx] = pc2;
      } else atomicExch(b.status, ORBX_ERR_CAPACITY);
      if (s > g.iniTh) atomicOr(cellFlags + cell, 1);
    }
    if (keepB) {
      const int idx = base + nA + __popc(mB & ((1u << p) - 1)), s = cc >> 16, cell = rowCell + (int)(ciB & 0x3fffu);
      if (idx < L.candCap) {
        PreCand pc2; pc2.x = (uint16_t)(xa + 1); pc2.y = (uint16_t)y; pc2.cell = (uint16_t)cell; pc2.S = (uint8_t)s; pc2.pad = 0;
        cand[idx] = pc2;
      } else atomicExch(b.status, ORBX_ERR_CAPACITY);
      if (s > g.iniTh) atomicOr(cellFlags + cell, 1);
    }
  }
}

}  // namespace orbx
