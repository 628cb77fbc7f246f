// k_fast_nms3 -- FAST-9/16 corner score + per-cell 3x3 NMS with in-CTA compaction (third version; SURVEY.md §9.3,
// reference call sites src/ORBextractor.cc:808,827 -> cv::FAST(cell, th, true)).
//   S(p) = max over the 16 arcs of 9 ring pixels of max(min(ring) - v, v - max(ring))
//   keypoint of its cell at threshold t  <=>  S > t  and  S strictly greater than S of the 8 neighbours inside the same cell
// What the profile of k_fast_nms said (profiles/r01b): ALU pipe 70 % busy -- every min/max form issues at 2 warp-inst/clk/SM
// (tools/microbench/minmax_rate.cu) -- 39 % of the instructions in the score loop, 40 % in a dense NMS pass, and 45 % of the
// stall samples on the un-unrolled staging loads.  Only ~10 % of the pixels of a frame pass the compass bound, but a 64-px
// warp row contains one of them 2 times out of 3, so skipping whole warp rows saved little.  This version:
//   0. stages a 70x70 window for a 62x62 output tile (halo overhead 27 % instead of 43 %), all loads of a warp issued
//      before the first use; words are stored at their natural 4-byte alignment (no funnel shift), the two u16-pair copies
//      (even / odd first column) swap roles with the parity of the window origin instead;
//   1. evaluates the 4-point compass bound for every pixel pair and appends the survivors to a shared-memory list
//      (ballot + one shared atomic per warp row);
//   2. computes the full 16-arc score only for list entries (dense warps again: VIMNMX3.U16x2, 82 per pair);
//   3. runs the NMS / emission only for list entries whose score is non-zero.
// Output: pre-candidate records (x, y, cell, S) with S > minTh in arbitrary order + per-cell "has S > iniTh" flags; the
// quadtree kernel orders them and applies the per-cell threshold fallback (src/ORBextractor.cc:825-828).
// Algorithmic bytes per launch: read every level once.
#pragma once
#include "dev_common.cuh"
#include "geom.h"

namespace orbx {

constexpr int kF3TileW = 62, kF3TileH = 62;   // NMS outputs per CTA
constexpr int kF3SW = 64, kF3SH = 64;         // score tile (one halo pixel each side)
constexpr int kF3WinH = kF3SH + 6;            // 70 window rows
constexpr int kF3Pitch = 38;                  // u16-pair words per window row (76 columns >= 3 + 70 + 1)
constexpr int kF3RowsPerWarp = (kF3WinH + 7) / 8;

struct F3Bases { const uint32_t* e; const uint32_t* o; };   // pair copies for even / odd dx, offsets folded in

template <int DX, int DY>
__device__ __forceinline__ uint32_t f3ring(const F3Bases& s, int idx) {   // idx = (sr + 3) * pitch + p
  constexpr bool odd = (DX & 1) != 0;
  if (odd) return s.o[idx + DY * kF3Pitch + (DX + 1) / 2];
  return s.e[idx + DY * kF3Pitch + DX / 2];
}

__global__ void __launch_bounds__(256) k_fast_nms3(const __grid_constant__ Geom g, const __grid_constant__ Batch b) {
  __shared__ __align__(16) uint32_t sA[kF3WinH * kF3Pitch];   // (col 2j, 2j+1) relative to the aligned window origin
  __shared__ __align__(16) uint32_t sB[kF3WinH * kF3Pitch];   // (col 2j+1, 2j+2)
  __shared__ uint32_t sS[kF3SH * 32 + 64];                    // 64x64 scores as u16 pairs, 32 words of slack each side
  __shared__ uint16_t sList[kF3SH * 32];                      // surviving pairs: sr * 32 + p
  __shared__ uint16_t sCol[kF3SW], sRow[kF3SH];               // cell info of the tile's columns / rows
  __shared__ int sN;

  const int tid = threadIdx.x, warp = tid >> 5, p = tid & 31;
  const int f = blockIdx.y;
  int level = 0;
  while (level + 1 < g.nlevels && (int)blockIdx.x >= g.lv[level + 1].tileBase) level++;
  const LevelGeom& L = g.lv[level];
  const int t = blockIdx.x - L.tileBase;
  const int ty = t / L.tilesX, tx = t - ty * L.tilesX;
  const int ox = kEdge + kF3TileW * tx, oy = kEdge + kF3TileH * ty;  // first output pixel of the tile
  const int wx0 = ox - 4, wy0 = oy - 4;                              // window origin (>= 15)
  const int a = wx0 & 3, gx0 = wx0 & ~3;
  const int maxX = L.w - kEdge, maxY = L.h - kEdge;                  // region = [kEdge, max)
  int sp;
  const uint8_t* src = level_ptr(g, b, level, f, sp);
  uint32_t* const S = sS + 32;

  if (tid == 0) sN = 0;
  if (tid >= 192) { sS[tid - 192 >= 32 ? kF3SH * 32 + tid - 192 : tid - 192] = 0; }   // slack words are read as (masked) neighbours: keep them <= 255 per half
  if (tid < kF3SW) {
    const int x = ox - 1 + tid;
    sCol[tid] = (x >= kEdge && x < maxX) ? b.info[L.colInfoOff + x] : (uint16_t)0;
  } else if (tid < kF3SW + kF3SH) {
    const int y = oy - 1 + (tid - kF3SW);
    sRow[tid - kF3SW] = (y >= kEdge && y < maxY) ? b.info[L.rowInfoOff + y] : (uint16_t)0;
  }

  // ---- 0. stage the window as two u16-pair copies (all loads first)
  {
    const int gx = min(gx0 + 4 * p, sp - 4);
    const uint8_t* colp = src + gx;
    uint32_t w[kF3RowsPerWarp];
#pragma unroll
    for (int k = 0; k < kF3RowsPerWarp; k++) {
      const int r = warp + 8 * k;
      const int gy = min(wy0 + r, L.h - 1);
      w[k] = 0;
      if (r < kF3WinH && p < 19) w[k] = __ldg(reinterpret_cast<const uint32_t*>(colp + (long long)gy * sp));
    }
#pragma unroll
    for (int k = 0; k < kF3RowsPerWarp; k++) {
      const int r = warp + 8 * k;
      const uint32_t u = w[k];
      const uint32_t un = __shfl_down_sync(ORBX_FULL, u, 1) & 0xffu;   // first byte of the next word
      if (r < kF3WinH && p < 19) {
        uint2 va, vb;
        va.x = __byte_perm(u, 0u, 0x4140); va.y = __byte_perm(u, 0u, 0x4342);
        vb.x = __byte_perm(u, 0u, 0x4241); vb.y = __byte_perm(u, un, 0x5453);
        *reinterpret_cast<uint2*>(&sA[r * kF3Pitch + 2 * p]) = va;
        *reinterpret_cast<uint2*>(&sB[r * kF3Pitch + 2 * p]) = vb;
      }
    }
  }
  // the pair of lane p starts at window column 2p+3, i.e. column c0 = 2p + 3 + a of the staged row:
  //   a odd  -> c0 even: centre / even dx from sA at p + (3+a)/2 + dx/2,   odd dx from sB at p + (a+1)/2 + (dx+1)/2
  //   a even -> c0 odd:  centre / even dx from sB at p + 1 + a/2 + dx/2,   odd dx from sA at p + a/2 + 1 + (dx+1)/2
  F3Bases sb;
  if (a & 1) { sb.e = sA + (3 + a) / 2; sb.o = sB + (a + 1) / 2; }
  else { sb.e = sB + 1 + a / 2; sb.o = sA + a / 2 + 1; }
  __syncthreads();

  // ---- 1. compass bound for every pair: every 9-arc holds one of (r0|r8) and one of (r4|r12)  =>  S <= Q
  const uint32_t kBias = 0x01000100u;
  const uint32_t thrB = (uint32_t)g.minTh * 0x00010001u + kBias;
  {
    const bool colNeeded = (ox - 1 + 2 * p) < maxX;
#pragma unroll 2
    for (int sr = warp; sr < kF3SH; sr += 8) {
      S[sr * 32 + p] = 0;
      if (oy - 1 + sr >= maxY) continue;   // warp-uniform: rows below the region are never read as neighbours of a kept pixel
      const int idx = (sr + 3) * kF3Pitch + p;
      const uint32_t v = f3ring<0, 0>(sb, idx);
      const uint32_t r0 = f3ring<0, 3>(sb, idx), r8 = f3ring<0, -3>(sb, idx);
      const uint32_t r4 = f3ring<3, 0>(sb, idx), r12 = f3ring<-3, 0>(sb, idx);
      const uint32_t hi = __vminu2(__vmaxu2(r0, r8), __vmaxu2(r4, r12));
      const uint32_t lo = __vmaxu2(__vminu2(r0, r8), __vminu2(r4, r12));
      const uint32_t q = __vmaxu2((hi + kBias) - v, (v + kBias) - lo);  // Q + 256 per half
      const bool pass = colNeeded && (__vmaxu2(q, thrB) != thrB);
      const uint32_t m = __ballot_sync(ORBX_FULL, pass);
      if (m) {
        int base = 0;
        if (p == 0) base = atomicAdd(&sN, __popc(m));
        base = __shfl_sync(ORBX_FULL, base, 0);
        if (pass) sList[base + __popc(m & ((1u << p) - 1))] = (uint16_t)(sr * 32 + p);
      }
    }
  }
  __syncthreads();
  const int n = sN;

  // ---- 2. full score of the surviving pairs
#pragma unroll 1
  for (int i = tid; i < n; i += 256) {
    const int e = sList[i];
    const int idx = ((e >> 5) + 3) * kF3Pitch + (e & 31);
    const uint32_t v = f3ring<0, 0>(sb, idx);
    uint32_t r[16];
    r[0] = f3ring<0, 3>(sb, idx);   r[1] = f3ring<1, 3>(sb, idx);    r[2] = f3ring<2, 2>(sb, idx);    r[3] = f3ring<3, 1>(sb, idx);
    r[4] = f3ring<3, 0>(sb, idx);   r[5] = f3ring<3, -1>(sb, idx);   r[6] = f3ring<2, -2>(sb, idx);   r[7] = f3ring<1, -3>(sb, idx);
    r[8] = f3ring<0, -3>(sb, idx);  r[9] = f3ring<-1, -3>(sb, idx);  r[10] = f3ring<-2, -2>(sb, idx); r[11] = f3ring<-3, -1>(sb, idx);
    r[12] = f3ring<-3, 0>(sb, idx); r[13] = f3ring<-3, 1>(sb, idx);  r[14] = f3ring<-2, 2>(sb, idx);  r[15] = f3ring<-1, 3>(sb, idx);
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
    S[e] = __vmaxu2(__vmaxu2((mxn + kBias) - v, (v + kBias) - mnx), kBias) - kBias;  // max(S, 0) per half
  }
  __syncthreads();

  // ---- 3. NMS inside the owning cell + emission of pre-candidates (S > minTh), list entries only
  const uint32_t thr2 = (uint32_t)g.minTh * 0x00010001u;
  PreCand* cand = b.cand + (long long)f * g.candSlab + L.candOff;
  int* candCount = b.candCount + f * ORBX_MAX_LEVELS + level;
  int* cellFlags = b.cellFlags + (long long)f * g.nCellsTotal + L.cellBase;
#pragma unroll 1
  for (int base0 = 0; base0 < n; base0 += 256) {
    const int i = base0 + tid;
    bool keepA = false, keepB = false;
    uint32_t cc = 0, ciA = 0, ciB = 0, ri = 0;
    int xa = 0, y = 0;
    if (i < n) {
      const int e = sList[i];
      const int sr = e >> 5, q = e & 31;
      cc = S[e];
      y = oy - 1 + sr; xa = ox - 1 + 2 * q;
      if (cc != 0 && sr >= 1 && sr <= kF3TileH && y < maxY) {
        const bool okA = (q >= 1) && (xa < maxX), okB = (q <= 30) && (xa + 1 < maxX);
        ciA = sCol[2 * q]; ciB = sCol[2 * q + 1]; ri = sRow[sr];
        // the left neighbour of a first-column pixel and the right neighbour of a last-column pixel belong to another cell
        const uint32_t maskL = ((ciA & 0x4000u) ? 0u : 0x0000ffffu) | ((ciB & 0x4000u) ? 0u : 0xffff0000u);
        const uint32_t maskR = ((ciA & 0x8000u) ? 0u : 0x0000ffffu) | ((ciB & 0x8000u) ? 0u : 0xffff0000u);
        const uint32_t valid = (okA ? 0x0000ffffu : 0u) | (okB ? 0xffff0000u : 0u);
        const uint32_t pc = S[e - 1], nc = S[e + 1];
        uint32_t nb = __vmaxu2(__funnelshift_r(pc, cc, 16) & maskL, __funnelshift_r(cc, nc, 16) & maskR);
        if (!(ri & 0x4000u)) {
          const uint32_t pu = S[e - 33], cu = S[e - 32], nu = S[e - 31];
          nb = __vimax3_u16x2(nb, cu, __vmaxu2(__funnelshift_r(pu, cu, 16) & maskL, __funnelshift_r(cu, nu, 16) & maskR));
        }
        if (!(ri & 0x8000u)) {
          const uint32_t pd = S[e + 31], cd = S[e + 32], nd = S[e + 33];
          nb = __vimax3_u16x2(nb, cd, __vmaxu2(__funnelshift_r(pd, cd, 16) & maskL, __funnelshift_r(cd, nd, 16) & maskR));
        }
        nb = __vmaxu2(nb, thr2);
        const uint32_t tt = ((cc & valid) + kBias) - nb;      // per half: > 0x100  <=>  S > max(neighbours, minTh)
        keepA = (tt & 0xffffu) > 0x100u; keepB = (tt >> 16) > 0x100u;
      }
    }
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
