// sm_100a kernels of the ORB extractor.  One launch covers a whole batch of frames (blockIdx.y / .z = frame); every
// kernel is integer / byte work bound by issue slots and HBM/L2, so: coalesced 32-bit accesses, shared-memory staging,
// u16x2 SIMD-in-register min/max (VIMNMX3.U16x2) for the FAST arcs, dp4a for the blur taps, warp shuffles for reductions.
// No tensor cores: nothing here is a dense floating-point contraction.
//
// Reference stages (wangyunpeng139/3_ORB_SLAM3_SelfNote):
//   k_pyr_resize   ORBextractor::ComputePyramid            src/ORBextractor.cc:1186-1219  (cv::resize INTER_LINEAR 8U)
//   k_fast_nms     ComputeKeyPointsOctTree cell loop        src/ORBextractor.cc:787-854   (cv::FAST 9/16 + NMS per cell)
//   k_quadtree     DistributeOctTree / DivideNode           src/ORBextractor.cc:479-761
//   k_blur7        GaussianBlur(7x7, 2, REFLECT_101)        src/ORBextractor.cc:1132-1141
//   k_assemble     operator() output ordering               src/ORBextractor.cc:1118,1164-1178
//   k_orient_desc  IC_Angle + computeOrbDescriptor          src/ORBextractor.cc:75-145, 470-477, 1059-1066
#pragma once
#include <cuda_runtime.h>
#include <float.h>

#include "dev_common.cuh"
#include "geom.h"

namespace orbx {

// ---------------------------------------------------------------------------------------------------------------
// Pyramid: level L from level L-1, OpenCV's fixed-point bilinear (Q11 weights).
// Thread = 4 destination pixels (one 32-bit store); block 32x8 threads = 128x8 pixels; grid.z = frame.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_pyr_resize(const __grid_constant__ Geom g, const __grid_constant__ Batch b, int level) {
  const LevelGeom& L = g.lv[level];
  const int x4 = (blockIdx.x * 32 + threadIdx.x) * 4;
  const int y = blockIdx.y * 8 + threadIdx.y;
  const int f = blockIdx.z;
  if (x4 >= L.w || y >= L.h) return;
  int sp;
  const uint8_t* src = level_ptr(g, b, level - 1, f, sp);
  const RsY ry = b.rsy[L.rsyOff + y];
  const uint8_t* r0 = src + (long long)ry.s0 * sp;
  const uint8_t* r1 = src + (long long)ry.s1 * sp;
  uint32_t out = 0;
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const RsX rx = b.rsx[L.rsxOff + x4 + i];  // table is padded to a multiple of 4 entries
    const int h0 = (int)__ldg(r0 + rx.s0) * rx.a0 + (int)__ldg(r0 + rx.s1) * rx.a1;
    const int h1 = (int)__ldg(r1 + rx.s0) * rx.a0 + (int)__ldg(r1 + rx.s1) * rx.a1;
    int v = ((((int)ry.b0 * (h0 >> 4)) >> 16) + (((int)ry.b1 * (h1 >> 4)) >> 16) + 2) >> 2;
    v = min(255, max(0, v));
    out |= (uint32_t)v << (8 * i);
  }
  uint8_t* dst = b.pyr + (long long)f * b.pyrFrameStride + L.pyrOff + (long long)y * L.pitch + x4;
  *reinterpret_cast<uint32_t*>(dst) = out;
}

// ---------------------------------------------------------------------------------------------------------------
// FAST-9/16 score + per-cell 3x3 NMS.
//   S(p) = max over the 16 arcs of 9 ring pixels of max(min(ring) - v, v - max(ring))  (SURVEY.md §9.3); a pixel is a
//   cell keypoint at threshold t iff S > t and S is strictly greater than S of its 8 neighbours that lie in the same
//   cell interior.  The ini/min threshold choice is per cell (src/ORBextractor.cc:825-828): this kernel emits every
//   local maximum with S > minTh as a "pre-candidate" and raises cellFlags[cell] when one has S > iniTh; the quadtree
//   kernel applies the per-cell choice when it loads its level.
// CTA = one 62x30 output tile of one level of one frame.  Window 70x38 pixels is staged in shared memory twice as
// u16 pairs -- copy A = (col 2j, 2j+1), copy B = (col 2j+1, 2j+2) -- so that every ring pixel of a horizontally adjacent
// pixel PAIR is one aligned 32-bit LDS and all arc min/max run as 2-wide VIMNMX3.U16x2.
// ---------------------------------------------------------------------------------------------------------------
constexpr int kWinH = kTileH + 8;   // 38 window rows
constexpr int kRawWords = 20;       // 80 raw bytes per row (70 + alignment slack)
constexpr int kPairWords = 36;      // 72 columns per row as pairs

template <int DX, int DY>
__device__ __forceinline__ uint32_t ring_ld(const uint32_t* sA, const uint32_t* sB, int sr, int p) {
  constexpr bool odd = (DX & 1) != 0;
  const int row = sr + 3 + DY;
  if (odd) return sA[row * kPairWords + p + (3 + DX) / 2];
  return sB[row * kPairWords + p + 1 + DX / 2];
}

// (first version, kept for A/B comparison through ORBX_FAST_V1=1; the tuned kernel is k_fast_nms in fast_nms.cuh)
__global__ void __launch_bounds__(256) k_fast_nms_v1(const __grid_constant__ Geom g, const __grid_constant__ Batch b) {
  __shared__ uint32_t sRaw[kWinH * kRawWords];
  __shared__ uint32_t sA[kWinH * kPairWords];
  __shared__ uint32_t sB[kWinH * kPairWords];
  __shared__ uint16_t sS[32 * 32];  // 64x32 scores as bytes

  const int tid = threadIdx.x;
  const int f = blockIdx.y;
  int level = 0;
  while (level + 1 < g.nlevels && (int)blockIdx.x >= g.lv[level + 1].tileBase) level++;
  const LevelGeom& L = g.lv[level];
  const int t = blockIdx.x - L.tileBase;
  const int tx = t % L.tilesX, ty = t / L.tilesX;
  const int ox = kEdge + kTileW * tx, oy = kEdge + kTileH * ty;  // first output pixel of the tile
  const int wx0 = ox - 4, wy0 = oy - 4;                          // window origin
  const int a = wx0 & 3, gx0 = wx0 - a;
  int sp;
  const uint8_t* src = level_ptr(g, b, level, f, sp);

  // stage 1: aligned 32-bit global loads of the window rows
  for (int i = tid; i < kWinH * kRawWords; i += 256) {
    const int r = i / kRawWords, k = i - r * kRawWords;
    const int gy = min(wy0 + r, L.h - 1);
    const int gx = min(gx0 + 4 * k, sp - 4);
    sRaw[i] = __ldg(reinterpret_cast<const uint32_t*>(src + (long long)gy * sp + gx));
  }
  __syncthreads();
  // stage 2: expand to the two u16-pair copies
  {
    const uint8_t* rawb = reinterpret_cast<const uint8_t*>(sRaw);
    for (int i = tid; i < kWinH * kPairWords; i += 256) {
      const int r = i / kPairWords, j = i - r * kPairWords;
      const int base = r * (kRawWords * 4) + 2 * j + a;
      const uint32_t b0 = rawb[base], b1 = rawb[base + 1], b2 = rawb[base + 2];
      sA[i] = b0 | (b1 << 16);
      sB[i] = b1 | (b2 << 16);
    }
  }
  __syncthreads();

  const int warp = tid >> 5, p = tid & 31;
  const uint32_t kBias = 0x01000100u;
  const uint32_t thr2 = (uint32_t)g.minTh * 0x00010001u;
  for (int sr = warp; sr < 32; sr += 8) {
    const int y = oy - 1 + sr, x0 = ox - 1 + 2 * p;
    const bool rowin = (y >= kEdge) && (y < L.h - kEdge);
    const bool in0 = rowin && (x0 >= kEdge) && (x0 < L.w - kEdge);
    const bool in1 = rowin && (x0 + 1 >= kEdge) && (x0 + 1 < L.w - kEdge);
    const uint32_t v = sB[(sr + 3) * kPairWords + p + 1];
    const uint32_t vb = v + kBias;
    uint32_t S2 = 0;
    // compass bound: every 9-arc holds one of (r0|r8) and one of (r4|r12)  =>  S <= Q
    const uint32_t r0 = ring_ld<0, 3>(sA, sB, sr, p), r8 = ring_ld<0, -3>(sA, sB, sr, p);
    const uint32_t r4 = ring_ld<3, 0>(sA, sB, sr, p), r12 = ring_ld<-3, 0>(sA, sB, sr, p);
    {
      const uint32_t hi = __vminu2(__vmaxu2(r0, r8), __vmaxu2(r4, r12));
      const uint32_t lo = __vmaxu2(__vminu2(r0, r8), __vminu2(r4, r12));
      const uint32_t q = __vmaxu2((hi + kBias) - v, vb - lo);             // biased by 256 per lane
      const uint32_t pass = __vmaxu2(q, thr2 + kBias) ^ (thr2 + kBias);  // != 0 iff some lane has Q > minTh
      const bool need = (pass != 0) && (in0 || in1);
      if (__any_sync(ORBX_FULL, need)) {
        const uint32_t r1 = ring_ld<1, 3>(sA, sB, sr, p), r2 = ring_ld<2, 2>(sA, sB, sr, p), r3 = ring_ld<3, 1>(sA, sB, sr, p);
        const uint32_t r5 = ring_ld<3, -1>(sA, sB, sr, p), r6 = ring_ld<2, -2>(sA, sB, sr, p), r7 = ring_ld<1, -3>(sA, sB, sr, p);
        const uint32_t r9 = ring_ld<-1, -3>(sA, sB, sr, p), r10 = ring_ld<-2, -2>(sA, sB, sr, p), r11 = ring_ld<-3, -1>(sA, sB, sr, p);
        const uint32_t r13 = ring_ld<-3, 1>(sA, sB, sr, p), r14 = ring_ld<-2, 2>(sA, sB, sr, p), r15 = ring_ld<-1, 3>(sA, sB, sr, p);
        const uint32_t r[16] = {r0, r1, r2, r3, r4, r5, r6, r7, r8, r9, r10, r11, r12, r13, r14, r15};
        uint32_t n3[16], x3[16];
#pragma unroll
        for (int k = 0; k < 16; k++) {
          n3[k] = __vimin3_u16x2(r[k], r[(k + 1) & 15], r[(k + 2) & 15]);
          x3[k] = __vimax3_u16x2(r[k], r[(k + 1) & 15], r[(k + 2) & 15]);
        }
        uint32_t n9[16], x9[16];
#pragma unroll
        for (int k = 0; k < 16; k++) {
          n9[k] = __vimin3_u16x2(n3[k], n3[(k + 3) & 15], n3[(k + 6) & 15]);
          x9[k] = __vimax3_u16x2(x3[k], x3[(k + 3) & 15], x3[(k + 6) & 15]);
        }
        uint32_t mxn = __vimax3_u16x2(n9[0], n9[1], n9[2]), mnx = __vimin3_u16x2(x9[0], x9[1], x9[2]);
#pragma unroll
        for (int k = 3; k < 15; k += 2) {
          mxn = __vimax3_u16x2(mxn, n9[k], n9[k + 1]);
          mnx = __vimin3_u16x2(mnx, x9[k], x9[k + 1]);
        }
        mxn = __vmaxu2(mxn, n9[15]);
        mnx = __vminu2(mnx, x9[15]);
        // S + 256 per lane; clamp at 0
        uint32_t s = __vmaxu2(__vmaxu2((mxn + kBias) - v, vb - mnx), kBias) - kBias;
        uint32_t s0 = s & 0xffffu, s1 = s >> 16;
        if (!in0 || s0 <= (uint32_t)g.minTh) s0 = 0;
        if (!in1 || s1 <= (uint32_t)g.minTh) s1 = 0;
        S2 = s0 | (s1 << 8);
      }
    }
    sS[sr * 32 + p] = (uint16_t)S2;
  }
  __syncthreads();

  // NMS inside the owning cell + emission
  const uint8_t* S8 = reinterpret_cast<const uint8_t*>(sS);
  PreCand* cand = b.cand + (long long)f * g.candSlab + L.candOff;
  int* candCount = b.candCount + f * ORBX_MAX_LEVELS + level;
  int* cellFlags = b.cellFlags + (long long)f * g.nCellsTotal + L.cellBase;
  for (int sr = 1 + warp; sr <= kTileH; sr += 8) {
    const uint32_t mine = sS[sr * 32 + p];
    const int y = oy - 1 + sr;
#pragma unroll
    for (int half = 0; half < 2; half++) {
      const int sc = 2 * p + half;
      const int s = (mine >> (8 * half)) & 0xff;
      bool keep = false;
      int x = ox - 1 + sc, cell = 0;
      if (s > 0 && sc >= 1 && sc <= kTileW) {
        const uint32_t ci = b.info[L.colInfoOff + x], ri = b.info[L.rowInfoOff + y];
        const bool fc = ci & 0x4000u, lc = ci & 0x8000u, fr = ri & 0x4000u, lr = ri & 0x8000u;
        const uint8_t* c = S8 + sr * 64 + sc;
        int m = 0;
        if (!fr) {
          if (!fc) m = max(m, (int)c[-65]);
          m = max(m, (int)c[-64]);
          if (!lc) m = max(m, (int)c[-63]);
        }
        if (!fc) m = max(m, (int)c[-1]);
        if (!lc) m = max(m, (int)c[1]);
        if (!lr) {
          if (!fc) m = max(m, (int)c[63]);
          m = max(m, (int)c[64]);
          if (!lc) m = max(m, (int)c[65]);
        }
        keep = s > m;
        cell = (int)(ri & 0x3fffu) * L.nCols + (int)(ci & 0x3fffu);
      }
      const uint32_t mask = __ballot_sync(ORBX_FULL, keep);
      if (mask) {
        int base = 0;
        const int leader = __ffs(mask) - 1;
        if (p == leader) base = atomicAdd(candCount, __popc(mask));
        base = __shfl_sync(ORBX_FULL, base, leader);
        if (keep) {
          const int idx = base + __popc(mask & ((1u << p) - 1));
          if (idx < L.candCap) {
            PreCand pc;
            pc.x = (uint16_t)x; pc.y = (uint16_t)y; pc.cell = (uint16_t)cell; pc.S = (uint8_t)s; pc.pad = 0;
            cand[idx] = pc;
          } else {
            atomicExch(b.status, ORBX_ERR_CAPACITY);
          }
          if (s > g.iniTh) atomicOr(cellFlags + cell, 1);
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Quadtree distribution: one CTA per (level, frame).  The reference's std::list algorithm is restated as data-parallel
// "walks": per-node child histograms (shared-memory atomics), exclusive scans for the new list positions (children are
// pushed to the FRONT in creation order => position = C-1-creationIndex; untouched nodes keep their order behind them),
// and for the final partial sweep a rank by (size desc, creation desc) with a prefix search for the break point
// (tie rule F5 = creation order).  Points keep only their node id; nothing is sorted or moved.
// ---------------------------------------------------------------------------------------------------------------
struct QBox { int16_t x0, y0, x1, y1; };

__device__ __forceinline__ int qt_child(const QBox bx, int x, int y) {
  const int mx = bx.x0 + ((bx.x1 - bx.x0 + 1) >> 1);  // UL.x + ceil((UR.x-UL.x)/2)
  const int my = bx.y0 + ((bx.y1 - bx.y0 + 1) >> 1);
  return (x < mx) ? ((y < my) ? 0 : 2) : ((y < my) ? 1 : 3);
}
__device__ __forceinline__ QBox qt_child_box(const QBox bx, int c) {
  const int mx = bx.x0 + ((bx.x1 - bx.x0 + 1) >> 1);
  const int my = bx.y0 + ((bx.y1 - bx.y0 + 1) >> 1);
  QBox r;
  r.x0 = (c & 1) ? mx : bx.x0; r.x1 = (c & 1) ? bx.x1 : mx;
  r.y0 = (c & 2) ? my : bx.y0; r.y1 = (c & 2) ? bx.y1 : my;
  return r;
}

// exclusive scan of data[0..n) in shared memory by the whole CTA (256 threads); returns the total.  Ends with a barrier.
__device__ int block_exscan(int* data, int n, int* warpTmp) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int chunk = (n + 255) >> 8;
  const int beg = min(n, tid * chunk), end = min(n, beg + chunk);
  int sum = 0;
  for (int i = beg; i < end; i++) sum += data[i];
  int incl = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(ORBX_FULL, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) warpTmp[warp] = incl;
  __syncthreads();
  int woff = 0, total = 0;
#pragma unroll
  for (int w = 0; w < 8; w++) {
    const int v = warpTmp[w];
    if (w < warp) woff += v;
    total += v;
  }
  int run = woff + incl - sum;
  for (int i = beg; i < end; i++) { const int v = data[i]; data[i] = run; run += v; }
  __syncthreads();
  return total;
}

__global__ void __launch_bounds__(256) k_quadtree(const __grid_constant__ Geom g, const __grid_constant__ Batch b, int* nCandOut) {
  extern __shared__ int qsm[];
  const int level = blockIdx.x, f = blockIdx.y, tid = threadIdx.x;
  const LevelGeom& L = g.lv[level];
  const int cap = L.nodeCap;
  // shared layout (ints): box[2][cap] (2 ints each), cnt[2][cap], cc[cap][4], cpos[cap][4], kk[cap], rr[cap], rk[cap], br[cap]
  QBox* box0 = reinterpret_cast<QBox*>(qsm);
  QBox* box1 = box0 + cap;
  int* cnt0 = reinterpret_cast<int*>(box1 + cap);
  int* cnt1 = cnt0 + cap;
  int* cc = cnt1 + cap;
  int* cpos = cc + 4 * cap;
  int* kk = cpos + 4 * cap;
  int* rr = kk + cap;
  int* rk = rr + cap;
  int* br = rk + cap;
  __shared__ int sh_n, sh_s, sh_E, sh_mode, sh_C, sh_rstar, sh_nc, sh_rootCnt[16], sh_rootPos[16], warpTmp[8];

  const long long slab = (long long)f * g.candSlab + L.candOff;
  uint32_t* qXY = b.qXY + slab;
  uint32_t* qOrd = b.qOrd + slab;
  uint16_t* qNode = b.qNode + slab;
  uint8_t* qResp = b.qResp + slab;
  uint8_t* qSel = b.qSel + slab;
  const PreCand* cand = b.cand + slab;
  const int* cellFlags = b.cellFlags + (long long)f * g.nCellsTotal + L.cellBase;
  const int npre = min(b.candCount[f * ORBX_MAX_LEVELS + level], L.candCap);
  const int N = L.quota;
  const int nIni = L.nIni;

  if (tid == 0) { sh_n = 0; sh_E = 0; }
  if (tid < 16) sh_rootCnt[tid] = 0;
  __syncthreads();
  // load + per-cell threshold choice (src/ORBextractor.cc:808-828) + root assignment (:563-568)
  for (int i = tid; i < npre; i += 256) {
    const PreCand pc = cand[i];
    const bool keep = (pc.S > g.iniTh) || !(cellFlags[pc.cell] & 1);
    if (keep) {
      const int idx = atomicAdd(&sh_n, 1);
      const int x = pc.x - kMinBorder, y = pc.y - kMinBorder;
      const int cellCol = pc.cell % L.nCols, cellRow = pc.cell / L.nCols;
      const int lx = pc.x - kEdge - cellCol * L.wCell, ly = pc.y - kEdge - cellRow * L.hCell;
      int root = (int)__fdiv_rn((float)x, L.hX);
      root = min(root, nIni - 1);
      qXY[idx] = (uint32_t)x | ((uint32_t)y << 16);
      qResp[idx] = (uint8_t)(pc.S - 1);
      qOrd[idx] = ((uint32_t)pc.cell << 12) | ((uint32_t)ly << 6) | (uint32_t)lx;
      qNode[idx] = (uint16_t)root;
      atomicAdd(&sh_rootCnt[root], 1);
    }
  }
  __syncthreads();
  const int n = sh_n;
  if (tid == 0) {
    if (nCandOut) nCandOut[f * ORBX_MAX_LEVELS + level] = n;
    const int H = L.h - 2 * kMinBorder;
    int s = 0;
    for (int i = 0; i < nIni; i++) {
      if (sh_rootCnt[i] > 0) {
        QBox bx;
        bx.x0 = (int16_t)(int)__fmul_rn(L.hX, (float)i);
        bx.x1 = (int16_t)(int)__fmul_rn(L.hX, (float)(i + 1));
        bx.y0 = 0; bx.y1 = (int16_t)H;
        box0[s] = bx; cnt0[s] = sh_rootCnt[i];
        sh_rootPos[i] = s++;
      } else sh_rootPos[i] = 0;
    }
    sh_s = s; sh_mode = 1; sh_C = 0;
  }
  __syncthreads();
  for (int i = tid; i < n; i += 256) qNode[i] = (uint16_t)sh_rootPos[qNode[i]];
  __syncthreads();

  int cur = 0;
  // mode 1 = full walk (phase 1), 2 = partial sweep (phase 2), 0 = finished
  while (sh_mode != 0) {
    const int mode = sh_mode, s = sh_s, Cprev = sh_C;
    QBox* bxC = cur ? box1 : box0; QBox* bxN = cur ? box0 : box1;
    int* cnC = cur ? cnt1 : cnt0;  int* cnN = cur ? cnt0 : cnt1;
    const int nScan = (mode == 1) ? s : Cprev;  // nodes that may be divided in this pass
    for (int i = tid; i < 4 * nScan; i += 256) cc[i] = 0;
    if (tid == 0) { sh_E = 0; sh_rstar = 0x7fffffff; sh_nc = 0; }
    __syncthreads();
    for (int i = tid; i < n; i += 256) {
      const int j = qNode[i];
      if (j < nScan && cnC[j] > 1) {
        const uint32_t xy = qXY[i];
        const int c = qt_child(bxC[j], (int)(xy & 0xffffu), (int)(xy >> 16));
        qSel[i] = (uint8_t)c;
        atomicAdd(&cc[4 * j + c], 1);
      }
    }
    __syncthreads();

    if (mode == 1) {
      int eLocal = 0;
      for (int j = tid; j < s; j += 256) {
        if (cnC[j] > 1) {
          int k = 0;
#pragma unroll
          for (int c = 0; c < 4; c++) { k += cc[4 * j + c] > 0; eLocal += cc[4 * j + c] > 1; }
          kk[j] = k; rr[j] = 0;
        } else { kk[j] = 0; rr[j] = 1; }
      }
      if (eLocal) atomicAdd(&sh_E, eLocal);
      __syncthreads();
      const int C = block_exscan(kk, s, warpTmp);
      const int R = block_exscan(rr, s, warpTmp);
      for (int j = tid; j < s; j += 256) {
        if (cnC[j] > 1) {
          int q = kk[j];
          const QBox pb = bxC[j];
#pragma unroll
          for (int c = 0; c < 4; c++) {
            const int m = cc[4 * j + c];
            if (m > 0) {
              const int pos = C - 1 - q; q++;
              bxN[pos] = qt_child_box(pb, c); cnN[pos] = m; cpos[4 * j + c] = pos;
            }
          }
        } else {
          const int pos = C + rr[j];
          bxN[pos] = bxC[j]; cnN[pos] = cnC[j]; cpos[4 * j] = pos;
        }
      }
      __syncthreads();
      for (int i = tid; i < n; i += 256) {
        const int j = qNode[i];
        qNode[i] = (uint16_t)((cnC[j] > 1) ? cpos[4 * j + qSel[i]] : cpos[4 * j]);
      }
      __syncthreads();
      if (tid == 0) {
        const int sNew = C + R;
        sh_s = sNew; sh_C = C;
        if (sNew >= N || sNew == s) sh_mode = 0;             // src/ORBextractor.cc:667
        else if (sNew + 3 * sh_E > N) sh_mode = 2;            // :671
        else sh_mode = 1;
      }
      cur ^= 1;
      __syncthreads();
    } else {
      // candidates = nodes created by the previous pass (front segment) holding more than one point
      for (int j = tid; j < Cprev; j += 256) {
        int k = 0;
        if (cnC[j] > 1) {
#pragma unroll
          for (int c = 0; c < 4; c++) k += cc[4 * j + c] > 0;
        }
        kk[j] = k;  // 0 for non-candidates
      }
      __syncthreads();
      // processing order: size descending, then creation descending == list position ascending
      for (int j = tid; j < Cprev; j += 256) {
        const int mj = cnC[j];
        if (mj > 1) {
          int r = 0;
          for (int i = 0; i < Cprev; i++) {
            const int mi = cnC[i];
            r += (mi > 1) && ((mi > mj) || (mi == mj && i < j));
          }
          rk[j] = r; br[r] = j;
          atomicAdd(&sh_nc, 1);
        } else rk[j] = -1;
      }
      __syncthreads();
      const int nc = sh_nc;
      for (int r = tid; r < nc; r += 256) rr[r] = kk[br[r]];  // children per candidate, in processing order
      __syncthreads();
      block_exscan(rr, nc, warpTmp);                           // rr[r] = creation index of the first child
      for (int r = tid; r < nc; r += 256) {
        const int incl = rr[r] + kk[br[r]] - (r + 1);          // list growth after processing ranks 0..r
        if (s + incl >= N) atomicMin(&sh_rstar, r);            // src/ORBextractor.cc:726-727 break
      }
      __syncthreads();
      const int nproc = (sh_rstar == 0x7fffffff) ? nc : sh_rstar + 1;
      const int Cn = nproc > 0 ? rr[nproc - 1] + kk[br[nproc - 1]] : 0;
      // survivors: every current node that is not processed keeps its relative order behind the new children
      for (int j = tid; j < s; j += 256) {
        const bool processed = (j < Cprev) && (rk[j] >= 0) && (rk[j] < nproc);
        cpos[4 * j + 1] = processed ? 0 : 1;  // temp flag; scanned below through kk2 = cc reuse
      }
      __syncthreads();
      // scan survivor flags (reuse kk for the flags; k values of processed candidates are still needed -> stash in cc[.][0..3] sums)
      for (int j = tid; j < s; j += 256) {
        const int flag = cpos[4 * j + 1];
        cpos[4 * j + 2] = (j < Cprev) ? kk[j] : 0;  // keep k_j
        kk[j] = flag;
      }
      __syncthreads();
      block_exscan(kk, s, warpTmp);
      for (int j = tid; j < s; j += 256) {
        const bool processed = cpos[4 * j + 1] == 0;
        if (processed) {
          int q = rr[rk[j]];
          const QBox pb = bxC[j];
          int posc[4];
#pragma unroll
          for (int c = 0; c < 4; c++) {
            const int m = cc[4 * j + c];
            posc[c] = 0;
            if (m > 0) {
              const int pos = Cn - 1 - q; q++;
              bxN[pos] = qt_child_box(pb, c); cnN[pos] = m; posc[c] = pos;
            }
          }
#pragma unroll
          for (int c = 0; c < 4; c++) cc[4 * j + c] = posc[c];   // cc now holds child positions for processed nodes
        } else {
          const int pos = Cn + kk[j];
          bxN[pos] = bxC[j]; cnN[pos] = cnC[j];
          cc[4 * j] = pos;
        }
      }
      __syncthreads();
      for (int i = tid; i < n; i += 256) {
        const int j = qNode[i];
        const bool processed = cpos[4 * j + 1] == 0;
        qNode[i] = (uint16_t)(processed ? cc[4 * j + qSel[i]] : cc[4 * j]);
      }
      __syncthreads();
      if (tid == 0) {
        const int sNew = Cn + (s - nproc);
        sh_s = sNew; sh_C = Cn;
        if (sNew >= N || sNew == s) sh_mode = 0;  // src/ORBextractor.cc:730-731
        else sh_mode = 2;
      }
      cur ^= 1;
      __syncthreads();
    }
  }

  // keep the best response per node, first in candidate order on ties (src/ORBextractor.cc:739-758)
  const int s = sh_s;
  int* best = cc;
  unsigned* bestOrd = reinterpret_cast<unsigned*>(cpos);
  for (int j = tid; j < s; j += 256) { best[j] = -1; bestOrd[j] = 0xffffffffu; }
  __syncthreads();
  for (int i = tid; i < n; i += 256) atomicMax(&best[qNode[i]], (int)qResp[i]);
  __syncthreads();
  for (int i = tid; i < n; i += 256)
    if ((int)qResp[i] == best[qNode[i]]) atomicMin(&bestOrd[qNode[i]], qOrd[i]);
  __syncthreads();
  LevelKp* out = b.lkp + (long long)f * g.kpSlab + L.kpOff;
  for (int i = tid; i < n; i += 256) {
    const int j = qNode[i];
    if ((int)qResp[i] == best[j] && qOrd[i] == bestOrd[j]) {
      if (j < L.kpCap) {
        const uint32_t xy = qXY[i];
        LevelKp kp;
        kp.x = (int16_t)((xy & 0xffffu) + kMinBorder);
        kp.y = (int16_t)((xy >> 16) + kMinBorder);
        kp.resp = qResp[i];
        out[j] = kp;
      }
    }
  }
  if (tid == 0) {
    if (s > L.kpCap) atomicExch(b.status, ORBX_ERR_CAPACITY);
    b.lkpCount[f * ORBX_MAX_LEVELS + level] = min(s, L.kpCap);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// 7x7 Gaussian blur, OpenCV's 8-bit fixed-point path (Q8.8 horizontal into u16, Q16.16 vertical, +0x8000 >> 16),
// BORDER_REFLECT_101 taken from the level itself.  CTA = 128x16 output tile; horizontal pass with
// dp4a over byte quads from shared memory, vertical pass with 32-bit IMADs; 32-bit stores.
// ---------------------------------------------------------------------------------------------------------------
constexpr int kBlurW = 128, kBlurH = 16;
constexpr int kBlurRawW = kBlurW + 8;  // bytes per staged row: x0-4 .. x0+131

__device__ __forceinline__ int refl101(int i, int n) {
  if (i < 0) i = -i;
  if (i >= n) i = 2 * (n - 1) - i;
  return max(0, min(n - 1, i));  // (only reached twice-reflected on degenerate tiny levels)
}

__global__ void __launch_bounds__(256) k_blur7(const __grid_constant__ Geom g, const __grid_constant__ Batch b) {
  __shared__ uint32_t sRaw[(kBlurH + 6) * (kBlurRawW / 4)];
  __shared__ uint32_t sHz[(kBlurH + 6) * (kBlurW / 2)];  // u16 pairs
  const int tid = threadIdx.x, f = blockIdx.y;
  int level = 0;
  while (level + 1 < g.nlevels && (int)blockIdx.x >= g.lv[level + 1].blurTileBase) level++;
  const LevelGeom& L = g.lv[level];
  const int t = blockIdx.x - L.blurTileBase;
  const int tx = t % L.blurTilesX, ty = t / L.blurTilesX;
  const int x0 = tx * kBlurW, y0 = ty * kBlurH;
  int sp;
  const uint8_t* src = level_ptr(g, b, level, f, sp);
  const bool interior = (x0 >= 4) && (x0 + kBlurW + 4 <= L.w) && ((sp & 3) == 0);
  if (interior) {
    for (int i = tid; i < (kBlurH + 6) * (kBlurRawW / 4); i += 256) {
      const int r = i / (kBlurRawW / 4), k = i - r * (kBlurRawW / 4);
      const int gy = refl101(y0 - 3 + r, L.h);
      sRaw[i] = __ldg(reinterpret_cast<const uint32_t*>(src + (long long)gy * sp + x0 - 4 + 4 * k));
    }
  } else {
    uint8_t* rawb = reinterpret_cast<uint8_t*>(sRaw);
    for (int i = tid; i < (kBlurH + 6) * kBlurRawW; i += 256) {
      const int r = i / kBlurRawW, c = i - r * kBlurRawW;
      const int gy = refl101(y0 - 3 + r, L.h), gx = refl101(x0 - 4 + c, L.w);
      rawb[i] = __ldg(src + (long long)gy * sp + gx);
    }
  }
  __syncthreads();
  const uint32_t tapLo = (uint32_t)g.taps[0] | ((uint32_t)g.taps[1] << 8) | ((uint32_t)g.taps[2] << 16) | ((uint32_t)g.taps[3] << 24);
  const uint32_t tapHi = (uint32_t)g.taps[4] | ((uint32_t)g.taps[5] << 8) | ((uint32_t)g.taps[6] << 16);
  // horizontal: task = 4 consecutive pixels of one staged row
  for (int i = tid; i < (kBlurH + 6) * (kBlurW / 4); i += 256) {
    const int r = i / (kBlurW / 4), q = i - r * (kBlurW / 4);
    const uint32_t* row = sRaw + r * (kBlurRawW / 4) + q;   // bytes x-4 .. x+7 of pixel quad x = x0+4q
    const uint32_t w0 = row[0], w1 = row[1], w2 = row[2];
    uint32_t hz[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
      // taps cover bytes (k+1 .. k+7) of the 12 staged bytes
      const uint32_t lo = (k == 3) ? w1 : __funnelshift_r(w0, w1, 8 * (k + 1));
      const uint32_t hi = (k == 3) ? w2 : __funnelshift_r(w1, w2, 8 * (k + 1));
      uint32_t acc = __dp4a(lo, tapLo, 0u);
      acc = __dp4a(hi, tapHi, acc);
      hz[k] = min(acc, 65535u);
    }
    sHz[r * (kBlurW / 2) + 2 * q] = hz[0] | (hz[1] << 16);
    sHz[r * (kBlurW / 2) + 2 * q + 1] = hz[2] | (hz[3] << 16);
  }
  __syncthreads();
  uint8_t* dst = b.blur + (long long)f * b.blurFrameStride + L.blurOff;
  for (int i = tid; i < kBlurH * (kBlurW / 4); i += 256) {
    const int r = i / (kBlurW / 4), q = i - r * (kBlurW / 4);
    const int x = x0 + 4 * q, y = y0 + r;
    if (x >= L.w || y >= L.h) continue;
    uint32_t a0 = 0, a1 = 0, a2 = 0, a3 = 0;
#pragma unroll
    for (int k = 0; k < 7; k++) {
      const uint32_t p0 = sHz[(r + k) * (kBlurW / 2) + 2 * q], p1 = sHz[(r + k) * (kBlurW / 2) + 2 * q + 1];
      const uint32_t tk = (uint32_t)g.taps[k];
      a0 += tk * (p0 & 0xffffu); a1 += tk * (p0 >> 16);
      a2 += tk * (p1 & 0xffffu); a3 += tk * (p1 >> 16);
    }
    const uint32_t o0 = min(255u, (a0 + 32768u) >> 16), o1 = min(255u, (a1 + 32768u) >> 16);
    const uint32_t o2 = min(255u, (a2 + 32768u) >> 16), o3 = min(255u, (a3 + 32768u) >> 16);
    *reinterpret_cast<uint32_t*>(dst + (long long)y * L.blurPitch + x) = o0 | (o1 << 8) | (o2 << 16) | (o3 << 24);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Output ordering of operator(): keypoints whose scaled x lies in [lap0, lap1] are written from the back, the others
// from the front, level-major (src/ORBextractor.cc:1118, 1164-1178).  One CTA per frame computes every level
// keypoint's final slot with two running exclusive scans.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_assemble(const __grid_constant__ Geom g, const __grid_constant__ Batch b) {
  __shared__ int warpTmp[8];
  __shared__ int flags[256];
  const int f = blockIdx.x, tid = threadIdx.x;
  int total = 0;
  for (int l = 0; l < g.nlevels; l++) total += b.lkpCount[f * ORBX_MAX_LEVELS + l];
  int monoRun = 0, stereoRun = 0;
  const float lap0 = (float)b.lap0, lap1 = (float)b.lap1;
  for (int l = 0; l < g.nlevels; l++) {
    const LevelGeom& L = g.lv[l];
    const int cnt = b.lkpCount[f * ORBX_MAX_LEVELS + l];
    const LevelKp* kp = b.lkp + (long long)f * g.kpSlab + L.kpOff;
    int* dest = b.lkpDest + (long long)f * g.kpSlab + L.kpOff;
    for (int base = 0; base < cnt; base += 256) {
      const int i = base + tid;
      int lap = 0;
      if (i < cnt) {
        const float xs = (l == 0) ? (float)kp[i].x : __fmul_rn((float)kp[i].x, L.scale);
        lap = (xs >= lap0 && xs <= lap1) ? 1 : 0;
      }
      flags[tid] = lap;
      __syncthreads();
      const int m = min(256, cnt - base);
      const int lapsHere = block_exscan(flags, 256, warpTmp);
      if (i < cnt) {
        const int lapBefore = flags[tid];
        dest[i] = lap ? (total - 1 - (stereoRun + lapBefore)) : (monoRun + (tid - lapBefore));
      }
      monoRun += m - lapsHere;
      stereoRun += lapsHere;
      __syncthreads();
    }
  }
  if (tid == 0) { b.counts[f] = total; b.mono[f] = monoRun; }
}

// ---------------------------------------------------------------------------------------------------------------
// Orientation (intensity centroid on the un-blurred level, cv::fastAtan2 in strict fp32) + steered BRIEF on the blurred
// level.  One warp per level keypoint: lanes = patch columns for the moments, lanes = descriptor bytes for rBRIEF.
// Float parity: no FMA contraction anywhere (__fmul_rn / __fadd_rn / __fdiv_rn), round-half-even (__float2int_rn),
// sin/cos = (float) of the double-precision value ("portable" mode).
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
  const float scale = (float)(180.0 / 3.1415926535897932384626433832795);
  const float p1 = __fmul_rn(0.9997878412794807f, scale), p3 = __fmul_rn(-0.3258083974640975f, scale);
  const float p5 = __fmul_rn(0.1555786518463281f, scale), p7 = __fmul_rn(-0.04432655554792128f, scale);
  const float ax = fabsf(x), ay = fabsf(y);
  float a, c, c2;
  if (ax >= ay) {
    c = __fdiv_rn(ay, __fadd_rn(ax, (float)DBL_EPSILON));
    c2 = __fmul_rn(c, c);
    a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
  } else {
    c = __fdiv_rn(ax, __fadd_rn(ay, (float)DBL_EPSILON));
    c2 = __fmul_rn(c, c);
    a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
  }
  if (x < 0) a = __fsub_rn(180.f, a);
  if (y < 0) a = __fsub_rn(360.f, a);
  return a;
}

__global__ void __launch_bounds__(256) k_orient_desc(const __grid_constant__ Geom g, const __grid_constant__ Batch b) {
  const int f = blockIdx.y, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int slot = blockIdx.x * 8 + warp;
  if (slot >= g.kpSlab) return;
  int level = 0;
  while (level + 1 < g.nlevels && slot >= g.lv[level + 1].kpOff) level++;
  const LevelGeom& L = g.lv[level];
  const int i = slot - L.kpOff;
  if (i >= b.lkpCount[f * ORBX_MAX_LEVELS + level]) return;
  const LevelKp kp = b.lkp[(long long)f * g.kpSlab + slot];
  const int dest = b.lkpDest[(long long)f * g.kpSlab + slot];
  int sp;
  const uint8_t* img = level_ptr(g, b, level, f, sp);
  // IC_Angle: m10 = sum u*I, m01 = sum v*I over the radius-15 disc
  int m10 = 0, m01 = 0;
  {
    const int u = lane - kHalfPatch;
    const int au = abs(u);
    const uint8_t* c = img + (long long)kp.y * sp + kp.x + u;
    if (lane < 31) {
#pragma unroll 1
      for (int v = -kHalfPatch; v <= kHalfPatch; v++) {
        if (au <= g.umax[abs(v)]) {
          const int val = __ldg(c + (long long)v * sp);
          m10 += u * val;
          m01 += v * val;
        }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      m10 += __shfl_xor_sync(ORBX_FULL, m10, o);
      m01 += __shfl_xor_sync(ORBX_FULL, m01, o);
    }
  }
  const float angle = fast_atan2_deg((float)m01, (float)m10);
  // rBRIEF
  const float factorPI = (float)(3.1415926535897932384626433832795 / 180.0);
  const float ang = __fmul_rn(angle, factorPI);
  const float ca = (float)cos((double)ang), sa = (float)sin((double)ang);
  const uint8_t* bl = b.blur + (long long)f * b.blurFrameStride + L.blurOff + (long long)kp.y * L.blurPitch + kp.x;
  const int bp = L.blurPitch;
  const int4* pat4 = reinterpret_cast<const int4*>(b.pattern) + lane * 2;
  const int4 pa = __ldg(pat4), pb = __ldg(pat4 + 1);
  const int pw[8] = {pa.x, pa.y, pa.z, pa.w, pb.x, pb.y, pb.z, pb.w};
  int val = 0;
#pragma unroll
  for (int j = 0; j < 8; j++) {
    const float x0 = (float)(int8_t)(pw[j] & 0xff), y0 = (float)(int8_t)((pw[j] >> 8) & 0xff);
    const float x1 = (float)(int8_t)((pw[j] >> 16) & 0xff), y1 = (float)(int8_t)((pw[j] >> 24) & 0xff);
    const int rr0 = __float2int_rn(__fadd_rn(__fmul_rn(x0, sa), __fmul_rn(y0, ca)));
    const int cc0 = __float2int_rn(__fsub_rn(__fmul_rn(x0, ca), __fmul_rn(y0, sa)));
    const int rr1 = __float2int_rn(__fadd_rn(__fmul_rn(x1, sa), __fmul_rn(y1, ca)));
    const int cc1 = __float2int_rn(__fsub_rn(__fmul_rn(x1, ca), __fmul_rn(y1, sa)));
    const int t0 = __ldg(bl + rr0 * bp + cc0), t1 = __ldg(bl + rr1 * bp + cc1);
    val |= (t0 < t1) << j;
  }
  b.desc[((long long)f * g.kpStride + dest) * 32 + lane] = (uint8_t)val;
  if (lane == 0) {
    orbx_keypoint o;
    o.x = (level == 0) ? (float)kp.x : __fmul_rn((float)kp.x, L.scale);
    o.y = (level == 0) ? (float)kp.y : __fmul_rn((float)kp.y, L.scale);
    o.size = L.kpSize;
    o.angle = angle;
    o.response = (float)kp.resp;
    o.octave = level;
    o.class_id = -1;
    b.kps[(long long)f * g.kpStride + dest] = o;
  }
}

}  // namespace orbx
