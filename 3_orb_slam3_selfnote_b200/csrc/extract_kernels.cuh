// sm_100a kernels of the ORB extractor.  One launch covers a whole batch of frames (blockIdx.y / .z = frame); every
// kernel is integer / byte work bound by issue slots and HBM/L2, so: coalesced 32-bit accesses, shared-memory staging,
// u16x2 SIMD-in-register min/max (VIMNMX3.U16x2) for the FAST arcs, dp4a for the blur taps, warp shuffles for reductions.
// No tensor cores: nothing here is a dense floating-point contraction.
//
// Reference stages (wangyunpeng139/3_ORB_SLAM3_SelfNote) and the file that holds each kernel:
//   k_pyr_resize4  ORBextractor::ComputePyramid            src/ORBextractor.cc:1186-1219  (cv::resize INTER_LINEAR 8U)   pyr_resize.cuh
//   k_fast_nms3    ComputeKeyPointsOctTree cell loop        src/ORBextractor.cc:787-854   (cv::FAST 9/16 + NMS per cell)  fast_nms3.cuh
//   k_quadtree     DistributeOctTree / DivideNode           src/ORBextractor.cc:479-761                                   this file
//   k_blur7v       GaussianBlur(7x7, 2, REFLECT_101)        src/ORBextractor.cc:1132-1141                                 blur7.cuh
//   k_assemble     operator() output ordering               src/ORBextractor.cc:1118,1164-1178                            this file
//   k_orient_desc2 IC_Angle + computeOrbDescriptor          src/ORBextractor.cc:75-145, 470-477, 1059-1066                orient_desc.cuh
#pragma once
#include <cuda_runtime.h>
#include <float.h>

#include "dev_common.cuh"
#include "geom.h"

namespace orbx {

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

}  // namespace orbx
