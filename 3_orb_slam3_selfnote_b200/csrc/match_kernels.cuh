// sm_100a kernels of the Hamming searches (ORBmatcher / Frame side of the frontend).
//   k_hamming_pairs      ORBmatcher::DescriptorDistance                      src/ORBmatcher.cc:2463-2483
//   k_build_grid         Frame::AssignFeaturesToGrid / PosInGrid             src/Frame.cc:434-465, 815-825
//   k_window_candidates  Frame::GetFeaturesInArea + the distance loop        src/Frame.cc:744-813, src/ORBmatcher.cc:85-121
//   k_resolve_*          the sequential claim / ratio / rotation logic       src/ORBmatcher.cc:44-214, 722-837, 2027-2413
//   k_stereo_*           Frame::ComputeStereoMatches                         src/Frame.cc:901-1079
//   k_knn2               cv::BFMatcher(NORM_HAMMING).knnMatch(k=2)           src/Frame.cc:1246
// Distances are 8 x (xor + popc) on 32-bit words, candidates of one query are spread over the lanes of a warp and reduced
// with shuffles on packed (distance, traversal order, index) keys so that "first minimum wins" ties resolve exactly as the
// reference's sequential loops do.  The searches whose result depends on earlier queries (keypoints
// claimed by earlier map points) are split in two: a fully parallel candidate/distance pass, then a parallel fixed-point
// iteration over the claims that reaches the sequential result (k_resolve_par, k_bow_resolve_par); SearchForInitialization,
// whose claims are not monotone (matches can be stolen back), replays the reference's order with one warp (k_resolve_init).
#pragma once
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <limits.h>

#include "dev_common.cuh"
#include "geom.h"

namespace orbx {

constexpr int kGridCols = 64, kGridRows = 48, kGridCells = kGridCols * kGridRows;  // include/Frame.h:38-39
constexpr unsigned long long kKeyNone = 0xffffffffffffffffull;

__device__ __forceinline__ int hamming8(const uint32_t* a, const uint32_t* b) {
  int d = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) d += __popc(a[i] ^ b[i]);
  return d;
}
__device__ __forceinline__ void load_desc(const uint8_t* p, uint32_t* out) {
  const uint4 lo = __ldg(reinterpret_cast<const uint4*>(p)), hi = __ldg(reinterpret_cast<const uint4*>(p) + 1);
  out[0] = lo.x; out[1] = lo.y; out[2] = lo.z; out[3] = lo.w; out[4] = hi.x; out[5] = hi.y; out[6] = hi.z; out[7] = hi.w;
}

__global__ void k_hamming_pairs(const uint8_t* a, const uint8_t* b, int n, int* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t x[8], y[8];
  load_desc(a + 32ll * i, x);
  load_desc(b + 32ll * i, y);
  out[i] = hamming8(x, y);
}

// ---- frame grid ---------------------------------------------------------------------------------------------------
struct GridFrame {
  const orbx_keypoint* kps;   // mvKeysUn
  const uint8_t* desc;        // mDescriptors
  const float* uRight;        // mvuRight or null
  int n;
  float minX, minY, invW, invH;
  int* cellStart;             // [kGridCells + 1]
  int* cellItems;             // [n] keypoint indices grouped by cell (cell = posX*48 + posY), arbitrary order inside a cell
};

__device__ int mblock_exscan(int* data, int n, int* warpTmp) {  // same contract as block_exscan (256 threads)
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int chunk = (n + 255) >> 8;
  const int beg = min(n, tid * chunk), end = min(n, beg + chunk);
  int sum = 0;
  for (int i = beg; i < end; i++) sum += data[i];
  int incl = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, o);
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

__device__ __forceinline__ int grid_cell_of(const GridFrame& F, float x, float y) {
  const int posX = (int)roundf(__fmul_rn(__fsub_rn(x, F.minX), F.invW));
  const int posY = (int)roundf(__fmul_rn(__fsub_rn(y, F.minY), F.invH));
  if (posX < 0 || posX >= kGridCols || posY < 0 || posY >= kGridRows) return -1;
  return posX * kGridRows + posY;
}

__global__ void __launch_bounds__(256) k_build_grid(const GridFrame* frames) {
  __shared__ int cnt[kGridCells];
  __shared__ int warpTmp[8];
  const GridFrame F = frames[blockIdx.x];
  const int tid = threadIdx.x;
  for (int i = tid; i < kGridCells; i += 256) cnt[i] = 0;
  __syncthreads();
  for (int i = tid; i < F.n; i += 256) {
    const int c = grid_cell_of(F, F.kps[i].x, F.kps[i].y);
    if (c >= 0) atomicAdd(&cnt[c], 1);
  }
  __syncthreads();
  const int total = mblock_exscan(cnt, kGridCells, warpTmp);
  for (int i = tid; i < kGridCells; i += 256) F.cellStart[i] = cnt[i];
  if (tid == 0) F.cellStart[kGridCells] = total;
  __syncthreads();
  for (int i = tid; i < F.n; i += 256) {
    const int c = grid_cell_of(F, F.kps[i].x, F.kps[i].y);
    if (c >= 0) F.cellItems[atomicAdd(&cnt[c], 1)] = i;
  }
}

// ---- window candidates ----------------------------------------------------------------------------------------------
struct QuerySet {            // one per frame
  int nq;
  const uint8_t* valid;      // query is searched at all
  const float* x; const float* y; const float* r;   // window centre / half size
  const int8_t* minL; const int8_t* maxL;            // GetFeaturesInArea level arguments
  const float* xr;           // right-image coordinate of the query (used when F.uRight != null and xrCheck)
  const uint8_t* desc;       // 32-byte descriptor per query
  int xrCheck;
  int gate;                  // Fuse chi-square gate (ORBmatcher.cc:1547-1575): uses xr as the projected right coordinate
  const float* invSigma2;    // mvInvLevelSigma2 (gate only)
  int* qOff; int* qCnt;      // per query: first entry / number of entries
  unsigned long long* entries; int entryCap; int* entryCursor;
};

// key layout: dist[52:44] | cell order[43:32] | keypoint index[31:8] | octave[7:0]
__device__ __forceinline__ unsigned long long make_key(int dist, int cell, int idx, int octave) {
  return ((unsigned long long)dist << 44) | ((unsigned long long)cell << 32) | ((unsigned long long)idx << 8) | (unsigned)(octave & 0xff);
}

__global__ void __launch_bounds__(256) k_window_candidates(const GridFrame* frames, const QuerySet* sets, int* status) {
  const GridFrame F = frames[blockIdx.y];
  const QuerySet Q = sets[blockIdx.y];
  const int lane = threadIdx.x & 31;
  const int i = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (i >= Q.nq) return;
  int nMinCellX = 0, nMaxCellX = -1, nMinCellY = 0, nMaxCellY = -1;
  const float x = Q.x[i], y = Q.y[i], r = Q.r[i];
  const int minLevel = Q.minL[i], maxLevel = Q.maxL[i];
  bool live = Q.valid[i] != 0;
  if (live) {  // Frame::GetFeaturesInArea, Frame.cc:755-777
    nMinCellX = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, F.minX), r), F.invW)));
    nMaxCellX = min(kGridCols - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, F.minX), r), F.invW)));
    nMinCellY = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, F.minY), r), F.invH)));
    nMaxCellY = min(kGridRows - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, F.minY), r), F.invH)));
    if (nMinCellX >= kGridCols || nMaxCellX < 0 || nMinCellY >= kGridRows || nMaxCellY < 0) live = false;
  }
  const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
  const float xr = (Q.xrCheck && Q.xr) ? Q.xr[i] : 0.f;
  const bool rightCheck = Q.xrCheck && F.uRight != nullptr;
  uint32_t qd[8];
  if (live) load_desc(Q.desc + 32ll * i, qd);
  int off = 0, total = 0;
  for (int pass = 0; pass < 2; pass++) {
    int written = 0;
    if (live) {
      for (int ix = nMinCellX; ix <= nMaxCellX; ix++) {
        const int c0 = ix * kGridRows + nMinCellY, c1 = ix * kGridRows + nMaxCellY;
        // cells of one grid column are consecutive in memory: walk their items as one range, tracking the cell id
        const int jb = F.cellStart[c0], je = F.cellStart[c1 + 1];
        for (int j0 = jb; j0 < je; j0 += 32) {
          const int j = j0 + lane;
          bool ok = false;
          int idx = 0, oct = 0, cell = 0;
          if (j < je) {
            idx = F.cellItems[j];
            const orbx_keypoint kp = F.kps[idx];
            oct = kp.octave;
            ok = true;
            if (bCheckLevels) {
              if (oct < minLevel) ok = false;
              if (maxLevel >= 0 && oct > maxLevel) ok = false;
            }
            const float distx = __fsub_rn(kp.x, x), disty = __fsub_rn(kp.y, y);
            if (!(fabsf(distx) < r && fabsf(disty) < r)) ok = false;
            if (ok && rightCheck) {
              const float ur = F.uRight[idx];
              if (ur > 0) {
                const float er = fabsf(__fsub_rn(xr, ur));
                if (er > r) ok = false;
              }
            }
            if (ok && Q.gate) {  // ORBmatcher.cc:1547-1575
              const float ex = __fsub_rn(x, kp.x), ey = __fsub_rn(y, kp.y);
              float e2 = __fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey));
              const float kpr = F.uRight ? F.uRight[idx] : -1.0f;
              double lim = 5.99;
              if (kpr >= 0) {
                const float er = __fsub_rn(Q.xr[i], kpr);
                e2 = __fadd_rn(e2, __fmul_rn(er, er));
                lim = 7.8;
              }
              if ((double)__fmul_rn(e2, Q.invSigma2[oct]) > lim) ok = false;
            }
            if (ok && pass == 1) {  // which cell of the column does item j belong to?
              cell = c0;
              while (F.cellStart[cell + 1] <= j) cell++;
            }
          }
          const unsigned m = __ballot_sync(0xffffffffu, ok);
          if (pass == 1 && ok) {
            uint32_t cd[8];
            load_desc(F.desc + 32ll * idx, cd);
            const int dist = hamming8(qd, cd);
            Q.entries[off + written + __popc(m & ((1u << lane) - 1))] = make_key(dist, cell, idx, oct);
          }
          written += __popc(m);
        }
      }
    }
    if (pass == 0) {
      total = written;
      if (lane == 0) {
        off = total > 0 ? atomicAdd(Q.entryCursor, total) : 0;
        if (off + total > Q.entryCap) { atomicExch(status, ORBX_ERR_CAPACITY); total = 0; }
        Q.qOff[i] = off; Q.qCnt[i] = total;
      }
      off = __shfl_sync(0xffffffffu, off, 0);
      total = __shfl_sync(0xffffffffu, total, 0);
      if (total == 0) break;
    }
  }
}

// ---- warp top-2 on keys ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void top2_insert(unsigned long long k, unsigned long long& k1, unsigned long long& k2) {
  const unsigned long long lo = min(k, k1), hi = max(k, k1);
  k1 = lo;
  k2 = min(k2, hi);
}
__device__ __forceinline__ void top2_warp(unsigned long long& k1, unsigned long long& k2) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long o1 = __shfl_xor_sync(0xffffffffu, k1, o), o2 = __shfl_xor_sync(0xffffffffu, k2, o);
    const unsigned long long lo = min(k1, o1), hi = max(k1, o1);
    k1 = lo;
    k2 = min(hi, min(k2, o2));
  }
}
__device__ __forceinline__ int key_dist(unsigned long long k) { return (int)(k >> 44); }
__device__ __forceinline__ int key_idx(unsigned long long k) { return (int)((k >> 8) & 0xffffffu); }
__device__ __forceinline__ int key_oct(unsigned long long k) { return (int)(k & 0xffu); }

constexpr int kMaxKpSmem = 16384;

// ORBmatcher::ComputeThreeMaxima, ORBmatcher.cc:2416-2458
__device__ void three_maxima(const int* hs, int L, int& ind1, int& ind2, int& ind3) {
  int max1 = 0, max2 = 0, max3 = 0;
  ind1 = ind2 = ind3 = -1;
  for (int i = 0; i < L; i++) {
    const int s = hs[i];
    if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
    else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
    else if (s > max3) { max3 = s; ind3 = i; }
  }
  if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
  else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) { ind3 = -1; }
}

__device__ __forceinline__ int rot_bin(float a1, float a2) {  // ORBmatcher.cc:2177-2183
  float rot = __fsub_rn(a1, a2);
  if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
  int bin = (int)roundf(__fmul_rn(rot, 1.0f / 30));
  if (bin == 30) bin = 0;
  return bin;
}

struct ResolveArgs {        // one per frame
  const uint8_t* observed;  // per query: MapPoint::Observations() > 0 (null = all observed)
  const float* angle;       // per query: source keypoint angle (rotation histogram)
  int* kpMp;                // in/out, F.mvpMapPoints as indices (-1 empty, -2 blocking, -3 non-blocking pre-existing)
  int* assigned;            // scratch [nq] (k_resolve_best: best index per query)
  int* bestDist;            // k_resolve_best only: best distance per query
  int* blk[2];              // parallel resolve: per keypoint, smallest query index whose (blocking) claim holds it ([n] each)
  int* chg;                 // parallel resolve: [2] "a choice changed" flags
  int* nmatches;            // out
  float nnratio;
  int distTh;
  int anyMpBlocks;
  int checkOrientation;
  const int* partner;       // k_resolve_fisheye: stereo partner of each keypoint in the other image (-1 none), or null
  const uint8_t* gateValid; // k_resolve_par: query i is searched only if the sibling job's query i was valid and had candidates
  const int* gateCnt;       //   (right-camera pass of SearchByProjection(Frame, Frame), ORBmatcher.cc:2077-2124 `continue`s); or null
};

// SearchByProjection(Frame&, vector<MapPoint*>&, ...), ORBmatcher.cc:44-142, as a parallel fixed point.  The reference's loop is sequential only through the claims: map
// point i cannot take a keypoint that an earlier map point j < i with observations already took.  Let choice[i] be the keypoint
// map point i ends up with; choice[i] is a function of {choice[j] : j < i} alone, so the sequential result is the unique fixed
// point of  choice <- F(choice), and iterating F from "no claims" reaches it after at most (longest dependency chain) + 1
// rounds -- a handful in practice, because most windows do not compete.  One round is fully parallel: warp = map point, the
// per-keypoint state is blk[k] = smallest index of an observed map point currently choosing k (atomicMin), and k is blocked for
// i iff blk[k] < i.  A cluster of 8 CTAs works on one frame (128 warps), two cluster barriers per round.  Afterwards
// kpMp[k] = the largest index that chose k (later map points overwrite earlier ones without observations, ORBmatcher.cc:89-91).
constexpr int kParCluster = 8, kParThreads = 512;
// mode 0: SearchByProjection(Frame&, vector<MapPoint*>&) (best / second with the level rule, TH_HIGH, ratio)
// mode 1: the best-1 window searches (best-1 rule: distTh, anyMpBlocks, rotation histogram afterwards)
__global__ void __cluster_dims__(kParCluster, 1, 1) __launch_bounds__(kParThreads)
k_resolve_par(const GridFrame* frames, const QuerySet* sets, const ResolveArgs* args, int mode) {
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  const GridFrame F = frames[blockIdx.y];
  const QuerySet Q = sets[blockIdx.y];
  const ResolveArgs A = args[blockIdx.y];
  const int lane = threadIdx.x & 31;
  const int gthread = blockIdx.x * kParThreads + threadIdx.x, nthreads = kParCluster * kParThreads;
  const int gwarp = gthread >> 5, nwarps = nthreads >> 5;
  const bool anyBlocks = mode == 1 && A.anyMpBlocks;
  auto obs = [&](int i) { return A.observed ? A.observed[i] != 0 : true; };
  auto pre = [&](int k) {   // keypoints that block everybody before the call
    const int s = A.kpMp[k];
    const bool bl = anyBlocks ? (s != -1) : ((s == -2) || (s >= 0 && obs(s)));
    return bl ? -1 : INT_MAX;
  };
  for (int k = gthread; k < F.n; k += nthreads) { A.blk[0][k] = pre(k); A.blk[1][k] = pre(k); }
  for (int i = gthread; i < Q.nq; i += nthreads) A.assigned[i] = -1;
  if (gthread < 2) A.chg[gthread] = 0;
  if (gthread == 0) *A.nmatches = 0;
  cluster.sync();
  for (int it = 0; it <= Q.nq; it++) {
    const int* cur = A.blk[it & 1];
    int* nxt = A.blk[(it & 1) ^ 1];
    bool changed = false;
    for (int i = gwarp; i < Q.nq; i += nwarps) {
      const int c = (A.gateValid && (!A.gateValid[i] || A.gateCnt[i] == 0)) ? 0 : Q.qCnt[i];
      int choice = -1;
      if (c > 0) {
        const unsigned long long* e = Q.entries + Q.qOff[i];
        unsigned long long k1 = kKeyNone, k2 = kKeyNone;
        for (int j = lane; j < c; j += 32) {
          const unsigned long long k = e[j];
          if (!(__ldcg(cur + key_idx(k)) < i)) top2_insert(k, k1, k2);
        }
        top2_warp(k1, k2);
        if (k1 != kKeyNone) {
          const int bestDist = key_dist(k1);
          if (mode == 0) {
            const int bestLevel = key_oct(k1);
            const int bestDist2 = (k2 == kKeyNone) ? 256 : key_dist(k2), bestLevel2 = (k2 == kKeyNone) ? -1 : key_oct(k2);
            if (bestDist <= 100) {  // TH_HIGH
              const float lim = __fmul_rn(A.nnratio, (float)bestDist2);
              if (!(bestLevel == bestLevel2 && (float)bestDist > lim)) choice = key_idx(k1);
            }
          } else if (bestDist <= A.distTh) choice = key_idx(k1);
        }
      }
      if (lane == 0) {
        if (choice != A.assigned[i]) { A.assigned[i] = choice; changed = true; }
        if (choice >= 0 && (anyBlocks || obs(i))) atomicMin(nxt + choice, i);
      }
    }
    if (changed) atomicExch(A.chg + (it & 1), 1);
    cluster.sync();
    const int any = __ldcg(A.chg + (it & 1));
    // nobody reads `cur` any more: it becomes the target of the next round
    for (int k = gthread; k < F.n; k += nthreads) A.blk[it & 1][k] = pre(k);
    if (gthread == 0) A.chg[(it & 1) ^ 1] = 0;
    cluster.sync();
    if (!any) break;
  }
  // the fixed point is reached.  blk[0] becomes "largest index that chose k", blk[1][0..29] the rotation histogram
  for (int k = gthread; k < F.n; k += nthreads) A.blk[0][k] = -1;
  if (gthread < 32) A.blk[1][gthread] = 0;
  cluster.sync();
  const bool ori = mode == 1 && A.checkOrientation;
  int mine = 0;
  for (int i = gthread; i < Q.nq; i += nthreads) {
    const int ch = __ldcg(A.assigned + i);
    if (ch >= 0) {
      atomicMax(A.blk[0] + ch, i);
      mine++;
      if (ori) {
        const int bin = rot_bin(A.angle[i], F.kps[ch].angle);
        atomicAdd(A.blk[1] + bin, 1);
        A.assigned[i] = ch | (bin << 24);
      }
    }
  }
  cluster.sync();
  for (int k = gthread; k < F.n; k += nthreads) {
    const int w = __ldcg(A.blk[0] + k);
    if (w >= 0) A.kpMp[k] = w;
  }
  if (ori) {
    cluster.sync();   // winners are in place before the histogram filter clears some of them
    int hist[30];
#pragma unroll
    for (int bI = 0; bI < 30; bI++) hist[bI] = __ldcg(A.blk[1] + bI);
    int i1, i2, i3;
    three_maxima(hist, 30, i1, i2, i3);
    for (int i = gthread; i < Q.nq; i += nthreads) {
      const int a = A.assigned[i];   // written by this thread above
      if (a >= 0) {
        const int bin = a >> 24;
        if (bin != i1 && bin != i2 && bin != i3) { A.kpMp[a & 0xffffff] = -1; mine--; }
      }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
  if (lane == 0 && mine) atomicAdd(A.nmatches, mine);
}

// SearchByProjection(Frame& Cur, const Frame& Last, ...) on a fisheye-stereo current frame (ORBmatcher.cc:2027-2289): the left and the
// right image are resolved as two k_resolve_par problems (claims never cross the images), but the rotation histogram is ONE
// (:2176-2186 and :2240-2252 push into the same rotHist): this kernel runs after both, bins every match of both images, keeps
// the three maxima and clears the rest (:2264-2283).  *args[0].nmatches receives the total.
__global__ void __launch_bounds__(256) k_rotfilter_pair(const GridFrame* frames, const QuerySet* sets, const ResolveArgs* args, int checkOrientation) {
  __shared__ int hist[30];
  __shared__ int keep[3];
  __shared__ int removedAll;
  const int tid = threadIdx.x;
  if (tid < 30) hist[tid] = 0;
  if (tid == 0) removedAll = 0;
  __syncthreads();
  if (checkOrientation) {
    for (int j = 0; j < 2; j++) {
      const ResolveArgs A = args[j];
      for (int i = tid; i < sets[j].nq; i += 256) {
        const int ch = A.assigned[i];
        if (ch >= 0) atomicAdd(&hist[rot_bin(A.angle[i], frames[j].kps[ch].angle)], 1);
      }
    }
    __syncthreads();
    if (tid == 0) { int i1, i2, i3; three_maxima(hist, 30, i1, i2, i3); keep[0] = i1; keep[1] = i2; keep[2] = i3; }
    __syncthreads();
    int removed = 0;
    for (int j = 0; j < 2; j++) {
      const ResolveArgs A = args[j];
      for (int i = tid; i < sets[j].nq; i += 256) {
        const int ch = A.assigned[i];
        if (ch < 0) continue;
        const int bin = rot_bin(A.angle[i], frames[j].kps[ch].angle);
        if (bin != keep[0] && bin != keep[1] && bin != keep[2]) { A.kpMp[ch] = -1; removed++; }
      }
    }
    atomicAdd(&removedAll, removed);
    __syncthreads();
  }
  if (tid == 0) *args[0].nmatches = *args[0].nmatches + *args[1].nmatches - removedAll;
}

struct InitArgs {  // SearchForInitialization state, one per frame pair
  const orbx_keypoint* kps1;   // F1.mvKeysUn
  float* prevMatched;          // [n1][2] in/out
  int* matches12;              // [n1] out
  int* matches21;              // [n2] scratch
  int* matchedDist;            // [n2] scratch
  int* bins;                   // [n1] scratch
  int* nmatches;
  float nnratio;
  int checkOrientation;
};

// ORBmatcher::SearchForInitialization, ORBmatcher.cc:722-837.
__global__ void __launch_bounds__(32) k_resolve_init(const GridFrame* frames, const QuerySet* sets, const InitArgs* args) {
  __shared__ int hist[30];
  const GridFrame F2 = frames[blockIdx.x];
  const QuerySet Q = sets[blockIdx.x];
  const InitArgs A = args[blockIdx.x];
  const int lane = threadIdx.x;
  volatile int* matchedDist = A.matchedDist;
  volatile int* matches21 = A.matches21;
  volatile int* matches12 = A.matches12;
  for (int k = lane; k < F2.n; k += 32) { matchedDist[k] = INT_MAX; matches21[k] = -1; }
  for (int k = lane; k < Q.nq; k += 32) { matches12[k] = -1; A.bins[k] = -1; }
  if (lane < 30) hist[lane] = 0;
  __syncwarp();
  int nmatches = 0;
  for (int i1 = 0; i1 < Q.nq; i1++) {
    const int c = Q.qCnt[i1];
    if (c == 0) continue;
    const unsigned long long* e = Q.entries + Q.qOff[i1];
    unsigned long long k1 = kKeyNone, k2 = kKeyNone;
    for (int j = lane; j < c; j += 32) {
      const unsigned long long k = e[j];
      if (matchedDist[key_idx(k)] <= key_dist(k)) continue;
      top2_insert(k, k1, k2);
    }
    top2_warp(k1, k2);
    if (k1 == kKeyNone) continue;
    const int bestDist = key_dist(k1), bestIdx2 = key_idx(k1);
    const float bestDist2 = (k2 == kKeyNone) ? (float)INT_MAX : (float)key_dist(k2);
    if (bestDist <= 50 && (float)bestDist < __fmul_rn(bestDist2, A.nnratio)) {  // TH_LOW
      const int prev = matches21[bestIdx2];
      if (lane == 0) {
        if (prev >= 0) matches12[prev] = -1;
        matches12[i1] = bestIdx2;
        matches21[bestIdx2] = i1;
        matchedDist[bestIdx2] = bestDist;
        if (A.checkOrientation) {
          const int bin = rot_bin(A.kps1[i1].angle, F2.kps[bestIdx2].angle);
          hist[bin]++;
          A.bins[i1] = bin;
        }
      }
      if (prev >= 0) nmatches--;
      nmatches++;
      __syncwarp();
    }
  }
  __syncwarp();
  if (A.checkOrientation) {
    int i1, i2, i3;
    three_maxima(hist, 30, i1, i2, i3);
    int removed = 0;
    for (int i = lane; i < Q.nq; i += 32) {
      const int bin = ((volatile int*)A.bins)[i];
      if (bin >= 0 && bin != i1 && bin != i2 && bin != i3 && matches12[i] >= 0) { matches12[i] = -1; removed++; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) removed += __shfl_xor_sync(0xffffffffu, removed, o);
    nmatches -= removed;
  }
  __syncwarp();
  for (int i = lane; i < Q.nq; i += 32) {
    const int m = matches12[i];
    if (m >= 0) { A.prevMatched[2 * i] = F2.kps[m].x; A.prevMatched[2 * i + 1] = F2.kps[m].y; }
  }
  if (lane == 0) *A.nmatches = nmatches;
}

// ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, ...) on a fisheye-STEREO frame (F.Nleft != -1), ORBmatcher.cc:44-211:
// per map point a left-camera pass over the left grid and a right-camera pass over mGridRight, both claiming F.mvpMapPoints.
// A match also hands the map point to the keypoint's stereo partner in the other image (mvLeftToRightMatch /
// mvRightToLeftMatch, :131-135 and :196-200) WITHOUT asking who held it, so a blocking holder can be replaced by a
// non-blocking one and the claims are not monotone: the reference's order is replayed by one warp (lanes = window candidates),
// the per-keypoint "held by a map point with observations" bit lives in shared memory.  frames / sets / args [0] = left image,
// [1] = right image; both query sets have one entry per map point.  The left pass's `continue` on a failed ratio test (:121-122)
// skips the map point's right pass, as in the reference.
__global__ void __launch_bounds__(32) k_resolve_fisheye(const GridFrame* frames, const QuerySet* sets, const ResolveArgs* args) {
  extern __shared__ uint8_t sBlocked[];   // [nL + nR]
  const GridFrame FL = frames[0], FR = frames[1];
  const QuerySet QL = sets[0], QR = sets[1];
  const ResolveArgs AL = args[0], AR = args[1];
  const int lane = threadIdx.x, nL = FL.n, nR = FR.n;
  auto obs = [&](int i) { return AL.observed ? AL.observed[i] != 0 : true; };
  for (int k = lane; k < nL + nR; k += 32) {
    const int s = k < nL ? AL.kpMp[k] : AR.kpMp[k - nL];
    sBlocked[k] = (s == -2) || (s >= 0 && obs(s));
  }
  __syncwarp();
  int nmatches = 0;
  for (int i = 0; i < QL.nq; i++) {
    const uint8_t ob = obs(i) ? 1 : 0;
    bool skipRight = false;
    if (QL.valid[i] && QL.qCnt[i] > 0) {
      const unsigned long long* e = QL.entries + QL.qOff[i];
      const int c = QL.qCnt[i];
      unsigned long long k1 = kKeyNone, k2 = kKeyNone;
      for (int j = lane; j < c; j += 32) {
        const unsigned long long k = e[j];
        if (!sBlocked[key_idx(k)]) top2_insert(k, k1, k2);
      }
      top2_warp(k1, k2);
      if (k1 != kKeyNone && key_dist(k1) <= 100) {   // TH_HIGH
        const int bestDist = key_dist(k1), bestLevel = key_oct(k1);
        const int bestDist2 = (k2 == kKeyNone) ? 256 : key_dist(k2), bestLevel2 = (k2 == kKeyNone) ? -1 : key_oct(k2);
        if (bestLevel == bestLevel2 && (float)bestDist > __fmul_rn(AL.nnratio, (float)bestDist2)) skipRight = true;
        else {
          const int best = key_idx(k1);
          const int p = AL.partner ? AL.partner[best] : -1;
          if (lane == 0) {
            AL.kpMp[best] = i; sBlocked[best] = ob;
            if (p != -1) { AR.kpMp[p] = i; sBlocked[nL + p] = ob; }
          }
          nmatches += (p != -1) ? 2 : 1;
          __syncwarp();
        }
      }
    }
    if (skipRight) continue;
    if (QR.valid[i] && QR.qCnt[i] > 0) {
      const unsigned long long* e = QR.entries + QR.qOff[i];
      const int c = QR.qCnt[i];
      unsigned long long k1 = kKeyNone, k2 = kKeyNone;
      for (int j = lane; j < c; j += 32) {
        const unsigned long long k = e[j];
        if (!sBlocked[nL + key_idx(k)]) top2_insert(k, k1, k2);
      }
      top2_warp(k1, k2);
      if (k1 != kKeyNone && key_dist(k1) <= 100) {
        const int bestDist = key_dist(k1), bestLevel = key_oct(k1);
        const int bestDist2 = (k2 == kKeyNone) ? 256 : key_dist(k2), bestLevel2 = (k2 == kKeyNone) ? -1 : key_oct(k2);
        if (!(bestLevel == bestLevel2 && (float)bestDist > __fmul_rn(AL.nnratio, (float)bestDist2))) {
          const int best = key_idx(k1);
          const int p = AR.partner ? AR.partner[best] : -1;
          if (lane == 0) {
            if (p != -1) { AL.kpMp[p] = i; sBlocked[p] = ob; }
            AR.kpMp[best] = i; sBlocked[nL + best] = ob;
          }
          nmatches += (p != -1) ? 2 : 1;
          __syncwarp();
        }
      }
    }
  }
  if (lane == 0) { *AL.nmatches = nmatches; *AR.nmatches = nmatches; }
}

// The same search as a parallel fixed point, for the common case in which every map point that takes part has observations (then a
// keypoint, once written -- directly or through the partner hand-over -- stays blocked for every later query, and the claims are
// monotone again).  Queries are ordered as the reference walks them: q = 2 i (left pass of map point i), q = 2 i + 1 (its right pass);
// blk[side][k] = smallest q that writes keypoint k of that image (its own choice or the stereo partner of its choice); k is blocked
// for q iff blk < q; the right pass of i is skipped when its left pass failed the ratio test (a per-point flag of the same iteration
// scheme).  choice[q] depends only on {choice[q'], skip : q' < q}, so iterating from "nothing written" reaches the sequential result.
// One cluster of 8 CTAs, warp = query.  The host picks this kernel when no participating point has observed == 0.
__global__ void __cluster_dims__(kParCluster, 1, 1) __launch_bounds__(kParThreads)
k_resolve_fisheye_par(const GridFrame* frames, const QuerySet* sets, const ResolveArgs* args) {
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  const GridFrame F[2] = {frames[0], frames[1]};
  const QuerySet Q[2] = {sets[0], sets[1]};
  const ResolveArgs A[2] = {args[0], args[1]};
  const int lane = threadIdx.x & 31;
  const int gthread = blockIdx.x * kParThreads + threadIdx.x, nthreads = kParCluster * kParThreads;
  const int gwarp = gthread >> 5, nwarps = nthreads >> 5;
  const int nmp = Q[0].nq;
  int* const skip = A[1].bestDist;   // per map point: its left pass failed the ratio test (right pass skipped); scratch [nmp]
  auto pre = [&](int s, int k) { const int v = A[s].kpMp[k]; return (v == -2 || v >= 0) ? -1 : INT_MAX; };   // (every index written earlier is an observed point)
  for (int s = 0; s < 2; s++)
    for (int k = gthread; k < F[s].n; k += nthreads) { A[s].blk[0][k] = pre(s, k); A[s].blk[1][k] = pre(s, k); }
  for (int i = gthread; i < nmp; i += nthreads) { A[0].assigned[i] = -1; A[1].assigned[i] = -1; skip[i] = 0; }
  if (gthread < 2) A[0].chg[gthread] = 0;
  cluster.sync();
  for (int it = 0; it <= 2 * nmp; it++) {
    const int cb = it & 1;
    bool changed = false;
    for (int q = gwarp; q < 2 * nmp; q += nwarps) {
      const int s = q & 1, i = q >> 1;
      int c = Q[s].valid[i] ? Q[s].qCnt[i] : 0;
      if (s == 1 && __ldcg(skip + i)) c = 0;
      int choice = -1, sk = 0;
      if (c > 0) {
        const unsigned long long* e = Q[s].entries + Q[s].qOff[i];
        const int* cur = A[s].blk[cb];
        unsigned long long k1 = kKeyNone, k2 = kKeyNone;
        for (int j = lane; j < c; j += 32) {
          const unsigned long long k = e[j];
          if (!(__ldcg(cur + key_idx(k)) < q)) top2_insert(k, k1, k2);
        }
        top2_warp(k1, k2);
        if (k1 != kKeyNone && key_dist(k1) <= 100) {   // TH_HIGH
          const int bestDist = key_dist(k1), bestLevel = key_oct(k1);
          const int bestDist2 = (k2 == kKeyNone) ? 256 : key_dist(k2), bestLevel2 = (k2 == kKeyNone) ? -1 : key_oct(k2);
          if (bestLevel == bestLevel2 && (float)bestDist > __fmul_rn(A[0].nnratio, (float)bestDist2)) sk = 1;
          else choice = key_idx(k1);
        }
      }
      if (lane == 0) {
        if (choice != A[s].assigned[i]) { A[s].assigned[i] = choice; changed = true; }
        if (s == 0 && sk != skip[i]) { skip[i] = sk; changed = true; }
        if (choice >= 0) {
          atomicMin(A[s].blk[cb ^ 1] + choice, q);
          const int p = A[s].partner ? A[s].partner[choice] : -1;
          if (p != -1) atomicMin(A[s ^ 1].blk[cb ^ 1] + p, q);
        }
      }
    }
    if (changed) atomicExch(A[0].chg + cb, 1);
    cluster.sync();
    const int any = __ldcg(A[0].chg + cb);
    for (int s = 0; s < 2; s++)
      for (int k = gthread; k < F[s].n; k += nthreads) A[s].blk[cb][k] = pre(s, k);
    if (gthread == 0) A[0].chg[cb ^ 1] = 0;
    cluster.sync();
    if (!any) break;
  }
  // fixed point reached: the holder of a keypoint is its LAST writer (the hand-over overwrites), :131-135, :196-203
  for (int s = 0; s < 2; s++)
    for (int k = gthread; k < F[s].n; k += nthreads) A[s].blk[0][k] = -1;
  if (gthread == 0) { *A[0].nmatches = 0; *A[1].nmatches = 0; }
  cluster.sync();
  int mine = 0;
  for (int q = gthread; q < 2 * nmp; q += nthreads) {
    const int s = q & 1, i = q >> 1;
    const int ch = __ldcg(A[s].assigned + i);
    if (ch < 0) continue;
    atomicMax(A[s].blk[0] + ch, q);
    mine++;
    const int p = A[s].partner ? A[s].partner[ch] : -1;
    if (p != -1) { atomicMax(A[s ^ 1].blk[0] + p, q); mine++; }
  }
  cluster.sync();
  for (int s = 0; s < 2; s++)
    for (int k = gthread; k < F[s].n; k += nthreads) {
      const int w = __ldcg(A[s].blk[0] + k);
      if (w >= 0) A[s].kpMp[k] = w >> 1;
    }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
  if (lane == 0 && mine) { atomicAdd(A[0].nmatches, mine); }
}

// ---- brute-force 2-NN ----------------------------------------------------------------------------------------------------
// thread = one query (descriptor in 8 registers); train descriptors stream through shared memory in tiles of 128 and are
// read as broadcast LDS.128; best/second inside one launch are packed 32-bit keys (dist << 23 | local train index) so the
// update is three integer min/max; lower index wins ties exactly like cv::BFMatcher.  The running state across launches
// (successive train chunks of the all-pairs sweep arriving over NCCL) is a pair of 64-bit keys (dist << 32 | global index).
// `exclude` (optional, per query) removes one contiguous global train range (a keyframe is not matched against itself);
// only the few tiles that overlap the block's exclusion ranges take the checked path.
__global__ void __launch_bounds__(128) k_knn2(const uint8_t* q, int nq, const uint8_t* t, int nt, long long trainBase, const int2* exclude,
                                              unsigned long long* key0, unsigned long long* key1, int init) {
  __shared__ uint4 tile[128 * 2];
  __shared__ int sExMin, sExMax;
  const int i = blockIdx.x * 128 + threadIdx.x;
  uint32_t qd[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  uint32_t k0 = 0xffffffffu, k1 = 0xffffffffu;
  int ex0 = 0, ex1 = 0;  // local (chunk-relative) exclusion range
  if (threadIdx.x == 0) { sExMin = INT_MAX; sExMax = INT_MIN; }
  __syncthreads();
  if (i < nq) {
    load_desc(q + 32ll * i, qd);
    if (exclude) {
      const int2 ex = exclude[i];
      ex0 = (int)max(0ll, min((long long)nt, (long long)ex.x - trainBase));
      ex1 = (int)max(0ll, min((long long)nt, (long long)ex.y - trainBase));
      if (ex1 > ex0) { atomicMin(&sExMin, ex0); atomicMax(&sExMax, ex1); }
    }
  }
  __syncthreads();
  const int bExMin = sExMin, bExMax = sExMax;
  for (int base = 0; base < nt; base += 128) {
    const int m = min(128, nt - base);
    __syncthreads();
    for (int w = threadIdx.x; w < 2 * m; w += 128) tile[w] = __ldg(reinterpret_cast<const uint4*>(t + 32ll * base) + w);
    __syncthreads();
    if (i >= nq) continue;
    const bool checked = (base < bExMax) && (base + m > bExMin);
    if (!checked) {
#pragma unroll 4
      for (int j = 0; j < m; j++) {
        const uint4 a = tile[2 * j], b = tile[2 * j + 1];
        const int d = __popc(qd[0] ^ a.x) + __popc(qd[1] ^ a.y) + __popc(qd[2] ^ a.z) + __popc(qd[3] ^ a.w) + __popc(qd[4] ^ b.x) +
                      __popc(qd[5] ^ b.y) + __popc(qd[6] ^ b.z) + __popc(qd[7] ^ b.w);
        const uint32_t k = ((uint32_t)d << 23) | (uint32_t)(base + j);
        const uint32_t lo = min(k, k0), hi = max(k, k0);
        k0 = lo;
        k1 = min(k1, hi);
      }
    } else {
      for (int j = 0; j < m; j++) {
        const int lj = base + j;
        if (lj >= ex0 && lj < ex1) continue;
        const uint4 a = tile[2 * j], b = tile[2 * j + 1];
        const int d = __popc(qd[0] ^ a.x) + __popc(qd[1] ^ a.y) + __popc(qd[2] ^ a.z) + __popc(qd[3] ^ a.w) + __popc(qd[4] ^ b.x) +
                      __popc(qd[5] ^ b.y) + __popc(qd[6] ^ b.z) + __popc(qd[7] ^ b.w);
        const uint32_t k = ((uint32_t)d << 23) | (uint32_t)lj;
        const uint32_t lo = min(k, k0), hi = max(k, k0);
        k0 = lo;
        k1 = min(k1, hi);
      }
    }
  }
  if (i < nq) {
    unsigned long long g0 = kKeyNone, g1 = kKeyNone;
    if (!init) { g0 = key0[i]; g1 = key1[i]; }
    if (k0 != 0xffffffffu) top2_insert(((unsigned long long)(k0 >> 23) << 32) | (unsigned long long)(trainBase + (k0 & 0x7fffffu)), g0, g1);
    if (k1 != 0xffffffffu) top2_insert(((unsigned long long)(k1 >> 23) << 32) | (unsigned long long)(trainBase + (k1 & 0x7fffffu)), g0, g1);
    key0[i] = g0; key1[i] = g1;
  }
}

// ---- stereo -----------------------------------------------------------------------------------------------------------------
struct ScaleTab { float scale[ORBX_MAX_LEVELS]; float invScale[ORBX_MAX_LEVELS]; };

struct StereoOut {   // [nframes][kpStride]
  float* uRight; float* depth; int* bestDist; int* bestIdx; int* sad;
};

// ---- row table (Frame.cc:911-928) -------------------------------------------------------------------------------------------
// The reference lists every right keypoint iR in the rows floor(y - r) .. ceil(y + r), r = 2 * scale[octave], and a left
// keypoint only looks at the list of its own row.  k_stereo_rowtable builds that table per frame in CSR form (one CTA per
// frame: shared-memory histogram over the image rows, exclusive scan, fill); an item carries everything the gates need
// (index, octave, x) so k_stereo_hamming2 touches a right keypoint's record only for its 32-byte descriptor.  ~27 items per
// row instead of a scan over all ~1200 right keypoints per left keypoint.
struct RowTable {
  int* rowStart;   // [nframes][rows + 1]
  uint2* items;    // [nframes][itemCap]   x = iR | octave << 16, y = float bits of kp.x
  int rows, itemCap;
};

__global__ void __launch_bounds__(256) k_stereo_rowtable(const __grid_constant__ Geom g, const __grid_constant__ Batch bR,
                                                         const __grid_constant__ ScaleTab tab, RowTable rt) {
  extern __shared__ int srow[];   // [rows] counts -> starts, [rows] fill cursors
  __shared__ int warpTmp[8];
  const int f = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int rows = rt.rows;
  int* cnt = srow;
  int* cur = srow + rows;
  const int nR = bR.counts[f];
  const long long o = (long long)f * g.kpStride;
  for (int i = tid; i < rows; i += 256) { cnt[i] = 0; cur[i] = 0; }
  __syncthreads();
  for (int iR = tid; iR < nR; iR += 256) {
    const orbx_keypoint kr = bR.kps[o + iR];
    const float r = __fmul_rn(2.0f, tab.scale[kr.octave]);
    const int maxr = min((int)ceilf(__fadd_rn(kr.y, r)), rows - 1), minr = max((int)floorf(__fsub_rn(kr.y, r)), 0);
    for (int y = minr; y <= maxr; y++) atomicAdd(&cnt[y], 1);
  }
  __syncthreads();
  // exclusive scan of cnt[0..rows): thread = contiguous chunk
  const int chunk = (rows + 255) >> 8;
  const int beg = min(rows, tid * chunk), end = min(rows, beg + chunk);
  int sum = 0;
  for (int i = beg; i < end; i++) sum += cnt[i];
  int incl = sum;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += v;
  }
  if (lane == 31) warpTmp[warp] = incl;
  __syncthreads();
  int woff = 0;
#pragma unroll
  for (int w = 0; w < 8; w++) if (w < warp) woff += warpTmp[w];
  int run = woff + incl - sum;
  int* rowStart = rt.rowStart + (long long)f * (rows + 1);
  for (int i = beg; i < end; i++) { const int v = cnt[i]; cnt[i] = run; rowStart[i] = min(run, rt.itemCap); run += v; }
  if (tid == 255) rowStart[rows] = min(run, rt.itemCap);
  __syncthreads();
  uint2* items = rt.items + (long long)f * rt.itemCap;
  for (int iR = tid; iR < nR; iR += 256) {
    const orbx_keypoint kr = bR.kps[o + iR];
    const float r = __fmul_rn(2.0f, tab.scale[kr.octave]);
    const int maxr = min((int)ceilf(__fadd_rn(kr.y, r)), rows - 1), minr = max((int)floorf(__fsub_rn(kr.y, r)), 0);
    const uint2 it = make_uint2((uint32_t)iR | ((uint32_t)kr.octave << 16), __float_as_uint(kr.x));
    for (int y = minr; y <= maxr; y++) {
      const int pos = cnt[y] + atomicAdd(&cur[y], 1);
      if (pos < rt.itemCap) items[pos] = it;
    }
  }
}

// Hamming stage of Frame::ComputeStereoMatches (Frame.cc:936-984) over the row table: one warp per left keypoint, lanes = items of
// its row.  Ties in distance go to the lower iR, as the reference's strict `dist < bestDist` over an iR-ascending list does.
__global__ void __launch_bounds__(256) k_stereo_hamming2(const __grid_constant__ Geom g, const __grid_constant__ Batch bL,
                                                         const __grid_constant__ Batch bR, float maxD, RowTable rt, StereoOut out) {
  const int f = blockIdx.y, lane = threadIdx.x & 31;
  const int iL = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int nL = bL.counts[f];
  if (iL >= nL) return;
  const long long o = (long long)f * g.kpStride;
  const orbx_keypoint kl = bL.kps[o + iL];
  uint32_t dl[8];
  load_desc(bL.desc + (o + iL) * 32, dl);   // independent of the row table: in flight while the row bounds and items arrive
  const int levelL = kl.octave;
  const float vL = kl.y, uL = kl.x;
  const int row = (int)vL;
  const float minU = __fsub_rn(uL, maxD), maxU = uL;
  uint32_t key = 100u << 16;  // TH_HIGH: only dist < 100 replaces it
  if (!(maxU < 0) && row >= 0 && row < rt.rows) {
    const int* rowStart = rt.rowStart + (long long)f * (rt.rows + 1);
    const int beg = rowStart[row], end = rowStart[row + 1];
    const uint2* items = rt.items + (long long)f * rt.itemCap;
    for (int j = beg + lane; j < end; j += 32) {
      const uint2 it = items[j];
      const int oct = (int)(it.x >> 16), iR = (int)(it.x & 0xffffu);
      const float uR = __uint_as_float(it.y);
      if (oct < levelL - 1 || oct > levelL + 1) continue;
      if (uR >= minU && uR <= maxU) {
        uint32_t dr[8];
        load_desc(bR.desc + (o + iR) * 32, dr);
        key = min(key, ((uint32_t)hamming8(dl, dr) << 16) | (uint32_t)iR);
      }
    }
  }
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) key = min(key, __shfl_xor_sync(0xffffffffu, key, s));
  if (lane == 0) {
    const int d = key >> 16;
    out.bestDist[o + iL] = d;
    out.bestIdx[o + iL] = d < 100 ? (int)(key & 0xffffu) : -1;
    out.uRight[o + iL] = -1.0f; out.depth[o + iL] = -1.0f; out.sad[o + iL] = -1;
  }
}

// Sub-pixel refinement (Frame.cc:987-1062): 11x11 centre-subtracted SAD over 11 shifts on the keypoint's pyramid level, parabola fit.
__global__ void __launch_bounds__(256) k_stereo_sad(const __grid_constant__ Geom g, const __grid_constant__ Batch bL,
                                                    const __grid_constant__ Batch bR, const __grid_constant__ ScaleTab tab,
                                                    float maxD, float mbf, StereoOut out) {
  const int f = blockIdx.y, lane = threadIdx.x & 31;
  const int iL = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (iL >= bL.counts[f]) return;
  const long long o = (long long)f * g.kpStride;
  if (out.bestDist[o + iL] >= 75) return;  // thOrbDist = (TH_HIGH+TH_LOW)/2
  const int bestIdxR = out.bestIdx[o + iL];
  const orbx_keypoint kl = bL.kps[o + iL];
  const int levelL = kl.octave;
  const float uL = kl.x, vL = kl.y;
  const float uR0 = bR.kps[o + bestIdxR].x;
  const float sf = tab.invScale[levelL];
  const float scaleduL = roundf(__fmul_rn(uL, sf)), scaledvL = roundf(__fmul_rn(vL, sf)), scaleduR0 = roundf(__fmul_rn(uR0, sf));
  const int w = 5, L = 5;
  const LevelGeom& G = g.lv[levelL];
  int spL, spR;
  const uint8_t* imL = level_ptr(g, bL, levelL, f, spL);
  const uint8_t* imR = level_ptr(g, bR, levelL, f, spR);
  const int y0 = (int)(scaledvL - w), xl0 = (int)(scaleduL - w);
  if (y0 < 0 || y0 + 2 * w + 1 > G.h || xl0 < 0 || xl0 + 2 * w + 1 > G.w) return;  // the reference throws cv::Exception here
  const float iniu = __fsub_rn(__fadd_rn(scaleduR0, (float)L), (float)w);
  const float endu = __fadd_rn(__fadd_rn(__fadd_rn(scaleduR0, (float)L), (float)w), 1.0f);
  if (iniu < 0 || endu >= (float)G.w) return;
  const int xrBase = (int)scaleduR0 - w;  // leftmost column of the incR = 0 window
  if (xrBase - L < 0 || xrBase + L + 2 * w + 1 > G.w) return;  // cv::Exception in the reference
  const int cL = imL[(long long)(y0 + w) * spL + xl0 + w];
  int il[4], py[4], px[4];
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const int p = lane + 32 * k;
    py[k] = p / 11; px[k] = p - py[k] * 11;
    il[k] = (p < 121) ? (int)imL[(long long)(y0 + py[k]) * spL + xl0 + px[k]] - cL : 0;
  }
  // The 11 shifted windows overlap in an 11 x 21 strip of the right image: stage it once per warp in shared memory (8 byte
  // loads per lane, all in flight together) instead of 44 dependent one-byte global loads per lane.  Every exit above is
  // warp-uniform (it depends on iL only), so a warp-level barrier is enough.
  __shared__ uint8_t sR[8][11 * 24];
  uint8_t* strip = sR[threadIdx.x >> 5];
#pragma unroll
  for (int k = 0; k < 8; k++) {
    const int p = lane + 32 * k;
    if (p < 231) {
      const int r = p / 21, c = p - r * 21;
      strip[r * 24 + c] = imR[(long long)(y0 + r) * spR + xrBase - L + c];
    }
  }
  __syncwarp();
  int off[4];
#pragma unroll
  for (int k = 0; k < 4; k++) off[k] = py[k] * 24 + px[k];
  int bestDistS = INT_MAX, bestincR = 0;
  float vDists[11];
#pragma unroll
  for (int incR = -5; incR <= 5; incR++) {
    const int sh = incR + L;                      // column of the window's left edge inside the strip
    const int cR = strip[w * 24 + sh + w];
    int sum = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int p = lane + 32 * k;
      if (p < 121) sum += abs(il[k] - ((int)strip[off[k] + sh] - cR));
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, s);
    const float dist = (float)sum;
    if (dist < (float)bestDistS) { bestDistS = (int)dist; bestincR = incR; }
    vDists[incR + 5] = dist;
  }
  if (bestincR == -L || bestincR == L) return;
  float dist1 = 0, dist2 = 0, dist3 = 0;
#pragma unroll
  for (int k = 1; k < 10; k++)
    if (k == bestincR + 5) { dist1 = vDists[k - 1]; dist2 = vDists[k]; dist3 = vDists[k + 1]; }
  const float deltaR = __fdiv_rn(__fsub_rn(dist1, dist3), __fmul_rn(2.0f, __fsub_rn(__fadd_rn(dist1, dist3), __fmul_rn(2.0f, dist2))));
  if (deltaR < -1 || deltaR > 1) return;
  float bestuR = __fmul_rn(tab.scale[levelL], __fadd_rn(__fadd_rn(scaleduR0, (float)bestincR), deltaR));
  float disparity = __fsub_rn(uL, bestuR);
  if (disparity >= 0 && disparity < maxD) {
    if (disparity <= 0) { disparity = 0.01f; bestuR = (float)((double)uL - 0.01); }
    if (lane == 0) {
      out.depth[o + iL] = __fdiv_rn(mbf, disparity);
      out.uRight[o + iL] = bestuR;
      out.sad[o + iL] = bestDistS;
    }
  }
}

// Outlier cut (Frame.cc:1065-1078): median of the accepted SAD values by (sad, iL) order, drop everything >= 1.5*1.4*median.
__global__ void __launch_bounds__(256) k_stereo_median(const __grid_constant__ Geom g, const __grid_constant__ Batch bL, StereoOut out) {
  extern __shared__ uint32_t skey[];   // accepted matches as (sad << 16 | iL) keys: distinct, ordered like the reference's sorted (dist, iL) pairs
  __shared__ int sN, sMedian;
  const int f = blockIdx.x, tid = threadIdx.x, lane = tid & 31;
  const int nL = bL.counts[f];
  const long long o = (long long)f * g.kpStride;
  if (tid == 0) { sN = 0; sMedian = -1; }
  __syncthreads();
  for (int base = 0; base < nL; base += 256) {
    const int i = base + tid;
    const int v = (i < nL) ? out.sad[o + i] : -1;     // 11x11 SAD <= 121 * 510 < 65536
    const uint32_t m = __ballot_sync(0xffffffffu, v >= 0);
    int wbase = 0;
    if (lane == 0 && m) wbase = atomicAdd(&sN, __popc(m));
    wbase = __shfl_sync(0xffffffffu, wbase, 0);
    if (v >= 0) skey[wbase + __popc(m & ((1u << lane) - 1))] = ((uint32_t)v << 16) | (uint32_t)i;
  }
  __syncthreads();
  const int n = sN;
  if (n == 0) return;
  const int target = n / 2;
  for (int i = tid; i < n; i += 256) {
    const uint32_t k = skey[i];
    int rank = 0;
    for (int j = 0; j < n; j++) rank += skey[j] < k;
    if (rank == target) sMedian = (int)(k >> 16);
  }
  __syncthreads();
  const float thDist = __fmul_rn(__fmul_rn(1.5f, 1.4f), (float)sMedian);
  for (int i = tid; i < n; i += 256) {
    const uint32_t k = skey[i];
    if (!((float)(k >> 16) < thDist)) { out.uRight[o + (k & 0xffffu)] = -1.0f; out.depth[o + (k & 0xffffu)] = -1.0f; }
  }
}

// ---- Fuse / SearchBySim3 search loops, SearchByBoW, SearchForTriangulation ------------------------------
// Independent window argmin (ORBmatcher.cc:1524-1590, 1728-1756, 1866-1895, 1946-1975): warp = query over its candidate list;
// the key order (distance, grid traversal order) reproduces "first strict minimum wins".
__global__ void __launch_bounds__(256) k_resolve_best(const QuerySet* sets, const ResolveArgs* args) {
  const QuerySet Q = sets[blockIdx.y];
  const ResolveArgs A = args[blockIdx.y];
  const int lane = threadIdx.x & 31;
  const int i = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (i >= Q.nq) return;
  const int c = Q.qCnt[i];
  unsigned long long k1 = kKeyNone;
  const unsigned long long* e = Q.entries + Q.qOff[i];
  for (int j = lane; j < c; j += 32) k1 = min(k1, e[j]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) k1 = min(k1, __shfl_xor_sync(0xffffffffu, k1, o));
  if (lane == 0) {
    A.assigned[i] = (k1 == kKeyNone) ? -1 : key_idx(k1);
    A.bestDist[i] = (k1 == kKeyNone) ? 256 : key_dist(k1);
  }
}

// One side of a bag-of-words bucketed search: the features of a keyframe / frame and its DBoW2::FeatureVector as CSR.
struct BowSide {
  const orbx_keypoint* kps; const uint8_t* desc; int n;
  const unsigned* indices;      // concatenated per-node feature index lists
  const uint8_t* elig;          // static eligibility per feature (may be null = all)
  const uint8_t* stereo;        // mvuRight >= 0 (triangulation only, may be null)
};
struct BowWork { int p1; int b0; int b1; int off; };   // position in side 1's index list; [b0,b1) = matching node's range in side 2; entry offset
struct BowArgs {
  BowSide s1, s2;
  const BowWork* work; int nwork;
  unsigned long long* entries;  // [sum (b1 - b0)]
  int* matches12;               // [n1]
  int* nmatches;
  int maxDist; float nnratio; int checkOrientation;
  int* blk[2]; int* chg; int* choice;   // parallel resolve scratch: [n2] x 2, [2], [nwork]
  // fisheye-stereo frame on side 2 (F.Nleft != -1, ORBmatcher.cc:338-431): features >= nLeft2 belong to the right image and are
  // matched separately (best only, no ratio test); nLeft2 < 0 otherwise
  int nLeft2; int* matches12R; int* choiceR;
  // triangulation gates (ORBmatcher.cc:1087-1146, Pinhole.cpp:150-164)
  float F12[9]; float epx, epy; const float* scaleFactors2; const float* levelSigma2; int coarse;
};
// bow key: dist[63:40] | position in the node list [39:20] | feature index [19:0]
__device__ __forceinline__ unsigned long long bow_key(int dist, int pos, int idx) {
  return ((unsigned long long)dist << 40) | ((unsigned long long)pos << 20) | (unsigned)idx;
}

// distance pass of both SearchByBoW overloads: warp = one side-1 feature of a common node, lanes = the node's side-2 features
__global__ void __launch_bounds__(256) k_bow_distances(const __grid_constant__ BowArgs A) {
  const int lane = threadIdx.x & 31;
  const int w = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (w >= A.nwork) return;
  const BowWork W = A.work[w];
  const int idx1 = (int)A.s1.indices[W.p1];
  const bool live = (unsigned)idx1 < (unsigned)A.s1.n && (!A.s1.elig || A.s1.elig[idx1]);
  uint32_t d1[8];
  if (live) load_desc(A.s1.desc + 32ll * idx1, d1);
  for (int p = W.b0 + lane; p < W.b1; p += 32) {
    unsigned long long key = kKeyNone;
    const int idx2 = (int)A.s2.indices[p];
    if (live && (unsigned)idx2 < (unsigned)A.s2.n && (!A.s2.elig || A.s2.elig[idx2])) {
      uint32_t d2[8];
      load_desc(A.s2.desc + 32ll * idx2, d2);
      key = bow_key(hamming8(d1, d2), p - W.b0, idx2);
    }
    A.entries[W.off + (p - W.b0)] = key;
  }
}

// Replay of the reference's order (node by node, then the node's list) with claiming of side-2 features, as a parallel fixed
// point (see k_resolve_par): a side-2 feature claimed by work item w blocks every later
// work item; choice[w] depends only on the choices of earlier work items, so iterating from "nothing claimed" converges to
// the sequential result.  Cluster of 8 CTAs per problem, warp = work item.
__global__ void __cluster_dims__(kParCluster, 1, 1) __launch_bounds__(kParThreads) k_bow_resolve_par(const __grid_constant__ BowArgs A) {
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  const int lane = threadIdx.x & 31;
  const int gthread = blockIdx.x * kParThreads + threadIdx.x, nthreads = kParCluster * kParThreads;
  const int gwarp = gthread >> 5, nwarps = nthreads >> 5;
  const bool fish = A.nLeft2 >= 0;
  for (int k = gthread; k < A.s2.n; k += nthreads) { A.blk[0][k] = INT_MAX; A.blk[1][k] = INT_MAX; }
  for (int k = gthread; k < A.s1.n; k += nthreads) { A.matches12[k] = -1; if (fish) A.matches12R[k] = -1; }
  for (int w = gthread; w < A.nwork; w += nthreads) { A.choice[w] = -1; if (fish) A.choiceR[w] = -1; }
  if (gthread < 2) A.chg[gthread] = 0;
  cluster.sync();
  for (int it = 0; it <= A.nwork; it++) {
    const int* cur = A.blk[it & 1];
    int* nxt = A.blk[(it & 1) ^ 1];
    bool changed = false;
    for (int w = gwarp; w < A.nwork; w += nwarps) {
      const BowWork W = A.work[w];
      const int idx1 = (int)A.s1.indices[W.p1];
      int choice = -1, choiceR = -1;
      if ((unsigned)idx1 < (unsigned)A.s1.n && (!A.s1.elig || A.s1.elig[idx1])) {
        unsigned long long k1 = kKeyNone, k2 = kKeyNone, r1 = kKeyNone;
        const int c = W.b1 - W.b0;
        for (int j = lane; j < c; j += 32) {
          const unsigned long long k = A.entries[W.off + j];
          if (k != kKeyNone && !(__ldcg(cur + (int)(k & 0xfffffu)) < w)) {
            if (fish && (int)(k & 0xfffffu) >= A.nLeft2) r1 = min(r1, k);
            else top2_insert(k, k1, k2);
          }
        }
        top2_warp(k1, k2);
        if (fish) {
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) r1 = min(r1, __shfl_xor_sync(0xffffffffu, r1, o));
        }
        if (k1 != kKeyNone) {
          const int bestDist1 = (int)(k1 >> 40);
          const int bestDist2 = (k2 == kKeyNone) ? 256 : (int)(k2 >> 40);
          if (bestDist1 <= A.maxDist) {
            if ((float)bestDist1 < __fmul_rn(A.nnratio, (float)bestDist2)) choice = (int)(k1 & 0xfffffu);
            // ORBmatcher.cc:400-403: inside `bestDist1 <= TH_LOW`, whatever the left ratio test said; "|| true" voids the right ratio test
            if (fish && r1 != kKeyNone && (int)(r1 >> 40) <= A.maxDist) choiceR = (int)(r1 & 0xfffffu);
          }
        }
      }
      if (lane == 0) {
        if (choice != A.choice[w]) { A.choice[w] = choice; changed = true; }
        if (choice >= 0) atomicMin(nxt + choice, w);
        if (fish) {
          if (choiceR != A.choiceR[w]) { A.choiceR[w] = choiceR; changed = true; }
          if (choiceR >= 0) atomicMin(nxt + choiceR, w);
        }
      }
    }
    if (changed) atomicExch(A.chg + (it & 1), 1);
    cluster.sync();
    const int any = __ldcg(A.chg + (it & 1));
    for (int k = gthread; k < A.s2.n; k += nthreads) A.blk[it & 1][k] = INT_MAX;
    if (gthread == 0) A.chg[(it & 1) ^ 1] = 0;
    cluster.sync();
    if (!any) break;
  }
  for (int w = gthread; w < A.nwork; w += nthreads) {
    const int idx1 = (int)A.s1.indices[A.work[w].p1];   // a feature sits in one node: one work item per idx1
    const int ch = __ldcg(A.choice + w);
    if (ch >= 0) A.matches12[idx1] = ch;
    if (fish) { const int chR = __ldcg(A.choiceR + w); if (chR >= 0) A.matches12R[idx1] = chR; }
  }
}

// SearchForTriangulation (ORBmatcher.cc:1046-1175): no claiming in the reference (vbMatched2 is never set), so every side-1
// feature is independent: last minimum within dist <= TH_LOW among the gated candidates (":dist>bestDist -> continue").
__global__ void __launch_bounds__(256) k_bow_triangulate(const __grid_constant__ BowArgs A) {
  const int lane = threadIdx.x & 31;
  const int w = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (w >= A.nwork) return;
  const BowWork W = A.work[w];
  const int idx1 = (int)A.s1.indices[W.p1];
  if ((unsigned)idx1 >= (unsigned)A.s1.n || !A.s1.elig[idx1]) return;
  const bool bStereo1 = A.s1.stereo && A.s1.stereo[idx1];
  const orbx_keypoint kp1 = A.s1.kps[idx1];
  uint32_t d1[8];
  load_desc(A.s1.desc + 32ll * idx1, d1);
  // epipolar line of kp1 in the second image
  const float a = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, A.F12[0]), __fmul_rn(kp1.y, A.F12[3])), A.F12[6]);
  const float b = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, A.F12[1]), __fmul_rn(kp1.y, A.F12[4])), A.F12[7]);
  const float c = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, A.F12[2]), __fmul_rn(kp1.y, A.F12[5])), A.F12[8]);
  const float den = __fadd_rn(__fmul_rn(a, a), __fmul_rn(b, b));
  unsigned long long best = kKeyNone;
  for (int p = W.b0 + lane; p < W.b1; p += 32) {
    const int idx2 = (int)A.s2.indices[p];
    if ((unsigned)idx2 >= (unsigned)A.s2.n || !A.s2.elig[idx2]) continue;
    uint32_t d2[8];
    load_desc(A.s2.desc + 32ll * idx2, d2);
    const int dist = hamming8(d1, d2);
    if (dist > 50) continue;  // TH_LOW
    const orbx_keypoint kp2 = A.s2.kps[idx2];
    const bool bStereo2 = A.s2.stereo && A.s2.stereo[idx2];
    if (!bStereo1 && !bStereo2) {
      const float distex = __fsub_rn(A.epx, kp2.x), distey = __fsub_rn(A.epy, kp2.y);
      if (__fadd_rn(__fmul_rn(distex, distex), __fmul_rn(distey, distey)) < __fmul_rn(100.0f, A.scaleFactors2[kp2.octave])) continue;
    }
    bool ok = A.coarse != 0;
    if (den != 0.0f) {
      const float num = __fadd_rn(__fadd_rn(__fmul_rn(a, kp2.x), __fmul_rn(b, kp2.y)), c);
      const float dsqr = __fdiv_rn(__fmul_rn(num, num), den);
      if ((double)dsqr < 3.84 * (double)A.levelSigma2[kp2.octave]) ok = true;
    }
    if (!ok) continue;
    best = min(best, bow_key(dist, 0xfffff - (p - W.b0), idx2));   // later list position wins ties
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
  if (lane == 0 && best != kKeyNone) A.matches12[idx1] = (int)(best & 0xfffffu);
}

// rotation-consistency filter over matches12 (ORBmatcher.cc:453-470, 957-975, 1190-1209) + final count: one CTA
__global__ void __launch_bounds__(256) k_bow_rotfilter(const __grid_constant__ BowArgs A) {
  __shared__ int hist[30];
  __shared__ int keep[3];
  __shared__ int total;
  const int tid = threadIdx.x;
  if (tid < 30) hist[tid] = 0;
  if (tid == 0) total = 0;
  __syncthreads();
  const int nsets = A.nLeft2 >= 0 ? 2 : 1;   // fisheye-stereo frame: the right-image matches vote and are filtered too (:405-424)
  int mine = 0;
  if (A.checkOrientation) {
    for (int s = 0; s < nsets; s++) {
      const int* mm = s ? A.matches12R : A.matches12;
      for (int i = tid; i < A.s1.n; i += 256) {
        const int m = mm[i];
        if (m >= 0) atomicAdd(&hist[rot_bin(A.s1.kps[i].angle, A.s2.kps[m].angle)], 1);
      }
    }
    __syncthreads();
    if (tid == 0) { int i1, i2, i3; three_maxima(hist, 30, i1, i2, i3); keep[0] = i1; keep[1] = i2; keep[2] = i3; }
    __syncthreads();
  }
  for (int s = 0; s < nsets; s++) {
    int* mm = s ? A.matches12R : A.matches12;
    for (int i = tid; i < A.s1.n; i += 256) {
      const int m = mm[i];
      if (m < 0) continue;
      if (A.checkOrientation) {
        const int bin = rot_bin(A.s1.kps[i].angle, A.s2.kps[m].angle);
        if (bin != keep[0] && bin != keep[1] && bin != keep[2]) { mm[i] = -1; continue; }
      }
      mine++;
    }
  }
  atomicAdd(&total, mine);
  __syncthreads();
  if (tid == 0) *A.nmatches = total;
}

// ---- MapPoint::ComputeDistinctiveDescriptors, DBoW2 tree descent ------------------------------------------
// MapPoint.cc:385-428 for a batch of map points: warp = map point.  For every row i the lanes spread over the other
// observations, distances go into a per-warp 257-bin histogram in shared memory, the median (sorted row [0.5*(N-1)]) is the
// first bin whose running count exceeds that rank; the row with the least median wins, first row on ties.
__global__ void __launch_bounds__(256) k_distinctive(const uint8_t* desc, const int* start, int npoints, int* bestIdx, int* bestMedian) {
  __shared__ int hist[8][288];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int p = blockIdx.x * 8 + warp;
  if (p >= npoints) return;
  const int s0 = start[p], N = start[p + 1] - s0;
  if (N <= 0) { if (lane == 0) { bestIdx[p] = -1; bestMedian[p] = INT_MAX; } return; }
  const uint8_t* d = desc + 32ll * s0;
  const int rank = (int)(0.5 * (double)(N - 1));
  int* h = hist[warp];
  int best = INT_MAX, bestI = 0;
  for (int i = 0; i < N; i++) {
#pragma unroll
    for (int b = 0; b < 9; b++) h[lane * 9 + b] = 0;
    __syncwarp();
    uint32_t di[8];
    load_desc(d + 32ll * i, di);
    for (int j = lane; j < N; j += 32) {
      uint32_t dj[8];
      load_desc(d + 32ll * j, dj);
      atomicAdd(&h[(i == j) ? 0 : hamming8(di, dj)], 1);
    }
    __syncwarp();
    // lane owns bins [9*lane, 9*lane+9): exclusive prefix of the lane sums, then walk the own bins
    int c[9], sum = 0;
#pragma unroll
    for (int b = 0; b < 9; b++) { c[b] = h[lane * 9 + b]; sum += c[b]; }
    int incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    int run = incl - sum, median = INT_MAX;
#pragma unroll
    for (int b = 0; b < 9; b++) {
      if (median == INT_MAX && run <= rank && rank < run + c[b]) median = lane * 9 + b;
      run += c[b];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) median = min(median, __shfl_xor_sync(0xffffffffu, median, o));
    if (median < best) { best = median; bestI = i; }
    __syncwarp();
  }
  if (lane == 0) { bestIdx[p] = bestI; bestMedian[p] = best; }
}

// TemplatedVocabulary::transform(feature, word_id, weight, nid, levelsup), TemplatedVocabulary.h:1218-1259: warp = feature, lanes =
// the children of the current node (k = 10 for ORBvoc), first strict minimum wins (key = distance | child position).
struct VocabDev {
  const int* childStart; const int* childIds; const uint8_t* nodeDesc; const double* nodeWeight; const int* nodeWordId;
  int nnodes, L;
};
__global__ void __launch_bounds__(256) k_vocab_transform(const __grid_constant__ VocabDev V, const uint8_t* desc, int n, int levelsup,
                                                         int* wordId, int* nodeId, double* weight) {
  const int lane = threadIdx.x & 31;
  const int f = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (f >= n) return;
  uint32_t d[8];
  load_desc(desc + 32ll * f, d);
  const int nidLevel = V.L - levelsup;
  int nid = 0, finalId = 0, level = 0;
  for (;;) {
    const int c0 = V.childStart[finalId], c1 = V.childStart[finalId + 1];
    if (c1 <= c0 || level >= 64) break;   // leaf (the host validated the depth)
    ++level;
    unsigned best = 0xffffffffu;
    for (int c = c0 + lane; c < c1; c += 32) {
      uint32_t cd[8];
      load_desc(V.nodeDesc + 32ll * V.childIds[c], cd);
      best = min(best, ((unsigned)hamming8(d, cd) << 20) | (unsigned)(c - c0));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
    finalId = V.childIds[c0 + (int)(best & 0xfffffu)];
    if (level == nidLevel) nid = finalId;
  }
  if (lane == 0) { wordId[f] = V.nodeWordId[finalId]; nodeId[f] = nid; weight[f] = V.nodeWeight[finalId]; }
}

}  // namespace orbx
