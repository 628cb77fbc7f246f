// Host side of the matcher: handle, device arena, query preparation (the scalar per-query arithmetic the reference does in
// its loop prologues) and the extern "C" entry points of include/orbx.h.  All distance / search work is in match_kernels.cuh.
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "frustum.cuh"
#include "knn2_tc.cuh"
#include "match_kernels.cuh"
#include "orbx_internal.h"
#include "kb8_triangulate.cuh"   // after match_kernels.cuh: uses BowArgs, load_desc, hamming8, bow_key

using namespace orbx;

struct orbx_matcher {
  int device = 0;
  cudaStream_t stream = nullptr;
  uint8_t* arena = nullptr;
  size_t arenaCap = 0, arenaUsed = 0;
  int* dStatus = nullptr;
  // all-pairs fold: int8 tile images of the current train chunk for k_knn2_tc (grown on demand), and which kernel folds
  uint8_t* tcWs = nullptr; size_t tcWsBytes = 0;
  int knn2Kernel = ORBX_KNN2_AUTO;
  // stereo results of the last orbx_stereo_match_batch
  float* sURight = nullptr; float* sDepth = nullptr; int* sBestDist = nullptr; int* sBestIdx = nullptr; int* sSad = nullptr;
  size_t sCap = 0, sHalf = 0;
  unsigned long long sEpoch = 0;   // bumped whenever a stereo buffer is reallocated (invalidates cached graphs)
  int sFrames = 0, sStride = 0, sLast = 0;
  // row table of the right keypoints (Frame.cc:911-928), two halves like the results
  int* sRowStart = nullptr; uint2* sRowItems = nullptr;
  size_t sRowStartCap = 0, sRowItemsCap = 0;
  // CUDA graph of one chunk of orbx_stereo_frames_host (extract L + extract R + stereo kernels), one per slot: the call is
  // bound by the host's launch rate otherwise (~30 launches with 2 KB of parameters each per chunk)
  struct ChunkGraph {
    cudaGraphExec_t execL = nullptr, execR = nullptr, execS = nullptr;   // extract left / extract right / stereo kernels
    int nb = 0; float mb = 0, mbf = 0;
    const orbx_extractor* exL = nullptr; const orbx_extractor* exR = nullptr;
    unsigned long long epochL = 0, epochR = 0, epochM = 0;
    void reset() {
      for (cudaGraphExec_t* e : {&execL, &execR, &execS}) { if (*e) cudaGraphExecDestroy(*e); *e = nullptr; }
    }
  } chunkGraph[2][4];    // [slot][entry]: a call's chunks come in up to three sizes (ramped first, full, tail)
  int chunkGraphNext[2] = {0, 0};
  // orbx_stereo_frames_host pipeline: one upload stream feeds both slots in issue order; per slot: images uploaded (left, right),
  // right branch joined, input buffers free again (both extractions done), slot's downloads done
  cudaStream_t upStream = nullptr;
  cudaEvent_t evUpL[2] = {}, evUpR[2] = {}, evJoin[2] = {}, evInFree[2] = {}, evSlotDone[2] = {};
  // pinned staging of run_window_jobs: host inputs are packed here and uploaded from page-locked memory (asynchronous, ~2 us per
  // copy) instead of ~15 pageable cudaMemcpyAsync per job that each block the host; results come back the same way.  Grow-only,
  // sized by the previous call's demand; a call that does not fit falls back to direct copies.
  uint8_t* hPin = nullptr; size_t hPinCap = 0, hPinWant = 0;
  uint8_t* dStage = nullptr;   // device twin of hPin: one copy each way per call, k_stage_copy moves the pieces to / from the arena
  int* hStat = nullptr;    // pinned: device status words of every chunk of a call (2 per chunk)
  static constexpr int kStatChunks = 256;
};

int orbx::matcher_device(const orbx_matcher* m) { return m->device; }

namespace {

struct StageCopy { unsigned long long stageOff, arenaOff; unsigned bytes, pad; };   // bytes: multiple of 16

// grid (pieces, 2): moves piece blockIdx.x between the staging buffer and the arena in 16-byte words
__global__ void __launch_bounds__(256) k_stage_copy(uint8_t* stage, uint8_t* arena, const StageCopy* list, int toArena) {
  const StageCopy c = list[blockIdx.x];
  const uint4* src = reinterpret_cast<const uint4*>(toArena ? stage + c.stageOff : arena + c.arenaOff);
  uint4* dst = reinterpret_cast<uint4*>(toArena ? arena + c.arenaOff : stage + c.stageOff);
  for (unsigned i = blockIdx.y * 256 + threadIdx.x; i < (c.bytes >> 4); i += 256 * gridDim.y) dst[i] = src[i];
}

struct ArenaPlan {  // two-phase bump allocation: plan sizes, ensure capacity, then carve
  size_t total = 0;
  size_t add(size_t bytes) { size_t off = total; total += (bytes + 255) & ~(size_t)255; return off; }
};

int arena_ensure(orbx_matcher* m, size_t bytes) {
  if (bytes <= m->arenaCap) return ORBX_OK;
  std::lock_guard<std::recursive_mutex> setupLock(setup_mutex());
  CK(cudaStreamSynchronize(m->stream));
  if (m->arena) CK(cudaFree(m->arena));
  m->arena = nullptr;
  m->arenaCap = 0;
  size_t cap = std::max(bytes + bytes / 4, (size_t)1 << 20);
  CK(cudaMalloc(&m->arena, cap));
  m->arenaCap = cap;
  return ORBX_OK;
}

template <class T>
T* at(orbx_matcher* m, size_t off) { return reinterpret_cast<T*>(m->arena + off); }

int check_view(const orbx_frame_view* F, bool needScale) {
  if (!F || F->n < 0 || (F->n > 0 && (!F->keypoints || !F->descriptors))) { set_error("bad frame view"); return ORBX_ERR_ARG; }
  if (needScale && (!F->scale_factors || F->nlevels < 1)) { set_error("frame view without scale factors"); return ORBX_ERR_ARG; }
  if (!(F->max_x > F->min_x) || !(F->max_y > F->min_y)) { set_error("frame view with empty bounds"); return ORBX_ERR_ARG; }
  if (F->n > kMaxKpSmem) { set_error("more than %d keypoints per frame is unsupported by the resolve kernels", kMaxKpSmem); return ORBX_ERR_UNSUPPORTED; }
  return ORBX_OK;
}

// One window-search problem: the frame, its queries (already reduced to x / y / r / level range by the caller) and where
// results go.  Everything is a host pointer; run_window_search stages it, builds the grid, lists candidates and resolves.
struct HostQueries {
  int nq = 0;
  std::vector<uint8_t> valid;
  std::vector<float> x, y, r, xr;
  std::vector<int8_t> minL, maxL;
  const uint8_t* desc = nullptr;  // nq x 32
  bool xrCheck = false;
  bool gate = false; const float* invSigma2 = nullptr; int nSigma = 0;   // Fuse chi-square gate (xr = projected right coordinate)
};

enum ResolveKind { RESOLVE_NONE, RESOLVE_MAPPOINTS, RESOLVE_WINDOW, RESOLVE_INIT, RESOLVE_BEST, RESOLVE_FISHEYE, RESOLVE_WINDOW_PAIR };

struct WindowJob {
  // frame (host, or device when devKps != null)
  const orbx_frame_view* F = nullptr;
  const orbx_keypoint* devKps = nullptr; const uint8_t* devDesc = nullptr; const int* devCount = nullptr; int devN = 0;
  float minX = 0, minY = 0, maxX = 0, maxY = 0;
  HostQueries q;
  // resolve inputs
  const uint8_t* observed = nullptr; const float* angle = nullptr;
  int* kpMp = nullptr; int kpMpLen = 0;  // host in/out
  float nnratio = 0; int distTh = 100; int anyMpBlocks = 0; int checkOrientation = 0;
  const int* partner = nullptr;   // RESOLVE_FISHEYE: mvLeftToRightMatch (left job) / mvRightToLeftMatch (right job)
  int gateJob = -1;               // RESOLVE_WINDOW_PAIR: this job's query i runs only if job gateJob's query i was valid with candidates
  // init search
  const orbx_frame_view* F1 = nullptr; float* prevMatched = nullptr; int* matches12 = nullptr;
  int* nmatches = nullptr;
  // independent argmin (Fuse / SearchBySim3)
  int* bestIdx = nullptr; int* bestDist = nullptr;
  // features_in_area
  std::vector<unsigned long long>* entriesOut = nullptr;
  // projection prologue on the device (k_project_points): the queries are derived from 3-D points and a pose instead of q.x / q.y / ...
  const orbx_projection* proj = nullptr; const orbx_projection_points* pts = nullptr; const orbx_projection_debug* dbg = nullptr;
};

// ---- projection prologue of the searches that start from map points and a pose ------------------------------------------------
// One thread per point, the reference's float / double evaluation order spelled out (no contraction):
//   x_c = R x_w + t        cv::gemm small-matrix path: ((a0*b0 + a1*b1) + a2*b2) in float, + t as one float add   (ORBmatcher.cc:2072)
//   depth test / invz      per depth_mode / invz_mode (:2077, :1490-1493, :1850-1852)
//   project                Pinhole::project / KannalaBrandt8::project (proj_mode 0) or fx * (x * invz) + cx (proj_mode 1, :1853-1858)
//   bounds                 Frame float bounds (:2084-2087) or KeyFrame::IsInImage (:540, src/KeyFrame.cc:844-847)
//   dist, range, normal    cv::norm / Mat::dot accumulate in double; 0.8f * mfMinDistance, 1.2f * mfMaxDistance (MapPoint.cc:555-565)
//   level                  given octave (:2089) or MapPoint::PredictScale (MapPoint.cc:570-602); radius = th * mvScaleFactors[level]
struct ProjectArgs {
  orbx_projection P;
  int n, nlevels;
  const float* pos; const float* normal; const float* maxD; const float* minD; const int* octave; const uint8_t* validIn;
  const float* scale;
  uint8_t* valid; float* x; float* y; float* r; int8_t* minL; int8_t* maxL; float* xr; int* level;
};

__global__ void __launch_bounds__(256) k_project_points(const __grid_constant__ ProjectArgs A) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= A.n) return;
  const orbx_projection& P = A.P;
  A.valid[i] = 0; A.x[i] = -1.f; A.y[i] = -1.f; A.r[i] = 0.f; A.minL[i] = 0; A.maxL[i] = 0; A.level[i] = -1;
  if (A.xr) A.xr[i] = 0.f;
  if (A.validIn && !A.validIn[i]) return;
  const float p0 = A.pos[3 * i], p1 = A.pos[3 * i + 1], p2 = A.pos[3 * i + 2];
  float c[3];
#pragma unroll
  for (int k = 0; k < 3; ++k)
    c[k] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(P.R[3 * k], p0), __fmul_rn(P.R[3 * k + 1], p1)), __fmul_rn(P.R[3 * k + 2], p2)), P.t[k]);
  if (P.chain) {
    float d[3];
#pragma unroll
    for (int k = 0; k < 3; ++k)
      d[k] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(P.R2[3 * k], c[0]), __fmul_rn(P.R2[3 * k + 1], c[1])), __fmul_rn(P.R2[3 * k + 2], c[2])), P.t2[k]);
    c[0] = d[0]; c[1] = d[1]; c[2] = d[2];
  }
  const float z = c[2];
  const float invz = P.invz_mode == 0 ? (float)(1.0 / (double)z) : __fdiv_rn(1.0f, z);
  if (P.depth_mode == 0) { if (invz < 0.f) return; }
  else if (P.depth_mode == 1) { if (z < 0.0f) return; }
  float u, v;
  if (P.proj_mode == 1) {
    u = __fadd_rn(__fmul_rn(P.fx, __fmul_rn(c[0], invz)), P.cx);
    v = __fadd_rn(__fmul_rn(P.fy, __fmul_rn(c[1], invz)), P.cy);
  } else if (P.camera_model == ORBX_CAMERA_KB8) {   // KannalaBrandt8.cpp:29-45, same policy as frustum_point (frustum.cuh)
    const float x2y2 = __fadd_rn(__fmul_rn(c[0], c[0]), __fmul_rn(c[1], c[1]));
    const float theta = (float)atan2((double)sqrtf(x2y2), (double)z);
    const float psi = (float)atan2((double)c[1], (double)c[0]);
    const float t2 = __fmul_rn(theta, theta), t3 = __fmul_rn(theta, t2), t5 = __fmul_rn(t3, t2), t7 = __fmul_rn(t5, t2), t9 = __fmul_rn(t7, t2);
    const float rr = __fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(theta, __fmul_rn(P.kb[0], t3)), __fmul_rn(P.kb[1], t5)), __fmul_rn(P.kb[2], t7)), __fmul_rn(P.kb[3], t9));
    u = (float)__dadd_rn(__dmul_rn((double)__fmul_rn(P.fx, rr), cos((double)psi)), (double)P.cx);
    v = (float)__dadd_rn(__dmul_rn((double)__fmul_rn(P.fy, rr), sin((double)psi)), (double)P.cy);
  } else {
    u = __fadd_rn(__fdiv_rn(__fmul_rn(P.fx, c[0]), z), P.cx);
    v = __fadd_rn(__fdiv_rn(__fmul_rn(P.fy, c[1]), z), P.cy);
  }
  A.x[i] = u; A.y[i] = v;
  if (P.bounds_mode == 0) { if (u < P.min_x || u > P.max_x || v < P.min_y || v > P.max_y) return; }
  else if (P.bounds_mode == 1) { if (!(u >= P.min_x && u < P.max_x && v >= P.min_y && v < P.max_y)) return; }
  if (A.xr) A.xr[i] = __fsub_rn(u, __fmul_rn(P.bf, invz));
  float dist = 0.f;
  float o0 = 0.f, o1 = 0.f, o2 = 0.f;
  if (P.range_mode || P.normal_check || P.level_mode == 1) {
    if (P.dist_mode == 1) { o0 = c[0]; o1 = c[1]; o2 = c[2]; }
    else { o0 = __fsub_rn(p0, P.Ow[0]); o1 = __fsub_rn(p1, P.Ow[1]); o2 = __fsub_rn(p2, P.Ow[2]); }
    dist = (float)sqrt((double)o0 * o0 + (double)o1 * o1 + (double)o2 * o2);
  }
  if (P.range_mode) {
    const float maxDistance = __fmul_rn(1.2f, A.maxD[i]), minDistance = __fmul_rn(0.8f, A.minD[i]);
    if (dist < minDistance || dist > maxDistance) return;
  }
  if (P.normal_check) {
    const double dot = (double)o0 * A.normal[3 * i] + (double)o1 * A.normal[3 * i + 1] + (double)o2 * A.normal[3 * i + 2];
    if (dot < 0.5 * (double)dist) return;
  }
  int lvl;
  if (P.level_mode == 1) {   // MapPoint::PredictScale
    const float ratio = __fdiv_rn(A.maxD[i], dist);
    const float lg = (float)log((double)ratio);
    lvl = (int)ceilf(__fdiv_rn(lg, P.log_scale_factor));
    if (lvl < 0) lvl = 0;
    else if (lvl >= P.n_scale_levels) lvl = P.n_scale_levels - 1;
  } else lvl = A.octave[i];
  A.level[i] = lvl;
  if (lvl < 0 || lvl >= A.nlevels || lvl > 126) return;
  A.r[i] = __fmul_rn(P.th, A.scale[lvl]);
  int8_t lo, hi;
  if (P.window_mode == 1) { lo = (int8_t)lvl; hi = -1; }
  else if (P.window_mode == 2) { lo = 0; hi = (int8_t)lvl; }
  else if (P.window_mode == 3) { lo = (int8_t)(lvl - 1); hi = (int8_t)lvl; }
  else { lo = (int8_t)(lvl - 1); hi = (int8_t)(lvl + 1); }
  A.minL[i] = lo; A.maxL[i] = hi;
  A.valid[i] = 1;
}

int run_window_jobs(orbx_matcher* m, std::vector<WindowJob>& jobs, ResolveKind kind) {
  const int nj = (int)jobs.size();
  if (nj == 0) return ORBX_OK;
  CK(cudaSetDevice(m->device));
  // ---- plan the arena
  ArenaPlan plan;
  struct Off { size_t kps, desc, uRight, cellStart, cellItems, valid, x, y, r, xr, minL, maxL, qdesc, qOff, qCnt, entries, cursor,
               observed, angle, kpMp, assigned, nmatches, kps1, prevMatched, m12, m21, mdist, bins, sigma, blk0, blk1, chg, partner,
               pPos, pNormal, pMaxD, pMinD, pOct, pValid, pLevel, pScale; };
  std::vector<Off> off(nj);
  std::vector<int> entryCap(nj);
  for (int j = 0; j < nj; j++) {
    WindowJob& J = jobs[j];
    const int n = J.devKps ? J.devN : J.F->n;
    const int nq = J.q.nq;
    Off& o = off[j];
    o.kps = J.devKps ? 0 : plan.add((size_t)std::max(n, 1) * sizeof(orbx_keypoint));
    o.desc = J.devKps ? 0 : plan.add((size_t)std::max(n, 1) * 32);
    o.uRight = (!J.devKps && J.F->u_right) ? plan.add((size_t)std::max(n, 1) * 4) : (size_t)-1;
    o.cellStart = plan.add((kGridCells + 1) * sizeof(int));
    o.cellItems = plan.add((size_t)std::max(n, 1) * sizeof(int));
    o.valid = plan.add(std::max(nq, 1));
    o.x = plan.add((size_t)std::max(nq, 1) * 4); o.y = plan.add((size_t)std::max(nq, 1) * 4); o.r = plan.add((size_t)std::max(nq, 1) * 4);
    o.xr = (J.q.xrCheck || (J.q.gate && !J.q.xr.empty()) || (J.proj && J.proj->want_xr)) ? plan.add((size_t)std::max(nq, 1) * 4) : (size_t)-1;
    if (J.proj) {
      o.pPos = plan.add((size_t)std::max(nq, 1) * 12);
      o.pNormal = J.pts->normal ? plan.add((size_t)std::max(nq, 1) * 12) : (size_t)-1;
      o.pMaxD = J.pts->max_distance ? plan.add((size_t)std::max(nq, 1) * 4) : (size_t)-1;
      o.pMinD = J.pts->min_distance ? plan.add((size_t)std::max(nq, 1) * 4) : (size_t)-1;
      o.pOct = J.pts->octave ? plan.add((size_t)std::max(nq, 1) * 4) : (size_t)-1;
      o.pValid = J.pts->valid ? plan.add(std::max(nq, 1)) : (size_t)-1;
      o.pLevel = plan.add((size_t)std::max(nq, 1) * 4);
      o.pScale = plan.add((size_t)std::max(J.F->nlevels, 1) * 4);
    }
    o.sigma = J.q.gate ? plan.add((size_t)std::max(J.q.nSigma, 1) * 4) : (size_t)-1;
    if (kind == RESOLVE_BEST || kind == RESOLVE_FISHEYE) o.m12 = plan.add((size_t)std::max(nq, 1) * 4);
    o.minL = plan.add(std::max(nq, 1)); o.maxL = plan.add(std::max(nq, 1));
    o.qdesc = plan.add((size_t)std::max(nq, 1) * 32);
    o.qOff = plan.add((size_t)std::max(nq, 1) * 4); o.qCnt = plan.add((size_t)std::max(nq, 1) * 4);
    // candidate capacity: windows hold a few dozen keypoints; size generously and report overflow instead of truncating
    entryCap[j] = (int)std::min<long long>((long long)nq * std::max(n, 1), std::max<long long>(1 << 16, (long long)nq * 256));
    o.entries = plan.add((size_t)std::max(entryCap[j], 1) * 8);
    o.cursor = plan.add(4);
    o.observed = J.observed ? plan.add(std::max(nq, 1)) : (size_t)-1;
    o.angle = J.angle ? plan.add((size_t)std::max(nq, 1) * 4) : (size_t)-1;
    o.kpMp = J.kpMp ? plan.add((size_t)std::max(n, 1) * 4) : (size_t)-1;
    o.assigned = plan.add((size_t)std::max(nq, 1) * 4);
    o.nmatches = plan.add(4);
    o.blk0 = plan.add((size_t)std::max(n, 32) * 4); o.blk1 = plan.add((size_t)std::max(n, 32) * 4); o.chg = plan.add(8);
    o.partner = J.partner ? plan.add((size_t)std::max(n, 1) * 4) : (size_t)-1;
    if (kind == RESOLVE_INIT) {
      o.kps1 = plan.add((size_t)std::max(nq, 1) * sizeof(orbx_keypoint));
      o.prevMatched = plan.add((size_t)std::max(nq, 1) * 8);
      o.m12 = plan.add((size_t)std::max(nq, 1) * 4);
      o.m21 = plan.add((size_t)std::max(n, 1) * 4);
      o.mdist = plan.add((size_t)std::max(n, 1) * 4);
      o.bins = plan.add((size_t)std::max(nq, 1) * 4);
    }
  }
  const size_t oFrames = plan.add(nj * sizeof(GridFrame)), oSets = plan.add(nj * sizeof(QuerySet));
  const size_t oArgs = plan.add(nj * std::max(sizeof(ResolveArgs), sizeof(InitArgs)));
  int rc = arena_ensure(m, plan.total);
  if (rc != ORBX_OK) return rc;
  cudaStream_t st = m->stream;
  if (m->hPinWant > m->hPinCap) {   // the previous call outgrew the staging buffer (the stream is idle between calls)
    std::lock_guard<std::recursive_mutex> setupLock(setup_mutex());
    if (m->hPin) cudaFreeHost(m->hPin);
    m->hPin = nullptr; m->hPinCap = 0;
    const size_t want = m->hPinWant + m->hPinWant / 4;
    if (m->dStage) cudaFree(m->dStage);
    m->dStage = nullptr;
    if (cudaHostAlloc((void**)&m->hPin, want, cudaHostAllocDefault) == cudaSuccess && cudaMalloc((void**)&m->dStage, want) == cudaSuccess) m->hPinCap = want;
    else { cudaGetLastError(); if (m->hPin) cudaFreeHost(m->hPin); m->hPin = nullptr; }
  }
  size_t pinCur = 0, pinNeed = 0;
  auto pin = [&](size_t bytes) -> uint8_t* {   // 64-byte aligned slice of the staging buffer, or null when it is full
    const size_t b = (bytes + 63) & ~(size_t)63;
    pinNeed += b;
    if (pinCur + b > m->hPinCap) return nullptr;
    uint8_t* r = m->hPin + pinCur;
    pinCur += b;
    return r;
  };
  // Inputs are packed into the page-locked buffer and recorded; flush_up() sends the whole pack with ONE copy and lets
  // k_stage_copy scatter the pieces to their arena offsets (every arena region is padded to 256 bytes, every slice to 64, so
  // whole 16-byte words can be moved).  Results travel the other way in flush_down().  Pieces that do not fit the buffer
  // (first call, or a call larger than any before) use direct copies.
  std::vector<StageCopy> upList, downList;
  auto up = [&](size_t o, const void* src, size_t bytes) -> cudaError_t {
    if (bytes == 0 || !src) return cudaSuccess;
    if (uint8_t* h = pin(bytes)) {
      memcpy(h, src, bytes);
      upList.push_back({(unsigned long long)(h - m->hPin), (unsigned long long)o, (unsigned)((bytes + 15) & ~(size_t)15), 0u});
      return cudaSuccess;
    }
    return cudaMemcpyAsync(m->arena + o, src, bytes, cudaMemcpyHostToDevice, st);
  };
  auto flush_up = [&]() -> cudaError_t {
    if (upList.empty()) return cudaSuccess;
    const size_t lb = upList.size() * sizeof(StageCopy);
    uint8_t* hl = pin(lb);
    if (!hl) {   // no room for the list itself: send the pieces one by one from the page-locked slices
      for (const StageCopy& c : upList) {
        cudaError_t e = cudaMemcpyAsync(m->arena + c.arenaOff, m->hPin + c.stageOff, c.bytes, cudaMemcpyHostToDevice, st);
        if (e != cudaSuccess) return e;
      }
      return cudaSuccess;
    }
    memcpy(hl, upList.data(), lb);
    cudaError_t e = cudaMemcpyAsync(m->dStage, m->hPin, pinCur, cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return e;
    k_stage_copy<<<dim3((unsigned)upList.size(), 2), 256, 0, st>>>(m->dStage, m->arena, reinterpret_cast<const StageCopy*>(m->dStage + (hl - m->hPin)), 1);
    return cudaGetLastError();
  };
  struct Pending { void* dst; const void* src; size_t bytes; };
  std::vector<Pending> pending;   // results parked in the staging buffer, handed to the caller after the final synchronisation
  auto down = [&](void* dst, size_t o, size_t bytes) -> cudaError_t {
    if (bytes == 0 || !dst) return cudaSuccess;
    if (uint8_t* h = pin(bytes)) {
      pending.push_back({dst, h, bytes});
      downList.push_back({(unsigned long long)(h - m->hPin), (unsigned long long)o, (unsigned)((bytes + 15) & ~(size_t)15), 0u});
      return cudaSuccess;
    }
    return cudaMemcpyAsync(dst, m->arena + o, bytes, cudaMemcpyDeviceToHost, st);
  };
  auto flush_down = [&]() -> cudaError_t {
    if (downList.empty()) return cudaSuccess;
    const size_t first = downList.front().stageOff;
    const size_t lb = downList.size() * sizeof(StageCopy);
    uint8_t* hl = pin(lb);
    if (!hl) {
      for (const StageCopy& c : downList) {
        cudaError_t e = cudaMemcpyAsync(m->hPin + c.stageOff, m->arena + c.arenaOff, c.bytes, cudaMemcpyDeviceToHost, st);
        if (e != cudaSuccess) return e;
      }
      return cudaSuccess;
    }
    memcpy(hl, downList.data(), lb);
    cudaError_t e = cudaMemcpyAsync(m->dStage + (hl - m->hPin), hl, lb, cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return e;
    k_stage_copy<<<dim3((unsigned)downList.size(), 2), 256, 0, st>>>(m->dStage, m->arena, reinterpret_cast<const StageCopy*>(m->dStage + (hl - m->hPin)), 0);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    return cudaMemcpyAsync(m->hPin + first, m->dStage + first, (size_t)(hl - m->hPin) - first, cudaMemcpyDeviceToHost, st);
  };
  std::vector<GridFrame> frames(nj);
  std::vector<QuerySet> sets(nj);
  std::vector<ResolveArgs> rargs(nj);
  std::vector<InitArgs> iargs(nj);
  int maxQ = 0, maxN = 0;
  for (int j = 0; j < nj; j++) {
    WindowJob& J = jobs[j];
    const Off& o = off[j];
    const int nq = J.q.nq;
    GridFrame& G = frames[j];
    if (J.devKps) {
      G.kps = J.devKps; G.desc = J.devDesc; G.uRight = nullptr; G.n = J.devN;
    } else {
      CK(up(o.kps, J.F->keypoints, (size_t)J.F->n * sizeof(orbx_keypoint)));
      CK(up(o.desc, J.F->descriptors, (size_t)J.F->n * 32));
      if (J.F->u_right) CK(up(o.uRight, J.F->u_right, (size_t)J.F->n * 4));
      G.kps = at<orbx_keypoint>(m, o.kps); G.desc = at<uint8_t>(m, o.desc);
      G.uRight = J.F->u_right ? at<float>(m, o.uRight) : nullptr;
      G.n = J.F->n;
    }
    G.minX = J.minX; G.minY = J.minY;
    G.invW = (float)kGridCols / (J.maxX - J.minX);  // Frame.cc:160-161
    G.invH = (float)kGridRows / (J.maxY - J.minY);
    G.cellStart = at<int>(m, o.cellStart); G.cellItems = at<int>(m, o.cellItems);
    maxN = std::max(maxN, G.n);
    QuerySet& Q = sets[j];
    Q.nq = nq;
    if (J.proj) {   // the queries are computed on the device from the points and the pose
      CK(up(o.pPos, J.pts->world_pos, (size_t)nq * 12));
      if (J.pts->normal) CK(up(o.pNormal, J.pts->normal, (size_t)nq * 12));
      if (J.pts->max_distance) CK(up(o.pMaxD, J.pts->max_distance, (size_t)nq * 4));
      if (J.pts->min_distance) CK(up(o.pMinD, J.pts->min_distance, (size_t)nq * 4));
      if (J.pts->octave) CK(up(o.pOct, J.pts->octave, (size_t)nq * 4));
      if (J.pts->valid) CK(up(o.pValid, J.pts->valid, nq));
      CK(up(o.pScale, J.F->scale_factors, (size_t)J.F->nlevels * 4));
    } else {
      CK(up(o.valid, J.q.valid.data(), nq)); CK(up(o.x, J.q.x.data(), (size_t)nq * 4)); CK(up(o.y, J.q.y.data(), (size_t)nq * 4));
      CK(up(o.r, J.q.r.data(), (size_t)nq * 4)); CK(up(o.minL, J.q.minL.data(), nq)); CK(up(o.maxL, J.q.maxL.data(), nq));
      if (o.xr != (size_t)-1) CK(up(o.xr, J.q.xr.data(), (size_t)nq * 4));
    }
    CK(up(o.qdesc, J.q.desc, (size_t)nq * 32));
    if (J.q.gate) CK(up(o.sigma, J.q.invSigma2, (size_t)J.q.nSigma * 4));
    Q.valid = at<uint8_t>(m, o.valid); Q.x = at<float>(m, o.x); Q.y = at<float>(m, o.y); Q.r = at<float>(m, o.r);
    Q.minL = at<int8_t>(m, o.minL); Q.maxL = at<int8_t>(m, o.maxL); Q.desc = at<uint8_t>(m, o.qdesc);
    Q.xr = o.xr != (size_t)-1 ? at<float>(m, o.xr) : nullptr; Q.xrCheck = J.q.xrCheck ? 1 : 0;
    Q.gate = J.q.gate ? 1 : 0; Q.invSigma2 = J.q.gate ? at<float>(m, o.sigma) : nullptr;
    Q.qOff = at<int>(m, o.qOff); Q.qCnt = at<int>(m, o.qCnt);
    Q.entries = at<unsigned long long>(m, o.entries); Q.entryCap = entryCap[j]; Q.entryCursor = at<int>(m, o.cursor);
    CK(cudaMemsetAsync(m->arena + o.cursor, 0, 4, st));
    maxQ = std::max(maxQ, nq);
    if (kind == RESOLVE_MAPPOINTS || kind == RESOLVE_WINDOW || kind == RESOLVE_FISHEYE || kind == RESOLVE_WINDOW_PAIR) {
      ResolveArgs& A = rargs[j];
      A.gateValid = J.gateJob >= 0 ? at<uint8_t>(m, off[J.gateJob].valid) : nullptr;
      A.gateCnt = J.gateJob >= 0 ? at<int>(m, off[J.gateJob].qCnt) : nullptr;
      if (J.partner) CK(up(o.partner, J.partner, (size_t)G.n * 4));
      A.partner = J.partner ? at<int>(m, o.partner) : nullptr;
      if (J.observed) CK(up(o.observed, J.observed, nq));
      if (J.angle) CK(up(o.angle, J.angle, (size_t)nq * 4));
      CK(up(o.kpMp, J.kpMp, (size_t)G.n * 4));
      A.observed = J.observed ? at<uint8_t>(m, o.observed) : nullptr;
      A.angle = J.angle ? at<float>(m, o.angle) : nullptr;
      A.kpMp = at<int>(m, o.kpMp); A.assigned = at<int>(m, o.assigned); A.nmatches = at<int>(m, o.nmatches);
      A.nnratio = J.nnratio; A.distTh = J.distTh; A.anyMpBlocks = J.anyMpBlocks;
      A.checkOrientation = kind == RESOLVE_WINDOW_PAIR ? 0 : J.checkOrientation;   // pair: one histogram over both images, k_rotfilter_pair
      A.blk[0] = at<int>(m, o.blk0); A.blk[1] = at<int>(m, o.blk1); A.chg = at<int>(m, o.chg);
      if (kind == RESOLVE_FISHEYE) A.bestDist = at<int>(m, o.m12);   // k_resolve_fisheye_par: per-point scratch
    } else if (kind == RESOLVE_BEST) {
      ResolveArgs& A = rargs[j];
      A = ResolveArgs{};
      A.assigned = at<int>(m, o.assigned); A.bestDist = at<int>(m, o.m12);
    } else if (kind == RESOLVE_INIT) {
      InitArgs& A = iargs[j];
      CK(up(o.kps1, J.F1->keypoints, (size_t)nq * sizeof(orbx_keypoint)));
      CK(up(o.prevMatched, J.prevMatched, (size_t)nq * 8));
      A.kps1 = at<orbx_keypoint>(m, o.kps1); A.prevMatched = at<float>(m, o.prevMatched); A.matches12 = at<int>(m, o.m12);
      A.matches21 = at<int>(m, o.m21); A.matchedDist = at<int>(m, o.mdist); A.bins = at<int>(m, o.bins);
      A.nmatches = at<int>(m, o.nmatches); A.nnratio = J.nnratio; A.checkOrientation = J.checkOrientation;
    }
  }
  CK(up(oFrames, frames.data(), nj * sizeof(GridFrame)));
  CK(up(oSets, sets.data(), nj * sizeof(QuerySet)));
  if (kind == RESOLVE_INIT) CK(up(oArgs, iargs.data(), nj * sizeof(InitArgs)));
  else CK(up(oArgs, rargs.data(), nj * sizeof(ResolveArgs)));
  CK(flush_up());
  CK(cudaMemsetAsync(m->dStatus, 0, sizeof(int), st));
  const GridFrame* dFrames = at<GridFrame>(m, oFrames);
  const QuerySet* dSets = at<QuerySet>(m, oSets);
  for (int j = 0; j < nj; j++) {
    const WindowJob& J = jobs[j];
    if (!J.proj || J.q.nq == 0) continue;
    const Off& o = off[j];
    ProjectArgs A{};
    A.P = *J.proj; A.n = J.q.nq; A.nlevels = J.F->nlevels;
    A.pos = at<float>(m, o.pPos); A.normal = o.pNormal != (size_t)-1 ? at<float>(m, o.pNormal) : nullptr;
    A.maxD = o.pMaxD != (size_t)-1 ? at<float>(m, o.pMaxD) : nullptr; A.minD = o.pMinD != (size_t)-1 ? at<float>(m, o.pMinD) : nullptr;
    A.octave = o.pOct != (size_t)-1 ? at<int>(m, o.pOct) : nullptr; A.validIn = o.pValid != (size_t)-1 ? at<uint8_t>(m, o.pValid) : nullptr;
    A.scale = at<float>(m, o.pScale);
    A.valid = at<uint8_t>(m, o.valid); A.x = at<float>(m, o.x); A.y = at<float>(m, o.y); A.r = at<float>(m, o.r);
    A.minL = at<int8_t>(m, o.minL); A.maxL = at<int8_t>(m, o.maxL); A.xr = o.xr != (size_t)-1 ? at<float>(m, o.xr) : nullptr;
    A.level = at<int>(m, o.pLevel);
    k_project_points<<<(A.n + 255) / 256, 256, 0, st>>>(A);
  }
  k_build_grid<<<nj, 256, 0, st>>>(dFrames);
  if (maxQ > 0) k_window_candidates<<<dim3((maxQ + 7) / 8, nj), 256, 0, st>>>(dFrames, dSets, m->dStatus);
  if (kind == RESOLVE_MAPPOINTS) {
    k_resolve_par<<<dim3(kParCluster, nj), kParThreads, 0, st>>>(dFrames, dSets, at<ResolveArgs>(m, oArgs), 0);
  } else if (kind == RESOLVE_WINDOW) {
    k_resolve_par<<<dim3(kParCluster, nj), kParThreads, 0, st>>>(dFrames, dSets, at<ResolveArgs>(m, oArgs), 1);
  }
  else if (kind == RESOLVE_WINDOW_PAIR) {   // both images with the rotation filter deferred, then the shared histogram
    k_resolve_par<<<dim3(kParCluster, nj), kParThreads, 0, st>>>(dFrames, dSets, at<ResolveArgs>(m, oArgs), 1);
    k_rotfilter_pair<<<1, 256, 0, st>>>(dFrames, dSets, at<ResolveArgs>(m, oArgs), jobs[0].checkOrientation);
  }
  else if (kind == RESOLVE_FISHEYE) {
    // claims are monotone when every participating map point has observations (the normal case: local-map points) -> cluster fixed point;
    // otherwise the partner hand-over can replace a blocking holder by a non-blocking one -> one-warp replay of the reference's order
    bool allObserved = true;
    const uint8_t* ob = jobs[0].observed;
    if (ob)
      for (int i = 0; i < jobs[0].q.nq && allObserved; i++)
        if ((jobs[0].q.valid[i] || jobs[1].q.valid[i]) && !ob[i]) allObserved = false;
    if (allObserved) k_resolve_fisheye_par<<<dim3(kParCluster, 1), kParThreads, 0, st>>>(dFrames, dSets, at<ResolveArgs>(m, oArgs));
    else k_resolve_fisheye<<<1, 32, (size_t)(frames[0].n + frames[1].n), st>>>(dFrames, dSets, at<ResolveArgs>(m, oArgs));
  }
  else if (kind == RESOLVE_INIT) k_resolve_init<<<nj, 32, 0, st>>>(dFrames, dSets, at<InitArgs>(m, oArgs));
  else if (kind == RESOLVE_BEST && maxQ > 0) k_resolve_best<<<dim3((maxQ + 7) / 8, nj), 256, 0, st>>>(dSets, at<ResolveArgs>(m, oArgs));
  CK(cudaGetLastError());
  // ---- results
  int status = 0;
  CK(cudaMemcpyAsync(&status, m->dStatus, sizeof(int), cudaMemcpyDeviceToHost, st));
  for (int j = 0; j < nj; j++) {
    WindowJob& J = jobs[j];
    const Off& o = off[j];
    if (kind == RESOLVE_MAPPOINTS || kind == RESOLVE_WINDOW || kind == RESOLVE_FISHEYE || kind == RESOLVE_WINDOW_PAIR) {
      CK(down(J.kpMp, o.kpMp, (size_t)std::min(frames[j].n, J.kpMpLen) * 4));
      CK(down(J.nmatches, o.nmatches, 4));
    } else if (kind == RESOLVE_INIT) {
      CK(down(J.matches12, o.m12, (size_t)J.q.nq * 4));
      CK(down(J.prevMatched, o.prevMatched, (size_t)J.q.nq * 8));
      CK(down(J.nmatches, o.nmatches, 4));
    } else if (kind == RESOLVE_BEST) {
      CK(down(J.bestIdx, o.assigned, (size_t)J.q.nq * 4));
      CK(down(J.bestDist, o.m12, (size_t)J.q.nq * 4));
    }
    if (J.proj && J.dbg) {   // what the prologue produced (tests compare it with the oracle)
      const int nq = J.q.nq;
      CK(down(J.dbg->valid, o.valid, nq)); CK(down(J.dbg->u, o.x, (size_t)nq * 4)); CK(down(J.dbg->v, o.y, (size_t)nq * 4));
      CK(down(J.dbg->radius, o.r, (size_t)nq * 4)); CK(down(J.dbg->level, o.pLevel, (size_t)nq * 4));
      if (o.xr != (size_t)-1) CK(down(J.dbg->xr, o.xr, (size_t)nq * 4));
    }
    if (kind == RESOLVE_NONE && J.entriesOut) {
      int cnt = 0;
      CK(cudaMemcpyAsync(&cnt, m->arena + o.qCnt, 4, cudaMemcpyDeviceToHost, st));
      CK(cudaStreamSynchronize(st));
      J.entriesOut->resize(cnt);
      if (cnt) CK(cudaMemcpyAsync(J.entriesOut->data(), m->arena + o.entries, (size_t)cnt * 8, cudaMemcpyDeviceToHost, st));
    }
  }
  CK(flush_down());
  CK(cudaStreamSynchronize(st));
  for (const Pending& P : pending) memcpy(P.dst, P.src, P.bytes);
  m->hPinWant = std::max(m->hPinWant, pinNeed);
  if (status != 0) { set_error("candidate list capacity exceeded"); return status; }
  return ORBX_OK;
}

void set_bounds(WindowJob& J, const orbx_frame_view* F) { J.minX = F->min_x; J.minY = F->min_y; J.maxX = F->max_x; J.maxY = F->max_y; }

// ORBmatcher.cc:52-73 + RadiusByViewingCos :216-222
void prep_mappoint_queries(HostQueries& q, int nmp, const uint8_t* inView, const float* projX, const float* projY, const int* level,
                           const float* viewCos, const uint8_t* mpDesc, const float* projXR, float th, const float* scaleFactors,
                           int nlevels, bool haveRight) {
  q.nq = nmp; q.valid.resize(nmp); q.x.resize(nmp); q.y.resize(nmp); q.r.resize(nmp); q.minL.resize(nmp); q.maxL.resize(nmp);
  q.desc = mpDesc;
  q.xrCheck = haveRight && projXR;
  if (q.xrCheck) q.xr.assign(projXR, projXR + nmp);
  const bool bFactor = th != 1.0;
  for (int i = 0; i < nmp; i++) {
    const int lvl = level[i];
    const bool ok = inView[i] != 0 && lvl >= 0 && lvl < nlevels;
    q.valid[i] = ok;
    float r = (viewCos[i] > 0.998) ? 2.5f : 4.0f;
    if (bFactor) r *= th;
    q.x[i] = projX[i]; q.y[i] = projY[i];
    q.r[i] = ok ? r * scaleFactors[lvl] : 0.f;
    q.minL[i] = (int8_t)(lvl - 1); q.maxL[i] = (int8_t)lvl;
  }
}

}  // namespace

extern "C" {

int orbx_matcher_create(int device, orbx_matcher** out) {
  if (!out) return ORBX_ERR_ARG;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
    set_error("orbx_matcher_create: no usable CUDA device %d (this library has no CPU fallback)", device);
    return ORBX_ERR_CUDA;
  }
  CK(cudaSetDevice(device));
  orbx_matcher* m = new orbx_matcher();
  m->device = device;
  bool evOk = true;
  for (int i = 0; i < 2; i++)
    for (cudaEvent_t* e : {&m->evUpL[i], &m->evUpR[i], &m->evJoin[i], &m->evInFree[i], &m->evSlotDone[i]})
      evOk = evOk && cudaEventCreateWithFlags(e, cudaEventDisableTiming) == cudaSuccess;
  evOk = evOk && cudaStreamCreateWithFlags(&m->upStream, cudaStreamNonBlocking) == cudaSuccess &&
         cudaHostAlloc(&m->hStat, 2 * orbx_matcher::kStatChunks * sizeof(int), cudaHostAllocDefault) == cudaSuccess;
  if (!evOk || cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking) != cudaSuccess || cudaMalloc(&m->dStatus, sizeof(int)) != cudaSuccess) {
    delete m; set_error("matcher resources"); return ORBX_ERR_CUDA;
  }
  *out = m;
  return ORBX_OK;
}

void orbx_matcher_destroy(orbx_matcher* m) {
  if (!m) return;
  cudaSetDevice(m->device);
  cudaStreamSynchronize(m->stream);
  void* ptrs[] = {m->arena, m->tcWs, m->dStatus, m->sURight, m->sDepth, m->sBestDist, m->sBestIdx, m->sSad, m->sRowStart, m->sRowItems};
  for (void* p : ptrs) if (p) cudaFree(p);
  for (auto& slot : m->chunkGraph) for (auto& cg : slot) cg.reset();
  for (int i = 0; i < 2; i++)
    for (cudaEvent_t e : {m->evUpL[i], m->evUpR[i], m->evJoin[i], m->evInFree[i], m->evSlotDone[i]}) if (e) cudaEventDestroy(e);
  if (m->upStream) cudaStreamDestroy(m->upStream);
  if (m->hStat) cudaFreeHost(m->hStat);
  if (m->hPin) cudaFreeHost(m->hPin);
  if (m->dStage) cudaFree(m->dStage);
  cudaStreamDestroy(m->stream);
  delete m;
}

int orbx_hamming_batch(orbx_matcher* m, const uint8_t* a, const uint8_t* b, int n, int* dist) {
  if (!m || n < 0 || (n > 0 && (!a || !b || !dist))) return ORBX_ERR_ARG;
  if (n == 0) return ORBX_OK;
  CK(cudaSetDevice(m->device));
  ArenaPlan plan;
  const size_t oa = plan.add((size_t)n * 32), ob = plan.add((size_t)n * 32), od = plan.add((size_t)n * 4);
  int rc = arena_ensure(m, plan.total);
  if (rc != ORBX_OK) return rc;
  CK(cudaMemcpyAsync(m->arena + oa, a, (size_t)n * 32, cudaMemcpyHostToDevice, m->stream));
  CK(cudaMemcpyAsync(m->arena + ob, b, (size_t)n * 32, cudaMemcpyHostToDevice, m->stream));
  k_hamming_pairs<<<(n + 255) / 256, 256, 0, m->stream>>>(m->arena + oa, m->arena + ob, n, at<int>(m, od));
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(dist, m->arena + od, (size_t)n * 4, cudaMemcpyDeviceToHost, m->stream));
  CK(cudaStreamSynchronize(m->stream));
  return ORBX_OK;
}

int orbx_features_in_area(orbx_matcher* m, const orbx_frame_view* F, float x, float y, float r, int minLevel, int maxLevel, int* indices,
                          int capacity, int* n) {
  if (!m || !n) return ORBX_ERR_ARG;
  int rc = check_view(F, false);
  if (rc != ORBX_OK) return rc;
  std::vector<WindowJob> jobs(1);
  WindowJob& J = jobs[0];
  J.F = F; set_bounds(J, F);
  static const uint8_t zeros[32] = {0};
  J.q.nq = 1; J.q.valid = {1}; J.q.x = {x}; J.q.y = {y}; J.q.r = {r};
  J.q.minL = {(int8_t)std::max(-128, std::min(127, minLevel))}; J.q.maxL = {(int8_t)std::max(-128, std::min(127, maxLevel))};
  J.q.desc = zeros;
  std::vector<unsigned long long> entries;
  J.entriesOut = &entries;
  rc = run_window_jobs(m, jobs, RESOLVE_NONE);
  if (rc != ORBX_OK) return rc;
  // order of the reference's vIndices: grid column, grid row, insertion (= keypoint index)
  std::sort(entries.begin(), entries.end(), [](unsigned long long a, unsigned long long b) { return (a & 0xfffffffffffull) < (b & 0xfffffffffffull); });
  *n = (int)entries.size();
  for (int i = 0; i < (int)entries.size() && i < capacity; i++) indices[i] = (int)((entries[i] >> 8) & 0xffffff);
  return ORBX_OK;
}

int orbx_search_by_projection_mappoints(orbx_matcher* m, const orbx_frame_view* F, int nmp, const uint8_t* inView, const float* projX,
                                        const float* projY, const int* level, const float* viewCos, const uint8_t* mpDesc,
                                        const uint8_t* observed, const float* projXR, float th, float nnratio, int* kpMp, int* nmatches) {
  if (!m || nmp < 0 || !kpMp || !nmatches || (nmp > 0 && (!inView || !projX || !projY || !level || !viewCos || !mpDesc))) return ORBX_ERR_ARG;
  int rc = check_view(F, true);
  if (rc != ORBX_OK) return rc;
  std::vector<WindowJob> jobs(1);
  WindowJob& J = jobs[0];
  J.F = F; set_bounds(J, F);
  prep_mappoint_queries(J.q, nmp, inView, projX, projY, level, viewCos, mpDesc, projXR, th, F->scale_factors, F->nlevels, F->u_right != nullptr);
  J.observed = observed; J.kpMp = kpMp; J.kpMpLen = F->n; J.nnratio = nnratio; J.nmatches = nmatches;
  return run_window_jobs(m, jobs, RESOLVE_MAPPOINTS);
}

int orbx_search_by_projection_mappoints_stereo(orbx_matcher* m, const orbx_frame_view* left, const orbx_frame_view* right, int nmp,
                                               const orbx_mappoint_track* trackLeft, const orbx_mappoint_track* trackRight, const uint8_t* mpDesc,
                                               const uint8_t* observed, const int* leftToRight, const int* rightToLeft, float th, float nnratio,
                                               int* kpMp, int* nmatches) {
  if (!m || nmp < 0 || !kpMp || !nmatches || !trackLeft || !trackRight || !left || !right) return ORBX_ERR_ARG;
  for (const orbx_mappoint_track* t : {trackLeft, trackRight})
    if (nmp > 0 && (!t->in_view || !t->proj_x || !t->proj_y || !t->level || !t->view_cos)) return ORBX_ERR_ARG;
  if (nmp > 0 && !mpDesc) return ORBX_ERR_ARG;
  if ((left->n > 0 && !leftToRight) || (right->n > 0 && !rightToLeft)) return ORBX_ERR_ARG;
  int rc = check_view(left, true);
  if (rc != ORBX_OK) return rc;
  if ((rc = check_view(right, true)) != ORBX_OK) return rc;
  if (left->n + right->n > 2 * kMaxKpSmem) { set_error("more than %d keypoints in the two images", 2 * kMaxKpSmem); return ORBX_ERR_CAPACITY; }
  for (int k = 0; k < left->n; k++)
    if (leftToRight[k] < -1 || leftToRight[k] >= right->n) { set_error("leftToRight[%d] out of range", k); return ORBX_ERR_ARG; }
  for (int k = 0; k < right->n; k++)
    if (rightToLeft[k] < -1 || rightToLeft[k] >= left->n) { set_error("rightToLeft[%d] out of range", k); return ORBX_ERR_ARG; }
  std::vector<WindowJob> jobs(2);
  const orbx_frame_view* views[2] = {left, right};
  const orbx_mappoint_track* tracks[2] = {trackLeft, trackRight};
  for (int s = 0; s < 2; s++) {
    WindowJob& J = jobs[s];
    J.F = views[s]; set_bounds(J, views[s]);
    // the left window is scaled by th (ORBmatcher.cc:70-71), the right one is not (:147-150); no mvuRight check on such frames (:92)
    prep_mappoint_queries(J.q, nmp, tracks[s]->in_view, tracks[s]->proj_x, tracks[s]->proj_y, tracks[s]->level, tracks[s]->view_cos, mpDesc, nullptr,
                          s == 0 ? th : 1.0f, views[s]->scale_factors, views[s]->nlevels, false);
    J.observed = observed; J.nnratio = nnratio; J.nmatches = s == 0 ? nmatches : nullptr;
    J.kpMp = kpMp + (s == 0 ? 0 : left->n); J.kpMpLen = views[s]->n;
    J.partner = s == 0 ? leftToRight : rightToLeft;
  }
  return run_window_jobs(m, jobs, RESOLVE_FISHEYE);
}

int orbx_search_by_projection_mappoints_batch(orbx_matcher* m, orbx_extractor* ex, int nframes, int nmp, const uint8_t* inView,
                                              const float* projX, const float* projY, const int* level, const float* viewCos,
                                              const uint8_t* mpDesc, const uint8_t* observed, float th, float nnratio, int* kpMp,
                                              int capPerFrame, int* nmatches) {
  if (!m || !ex || !ex->haveGeom || nframes < 1 || nmp < 0 || !kpMp || !nmatches) return ORBX_ERR_ARG;
  Slot& s = ex->slot[ex->lastSlot];
  if (nframes > s.lastFrames) { set_error("extractor holds %d frames", s.lastFrames); return ORBX_ERR_ARG; }
  CK(cudaSetDevice(m->device));
  CK(cudaDeviceSynchronize());  // the extractor batch may still be in flight on another stream
  std::vector<int> counts(nframes);
  CK(cudaMemcpy(counts.data(), s.counts, nframes * sizeof(int), cudaMemcpyDeviceToHost));
  std::vector<WindowJob> jobs(nframes);
  for (int f = 0; f < nframes; f++) {
    WindowJob& J = jobs[f];
    if (counts[f] > capPerFrame || counts[f] > kMaxKpSmem) { set_error("frame %d: %d keypoints exceed the capacity", f, counts[f]); return ORBX_ERR_CAPACITY; }
    J.devKps = s.kps + (size_t)f * ex->g.kpStride; J.devDesc = s.desc + (size_t)f * ex->g.kpStride * 32; J.devN = counts[f];
    J.minX = 0.f; J.minY = 0.f; J.maxX = (float)ex->g.width; J.maxY = (float)ex->g.height;  // Frame.cc:894-897 (no distortion)
    const size_t o = (size_t)f * nmp;
    prep_mappoint_queries(J.q, nmp, inView + o, projX + o, projY + o, level + o, viewCos + o, mpDesc + o * 32, nullptr, th, ex->scale.data(),
                          ex->p.nlevels, false);
    J.observed = observed ? observed + o : nullptr;
    J.kpMp = kpMp + (size_t)f * capPerFrame; J.kpMpLen = counts[f]; J.nnratio = nnratio; J.nmatches = nmatches + f;
  }
  return run_window_jobs(m, jobs, RESOLVE_MAPPOINTS);
}

// ---- Tracking::SearchLocalPoints over a batch of frames, device-resident ---------------------------
// The local map (positions, normals, distance ranges, descriptors) is uploaded once; per frame only its pose travels.
// k_local_point_queries runs Frame::isInFrustum for every (frame, map point) pair and writes the window queries of
// SearchByProjection straight into the arena the candidate / resolve kernels read -- no per-frame host staging, one
// synchronisation per batch.
}  // extern "C"

struct orbx_local_map {
  int device = 0, nmp = 0;
  float* pos = nullptr; float* normal = nullptr; float* maxDist = nullptr; float* minDist = nullptr;
  uint8_t* desc = nullptr; uint8_t* observed = nullptr;
};

namespace {
struct LocalPointArgs {
  const orbx_frustum* cams; const float* pos; const float* normal; const float* maxDist; const float* minDist;
  int nmp, nlevels, kpStride, cap;
  float th; int bFactor;
  const float* scale;
  GridFrame* frames; const QuerySet* sets; const int* counts; int* status;
};

// grid (ceil(nmp / 256), nframes): ORBmatcher.cc:52-73 (radius by viewing cosine, predicted level window) on top of isInFrustum
__global__ void __launch_bounds__(256) k_local_point_queries(const __grid_constant__ LocalPointArgs A) {
  __shared__ orbx_frustum cam;
  const int f = blockIdx.y;
  for (int i = threadIdx.x; i < (int)(sizeof(orbx_frustum) / 4); i += 256) reinterpret_cast<int*>(&cam)[i] = reinterpret_cast<const int*>(A.cams + f)[i];
  __syncthreads();
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    const int n = A.counts[f];
    if (n > A.cap || n > kMaxKpSmem) atomicExch(A.status, ORBX_ERR_CAPACITY);
    A.frames[f].n = min(n, min(A.cap, kMaxKpSmem));
  }
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= A.nmp) return;
  const orbx::FrustumOut o = orbx::frustum_point(cam, A.pos[3 * i], A.pos[3 * i + 1], A.pos[3 * i + 2], A.normal[3 * i], A.normal[3 * i + 1],
                                                 A.normal[3 * i + 2], A.maxDist[i], A.minDist[i]);
  const QuerySet Q = A.sets[f];
  const int lvl = o.level;
  const bool ok = o.inView != 0 && lvl >= 0 && lvl < A.nlevels;
  float r = (o.viewCos > 0.998) ? 2.5f : 4.0f;
  if (A.bFactor) r = __fmul_rn(r, A.th);
  const_cast<uint8_t*>(Q.valid)[i] = ok ? 1 : 0;
  const_cast<float*>(Q.x)[i] = o.projX; const_cast<float*>(Q.y)[i] = o.projY;
  const_cast<float*>(Q.r)[i] = ok ? __fmul_rn(r, A.scale[lvl]) : 0.f;
  const_cast<int8_t*>(Q.minL)[i] = (int8_t)(lvl - 1); const_cast<int8_t*>(Q.maxL)[i] = (int8_t)lvl;
}
}  // namespace

extern "C" {

int orbx_local_map_create(orbx_matcher* m, int nmp, const float* world_pos, const float* normal, const float* max_dist,
                          const float* min_dist, const uint8_t* desc, const uint8_t* observed, orbx_local_map** out) {
  if (!m || !out || nmp < 1 || !world_pos || !normal || !max_dist || !min_dist || !desc) { set_error("local map: bad argument"); return ORBX_ERR_ARG; }
  *out = nullptr;
  std::lock_guard<std::recursive_mutex> setupLock(setup_mutex());
  CK(cudaSetDevice(m->device));
  orbx_local_map* L = new (std::nothrow) orbx_local_map();
  if (!L) return ORBX_ERR_ARG;
  L->device = m->device; L->nmp = nmp;
  auto put = [&](void** d, const void* h, size_t bytes) -> cudaError_t {
    cudaError_t e = cudaMalloc(d, bytes);
    return e != cudaSuccess ? e : cudaMemcpy(*d, h, bytes, cudaMemcpyHostToDevice);
  };
  cudaError_t e = put((void**)&L->pos, world_pos, (size_t)nmp * 12);
  if (e == cudaSuccess) e = put((void**)&L->normal, normal, (size_t)nmp * 12);
  if (e == cudaSuccess) e = put((void**)&L->maxDist, max_dist, (size_t)nmp * 4);
  if (e == cudaSuccess) e = put((void**)&L->minDist, min_dist, (size_t)nmp * 4);
  if (e == cudaSuccess) e = put((void**)&L->desc, desc, (size_t)nmp * 32);
  if (e == cudaSuccess && observed) e = put((void**)&L->observed, observed, (size_t)nmp);
  if (e != cudaSuccess) {
    set_error("local map upload: %s", cudaGetErrorString(e));
    orbx_local_map_destroy(L);
    return ORBX_ERR_CUDA;
  }
  *out = L;
  return ORBX_OK;
}

void orbx_local_map_destroy(orbx_local_map* L) {
  if (!L) return;
  std::lock_guard<std::recursive_mutex> setupLock(setup_mutex());
  cudaSetDevice(L->device);
  cudaFree(L->pos); cudaFree(L->normal); cudaFree(L->maxDist); cudaFree(L->minDist); cudaFree(L->desc); cudaFree(L->observed);
  delete L;
}

int orbx_search_local_points_batch(orbx_matcher* m, orbx_extractor* ex, int nframes, const orbx_frustum* frames, const orbx_local_map* map,
                                   float th, float nnratio, int* kpMp, int capPerFrame, int* nmatches, void* stream) {
  if (!m || !ex || !ex->haveGeom || nframes < 1 || !frames || !map || !kpMp || !nmatches || capPerFrame < 1) { set_error("search_local_points: bad argument"); return ORBX_ERR_ARG; }
  if (map->device != m->device) { set_error("local map lives on another device"); return ORBX_ERR_ARG; }
  Slot& s = ex->slot[ex->lastSlot];
  if (nframes > s.lastFrames) { set_error("extractor holds %d frames", s.lastFrames); return ORBX_ERR_ARG; }
  CK(cudaSetDevice(m->device));
  const int nq = map->nmp, kpStride = ex->g.kpStride, nlevels = ex->p.nlevels;
  const int nmax = std::min(kpStride, kMaxKpSmem);
  // ---- arena: per-frame regions are laid out frame after frame so that the clears and the downloads are single calls
  ArenaPlan plan;
  const size_t szCell = ((kGridCells + 1) * sizeof(int) + 255) & ~(size_t)255, szItems = ((size_t)nmax * 4 + 255) & ~(size_t)255;
  const size_t szQ1 = ((size_t)nq + 255) & ~(size_t)255, szQ4 = ((size_t)nq * 4 + 255) & ~(size_t)255;
  const int entryCap = (int)std::min<long long>((long long)nq * nmax, std::max<long long>(1 << 16, (long long)nq * 256));
  const size_t szEntries = (size_t)entryCap * 8;
  const size_t oCell = plan.add(szCell * nframes), oItems = plan.add(szItems * nframes);
  const size_t oValid = plan.add(szQ1 * nframes), oMinL = plan.add(szQ1 * nframes), oMaxL = plan.add(szQ1 * nframes);
  const size_t oX = plan.add(szQ4 * nframes), oY = plan.add(szQ4 * nframes), oR = plan.add(szQ4 * nframes);
  const size_t oQOff = plan.add(szQ4 * nframes), oQCnt = plan.add(szQ4 * nframes), oAssigned = plan.add(szQ4 * nframes);
  const size_t oEntries = plan.add(szEntries * nframes), oCursor = plan.add((size_t)nframes * 4);
  const size_t oKpMp = plan.add((size_t)nframes * kpStride * 4), oNm = plan.add((size_t)nframes * 4);
  const size_t oBlk0 = plan.add(szItems * nframes), oBlk1 = plan.add(szItems * nframes), oChg = plan.add((size_t)nframes * 8);
  const size_t oScale = plan.add((size_t)nlevels * 4), oCams = plan.add((size_t)nframes * sizeof(orbx_frustum));
  const size_t oFrames = plan.add(nframes * sizeof(GridFrame)), oSets = plan.add(nframes * sizeof(QuerySet)), oArgs = plan.add(nframes * sizeof(ResolveArgs));
  int rc = arena_ensure(m, plan.total);
  if (rc != ORBX_OK) return rc;
  cudaStream_t st = stream ? (cudaStream_t)stream : m->stream;
  if (!stream) CK(cudaDeviceSynchronize());  // the extractor batch may be in flight on another stream
  std::vector<GridFrame> G(nframes);
  std::vector<QuerySet> Q(nframes);
  std::vector<ResolveArgs> R(nframes);
  for (int f = 0; f < nframes; f++) {
    const orbx_frustum& C = frames[f];
    if (!(C.max_x > C.min_x) || !(C.max_y > C.min_y)) { set_error("frame %d: empty image bounds", f); return ORBX_ERR_ARG; }
    G[f].kps = s.kps + (size_t)f * kpStride; G[f].desc = s.desc + (size_t)f * kpStride * 32; G[f].uRight = nullptr; G[f].n = 0;  // n: device
    G[f].minX = C.min_x; G[f].minY = C.min_y;
    G[f].invW = (float)kGridCols / (C.max_x - C.min_x); G[f].invH = (float)kGridRows / (C.max_y - C.min_y);  // Frame.cc:160-161
    G[f].cellStart = at<int>(m, oCell + f * szCell); G[f].cellItems = at<int>(m, oItems + f * szItems);
    QuerySet& q = Q[f];
    q = QuerySet{};
    q.nq = nq;
    q.valid = at<uint8_t>(m, oValid + f * szQ1); q.minL = at<int8_t>(m, oMinL + f * szQ1); q.maxL = at<int8_t>(m, oMaxL + f * szQ1);
    q.x = at<float>(m, oX + f * szQ4); q.y = at<float>(m, oY + f * szQ4); q.r = at<float>(m, oR + f * szQ4);
    q.xr = nullptr; q.desc = map->desc; q.xrCheck = 0; q.gate = 0; q.invSigma2 = nullptr;
    q.qOff = at<int>(m, oQOff + f * szQ4); q.qCnt = at<int>(m, oQCnt + f * szQ4);
    q.entries = at<unsigned long long>(m, oEntries + f * szEntries); q.entryCap = entryCap; q.entryCursor = at<int>(m, oCursor + (size_t)f * 4);
    ResolveArgs& a = R[f];
    a = ResolveArgs{};
    a.observed = map->observed; a.angle = nullptr;
    a.kpMp = at<int>(m, oKpMp + (size_t)f * kpStride * 4); a.assigned = at<int>(m, oAssigned + f * szQ4);
    a.nmatches = at<int>(m, oNm + (size_t)f * 4);
    a.blk[0] = at<int>(m, oBlk0 + f * szItems); a.blk[1] = at<int>(m, oBlk1 + f * szItems); a.chg = at<int>(m, oChg + (size_t)f * 8);
    a.nnratio = nnratio; a.distTh = 100; a.anyMpBlocks = 0; a.checkOrientation = 0;
  }
  CK(cudaMemcpyAsync(m->arena + oFrames, G.data(), nframes * sizeof(GridFrame), cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(m->arena + oSets, Q.data(), nframes * sizeof(QuerySet), cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(m->arena + oArgs, R.data(), nframes * sizeof(ResolveArgs), cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(m->arena + oCams, frames, nframes * sizeof(orbx_frustum), cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(m->arena + oScale, ex->scale.data(), (size_t)nlevels * 4, cudaMemcpyHostToDevice, st));
  CK(cudaMemsetAsync(m->arena + oCursor, 0, (size_t)nframes * 4, st));
  CK(cudaMemsetAsync(m->arena + oKpMp, 0xff, (size_t)nframes * kpStride * 4, st));  // mvpMapPoints all empty (-1)
  CK(cudaMemsetAsync(m->dStatus, 0, sizeof(int), st));
  LocalPointArgs A{};
  A.cams = at<orbx_frustum>(m, oCams); A.pos = map->pos; A.normal = map->normal; A.maxDist = map->maxDist; A.minDist = map->minDist;
  A.nmp = nq; A.nlevels = nlevels; A.kpStride = kpStride; A.cap = std::min(capPerFrame, nmax);
  A.th = th; A.bFactor = th != 1.0f; A.scale = at<float>(m, oScale);
  A.frames = at<GridFrame>(m, oFrames); A.sets = at<QuerySet>(m, oSets); A.counts = s.counts; A.status = m->dStatus;
  k_local_point_queries<<<dim3((nq + 255) / 256, nframes), 256, 0, st>>>(A);
  k_build_grid<<<nframes, 256, 0, st>>>(A.frames);
  k_window_candidates<<<dim3((nq + 7) / 8, nframes), 256, 0, st>>>(A.frames, A.sets, m->dStatus);
  k_resolve_par<<<dim3(kParCluster, nframes), kParThreads, 0, st>>>(A.frames, A.sets, at<ResolveArgs>(m, oArgs), 0);
  CK(cudaGetLastError());
  int status = 0;
  CK(cudaMemcpyAsync(&status, m->dStatus, sizeof(int), cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpy2DAsync(kpMp, (size_t)capPerFrame * 4, m->arena + oKpMp, (size_t)kpStride * 4, (size_t)std::min(capPerFrame, kpStride) * 4, nframes,
                       cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(nmatches, m->arena + oNm, (size_t)nframes * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  if (status != 0) { set_error(status == ORBX_ERR_CAPACITY ? "a frame's keypoints or candidate list exceed the capacity" : "search_local_points failed"); return status; }
  return ORBX_OK;
}

int orbx_search_by_projection_window(orbx_matcher* m, const orbx_frame_view* F, int n, const uint8_t* valid, const float* u, const float* v,
                                     const int* level, const float* ur, const float* angle, const uint8_t* srcDesc, const uint8_t* observed,
                                     float th, int mode, int distTh, int anyMpBlocks, int checkOrientation, int* kpMp, int* nmatches) {
  if (!m || n < 0 || !kpMp || !nmatches || (n > 0 && (!valid || !u || !v || !level || !srcDesc)) || (checkOrientation && n > 0 && !angle)) return ORBX_ERR_ARG;
  int rc = check_view(F, true);
  if (rc != ORBX_OK) return rc;
  std::vector<WindowJob> jobs(1);
  WindowJob& J = jobs[0];
  J.F = F; set_bounds(J, F);
  HostQueries& q = J.q;
  q.nq = n; q.valid.resize(n); q.x.assign(u, u + n); q.y.assign(v, v + n); q.r.resize(n); q.minL.resize(n); q.maxL.resize(n);
  q.desc = srcDesc;
  q.xrCheck = ur != nullptr && F->u_right != nullptr;
  if (q.xrCheck) q.xr.assign(ur, ur + n);
  for (int i = 0; i < n; i++) {  // ORBmatcher.cc:2105-2118 / :2340-2342
    const int l = level[i];
    const bool ok = valid[i] != 0 && l >= 0 && l < F->nlevels;
    q.valid[i] = ok;
    q.r[i] = ok ? th * F->scale_factors[l] : 0.f;
    if (mode == 1) { q.minL[i] = (int8_t)l; q.maxL[i] = -1; }
    else if (mode == 2) { q.minL[i] = 0; q.maxL[i] = (int8_t)l; }
    else if (mode == 3) { q.minL[i] = (int8_t)(l - 1); q.maxL[i] = (int8_t)l; }   // ORBmatcher.cc:573-576, 689-692
    else { q.minL[i] = (int8_t)(l - 1); q.maxL[i] = (int8_t)(l + 1); }
  }
  J.observed = observed; J.angle = angle; J.kpMp = kpMp; J.kpMpLen = F->n; J.distTh = distTh; J.anyMpBlocks = anyMpBlocks;
  J.checkOrientation = checkOrientation; J.nmatches = nmatches;
  return run_window_jobs(m, jobs, RESOLVE_WINDOW);
}

int orbx_search_window_best(orbx_matcher* m, const orbx_frame_view* F, int n, const uint8_t* valid, const float* u, const float* v,
                            const float* radius, const int* level, const float* ur, const uint8_t* desc, int gate,
                            const float* invLevelSigma2, int* bestIdx, int* bestDist) {
  if (!m || n < 0 || (n > 0 && (!valid || !u || !v || !radius || !level || !desc || !bestIdx || !bestDist))) return ORBX_ERR_ARG;
  int rc = check_view(F, false);
  if (rc != ORBX_OK) return rc;
  if (gate && (!invLevelSigma2 || F->nlevels < 1 || (F->u_right && !ur))) { set_error("chi-square gate needs inv_level_sigma2 (and ur for stereo keyframes)"); return ORBX_ERR_ARG; }
  if (n == 0) return ORBX_OK;
  std::vector<WindowJob> jobs(1);
  WindowJob& J = jobs[0];
  J.F = F; set_bounds(J, F);
  HostQueries& q = J.q;
  q.nq = n; q.valid.assign(valid, valid + n); q.x.assign(u, u + n); q.y.assign(v, v + n); q.r.assign(radius, radius + n);
  q.minL.resize(n); q.maxL.resize(n);
  q.desc = desc;
  q.gate = gate != 0; q.invSigma2 = invLevelSigma2; q.nSigma = gate ? F->nlevels : 0;
  if (q.gate && ur) q.xr.assign(ur, ur + n);
  for (int i = 0; i < n; i++) {  // ORBmatcher.cc:1535-1537: kpLevel in [nPredictedLevel-1, nPredictedLevel]
    const int l = level[i];
    if (l < 0 || l > 127) q.valid[i] = 0;
    q.minL[i] = (int8_t)(l - 1); q.maxL[i] = (int8_t)l;
  }
  J.bestIdx = bestIdx; J.bestDist = bestDist;
  return run_window_jobs(m, jobs, RESOLVE_BEST);
}

static int check_projection(const orbx_frame_view* F, const orbx_projection* P, const orbx_projection_points* pts) {
  if (!P || !pts || pts->n < 0) return ORBX_ERR_ARG;
  if (pts->n > 0 && (!pts->world_pos || !pts->desc)) { set_error("projection: world_pos / desc missing"); return ORBX_ERR_ARG; }
  if (P->level_mode == 0 && pts->n > 0 && !pts->octave) { set_error("projection: level_mode 0 needs the source octaves"); return ORBX_ERR_ARG; }
  if ((P->range_mode || P->level_mode == 1) && pts->n > 0 && (!pts->max_distance || !pts->min_distance)) {
    set_error("projection: the distance range / PredictScale need mfMaxDistance and mfMinDistance"); return ORBX_ERR_ARG;
  }
  if (P->normal_check && pts->n > 0 && !pts->normal) { set_error("projection: the viewing-angle test needs the normals"); return ORBX_ERR_ARG; }
  if (P->level_mode == 1 && (!(P->log_scale_factor > 0.f) || P->n_scale_levels < 1)) { set_error("projection: PredictScale needs log_scale_factor / n_scale_levels"); return ORBX_ERR_ARG; }
  if (F->nlevels < 1 || !F->scale_factors) { set_error("projection: the frame view needs its scale factors"); return ORBX_ERR_ARG; }
  if (P->camera_model != ORBX_CAMERA_PINHOLE && P->camera_model != ORBX_CAMERA_KB8) return ORBX_ERR_ARG;
  return ORBX_OK;
}

int orbx_search_projected_window(orbx_matcher* m, const orbx_frame_view* F, const orbx_projection* P, const orbx_projection_points* pts,
                                 const float* angle, const uint8_t* observed, int distTh, int anyMpBlocks, int checkOrientation, int* kpMp,
                                 int* nmatches, const orbx_projection_debug* dbg) {
  if (!m || !kpMp || !nmatches) return ORBX_ERR_ARG;
  int rc = check_view(F, true);
  if (rc != ORBX_OK) return rc;
  if ((rc = check_projection(F, P, pts)) != ORBX_OK) return rc;
  if (checkOrientation && pts->n > 0 && !angle) return ORBX_ERR_ARG;
  std::vector<WindowJob> jobs(1);
  WindowJob& J = jobs[0];
  J.F = F; set_bounds(J, F);
  J.proj = P; J.pts = pts; J.dbg = dbg;
  J.q.nq = pts->n; J.q.desc = pts->desc;
  J.q.xrCheck = P->want_xr != 0 && F->u_right != nullptr;
  J.observed = observed; J.angle = angle; J.kpMp = kpMp; J.kpMpLen = F->n;
  J.nnratio = 0.f; J.distTh = distTh; J.anyMpBlocks = anyMpBlocks; J.checkOrientation = checkOrientation;
  J.nmatches = nmatches;
  *nmatches = 0;
  return run_window_jobs(m, jobs, RESOLVE_WINDOW);
}

int orbx_search_projected_window_stereo(orbx_matcher* m, const orbx_frame_view* left, const orbx_frame_view* right, const orbx_projection* projLeft,
                                        const orbx_projection* projRight, const orbx_projection_points* pts, const float* angle,
                                        const uint8_t* observed, int distTh, int checkOrientation, int* kpMp, int* nmatches) {
  if (!m || !kpMp || !nmatches || !left || !right) return ORBX_ERR_ARG;
  const orbx_frame_view* views[2] = {left, right};
  const orbx_projection* projs[2] = {projLeft, projRight};
  for (int s = 0; s < 2; s++) {
    int rc = check_view(views[s], true);
    if (rc != ORBX_OK) return rc;
    if ((rc = check_projection(views[s], projs[s], pts)) != ORBX_OK) return rc;
  }
  if (checkOrientation && pts->n > 0 && !angle) return ORBX_ERR_ARG;
  std::vector<WindowJob> jobs(2);
  for (int s = 0; s < 2; s++) {
    WindowJob& J = jobs[s];
    J.F = views[s]; set_bounds(J, views[s]);
    J.proj = projs[s]; J.pts = pts;
    J.q.nq = pts->n; J.q.desc = pts->desc;
    J.observed = observed; J.angle = angle;
    J.kpMp = kpMp + (s == 0 ? 0 : left->n); J.kpMpLen = views[s]->n;
    J.distTh = distTh; J.anyMpBlocks = 0; J.checkOrientation = checkOrientation;
    J.nmatches = s == 0 ? nmatches : nullptr;
    J.gateJob = s == 0 ? -1 : 0;
  }
  *nmatches = 0;
  return run_window_jobs(m, jobs, RESOLVE_WINDOW_PAIR);
}

int orbx_search_projected_best(orbx_matcher* m, const orbx_frame_view* F, const orbx_projection* P, const orbx_projection_points* pts, int gate,
                               const float* invLevelSigma2, int* bestIdx, int* bestDist, const orbx_projection_debug* dbg) {
  if (!m || !bestIdx || !bestDist) return ORBX_ERR_ARG;
  int rc = check_view(F, false);
  if (rc != ORBX_OK) return rc;
  if ((rc = check_projection(F, P, pts)) != ORBX_OK) return rc;
  if (gate && (!invLevelSigma2 || (F->u_right && !P->want_xr))) { set_error("chi-square gate needs inv_level_sigma2 (and want_xr for stereo keyframes)"); return ORBX_ERR_ARG; }
  if (pts->n == 0) return ORBX_OK;
  std::vector<WindowJob> jobs(1);
  WindowJob& J = jobs[0];
  J.F = F; set_bounds(J, F);
  J.proj = P; J.pts = pts; J.dbg = dbg;
  J.q.nq = pts->n; J.q.desc = pts->desc;
  J.q.gate = gate != 0; J.q.invSigma2 = invLevelSigma2; J.q.nSigma = gate ? F->nlevels : 0;
  J.bestIdx = bestIdx; J.bestDist = bestDist;
  return run_window_jobs(m, jobs, RESOLVE_BEST);
}

namespace {
// Common node ids of two FeatureVectors in ascending order, as the reference's merge loop visits them (ORBmatcher.cc:289-299,
// 440-451): one work item per (side-1 list position, side-2 node range).
int bow_check(const orbx_featvec* fv, int n) {
  if (!fv || fv->nnodes < 0 || (fv->nnodes > 0 && (!fv->node_ids || !fv->node_start || !fv->indices))) return ORBX_ERR_ARG;
  for (int i = 0; i < fv->nnodes; i++) {
    if (fv->node_start[i + 1] < fv->node_start[i]) return ORBX_ERR_ARG;
    if (i > 0 && fv->node_ids[i] <= fv->node_ids[i - 1]) return ORBX_ERR_ARG;
  }
  (void)n;
  return ORBX_OK;
}

struct BowProblem {
  int n1, n2;
  const orbx_keypoint* kps1; const uint8_t* desc1; const orbx_featvec* fv1; const uint8_t* elig1; const uint8_t* stereo1;
  const orbx_keypoint* kps2; const uint8_t* desc2; const orbx_featvec* fv2; const uint8_t* elig2; const uint8_t* stereo2;
  int maxDist; float nnratio; int checkOrientation;
  bool triangulate; const float* F12; float epx, epy; const float* scaleFactors2; const float* levelSigma2; int nlevels; int coarse;
  int nLeft2 = -1; int* matches12R = nullptr;   // side 2 is a fisheye-stereo frame: features >= nLeft2 are the right image's
  // KannalaBrandt8 gate of SearchForTriangulation (kb8 != null): pair records, left counts of both keyframes, pKF1's sigma table
  const orbx_kb8_pair* kb8 = nullptr; int triLeft1 = -1, triLeft2 = -1; const float* levelSigma1 = nullptr;
};

int run_bow(orbx_matcher* m, const BowProblem& P, int* matches12, int* nmatches) {
  if (!m || !matches12 || !nmatches || P.n1 < 0 || P.n2 < 0) return ORBX_ERR_ARG;
  if ((P.n1 > 0 && (!P.kps1 || !P.desc1)) || (P.n2 > 0 && (!P.kps2 || !P.desc2))) return ORBX_ERR_ARG;
  if (bow_check(P.fv1, P.n1) != ORBX_OK || bow_check(P.fv2, P.n2) != ORBX_OK) { set_error("bad feature vector (node ids must ascend)"); return ORBX_ERR_ARG; }
  if (P.n1 >= (1 << 20) || P.n2 >= (1 << 20)) { set_error("more than 2^20 features per keyframe is unsupported"); return ORBX_ERR_UNSUPPORTED; }
  const bool fish = P.nLeft2 >= 0;
  if (fish && (!P.matches12R || P.nLeft2 > P.n2 || P.triangulate)) return ORBX_ERR_ARG;
  *nmatches = 0;
  for (int i = 0; i < P.n1; i++) { matches12[i] = -1; if (fish) P.matches12R[i] = -1; }
  if (P.n1 == 0 || P.n2 == 0) return ORBX_OK;
  CK(cudaSetDevice(m->device));
  std::vector<BowWork> work;
  long long nEntries = 0;
  {
    const orbx_featvec &a = *P.fv1, &b = *P.fv2;
    int ia = 0, ib = 0;
    while (ia < a.nnodes && ib < b.nnodes) {
      if (a.node_ids[ia] == b.node_ids[ib]) {
        const int b0 = b.node_start[ib], b1 = b.node_start[ib + 1];
        if (b1 - b0 >= (1 << 20)) { set_error("vocabulary node with 2^20 features"); return ORBX_ERR_UNSUPPORTED; }
        for (int p1 = a.node_start[ia]; p1 < a.node_start[ia + 1]; p1++) {
          work.push_back(BowWork{p1, b0, b1, (int)nEntries});
          nEntries += b1 - b0;
          if (nEntries > (1ll << 30)) { set_error("candidate list capacity exceeded"); return ORBX_ERR_CAPACITY; }
        }
        ia++; ib++;
      } else if (a.node_ids[ia] < b.node_ids[ib]) ia++;
      else ib++;
    }
  }
  const int nw = (int)work.size();
  if (nw == 0) return ORBX_OK;
  const int ni1 = P.fv1->node_start[P.fv1->nnodes], ni2 = P.fv2->node_start[P.fv2->nnodes];
  ArenaPlan plan;
  const size_t oK1 = plan.add((size_t)P.n1 * sizeof(orbx_keypoint)), oD1 = plan.add((size_t)P.n1 * 32), oI1 = plan.add((size_t)std::max(ni1, 1) * 4);
  const size_t oK2 = plan.add((size_t)P.n2 * sizeof(orbx_keypoint)), oD2 = plan.add((size_t)P.n2 * 32), oI2 = plan.add((size_t)std::max(ni2, 1) * 4);
  const size_t oE1 = P.elig1 ? plan.add(P.n1) : (size_t)-1, oE2 = P.elig2 ? plan.add(P.n2) : (size_t)-1;
  const size_t oS1 = P.stereo1 ? plan.add(P.n1) : (size_t)-1, oS2 = P.stereo2 ? plan.add(P.n2) : (size_t)-1;
  const size_t oW = plan.add((size_t)nw * sizeof(BowWork));
  const size_t oEnt = P.triangulate ? (size_t)-1 : plan.add((size_t)std::max<long long>(nEntries, 1) * 8);
  const size_t oM = plan.add((size_t)P.n1 * 4), oN = plan.add(4);
  const size_t oB0 = plan.add((size_t)P.n2 * 4), oB1 = plan.add((size_t)P.n2 * 4), oChg = plan.add(8), oCh = plan.add((size_t)nw * 4);
  const size_t oSF = P.triangulate ? plan.add((size_t)P.nlevels * 4) : (size_t)-1, oLS = P.triangulate ? plan.add((size_t)P.nlevels * 4) : (size_t)-1;
  const size_t oMR = fish ? plan.add((size_t)P.n1 * 4) : (size_t)-1, oChR = fish ? plan.add((size_t)nw * 4) : (size_t)-1;
  const size_t oLS1 = P.kb8 ? plan.add((size_t)P.nlevels * 4) : (size_t)-1;
  int rc = arena_ensure(m, plan.total);
  if (rc != ORBX_OK) return rc;
  cudaStream_t st = m->stream;
  auto up = [&](size_t o, const void* src, size_t bytes) -> cudaError_t {
    if (bytes == 0 || !src || o == (size_t)-1) return cudaSuccess;
    return cudaMemcpyAsync(m->arena + o, src, bytes, cudaMemcpyHostToDevice, st);
  };
  CK(up(oK1, P.kps1, (size_t)P.n1 * sizeof(orbx_keypoint))); CK(up(oD1, P.desc1, (size_t)P.n1 * 32)); CK(up(oI1, P.fv1->indices, (size_t)ni1 * 4));
  CK(up(oK2, P.kps2, (size_t)P.n2 * sizeof(orbx_keypoint))); CK(up(oD2, P.desc2, (size_t)P.n2 * 32)); CK(up(oI2, P.fv2->indices, (size_t)ni2 * 4));
  CK(up(oE1, P.elig1, P.n1)); CK(up(oE2, P.elig2, P.n2)); CK(up(oS1, P.stereo1, P.n1)); CK(up(oS2, P.stereo2, P.n2));
  CK(up(oW, work.data(), (size_t)nw * sizeof(BowWork)));
  CK(up(oSF, P.scaleFactors2, (size_t)P.nlevels * 4)); CK(up(oLS, P.levelSigma2, (size_t)P.nlevels * 4));
  CK(up(oLS1, P.levelSigma1, (size_t)P.nlevels * 4));
  BowArgs A{};
  A.s1 = BowSide{at<orbx_keypoint>(m, oK1), at<uint8_t>(m, oD1), P.n1, at<unsigned>(m, oI1), P.elig1 ? at<uint8_t>(m, oE1) : nullptr,
                 P.stereo1 ? at<uint8_t>(m, oS1) : nullptr};
  A.s2 = BowSide{at<orbx_keypoint>(m, oK2), at<uint8_t>(m, oD2), P.n2, at<unsigned>(m, oI2), P.elig2 ? at<uint8_t>(m, oE2) : nullptr,
                 P.stereo2 ? at<uint8_t>(m, oS2) : nullptr};
  A.work = at<BowWork>(m, oW); A.nwork = nw;
  A.entries = P.triangulate ? nullptr : at<unsigned long long>(m, oEnt);
  A.matches12 = at<int>(m, oM); A.nmatches = at<int>(m, oN);
  A.maxDist = P.maxDist; A.nnratio = P.nnratio; A.checkOrientation = P.checkOrientation;
  A.blk[0] = at<int>(m, oB0); A.blk[1] = at<int>(m, oB1); A.chg = at<int>(m, oChg); A.choice = at<int>(m, oCh);
  A.nLeft2 = P.nLeft2; A.matches12R = fish ? at<int>(m, oMR) : nullptr; A.choiceR = fish ? at<int>(m, oChR) : nullptr;
  if (P.triangulate) {
    if (P.F12) memcpy(A.F12, P.F12, sizeof(A.F12));
    A.epx = P.epx; A.epy = P.epy; A.scaleFactors2 = at<float>(m, oSF); A.levelSigma2 = at<float>(m, oLS); A.coarse = P.coarse;
    CK(cudaMemsetAsync(m->arena + oM, 0xff, (size_t)P.n1 * 4, st));
    if (P.kb8) {
      Kb8TriArgs T{};
      static_assert(sizeof(Kb8PairDev) == sizeof(orbx_kb8_pair), "orbx_kb8_pair layout");
      memcpy(T.pair, P.kb8, sizeof(orbx_kb8_pair) * (P.triLeft1 != -1 ? 4 : 1));
      T.nLeft1 = P.triLeft1; T.nLeft2 = P.triLeft2; T.levelSigma1 = at<float>(m, oLS1);
      k_bow_triangulate_kb8<<<(nw + 7) / 8, 256, 0, st>>>(A, T);
    } else {
      k_bow_triangulate<<<(nw + 7) / 8, 256, 0, st>>>(A);
    }
  } else {
    k_bow_distances<<<(nw + 7) / 8, 256, 0, st>>>(A);
    k_bow_resolve_par<<<kParCluster, kParThreads, 0, st>>>(A);
  }
  k_bow_rotfilter<<<1, 256, 0, st>>>(A);
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(matches12, m->arena + oM, (size_t)P.n1 * 4, cudaMemcpyDeviceToHost, st));
  if (fish) CK(cudaMemcpyAsync(P.matches12R, m->arena + oMR, (size_t)P.n1 * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(nmatches, m->arena + oN, 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return ORBX_OK;
}
}  // namespace

int orbx_search_by_bow(orbx_matcher* m, int n1, const orbx_keypoint* kps1, const uint8_t* desc1, const orbx_featvec* fv1, const uint8_t* elig1,
                       int n2, const orbx_keypoint* kps2, const uint8_t* desc2, const orbx_featvec* fv2, const uint8_t* elig2, int maxDist,
                       float nnratio, int checkOrientation, int* matches12, int* nmatches) {
  BowProblem P{};
  P.n1 = n1; P.kps1 = kps1; P.desc1 = desc1; P.fv1 = fv1; P.elig1 = elig1;
  P.n2 = n2; P.kps2 = kps2; P.desc2 = desc2; P.fv2 = fv2; P.elig2 = elig2;
  P.maxDist = maxDist; P.nnratio = nnratio; P.checkOrientation = checkOrientation; P.triangulate = false;
  return run_bow(m, P, matches12, nmatches);
}

int orbx_search_by_bow_stereo(orbx_matcher* m, int n1, const orbx_keypoint* kps1, const uint8_t* desc1, const orbx_featvec* fv1,
                              const uint8_t* elig1, int n2, int nLeft2, const orbx_keypoint* kps2, const uint8_t* desc2, const orbx_featvec* fv2,
                              int maxDist, float nnratio, int checkOrientation, int* matches12Left, int* matches12Right, int* nmatches) {
  if (nLeft2 < 0 || nLeft2 > n2 || !matches12Right) return ORBX_ERR_ARG;
  BowProblem P{};
  P.n1 = n1; P.kps1 = kps1; P.desc1 = desc1; P.fv1 = fv1; P.elig1 = elig1;
  P.n2 = n2; P.kps2 = kps2; P.desc2 = desc2; P.fv2 = fv2; P.elig2 = nullptr;
  P.maxDist = maxDist; P.nnratio = nnratio; P.checkOrientation = checkOrientation; P.triangulate = false;
  P.nLeft2 = nLeft2; P.matches12R = matches12Right;
  return run_bow(m, P, matches12Left, nmatches);
}

int orbx_search_for_triangulation(orbx_matcher* m, int n1, const orbx_keypoint* kps1, const uint8_t* desc1, const orbx_featvec* fv1,
                                  const uint8_t* elig1, const uint8_t* stereo1, int n2, const orbx_keypoint* kps2, const uint8_t* desc2,
                                  const orbx_featvec* fv2, const uint8_t* elig2, const uint8_t* stereo2, const float* F12, float epx,
                                  float epy, const float* scaleFactors2, const float* levelSigma2_2, int nlevels, int coarse,
                                  int checkOrientation, int* matches12, int* nmatches) {
  if (!F12 || !scaleFactors2 || !levelSigma2_2 || nlevels < 1 || !elig1 || !elig2) return ORBX_ERR_ARG;
  for (int i = 0; i < n2 && kps2; i++)
    if (kps2[i].octave < 0 || kps2[i].octave >= nlevels) { set_error("keypoint octave outside the level tables"); return ORBX_ERR_ARG; }
  BowProblem P{};
  P.n1 = n1; P.kps1 = kps1; P.desc1 = desc1; P.fv1 = fv1; P.elig1 = elig1; P.stereo1 = stereo1;
  P.n2 = n2; P.kps2 = kps2; P.desc2 = desc2; P.fv2 = fv2; P.elig2 = elig2; P.stereo2 = stereo2;
  P.checkOrientation = checkOrientation; P.triangulate = true;
  P.F12 = F12; P.epx = epx; P.epy = epy; P.scaleFactors2 = scaleFactors2; P.levelSigma2 = levelSigma2_2; P.nlevels = nlevels; P.coarse = coarse;
  return run_bow(m, P, matches12, nmatches);
}

int orbx_search_for_triangulation_kb8(orbx_matcher* m, int n1, const orbx_keypoint* kps1, const uint8_t* desc1, const orbx_featvec* fv1,
                                      const uint8_t* elig1, const uint8_t* stereo1, int nLeft1, int n2, const orbx_keypoint* kps2,
                                      const uint8_t* desc2, const orbx_featvec* fv2, const uint8_t* elig2, const uint8_t* stereo2,
                                      int nLeft2, const orbx_kb8_pair* pairs, int npairs, float epx, float epy,
                                      const float* scaleFactors2, const float* levelSigma2_1, const float* levelSigma2_2, int nlevels,
                                      int coarse, int checkOrientation, int* matches12, int* nmatches) {
  if (!pairs || !scaleFactors2 || !levelSigma2_1 || !levelSigma2_2 || nlevels < 1 || !elig1 || !elig2) return ORBX_ERR_ARG;
  const bool two = nLeft1 != -1;
  if ((nLeft2 != -1) != two || npairs != (two ? 4 : 1)) { set_error("both keyframes carry a second camera (4 pair records) or neither does (1)"); return ORBX_ERR_ARG; }
  if (two && (nLeft1 < 0 || nLeft1 > n1 || nLeft2 < 0 || nLeft2 > n2)) return ORBX_ERR_ARG;
  for (int i = 0; i < n1 && kps1; i++)
    if (kps1[i].octave < 0 || kps1[i].octave >= nlevels) { set_error("keypoint octave outside the level tables"); return ORBX_ERR_ARG; }
  for (int i = 0; i < n2 && kps2; i++)
    if (kps2[i].octave < 0 || kps2[i].octave >= nlevels) { set_error("keypoint octave outside the level tables"); return ORBX_ERR_ARG; }
  BowProblem P{};
  P.n1 = n1; P.kps1 = kps1; P.desc1 = desc1; P.fv1 = fv1; P.elig1 = elig1; P.stereo1 = two ? nullptr : stereo1;
  P.n2 = n2; P.kps2 = kps2; P.desc2 = desc2; P.fv2 = fv2; P.elig2 = elig2; P.stereo2 = two ? nullptr : stereo2;
  P.checkOrientation = checkOrientation; P.triangulate = true;
  P.F12 = nullptr; P.epx = epx; P.epy = epy; P.scaleFactors2 = scaleFactors2; P.levelSigma2 = levelSigma2_2; P.nlevels = nlevels; P.coarse = coarse;
  P.kb8 = pairs; P.triLeft1 = nLeft1; P.triLeft2 = nLeft2; P.levelSigma1 = levelSigma2_1;
  return run_bow(m, P, matches12, nmatches);
}

int orbx_kb8_triangulate_matches(orbx_matcher* m, const orbx_kb8_pair* pair, int n, const float* xy1, const float* xy2, const float* sigma1,
                                 const float* sigma2, float* z1, float* p3D) {
  if (!m || !pair || n < 0 || (n > 0 && (!xy1 || !xy2 || !sigma1 || !sigma2 || !z1 || !p3D))) return ORBX_ERR_ARG;
  if (n == 0) return ORBX_OK;
  CK(cudaSetDevice(m->device));
  ArenaPlan plan;
  const size_t oA = plan.add((size_t)n * 8), oB = plan.add((size_t)n * 8), oS1 = plan.add((size_t)n * 4), oS2 = plan.add((size_t)n * 4);
  const size_t oZ = plan.add((size_t)n * 4), oP = plan.add((size_t)n * 12);
  int rc = arena_ensure(m, plan.total);
  if (rc != ORBX_OK) return rc;
  cudaStream_t st = m->stream;
  CK(cudaMemcpyAsync(m->arena + oA, xy1, (size_t)n * 8, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(m->arena + oB, xy2, (size_t)n * 8, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(m->arena + oS1, sigma1, (size_t)n * 4, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(m->arena + oS2, sigma2, (size_t)n * 4, cudaMemcpyHostToDevice, st));
  Kb8PairDev G;
  memcpy(&G, pair, sizeof(G));
  k_kb8_triangulate_pairs<<<(n + 127) / 128, 128, 0, st>>>(G, n, at<float>(m, oA), at<float>(m, oB), at<float>(m, oS1), at<float>(m, oS2),
                                                          at<float>(m, oZ), at<float>(m, oP));
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(z1, m->arena + oZ, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(p3D, m->arena + oP, (size_t)n * 12, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return ORBX_OK;
}

int orbx_distinctive_descriptors(orbx_matcher* m, const uint8_t* desc, const int* start, int npoints, int* bestIdx, int* bestMedian) {
  if (!m || npoints < 0 || (npoints > 0 && (!start || !bestIdx || !bestMedian))) return ORBX_ERR_ARG;
  if (npoints == 0) return ORBX_OK;
  for (int p = 0; p < npoints; p++)
    if (start[p + 1] < start[p] || start[p] < 0) { set_error("descriptor ranges must ascend"); return ORBX_ERR_ARG; }
  const int total = start[npoints];
  if (total > 0 && !desc) return ORBX_ERR_ARG;
  CK(cudaSetDevice(m->device));
  ArenaPlan plan;
  const size_t oD = plan.add((size_t)std::max(total, 1) * 32), oS = plan.add((size_t)(npoints + 1) * 4);
  const size_t oI = plan.add((size_t)npoints * 4), oM = plan.add((size_t)npoints * 4);
  int rc = arena_ensure(m, plan.total);
  if (rc != ORBX_OK) return rc;
  cudaStream_t st = m->stream;
  if (total > 0) CK(cudaMemcpyAsync(m->arena + oD, desc, (size_t)total * 32, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(m->arena + oS, start, (size_t)(npoints + 1) * 4, cudaMemcpyHostToDevice, st));
  k_distinctive<<<(npoints + 7) / 8, 256, 0, st>>>(at<uint8_t>(m, oD), at<int>(m, oS), npoints, at<int>(m, oI), at<int>(m, oM));
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(bestIdx, m->arena + oI, (size_t)npoints * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(bestMedian, m->arena + oM, (size_t)npoints * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return ORBX_OK;
}

struct orbx_vocabulary {
  int device = 0;
  uint8_t* mem = nullptr;
  VocabDev dev{};
};

int orbx_vocabulary_create(int device, int nnodes, int L, const int* childStart, const int* childIds, const uint8_t* nodeDesc,
                           const double* nodeWeight, const int* nodeWordId, orbx_vocabulary** out) {
  if (!out || nnodes < 2 || L < 1 || L > 63 || !childStart || !childIds || !nodeDesc || !nodeWeight || !nodeWordId) return ORBX_ERR_ARG;
  *out = nullptr;
  // structure checks: child ranges ascend, ids are in range and point "down" (every non-root node has exactly one parent), root not a leaf,
  // depth <= L so the descent terminates
  if (childStart[0] != 0 || childStart[1] <= 0) { set_error("vocabulary root has no children"); return ORBX_ERR_ARG; }
  const int nchild = childStart[nnodes];
  if (nchild != nnodes - 1) { set_error("vocabulary is not a tree (%d child links for %d nodes)", nchild, nnodes); return ORBX_ERR_ARG; }
  std::vector<int> depth(nnodes, -1);
  depth[0] = 0;
  std::vector<int> stack(1, 0);
  int visited = 0;
  while (!stack.empty()) {
    const int id = stack.back(); stack.pop_back();
    visited++;
    if (childStart[id + 1] < childStart[id]) { set_error("vocabulary child ranges must ascend"); return ORBX_ERR_ARG; }
    for (int c = childStart[id]; c < childStart[id + 1]; c++) {
      const int ch = childIds[c];
      if (ch <= 0 || ch >= nnodes || depth[ch] != -1) { set_error("vocabulary child id %d invalid or shared", ch); return ORBX_ERR_ARG; }
      depth[ch] = depth[id] + 1;
      if (depth[ch] > L) { set_error("vocabulary deeper than L = %d", L); return ORBX_ERR_ARG; }
      stack.push_back(ch);
    }
  }
  if (visited != nnodes) { set_error("vocabulary has unreachable nodes"); return ORBX_ERR_ARG; }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) { set_error("no usable CUDA device %d", device); return ORBX_ERR_CUDA; }
  CK(cudaSetDevice(device));
  orbx_vocabulary* v = new orbx_vocabulary();
  v->device = device;
  ArenaPlan plan;
  const size_t oCS = plan.add((size_t)(nnodes + 1) * 4), oCI = plan.add((size_t)nchild * 4), oND = plan.add((size_t)nnodes * 32);
  const size_t oNW = plan.add((size_t)nnodes * 8), oWI = plan.add((size_t)nnodes * 4);
  if (cudaMalloc(&v->mem, plan.total) != cudaSuccess) { delete v; set_error("cudaMalloc of the vocabulary failed"); return ORBX_ERR_CUDA; }
  cudaMemcpy(v->mem + oCS, childStart, (size_t)(nnodes + 1) * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(v->mem + oCI, childIds, (size_t)nchild * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(v->mem + oND, nodeDesc, (size_t)nnodes * 32, cudaMemcpyHostToDevice);
  cudaMemcpy(v->mem + oNW, nodeWeight, (size_t)nnodes * 8, cudaMemcpyHostToDevice);
  if (cudaMemcpy(v->mem + oWI, nodeWordId, (size_t)nnodes * 4, cudaMemcpyHostToDevice) != cudaSuccess) {
    cudaFree(v->mem); delete v; set_error("vocabulary upload failed"); return ORBX_ERR_CUDA;
  }
  v->dev.childStart = reinterpret_cast<int*>(v->mem + oCS); v->dev.childIds = reinterpret_cast<int*>(v->mem + oCI);
  v->dev.nodeDesc = v->mem + oND; v->dev.nodeWeight = reinterpret_cast<double*>(v->mem + oNW);
  v->dev.nodeWordId = reinterpret_cast<int*>(v->mem + oWI);
  v->dev.nnodes = nnodes; v->dev.L = L;
  *out = v;
  return ORBX_OK;
}

void orbx_vocabulary_destroy(orbx_vocabulary* v) {
  if (!v) return;
  cudaSetDevice(v->device);
  if (v->mem) cudaFree(v->mem);
  delete v;
}

int orbx_vocabulary_transform(orbx_matcher* m, const orbx_vocabulary* v, const uint8_t* desc, int n, int levelsup, int* wordId, int* nodeId,
                              double* weight) {
  if (!m || !v || n < 0 || (n > 0 && (!desc || !wordId || !nodeId || !weight))) return ORBX_ERR_ARG;
  if (v->device != m->device) { set_error("vocabulary and matcher live on different devices"); return ORBX_ERR_ARG; }
  if (n == 0) return ORBX_OK;
  CK(cudaSetDevice(m->device));
  ArenaPlan plan;
  const size_t oD = plan.add((size_t)n * 32), oW = plan.add((size_t)n * 4), oN = plan.add((size_t)n * 4), oWt = plan.add((size_t)n * 8);
  int rc = arena_ensure(m, plan.total);
  if (rc != ORBX_OK) return rc;
  cudaStream_t st = m->stream;
  CK(cudaMemcpyAsync(m->arena + oD, desc, (size_t)n * 32, cudaMemcpyHostToDevice, st));
  k_vocab_transform<<<(n + 7) / 8, 256, 0, st>>>(v->dev, at<uint8_t>(m, oD), n, levelsup, at<int>(m, oW), at<int>(m, oN), at<double>(m, oWt));
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(wordId, m->arena + oW, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(nodeId, m->arena + oN, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(weight, m->arena + oWt, (size_t)n * 8, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return ORBX_OK;
}

int orbx_search_for_initialization(orbx_matcher* m, const orbx_frame_view* F1, const orbx_frame_view* F2, float* prevMatched, int windowSize,
                                   float nnratio, int checkOrientation, int* matches12, int* nmatches) {
  if (!m || !F1 || !prevMatched || !matches12 || !nmatches) return ORBX_ERR_ARG;
  int rc = check_view(F2, false);
  if (rc != ORBX_OK) return rc;
  if (F1->n < 0 || (F1->n > 0 && (!F1->keypoints || !F1->descriptors))) return ORBX_ERR_ARG;
  std::vector<WindowJob> jobs(1);
  WindowJob& J = jobs[0];
  J.F = F2; set_bounds(J, F2);
  HostQueries& q = J.q;
  const int n = F1->n;
  q.nq = n; q.valid.resize(n); q.x.resize(n); q.y.resize(n); q.r.assign(n, (float)windowSize); q.minL.resize(n); q.maxL.resize(n);
  q.desc = F1->descriptors;
  for (int i = 0; i < n; i++) {  // ORBmatcher.cc:736-742
    const int level1 = F1->keypoints[i].octave;
    q.valid[i] = !(level1 > 0);
    q.x[i] = prevMatched[2 * i]; q.y[i] = prevMatched[2 * i + 1];
    q.minL[i] = (int8_t)level1; q.maxL[i] = (int8_t)level1;
  }
  J.F1 = F1; J.prevMatched = prevMatched; J.matches12 = matches12; J.nnratio = nnratio; J.checkOrientation = checkOrientation;
  J.nmatches = nmatches;
  return run_window_jobs(m, jobs, RESOLVE_INIT);
}

int orbx_matcher_set_knn2_kernel(orbx_matcher* m, int kind) {
  if (!m || kind < ORBX_KNN2_AUTO || kind > ORBX_KNN2_TENSOR) return ORBX_ERR_ARG;
  m->knn2Kernel = kind;
  return ORBX_OK;
}

int orbx_knn2_chunk_device(orbx_matcher* m, const uint8_t* dQ, int nq, const uint8_t* dT, int nt, long long trainBase, const int* dExclude,
                           unsigned long long* dKey0, unsigned long long* dKey1, int init, void* stream) {
  if (!m || nq < 0 || nt < 0 || !dKey0 || !dKey1) return ORBX_ERR_ARG;
  if (nq == 0) return ORBX_OK;
  if (nt > (1 << 23)) { set_error("train chunk larger than 2^23 rows"); return ORBX_ERR_ARG; }
  CK(cudaSetDevice(m->device));
  cudaStream_t st = stream ? (cudaStream_t)stream : m->stream;
  // k_knn2_tc pays one expansion pass over the train chunk and a per-CTA prologue: it wins from ~1e8 pairs on; below that (and for
  // nt = 0, where only the init matters) the POPC kernel folds.  Both write identical keys.
  const bool tensor = nt > 0 && (m->knn2Kernel == ORBX_KNN2_TENSOR || (m->knn2Kernel == ORBX_KNN2_AUTO && (long long)nq * nt >= (1ll << 26)));
  if (tensor) {
    const size_t need = knn2_tc_workspace_bytes(nt);
    if (need > m->tcWsBytes) {   // cudaFree waits for every kernel still reading the old buffer
      if (m->tcWs) CK(cudaFree(m->tcWs));
      m->tcWs = nullptr; m->tcWsBytes = 0;
      CK(cudaMalloc((void**)&m->tcWs, need));
      m->tcWsBytes = need;
    }
    CK(knn2_tc_launch(dQ, nq, dT, nt, trainBase, reinterpret_cast<const int2*>(dExclude), dKey0, dKey1, init, m->tcWs, st));
    return ORBX_OK;
  }
  k_knn2<<<(nq + 127) / 128, 128, 0, st>>>(dQ, nq, dT, nt, trainBase, reinterpret_cast<const int2*>(dExclude), dKey0, dKey1, init);
  CK(cudaGetLastError());
  return ORBX_OK;
}

int orbx_knn2(orbx_matcher* m, const uint8_t* q, int nq, const uint8_t* t, int nt, int* idx, int* dist, uint8_t* good) {
  if (!m || nq < 0 || nt < 0 || (nq > 0 && (!q || !idx || !dist)) || (nt > 0 && !t)) return ORBX_ERR_ARG;
  if (nq == 0) return ORBX_OK;
  CK(cudaSetDevice(m->device));
  ArenaPlan plan;
  const size_t oq = plan.add((size_t)nq * 32), ot = plan.add((size_t)std::max(nt, 1) * 32), k0 = plan.add((size_t)nq * 8), k1 = plan.add((size_t)nq * 8);
  int rc = arena_ensure(m, plan.total);
  if (rc != ORBX_OK) return rc;
  cudaStream_t st = m->stream;
  CK(cudaMemcpyAsync(m->arena + oq, q, (size_t)nq * 32, cudaMemcpyHostToDevice, st));
  if (nt > 0) CK(cudaMemcpyAsync(m->arena + ot, t, (size_t)nt * 32, cudaMemcpyHostToDevice, st));
  int init = 1;
  if (nt == 0) CK(cudaMemsetAsync(m->arena + k0, 0xff, (size_t)nq * 16 + 256, st));
  for (int base = 0; base < nt; base += (1 << 23)) {
    const int n = std::min(1 << 23, nt - base);
    rc = orbx_knn2_chunk_device(m, m->arena + oq, nq, m->arena + ot + (size_t)base * 32, n, base, nullptr, at<unsigned long long>(m, k0),
                                at<unsigned long long>(m, k1), init, st);
    if (rc != ORBX_OK) return rc;
    init = 0;
  }
  std::vector<unsigned long long> h0(nq), h1(nq);
  CK(cudaMemcpyAsync(h0.data(), m->arena + k0, (size_t)nq * 8, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(h1.data(), m->arena + k1, (size_t)nq * 8, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  for (int i = 0; i < nq; i++) {
    const bool a = h0[i] != kKeyNone, b = h1[i] != kKeyNone;
    idx[2 * i] = a ? (int)(h0[i] & 0xffffffffu) : -1; dist[2 * i] = a ? (int)(h0[i] >> 32) : -1;
    idx[2 * i + 1] = b ? (int)(h1[i] & 0xffffffffu) : -1; dist[2 * i + 1] = b ? (int)(h1[i] >> 32) : -1;
    if (good) good[i] = b && ((float)dist[2 * i] < (float)dist[2 * i + 1] * 0.7);  // Frame.cc:1253
  }
  return ORBX_OK;
}

// Allocations of the stereo stage for the two extractors' geometry (never called while a stream is capturing).
static int stereo_prepare(orbx_matcher* m, orbx_extractor* exL, orbx_extractor* exR) {
  std::lock_guard<std::recursive_mutex> setupLock(setup_mutex());
  const Geom& g = exL->g;
  const int nfCap = std::max(exL->p.max_batch, exR->p.max_batch);
  const size_t half = (size_t)nfCap * g.kpStride;
  if (2 * half > m->sCap) {
    CK(cudaDeviceSynchronize());
    void** ptrs[] = {(void**)&m->sURight, (void**)&m->sDepth, (void**)&m->sBestDist, (void**)&m->sBestIdx, (void**)&m->sSad};
    for (void** p : ptrs) { if (*p) CK(cudaFree(*p)); *p = nullptr; CK(cudaMalloc(p, 2 * half * 4)); }
    m->sCap = 2 * half;
    m->sEpoch++;
  }
  {
    // a right keypoint is listed in at most 2 * ceil(r) + 2 rows, r = 2 * scale[octave]
    const int maxRows = 2 * (int)std::ceil(2.0f * exL->scale[g.nlevels - 1]) + 2;
    const size_t itemCap = (size_t)g.kpStride * maxRows, needStart = 2 * (size_t)nfCap * (g.height + 1), needItems = 2 * (size_t)nfCap * itemCap;
    if (needStart > m->sRowStartCap || needItems > m->sRowItemsCap) {
      CK(cudaDeviceSynchronize());
      if (m->sRowStart) CK(cudaFree(m->sRowStart));
      if (m->sRowItems) CK(cudaFree(m->sRowItems));
      m->sRowStart = nullptr; m->sRowItems = nullptr;
      CK(cudaMalloc(&m->sRowStart, needStart * sizeof(int)));
      CK(cudaMalloc(&m->sRowItems, needItems * sizeof(uint2)));
      m->sRowStartCap = needStart; m->sRowItemsCap = needItems;
      m->sEpoch++;
    }
    const size_t rtSmem = 2 * (size_t)g.height * sizeof(int);
    if (rtSmem > 48 * 1024) CK(cudaFuncSetAttribute(k_stereo_rowtable, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rtSmem));
  }
  return ORBX_OK;
}

// Enqueue ComputeStereoMatches for the batches held by slot `si` of both extractors; results go to half `si` of the matcher's buffers.
static int stereo_launch(orbx_matcher* m, orbx_extractor* exL, orbx_extractor* exR, int si, float mb, float mbf, cudaStream_t st, bool prepared = false) {
  Slot& sL = exL->slot[si];
  Slot& sR = exR->slot[si];
  const Geom& g = exL->g;
  if (exR->g.width != g.width || exR->g.height != g.height || exR->g.nlevels != g.nlevels || exR->g.kpStride != g.kpStride ||
      sL.lastFrames != sR.lastFrames || sL.lastFrames < 1 || exL->p.scale_factor != exR->p.scale_factor) {
    set_error("left / right extractors must hold batches of the same geometry and size");
    return ORBX_ERR_ARG;
  }
  if (g.kpStride > 65535) { set_error("more than 65535 keypoints per frame"); return ORBX_ERR_UNSUPPORTED; }
  if (!prepared) { const int rc = stereo_prepare(m, exL, exR); if (rc != ORBX_OK) return rc; }
  const int nf = sL.lastFrames;
  const int nfCap = std::max(exL->p.max_batch, exR->p.max_batch);
  const size_t half = (size_t)nfCap * g.kpStride;
  m->sFrames = nf; m->sStride = g.kpStride; m->sHalf = half; m->sLast = si;
  ScaleTab tab{};
  for (int l = 0; l < g.nlevels; l++) { tab.scale[l] = exL->scale[l]; tab.invScale[l] = exL->invScale[l]; }
  Geom gg = g;
  const float maxD = mbf / mb;  // Frame.cc:931-933 (minZ = mb, minD = 0)
  const size_t o = (size_t)si * half;
  StereoOut out{m->sURight + o, m->sDepth + o, m->sBestDist + o, m->sBestIdx + o, m->sSad + o};
  dim3 grid((g.kpStride + 7) / 8, nf);
  {
    const int maxRows = 2 * (int)std::ceil(2.0f * exL->scale[g.nlevels - 1]) + 2;
    const size_t itemCap = (size_t)g.kpStride * maxRows;
    RowTable rt{m->sRowStart + (size_t)si * nfCap * (g.height + 1), m->sRowItems + (size_t)si * nfCap * itemCap, g.height, (int)itemCap};
    k_stereo_rowtable<<<nf, 256, 2 * (size_t)g.height * sizeof(int), st>>>(gg, sR.last, tab, rt);
    k_stereo_hamming2<<<grid, 256, 0, st>>>(gg, sL.last, sR.last, maxD, rt, out);
  }
  k_stereo_sad<<<grid, 256, 0, st>>>(gg, sL.last, sR.last, tab, maxD, mbf, out);
  k_stereo_median<<<nf, 256, (size_t)g.kpStride * sizeof(int), st>>>(gg, sL.last, out);
  CK(cudaGetLastError());
  return ORBX_OK;
}

int orbx_stereo_match_batch(orbx_matcher* m, orbx_extractor* exL, orbx_extractor* exR, float mb, float mbf, void* stream) {
  if (!m || !exL || !exR || !exL->haveGeom || !exR->haveGeom) return ORBX_ERR_ARG;
  if (exL->lastSlot != exR->lastSlot) { set_error("left / right extractors last ran in different slots"); return ORBX_ERR_ARG; }
  CK(cudaSetDevice(m->device));
  cudaStream_t st = stream ? (cudaStream_t)stream : m->stream;
  for (orbx_extractor* e : {exL, exR}) {   // the batches may have been enqueued on other streams (or the handles' own): wait for them on the device
    Slot& s = e->slot[e->lastSlot];
    if (s.evBatchValid && s.lastStream != st) CK(cudaStreamWaitEvent(st, s.evBatch, 0));
  }
  return stereo_launch(m, exL, exR, exL->lastSlot, mb, mbf, st);
}

// Whole stereo frames from host buffers: per chunk H2D(left,right) -> both extractions -> stereo matching -> D2H, double-buffered
// over the two slots of the extractor handles so copies overlap kernels (the e2e form of Frame::Frame(stereo), src/Frame.cc:89-187).
int orbx_stereo_frames_host(orbx_matcher* m, orbx_extractor* exL, orbx_extractor* exR, const uint8_t* imagesL, const uint8_t* imagesR,
                            int nframes, int width, int height, int stride, size_t frameStride, float mb, float mbf, orbx_keypoint* kpsL,
                            uint8_t* descL, int* countsL, orbx_keypoint* kpsR, uint8_t* descR, int* countsR, float* uRight, float* depth,
                            int capPerFrame) {
  if (!m || !exL || !exR) return ORBX_ERR_ARG;
  if (!imagesL || !imagesR || width <= 0 || height <= 0 || nframes <= 0) return ORBX_EMPTY_IMAGE;
  if (!kpsL || !descL || !countsL || !countsR || !uRight || !depth || capPerFrame <= 0 || stride < width) return ORBX_ERR_ARG;
  if ((kpsR == nullptr) != (descR == nullptr)) return ORBX_ERR_ARG;   // the right image's records are downloaded together or not at all
  int rc = extractor_build_geom(exL, width, height);
  if (rc != ORBX_OK) return rc;
  rc = extractor_build_geom(exR, width, height);
  if (rc != ORBX_OK) return rc;
  CK(cudaSetDevice(m->device));
  const Geom& g = exL->g;
  const int B = std::min(exL->p.max_batch, exR->p.max_batch);
  const int ncopy = std::min(capPerFrame, g.kpStride);
  rc = stereo_prepare(m, exL, exR);
  if (rc != ORBX_OK) return rc;
  const bool useGraph = !exL->profile && !exR->profile;
  // ORBX_TRACE=1 (development): device timeline of the call, one line per chunk on stderr
  static const bool trace = [] { const char* e = getenv("ORBX_TRACE"); return e && e[0] == '1'; }();
  std::vector<cudaEvent_t> tev;
  auto tmark = [&](cudaStream_t cs) { if (trace) { cudaEvent_t e; cudaEventCreate(&e); cudaEventRecord(e, cs); tev.push_back(e); } };
  tmark(exL->slot[0].stream);
  // Chunk schedule: the call's latency is upload(first chunk, left) + kernels + download(last chunk), and a chunk's kernel time
  // has a fixed part (~0.1 ms of short dependent kernels), so the first chunk is a quarter of the others.
  const int first = (nframes > B / 4 && B >= 8) ? B / 4 : B;
  // on every return below (errors included) nothing may still be copying into the caller's buffers or hold the slot events open
  struct Drain {
    orbx_matcher* m; orbx_extractor* a; orbx_extractor* b;
    ~Drain() { for (int k = 0; k < 2; k++) { cudaStreamSynchronize(a->slot[k].stream); cudaStreamSynchronize(b->slot[k].stream); } cudaStreamSynchronize(m->upStream); }
  } drain{m, exL, exR};
  int chunk = 0;
  for (int f0 = 0; f0 < nframes; chunk++) {
    const int nb = std::min(chunk == 0 ? first : B, nframes - f0), si = chunk & 1;
    cudaStream_t st = exL->slot[si].stream, stR = exR->slot[si].stream;
    orbx_extractor* ex[2] = {exL, exR};
    const uint8_t* img[2] = {imagesL, imagesR};
    auto upload = [&](int e, cudaStream_t cs) -> int {
      Slot& s = ex[e]->slot[si];
      const uint8_t* src = img[e] + (size_t)f0 * frameStride;
      if (frameStride == (size_t)stride * height && stride == ex[e]->inPitch)
        CK(cudaMemcpyAsync(s.in, src, (size_t)stride * height * nb, cudaMemcpyHostToDevice, cs));
      else if (frameStride == (size_t)stride * height)
        CK(cudaMemcpy2DAsync(s.in, ex[e]->inPitch, src, stride, width, (size_t)height * nb, cudaMemcpyHostToDevice, cs));
      else
        for (int f = 0; f < nb; f++)
          CK(cudaMemcpy2DAsync(s.in + (size_t)f * ex[e]->inPitch * height, ex[e]->inPitch, src + (size_t)f * frameStride, stride, width,
                               height, cudaMemcpyHostToDevice, cs));
      return ORBX_OK;
    };
    auto extract = [&](int e, cudaStream_t cs) -> int {
      return extractor_run_batch(ex[e], si, ex[e]->slot[si].in, nb, ex[e]->inPitch, (long long)ex[e]->inPitch * height, 0, 0, cs);
    };
    if (!useGraph) {   // direct launches on one stream: the left extraction overlaps the right upload
      for (int e = 0; e < 2; e++) {
        if ((rc = upload(e, st)) != ORBX_OK) return rc;
        if ((rc = extract(e, st)) != ORBX_OK) return rc;
      }
      if ((rc = stereo_launch(m, exL, exR, si, mb, mbf, st, true)) != ORBX_OK) return rc;
    } else {
      // Three graphs per chunk.  The two extractions are independent branches, like the reference's two threads
      // (src/Frame.cc:120-123): left on the slot stream of exL, right on the slot stream of exR; the right upload is ordered
      // behind the left one so the left extraction starts as soon as its own images have arrived.  Everything that used this
      // slot before has completed (the host waits for a slot's previous chunk before reusing it).
      auto matches = [&](const orbx_matcher::ChunkGraph& c) {
        return c.execL && c.nb == nb && c.mb == mb && c.mbf == mbf && c.exL == exL && c.exR == exR && c.epochL == exL->geomEpoch &&
               c.epochR == exR->geomEpoch && c.epochM == m->sEpoch;
      };
      int ci = -1;
      for (int k = 0; k < 4; k++) if (matches(m->chunkGraph[si][k])) ci = k;
      const bool hit = ci >= 0;
      if (!hit) { ci = m->chunkGraphNext[si]; m->chunkGraphNext[si] = (ci + 1) & 3; }
      orbx_matcher::ChunkGraph& cg = m->chunkGraph[si][ci];
      if (!hit) {
        std::lock_guard<std::recursive_mutex> setupLock(setup_mutex());
        cg.reset();
        auto capture = [&](cudaStream_t cs, cudaGraphExec_t* out, auto&& body) -> int {
          cudaGraph_t graph = nullptr;
          CK(cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal));
          const int brc = body(cs);
          const cudaError_t ce = cudaStreamEndCapture(cs, &graph);
          if (brc != ORBX_OK) { if (graph) cudaGraphDestroy(graph); return brc; }
          CK(ce);
          CK(cudaGraphInstantiate(out, graph, 0));
          CK(cudaGraphDestroy(graph));
          return ORBX_OK;
        };
        if ((rc = capture(st, &cg.execL, [&](cudaStream_t cs) { return extract(0, cs); })) != ORBX_OK) return rc;
        if ((rc = capture(stR, &cg.execR, [&](cudaStream_t cs) { return extract(1, cs); })) != ORBX_OK) return rc;
        if ((rc = capture(st, &cg.execS, [&](cudaStream_t cs) { return stereo_launch(m, exL, exR, si, mb, mbf, cs, true); })) != ORBX_OK) return rc;
        cg.nb = nb; cg.mb = mb; cg.mbf = mbf; cg.exL = exL; cg.exR = exR;
        cg.epochL = exL->geomEpoch; cg.epochR = exR->geomEpoch; cg.epochM = m->sEpoch;
      } else {
        // same bookkeeping as the captured calls made (the batch descriptors kept in the slots are unchanged)
        for (int e = 0; e < 2; e++) { ex[e]->slot[si].lastFrames = nb; ex[e]->lastSlot = si; ex[e]->slot[si].evValid = false; }
        m->sFrames = nb; m->sStride = g.kpStride; m->sHalf = (size_t)std::max(exL->p.max_batch, exR->p.max_batch) * g.kpStride; m->sLast = si;
      }
      // uploads: one stream for the whole call, so the link carries L0 R0 L1 R1 ... in that order without waiting for kernels
      tmark(m->upStream);
      CK(cudaStreamWaitEvent(m->upStream, m->evInFree[si], 0));    // the slot's previous extractions have consumed their inputs
      if ((rc = upload(0, m->upStream)) != ORBX_OK) return rc;
      CK(cudaEventRecord(m->evUpL[si], m->upStream));
      tmark(m->upStream);
      if ((rc = upload(1, m->upStream)) != ORBX_OK) return rc;
      CK(cudaEventRecord(m->evUpR[si], m->upStream));
      tmark(m->upStream);
      // right branch (behind the slot's previous downloads, which ran on the left stream)
      CK(cudaStreamWaitEvent(stR, m->evSlotDone[si], 0));
      CK(cudaStreamWaitEvent(stR, m->evUpR[si], 0));
      CK(cudaGraphLaunch(cg.execR, stR));
      CK(cudaEventRecord(m->evJoin[si], stR));
      tmark(stR);
      // left branch, join, stereo
      CK(cudaStreamWaitEvent(st, m->evUpL[si], 0));
      CK(cudaGraphLaunch(cg.execL, st));
      tmark(st);
      CK(cudaStreamWaitEvent(st, m->evJoin[si], 0));
      CK(cudaGraphLaunch(cg.execS, st));
      // the staged images are level 0 of both pyramids, which the SAD refinement still reads (Frame.cc:998-1017): the slot's
      // input buffers are free only after the stereo kernels
      CK(cudaEventRecord(m->evInFree[si], st));
      tmark(st);
    }
    orbx_keypoint* kd[2] = {kpsL, kpsR}; uint8_t* dd[2] = {descL, descR}; int* cd[2] = {countsL, countsR};
    // downloads: rows of `ncopy` records per frame; one linear copy when the caller's capacity equals the device stride
    auto download = [&](void* dst, const void* src, size_t rec) -> int {
      if (capPerFrame == g.kpStride) CK(cudaMemcpyAsync(dst, src, rec * g.kpStride * nb, cudaMemcpyDeviceToHost, st));
      else CK(cudaMemcpy2DAsync(dst, rec * capPerFrame, src, rec * g.kpStride, rec * ncopy, nb, cudaMemcpyDeviceToHost, st));
      return ORBX_OK;
    };
    for (int e = 0; e < 2; e++) {
      Slot& s = ex[e]->slot[si];
      CK(cudaMemcpyAsync(cd[e] + f0, s.counts, nb * sizeof(int), cudaMemcpyDeviceToHost, st));
      if (kd[e]) {
        if ((rc = download(kd[e] + (size_t)f0 * capPerFrame, s.kps, sizeof(orbx_keypoint))) != ORBX_OK) return rc;
        if ((rc = download(dd[e] + (size_t)f0 * capPerFrame * 32, s.desc, 32)) != ORBX_OK) return rc;
      }
      CK(cudaMemcpyAsync(s.hStatus, s.status, sizeof(int), cudaMemcpyDeviceToHost, st));
      CK(cudaMemcpyAsync(m->hStat + 2 * (chunk % orbx_matcher::kStatChunks) + e, s.status, sizeof(int), cudaMemcpyDeviceToHost, st));
    }
    const size_t o = (size_t)si * m->sHalf;
    if ((rc = download(uRight + (size_t)f0 * capPerFrame, m->sURight + o, 4)) != ORBX_OK) return rc;
    if ((rc = download(depth + (size_t)f0 * capPerFrame, m->sDepth + o, 4)) != ORBX_OK) return rc;
    tmark(st);
    CK(cudaEventRecord(m->evSlotDone[si], st));
    f0 += nb;
    const bool wrap = (chunk + 1) % orbx_matcher::kStatChunks == 0;
    if (!useGraph) {
      if (chunk >= 1) CK(cudaStreamSynchronize(exL->slot[si ^ 1].stream));  // one stream per slot: at most two chunks in flight
    } else if (wrap || f0 >= nframes) {
      // graph pipeline: slot reuse is ordered by events on the device, the host only waits at the end (or when the status ring wraps)
      for (int k = 0; k < 2; k++) CK(cudaStreamSynchronize(exL->slot[k].stream));
      const int used = wrap ? orbx_matcher::kStatChunks : (chunk % orbx_matcher::kStatChunks) + 1;
      for (int k = 0; k < 2 * used; k++)
        if (m->hStat[k] != 0) { set_error("device reported status %d (capacity)", m->hStat[k]); return m->hStat[k]; }
    }
  }
  for (int si = 0; si < 2; si++) CK(cudaStreamSynchronize(exL->slot[si].stream));
  if (trace && useGraph) {
    static const char* kName[7] = {"begin", "upL", "upR", "extR", "extL", "stereo", "d2h"};
    for (size_t c = 0; 1 + 7 * (c + 1) <= tev.size(); c++) {
      fprintf(stderr, "[orbx trace] chunk %zu:", c);
      for (int k = 0; k < 7; k++) { float ms = 0; cudaEventElapsedTime(&ms, tev[0], tev[1 + 7 * c + k]); fprintf(stderr, " %s %.3f", kName[k], ms); }
      fprintf(stderr, "\n");
    }
  }
  for (cudaEvent_t e : tev) cudaEventDestroy(e);
  for (int si = 0; si < 2; si++)
    for (orbx_extractor* e : {exL, exR})
      if (e->slot[si].lastFrames > 0 && *e->slot[si].hStatus != 0) { set_error("device reported status %d (capacity)", *e->slot[si].hStatus); return *e->slot[si].hStatus; }
  for (int f = 0; f < nframes; f++)
    if (countsL[f] > capPerFrame || countsR[f] > capPerFrame) { set_error("frame %d exceeds the output capacity", f); return ORBX_ERR_CAPACITY; }
  return ORBX_OK;
}

int orbx_stereo_results_device(orbx_matcher* m, const float** uRight, const float** depth, int* stride) {
  if (!m || m->sFrames < 1) return ORBX_ERR_ARG;
  if (uRight) *uRight = m->sURight + (size_t)m->sLast * m->sHalf;
  if (depth) *depth = m->sDepth + (size_t)m->sLast * m->sHalf;
  if (stride) *stride = m->sStride;
  return ORBX_OK;
}

int orbx_stereo_results_host(orbx_matcher* m, float* uRight, float* depth, int* bestDist, int* bestIdx, int capPerFrame) {
  if (!m || m->sFrames < 1 || capPerFrame < 1) return ORBX_ERR_ARG;
  CK(cudaSetDevice(m->device));
  CK(cudaDeviceSynchronize());
  const size_t w = (size_t)std::min(capPerFrame, m->sStride) * 4;
  const size_t o = (size_t)m->sLast * m->sHalf;
  struct { void* dst; const void* src; } c[] = {{uRight, m->sURight + o}, {depth, m->sDepth + o}, {bestDist, m->sBestDist + o}, {bestIdx, m->sBestIdx + o}};
  for (auto& e : c)
    if (e.dst) CK(cudaMemcpy2D(e.dst, (size_t)capPerFrame * 4, e.src, (size_t)m->sStride * 4, w, m->sFrames, cudaMemcpyDeviceToHost));
  return ORBX_OK;
}


// ---- orbx_knn2_device_static: the 2-NN fold without any allocation ---------------------------------------------------------------
// No counterpart in the reference (see include/orbx.h).  orbx_knn2 keeps its running
// (best, second) keys in the matcher's arena (cudaMalloc on growth); this variant keeps them in a file-scope device array that
// every GPU loading the library holds until exit: ORBX_KNN2_STATIC_MAX_QUERIES x 2 keys x 8 bytes = 256 KB.  The array's name is a
// host-side stand-in in host code: its device address is fetched once per device with cudaGetSymbolAddress (that device current)
// and only that pointer is handed to kernels.
static __device__ __align__(256) unsigned long long g_knn2_static_keys[2 * ORBX_KNN2_STATIC_MAX_QUERIES];

// keys -> idx / dist / good exactly as the host loop of orbx_knn2 does (Frame.cc:1253: float distance against a double 0.7)
__global__ void __launch_bounds__(128) k_knn2_unpack(const unsigned long long* key0, const unsigned long long* key1, int nq, int* idx, int* dist,
                                                      uint8_t* good) {
  const int i = blockIdx.x * 128 + threadIdx.x;
  if (i >= nq) return;
  const unsigned long long h0 = key0[i], h1 = key1[i];
  const bool a = h0 != kKeyNone, b = h1 != kKeyNone;
  const int d0 = a ? (int)(h0 >> 32) : -1, d1 = b ? (int)(h1 >> 32) : -1;
  idx[2 * i] = a ? (int)(h0 & 0xffffffffu) : -1; dist[2 * i] = d0;
  idx[2 * i + 1] = b ? (int)(h1 & 0xffffffffu) : -1; dist[2 * i + 1] = d1;
  if (good) good[i] = b && ((double)(float)d0 < __dmul_rn((double)(float)d1, 0.7));
}

static int knn2_static_scratch(unsigned long long** out) {
  constexpr int kMaxDevices = 64;
  static std::mutex mu;
  static unsigned long long* ptrs[kMaxDevices] = {};
  int dev = -1;
  CK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) { set_error("device ordinal %d outside the static scratch table", dev); return ORBX_ERR_UNSUPPORTED; }
  std::lock_guard<std::mutex> lock(mu);
  if (!ptrs[dev]) {
    unsigned long long* p = nullptr;
    CK(cudaGetSymbolAddress((void**)&p, g_knn2_static_keys));
    if (!p) { set_error("cudaGetSymbolAddress returned a null scratch address"); return ORBX_ERR_CUDA; }
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return ORBX_OK;
}

int orbx_knn2_device_static(const uint8_t* dQ, int nq, const uint8_t* dT, int nt, int* dIdx, int* dDist, uint8_t* dGood, void* stream) {
  if (nq < 0 || nt < 0 || (nq > 0 && (!dQ || !dIdx || !dDist)) || (nt > 0 && !dT)) return ORBX_ERR_ARG;
  if (nq > ORBX_KNN2_STATIC_MAX_QUERIES) {
    set_error("orbx_knn2_device_static: %d queries exceed the static scratch (ORBX_KNN2_STATIC_MAX_QUERIES = %d); use orbx_knn2", nq,
              ORBX_KNN2_STATIC_MAX_QUERIES);
    return ORBX_ERR_CAPACITY;
  }
  if (nq == 0) return ORBX_OK;
  unsigned long long* keys = nullptr;
  const int rc = knn2_static_scratch(&keys);
  if (rc != ORBX_OK) return rc;
  unsigned long long *key0 = keys, *key1 = keys + ORBX_KNN2_STATIC_MAX_QUERIES;
  cudaStream_t st = (cudaStream_t)stream;
  // the array keeps the previous call's keys: the first fold runs with init = 1, which overwrites every key this call reads
  int init = 1, base = 0;
  do {
    const int n = std::min(1 << 23, nt - base);
    k_knn2<<<(nq + 127) / 128, 128, 0, st>>>(dQ, nq, dT ? dT + (size_t)base * 32 : nullptr, n, base, nullptr, key0, key1, init);
    CK(cudaGetLastError());
    init = 0; base += n;
  } while (base < nt);
  k_knn2_unpack<<<(nq + 127) / 128, 128, 0, st>>>(key0, key1, nq, dIdx, dDist, dGood);
  CK(cudaGetLastError());
  return ORBX_OK;
}

}  // extern "C"
