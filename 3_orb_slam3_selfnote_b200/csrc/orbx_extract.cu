// Host side of the extractor: handle, geometry/tables, workspaces, stream orchestration, and the extern "C" entry points
// declared in include/orbx.h.  All pixel / keypoint arithmetic happens in the kernels of extract_kernels.cuh; this file
// only sizes buffers, fills the small lookup tables and enqueues launches.  There is no CPU fallback.
#include <cuda.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <mutex>
#include <string>
#include <vector>

#include "extract_kernels.cuh"
#include "fast_nms3.cuh"
#include "pyr_resize.cuh"
#include "blur7.cuh"
#include "orient_desc.cuh"
#include "orbx_internal.h"

namespace orbx {

static const int8_t kPatternHost[1024] = {
#include "orb_pattern.inc"
};

thread_local std::string g_err;
void set_error(const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_err = buf;
}

static inline int round_up(int v, int a) { return (v + a - 1) / a * a; }
static inline int cv_round_f(float v) { return (int)lrintf(v); }
static inline int cv_floor_d(double v) { int i = (int)v; return i - (i > v); }

}  // namespace orbx

using namespace orbx;

static void free_geom(orbx_extractor* h) {
  cudaSetDevice(h->p.device);
  for (Slot& s : h->slot) {
    if (s.stream) cudaStreamSynchronize(s.stream);
    void* ptrs[] = {s.in, s.pyr, s.blur, s.cand, s.zero, s.lkp, s.lkpCount, s.lkpDest, s.nCand, s.qXY, s.qOrd, s.qNode, s.qResp, s.qSel,
                    s.kps, s.desc, s.counts, s.mono};
    for (void* q : ptrs) if (q) cudaFree(q);
    if (s.hStatus) cudaFreeHost(s.hStatus);
    cudaStream_t st = s.stream;
    cudaEvent_t eb = s.evBatch;
    for (cudaEvent_t e : s.ev) if (e) cudaEventDestroy(e);
    s = Slot();
    s.stream = st; s.evBatch = eb;
  }
  if (h->dRsx) cudaFree(h->dRsx);
  if (h->dRsy) cudaFree(h->dRsy);
  if (h->dInfo) cudaFree(h->dInfo);
  if (h->dTiles) cudaFree(h->dTiles);
  h->dRsx = nullptr; h->dRsy = nullptr; h->dInfo = nullptr; h->dTiles = nullptr;
  h->haveGeom = false;
}

// Build level sizes, cell grids, quadtree roots, tile tables and the lookup tables for one image size.
std::recursive_mutex& orbx::setup_mutex() { static std::recursive_mutex m; return m; }

int orbx::extractor_build_geom(orbx_extractor* h, int width, int height) {
  if (h->haveGeom && h->g.width == width && h->g.height == height) return ORBX_OK;   // steady state: no lock
  std::lock_guard<std::recursive_mutex> setupLock(setup_mutex());
  if (h->haveGeom && h->g.width == width && h->g.height == height) return ORBX_OK;
  if (h->haveGeom) free_geom(h);
  h->geomEpoch++;
  h->rs3ok = true;
  CK(cudaSetDevice(h->p.device));
  if (width > 8192 || height > 8192) { set_error("image larger than 8192 px is unsupported"); return ORBX_ERR_UNSUPPORTED; }
  Geom& g = h->g;
  memset(&g, 0, sizeof g);
  const int nl = h->p.nlevels;
  g.nlevels = nl; g.width = width; g.height = height;
  g.iniTh = h->p.ini_th_fast; g.minTh = h->p.min_th_fast;
  static const int taps[2][7] = {{18, 34, 49, 55, 49, 34, 18}, {18, 34, 48, 56, 48, 34, 18}};
  for (int i = 0; i < 7; i++) g.taps[i] = taps[h->p.blur_taps][i];
  for (int i = 0; i < 16; i++) g.umax[i] = h->umax[i];

  std::vector<RsX> rsx; std::vector<RsY> rsy; std::vector<uint16_t> info;
  int rsBoxW = 0, rsBoxH = 0;
  int pyrOff = 0, blurOff = 0, cellBase = 0, candOff = 0, kpOff = 0, tileBase = 0, blurTileBase = 0, maxNodeCap = 0;
  h->inPitch = round_up(width, 4);  // dense when width % 4 == 0: host frames then upload as one linear copy
  for (int l = 0; l < nl; l++) {
    LevelGeom& L = g.lv[l];
    // level size, src/ORBextractor.cc:1192-1193
    L.w = cv_round_f((float)width * h->invScale[l]);
    L.h = cv_round_f((float)height * h->invScale[l]);
    if (L.w < 1 || L.h < 1) { set_error("level %d is empty for a %dx%d image", l, width, height); return ORBX_ERR_UNSUPPORTED; }
    L.pitch = (l == 0) ? h->inPitch : round_up(L.w, 64);
    L.pyrOff = pyrOff;
    if (l > 0) pyrOff += L.pitch * L.h;
    L.blurPitch = round_up(L.w, 64);
    L.blurOff = blurOff;
    blurOff += L.blurPitch * L.h;
    L.scale = h->scale[l];
    L.kpSize = (float)(int)(kPatch * h->scale[l]);  // src/ORBextractor.cc:862
    L.quota = h->quota[l];
    // cell grid, src/ORBextractor.cc:771-785
    const int minB = kMinBorder, maxBX = L.w - kEdge + 3, maxBY = L.h - kEdge + 3;
    const float fw = (float)(maxBX - minB), fh = (float)(maxBY - minB);
    L.nCols = fw > 0 ? (int)(fw / 30.f) : 0;
    L.nRows = fh > 0 ? (int)(fh / 30.f) : 0;
    if (L.nCols > 0 && L.nRows > 0) {
      L.wCell = (int)ceilf(fw / L.nCols);
      L.hCell = (int)ceilf(fh / L.nRows);
    } else { L.nCols = L.nRows = 0; L.wCell = L.hCell = 1; }
    if (L.wCell > 63 || L.hCell > 63) { set_error("cell larger than 63 px"); return ORBX_ERR_UNSUPPORTED; }
    L.cellBase = cellBase;
    cellBase += L.nCols * L.nRows;
    if (L.nCols * L.nRows > 65535) { set_error("too many cells"); return ORBX_ERR_UNSUPPORTED; }
    // quadtree roots, src/ORBextractor.cc:541-543
    const int W = maxBX - minB, H = maxBY - minB;
    if (L.nCols > 0) {
      L.nIni = (int)roundf((float)W / H);
      if (L.nIni < 1) { set_error("aspect ratio below 0.5 is unsupported (the reference divides by zero, ORBextractor.cc:541-543)"); return ORBX_ERR_UNSUPPORTED; }
      if (L.nIni > 16) { set_error("aspect ratio above 16 is unsupported"); return ORBX_ERR_UNSUPPORTED; }
      L.hX = (float)W / L.nIni;
    } else { L.nIni = 1; L.hX = 1.f; }
    L.nodeCap = round_up(std::max(L.quota + 3, 4 * L.nIni) + 4, 8);
    maxNodeCap = std::max(maxNodeCap, L.nodeCap);
    L.kpCap = L.nodeCap; L.kpOff = kpOff; kpOff += L.kpCap;
    L.candCap = round_up(std::max(1024, L.w * L.h / 8), 64);
    L.candOff = candOff; candOff += L.candCap;
    // FAST tiles over the region [19, w-19) x [19, h-19)
    const int rw = L.w - 2 * kEdge, rh = L.h - 2 * kEdge;
    const int ftw = kF3TileW, fth = kF3TileH;
    L.tilesX = (rw > 0 && rh > 0 && L.nCols > 0) ? (rw + ftw - 1) / ftw : 0;
    L.tilesY = (rw > 0 && rh > 0 && L.nCols > 0) ? (rh + fth - 1) / fth : 0;
    L.tileBase = tileBase; tileBase += L.tilesX * L.tilesY;
    const int btw = kBvW, bth = kBvH;
    L.blurTilesX = (L.w + btw - 1) / btw; L.blurTilesY = (L.h + bth - 1) / bth;
    L.blurTileBase = blurTileBase; blurTileBase += L.blurTilesX * L.blurTilesY;
    // column / row cell info: cell index | first<<14 | last<<15
    L.colInfoOff = (int)info.size();
    for (int x = 0; x < L.w; x++) {
      uint16_t v = 0;
      if (x >= kEdge && x < L.w - kEdge && L.nCols > 0) {
        const int c = (x - kEdge) / L.wCell, r = (x - kEdge) % L.wCell;
        v = (uint16_t)c;
        if (r == 0) v |= 0x4000;
        if (r == L.wCell - 1 || x == L.w - kEdge - 1) v |= 0x8000;
      }
      info.push_back(v);
    }
    L.rowInfoOff = (int)info.size();
    for (int y = 0; y < L.h; y++) {
      uint16_t v = 0;
      if (y >= kEdge && y < L.h - kEdge && L.nRows > 0) {
        const int c = (y - kEdge) / L.hCell, r = (y - kEdge) % L.hCell;
        v = (uint16_t)c;
        if (r == 0) v |= 0x4000;
        if (r == L.hCell - 1 || y == L.h - kEdge - 1) v |= 0x8000;
      }
      info.push_back(v);
    }
    // resize tables (cv::resize INTER_LINEAR 8U) from level l-1
    if (l > 0) {
      const LevelGeom& S = g.lv[l - 1];
      const double scale_x = 1. / ((double)L.w / S.w), scale_y = 1. / ((double)L.h / S.h);
      L.rsxOff = (int)rsx.size();
      const int wpad = round_up(L.w, 4);
      for (int dx0 = 0; dx0 < wpad; dx0++) {
        const int dx = std::min(dx0, L.w - 1);
        float fx = (float)((dx + 0.5) * scale_x - 0.5);
        int sx = cv_floor_d(fx);
        fx -= sx;
        if (sx < 0) { fx = 0; sx = 0; }
        if (sx >= S.w - 1) { fx = 0; sx = S.w - 1; }
        RsX e;
        e.s0 = (int16_t)sx; e.s1 = (int16_t)std::min(sx + 1, S.w - 1);
        e.a0 = (int16_t)std::min(32767, std::max(-32768, cv_round_f((1.f - fx) * 2048.f)));
        e.a1 = (int16_t)std::min(32767, std::max(-32768, cv_round_f(fx * 2048.f)));
        rsx.push_back(e);
      }
      L.rsyOff = (int)rsy.size();
      for (int dy = 0; dy < L.h; dy++) {
        float fy = (float)((dy + 0.5) * scale_y - 0.5);
        int sy = cv_floor_d(fy);
        fy -= sy;
        RsY e;
        e.s0 = (int16_t)std::min(std::max(sy, 0), S.h - 1); e.s1 = (int16_t)std::min(std::max(sy + 1, 0), S.h - 1);
        e.b0 = (int16_t)std::min(32767, std::max(-32768, cv_round_f((1.f - fy) * 2048.f)));
        e.b1 = (int16_t)std::min(32767, std::max(-32768, cv_round_f(fy * 2048.f)));
        rsy.push_back(e);
      }
      // k_pyr_resize4 covers the taps of 4 adjacent destination columns with 8 source bytes
      for (int x0 = 0; x0 < wpad; x0 += 4)
        if (rsx[L.rsxOff + x0 + 3].s1 - rsx[L.rsxOff + x0].s0 > 7 || (L.rsxOff & 1)) h->rs3ok = false;
      // staged source window of k_pyr_resize2: maxima over the level's tiles
      const RsX* tx = rsx.data() + L.rsxOff;
      const RsY* ty = rsy.data() + L.rsyOff;
      for (int x0 = 0; x0 < L.w; x0 += kRsTileW) {
        const int xl = std::min(x0 + kRsTileW, wpad) - 1;
        const int gx0 = tx[x0].s0 & ~3;
        g.rsWinWords = std::max(g.rsWinWords, ((tx[xl].s1 - gx0) >> 2) + 1);
      }
      for (int y0 = 0; y0 < L.h; y0 += kRsTileH) {
        const int yl = std::min(y0 + kRsTileH, L.h) - 1;
        g.rsWinRows = std::max(g.rsWinRows, ty[yl].s1 - ty[y0].s0 + 1);
      }
      // source window of k_pyr_resize4t's 128 x 64 destination tiles: from the 16-byte boundary at or before the first tap of the tile to
      // the last of the three words its last 4-column group loads; rows from the first row's s0 to the last row's s1
      for (int x0 = 0; x0 < L.w; x0 += kRs4TileW) {
        const int xg = ((std::min(x0 + kRs4TileW, wpad) - 1) & ~3);
        rsBoxW = std::max(rsBoxW, ((tx[xg].s0 & ~3) + 12) - (tx[x0].s0 & ~15));
      }
      for (int y0 = 0; y0 < L.h; y0 += 8 * kRs4Rows) {
        const int yl = std::min(y0 + 8 * kRs4Rows, L.h) - 1;
        rsBoxH = std::max(rsBoxH, ty[yl].s1 - ty[y0].s0 + 1);
      }
    }
  }
  rsBoxW = round_up(std::max(rsBoxW, 16), 16);
  h->rsBoxW = (rsBoxW <= 256 && rsBoxH <= 256 && rsBoxH >= 1) ? rsBoxW : 0;
  h->rsBoxH = h->rsBoxW ? rsBoxH : 0;
  g.rsWinWords = round_up(std::max(g.rsWinWords, 4), 4);
  g.rsWinRows = std::max(g.rsWinRows, 1);
  h->rsSmem = (size_t)g.rsWinRows * (g.rsWinWords + kRsTileW) * 4;
  if (h->rsSmem > 200 * 1024) { set_error("scale factor too large for the resize kernel's shared memory"); return ORBX_ERR_UNSUPPORTED; }
  CK(cudaFuncSetAttribute(k_pyr_resize2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->rsSmem));
  g.nCellsTotal = cellBase; g.nTilesTotal = tileBase; g.nBlurTilesTotal = blurTileBase;
  g.candSlab = candOff; g.kpSlab = kpOff; g.kpStride = kpOff;
  h->pyrFrame = round_up(std::max(pyrOff, 256), 256);
  h->blurFrame = round_up(blurOff, 256);
  h->qtSmem = (size_t)maxNodeCap * 18 * sizeof(int);
  if (h->qtSmem > 200 * 1024) { set_error("nfeatures too large for the quadtree kernel's shared memory"); return ORBX_ERR_UNSUPPORTED; }
  CK(cudaFuncSetAttribute(k_quadtree, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->qtSmem));

  // per-CTA tile tables (level | tile x << 4 | tile y << 18): the kernels that cover all levels in one launch look their tile up
  // instead of searching the level list and dividing (~50 instructions per warp in profiles/r01d)
  std::vector<uint32_t> tiles;
  for (int l = 0; l < nl; l++)
    for (int t = 0; t < g.lv[l].tilesX * g.lv[l].tilesY; t++)
      tiles.push_back((uint32_t)l | ((uint32_t)(t % g.lv[l].tilesX) << 4) | ((uint32_t)(t / g.lv[l].tilesX) << 18));
  h->blurTilesOff = (int)tiles.size();
  for (int l = 0; l < nl; l++)
    for (int t = 0; t < g.lv[l].blurTilesX * g.lv[l].blurTilesY; t++)
      tiles.push_back((uint32_t)l | ((uint32_t)(t % g.lv[l].blurTilesX) << 4) | ((uint32_t)(t / g.lv[l].blurTilesX) << 18));
  CK(cudaMalloc(&h->dTiles, std::max<size_t>(1, tiles.size()) * sizeof(uint32_t)));
  if (!tiles.empty()) CK(cudaMemcpy(h->dTiles, tiles.data(), tiles.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
  CK(cudaMalloc(&h->dRsx, std::max<size_t>(1, rsx.size()) * sizeof(RsX)));
  CK(cudaMalloc(&h->dRsy, std::max<size_t>(1, rsy.size()) * sizeof(RsY)));
  CK(cudaMalloc(&h->dInfo, info.size() * sizeof(uint16_t)));
  if (!rsx.empty()) CK(cudaMemcpy(h->dRsx, rsx.data(), rsx.size() * sizeof(RsX), cudaMemcpyHostToDevice));
  if (!rsy.empty()) CK(cudaMemcpy(h->dRsy, rsy.data(), rsy.size() * sizeof(RsY), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(h->dInfo, info.data(), info.size() * sizeof(uint16_t), cudaMemcpyHostToDevice));

  const size_t B = (size_t)h->p.max_batch;
  for (Slot& s : h->slot) {
    CK(cudaMalloc(&s.in, B * h->inPitch * height));
    CK(cudaMalloc(&s.pyr, B * h->pyrFrame + 64));   // + slack: k_pyr_resize4 reads up to 11 bytes past a row's last tap
    CK(cudaMalloc(&s.blur, B * h->blurFrame));
    CK(cudaMalloc(&s.cand, B * g.candSlab * sizeof(PreCand)));
    s.zeroBytes = (B * ORBX_MAX_LEVELS + B * g.nCellsTotal + 16) * sizeof(int);
    CK(cudaMalloc(&s.zero, s.zeroBytes));
    s.candCount = s.zero; s.cellFlags = s.zero + B * ORBX_MAX_LEVELS; s.status = s.cellFlags + B * g.nCellsTotal;
    CK(cudaMalloc(&s.lkp, B * g.kpSlab * sizeof(LevelKp)));
    CK(cudaMalloc(&s.lkpCount, B * ORBX_MAX_LEVELS * sizeof(int)));
    CK(cudaMalloc(&s.nCand, B * ORBX_MAX_LEVELS * sizeof(int)));
    CK(cudaMalloc(&s.lkpDest, B * g.kpSlab * sizeof(int)));
    CK(cudaMalloc(&s.qXY, B * g.candSlab * sizeof(uint32_t)));
    CK(cudaMalloc(&s.qOrd, B * g.candSlab * sizeof(uint32_t)));
    CK(cudaMalloc(&s.qNode, B * g.candSlab * sizeof(uint16_t)));
    CK(cudaMalloc(&s.qResp, B * g.candSlab));
    CK(cudaMalloc(&s.qSel, B * g.candSlab));
    CK(cudaMalloc(&s.kps, B * g.kpStride * sizeof(orbx_keypoint)));
    CK(cudaMalloc(&s.desc, B * g.kpStride * 32));
    CK(cudaMalloc(&s.counts, B * sizeof(int)));
    CK(cudaMalloc(&s.mono, B * sizeof(int)));
    CK(cudaMallocHost(&s.hStatus, sizeof(int)));
    s.lastFrames = 0;
  }
  h->haveGeom = true;
  return ORBX_OK;
}

// Enqueue one device batch on `st` using the workspaces of slot `si`.
namespace {
// Tensor maps of the FAST kernel's TMA staging (fast_nms3.cuh): one per level over (x, y, frame) of the un-blurred level, u8, box
// 96 x 70 x 1, no swizzle, zero fill outside.  cuTensorMapEncodeTiled comes from the driver through the runtime (no -lcuda).
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q = cudaDriverEntryPointSymbolNotFound;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) { cudaGetLastError(); p = nullptr; }
    return reinterpret_cast<EncodeTiledFn>(p);
  }();
  return fn;
}
// false when the level cannot be described (base or strides not 16-byte aligned, or no driver entry point): the caller then launches the per-lane path
bool encode_level_map(void* out, const uint8_t* base, int w, int h, int pitch, long long frameStride, int nframes, int boxW = orbx::kF3RawW, int boxH = orbx::kF3WinH) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn || ((uintptr_t)base & 15) || (pitch & 15) || (frameStride & 15) || w < 1 || h < 1 || nframes < 1) return false;
  static_assert(sizeof(CUtensorMap) == 128, "CUtensorMap size");
  const cuuint64_t dims[3] = {(cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)nframes};
  const cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)(nframes > 1 ? frameStride : (long long)pitch * h)};
  const cuuint32_t box[3] = {(cuuint32_t)boxW, (cuuint32_t)boxH, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  CUtensorMap m;
  if (fn(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<uint8_t*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
         CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) return false;
  memcpy(out, &m, sizeof m);
  return true;
}
}  // namespace

int orbx::extractor_run_batch(orbx_extractor* h, int si, const uint8_t* dIn, int nframes, int pitch, long long frameStride, int lap0, int lap1,
                     cudaStream_t st) {
  // the calling host thread may never have selected this handle's device (a worker thread of a rank that owns GPU 1 starts on GPU 0:
  // launches into the handle's streams then fail with "invalid resource handle")
  CK(cudaSetDevice(h->p.device));
  Slot& s = h->slot[si];
  const Geom& g = h->g;
  Geom gg = g;
  gg.lv[0].pitch = pitch;
  Batch b{};
  b.in = dIn; b.inFrameStride = frameStride; b.inPitch = pitch; b.nframes = nframes; b.lap0 = lap0; b.lap1 = lap1;
  b.pyr = s.pyr; b.pyrFrameStride = h->pyrFrame; b.blur = s.blur; b.blurFrameStride = h->blurFrame;
  b.cand = s.cand; b.candCount = s.candCount; b.cellFlags = s.cellFlags;
  b.lkp = s.lkp; b.lkpCount = s.lkpCount; b.lkpDest = s.lkpDest;
  b.qXY = s.qXY; b.qOrd = s.qOrd; b.qNode = s.qNode; b.qResp = s.qResp; b.qSel = s.qSel;
  b.kps = s.kps; b.desc = s.desc; b.counts = s.counts; b.mono = s.mono; b.status = s.status;
  b.rsx = h->dRsx; b.rsy = h->dRsy; b.info = h->dInfo; b.fastTiles = h->dTiles; b.blurTiles = h->dTiles + h->blurTilesOff; b.pattern = h->dPattern; b.patternF = reinterpret_cast<const float4*>(h->dPattern + 1024);
  s.last = b; s.lastFrames = nframes;
  h->lastSlot = si;

  const bool prof = h->profile;
  if (prof && !s.ev[0])
    for (auto& e : s.ev) CK(cudaEventCreate(&e));
  s.evValid = false;
  auto mark = [&](int i) { if (prof) cudaEventRecord(s.ev[i], st); };
  mark(0);
  CK(cudaMemsetAsync(s.zero, 0, s.zeroBytes, st));
  mark(1);
  // TMA-fed strips when every source level has a tensor map with the window box (encoded once per slot and geometry, level 0 per call)
  bool rsTma = false;
  if (h->rs3ok && h->rsBoxW > 0 && g.nlevels > 1 && g.nlevels <= kFastMapLevels) {
    if (s.rmapsEpoch != h->geomEpoch + 1) {
      s.rmapsOk = true;
      for (int l = 1; l + 1 < g.nlevels && s.rmapsOk; l++)
        s.rmapsOk = encode_level_map(s.rmaps.m[l], s.pyr + g.lv[l].pyrOff, g.lv[l].w, g.lv[l].h, g.lv[l].pitch, h->pyrFrame, h->p.max_batch, h->rsBoxW, h->rsBoxH);
      s.rmapsEpoch = h->geomEpoch + 1;
    }
    rsTma = s.rmapsOk && encode_level_map(s.rmaps.m[0], dIn, g.lv[0].w, g.lv[0].h, pitch, frameStride, nframes, h->rsBoxW, h->rsBoxH);
  }
  for (int l = 1; l < g.nlevels; l++) {
    const LevelGeom& L = g.lv[l];
    if (rsTma) {
      dim3 grid((L.w + kRs4TileW - 1) / kRs4TileW, (L.h + 8 * kRs4Rows - 1) / (8 * kRs4Rows), nframes);
      k_pyr_resize4t<<<grid, 256, (size_t)h->rsBoxW * h->rsBoxH, st>>>(gg, b, s.rmaps, l, h->rsBoxW, h->rsBoxH);
    } else if (h->rs3ok) {
      dim3 grid((L.w + kRs4TileW - 1) / kRs4TileW, (L.h + 8 * kRs4Rows - 1) / (8 * kRs4Rows), nframes);
      if (l == 1) k_pyr_resize4<true><<<grid, 256, 0, st>>>(gg, b, l);     // source = the caller's image: no bytes may be read past its rows
      else k_pyr_resize4<false><<<grid, 256, 0, st>>>(gg, b, l);
    } else {   // scale factor > 2: a 4-column group spans more than 8 source bytes -> staged-window kernel
      dim3 grid((L.w + kRsTileW - 1) / kRsTileW, (L.h + kRsTileH - 1) / kRsTileH, nframes);
      k_pyr_resize2<<<grid, 256, h->rsSmem, st>>>(gg, b, l);
    }
  }
  mark(2);
  if (g.nTilesTotal > 0) {
    // TMA staging when every level can be described by a tensor map (16-byte aligned base / pitch / frame stride: always for the staged
    // host path and the usual image widths); the maps of the slab levels are encoded once per slot and geometry, level 0 per call
    if (s.mapsEpoch != h->geomEpoch + 1) {
      s.mapsOk = g.nlevels <= kFastMapLevels;
      for (int l = 1; l < g.nlevels && s.mapsOk; l++)
        s.mapsOk = encode_level_map(s.maps.m[l], s.pyr + g.lv[l].pyrOff, g.lv[l].w, g.lv[l].h, g.lv[l].pitch, h->pyrFrame, h->p.max_batch);
      s.omapsOk = s.mapsOk;
      for (int l = 0; l < g.nlevels && s.omapsOk; l++) {
        s.omapsOk = encode_level_map(s.omaps.blur[l], s.blur + g.lv[l].blurOff, g.lv[l].w, g.lv[l].h, g.lv[l].blurPitch, h->blurFrame, h->p.max_batch, kOdBoxW, kOdRows);
        if (l > 0 && s.omapsOk)
          s.omapsOk = encode_level_map(s.omaps.raw[l], s.pyr + g.lv[l].pyrOff, g.lv[l].w, g.lv[l].h, g.lv[l].pitch, h->pyrFrame, h->p.max_batch, kOmBoxW, kPatch);
      }
      s.mapsEpoch = h->geomEpoch + 1;
    }
    const bool tma = s.mapsOk && encode_level_map(s.maps.m[0], dIn, g.lv[0].w, g.lv[0].h, pitch, frameStride, nframes);
    if (tma) k_fast_nms3<true><<<dim3(g.nTilesTotal, nframes), 256, 0, st>>>(gg, b, s.maps);
    else k_fast_nms3<false><<<dim3(g.nTilesTotal, nframes), 256, 0, st>>>(gg, b, s.maps);
  }
  mark(3);
  k_quadtree<<<dim3(g.nlevels, nframes), 256, h->qtSmem, st>>>(gg, b, s.nCand);
  mark(4);
  k_blur7v<<<dim3(g.nBlurTilesTotal, nframes), 256, 0, st>>>(gg, b);
  mark(5);
  k_assemble<<<nframes, 256, 0, st>>>(gg, b);
  mark(6);
  if (g.nTilesTotal > 0 && s.omapsOk && encode_level_map(s.omaps.raw[0], dIn, g.lv[0].w, g.lv[0].h, pitch, frameStride, nframes, kOmBoxW, kPatch))
    k_orient_desc2<true><<<dim3((g.kpSlab + 7) / 8, nframes), 256, 0, st>>>(gg, b, s.omaps);
  else k_orient_desc2<false><<<dim3((g.kpSlab + 7) / 8, nframes), 256, 0, st>>>(gg, b, s.omaps);
  mark(7);
  s.evValid = prof;
  CK(cudaGetLastError());
  {
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    s.evBatchValid = false;
    s.lastStream = st;
    if (cudaStreamIsCapturing(st, &cap) == cudaSuccess && cap == cudaStreamCaptureStatusNone) {
      if (!s.evBatch) CK(cudaEventCreateWithFlags(&s.evBatch, cudaEventDisableTiming));
      CK(cudaEventRecord(s.evBatch, st));
      s.evBatchValid = true;
    }
  }
  return ORBX_OK;
}

extern "C" {

const char* orbx_last_error(void) { return g_err.c_str(); }

int orbx_host_alloc(size_t bytes, void** ptr) {
  if (!ptr || bytes == 0) return ORBX_ERR_ARG;
  *ptr = nullptr;
  CK(cudaHostAlloc(ptr, bytes, cudaHostAllocPortable));
  return ORBX_OK;
}
int orbx_host_free(void* ptr) {
  if (!ptr) return ORBX_OK;
  CK(cudaFreeHost(ptr));
  return ORBX_OK;
}
const char* orbx_version(void) { return "orbx 0.1 (sm_100a)"; }

int orbx_create(const orbx_params* p, orbx_extractor** out) {
  if (!p || !out) return ORBX_ERR_ARG;
  if (p->nfeatures < 1 || p->nlevels < 1 || p->nlevels > ORBX_MAX_LEVELS || !(p->scale_factor > 1.0f) || p->min_th_fast < 1 ||
      p->ini_th_fast < p->min_th_fast || p->ini_th_fast > 254 || p->max_batch < 1 || p->max_batch > 65535 ||
      (p->blur_taps != ORBX_TAPS_OCV346 && p->blur_taps != ORBX_TAPS_OCV4)) {
    set_error("orbx_create: bad parameters");
    return ORBX_ERR_ARG;
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= p->device || p->device < 0) {
    set_error("orbx_create: no usable CUDA device %d (this library has no CPU fallback)", p->device);
    return ORBX_ERR_CUDA;
  }
  CK(cudaSetDevice(p->device));
  orbx_extractor* h = new orbx_extractor();
  h->p = *p;
  const int nl = p->nlevels;
  // ORBextractor::ORBextractor, src/ORBextractor.cc:408-468
  h->scaleFactorD = (double)p->scale_factor;  // `double scaleFactor` member initialised from the float argument
  h->scale.resize(nl); h->invScale.resize(nl); h->sigma2.resize(nl); h->invSigma2.resize(nl); h->quota.resize(nl);
  h->scale[0] = 1.0f; h->sigma2[0] = 1.0f;
  for (int i = 1; i < nl; i++) {
    h->scale[i] = (float)(h->scale[i - 1] * h->scaleFactorD);
    h->sigma2[i] = h->scale[i] * h->scale[i];
  }
  for (int i = 0; i < nl; i++) { h->invScale[i] = 1.0f / h->scale[i]; h->invSigma2[i] = 1.0f / h->sigma2[i]; }
  float factor = (float)(1.0f / h->scaleFactorD);
  float nDesired = p->nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nl));
  int sum = 0;
  for (int l = 0; l < nl - 1; l++) {
    h->quota[l] = cv_round_f(nDesired);
    sum += h->quota[l];
    nDesired *= factor;
  }
  h->quota[nl - 1] = std::max(p->nfeatures - sum, 0);
  {  // umax, src/ORBextractor.cc:452-467
    int v, v0, vmax = cv_floor_d(kHalfPatch * sqrt(2.f) / 2 + 1);
    int vmin = (int)ceil(kHalfPatch * sqrt(2.f) / 2);
    const double hp2 = kHalfPatch * kHalfPatch;
    for (v = 0; v <= vmax; ++v) h->umax[v] = (int)lrint(sqrt(hp2 - v * v));
    for (v = kHalfPatch, v0 = 0; v >= vmin; --v) {
      while (h->umax[v0] == h->umax[v0 + 1]) ++v0;
      h->umax[v] = v0;
      ++v0;
    }
  }
  for (Slot& s : h->slot)
    if (cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking) != cudaSuccess) {
      delete h; set_error("stream creation failed"); return ORBX_ERR_CUDA;
    }
  float patF[1024];   // pair 8 * lane + j  ->  record j * 32 + lane (k_orient_desc2 reads [j][lane])
  for (int pr = 0; pr < 256; pr++)
    for (int k = 0; k < 4; k++) patF[((pr & 7) * 32 + (pr >> 3)) * 4 + k] = (float)kPatternHost[pr * 4 + k];
  for (int v = 0; v <= kHalfPatch; v++)
    if (h->umax[v] != od_umax(v)) { delete h; set_error("umax table differs from the compiled-in one"); return ORBX_ERR_UNSUPPORTED; }
  if (cudaMalloc(&h->dPattern, 1024 + sizeof(patF)) != cudaSuccess ||
      cudaMemcpy(h->dPattern, kPatternHost, 1024, cudaMemcpyHostToDevice) != cudaSuccess ||
      cudaMemcpy(h->dPattern + 1024, patF, sizeof(patF), cudaMemcpyHostToDevice) != cudaSuccess) {
    delete h; set_error("pattern upload failed"); return ORBX_ERR_CUDA;
  }
  *out = h;
  return ORBX_OK;
}

void orbx_destroy(orbx_extractor* h) {
  if (!h) return;
  free_geom(h);
  for (Slot& s : h->slot) {
    if (s.stream) cudaStreamDestroy(s.stream);
    if (s.evBatch) cudaEventDestroy(s.evBatch);
  }
  if (h->dPattern) cudaFree(h->dPattern);
  if (h->upStream) cudaStreamDestroy(h->upStream);
  for (int k = 0; k < 2; k++) { if (h->evUp[k]) cudaEventDestroy(h->evUp[k]); if (h->evInFree[k]) cudaEventDestroy(h->evInFree[k]); }
  if (h->hRing) cudaFreeHost(h->hRing);
  if (h->lat.exec) cudaGraphExecDestroy(h->lat.exec);
  if (h->lat.hIn) cudaFreeHost(h->lat.hIn);
  if (h->lat.hOut) cudaFreeHost(h->lat.hOut);
  delete h;
}

int orbx_get_levels(const orbx_extractor* h) { return h ? h->p.nlevels : ORBX_ERR_ARG; }
float orbx_get_scale_factor(const orbx_extractor* h) { return h ? (float)h->scaleFactorD : 0.f; }
int orbx_get_scale_tables(const orbx_extractor* h, float* scale, float* inv, float* s2, float* is2) {
  if (!h) return ORBX_ERR_ARG;
  for (int i = 0; i < h->p.nlevels; i++) {
    if (scale) scale[i] = h->scale[i];
    if (inv) inv[i] = h->invScale[i];
    if (s2) s2[i] = h->sigma2[i];
    if (is2) is2[i] = h->invSigma2[i];
  }
  return ORBX_OK;
}
int orbx_get_features_per_level(const orbx_extractor* h, int* quota) {
  if (!h || !quota) return ORBX_ERR_ARG;
  for (int i = 0; i < h->p.nlevels; i++) quota[i] = h->quota[i];
  return ORBX_OK;
}
int orbx_max_keypoints(const orbx_extractor* h, int width, int height) {
  if (!h || width <= 0 || height <= 0) return ORBX_ERR_ARG;
  int rc = extractor_build_geom(const_cast<orbx_extractor*>(h), width, height);
  if (rc != ORBX_OK) return rc;
  return h->g.kpStride;
}
int orbx_set_profiling(orbx_extractor* h, int enable) {
  if (!h) return ORBX_ERR_ARG;
  h->profile = enable != 0;
  return ORBX_OK;
}
int orbx_stage_times(orbx_extractor* h, float* ms, const char** names) {
  static const char* kNames[7] = {"memset", "k_pyr_resize(x levels-1)", "k_fast_nms", "k_quadtree", "k_blur7", "k_assemble", "k_orient_desc"};
  if (!h || !h->haveGeom) return ORBX_ERR_ARG;
  Slot& s = h->slot[h->lastSlot];
  if (!s.evValid) { set_error("no profiled batch (call orbx_set_profiling(h,1) and run a batch)"); return ORBX_ERR_ARG; }
  CK(cudaEventSynchronize(s.ev[7]));
  for (int i = 0; i < 7; i++) {
    if (ms) CK(cudaEventElapsedTime(&ms[i], s.ev[i], s.ev[i + 1]));
    if (names) names[i] = kNames[i];
  }
  return 7;
}
int orbx_launches_per_batch(const orbx_extractor* h) {
  if (!h || !h->haveGeom) return ORBX_ERR_ARG;
  return (h->g.nlevels - 1) + 5;  // resize per level + fast + quadtree + blur + assemble + orient/desc (memset excluded)
}

int orbx_extract_batch_device(orbx_extractor* h, const uint8_t* dImages, int nframes, int width, int height, int pitch,
                              size_t frameStride, int lap0, int lap1, void* stream) {
  return orbx_extract_batch_device_slot(h, 0, dImages, nframes, width, height, pitch, frameStride, lap0, lap1, stream);
}

int orbx_extract_batch_device_slot(orbx_extractor* h, int slot, const uint8_t* dImages, int nframes, int width, int height, int pitch,
                                   size_t frameStride, int lap0, int lap1, void* stream) {
  if (!h || slot < 0 || slot > 1) return ORBX_ERR_ARG;
  if (!dImages || width <= 0 || height <= 0 || nframes <= 0) return ORBX_EMPTY_IMAGE;
  if (nframes > h->p.max_batch) { set_error("nframes %d > max_batch %d", nframes, h->p.max_batch); return ORBX_ERR_ARG; }
  if (pitch < width || (pitch & 3) || ((uintptr_t)dImages & 3) || (frameStride & 3)) {
    set_error("device frames must be 4-byte aligned with a pitch that is a multiple of 4");
    return ORBX_ERR_ARG;
  }
  int rc = extractor_build_geom(h, width, height);
  if (rc != ORBX_OK) return rc;
  cudaStream_t st = stream ? (cudaStream_t)stream : h->slot[slot].stream;
  return extractor_run_batch(h, slot, dImages, nframes, pitch, (long long)frameStride, lap0, lap1, st);
}

int orbx_batch_results_device(orbx_extractor* h, const orbx_keypoint** kps, const uint8_t** desc, const int** counts, const int** mono,
                              const int** status, int* kpStride) {
  if (!h || !h->haveGeom) return ORBX_ERR_ARG;
  Slot& s = h->slot[h->lastSlot];
  if (kps) *kps = s.kps;
  if (desc) *desc = s.desc;
  if (counts) *counts = s.counts;
  if (mono) *mono = s.mono;
  if (status) *status = s.status;
  if (kpStride) *kpStride = h->g.kpStride;
  return ORBX_OK;
}

int orbx_batch_results_host(orbx_extractor* h, orbx_keypoint* kps, uint8_t* desc, int capPerFrame, int* counts, int* mono) {
  if (!h || !h->haveGeom) return ORBX_ERR_ARG;
  Slot& s = h->slot[h->lastSlot];
  const int nb = s.lastFrames;
  if (nb <= 0) return ORBX_ERR_ARG;
  const Geom& g = h->g;
  CK(cudaSetDevice(h->p.device));
  CK(cudaDeviceSynchronize());
  const int ncopy = std::min(capPerFrame, g.kpStride);
  if (counts) CK(cudaMemcpy(counts, s.counts, nb * sizeof(int), cudaMemcpyDeviceToHost));
  if (mono) CK(cudaMemcpy(mono, s.mono, nb * sizeof(int), cudaMemcpyDeviceToHost));
  if (kps) CK(cudaMemcpy2D(kps, (size_t)capPerFrame * sizeof(orbx_keypoint), s.kps, (size_t)g.kpStride * sizeof(orbx_keypoint),
                           (size_t)ncopy * sizeof(orbx_keypoint), nb, cudaMemcpyDeviceToHost));
  if (desc) CK(cudaMemcpy2D(desc, (size_t)capPerFrame * 32, s.desc, (size_t)g.kpStride * 32, (size_t)ncopy * 32, nb, cudaMemcpyDeviceToHost));
  int st = 0;
  CK(cudaMemcpy(&st, s.status, sizeof(int), cudaMemcpyDeviceToHost));
  if (st != 0) set_error("device reported status %d (capacity)", st);
  return st;
}

int orbx_extract_batch_host(orbx_extractor* h, const uint8_t* images, int nframes, int width, int height, int stride, size_t frameStride,
                            int lap0, int lap1, orbx_keypoint* kps, uint8_t* desc, int capPerFrame, int* counts, int* mono) {
  if (!h) return ORBX_ERR_ARG;
  if (!images || width <= 0 || height <= 0 || nframes <= 0) return ORBX_EMPTY_IMAGE;
  if (!kps || !desc || !counts || !mono || capPerFrame <= 0 || stride < width) return ORBX_ERR_ARG;
  CK(cudaSetDevice(h->p.device));   // every call: the steady state below creates nothing, but its copies / events / launches need the handle's device current
  int rc = extractor_build_geom(h, width, height);
  if (rc != ORBX_OK) return rc;
  const Geom& g = h->g;
  const int B = h->p.max_batch;
  const int ncopy = std::min(capPerFrame, g.kpStride);
  // Latency path: a call that fits one small chunk (the reference's per-frame operator(), src/Frame.cc:469-483) pays for launches
  // and for synchronous pageable copies, not for arithmetic.  Frames go through a page-locked buffer and upload + the 12 kernels +
  // the downloads are ONE graph launch; results are handed out from page-locked memory after a single synchronisation.
  const size_t inBytes = (size_t)h->inPitch * height * nframes;
  if (nframes <= B && inBytes <= (size_t)(4 << 20) && !h->profile) {
    orbx_extractor::LatencyPath& L = h->lat;
    Slot& s = h->slot[0];
    const size_t kpB = (size_t)nframes * g.kpStride * sizeof(orbx_keypoint), dB = (size_t)nframes * g.kpStride * 32;
    const size_t oCounts = 0, oMono = (size_t)nframes * 4, oStatus = oMono + (size_t)nframes * 4, oKps = (oStatus + 4 + 63) & ~(size_t)63, oDesc = oKps + kpB;
    const size_t outBytes = oDesc + dB;
    const bool hit = L.exec && L.nb == nframes && L.lap0 == lap0 && L.lap1 == lap1 && L.epoch == h->geomEpoch && L.inBytes >= inBytes && L.outBytes >= outBytes;
    if (!hit) {
      std::lock_guard<std::recursive_mutex> setupLock(setup_mutex());
      CK(cudaSetDevice(h->p.device));
      for (Slot& o : h->slot) CK(cudaStreamSynchronize(o.stream));
      if (L.exec) { cudaGraphExecDestroy(L.exec); L.exec = nullptr; }
      if (L.inBytes < inBytes) { if (L.hIn) cudaFreeHost(L.hIn); L.hIn = nullptr; L.inBytes = 0; CK(cudaHostAlloc((void**)&L.hIn, inBytes, cudaHostAllocDefault)); L.inBytes = inBytes; }
      if (L.outBytes < outBytes) { if (L.hOut) cudaFreeHost(L.hOut); L.hOut = nullptr; L.outBytes = 0; CK(cudaHostAlloc((void**)&L.hOut, outBytes, cudaHostAllocDefault)); L.outBytes = outBytes; }
      cudaGraph_t graph = nullptr;
      CK(cudaStreamBeginCapture(s.stream, cudaStreamCaptureModeThreadLocal));
      cudaError_t ce = cudaMemcpyAsync(s.in, L.hIn, inBytes, cudaMemcpyHostToDevice, s.stream);
      int brc = ce == cudaSuccess ? extractor_run_batch(h, 0, s.in, nframes, h->inPitch, (long long)h->inPitch * height, lap0, lap1, s.stream) : ORBX_ERR_CUDA;
      if (brc == ORBX_OK) {
        const struct { size_t off; const void* src; size_t bytes; } cp[] = {{oCounts, s.counts, (size_t)nframes * 4}, {oMono, s.mono, (size_t)nframes * 4},
                                                                             {oStatus, s.status, 4}, {oKps, s.kps, kpB}, {oDesc, s.desc, dB}};
        for (const auto& c : cp)
          if (ce == cudaSuccess) ce = cudaMemcpyAsync(L.hOut + c.off, c.src, c.bytes, cudaMemcpyDeviceToHost, s.stream);
      }
      const cudaError_t ee = cudaStreamEndCapture(s.stream, &graph);
      if (brc != ORBX_OK) { if (graph) cudaGraphDestroy(graph); return brc; }
      if (ce != cudaSuccess || ee != cudaSuccess) { if (graph) cudaGraphDestroy(graph); set_error("latency-path capture failed: %s", cudaGetErrorString(ce != cudaSuccess ? ce : ee)); return ORBX_ERR_CUDA; }
      CK(cudaGraphInstantiate(&L.exec, graph, 0));
      CK(cudaGraphDestroy(graph));
      L.nb = nframes; L.lap0 = lap0; L.lap1 = lap1; L.epoch = h->geomEpoch;
    } else {
      s.lastFrames = nframes; h->lastSlot = 0; s.evValid = false;   // the bookkeeping the captured call made
    }
    for (int f = 0; f < nframes; f++) {   // rows into the staged layout [f][height][inPitch]
      const uint8_t* src = images + (size_t)f * frameStride;
      uint8_t* dst = L.hIn + (size_t)f * h->inPitch * height;
      if (stride == h->inPitch) memcpy(dst, src, (size_t)stride * height);
      else for (int y = 0; y < height; y++) memcpy(dst + (size_t)y * h->inPitch, src + (size_t)y * stride, width);
    }
    CK(cudaGraphLaunch(L.exec, s.stream));
    CK(cudaStreamSynchronize(s.stream));
    const int st = *reinterpret_cast<const int*>(L.hOut + oStatus);
    if (st != 0) { set_error("device reported status %d (capacity)", st); return st; }
    memcpy(counts, L.hOut + oCounts, (size_t)nframes * 4);
    memcpy(mono, L.hOut + oMono, (size_t)nframes * 4);
    for (int f = 0; f < nframes; f++) {
      if (counts[f] > capPerFrame) { set_error("frame %d has %d keypoints, capacity %d", f, counts[f], capPerFrame); return ORBX_ERR_CAPACITY; }
      const int c = std::min(counts[f], ncopy);
      memcpy(kps + (size_t)f * capPerFrame, L.hOut + oKps + (size_t)f * g.kpStride * sizeof(orbx_keypoint), (size_t)c * sizeof(orbx_keypoint));
      memcpy(desc + (size_t)f * capPerFrame * 32, L.hOut + oDesc + (size_t)f * g.kpStride * 32, (size_t)c * 32);
    }
    return ORBX_OK;
  }
  // pipelined path (many frames).  Uploads run back to back on their own stream; a slot's kernels wait for their upload on the device and
  // release the slot's input buffer with an event when they are done with it (level 0 is read until the descriptor stage); downloads follow
  // the kernels on the slot stream; the host only waits at the end (or when the status ring wraps).
  if (!h->upStream) {
    std::lock_guard<std::recursive_mutex> setupLock(setup_mutex());
    CK(cudaSetDevice(h->p.device));
    CK(cudaStreamCreateWithFlags(&h->upStream, cudaStreamNonBlocking));
    for (int k = 0; k < 2; k++) {
      CK(cudaEventCreateWithFlags(&h->evUp[k], cudaEventDisableTiming));
      CK(cudaEventCreateWithFlags(&h->evInFree[k], cudaEventDisableTiming));
    }
    CK(cudaHostAlloc((void**)&h->hRing, orbx_extractor::kRing * sizeof(int), cudaHostAllocDefault));
  }
  // on every return below, nothing may still be writing into the caller's buffers
  struct Drain { orbx_extractor* h; ~Drain() { cudaStreamSynchronize(h->upStream); for (Slot& s : h->slot) cudaStreamSynchronize(s.stream); } } drain{h};
  auto check_ring = [&](int used) -> int {
    for (Slot& s : h->slot) CK(cudaStreamSynchronize(s.stream));
    for (int k = 0; k < used; k++)
      if (h->hRing[k] != 0) { set_error("device reported status %d (capacity)", h->hRing[k]); return h->hRing[k]; }
    return ORBX_OK;
  };
  int chunk = 0;
  for (int f0 = 0; f0 < nframes; f0 += B, chunk++) {
    const int nb = std::min(B, nframes - f0);
    const int si = chunk & 1;
    Slot& s = h->slot[si];
    cudaStream_t st = s.stream, up = h->upStream;
    CK(cudaStreamWaitEvent(up, h->evInFree[si], 0));   // the slot's previous kernels no longer read its input buffer
    const uint8_t* src = images + (size_t)f0 * frameStride;
    if (frameStride == (size_t)stride * height && stride == h->inPitch) {
      CK(cudaMemcpyAsync(s.in, src, (size_t)stride * height * nb, cudaMemcpyHostToDevice, up));
    } else if (frameStride == (size_t)stride * height) {
      CK(cudaMemcpy2DAsync(s.in, h->inPitch, src, stride, width, (size_t)height * nb, cudaMemcpyHostToDevice, up));
    } else {
      for (int f = 0; f < nb; f++)
        CK(cudaMemcpy2DAsync(s.in + (size_t)f * h->inPitch * height, h->inPitch, src + (size_t)f * frameStride, stride, width, height,
                             cudaMemcpyHostToDevice, up));
    }
    CK(cudaEventRecord(h->evUp[si], up));
    CK(cudaStreamWaitEvent(st, h->evUp[si], 0));
    // (the slot's previous chunk was enqueued on the same stream, so its D2H copies finish before the result buffers are rewritten)
    rc = extractor_run_batch(h, si, s.in, nb, h->inPitch, (long long)h->inPitch * height, lap0, lap1, st);
    if (rc != ORBX_OK) return rc;
    CK(cudaEventRecord(h->evInFree[si], st));
    CK(cudaMemcpyAsync(counts + f0, s.counts, nb * sizeof(int), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(mono + f0, s.mono, nb * sizeof(int), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpy2DAsync(kps + (size_t)f0 * capPerFrame, (size_t)capPerFrame * sizeof(orbx_keypoint), s.kps,
                         (size_t)g.kpStride * sizeof(orbx_keypoint), (size_t)ncopy * sizeof(orbx_keypoint), nb, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpy2DAsync(desc + (size_t)f0 * capPerFrame * 32, (size_t)capPerFrame * 32, s.desc, (size_t)g.kpStride * 32, (size_t)ncopy * 32, nb,
                         cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(h->hRing + chunk % orbx_extractor::kRing, s.status, sizeof(int), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(s.hStatus, s.status, sizeof(int), cudaMemcpyDeviceToHost, st));
    if ((chunk + 1) % orbx_extractor::kRing == 0 && (rc = check_ring(orbx_extractor::kRing)) != ORBX_OK) return rc;
  }
  if ((rc = check_ring(chunk % orbx_extractor::kRing)) != ORBX_OK) return rc;
  for (int f = 0; f < nframes; f++)
    if (counts[f] > capPerFrame) { set_error("frame %d has %d keypoints, capacity %d", f, counts[f], capPerFrame); return ORBX_ERR_CAPACITY; }
  return ORBX_OK;
}

int orbx_extract(orbx_extractor* h, const uint8_t* image, int width, int height, int stride, int lap0, int lap1, orbx_keypoint* kps,
                 uint8_t* desc, int capacity, int* n, int* monoIndex) {
  if (!h) return ORBX_ERR_ARG;
  if (!image || width <= 0 || height <= 0) return ORBX_EMPTY_IMAGE;
  int cnt = 0, mono = 0;
  int rc = orbx_extract_batch_host(h, image, 1, width, height, stride, (size_t)stride * height, lap0, lap1, kps, desc, capacity, &cnt, &mono);
  if (n) *n = cnt;
  if (monoIndex) *monoIndex = mono;
  return rc;
}

static int level_view(orbx_extractor* h, int frame, int level, const uint8_t** ptr, int* w, int* hh, int* pitch) {
  if (!h || !h->haveGeom) return ORBX_ERR_ARG;
  Slot& s = h->slot[h->lastSlot];
  if (frame < 0 || frame >= s.lastFrames || level < 0 || level >= h->g.nlevels) return ORBX_ERR_ARG;
  const LevelGeom& L = h->g.lv[level];
  if (level == 0) { *ptr = s.last.in + (long long)frame * s.last.inFrameStride; *pitch = s.last.inPitch; }
  else { *ptr = s.pyr + (long long)frame * h->pyrFrame + L.pyrOff; *pitch = L.pitch; }
  *w = L.w; *hh = L.h;
  return ORBX_OK;
}

int orbx_pyramid_level_device(orbx_extractor* h, int frame, int level, const uint8_t** dPtr, int* w, int* hh, int* pitch) {
  const uint8_t* p; int ww, hv, pp;
  int rc = level_view(h, frame, level, &p, &ww, &hv, &pp);
  if (rc != ORBX_OK) return rc;
  if (dPtr) *dPtr = p; if (w) *w = ww; if (hh) *hh = hv; if (pitch) *pitch = pp;
  return ORBX_OK;
}

int orbx_pyramid_level(orbx_extractor* h, int frame, int level, uint8_t* dst, int dstStride, int* w, int* hh) {
  const uint8_t* p; int ww, hv, pp;
  int rc = level_view(h, frame, level, &p, &ww, &hv, &pp);
  if (rc != ORBX_OK) return rc;
  if (w) *w = ww; if (hh) *hh = hv;
  if (dst) {
    if (dstStride < ww) return ORBX_ERR_ARG;
    CK(cudaSetDevice(h->p.device));
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy2D(dst, dstStride, p, pp, ww, hv, cudaMemcpyDeviceToHost));
  }
  return ORBX_OK;
}

int orbx_debug_blurred_level(orbx_extractor* h, int frame, int level, uint8_t* dst, int dstStride) {
  if (!h || !h->haveGeom || !dst) return ORBX_ERR_ARG;
  Slot& s = h->slot[h->lastSlot];
  if (frame < 0 || frame >= s.lastFrames || level < 0 || level >= h->g.nlevels) return ORBX_ERR_ARG;
  const LevelGeom& L = h->g.lv[level];
  CK(cudaSetDevice(h->p.device));
  CK(cudaDeviceSynchronize());
  CK(cudaMemcpy2D(dst, dstStride, s.blur + (long long)frame * h->blurFrame + L.blurOff, L.blurPitch, L.w, L.h, cudaMemcpyDeviceToHost));
  return ORBX_OK;
}

int orbx_debug_candidates(orbx_extractor* h, int frame, int level, int* xyr, int capacity, int* n) {
  if (!h || !h->haveGeom || !n) return ORBX_ERR_ARG;
  Slot& s = h->slot[h->lastSlot];
  if (frame < 0 || frame >= s.lastFrames || level < 0 || level >= h->g.nlevels) return ORBX_ERR_ARG;
  const LevelGeom& L = h->g.lv[level];
  CK(cudaSetDevice(h->p.device));
  CK(cudaDeviceSynchronize());
  int cnt = 0;
  CK(cudaMemcpy(&cnt, s.nCand + frame * ORBX_MAX_LEVELS + level, sizeof(int), cudaMemcpyDeviceToHost));
  *n = cnt;
  const int m = std::min(cnt, capacity);
  if (m > 0 && xyr) {
    std::vector<uint32_t> xy(m);
    std::vector<uint8_t> r(m);
    const long long off = (long long)frame * h->g.candSlab + L.candOff;
    CK(cudaMemcpy(xy.data(), s.qXY + off, m * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(r.data(), s.qResp + off, m, cudaMemcpyDeviceToHost));
    for (int i = 0; i < m; i++) { xyr[3 * i] = xy[i] & 0xffff; xyr[3 * i + 1] = xy[i] >> 16; xyr[3 * i + 2] = r[i]; }
  }
  return ORBX_OK;
}

int orbx_debug_level_keypoints(orbx_extractor* h, int frame, int level, orbx_keypoint* out, int capacity, int* n) {
  if (!h || !h->haveGeom || !n) return ORBX_ERR_ARG;
  Slot& s = h->slot[h->lastSlot];
  if (frame < 0 || frame >= s.lastFrames || level < 0 || level >= h->g.nlevels) return ORBX_ERR_ARG;
  const LevelGeom& L = h->g.lv[level];
  CK(cudaSetDevice(h->p.device));
  CK(cudaDeviceSynchronize());
  int cnt = 0;
  CK(cudaMemcpy(&cnt, s.lkpCount + frame * ORBX_MAX_LEVELS + level, sizeof(int), cudaMemcpyDeviceToHost));
  *n = cnt;
  const int m = std::min(cnt, capacity);
  if (m > 0 && out) {
    std::vector<LevelKp> lk(m);
    std::vector<int> dest(m);
    const long long off = (long long)frame * h->g.kpSlab + L.kpOff;
    CK(cudaMemcpy(lk.data(), s.lkp + off, m * sizeof(LevelKp), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(dest.data(), s.lkpDest + off, m * sizeof(int), cudaMemcpyDeviceToHost));
    for (int i = 0; i < m; i++) {
      orbx_keypoint k;
      CK(cudaMemcpy(&k, s.kps + (long long)frame * h->g.kpStride + dest[i], sizeof k, cudaMemcpyDeviceToHost));
      k.x = lk[i].x; k.y = lk[i].y;
      out[i] = k;
    }
  }
  return ORBX_OK;
}

int orbx_hamming(const uint8_t* a, const uint8_t* b) {
  int d = 0;
  for (int i = 0; i < 4; i++) {
    uint64_t x, y;
    memcpy(&x, a + 8 * i, 8);
    memcpy(&y, b + 8 * i, 8);
    d += __builtin_popcountll(x ^ y);
  }
  return d;
}

}  // extern "C"
