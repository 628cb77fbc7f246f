// Internal definitions shared by the translation units of liborbx.so (not part of the C-ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdarg.h>

#include <mutex>
#include <vector>

#include "geom.h"

// tensor maps of the FAST kernel (fast_nms3.cuh), passed as a kernel parameter: 8 levels (the reference's setting) keep the parameter block
// below the classic 4 KB; extractors with more levels use the per-lane staging path
namespace orbx {
constexpr int kFastMapLevels = 8;
struct FastMaps { alignas(64) unsigned char m[kFastMapLevels][128]; };
struct OrientMaps { alignas(64) unsigned char raw[kFastMapLevels][128]; alignas(64) unsigned char blur[kFastMapLevels][128]; };   // k_orient_desc2's patch boxes
}

namespace orbx {
void set_error(const char* fmt, ...);
// Process-wide lock around the rare host-side set-up sections of every handle (allocation / free, function attributes,
// lookup-table uploads on the NULL stream, CUDA graph capture): several host threads may drive their own handles concurrently
// (include/orbx.h), and a capture must not overlap another handle's allocation phase.  The steady state never takes it.
std::recursive_mutex& setup_mutex();

#define CK(call)                                                                                         \
  do {                                                                                                   \
    cudaError_t e_ = (call);                                                                             \
    if (e_ != cudaSuccess) {                                                                             \
      set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_));                    \
      return ORBX_ERR_CUDA;                                                                              \
    }                                                                                                    \
  } while (0)

struct Slot {
  cudaStream_t stream = nullptr;
  uint8_t* in = nullptr;  // staged level-0 frames for the host path [max_batch][h][inPitch]
  uint8_t* pyr = nullptr;
  uint8_t* blur = nullptr;
  PreCand* cand = nullptr;
  int* zero = nullptr;  // candCount | cellFlags | status  (cleared per batch)
  size_t zeroBytes = 0;
  int* candCount = nullptr; int* cellFlags = nullptr; int* status = nullptr;
  LevelKp* lkp = nullptr; int* lkpCount = nullptr; int* lkpDest = nullptr; int* nCand = nullptr;
  uint32_t* qXY = nullptr; uint32_t* qOrd = nullptr; uint16_t* qNode = nullptr; uint8_t* qResp = nullptr; uint8_t* qSel = nullptr;
  orbx_keypoint* kps = nullptr; uint8_t* desc = nullptr; int* counts = nullptr; int* mono = nullptr;
  int* hStatus = nullptr;  // pinned
  Batch last{};            // descriptor of the last batch run in this slot
  int lastFrames = 0;
  cudaEvent_t ev[8] = {};  // stage boundaries (profiling mode only)
  bool evValid = false;
  FastMaps maps{};         // TMA descriptors of the un-blurred levels (levels >= 1 valid while mapsEpoch == geomEpoch + 1)
  unsigned long long mapsEpoch = 0; bool mapsOk = false;
  OrientMaps omaps{}; bool omapsOk = false;
  FastMaps rmaps{}; bool rmapsOk = false; unsigned long long rmapsEpoch = 0;   // k_pyr_resize4t: source windows of the levels 0 .. nlevels-2
  // completion of the last batch enqueued in this slot outside a graph capture: consumers on another stream (orbx_stereo_match_batch)
  // wait for it, so a NULL / different stream there cannot race the extraction
  cudaEvent_t evBatch = nullptr; cudaStream_t lastStream = nullptr; bool evBatchValid = false;
};

}  // namespace orbx

struct orbx_extractor {
  typedef orbx::Geom Geom; typedef orbx::RsX RsX; typedef orbx::RsY RsY; typedef orbx::Slot Slot;
  orbx_params p{};
  double scaleFactorD = 0;
  std::vector<float> scale, invScale, sigma2, invSigma2;
  std::vector<int> quota;
  int umax[16];
  // geometry-dependent state
  bool haveGeom = false;
  unsigned long long geomEpoch = 0;  // bumped whenever the geometry (and with it every workspace pointer) is rebuilt
  Geom g{};
  int inPitch = 0;  // pitch of the staged input
  long long pyrFrame = 0, blurFrame = 0;
  RsX* dRsx = nullptr; RsY* dRsy = nullptr; uint16_t* dInfo = nullptr; uint32_t* dTiles = nullptr; int blurTilesOff = 0; int8_t* dPattern = nullptr;
  size_t qtSmem = 0;
  Slot slot[2];
  int lastSlot = 0;
  bool profile = false;
  int rsBoxW = 0, rsBoxH = 0;   // k_pyr_resize4t: source window of a 128 x 64 destination tile (0 = does not fit a TMA box)
  bool rs3ok = true;     // k_pyr_resize4 applies to this geometry (scale factor <= 2); otherwise k_pyr_resize2
  size_t rsSmem = 0;
  // latency path of the host entry points (one small chunk, e.g. the reference's per-frame operator() call): page-locked staging
  // for the frames and the results, and ONE CUDA graph holding upload + kernels + downloads
  struct LatencyPath {
    uint8_t* hIn = nullptr; size_t inBytes = 0;
    uint8_t* hOut = nullptr; size_t outBytes = 0;
    cudaGraphExec_t exec = nullptr;
    int nb = 0, lap0 = 0, lap1 = 0; unsigned long long epoch = 0;
  } lat;
  // pipelined host path (many chunks per call): uploads run back to back on their own stream, a slot's kernels wait for its upload on
  // the device, its input buffer is released by an event after the kernels; chunk status words land in a page-locked ring
  cudaStream_t upStream = nullptr; cudaEvent_t evUp[2] = {nullptr, nullptr}, evInFree[2] = {nullptr, nullptr};
  int* hRing = nullptr; static constexpr int kRing = 32;
};


struct orbx_matcher;
namespace orbx {
int matcher_device(const orbx_matcher* m);   // defined in orbx_match.cu
// enqueue one extractor batch (defined in orbx_extract.cu)
int extractor_run_batch(orbx_extractor* h, int slot, const uint8_t* dIn, int nframes, int pitch, long long frameStride, int lap0, int lap1,
                        cudaStream_t st);
int extractor_build_geom(orbx_extractor* h, int width, int height);
}  // namespace orbx
