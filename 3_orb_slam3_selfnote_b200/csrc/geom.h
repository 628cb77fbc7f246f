// Geometry / batch descriptors shared by host code and kernels (passed BY VALUE as __grid_constant__ kernel parameters,
// so several handles with different geometry can run concurrently -- no __constant__ globals).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/orbx.h"

namespace orbx {

constexpr int kEdge = 19;         // EDGE_THRESHOLD, src/ORBextractor.cc:72
constexpr int kMinBorder = 16;    // EDGE_THRESHOLD-3, src/ORBextractor.cc:771
constexpr int kHalfPatch = 15;    // HALF_PATCH_SIZE
constexpr int kPatch = 31;        // PATCH_SIZE

struct LevelGeom {
  int w, h, pitch;            // un-blurred level (pitch: bytes per row inside the pyramid slab; level 0 uses the input pitch)
  int pyrOff;                 // byte offset of the level inside a frame's pyramid slab (levels >= 1)
  int blurOff, blurPitch;     // same for the blurred slab (all levels)
  int nCols, nRows, wCell, hCell, cellBase;  // FAST cell grid (src/ORBextractor.cc:779-785), cellBase = first cell id of the level
  int quota;                  // mnFeaturesPerLevel[level]
  int nIni;                   // quadtree roots (src/ORBextractor.cc:541)
  float hX;                   // root width
  int candCap, candOff;       // pre-candidate records of the level inside a frame's candidate slab
  int kpCap, kpOff;           // level keypoints (after the quadtree) inside a frame's level-keypoint slab
  int nodeCap;                // quadtree list capacity
  float scale;                // mvScaleFactor[level]
  float kpSize;               // (float)(int)(PATCH_SIZE*scale)
  int rsxOff, rsyOff;         // first entry of the level's resize tables (destination column / row)
  int colInfoOff, rowInfoOff; // first entry of the level's column / row cell-info tables
  int tilesX, tilesY, tileBase;  // FAST tiles
  int blurTilesX, blurTilesY, blurTileBase;
};

struct Geom {
  int nlevels, width, height;
  int iniTh, minTh;
  int nCellsTotal, nTilesTotal, nBlurTilesTotal;
  int candSlab;    // records per frame
  int kpSlab;      // level keypoints per frame (sum kpCap)
  int kpStride;    // final keypoints per frame (capacity of the output arrays)
  int rsWinWords, rsWinRows;  // k_pyr_resize2: staged source window, maxima over all tiles of all levels (words per row, rows)
  int taps[7];     // blur taps (Q8.8)
  int umax[16];
  LevelGeom lv[ORBX_MAX_LEVELS];
};

// resize tables: per destination column / row
struct RsX { int16_t s0, s1, a0, a1; };  // source columns (s1 clamped) and Q11 weights
struct RsY { int16_t s0, s1, b0, b1; };

// pre-candidate record written by the FAST/NMS kernel: a 3x3-local-max pixel with S > minTh
struct PreCand { uint16_t x, y; uint16_t cell; uint8_t S; uint8_t pad; };  // x,y: level coordinates; cell: level-local cell id

// level keypoint written by the quadtree kernel, list order
struct LevelKp { int16_t x, y; int32_t resp; };

struct Batch {
  // inputs (level 0)
  const uint8_t* in;
  long long inFrameStride;
  int inPitch;
  int nframes;
  int lap0, lap1;
  // per-frame slabs
  uint8_t* pyr;        long long pyrFrameStride;     // levels 1..
  uint8_t* blur;       long long blurFrameStride;    // levels 0..
  PreCand* cand;                                     // [frame][candSlab]
  int* candCount;                                    // [frame][ORBX_MAX_LEVELS]
  int* cellFlags;                                    // [frame][nCellsTotal]
  LevelKp* lkp;                                      // [frame][kpSlab]
  int* lkpCount;                                     // [frame][ORBX_MAX_LEVELS]
  int* lkpDest;                                      // [frame][kpSlab]  final slot of every level keypoint
  // quadtree scratch, [frame][candSlab]
  uint32_t* qXY; uint32_t* qOrd; uint16_t* qNode; uint8_t* qResp; uint8_t* qSel;
  // outputs
  orbx_keypoint* kps;  uint8_t* desc;                // [frame][kpStride]
  int* counts; int* mono;                            // [frame]
  int* status;                                       // single word, 0 = ok
  // tables
  const RsX* rsx; const RsY* rsy;                    // indexed by LevelGeom::rsxOff + x / rsyOff + y
  const uint16_t* info;                              // cell index | first<<14 | last<<15, indexed by colInfoOff + x / rowInfoOff + y
  const uint32_t* fastTiles; const uint32_t* blurTiles;  // per CTA of the all-level launches: level | tile x << 4 | tile y << 18
  const int8_t* pattern;                             // 256 x (x0,y0,x1,y1)
  const float4* patternF;                            // the same as floats (k_orient_desc2)
};

}  // namespace orbx
