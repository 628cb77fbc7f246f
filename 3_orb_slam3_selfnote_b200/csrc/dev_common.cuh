// Small device helpers shared by the extractor and matcher kernels.
#pragma once
#include <cuda_runtime.h>

#include "geom.h"

namespace orbx {

#define ORBX_FULL 0xffffffffu

// un-blurred pyramid level `level` of frame `frame`: level 0 is the caller's input, levels >= 1 live in the pyramid slab
__device__ __forceinline__ const uint8_t* level_ptr(const Geom& g, const Batch& b, int level, int frame, int& pitch) {
  if (level == 0) {
    pitch = b.inPitch;
    return b.in + (long long)frame * b.inFrameStride;
  }
  pitch = g.lv[level].pitch;
  return b.pyr + (long long)frame * b.pyrFrameStride + g.lv[level].pyrOff;
}

}  // namespace orbx
