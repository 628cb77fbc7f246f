// Kernels of the pre-processing step in front of ORBextractor::operator(): colour -> gray,
// stereo rectification (remap) and CLAHE, each bit-exact with the OpenCV 8-bit path the reference calls
// (src/Tracking.cc:989-1014, Examples/Stereo/stereo_euroc.cc:166-167, Examples/Monocular/mono_tum_vi.cc:101-109).
// All three are streaming byte kernels (HBM-bound): thread = 4 adjacent output pixels, 32-bit stores, frames batched in
// blockIdx.z / .y.  No floating-point contraction anywhere (--fmad=false, explicit _rn intrinsics in the CLAHE blend).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace orbx {

struct GrayArgs {
  const uint8_t* src; uint8_t* dst;
  int w, h, sstride, dstride;
  long long sfs, dfs;
  unsigned c0, c1, c2, half;
  int shift;
  int vec;  // rows and bases aligned for the word path (4 B for 3 channels, 16 B for 4 channels; dst 4 B)
};

__device__ __forceinline__ unsigned gray_px(unsigned a, unsigned b, unsigned c, const GrayArgs& A) {
  return (a * A.c0 + b * A.c1 + c * A.c2 + A.half) >> A.shift;
}

template <int CH>
__global__ void __launch_bounds__(256) k_cvt_gray(const __grid_constant__ GrayArgs A) {
  const int x4 = (blockIdx.x * 256 + threadIdx.x) * 4, y = blockIdx.y;
  if (x4 >= A.w) return;
  const uint8_t* s = A.src + (long long)blockIdx.z * A.sfs + (long long)y * A.sstride + (long long)x4 * CH;
  uint8_t* d = A.dst + (long long)blockIdx.z * A.dfs + (long long)y * A.dstride + x4;
  if (A.vec && x4 + 4 <= A.w) {
    unsigned g0, g1, g2, g3;
    if (CH == 3) {
      const unsigned* p = reinterpret_cast<const unsigned*>(s);
      unsigned w0 = __ldg(p), w1 = __ldg(p + 1), w2 = __ldg(p + 2);
      g0 = gray_px(w0 & 255u, (w0 >> 8) & 255u, (w0 >> 16) & 255u, A);
      g1 = gray_px(w0 >> 24, w1 & 255u, (w1 >> 8) & 255u, A);
      g2 = gray_px((w1 >> 16) & 255u, w1 >> 24, w2 & 255u, A);
      g3 = gray_px((w2 >> 8) & 255u, (w2 >> 16) & 255u, w2 >> 24, A);
    } else {
      uint4 q = __ldg(reinterpret_cast<const uint4*>(s));
      g0 = gray_px(q.x & 255u, (q.x >> 8) & 255u, (q.x >> 16) & 255u, A);
      g1 = gray_px(q.y & 255u, (q.y >> 8) & 255u, (q.y >> 16) & 255u, A);
      g2 = gray_px(q.z & 255u, (q.z >> 8) & 255u, (q.z >> 16) & 255u, A);
      g3 = gray_px(q.w & 255u, (q.w >> 8) & 255u, (q.w >> 16) & 255u, A);
    }
    *reinterpret_cast<unsigned*>(d) = g0 | (g1 << 8) | (g2 << 16) | (g3 << 24);
  } else {
    const int n = min(4, A.w - x4);
    for (int i = 0; i < n; ++i) d[i] = (uint8_t)gray_px(s[i * CH], s[i * CH + 1], s[i * CH + 2], A);
  }
}

// ---- remap ---------------------------------------------------------------------------------------------------------------
// The float maps are converted ONCE (orbx_remap_create) to OpenCV's fixed-point form s = cvRound(map * 32) and packed as
// (sy + 32) << 16 | (sx + 32) after clamping s to [-32, size * 32]: every clamped value still reads only out-of-image taps with
// non-zero weight, so the result is unchanged and a table entry is 4 bytes instead of the 8 bytes of the two float maps.
struct RemapArgs {
  const uint32_t* table; int tpitch;  // entries per row (multiple of 4)
  const uint8_t* src; uint8_t* dst;
  int sw, sh, dw, dh, sstride, dstride;
  long long sfs, dfs;
  int vecDst;
};

__device__ __forceinline__ unsigned remap_px(const uint8_t* __restrict__ S, uint32_t t, const RemapArgs& A) {
  const int sx = (int)(t & 0xffffu) - 32, sy = (int)(t >> 16) - 32;
  const int ix = sx >> 5, iy = sy >> 5, fx = sx & 31, fy = sy & 31;
  unsigned p00, p01, p10, p11;
  if ((unsigned)ix < (unsigned)(A.sw - 1) && (unsigned)iy < (unsigned)(A.sh - 1)) {
    const uint8_t* r = S + (long long)iy * A.sstride + ix;
    p00 = __ldg(r); p01 = __ldg(r + 1); p10 = __ldg(r + A.sstride); p11 = __ldg(r + A.sstride + 1);
  } else {
    const bool x0 = (unsigned)ix < (unsigned)A.sw, x1 = (unsigned)(ix + 1) < (unsigned)A.sw;
    const bool y0 = (unsigned)iy < (unsigned)A.sh, y1 = (unsigned)(iy + 1) < (unsigned)A.sh;
    const uint8_t* r = S + (long long)iy * A.sstride + ix;
    p00 = (x0 && y0) ? __ldg(r) : 0u; p01 = (x1 && y0) ? __ldg(r + 1) : 0u;
    p10 = (x0 && y1) ? __ldg(r + A.sstride) : 0u; p11 = (x1 && y1) ? __ldg(r + A.sstride + 1) : 0u;
  }
  // weights (32-fx)(32-fy)*32 ... scaled by 2^15 in OpenCV's table: factor the common 32 out, (sum*32 + 2^14) >> 15 == (sum + 512) >> 10
  const unsigned top = p00 * (32 - fx) + p01 * fx, bot = p10 * (32 - fx) + p11 * fx;
  return (top * (32 - fy) + bot * fy + 512u) >> 10;
}

__global__ void __launch_bounds__(256) k_remap(const __grid_constant__ RemapArgs A) {
  const int x4 = (blockIdx.x * 256 + threadIdx.x) * 4, y = blockIdx.y;
  if (x4 >= A.dw) return;
  const uint8_t* S = A.src + (long long)blockIdx.z * A.sfs;
  uint8_t* d = A.dst + (long long)blockIdx.z * A.dfs + (long long)y * A.dstride + x4;
  const uint4 t = __ldg(reinterpret_cast<const uint4*>(A.table + (long long)y * A.tpitch + x4));
  if (A.vecDst && x4 + 4 <= A.dw) {
    unsigned v = remap_px(S, t.x, A) | (remap_px(S, t.y, A) << 8) | (remap_px(S, t.z, A) << 16) | (remap_px(S, t.w, A) << 24);
    *reinterpret_cast<unsigned*>(d) = v;
  } else {
    const uint32_t tt[4] = {t.x, t.y, t.z, t.w};
    const int n = min(4, A.dw - x4);
    for (int i = 0; i < n; ++i) d[i] = (uint8_t)remap_px(S, tt[i], A);
  }
}

// ---- CLAHE ---------------------------------------------------------------------------------------------------------------
struct ClaheArgs {
  const uint8_t* src; uint8_t* dst; uint8_t* lut;  // lut [frame][tilesY*tilesX][256]
  int w, h, sstride, dstride;
  long long sfs, dfs;
  int tilesX, tilesY, tw, th, clip;
  float lutScale, invTw, invTh;
  int vec, rows;
};

__device__ __forceinline__ int reflect101_dev(int p, int len) {
  if (len == 1) return 0;
  while (p < 0 || p >= len) p = p < 0 ? -p : 2 * len - 2 - p;
  return p;
}

// CTA = one tile of one frame: per-warp shared histograms, clip + redistribute, 256-wide inclusive scan, LUT.
__global__ void __launch_bounds__(256) k_clahe_lut(const __grid_constant__ ClaheArgs A) {
  __shared__ int wh[8][256];
  __shared__ int wsum[8];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int tile = blockIdx.x, tx = tile % A.tilesX, ty = tile / A.tilesX;
  const uint8_t* S = A.src + (long long)blockIdx.y * A.sfs;
  for (int i = tid; i < 8 * 256; i += 256) (&wh[0][0])[i] = 0;
  __syncthreads();
  const int total = A.tw * A.th;
  const bool inside = (tx + 1) * A.tw <= A.w && (ty + 1) * A.th <= A.h;
  for (int i = tid; i < total; i += 256) {
    int yy = ty * A.th + i / A.tw, xx = tx * A.tw + i % A.tw;
    if (!inside) { yy = reflect101_dev(yy, A.h); xx = reflect101_dev(xx, A.w); }
    atomicAdd(&wh[warp][__ldg(S + (long long)yy * A.sstride + xx)], 1);
  }
  __syncthreads();
  int hv = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) hv += wh[k][tid];
  if (A.clip > 0) {
    int ex = hv > A.clip ? hv - A.clip : 0;
    hv = min(hv, A.clip);
    for (int o = 16; o; o >>= 1) ex += __shfl_xor_sync(0xffffffffu, ex, o);
    if (lane == 0) wsum[warp] = ex;
    __syncthreads();
    int clipped = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) clipped += wsum[k];
    __syncthreads();
    const int batch = clipped >> 8;
    int residual = clipped - (batch << 8);
    hv += batch;
    if (residual != 0) {
      const int step = max(256 / residual, 1);
      if (tid % step == 0 && tid / step < residual) hv++;
    }
  }
  int s = hv;
  for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, s, o); if (lane >= o) s += t; }
  if (lane == 31) wsum[warp] = s;
  __syncthreads();
  for (int k = 0; k < warp; ++k) s += wsum[k];
  int v = __float2int_rn(__fmul_rn((float)s, A.lutScale));
  A.lut[((long long)blockIdx.y * A.tilesX * A.tilesY + tile) * 256 + tid] = (uint8_t)min(max(v, 0), 255);
}

__device__ __forceinline__ unsigned clahe_px(const uint8_t* L1, const uint8_t* L2, int x, unsigned v, float ya, float ya1, const ClaheArgs& A) {
  const float txf = __fsub_rn(__fmul_rn((float)x, A.invTw), 0.5f);
  int tx1 = __float2int_rd(txf), tx2 = tx1 + 1;
  const float xa = __fsub_rn(txf, (float)tx1), xa1 = __fsub_rn(1.0f, xa);
  tx1 = max(tx1, 0); tx2 = min(tx2, A.tilesX - 1);
  const int i1 = tx1 * 256 + v, i2 = tx2 * 256 + v;
  const float a = __fadd_rn(__fmul_rn((float)L1[i1], xa1), __fmul_rn((float)L1[i2], xa));
  const float b = __fadd_rn(__fmul_rn((float)L2[i1], xa1), __fmul_rn((float)L2[i2], xa));
  const int r = __float2int_rn(__fadd_rn(__fmul_rn(a, ya1), __fmul_rn(b, ya)));
  return (unsigned)min(max(r, 0), 255);
}

// CTA = A.rows image rows of one frame; the frame's tile LUTs are staged in shared memory once per CTA.
__global__ void __launch_bounds__(256) k_clahe_apply(const __grid_constant__ ClaheArgs A) {
  extern __shared__ __align__(16) uint8_t sl[];
  const int nlut = A.tilesX * A.tilesY * 256;
  const uint8_t* G = A.lut + (long long)blockIdx.y * nlut;
  for (int i = threadIdx.x * 16; i < nlut; i += 256 * 16) *reinterpret_cast<uint4*>(sl + i) = __ldg(reinterpret_cast<const uint4*>(G + i));
  __syncthreads();
  const int wq = (A.w + 3) >> 2, y0 = blockIdx.x * A.rows, nrows = min(A.rows, A.h - y0);
  const uint8_t* S = A.src + (long long)blockIdx.y * A.sfs;
  uint8_t* D = A.dst + (long long)blockIdx.y * A.dfs;
  for (int g = threadIdx.x; g < nrows * wq; g += 256) {
    const int y = y0 + g / wq, x4 = (g % wq) * 4;
    const float tyf = __fsub_rn(__fmul_rn((float)y, A.invTh), 0.5f);
    int ty1 = __float2int_rd(tyf), ty2 = ty1 + 1;
    const float ya = __fsub_rn(tyf, (float)ty1), ya1 = __fsub_rn(1.0f, ya);
    ty1 = max(ty1, 0); ty2 = min(ty2, A.tilesY - 1);
    const uint8_t* L1 = sl + ty1 * A.tilesX * 256;
    const uint8_t* L2 = sl + ty2 * A.tilesX * 256;
    const uint8_t* s = S + (long long)y * A.sstride + x4;
    uint8_t* d = D + (long long)y * A.dstride + x4;
    if (A.vec && x4 + 4 <= A.w) {
      const unsigned q = __ldg(reinterpret_cast<const unsigned*>(s));
      unsigned o = clahe_px(L1, L2, x4, q & 255u, ya, ya1, A) | (clahe_px(L1, L2, x4 + 1, (q >> 8) & 255u, ya, ya1, A) << 8) |
                   (clahe_px(L1, L2, x4 + 2, (q >> 16) & 255u, ya, ya1, A) << 16) | (clahe_px(L1, L2, x4 + 3, q >> 24, ya, ya1, A) << 24);
      *reinterpret_cast<unsigned*>(d) = o;
    } else {
      const int n = min(4, A.w - x4);
      for (int i = 0; i < n; ++i) d[i] = (uint8_t)clahe_px(L1, L2, x4 + i, s[i], ya, ya1, A);
    }
  }
}

}  // namespace orbx
