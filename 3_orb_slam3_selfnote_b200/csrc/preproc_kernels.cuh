// Kernels of the pre-processing step in front of ORBextractor::operator(): colour -> gray,
// stereo rectification (remap) and CLAHE, each bit-exact with the OpenCV 8-bit path the reference calls
// (src/Tracking.cc:989-1014, Examples/Stereo/stereo_euroc.cc:166-167, Examples/Monocular/mono_tum_vi.cc:101-109).
// All three are streaming byte kernels (HBM-bound): thread = 4 adjacent output pixels, 32-bit stores, frames batched in
// blockIdx.z / .y.  No floating-point contraction anywhere (--fmad=false, explicit _rn intrinsics in the CLAHE blend).
#pragma once
#include <cuda_runtime.h>
#include <limits.h>
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

// Thread = 4 adjacent pixels of GRAY_ROWS consecutive rows; every row's loads are issued before the first multiply so one
// thread keeps GRAY_ROWS x 12 (16) bytes in flight (one row per thread measured 56-67 % of the HBM peak: latency-bound).
#define ORBX_GRAY_ROWS 4
template <int CH>
__global__ void __launch_bounds__(256) k_cvt_gray(const __grid_constant__ GrayArgs A) {
  const int x4 = (blockIdx.x * 256 + threadIdx.x) * 4, y0 = blockIdx.y * ORBX_GRAY_ROWS;
  if (x4 >= A.w) return;
  const uint8_t* s = A.src + (long long)blockIdx.z * A.sfs + (long long)y0 * A.sstride + (long long)x4 * CH;
  uint8_t* d = A.dst + (long long)blockIdx.z * A.dfs + (long long)y0 * A.dstride + x4;
  const int nr = min(ORBX_GRAY_ROWS, A.h - y0);
  if (A.vec && x4 + 4 <= A.w) {
    unsigned q[ORBX_GRAY_ROWS][4];
#pragma unroll
    for (int r = 0; r < ORBX_GRAY_ROWS; ++r) {
      if (r < nr) {
        if (CH == 3) {
          const unsigned* p = reinterpret_cast<const unsigned*>(s + (long long)r * A.sstride);
          q[r][0] = __ldg(p); q[r][1] = __ldg(p + 1); q[r][2] = __ldg(p + 2); q[r][3] = 0;
        } else {
          const uint4 v = __ldg(reinterpret_cast<const uint4*>(s + (long long)r * A.sstride));
          q[r][0] = v.x; q[r][1] = v.y; q[r][2] = v.z; q[r][3] = v.w;
        }
      }
    }
#pragma unroll
    for (int r = 0; r < ORBX_GRAY_ROWS; ++r) {
      if (r < nr) {
        unsigned g0, g1, g2, g3;
        if (CH == 3) {
          const unsigned w0 = q[r][0], w1 = q[r][1], w2 = q[r][2];
          g0 = gray_px(w0 & 255u, (w0 >> 8) & 255u, (w0 >> 16) & 255u, A);
          g1 = gray_px(w0 >> 24, w1 & 255u, (w1 >> 8) & 255u, A);
          g2 = gray_px((w1 >> 16) & 255u, w1 >> 24, w2 & 255u, A);
          g3 = gray_px((w2 >> 8) & 255u, (w2 >> 16) & 255u, w2 >> 24, A);
        } else {
          g0 = gray_px(q[r][0] & 255u, (q[r][0] >> 8) & 255u, (q[r][0] >> 16) & 255u, A);
          g1 = gray_px(q[r][1] & 255u, (q[r][1] >> 8) & 255u, (q[r][1] >> 16) & 255u, A);
          g2 = gray_px(q[r][2] & 255u, (q[r][2] >> 8) & 255u, (q[r][2] >> 16) & 255u, A);
          g3 = gray_px(q[r][3] & 255u, (q[r][3] >> 8) & 255u, (q[r][3] >> 16) & 255u, A);
        }
        *reinterpret_cast<unsigned*>(d + (long long)r * A.dstride) = g0 | (g1 << 8) | (g2 << 16) | (g3 << 24);
      }
    }
  } else {
    const int n = min(4, A.w - x4);
    for (int r = 0; r < nr; ++r)
      for (int i = 0; i < n; ++i)
        d[(long long)r * A.dstride + i] = (uint8_t)gray_px(s[(long long)r * A.sstride + i * CH], s[(long long)r * A.sstride + i * CH + 1],
                                                          s[(long long)r * A.sstride + i * CH + 2], A);
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
  int vecDst, vecSrc;
};

// Thread = 4 adjacent destination pixels of one row, one table word (uint4) and 16 one-byte gathers.  ncu (profiles/
// r01f_preproc_ncu_summary.csv): bound by the L1 -- l1tex 74 %, 8 sectors per request, LSU-queue (lg_throttle) stalls -- not by
// HBM (DRAM 7 %).  Variants measured on the B200 per 256 frames of 752x480 and dropped: 4 rows per thread with branch-free
// clamped addresses 0.32 ms, 2 rows + warp-uniform interior fast path 0.27 ms, 1 row + fast path 0.25 ms, this one 0.19 ms:
// deeper per-thread load queues only lengthen the LSU queue.  The step that would change the bound is a shared-memory
// staged source tile per CTA -- built in round 1 (k_remap_tile, see git history / profiles/README.md) and measured: slower still
// (0.31 ms: table load -> box reduction -> staging -> taps are three dependent round trips separated by CTA barriers).
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
  int vec;
};

__device__ __forceinline__ int reflect101_dev(int p, int len) {
  if (len == 1) return 0;
  while (p < 0 || p >= len) p = p < 0 ? -p : 2 * len - 2 - p;
  return p;
}

// Warp = one tile of one frame (8 tiles per CTA): a private 256-bin histogram per warp, no CTA barrier anywhere.  Rows are
// taken four at a time so that four word loads per lane are in flight; then lane l owns bins 8l .. 8l+7 for the clip, the
// redistribution, the running sum and the 8 LUT bytes it stores as one 64-bit word.
__global__ void __launch_bounds__(256) k_clahe_lut_w(const __grid_constant__ ClaheArgs A, int ntiles /* tiles per frame */, int total /* x frames */) {
  __shared__ int wh[8][256];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int job = blockIdx.x * 8 + warp;
  if (job >= total) return;
  const int frame = job / ntiles, tile = job - frame * ntiles, tx = tile % A.tilesX, ty = tile / A.tilesX;
  const uint8_t* S = A.src + (long long)frame * A.sfs;
  int* H = wh[warp];
#pragma unroll
  for (int j = 0; j < 8; ++j) H[lane + 32 * j] = 0;
  __syncwarp();
  const bool inside = (tx + 1) * A.tw <= A.w && (ty + 1) * A.th <= A.h;
  if (inside && A.vec) {
    const int x0 = tx * A.tw, a0 = x0 & ~3, nw = (x0 + A.tw - a0 + 3) >> 2;
    const uint8_t* base = S + (long long)(ty * A.th) * A.sstride + a0;
    for (int w0 = 0; w0 < nw; w0 += 32) {
      const int wi = w0 + lane, xb = a0 + wi * 4;
      const bool on = wi < nw;
      const bool whole = on && xb + 4 <= A.w;
      unsigned lo = 0, hi = 0;   // byte range of this word inside [x0, x0 + tw)
      if (on) { lo = (unsigned)max(x0 - xb, 0); hi = (unsigned)min(x0 + A.tw - xb, 4); }
      for (int r = 0; r < A.th; r += 4) {
        unsigned q[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          q[k] = 0;
          if (on && r + k < A.th) {
            const uint8_t* pw = base + (long long)(r + k) * A.sstride + wi * 4;
            if (whole) q[k] = __ldg(reinterpret_cast<const unsigned*>(pw));
            else for (unsigned b = lo; b < hi; ++b) q[k] |= (unsigned)__ldg(pw + b) << (8 * b);
          }
        }
        // (electing one lane per bin with __match_any_sync before the atomic was measured at 0.93 ms per 256 frames against
        // 0.26 ms for the plain shared-memory atomics below: match.any costs far more than the serialisation it removes)
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (on && r + k < A.th)
            for (unsigned b = lo; b < hi; ++b) atomicAdd(&H[(q[k] >> (8 * b)) & 255u], 1);
      }
    }
  } else {
    const int npx = A.tw * A.th;
    for (int i = lane; i < npx; i += 32) {
      int yy = ty * A.th + i / A.tw, xx = tx * A.tw + i % A.tw;
      if (!inside) { yy = reflect101_dev(yy, A.h); xx = reflect101_dev(xx, A.w); }
      atomicAdd(&H[__ldg(S + (long long)yy * A.sstride + xx)], 1);
    }
  }
  __syncwarp();
  int hv[8], ex = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    hv[j] = H[lane * 8 + j];
    if (A.clip > 0 && hv[j] > A.clip) { ex += hv[j] - A.clip; hv[j] = A.clip; }
  }
  if (A.clip > 0) {
    for (int o = 16; o; o >>= 1) ex += __shfl_xor_sync(0xffffffffu, ex, o);
    const int batch = ex >> 8, residual = ex - (batch << 8);
    const int step = residual ? max(256 / residual, 1) : 1;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int i = lane * 8 + j;
      hv[j] += batch;
      if (residual != 0 && i % step == 0 && i / step < residual) hv[j]++;
    }
  }
  int local = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) { local += hv[j]; hv[j] = local; }
  int incl = local;
  for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
  const int before = incl - local;
  unsigned w0 = 0, w1 = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int v = min(max(__float2int_rn(__fmul_rn((float)(before + hv[j]), A.lutScale)), 0), 255);
    if (j < 4) w0 |= (unsigned)v << (8 * j); else w1 |= (unsigned)v << (8 * (j - 4));
  }
  *reinterpret_cast<uint2*>(A.lut + ((long long)frame * ntiles + tile) * 256 + lane * 8) = make_uint2(w0, w1);
}

struct ClaheCol { float xa, xa1; int i1, i2; };  // per-column constants of the blend (clahe.cpp's ind1_p / ind2_p / xa_p / xa1_p)

__device__ __forceinline__ ClaheCol clahe_col(int x, const ClaheArgs& A) {
  const float txf = __fsub_rn(__fmul_rn((float)x, A.invTw), 0.5f);
  int tx1 = __float2int_rd(txf), tx2 = tx1 + 1;
  ClaheCol c;
  c.xa = __fsub_rn(txf, (float)tx1); c.xa1 = __fsub_rn(1.0f, c.xa);
  c.i1 = max(tx1, 0) * 256; c.i2 = min(tx2, A.tilesX - 1) * 256;
  return c;
}

// u8 -> float and cvRound(float) -> int through the 2^23 magic numbers (exact for these ranges): the conversion
// instructions (I2F / F2I) run at a quarter of the FP32 rate and there are five of them per pixel otherwise.
__device__ __forceinline__ float u8_to_float(unsigned v) { return __fsub_rn(__uint_as_float(0x4B000000u | v), 8388608.0f); }

__device__ __forceinline__ unsigned clahe_px(const uint8_t* L1, const uint8_t* L2, const ClaheCol& c, unsigned v, float ya, float ya1) {
  const float a = __fadd_rn(__fmul_rn(u8_to_float(L1[c.i1 + v]), c.xa1), __fmul_rn(u8_to_float(L1[c.i2 + v]), c.xa));
  const float b = __fadd_rn(__fmul_rn(u8_to_float(L2[c.i1 + v]), c.xa1), __fmul_rn(u8_to_float(L2[c.i2 + v]), c.xa));
  const float res = __fadd_rn(__fmul_rn(a, ya1), __fmul_rn(b, ya));  // in [0, 255.001]: a blend of LUT bytes
  const int r = __float_as_int(__fadd_rn(res, 12582912.0f)) - 0x4B400000;  // round-half-even, = cvRound(res)
  return (unsigned)min(max(r, 0), 255);
}

#define ORBX_CLAHE_ROWS 16
// CTA = ORBX_CLAHE_ROWS image rows of one frame; the frame's tile LUTs are staged in shared memory once per CTA.  Thread = a
// group of 4 adjacent columns walking down the rows: the column constants stay in registers, the row constants come from
// a small shared table, the next row's word is loaded while the current one is blended.
__global__ void __launch_bounds__(256) k_clahe_apply(const __grid_constant__ ClaheArgs A) {
  extern __shared__ __align__(16) uint8_t sl[];
  __shared__ float sYa[ORBX_CLAHE_ROWS], sYa1[ORBX_CLAHE_ROWS];
  __shared__ int sT1[ORBX_CLAHE_ROWS], sT2[ORBX_CLAHE_ROWS];
  const int nlut = A.tilesX * A.tilesY * 256;
  const uint8_t* G = A.lut + (long long)blockIdx.y * nlut;
  for (int i = threadIdx.x * 16; i < nlut; i += blockDim.x * 16) *reinterpret_cast<uint4*>(sl + i) = __ldg(reinterpret_cast<const uint4*>(G + i));
  const int wq = (A.w + 3) >> 2, y0 = blockIdx.x * ORBX_CLAHE_ROWS, nrows = min(ORBX_CLAHE_ROWS, A.h - y0);
  if (threadIdx.x < nrows) {
    const float tyf = __fsub_rn(__fmul_rn((float)(y0 + threadIdx.x), A.invTh), 0.5f);
    const int ty1 = __float2int_rd(tyf), ty2 = ty1 + 1;
    const float ya = __fsub_rn(tyf, (float)ty1);
    sYa[threadIdx.x] = ya; sYa1[threadIdx.x] = __fsub_rn(1.0f, ya);
    sT1[threadIdx.x] = max(ty1, 0) * A.tilesX * 256; sT2[threadIdx.x] = min(ty2, A.tilesY - 1) * A.tilesX * 256;
  }
  __syncthreads();
  const uint8_t* S = A.src + (long long)blockIdx.y * A.sfs + (long long)y0 * A.sstride;
  uint8_t* D = A.dst + (long long)blockIdx.y * A.dfs + (long long)y0 * A.dstride;
  for (int c = threadIdx.x; c < wq; c += blockDim.x) {
    const int x4 = c * 4;
    const ClaheCol c0 = clahe_col(x4, A), c1 = clahe_col(x4 + 1, A), c2 = clahe_col(x4 + 2, A), c3 = clahe_col(x4 + 3, A);
    if (A.vec && x4 + 4 <= A.w) {
      unsigned q = __ldg(reinterpret_cast<const unsigned*>(S + x4));
      for (int r = 0; r < nrows; ++r) {
        const unsigned cur = q;
        if (r + 1 < nrows) q = __ldg(reinterpret_cast<const unsigned*>(S + (long long)(r + 1) * A.sstride + x4));
        const uint8_t* L1 = sl + sT1[r];
        const uint8_t* L2 = sl + sT2[r];
        const float ya = sYa[r], ya1 = sYa1[r];
        *reinterpret_cast<unsigned*>(D + (long long)r * A.dstride + x4) =
            clahe_px(L1, L2, c0, cur & 255u, ya, ya1) | (clahe_px(L1, L2, c1, (cur >> 8) & 255u, ya, ya1) << 8) |
            (clahe_px(L1, L2, c2, (cur >> 16) & 255u, ya, ya1) << 16) | (clahe_px(L1, L2, c3, cur >> 24, ya, ya1) << 24);
      }
    } else {
      const ClaheCol cc[4] = {c0, c1, c2, c3};
      const int n = min(4, A.w - x4);
      for (int r = 0; r < nrows; ++r)
        for (int i = 0; i < n; ++i)
          D[(long long)r * A.dstride + x4 + i] = (uint8_t)clahe_px(sl + sT1[r], sl + sT2[r], cc[i], S[(long long)r * A.sstride + x4 + i], sYa[r], sYa1[r]);
    }
  }
}

}  // namespace orbx
