// C-ABI of the pre-processing step (include/orbx.h, "pre-processing" section): colour -> gray, remap, CLAHE.
// Host orchestration only; the arithmetic is in preproc_kernels.cuh.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <new>
#include <vector>

#include "../../include/orbx.h"
#include "orbx_internal.h"
#include "preproc_kernels.cuh"

using namespace orbx;

struct orbx_preproc {
  int device = 0;
  cudaStream_t stream = nullptr;
  uint8_t* dIn = nullptr; size_t inBytes = 0;
  uint8_t* dOut = nullptr; size_t outBytes = 0;
  uint8_t* dLut = nullptr; size_t lutBytes = 0;
  bool claheAttr = false;
};

struct orbx_remap_table {
  int device = 0;
  uint32_t* dTable = nullptr;
  int tpitch = 0, dw = 0, dh = 0, sw = 0, sh = 0;
};

namespace {
int grow(uint8_t** p, size_t* have, size_t need) {
  if (*have >= need) return ORBX_OK;
  std::lock_guard<std::recursive_mutex> lock(setup_mutex());
  if (*p) CK(cudaFree(*p));
  *p = nullptr; *have = 0;
  CK(cudaMalloc((void**)p, need));
  *have = need;
  return ORBX_OK;
}

// Resolves the (src, dst) pair of a call: device pointers are used in place; host images are staged densely through the
// handle's buffers (upload here, download in finish()).
struct Staging {
  orbx_preproc* p; int mem; cudaStream_t st;
  const uint8_t* hSrc; uint8_t* hDst;
  const uint8_t* dSrc; uint8_t* dDst;
  int n, srcRowBytes, srcRows, sstride, dstRowBytes, dstRows, dstride;
  long long sfs, dfs;

  int begin(orbx_preproc* p_, int mem_, void* stream, const uint8_t* src, int n_, int srcRowBytes_, int srcRows_, int& sstride_, long long& sfs_,
            uint8_t* dst, int dstRowBytes_, int dstRows_, int& dstride_, long long& dfs_) {
    p = p_; mem = mem_; n = n_; srcRowBytes = srcRowBytes_; srcRows = srcRows_; dstRowBytes = dstRowBytes_; dstRows = dstRows_;
    CK(cudaSetDevice(p->device));
    if (mem == ORBX_MEM_DEVICE) {
      st = stream ? (cudaStream_t)stream : p->stream;
      dSrc = src; dDst = dst;
      return ORBX_OK;
    }
    st = p->stream;
    hSrc = src; hDst = dst; sstride = sstride_; sfs = sfs_; dstride = dstride_; dfs = dfs_;
    const int sp = (srcRowBytes + 15) & ~15, dp = (dstRowBytes + 15) & ~15;
    int rc = grow(&p->dIn, &p->inBytes, (size_t)sp * srcRows * n);
    if (rc) return rc;
    rc = grow(&p->dOut, &p->outBytes, (size_t)dp * dstRows * n);
    if (rc) return rc;
    for (int f = 0; f < n; ++f)
      CK(cudaMemcpy2DAsync(p->dIn + (size_t)f * sp * srcRows, sp, hSrc + (long long)f * sfs, sstride, srcRowBytes, srcRows, cudaMemcpyHostToDevice, st));
    dSrc = p->dIn; dDst = p->dOut;
    sstride_ = sp; sfs_ = (long long)sp * srcRows; dstride_ = dp; dfs_ = (long long)dp * dstRows;
    return ORBX_OK;
  }
  int finish() {
    CK(cudaGetLastError());
    if (mem == ORBX_MEM_DEVICE) return ORBX_OK;
    const int dp = (dstRowBytes + 15) & ~15;
    for (int f = 0; f < n; ++f)
      CK(cudaMemcpy2DAsync(hDst + (long long)f * dfs, dstride, p->dOut + (size_t)f * dp * dstRows, dp, dstRowBytes, dstRows, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    return ORBX_OK;
  }
};

bool aligned(const void* p, long long a, long long b, int k) { return ((uintptr_t)p % k) == 0 && (a % k) == 0 && (b % k) == 0; }

int check_image(const void* src, const void* dst, int n, int w, int h, int mem) {
  if (!src || !dst) { set_error("null image pointer"); return ORBX_ERR_ARG; }
  if (n < 0 || n > 65535) { set_error("nframes out of range"); return ORBX_ERR_ARG; }
  if (w < 1 || h < 1) return ORBX_EMPTY_IMAGE;
  if (w > 8192 || h > 8192) { set_error("image larger than 8192 px is unsupported"); return ORBX_ERR_UNSUPPORTED; }
  if (mem != ORBX_MEM_HOST && mem != ORBX_MEM_DEVICE) { set_error("mem must be ORBX_MEM_HOST or ORBX_MEM_DEVICE"); return ORBX_ERR_ARG; }
  return ORBX_OK;
}
}  // namespace

namespace orbx {
// used by orbx_frame.cu (the Frame-side helpers share the handle's stream and staging buffers)
int preproc_device(const orbx_preproc* p) { return p->device; }
cudaStream_t preproc_stream(const orbx_preproc* p) { return p->stream; }
int preproc_scratch(orbx_preproc* p, size_t inBytes, size_t outBytes, uint8_t** dIn, uint8_t** dOut) {
  int rc = grow(&p->dIn, &p->inBytes, inBytes);
  if (rc) return rc;
  rc = grow(&p->dOut, &p->outBytes, outBytes);
  if (rc) return rc;
  *dIn = p->dIn; *dOut = p->dOut;
  return ORBX_OK;
}
}  // namespace orbx

extern "C" {

int orbx_preproc_create(int device, orbx_preproc** out) {
  if (!out) return ORBX_ERR_ARG;
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
    set_error("no usable CUDA device %d (there is no CPU fallback)", device);
    return ORBX_ERR_CUDA;
  }
  std::lock_guard<std::recursive_mutex> lock(setup_mutex());
  orbx_preproc* p = new (std::nothrow) orbx_preproc();
  if (!p) return ORBX_ERR_ARG;
  p->device = device;
  CK(cudaSetDevice(device));
  CK(cudaStreamCreateWithFlags(&p->stream, cudaStreamNonBlocking));
  *out = p;
  return ORBX_OK;
}

void orbx_preproc_destroy(orbx_preproc* p) {
  if (!p) return;
  std::lock_guard<std::recursive_mutex> lock(setup_mutex());
  cudaSetDevice(p->device);
  if (p->stream) { cudaStreamSynchronize(p->stream); cudaStreamDestroy(p->stream); }
  cudaFree(p->dIn); cudaFree(p->dOut); cudaFree(p->dLut);
  delete p;
}

int orbx_cvt_gray(orbx_preproc* p, const uint8_t* src, int nframes, int width, int height, int src_stride, long long src_frame_stride,
                  int channels, int rgb_order, int coeffs, uint8_t* dst, int dst_stride, long long dst_frame_stride, int mem, void* stream) {
  if (!p) return ORBX_ERR_ARG;
  int rc = check_image(src, dst, nframes, width, height, mem);
  if (rc) return rc;
  if (channels != 3 && channels != 4) { set_error("channels must be 3 or 4 (1-channel input needs no conversion, Tracking.cc:989)"); return ORBX_ERR_ARG; }
  if (coeffs != ORBX_GRAY_OCV346 && coeffs != ORBX_GRAY_OCV4) { set_error("unknown gray coefficient set"); return ORBX_ERR_ARG; }
  if (src_stride < width * channels || dst_stride < width) { set_error("stride shorter than a row"); return ORBX_ERR_ARG; }
  if (nframes == 0) return ORBX_OK;
  Staging S;
  rc = S.begin(p, mem, stream, src, nframes, width * channels, height, src_stride, src_frame_stride, dst, width, height, dst_stride, dst_frame_stride);
  if (rc) return rc;
  GrayArgs A{};
  A.src = S.dSrc; A.dst = S.dDst; A.w = width; A.h = height; A.sstride = src_stride; A.dstride = dst_stride;
  A.sfs = src_frame_stride; A.dfs = dst_frame_stride;
  const unsigned cr = coeffs == ORBX_GRAY_OCV4 ? 9798u : 4899u, cg = coeffs == ORBX_GRAY_OCV4 ? 19235u : 9617u, cb = coeffs == ORBX_GRAY_OCV4 ? 3735u : 1868u;
  A.shift = coeffs == ORBX_GRAY_OCV4 ? 15 : 14;
  A.half = 1u << (A.shift - 1);
  A.c0 = rgb_order ? cr : cb; A.c1 = cg; A.c2 = rgb_order ? cb : cr;
  A.vec = aligned(S.dSrc, src_stride, src_frame_stride, channels == 3 ? 4 : 16) && aligned(S.dDst, dst_stride, dst_frame_stride, 4);
  dim3 grid((width + 1023) / 1024, (height + ORBX_GRAY_ROWS - 1) / ORBX_GRAY_ROWS, nframes);
  if (channels == 3) k_cvt_gray<3><<<grid, 256, 0, S.st>>>(A);
  else k_cvt_gray<4><<<grid, 256, 0, S.st>>>(A);
  return S.finish();
}

int orbx_remap_create(orbx_preproc* p, const float* mapx, const float* mapy, int dst_width, int dst_height, int map_stride,
                      int src_width, int src_height, orbx_remap_table** out) {
  if (!p || !out || !mapx || !mapy) return ORBX_ERR_ARG;
  *out = nullptr;
  if (dst_width < 1 || dst_height < 1 || src_width < 1 || src_height < 1) return ORBX_EMPTY_IMAGE;
  if (map_stride < dst_width) { set_error("map stride shorter than a row"); return ORBX_ERR_ARG; }
  if (src_width > 2046 || src_height > 2046 || dst_width > 8192 || dst_height > 8192) {
    set_error("remap: source larger than 2046 px does not fit the packed 16.5 fixed-point table");
    return ORBX_ERR_UNSUPPORTED;
  }
  orbx_remap_table* t = new (std::nothrow) orbx_remap_table();
  if (!t) return ORBX_ERR_ARG;
  t->device = p->device; t->dw = dst_width; t->dh = dst_height; t->sw = src_width; t->sh = src_height;
  t->tpitch = (dst_width + 3) & ~3;
  std::vector<uint32_t> tab((size_t)t->tpitch * dst_height, 0u);
  for (int y = 0; y < dst_height; ++y)
    for (int x = 0; x < dst_width; ++x) {
      // cvRound(map * INTER_TAB_SIZE): round-half-even under the default rounding mode, INT_MIN for NaN / overflow like cvtss2si
      float fx = mapx[(size_t)y * map_stride + x] * 32.f, fy = mapy[(size_t)y * map_stride + x] * 32.f;
      long sx = (fx >= -2147483648.f && fx < 2147483648.f) ? lrintf(fx) : (long)INT32_MIN;
      long sy = (fy >= -2147483648.f && fy < 2147483648.f) ? lrintf(fy) : (long)INT32_MIN;
      sx = sx < -32 ? -32 : sx > (long)src_width * 32 ? (long)src_width * 32 : sx;
      sy = sy < -32 ? -32 : sy > (long)src_height * 32 ? (long)src_height * 32 : sy;
      tab[(size_t)y * t->tpitch + x] = ((uint32_t)(sy + 32) << 16) | (uint32_t)(sx + 32);
    }
  std::lock_guard<std::recursive_mutex> lock(setup_mutex());
  if (cudaSetDevice(p->device) != cudaSuccess || cudaMalloc((void**)&t->dTable, tab.size() * 4) != cudaSuccess ||
      cudaMemcpy(t->dTable, tab.data(), tab.size() * 4, cudaMemcpyHostToDevice) != cudaSuccess) {
    set_error("remap table upload failed: %s", cudaGetErrorString(cudaGetLastError()));
    cudaFree(t->dTable);
    delete t;
    return ORBX_ERR_CUDA;
  }
  *out = t;
  return ORBX_OK;
}

void orbx_remap_destroy(orbx_remap_table* t) {
  if (!t) return;
  std::lock_guard<std::recursive_mutex> lock(setup_mutex());
  cudaSetDevice(t->device);
  cudaFree(t->dTable);
  delete t;
}

int orbx_remap(orbx_preproc* p, const orbx_remap_table* t, const uint8_t* src, int nframes, int src_stride, long long src_frame_stride,
               uint8_t* dst, int dst_stride, long long dst_frame_stride, int mem, void* stream) {
  if (!p || !t) return ORBX_ERR_ARG;
  if (t->device != p->device) { set_error("remap table lives on another device"); return ORBX_ERR_ARG; }
  int rc = check_image(src, dst, nframes, t->dw, t->dh, mem);
  if (rc) return rc;
  if (src_stride < t->sw || dst_stride < t->dw) { set_error("stride shorter than a row"); return ORBX_ERR_ARG; }
  if ((long long)src_stride * t->sh > 0x7fffffffLL) { set_error("remap: source frame larger than 2 GB"); return ORBX_ERR_UNSUPPORTED; }
  if (nframes == 0) return ORBX_OK;
  Staging S;
  rc = S.begin(p, mem, stream, src, nframes, t->sw, t->sh, src_stride, src_frame_stride, dst, t->dw, t->dh, dst_stride, dst_frame_stride);
  if (rc) return rc;
  RemapArgs A{};
  A.table = t->dTable; A.tpitch = t->tpitch; A.src = S.dSrc; A.dst = S.dDst;
  A.sw = t->sw; A.sh = t->sh; A.dw = t->dw; A.dh = t->dh; A.sstride = src_stride; A.dstride = dst_stride;
  A.sfs = src_frame_stride; A.dfs = dst_frame_stride;
  A.vecDst = aligned(S.dDst, dst_stride, dst_frame_stride, 4);
  A.vecSrc = aligned(S.dSrc, src_stride, src_frame_stride, 4);
  k_remap<<<dim3((t->dw + 1023) / 1024, t->dh, nframes), 256, 0, S.st>>>(A);
  return S.finish();
}

int orbx_clahe(orbx_preproc* p, const uint8_t* src, int nframes, int width, int height, int src_stride, long long src_frame_stride,
               double clip_limit, int tiles_x, int tiles_y, uint8_t* dst, int dst_stride, long long dst_frame_stride, int mem, void* stream) {
  if (!p) return ORBX_ERR_ARG;
  int rc = check_image(src, dst, nframes, width, height, mem);
  if (rc) return rc;
  if (tiles_x < 1 || tiles_y < 1 || tiles_x * tiles_y > 512) { set_error("CLAHE tile grid must hold 1..512 tiles"); return ORBX_ERR_ARG; }
  if (src_stride < width || dst_stride < width) { set_error("stride shorter than a row"); return ORBX_ERR_ARG; }
  if (nframes == 0) return ORBX_OK;
  int ew = width, eh = height;  // clahe.cpp: pad BOTH directions as soon as one is not a multiple of the grid
  if (width % tiles_x != 0 || height % tiles_y != 0) { ew = width + tiles_x - width % tiles_x; eh = height + tiles_y - height % tiles_y; }
  Staging S;
  rc = S.begin(p, mem, stream, src, nframes, width, height, src_stride, src_frame_stride, dst, width, height, dst_stride, dst_frame_stride);
  if (rc) return rc;
  const size_t nlut = (size_t)tiles_x * tiles_y * 256;
  rc = grow(&p->dLut, &p->lutBytes, nlut * nframes);
  if (rc) return rc;
  ClaheArgs A{};
  A.src = S.dSrc; A.dst = S.dDst; A.lut = p->dLut; A.w = width; A.h = height; A.sstride = src_stride; A.dstride = dst_stride;
  A.sfs = src_frame_stride; A.dfs = dst_frame_stride;
  A.tilesX = tiles_x; A.tilesY = tiles_y; A.tw = ew / tiles_x; A.th = eh / tiles_y;
  const int total = A.tw * A.th;
  A.lutScale = (float)255 / total;
  A.clip = 0;
  if (clip_limit > 0.0) { A.clip = (int)(clip_limit * total / 256); if (A.clip < 1) A.clip = 1; }
  A.invTw = 1.0f / A.tw; A.invTh = 1.0f / A.th;
  A.vec = aligned(S.dSrc, src_stride, src_frame_stride, 4) && aligned(S.dDst, dst_stride, dst_frame_stride, 4);
  if (!p->claheAttr) {
    std::lock_guard<std::recursive_mutex> lock(setup_mutex());
    CK(cudaFuncSetAttribute(k_clahe_apply, cudaFuncAttributeMaxDynamicSharedMemorySize, 512 * 256));
    p->claheAttr = true;
  }
  const int ntiles = tiles_x * tiles_y;
  k_clahe_lut_w<<<(ntiles * nframes + 7) / 8, 256, 0, S.st>>>(A, ntiles, ntiles * nframes);
  const int wq = (width + 3) / 4, threads = wq >= 256 ? 256 : ((wq + 31) & ~31);
  k_clahe_apply<<<dim3((height + ORBX_CLAHE_ROWS - 1) / ORBX_CLAHE_ROWS, nframes), threads, nlut, S.st>>>(A);
  return S.finish();
}

}  // extern "C"
