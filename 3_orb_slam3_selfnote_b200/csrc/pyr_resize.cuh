// k_pyr_resize2 -- one pyramid level from the previous one, cv::resize(INTER_LINEAR) on 8U in OpenCV's fixed point
// (reference call site src/ORBextractor.cc:1201).  Tuned replacement of k_pyr_resize (extract_kernels.cuh),
// which issued 16 byte-wide LDGs per thread and was LSU/latency bound (profiles/r01: issue 30 %, DRAM 5 %).
//   CTA = 128x16 destination pixels of one frame.
//   1. the source window (<= ~156x21 px at scale 1.2) is staged with aligned 32-bit loads, coalesced along rows;
//   2. horizontal pass ONCE per source row (OpenCV's order: rows are interpolated horizontally first):
//      h = S[sx]*a0 + S[sx+1]*a1, kept as (h >> 4) in shared memory; a thread owns one destination column, so its
//      table entry (source columns + Q11 weights) lives in registers for all window rows;
//   3. vertical pass: thread = 4 destination pixels of one row, two LDS.128 (rows sy, sy+1), one 32-bit store:
//      v = ((b0*(h0>>4))>>16) + ((b1*(h1>>4))>>16) + 2 >> 2.
// Algorithmic bytes per launch: read level l-1 + write level l.
#pragma once
#include "dev_common.cuh"
#include "geom.h"
#include "orbx_internal.h"

namespace orbx {

constexpr int kRsTileW = 128, kRsTileH = 16;

__global__ void __launch_bounds__(256) k_pyr_resize2(const __grid_constant__ Geom g, const __grid_constant__ Batch b, int level) {
  extern __shared__ __align__(16) uint32_t rs_smem[];
  const LevelGeom& L = g.lv[level];
  const int winWords = g.rsWinWords, winRows = g.rsWinRows;
  uint32_t* sRaw = rs_smem;                                               // [winRows][winWords]  source bytes
  int* sH = reinterpret_cast<int*>(rs_smem + winRows * winWords);          // [winRows][128]       (h >> 4)
  const int tid = threadIdx.x, f = blockIdx.z;
  const int x0 = blockIdx.x * kRsTileW, y0 = blockIdx.y * kRsTileH;
  int sp;
  const uint8_t* src = level_ptr(g, b, level - 1, f, sp);
  const RsX* tx = b.rsx + L.rsxOff;
  const RsY* ty = b.rsy + L.rsyOff;
  const int wpad = (L.w + 3) & ~3;                                         // the column table is padded to a multiple of 4
  const int xl = min(x0 + kRsTileW, wpad) - 1, yl = min(y0 + kRsTileH, L.h) - 1;
  const int gx0 = (int)tx[x0].s0 & ~3;                                     // first staged source column (word aligned)
  const int nW = (((int)tx[xl].s1 - gx0) >> 2) + 1;                        // words per window row   (<= winWords)
  const int wy0 = ty[y0].s0;
  const int nR = (int)ty[yl].s1 - wy0 + 1;                                 // window rows            (<= winRows)

  // ---- 1. stage the window
  for (int r = tid >> 5; r < nR; r += 8) {
    const uint32_t* row = reinterpret_cast<const uint32_t*>(src + (long long)(wy0 + r) * sp + gx0);
    for (int k = tid & 31; k < nW; k += 32) sRaw[r * winWords + k] = __ldg(row + k);
  }
  __syncthreads();

  // ---- 2. horizontal pass: thread = destination column c, row group rg
  {
    const int c = tid & (kRsTileW - 1), rg = tid >> 7;
    if (x0 + c < wpad) {
      const RsX e = tx[x0 + c];
      const int o0 = e.s0 - gx0, o1 = e.s1 - gx0, a0 = e.a0, a1 = e.a1;
      const uint8_t* rawb = reinterpret_cast<const uint8_t*>(sRaw);
#pragma unroll 4
      for (int r = rg; r < nR; r += 2) {
        const uint8_t* rb = rawb + r * (winWords * 4);
        const int h = (int)rb[o0] * a0 + (int)rb[o1] * a1;
        sH[r * kRsTileW + c] = h >> 4;
      }
    }
  }
  __syncthreads();

  // ---- 3. vertical pass: task = 4 destination pixels
  uint8_t* dstBase = b.pyr + (long long)f * b.pyrFrameStride + L.pyrOff;
#pragma unroll
  for (int it = 0; it < (kRsTileW / 4) * kRsTileH / 256; it++) {
    const int task = tid + it * 256;
    const int q = task & 31, r = task >> 5;
    const int x = x0 + 4 * q, y = y0 + r;
    if (x >= L.w || y >= L.h) continue;
    const RsY e = ty[y];
    const int b0 = e.b0, b1 = e.b1;
    const int4 h0 = *reinterpret_cast<const int4*>(sH + (e.s0 - wy0) * kRsTileW + 4 * q);
    const int4 h1 = *reinterpret_cast<const int4*>(sH + (e.s1 - wy0) * kRsTileW + 4 * q);
    const int v0 = min(255, max(0, (((b0 * h0.x) >> 16) + ((b1 * h1.x) >> 16) + 2) >> 2));
    const int v1 = min(255, max(0, (((b0 * h0.y) >> 16) + ((b1 * h1.y) >> 16) + 2) >> 2));
    const int v2 = min(255, max(0, (((b0 * h0.z) >> 16) + ((b1 * h1.z) >> 16) + 2) >> 2));
    const int v3 = min(255, max(0, (((b0 * h0.w) >> 16) + ((b1 * h1.w) >> 16) + 2) >> 2));
    *reinterpret_cast<uint32_t*>(dstBase + (long long)y * L.pitch + x) = (uint32_t)v0 | ((uint32_t)v1 << 8) | ((uint32_t)v2 << 16) | ((uint32_t)v3 << 24);
  }
}



// ---------------------------------------------------------------------------------------------------------------------------
// k_pyr_resize4 -- register-resident version (k_pyr_resize2 executed ~70 thread-instructions per destination pixel, issue-bound
// at 82 %: partially filled staging loops, one shared-memory round trip for the horizontal pass, scalar vertical arithmetic).
//   thread = 4 destination columns (one 32-bit store) walking down a strip of kRs4Rows destination rows;
//   per source row: three aligned 32-bit loads cover the <= 8 source bytes the 4 columns touch (scale <= 2), two funnel shifts
//   line them up at the first source column, and per column one PRMT picks its two taps and one IDP.2A computes
//   S[sx]*a0 + S[sx+1]*a1 (the Q11 weights as a u16 pair); source rows are interpolated horizontally once and carried in
//   registers from one destination row to the next (the row tables are warp-uniform, so the reuse test does not diverge);
//   vertical: ((b0*(h0>>4))>>16) + ((b1*(h1>>4))>>16) as two IMAD.HI with the weights pre-shifted by 16, then (+2)>>2.
//   No clamp: h>>4 <= 32655 and b0+b1 <= 2049 bound the result by 255.
// Levels whose 4-column groups span more than 8 source bytes (scale factor > 2) use k_pyr_resize2.
// Its predecessor k_pyr_resize3 (round 1, 0.354 ms per 256 images of 752x480) spent 120 warp instructions per destination row
// where the arithmetic needs ~55 (profiles/r01f source page: 27 % IMAD -- register moves of the carried rows and 64-bit address
// arithmetic -- branches around the row-reuse tests, every load consumed right after issue).  This version (0.251 ms):
//   * the two carried horizontal rows alternate roles statically (row j keeps h[s0] where row j-1 kept h[s1]): the common case
//     "s0 advanced by one" computes ONE new source row and moves nothing; "advanced by two" computes both;
//   * the raw words of the next destination row's lower source row are loaded one row ahead;
//   * the row table of the strip is read with one 8-byte load per warp and broadcast by shuffles;
//   * levels >= 2 read a source whose rows have slack behind them (pyramid slab), so the two trailing words need no clamp.
// ---------------------------------------------------------------------------------------------------------------------------
constexpr int kRs4TileW = 128;
constexpr int kRs4Rows = 8;   // destination rows per warp (measured: 4 -> 0.284 ms, 8 -> 0.252, 12 -> 0.260, 16 -> 0.260)

// one strip: 128 destination columns (x = first column of this lane's 4) x kRs4Rows destination rows from ys on.
// kClamp: the source rows have no slack behind them (level 1 reads the caller's image).
// kSmem: the CTA's source window was fetched into shared memory by one TMA copy (k_pyr_resize4t): win = its first byte, (winX, winY) = the
// source coordinates of that byte, winPitch = bytes per window row; xIdle = a column of this CTA that lanes right of the level may read.
template <bool kClamp, bool kSmem = false>
__device__ __forceinline__ void rs4_strip(const Geom& g, const Batch& b, int level, int f, int x, int ys, int lane, const uint8_t* win = nullptr,
                                          int winX = 0, int winY = 0, int winPitch = 0, int xIdle = 0) {
  const LevelGeom& L = g.lv[level];
  const int wpad = (L.w + 3) & ~3;   // the column table is padded to a multiple of 4
  const bool active = x < wpad;      // lanes right of the level stay for the shuffles and skip loads / stores
  int sp;
  const uint8_t* src = level_ptr(g, b, level - 1, f, sp);
  if (kSmem) { src = win - ((long long)winY * winPitch + winX); sp = winPitch; }   // source (sx, sy) at src + sy * sp + sx, inside the window
  const int rows = min(kRs4Rows, L.h - ys);
  // row table of the strip: lane j holds entry ys + j (kSmem: lanes past the end repeat the last row, whose source rows are in the window)
  uint2 myRow = make_uint2(0u, 0u);
  if (lane < rows || kSmem) myRow = __ldg(reinterpret_cast<const uint2*>(b.rsy + L.rsyOff + ys + min(lane, rows - 1)));
  uint32_t sel[4], wt[4], sh;
  int off1 = 4, off2 = 8;
  const uint8_t* srcx = src;
  {
    const uint4* tx = reinterpret_cast<const uint4*>(b.rsx + L.rsxOff + (active ? x : xIdle));   // 4 entries of 8 bytes (lanes right of the level: column 0 / a column of the window)
    const uint4 p0 = __ldg(tx), p1 = __ldg(tx + 1);
    const uint32_t e[8] = {p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w};   // (s0 | s1 << 16), (a0 | a1 << 16) per column
    const int sx0 = (int)(e[0] & 0xffffu);
    const int base = sx0 & ~3;
    sh = (uint32_t)(sx0 & 3) * 8;
    srcx = src + base;
    if (kClamp && !kSmem) { off1 = min(base + 4, sp - 4) - base; off2 = min(base + 8, sp - 4) - base; }   // words past the row end hold no tap
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int s0 = (int)(e[2 * k] & 0xffffu), s1 = (int)(e[2 * k] >> 16);
      sel[k] = (uint32_t)min(s0 - sx0, 7) | ((uint32_t)min(s1 - sx0, 7) << 4);   // upper two bytes: don't care (byte 0)
      wt[k] = e[2 * k + 1];
    }
  }
  const unsigned usp = (unsigned)sp, upitch = (unsigned)L.pitch;
  auto load3 = [&](int srow, uint32_t (&w)[3]) {
    if (kSmem) {
      const uint8_t* p = srcx + srow * (int)usp;
      w[0] = *reinterpret_cast<const uint32_t*>(p);
      w[1] = *reinterpret_cast<const uint32_t*>(p + 4);
      w[2] = *reinterpret_cast<const uint32_t*>(p + 8);
    } else {
      const uint8_t* p = srcx + (unsigned long long)(unsigned)srow * usp;   // one IMAD.WIDE.U32
      w[0] = __ldg(reinterpret_cast<const uint32_t*>(p));
      w[1] = __ldg(reinterpret_cast<const uint32_t*>(p + off1));
      w[2] = __ldg(reinterpret_cast<const uint32_t*>(p + off2));
    }
  };
  auto prmt = [](uint32_t a, uint32_t b2, uint32_t sl) { uint32_t r; asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b2), "r"(sl)); return r; };
  auto hcalc = [&](const uint32_t (&w)[3], uint32_t (&h)[4]) {
    const uint32_t u0 = __funnelshift_r(w[0], w[1], sh), u1 = __funnelshift_r(w[1], w[2], sh);   // 8 bytes from the first source column on
#pragma unroll
    for (int k = 0; k < 4; k++) h[k] = __dp2a_lo(wt[k], prmt(u0, u1, sel[k]), 0u) >> 4;
  };
  uint8_t* dst = b.pyr + (long long)f * b.pyrFrameStride + L.pyrOff + x;
  asm volatile("" : "+l"(dst));      // keep both per-thread base pointers whole in registers: otherwise the uniform part is re-added
  if (!kSmem) asm volatile("" : "+l"(srcx));     // (IADD3 + IADD3.X) at every use
  uint32_t hh[2][4] = {{0, 0, 0, 0}, {0, 0, 0, 0}};
  int held0 = -1;                    // source row whose h is in hh[j & 1] when row j starts (= s1 of row j-1)
  // raw words of row j's s1 are loaded one row ahead (two rows ahead: 48 registers, 0.256 ms against 0.242), unconditionally: lanes right of the level read from column 0 and rows past
  // the strip's end read source row 0 (table entries of lanes >= rows are zero) -- valid addresses, results unused
  uint32_t wbuf[2][3];
  uint32_t e0 = __shfl_sync(ORBX_FULL, myRow.x, 0);
  load3((int)(e0 >> 16), wbuf[0]);
#pragma unroll
  for (int j = 0; j < kRs4Rows; j++) {
    if (j >= rows) break;
    const uint32_t eb = __shfl_sync(ORBX_FULL, myRow.y, j);
    const int s0 = (int)(e0 & 0xffffu), s1 = (int)(e0 >> 16);
    uint32_t eNext = 0;
    const uint32_t (&wcur)[3] = wbuf[j & 1];
    if (j + 1 < kRs4Rows) {
      eNext = __shfl_sync(ORBX_FULL, myRow.x, j + 1);
      load3((int)(eNext >> 16), wbuf[(j + 1) & 1]);
    }
    if (active) {
      uint32_t (&X)[4] = hh[j & 1];          // h[s0]: where the previous row kept h[s1]
      uint32_t (&Y)[4] = hh[(j & 1) ^ 1];
      hcalc(wcur, Y);
      if (held0 != s0) { uint32_t w0[3]; load3(s0, w0); hcalc(w0, X); }   // first row of the strip, or s0 advanced by two
      held0 = s1;
      const uint32_t B0 = eb << 16, B1 = eb & 0xffff0000u;
      uint32_t t[4];
#pragma unroll
      for (int k = 0; k < 4; k++) t[k] = __umulhi(B0, X[k]) + __umulhi(B1, Y[k]) + 2u;   // <= 1023
      // (t >> 2) of two values per shift: t0 | t1 << 16 shifted right by 2 leaves t0 >> 2 in byte 0 and t1 >> 2 in byte 2
      const uint32_t p01 = prmt(t[0], t[1], 0x5410u) >> 2, p23 = prmt(t[2], t[3], 0x5410u) >> 2;
      *reinterpret_cast<uint32_t*>(dst + (unsigned long long)(unsigned)(ys + j) * upitch) = prmt(p01, p23, 0x6420u);
    }
    e0 = eNext;
  }
}
template <bool kClamp>
__global__ void __launch_bounds__(256) k_pyr_resize4(const __grid_constant__ Geom g, const __grid_constant__ Batch b, int level) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int ys = (blockIdx.y * 8 + warp) * kRs4Rows;
  if (ys >= g.lv[level].h) return;   // warp-uniform
  rs4_strip<kClamp>(g, b, level, blockIdx.z, blockIdx.x * kRs4TileW + 4 * lane, ys, lane);
}
// k_pyr_resize4t -- the same strips fed from shared memory: the CTA's source window (128 x 64 destination pixels -> boxW x boxH source
// bytes, e.g. 192 x 80 at scale 1.2; the sizes are maxima over all tiles of all levels, computed on the host from the tables) arrives by ONE
// cp.async.bulk.tensor.3d issued by thread 0 (tensor map of level - 1 over x, y, frame; box start = the tile's first source column rounded
// down to 16 bytes, its first source row; zeros outside the level) and is waited for on an mbarrier.  The three word loads per source row
// become LDS with ~20x less latency than the L2 / HBM loads the register-resident kernel hides with one-row-ahead prefetching.
__global__ void __launch_bounds__(256) k_pyr_resize4t(const __grid_constant__ Geom g, const __grid_constant__ Batch b, const __grid_constant__ FastMaps tm,
                                                      int level, int boxW, int boxH) {
  extern __shared__ __align__(128) uint8_t rs_win[];
  __shared__ __align__(8) unsigned long long sBar;
  __shared__ int sOrg[2];
  const LevelGeom& L = g.lv[level];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int x0 = blockIdx.x * kRs4TileW, y0 = blockIdx.y * 8 * kRs4Rows, f = blockIdx.z;
  if (threadIdx.x == 0) {
    const int gx = (int)b.rsx[L.rsxOff + x0].s0 & ~15, wy = (int)b.rsy[L.rsyOff + y0].s0;
    sOrg[0] = gx; sOrg[1] = wy;
    const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&sBar);
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(bar));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(boxW * boxH) : "memory");
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 :: "r"((uint32_t)__cvta_generic_to_shared(rs_win)), "l"(reinterpret_cast<unsigned long long>(&tm.m[(level - 1) & (kFastMapLevels - 1)][0])),
                    "r"(gx), "r"(wy), "r"(f), "r"(bar) : "memory");
  }
  __syncthreads();
  {
    const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&sBar);
    uint32_t done = 0;
    while (!done)
      asm volatile("{\n\t.reg .pred P1;\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], 0;\n\tselp.u32 %0, 1, 0, P1;\n\t}" : "=r"(done) : "r"(bar) : "memory");
  }
  const int ys = y0 + warp * kRs4Rows;
  if (ys >= L.h) return;   // warp-uniform
  rs4_strip<false, true>(g, b, level, f, x0 + 4 * lane, ys, lane, rs_win, sOrg[0], sOrg[1], boxW, x0);
}
// Measured and dropped: one clustered launch for the small levels 4..7 of a frame (4 CTAs per frame, cluster barrier between the
// levels, sources re-read through L2): 0.262 - 0.302 ms against 0.251 for one launch per level -- the separate launches overlap
// their tails on the stream, the barrier-separated levels do not.
}  // namespace orbx
