// k_fast_nms3 -- FAST-9/16 corner score + per-cell 3x3 NMS with in-CTA compaction (third version; SURVEY.md §9.3,
// reference call sites src/ORBextractor.cc:808,827 -> cv::FAST(cell, th, true)).
//   S(p) = max over the 16 arcs of 9 ring pixels of max(min(ring) - v, v - max(ring))
//   keypoint of its cell at threshold t  <=>  S > t  and  S strictly greater than S of the 8 neighbours inside the same cell
// What the profile of k_fast_nms said (profiles/r01b): ALU pipe 70 % busy -- every min/max form issues at 2 warp-inst/clk/SM
// (tools/microbench/minmax_rate.cu) -- 39 % of the instructions in the score loop, 40 % in a dense NMS pass, and 45 % of the
// stall samples on the un-unrolled staging loads.  Only ~10 % of the pixels of a frame pass the compass bound, but a 64-px
// warp row contains one of them 2 times out of 3, so skipping whole warp rows saved little.  This version:
//   0. stages a 70x70 window for a 62x62 output tile (halo overhead 27 % instead of 43 %), all loads of a warp issued
//      before the first use; words are stored at their natural 4-byte alignment (no funnel shift), the two u16-pair copies
//      (even / odd first column) swap roles with the parity of the window origin instead;
//   1. evaluates the 4-point compass bound for every pixel pair and appends the survivors to a per-warp shared-memory list
//      (ballot + popc, no atomics, no CTA barrier before step 2: a warp scores the pairs of its own rows);
//   2. computes the full 16-arc score only for list entries (dense warps again: VIMNMX3.U16x2, 82 per pair);
//   3. runs the NMS only for list entries whose score is non-zero; kept pixels are gathered in shared memory and written
//      with one global atomic per CTA and coalesced 8-byte records.
// (profiles/r01c, first cut of this kernel: 38 % of the instructions were step 1 -- 63 per warp row, half of them the per-row
// shared atomic + shuffle and generic-pointer address arithmetic -- and 26 % step 3's divergent per-warp emission.)
// Output: pre-candidate records (x, y, cell, S) with S > minTh in arbitrary order + per-cell "has S > iniTh" flags; the
// quadtree kernel orders them and applies the per-cell threshold fallback (src/ORBextractor.cc:825-828).
// Algorithmic bytes per launch: read every level once.
#pragma once
#include "dev_common.cuh"
#include "geom.h"
#include "orbx_internal.h"

namespace orbx {

constexpr int kF3TileW = 62, kF3TileH = 62;   // NMS outputs per CTA
constexpr int kF3SW = 64, kF3SH = 64;         // score tile (one halo pixel each side)
constexpr int kF3WinH = kF3SH + 6;            // 70 window rows
constexpr int kF3Pitch = 38;                  // u16-pair words per window row (76 columns >= 3 + 70 + 1)
constexpr int kF3RowsPerWarp = (kF3WinH + 7) / 8;
constexpr int kF3Copy = kF3WinH * kF3Pitch;   // words per pair copy
constexpr int kF3SPitch = 32, kF3SSlack = 34; // score rows: 32 pair words (33 = one pad word: fewer conflicts in step 3, measured 3 % slower); slack before / after the tile
constexpr int kF3OutCap = 1024;               // kept pixels per tile: 3x3 local maxima, at most one per 2x2 block (961)
constexpr int kF3RawW = 96;                   // bytes per raw window row fetched by the TMA path: the box must start on a 16-byte boundary of the
                                              // row (measured: any other inner coordinate raises "illegal instruction", tools/microbench/
                                              // tma_window.cu), so up to 12 bytes of lead-in + 19 words + the byte lane 18 borrows
constexpr int kF3RawBytes = kF3WinH * kF3RawW;

// One tensor map per pyramid level over (x, y, frame) of the un-blurred level, encoded on the host (orbx_extract.cu): the TMA path
// of step 0 fetches a tile's 96 x 70 byte window with ONE cp.async.bulk.tensor instruction issued by one thread while the CTA
// sets up its cell tables -- no per-lane address arithmetic, row / column clamps or shuffles, and rows / columns outside the
// level arrive as zeros (they are never read for a pixel that can be kept).
// (struct FastMaps { alignas(64) unsigned char m[kFastMapLevels][128]; } lives in orbx_internal.h: CUtensorMap = 128 opaque bytes)

__device__ __forceinline__ uint32_t f3smem(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ring pixel pair (DX, DY) of the pair whose centre word is at eIdx (even dx) / oIdx (odd dx); both are word indices into sW
template <int DX, int DY>
__device__ __forceinline__ uint32_t f3ring(const uint32_t* sW, int eIdx, int oIdx) {
  constexpr bool odd = (DX & 1) != 0;
  if (odd) return sW[oIdx + DY * kF3Pitch + (DX + 1) / 2];
  return sW[eIdx + DY * kF3Pitch + DX / 2];
}

template <bool kTma>
__global__ void __launch_bounds__(256, 5) k_fast_nms3(const __grid_constant__ Geom g, const __grid_constant__ Batch b, const __grid_constant__ FastMaps tm) {
  // pair copy A = (col 2j, 2j+1) at [0, kF3Copy), copy B = (col 2j+1, 2j+2) at [kF3Copy, 2 kF3Copy), columns relative to the
  // aligned window origin; after step 2 the same memory holds the kept pixels of the tile
  __shared__ __align__(16) uint32_t sW[2 * kF3Copy];
  // 64x64 scores as u16 pairs (+1 pad word per row), slack words each side.  The TMA path lands the raw window here first (it is dead once
  // the pair copies are built, and the scores are not written before that): the kernel keeps its 34.5 KB, so five CTAs still leave room
  // for the other extractor's kernels on the SM (a separate 6.7 KB raw buffer cost 2.4 % of the two-stream step, measured)
  __shared__ __align__(128) uint32_t sS[kF3SH * kF3SPitch + 2 * kF3SSlack];
  static_assert(sizeof(uint32_t) * (kF3SH * kF3SPitch + 2 * kF3SSlack) >= kF3RawBytes, "the raw window must fit the score array");
  __shared__ uint16_t sList[8][kF3SH * 32 / 8];               // per warp: surviving pairs of its rows, sr * 32 + p
  __shared__ uint16_t sCol[kF3SW], sRow[kF3SH];               // cell info of the tile's columns / rows
  __shared__ uint32_t sMask[3][32];                           // per pair column: left / right neighbour masks, valid halves
  __shared__ int sNOut, sBase;
  uint8_t* const sRaw = reinterpret_cast<uint8_t*>(sS);         // TMA path: the raw window (see sS)
  __shared__ __align__(8) unsigned long long sBar;

  const int tid = threadIdx.x, warp = tid >> 5, p = tid & 31;
  const int f = blockIdx.y;
  const uint32_t tinfo = __ldg(b.fastTiles + blockIdx.x);
  const int level = tinfo & 15, tx = (tinfo >> 4) & 0x3fff, ty = tinfo >> 18;
  const LevelGeom& L = g.lv[level];
  const int ox = kEdge + kF3TileW * tx, oy = kEdge + kF3TileH * ty;  // first output pixel of the tile
  const int wx0 = ox - 4, wy0 = oy - 4;                              // window origin (>= 15)
  const int a = wx0 & 3, gx0 = wx0 & ~3;
  const int maxX = L.w - kEdge, maxY = L.h - kEdge;                  // region = [kEdge, max)
  int sp;
  const uint8_t* src = level_ptr(g, b, level, f, sp);
  uint32_t* const S = sS + kF3SSlack;   // score pair (sr, q) at S[sr * kF3SPitch + q]

  if (kTma && tid == 0) {
    const uint32_t bar = f3smem(&sBar);
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(bar));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(kF3RawBytes) : "memory");
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 :: "r"(f3smem(sRaw)), "l"(reinterpret_cast<unsigned long long>(&tm.m[level & (kFastMapLevels - 1)][0])), "r"(gx0 & ~15), "r"(wy0), "r"(f), "r"(bar) : "memory");
  }
  if (tid == 0) sNOut = 0;
  // slack and pad words are read as (masked) neighbours: they must stay <= 255 per half or the packed compare borrows
  if (!kTma && tid < kF3SSlack) { sS[tid] = 0; sS[kF3SSlack + kF3SH * kF3SPitch + tid] = 0; }   // (TMA path: after the raw window is consumed)
  if (tid < kF3SW) {
    const int x = ox - 1 + tid;
    sCol[tid] = (x >= kEdge && x < maxX) ? b.info[L.colInfoOff + x] : (uint16_t)0;
  } else if (tid < kF3SW + kF3SH) {
    const int y = oy - 1 + (tid - kF3SW);
    sRow[tid - kF3SW] = (y >= kEdge && y < maxY) ? b.info[L.rowInfoOff + y] : (uint16_t)0;
  }

  // ---- 0. stage the window as two u16-pair copies
  if (kTma) {
    __syncthreads();   // the barrier object is initialised (and the cell tables above are written) before anybody polls it
    {
      const uint32_t bar = f3smem(&sBar);
      uint32_t done = 0;
      while (!done)
        asm volatile("{\n\t.reg .pred P1;\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], 0;\n\tselp.u32 %0, 1, 0, P1;\n\t}" : "=r"(done) : "r"(bar) : "memory");
    }
    const uint32_t* raw = reinterpret_cast<const uint32_t*>(sRaw) + ((gx0 >> 2) & 3) + p;   // skip the lead-in words of the 16-byte aligned box
#pragma unroll
    for (int k = 0; k < kF3RowsPerWarp; k++) {
      const int r = warp + 8 * k;
      if (r < kF3WinH && p < 19) {
        const uint32_t u = raw[r * (kF3RawW / 4)];
        const uint32_t un = raw[r * (kF3RawW / 4) + 1] & 0xffu;   // first byte of the next word
        uint2 va, vb;
        va.x = __byte_perm(u, 0u, 0x4140); va.y = __byte_perm(u, 0u, 0x4342);
        vb.x = __byte_perm(u, 0u, 0x4241); vb.y = __byte_perm(u, un, 0x5453);
        *reinterpret_cast<uint2*>(&sW[r * kF3Pitch + 2 * p]) = va;
        *reinterpret_cast<uint2*>(&sW[kF3Copy + r * kF3Pitch + 2 * p]) = vb;
      }
    }
  } else {   // levels whose base / pitch is not 16-byte aligned (odd image widths): per-lane loads, all of a warp in flight before the first use
    const int gx = min(gx0 + 4 * p, sp - 4);
    const uint8_t* colp = src + gx;
    uint32_t w[kF3RowsPerWarp];
#pragma unroll
    for (int k = 0; k < kF3RowsPerWarp; k++) {
      const int r = warp + 8 * k;
      const int gy = min(wy0 + r, L.h - 1);
      w[k] = 0;
      if (r < kF3WinH && p < 19) w[k] = __ldg(reinterpret_cast<const uint32_t*>(colp + (long long)gy * sp));
    }
#pragma unroll
    for (int k = 0; k < kF3RowsPerWarp; k++) {
      const int r = warp + 8 * k;
      const uint32_t u = w[k];
      const uint32_t un = __shfl_down_sync(ORBX_FULL, u, 1) & 0xffu;   // first byte of the next word
      if (r < kF3WinH && p < 19) {
        uint2 va, vb;
        va.x = __byte_perm(u, 0u, 0x4140); va.y = __byte_perm(u, 0u, 0x4342);
        vb.x = __byte_perm(u, 0u, 0x4241); vb.y = __byte_perm(u, un, 0x5453);
        *reinterpret_cast<uint2*>(&sW[r * kF3Pitch + 2 * p]) = va;
        *reinterpret_cast<uint2*>(&sW[kF3Copy + r * kF3Pitch + 2 * p]) = vb;
      }
    }
  }
  // the pair of lane p starts at window column 2p+3, i.e. column c0 = 2p + 3 + a of the staged row:
  //   a odd  -> c0 even: centre / even dx from copy A at p + (3+a)/2 + dx/2,   odd dx from copy B at p + (a+1)/2 + (dx+1)/2
  //   a even -> c0 odd:  centre / even dx from copy B at p + 1 + a/2 + dx/2,   odd dx from copy A at p + a/2 + 1 + (dx+1)/2
  const int eBase = (a & 1) ? (3 + a) / 2 : kF3Copy + 1 + a / 2;
  const int oBase = (a & 1) ? kF3Copy + (a + 1) / 2 : a / 2 + 1;
  __syncthreads();
  if (kTma && tid < kF3SSlack) { sS[tid] = 0; sS[kF3SSlack + kF3SH * kF3SPitch + tid] = 0; }   // read first in step 3, two barriers from here
  if (tid < 32) {
    // the left neighbour of a first-column pixel and the right neighbour of a last-column pixel belong to another cell;
    // halves outside the tile's 62 output columns or outside the region are never kept
    const uint32_t ciA = sCol[2 * tid], ciB = sCol[2 * tid + 1];
    const int xa = ox - 1 + 2 * tid;
    const bool okA = (tid >= 1) && (xa < maxX), okB = (tid <= 30) && (xa + 1 < maxX);
    sMask[0][tid] = ((ciA & 0x4000u) ? 0u : 0x0000ffffu) | ((ciB & 0x4000u) ? 0u : 0xffff0000u);
    sMask[1][tid] = ((ciA & 0x8000u) ? 0u : 0x0000ffffu) | ((ciB & 0x8000u) ? 0u : 0xffff0000u);
    sMask[2][tid] = (okA ? 0x0000ffffu : 0u) | (okB ? 0xffff0000u : 0u);
  }

  // ---- 1. compass bound for every pair: every 9-arc holds one of (r0|r8) and one of (r4|r12)  =>  S <= Q
  // The 8 rows of a warp (sr = warp + 8k; dealt round-robin: 8 consecutive rows per warp would avoid some bank conflicts of the
  // gathers in steps 2 and 3 but measured 8 % slower -- busy rows cluster, the warps of a CTA finish unevenly) sit at constant
  // offsets from the first one, so the loop is fully unrolled over two base pointers (every LDS carries an immediate offset) and the
  // "row inside the region / column needed" tests are one precomputed bit per (lane, row) -- profiles/r02 showed 40 instructions per
  // row here, 14 of them addressing and loop control.
  const uint32_t kBias = 0x01000100u;
  const uint32_t thrB = (uint32_t)g.minTh * 0x00010001u + kBias;
  uint16_t* const myList = sList[warp];
  int cnt = 0;   // warp-uniform
  {
    // rows below the region are never read as neighbours of a kept pixel; columns right of it neither
    const int rowsIn = maxY - (oy - 1) - warp;                       // row k is inside  <=>  8k < rowsIn
    const uint32_t rowMask = rowsIn <= 0 ? 0u : (rowsIn >= 57 ? 0xffu : ((1u << ((rowsIn + 7) >> 3)) - 1u));
    const uint32_t need = ((ox - 1 + 2 * p) < maxX) ? rowMask : 0u;
    const uint32_t lt = (1u << p) - 1;
    const uint32_t* const pe = sW + eBase + (warp + 3) * kF3Pitch + p;
    const uint32_t* const po = sW + oBase + (warp + 3) * kF3Pitch + p;
    uint32_t* const Srow = S + warp * kF3SPitch + p;
    uint32_t entry = (uint32_t)(warp * 32 + p);
#pragma unroll
    for (int k = 0; k < kF3SH / 8; k++) {
      const int ro = 8 * k * kF3Pitch;
      Srow[8 * k * kF3SPitch] = 0;
      if (kF3SPitch > 32 && p == 0) Srow[8 * k * kF3SPitch + 32] = 0;
      const uint32_t v = pe[ro];
      const uint32_t r0 = pe[ro + 3 * kF3Pitch], r8 = pe[ro - 3 * kF3Pitch];
      const uint32_t r4 = po[ro + 2], r12 = po[ro - 1];               // dx = +3 / -3 live in the other pair copy at (dx + 1) / 2
      const uint32_t hi = __vminu2(__vmaxu2(r0, r8), __vmaxu2(r4, r12));
      const uint32_t lo = __vmaxu2(__vminu2(r0, r8), __vminu2(r4, r12));
      const uint32_t q = __vmaxu2((hi + kBias) - v, (v + kBias) - lo);  // Q + 256 per half
      const bool pass = ((need >> k) & 1u) && (__vmaxu2(q, thrB) != thrB);
      const uint32_t m = __ballot_sync(ORBX_FULL, pass);
      if (pass) myList[cnt + __popc(m & lt)] = (uint16_t)(entry + 256 * k);
      cnt += __popc(m);
    }
  }
  __syncwarp();

  // ---- 2. full score of the surviving pairs of this warp's rows
  // Entries whose two scores are both <= minTh can neither be kept nor beat a kept pixel: the list is compacted in place
  // (about half of the compass survivors go) so that step 3 runs over dense warps too.
  const uint32_t thr2 = (uint32_t)g.minTh * 0x00010001u;
  int cnt2 = 0;   // warp-uniform
#pragma unroll 1
  for (int i0 = 0; i0 < cnt; i0 += 32) {
    const int i = i0 + p;
    const int e = myList[min(i, cnt - 1)];
    const int idx = ((e >> 5) + 3) * kF3Pitch + (e & 31);
    const int eI = eBase + idx, oI = oBase + idx;
    const uint32_t v = f3ring<0, 0>(sW, eI, oI);
    uint32_t r[16];
    r[0] = f3ring<0, 3>(sW, eI, oI);   r[1] = f3ring<1, 3>(sW, eI, oI);    r[2] = f3ring<2, 2>(sW, eI, oI);    r[3] = f3ring<3, 1>(sW, eI, oI);
    r[4] = f3ring<3, 0>(sW, eI, oI);   r[5] = f3ring<3, -1>(sW, eI, oI);   r[6] = f3ring<2, -2>(sW, eI, oI);   r[7] = f3ring<1, -3>(sW, eI, oI);
    r[8] = f3ring<0, -3>(sW, eI, oI);  r[9] = f3ring<-1, -3>(sW, eI, oI);  r[10] = f3ring<-2, -2>(sW, eI, oI); r[11] = f3ring<-3, -1>(sW, eI, oI);
    r[12] = f3ring<-3, 0>(sW, eI, oI); r[13] = f3ring<-3, 1>(sW, eI, oI);  r[14] = f3ring<-2, 2>(sW, eI, oI);  r[15] = f3ring<-1, 3>(sW, eI, oI);
    uint32_t n3[16], x3[16];
#pragma unroll
    for (int k = 0; k < 16; k++) {
      n3[k] = __vimin3_u16x2(r[k], r[(k + 1) & 15], r[(k + 2) & 15]);
      x3[k] = __vimax3_u16x2(r[k], r[(k + 1) & 15], r[(k + 2) & 15]);
    }
    uint32_t mxn = 0, mnx = 0xffffffffu;
#pragma unroll
    for (int k = 0; k < 16; k += 2) {  // two 9-arcs per step: max3 / min3 absorb both
      const uint32_t na = __vimin3_u16x2(n3[k], n3[(k + 3) & 15], n3[(k + 6) & 15]);
      const uint32_t nb = __vimin3_u16x2(n3[k + 1], n3[(k + 4) & 15], n3[(k + 7) & 15]);
      const uint32_t xa = __vimax3_u16x2(x3[k], x3[(k + 3) & 15], x3[(k + 6) & 15]);
      const uint32_t xb = __vimax3_u16x2(x3[k + 1], x3[(k + 4) & 15], x3[(k + 7) & 15]);
      mxn = __vimax3_u16x2(mxn, na, nb);
      mnx = __vimin3_u16x2(mnx, xa, xb);
    }
    const uint32_t s2 = __vmaxu2(__vmaxu2((mxn + kBias) - v, (v + kBias) - mnx), kBias) - kBias;  // max(S, 0) per half
    const bool live = (i < cnt) && (__vmaxu2(s2, thr2) != thr2);
    if (i < cnt) S[e + (e >> 5) * (kF3SPitch - 32)] = s2;
    const uint32_t m = __ballot_sync(ORBX_FULL, live);
    __syncwarp();   // every lane has read its entry of this round before the slots are reused
    if (live) myList[cnt2 + __popc(m & ((1u << p) - 1))] = (uint16_t)e;
    cnt2 += __popc(m);
  }
  cnt = cnt2;
  __syncthreads();   // all scores written; the window copies are dead from here on

  // ---- 3. NMS inside the owning cell for list entries with a non-zero score; kept pixels go to the tile's output list
  uint2* const sOut = reinterpret_cast<uint2*>(sW);   // x | y << 16,  cell | S << 16   (= PreCand)
#pragma unroll 1
  for (int i0 = 0; i0 < cnt; i0 += 32) {
    const int i = i0 + p;
    bool keepA = false, keepB = false;
    uint32_t cc = 0, ciA = 0, ciB = 0, ri = 0;
    int xa = 0, y = 0;
    if (i < cnt) {
      const int e = myList[i];
      const int sr = e >> 5, q = e & 31;
      const uint32_t* Se = S + e + sr * (kF3SPitch - 32);
      cc = Se[0];
      y = oy - 1 + sr; xa = ox - 1 + 2 * q;
      if (cc != 0 && sr >= 1 && sr <= kF3TileH && y < maxY) {
        ciA = sCol[2 * q]; ciB = sCol[2 * q + 1]; ri = sRow[sr];
        const uint32_t maskL = sMask[0][q], maskR = sMask[1][q], valid = sMask[2][q];
        const uint32_t pc = Se[-1], nc = Se[1];
        uint32_t nb = __vmaxu2(__funnelshift_r(pc, cc, 16) & maskL, __funnelshift_r(cc, nc, 16) & maskR);
        if (!(ri & 0x4000u)) {
          const uint32_t pu = Se[-kF3SPitch - 1], cu = Se[-kF3SPitch], nu = Se[-kF3SPitch + 1];
          nb = __vimax3_u16x2(nb, cu, __vmaxu2(__funnelshift_r(pu, cu, 16) & maskL, __funnelshift_r(cu, nu, 16) & maskR));
        }
        if (!(ri & 0x8000u)) {
          const uint32_t pd = Se[kF3SPitch - 1], cd = Se[kF3SPitch], nd = Se[kF3SPitch + 1];
          nb = __vimax3_u16x2(nb, cd, __vmaxu2(__funnelshift_r(pd, cd, 16) & maskL, __funnelshift_r(cd, nd, 16) & maskR));
        }
        nb = __vmaxu2(nb, thr2);
        const uint32_t tt = ((cc & valid) + kBias) - nb;      // per half: > 0x100  <=>  S > max(neighbours, minTh)
        keepA = (tt & 0xffffu) > 0x100u; keepB = (tt >> 16) > 0x100u;
      }
    }
    const uint32_t mA = __ballot_sync(ORBX_FULL, keepA), mB = __ballot_sync(ORBX_FULL, keepB);
    if ((mA | mB) == 0) continue;
    const int nA = __popc(mA), total = nA + __popc(mB);
    int base = 0;
    if (p == 0) base = atomicAdd(&sNOut, total);
    base = __shfl_sync(ORBX_FULL, base, 0);
    const uint32_t rowCell = (ri & 0x3fffu) * (uint32_t)L.nCols, lt = (1u << p) - 1;
    if (keepA) {
      const int o = base + __popc(mA & lt);
      if (o < kF3OutCap) sOut[o] = make_uint2((uint32_t)xa | ((uint32_t)y << 16), (rowCell + (ciA & 0x3fffu)) | ((cc & 0xffu) << 16));
    }
    if (keepB) {
      const int o = base + nA + __popc(mB & lt);
      if (o < kF3OutCap) sOut[o] = make_uint2((uint32_t)(xa + 1) | ((uint32_t)y << 16), (rowCell + (ciB & 0x3fffu)) | (((cc >> 16) & 0xffu) << 16));
    }
  }
  __syncthreads();

  // ---- emission: one global reservation per tile, coalesced records, per-cell "has S > iniTh" flags
  const int nOut = min(sNOut, kF3OutCap);
  if (nOut == 0) return;
  if (tid == 0) sBase = atomicAdd(b.candCount + f * ORBX_MAX_LEVELS + level, nOut);
  __syncthreads();
  const int gbase = sBase;
  uint2* cand = reinterpret_cast<uint2*>(b.cand + (long long)f * g.candSlab + L.candOff);
  int* cellFlags = b.cellFlags + (long long)f * g.nCellsTotal + L.cellBase;
  for (int i = tid; i < nOut; i += 256) {
    const uint2 rec = sOut[i];
    if (gbase + i < L.candCap) cand[gbase + i] = rec;
    else atomicExch(b.status, ORBX_ERR_CAPACITY);
    if ((int)(rec.y >> 16) > g.iniTh) atomicOr(cellFlags + (rec.y & 0xffffu), 1);
  }
}

}  // namespace orbx
