// k_orient_desc2 -- IC_Angle + steered BRIEF, tuned replacement of k_orient_desc (extract_kernels.cuh).
// Reference: computeOrientation / IC_Angle (src/ORBextractor.cc:75-102, 470-477), computeOrbDescriptor (:105-145, 1059-1066).
// profiles/r01b: k_orient_desc executed 1 350 warp-instructions per keypoint, ~45 % of them in the rolled moment loop (umax
// lookup, predicate, 64-bit address arithmetic and a dependent byte load per row) and ~25 % unpacking the int8 pattern.
//   * the disc rows are unrolled with the compile-time umax table of a radius-15 patch (the constructor's table is checked
//     against it on the host), m10 = u * sum(row values), m01 = sum(v * value); the 31x31 raw patch is staged in shared
//     memory too, so the 31 row reads are byte LDS at immediate offsets instead of loads with 64-bit address arithmetic;
//   * the pattern is read as pre-converted float4 (x0, y0, x1, y1) records laid out [j][lane] (coalesced 512-B reads);
//   * the 512 descriptor samples were 1-byte gathers from L1 with up to 32 distinct sectors per instruction (the tag stage
//     serialises them: ~400 L1 cycles per keypoint, the kernel's bound).  The 37x37 blurred patch around the keypoint (pattern
//     radius 18.38 -> |offset| <= 18) is now staged per warp in shared memory with coalesced aligned 32-bit loads and the
//     gathers are byte LDS;
//   * arithmetic order, rounding and the double-precision sin/cos are unchanged (bit-exact angles and descriptors).
#pragma once
#include "dev_common.cuh"
#include "extract_kernels.cuh"
#include "geom.h"
#include "orbx_internal.h"

namespace orbx {

__host__ __device__ __forceinline__ constexpr int od_umax(int v) {
  // umax of HALF_PATCH_SIZE = 15 (src/ORBextractor.cc:452-467): 15,15,15,15,14,14,14,13,13,12,11,10,9,8,6,3
  return v < 4 ? 15 : v < 7 ? 14 : v < 9 ? 13 : v == 9 ? 12 : v == 10 ? 11 : v == 11 ? 10 : v == 12 ? 9 : v == 13 ? 8 : v == 14 ? 6 : 3;
}

constexpr int kOdR = 18;                    // largest |row| / |col| a rotated pattern point can round to
constexpr int kOdRows = 2 * kOdR + 1;        // 37
constexpr int kOdWords = 10;                 // 40-byte blurred-patch rows: 37 columns + up to 3 bytes of alignment slack
constexpr int kOmWords = 9;                  // 36-byte raw-patch rows:     31 columns + up to 3 bytes of alignment slack

// Stage ROWS rows of WORDS aligned 32-bit words into shared memory, three rows per warp instruction (lane = r3 * WORDS + word,
// so the destination index of step k is simply 3 * WORDS * k + lane); all loads are issued before the first store.
template <int ROWS, int WORDS>
__device__ __forceinline__ void od_stage(const uint8_t* rowPtr0, int pitch, uint32_t* dst, int lane) {
  constexpr int kSteps = (ROWS + 2) / 3;
  const int r3 = lane / WORDS, c = lane - r3 * WORDS;
  const uint32_t* ptr = reinterpret_cast<const uint32_t*>(rowPtr0 + (long long)r3 * pitch) + c;
  const long long step = 3ll * pitch;
  uint32_t wv[kSteps];
#pragma unroll
  for (int k = 0; k < kSteps; k++) {
    wv[k] = 0;
    if (lane < 3 * WORDS && 3 * k + r3 < ROWS) wv[k] = __ldg(ptr);
    ptr = reinterpret_cast<const uint32_t*>(reinterpret_cast<const uint8_t*>(ptr) + step);
  }
#pragma unroll
  for (int k = 0; k < kSteps; k++)
    if (lane < 3 * WORDS && 3 * k + r3 < ROWS) dst[3 * WORDS * k + lane] = wv[k];
}

// TMA staging (kTma): lane 0 of the warp fetches both patches with one cp.async.bulk.tensor.3d each -- boxes of 48 x 31 (raw) and 64 x 37
// (blurred) bytes starting on the 16-byte boundary at or before the patch's first column (the inner coordinate of a tensor copy must be a
// multiple of 16 bytes) -- onto a per-warp mbarrier; the 24 + 24 load / store instructions of od_stage and their L1 wavefronts disappear
// (the kernel is L1 / shared-memory bound).  Maps: OrientMaps, encoded on the host like FastMaps.
constexpr int kOmBoxW = 48, kOdBoxW = 64;
constexpr int kOmBoxBytes = (kOmBoxW * kPatch + 127) & ~127, kOdBoxBytes = (kOdBoxW * kOdRows + 127) & ~127;

template <bool kTma>
__global__ void __launch_bounds__(256) k_orient_desc2(const __grid_constant__ Geom g, const __grid_constant__ Batch b, const __grid_constant__ OrientMaps tm) {
  constexpr int kRawPitch = kTma ? kOmBoxW : kOmWords * 4, kBlurPitch = kTma ? kOdBoxW : kOdWords * 4;   // bytes per staged row
  __shared__ __align__(128) uint8_t sPatchB[8][kTma ? kOdBoxBytes : kOdRows * kOdWords * 4];   // blurred 37x37 patch (rBRIEF samples)
  __shared__ __align__(128) uint8_t sRawB[8][kTma ? kOmBoxBytes : kPatch * kOmWords * 4];       // un-blurred 31x31 patch (moments)
  __shared__ __align__(8) unsigned long long sBar[8];
  const int f = blockIdx.y, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int slot = blockIdx.x * 8 + warp;
  if (slot >= g.kpSlab) return;
  int level = 0;
  while (level + 1 < g.nlevels && slot >= g.lv[level + 1].kpOff) level++;
  const LevelGeom& L = g.lv[level];
  const int i = slot - L.kpOff;
  if (i >= b.lkpCount[f * ORBX_MAX_LEVELS + level]) return;
  const LevelKp kp = b.lkp[(long long)f * g.kpSlab + slot];
  const int dest = b.lkpDest[(long long)f * g.kpSlab + slot];
  const int px0 = kp.x - kOdR, mx0 = kp.x - kHalfPatch;
  int pa, ma;   // offset of the patch's first column inside the staged row
  if (kTma) {
    pa = px0 & 15; ma = mx0 & 15;
    if (lane == 0) {
      const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&sBar[warp]);
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(bar));
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(kOmBoxW * kPatch + kOdBoxW * kOdRows) : "memory");
      asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                   :: "r"((uint32_t)__cvta_generic_to_shared(sRawB[warp])), "l"(reinterpret_cast<unsigned long long>(&tm.raw[level & (kFastMapLevels - 1)][0])),
                      "r"(mx0 - ma), "r"(kp.y - kHalfPatch), "r"(f), "r"(bar) : "memory");
      asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                   :: "r"((uint32_t)__cvta_generic_to_shared(sPatchB[warp])), "l"(reinterpret_cast<unsigned long long>(&tm.blur[level & (kFastMapLevels - 1)][0])),
                      "r"(px0 - pa), "r"(kp.y - kOdR), "r"(f), "r"(bar) : "memory");
    }
    __syncwarp();   // the barrier is initialised before the other lanes poll it
    {
      const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&sBar[warp]);
      uint32_t done = 0;
      while (!done)
        asm volatile("{\n\t.reg .pred P1;\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], 0;\n\tselp.u32 %0, 1, 0, P1;\n\t}" : "=r"(done) : "r"(bar) : "memory");
    }
  } else {
    int sp;
    const uint8_t* img = level_ptr(g, b, level, f, sp);
    // both patches start at the aligned word holding their first column (level pitches and the input pitch are multiples of 4)
    const int bp = L.blurPitch;
    pa = px0 & 3; ma = mx0 & 3;
    od_stage<kPatch, kOmWords>(img + (long long)(kp.y - kHalfPatch) * sp + (mx0 - ma), sp, reinterpret_cast<uint32_t*>(sRawB[warp]), lane);
    od_stage<kOdRows, kOdWords>(b.blur + (long long)f * b.blurFrameStride + L.blurOff + (long long)(kp.y - kOdR) * bp + (px0 - pa), bp,
                                reinterpret_cast<uint32_t*>(sPatchB[warp]), lane);
  }
  __syncwarp();
  // IC_Angle: m10 = sum u*I, m01 = sum v*I over the radius-15 disc; lane = column u + 15 (lane 31 idles)
  int m10 = 0, m01 = 0;
  {
    const int u = lane - kHalfPatch;
    const int au = abs(u);
    const uint8_t* c = sRawB[warp] + ma + min(lane, kPatch - 1);
    int rowSum = 0;
#pragma unroll
    for (int r = 0; r < kPatch; r++) {
      const int v = r - kHalfPatch;
      const int raw = c[r * kRawPitch];                                   // always in the staged patch: select, don't branch
      const int val = (au <= od_umax(v < 0 ? -v : v)) ? raw : 0;
      rowSum += val;
      m01 += v * val;
    }
    m10 = u * rowSum;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      m10 += __shfl_xor_sync(ORBX_FULL, m10, o);
      m01 += __shfl_xor_sync(ORBX_FULL, m01, o);
    }
  }
  const float angle = fast_atan2_deg((float)m01, (float)m10);
  // rBRIEF
  const float factorPI = (float)(3.1415926535897932384626433832795 / 180.0);
  const float ang = __fmul_rn(angle, factorPI);
  // (float)cos((double)ang), (float)sin((double)ang) with one shared argument reduction: sincos() gives the same two floats for EVERY float in
  // [0, 6.3] (all 1 086 953 883 of them compared on the B200, tools/microbench/sincos_check.cu), and ang = angle * pi/180 lies in [0, 2 pi)
  double sd, cd;
  sincos((double)ang, &sd, &cd);
  const float ca = (float)cd, sa = (float)sd;
  const uint8_t* bl = sPatchB[warp] + kOdR * kBlurPitch + kOdR + pa;   // patch centre
  const float4* pat = b.patternF + lane;   // transposed: record j of lane l (pair 8l + j) at [j * 32 + l]
  int bits = 0;
#pragma unroll
  for (int j = 0; j < 8; j++) {
    const float4 q = __ldg(pat + 32 * j);
    const int rr0 = __float2int_rn(__fadd_rn(__fmul_rn(q.x, sa), __fmul_rn(q.y, ca)));
    const int cc0 = __float2int_rn(__fsub_rn(__fmul_rn(q.x, ca), __fmul_rn(q.y, sa)));
    const int rr1 = __float2int_rn(__fadd_rn(__fmul_rn(q.z, sa), __fmul_rn(q.w, ca)));
    const int cc1 = __float2int_rn(__fsub_rn(__fmul_rn(q.z, ca), __fmul_rn(q.w, sa)));
    const int t0 = bl[rr0 * kBlurPitch + cc0], t1 = bl[rr1 * kBlurPitch + cc1];
    bits |= (t0 < t1) << j;
  }
  b.desc[((long long)f * g.kpStride + dest) * 32 + lane] = (uint8_t)bits;
  if (lane == 0) {
    orbx_keypoint o;
    o.x = (level == 0) ? (float)kp.x : __fmul_rn((float)kp.x, L.scale);
    o.y = (level == 0) ? (float)kp.y : __fmul_rn((float)kp.y, L.scale);
    o.size = L.kpSize;
    o.angle = angle;
    o.response = (float)kp.resp;
    o.octave = level;
    o.class_id = -1;
    b.kps[(long long)f * g.kpStride + dest] = o;
  }
}

}  // namespace orbx
