// k_knn2_tc -- the 2-NN fold of the all-pairs keyframe sweep (BASELINE.json configs[4]) on the 5th-generation tensor cores.
// k_knn2 (match_kernels.cuh) runs at 0.97 of the POPC roof (16 POPC/clk/SM, 8 per 256-bit pair), so only another unit helps:
//   hamming(a, b) = (256 + dot(-a', b')) / 2   with a', b' the bits expanded to +1 / -1 int8,
// i.e. the distance matrix of 128 queries x 256 train descriptors is ONE 128 x 256 x 256 int8 GEMM tile: 8 tcgen05.mma
// kind::i8 (M 128, N 256, K 32 each) from shared-memory operands into a TMEM accumulator, 8.4 M MACs in ~1 100 clk against
// 32 768 x 8 POPC = 16 400 clk on the POPC pipe.
//   k_expand_i8   train bits -> int8 tile images in HBM, laid out exactly as the shared-memory operand ([tile][16-byte K chunk]
//                 [256 rows][16 B]: K-major, no swizzle, core matrix = 8 rows x 16 B), once per train chunk; every query CTA
//                 then pulls a tile with ONE 64 KB cp.async.bulk (no producer arithmetic, no tensor map).
//   k_knn2_tc     CTA = 128 queries (their negated int8 image stays in shared memory for the whole sweep), 6 warps:
//                 warp 4 = bulk-copy producer (2 x 64 KB ring), warp 5 = MMA issuer (one thread), warps 0-3 = epilogue: thread =
//                 TMEM lane = query row; two accumulator stages (2 x 256 TMEM columns) so tile n+1 multiplies while tile n is
//                 reduced.  The epilogue keeps a running (best, second) per row: per 32 columns a min3 tree (0.5 instructions per
//                 pair) against a threshold derived from the current second best; only a chunk that can change the top-2 takes
//                 the exact per-element path (key = distance << 23 | train index, ties to the lower index, own-keyframe range and
//                 the ragged last tile excluded there).  After ~2 ln(nt) early updates per row that path is almost never taken.
// Results are identical to k_knn2's (same keys, same tie rule): tests/test_match_gpu.py::test_knn2_tc_matches_popc_kernel.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace orbx {

// CTA shape: kTcH accumulator halves of 128 queries each against train tiles of kTcN = 256 / kTcH descriptors.  L2 -> shared-memory
// traffic is 256 B per train row per CTA pass, i.e. 2 / kTcH bytes per pair: with one half (128 queries x 256 train) the sweep ran
// at 1.97e12 pairs/s = 3.9 TB/s of L2 reads, the L2's limit, so more queries share a tile.
#ifndef ORBX_TC_H
#define ORBX_TC_H 2
#endif
constexpr int kTcH = ORBX_TC_H;
constexpr int kTcM = 128 * kTcH, kTcN = 256 / kTcH;
constexpr int kTcTileBytes = kTcN * 256;          // one train tile image
#ifndef ORBX_TC_STAGES
#define ORBX_TC_STAGES ((227 * 1024 - 128 * ORBX_TC_H * 256) / ((256 / ORBX_TC_H) * 256) > 6 ? 6 : (227 * 1024 - 128 * ORBX_TC_H * 256) / ((256 / ORBX_TC_H) * 256))
#endif
constexpr int kTcStages = ORBX_TC_STAGES;   // train-tile ring in shared memory (a bulk copy takes longer than one tile's MMAs: several must be in flight)
constexpr int kTcEpiWarps = 4 * kTcH;
constexpr int kTcThreads = 32 * (kTcEpiWarps + 2);

__host__ __device__ inline size_t knn2_tc_workspace_bytes(long long nt) { return (size_t)((nt + kTcN - 1) / kTcN) * kTcTileBytes; }

namespace tc {
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
// 4 bits -> 4 bytes of +1 (0x01) / -1 (0xff); neg swaps the two
__device__ __forceinline__ uint32_t spread4(uint32_t x, bool neg) {
  uint32_t w = ((x & 0xfu) * 0x00204081u) & 0x01010101u;
  if (neg) w ^= 0x01010101u;
  return w | ((w ^ 0x01010101u) * 0xffu);
}
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lboBytes, uint32_t sboBytes) {
  return (uint64_t)((saddr >> 4) & 0x3fffu) | ((uint64_t)((lboBytes >> 4) & 0x3fffu) << 16) | ((uint64_t)((sboBytes >> 4) & 0x3fffu) << 32) |
         (1ull << 46);   // descriptor version 1 (sm_100), layout type 0 = no swizzle
}
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(b)), "r"(count)); }
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
  asm volatile("{\n\t.reg .pred P1;\n\tLAB_WAIT:\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t@P1 bra DONE;\n\tbra LAB_WAIT;\n\tDONE:\n\t}"
               :: "r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(b)) : "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* b) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* b) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void umma_i8(uint32_t tmemD, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t}"
               :: "r"(tmemD), "l"(da), "l"(db), "r"(idesc), "r"(accumulate), "r"(0), "r"(0), "r"(0), "r"(0) : "memory");
}
#define ORBX_TC_V32(v) v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7], v[8], v[9], v[10], v[11], v[12], v[13], v[14], v[15], v[16], v[17], v[18], v[19], \
                       v[20], v[21], v[22], v[23], v[24], v[25], v[26], v[27], v[28], v[29], v[30], v[31]
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, int (&v)[32]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, "
               "%22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]),
                 "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]),
                 "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]),
                 "=r"(v[31])
               : "r"(taddr));
}
// wait for the outstanding TMEM loads; the registers are operands so that no use of them can be scheduled above the wait
__device__ __forceinline__ void tmem_ld_wait(int (&v)[32]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]), "+r"(v[8]), "+r"(v[9]), "+r"(v[10]),
                 "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15]), "+r"(v[16]), "+r"(v[17]), "+r"(v[18]), "+r"(v[19]), "+r"(v[20]),
                 "+r"(v[21]), "+r"(v[22]), "+r"(v[23]), "+r"(v[24]), "+r"(v[25]), "+r"(v[26]), "+r"(v[27]), "+r"(v[28]), "+r"(v[29]), "+r"(v[30]),
                 "+r"(v[31])
               :: "memory");
}
__device__ __forceinline__ void top2_insert64(unsigned long long k, unsigned long long& a, unsigned long long& b) {
  if (k < a) { b = a; a = k; } else if (k < b) b = k;
}
}  // namespace tc

// train bits [nt][32 B] -> tile images: tile T, chunk j (bits bytes 2j, 2j+1), row r at ((T * 16 + j) * 256 + r) * 16; rows >= nt are zero
__global__ void __launch_bounds__(256) k_expand_i8(const uint8_t* __restrict__ t, long long nt, uint8_t* __restrict__ out) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;     // (tile, j, r) with r fastest
  const long long ntiles = (nt + kTcN - 1) / kTcN;
  if (i >= ntiles * 16 * kTcN) return;
  const int r = (int)(i % kTcN), j = (int)((i / kTcN) % 16);
  const long long T = i / (16 * kTcN), row = T * kTcN + r;
  uint4 v = make_uint4(0u, 0u, 0u, 0u);
  if (row < nt) {
    const uint32_t b = reinterpret_cast<const uint16_t*>(t + 32 * row)[j];
    v.x = tc::spread4(b, false); v.y = tc::spread4(b >> 4, false); v.z = tc::spread4(b >> 8, false); v.w = tc::spread4(b >> 12, false);
  }
  reinterpret_cast<uint4*>(out)[i] = v;
}

__global__ void __launch_bounds__(kTcThreads, 1)
k_knn2_tc(const uint8_t* __restrict__ q, int nq, const uint8_t* __restrict__ texp, int nt, long long trainBase, const int2* __restrict__ exclude,
          unsigned long long* key0, unsigned long long* key1, int init) {
  extern __shared__ __align__(1024) uint8_t tc_smem[];
  uint8_t* sA = tc_smem;                                   // kTcM queries x 256 B, negated
  uint8_t* sB = tc_smem + kTcM * 256;                      // kTcStages x (kTcN train x 256 B)
  __shared__ __align__(8) uint64_t bFull[kTcStages], bEmpty[kTcStages], tFull[2], tEmpty[2];
  __shared__ uint32_t tmemBase;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int q0 = blockIdx.x * kTcM;
  const int ntiles = (nt + kTcN - 1) / kTcN;

  // ---- set-up: query operand, barriers, TMEM
  for (int i = tid; i < kTcM * 16; i += kTcThreads) {
    const int r = i % kTcM, j = i / kTcM;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (q0 + r < nq) {
      const uint32_t b = reinterpret_cast<const uint16_t*>(q + 32ll * (q0 + r))[j];
      v.x = tc::spread4(b, true); v.y = tc::spread4(b >> 4, true); v.z = tc::spread4(b >> 8, true); v.w = tc::spread4(b >> 12, true);
    }
    *reinterpret_cast<uint4*>(sA + ((size_t)j * kTcM + r) * 16) = v;
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy stores -> visible to the tensor core's async proxy
  if (tid == 0) {
    for (int s = 0; s < kTcStages; s++) { tc::mbar_init(&bFull[s], 1); tc::mbar_init(&bEmpty[s], 1); }
    for (int s = 0; s < 2; s++) { tc::mbar_init(&tFull[s], 1); tc::mbar_init(&tEmpty[s], kTcEpiWarps); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kTcEpiWarps + 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(tc::smem_u32(&tmemBase)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tb = tmemBase;

  if (warp == kTcEpiWarps) {
    // ---- producer: one 64 KB bulk copy per train tile into the 2-slot ring
    if (lane == 0) {
      for (int n = 0; n < ntiles; n++) {
        const int s = n % kTcStages, ph = (n / kTcStages) & 1;
        tc::mbar_wait(&bEmpty[s], ph ^ 1);
        tc::mbar_expect_tx(&bFull[s], kTcTileBytes);
        tc::bulk_g2s(sB + (size_t)s * kTcTileBytes, texp + (size_t)n * kTcTileBytes, kTcTileBytes, &bFull[s]);
      }
    }
  } else if (warp == kTcEpiWarps + 1) {
    // ---- MMA issuer: per half h, D_h[128 x kTcN] (TMEM stage s) = A_h[128 x 256] * B[kTcN x 256]^T, K = 32 bytes per instruction
    if (lane == 0) {
      // instruction descriptor: D = S32 (2 << 4), A = B = signed 8 bit (1 << 7, 1 << 10), both K-major, N >> 3 at bit 17, M >> 4 at bit 24
      const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(kTcN >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
      const uint32_t aBase = tc::smem_u32(sA), bBase = tc::smem_u32(sB);
      for (int n = 0; n < ntiles; n++) {
        const int s = n & 1, ph = (n >> 1) & 1;                          // accumulator stage
        const int bs = n % kTcStages, bph = (n / kTcStages) & 1;         // ring slot of the train tile
        tc::mbar_wait(&bFull[bs], bph);
        tc::mbar_wait(&tEmpty[s], ph ^ 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
        for (int h = 0; h < kTcH; h++) {
#pragma unroll
          for (int k = 0; k < 8; k++) {
            const uint64_t da = tc::umma_desc(aBase + (2 * k * kTcM + 128 * h) * 16, kTcM * 16, 128);
            const uint64_t db = tc::umma_desc(bBase + bs * kTcTileBytes + 2 * k * kTcN * 16, kTcN * 16, 128);
            tc::umma_i8(tb + s * 256 + h * kTcN, da, db, idesc, k > 0);
          }
        }
        tc::umma_commit(&bEmpty[bs]);   // the tile's shared-memory slot may be refilled once these MMAs have read it
        tc::umma_commit(&tFull[s]);     // ... and the accumulator stage is complete
      }
    }
  } else {
    // ---- epilogue: thread = TMEM lane = query row
    const int half = warp >> 2, row = half * 128 + (warp & 3) * 32 + lane, qi = q0 + row;
    const bool valid = qi < nq;
    unsigned long long g0 = 0xffffffffffffffffull, g1 = 0xffffffffffffffffull;
    if (valid && !init) { g0 = key0[qi]; g1 = key1[qi]; }
    const int hamCarry = (g1 == 0xffffffffffffffffull) ? 511 : (int)(g1 >> 32);
    int ex0 = 0, ex1 = 0;
    if (valid && exclude) {
      const int2 ex = exclude[qi];
      ex0 = (int)max(0ll, min((long long)nt, (long long)ex.x - trainBase));
      ex1 = (int)max(0ll, min((long long)nt, (long long)ex.y - trainBase));
    }
    uint32_t k0 = 0xffffffffu, k1 = 0xffffffffu;
    int thrDot = 2 * min(511, hamCarry + 1) - 256;   // a value (= 2 * distance - 256) below this can still enter the top-2
    auto process = [&](const int (&v)[32], int colBase) {
      int m = v[0];
#pragma unroll
      for (int i = 1; i < 32; i += 2) m = min(m, min(v[i], v[(i + 1) & 31]));   // (v[0] twice at the end: harmless)
      if (m >= thrDot) return;
#pragma unroll
      for (int i = 0; i < 32; i++) {
        if (v[i] < thrDot) {
          const int col = colBase + i;
          if (col < nt && (col < ex0 || col >= ex1)) {
            const uint32_t key = ((uint32_t)((v[i] + 256) >> 1) << 23) | (uint32_t)col;
            const uint32_t lo = min(key, k0), hi = max(key, k0);
            k0 = lo; k1 = min(k1, hi);
            const int hamLocal = (k1 == 0xffffffffu) ? 511 : (int)(k1 >> 23);
            thrDot = 2 * min(hamLocal, hamCarry + 1) - 256;
          }
        }
      }
    };
    for (int n = 0; n < ntiles; n++) {
      const int s = n & 1, ph = (n >> 1) & 1;
      tc::mbar_wait(&tFull[s], ph);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t taddr = tb + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)(s * 256 + half * kTcN);
      const int colBase = n * kTcN;
      int va[32], vb[32];
      tc::tmem_ld32(taddr, va);
#pragma unroll
      for (int c = 0; c < kTcN / 32; c += 2) {
        tc::tmem_ld_wait(va);
        tc::tmem_ld32(taddr + 32 * (c + 1), vb);
        process(va, colBase + 32 * c);
        tc::tmem_ld_wait(vb);
        if (c + 2 < kTcN / 32) tc::tmem_ld32(taddr + 32 * (c + 2), va);
        else {   // every load of the stage has landed: hand the accumulator back before the last reduction
          asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
          __syncwarp();
          if (lane == 0) tc::mbar_arrive(&tEmpty[s]);
        }
        process(vb, colBase + 32 * (c + 1));
      }
    }
    if (valid) {
      if (k0 != 0xffffffffu) tc::top2_insert64(((unsigned long long)(k0 >> 23) << 32) | (unsigned long long)(trainBase + (k0 & 0x7fffffu)), g0, g1);
      if (k1 != 0xffffffffu) tc::top2_insert64(((unsigned long long)(k1 >> 23) << 32) | (unsigned long long)(trainBase + (k1 & 0x7fffffu)), g0, g1);
      key0[qi] = g0; key1[qi] = g1;
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == kTcEpiWarps + 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tb), "r"(512));
}

// one fold: queries q[nq] against train t[nt] (nt <= 2^23), running keys in key0 / key1 (see k_knn2); workspace >= knn2_tc_workspace_bytes(nt)
inline cudaError_t knn2_tc_launch(const uint8_t* q, int nq, const uint8_t* t, int nt, long long trainBase, const int2* exclude, unsigned long long* key0,
                                  unsigned long long* key1, int init, void* workspace, cudaStream_t st) {
  if (nq <= 0 || nt <= 0) return cudaSuccess;
  if (nt > (1 << 23)) return cudaErrorInvalidValue;
  constexpr int smemBytes = kTcM * 256 + kTcStages * kTcTileBytes;
  cudaError_t e = cudaFuncSetAttribute(k_knn2_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, smemBytes);   // per device: set on every call
  if (e != cudaSuccess) return e;
  const long long ntiles = (nt + kTcN - 1) / kTcN;
  k_expand_i8<<<(unsigned)((ntiles * 16 * kTcN + 255) / 256), 256, 0, st>>>(t, nt, static_cast<uint8_t*>(workspace));
  k_knn2_tc<<<(nq + kTcM - 1) / kTcM, kTcThreads, smemBytes, st>>>(q, nq, static_cast<const uint8_t*>(workspace), nt, trainBase, exclude, key0, key1, init);
  return cudaGetLastError();
}

}  // namespace orbx
