// Step 1 of the tensor-core all-pairs fold: ONE 128 x 256 tile of Hamming distances through tcgen05.mma kind::i8.
// Bits are expanded to +1 / -1 int8 operands, so dot(a, b) = 256 - 2 * hamming(a, b).  Operands sit in shared memory in the
// K-major no-swizzle canonical layout ([16-byte K chunk][row][16 B]: core matrix = 8 rows x 16 B contiguous, SBO = 128 B,
// LBO = rows * 16 B); the accumulator is 128 lanes x 256 columns of TMEM, read back with tcgen05.ld.32x32b.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tc_hamming tc_hamming.cu && ./tc_hamming
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

constexpr int kM = 128, kN = 256;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// 4 bits -> 4 bytes of +1 (0x01) / -1 (0xff)
__device__ __forceinline__ uint32_t spread4(uint32_t x) {
  const uint32_t w = ((x & 0xfu) * 0x00204081u) & 0x01010101u;
  return w | ((w ^ 0x01010101u) * 0xffu);
}
// rows x 32 bytes of bits -> [16 chunks][rows][16 B] of int8
__device__ __forceinline__ void expand_tile(const uint8_t* bits, int rows, uint8_t* dst, int tid, int nthreads) {
  for (int i = tid; i < rows * 16; i += nthreads) {
    const int r = i % rows, j = i / rows;
    const uint32_t b = reinterpret_cast<const uint16_t*>(bits + 32 * r)[j];
    uint4 v;
    v.x = spread4(b); v.y = spread4(b >> 4); v.z = spread4(b >> 8); v.w = spread4(b >> 12);
    *reinterpret_cast<uint4*>(dst + ((size_t)j * rows + r) * 16) = v;
  }
}

__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return (uint64_t)((saddr >> 4) & 0x3fffu) | ((uint64_t)((lbo_bytes >> 4) & 0x3fffu) << 16) | ((uint64_t)((sbo_bytes >> 4) & 0x3fffu) << 32) |
         (1ull << 46);   // version 1 (Blackwell), layout type 0 = no swizzle
}

__global__ void __launch_bounds__(128) k_tc_tile(const uint8_t* q, const uint8_t* t, int* D) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sA = smem;                 // 128 x 256 B
  uint8_t* sB = smem + kM * 256;      // 256 x 256 B
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmemBase;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  expand_tile(q, kM, sA, tid, 128);
  expand_tile(t, kN, sB, tid, 128);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy stores -> visible to the tensor core's async proxy
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(smem_u32(&tmemBase)), "r"(256));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tb = tmemBase;
  if (tid == 0) {
    // instruction descriptor: D = S32 (2 << 4), A = B = signed 8 bit (1 << 7, 1 << 10), both K-major, N >> 3 at bit 17, M >> 4 at bit 24
    const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(kN >> 3) << 17) | ((uint32_t)(kM >> 4) << 24);
    const uint32_t aBase = smem_u32(sA), bBase = smem_u32(sB);
#pragma unroll
    for (int s = 0; s < 8; s++) {   // K = 32 bytes per instruction = two 16-byte chunks
      const uint64_t da = umma_desc(aBase + 2 * s * kM * 16, kM * 16, 128);
      const uint64_t db = umma_desc(bBase + 2 * s * kN * 16, kN * 16, 128);
      const uint32_t acc = s > 0;
      asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                   "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t}"
                   :: "r"(tb), "l"(da), "l"(db), "r"(idesc), "r"(acc), "r"(0), "r"(0), "r"(0), "r"(0) : "memory");
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(&bar)) : "memory");
  }
  // everyone waits for the accumulator
  {
    const uint32_t addr = smem_u32(&bar);
    asm volatile("{\n\t.reg .pred P1;\n\tWAIT_LOOP:\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t@P1 bra DONE;\n\tbra WAIT_LOOP;\n\tDONE:\n\t}"
                 :: "r"(addr), "r"(0) : "memory");
  }
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const int row = warp * 32 + lane;
  for (int c = 0; c < kN / 32; c++) {
    uint32_t v[32];
    const uint32_t taddr = tb + ((uint32_t)(warp * 32) << 16) + (uint32_t)(c * 32);
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, "
                 "%22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]),
                   "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]),
                   "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]),
                   "=r"(v[31])
                 : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; i++) D[row * kN + c * 32 + i] = (int)v[i];
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tb), "r"(256));
}

int main() {
  std::vector<uint8_t> q(kM * 32), t(kN * 32);
  srand(7);
  for (auto& x : q) x = rand() & 255;
  for (auto& x : t) x = rand() & 255;
  for (int i = 0; i < 32; i++) t[5 * 32 + i] = q[3 * 32 + i];   // an exact duplicate: distance 0
  uint8_t *dq, *dt; int* dD;
  CK(cudaMalloc(&dq, q.size())); CK(cudaMalloc(&dt, t.size())); CK(cudaMalloc(&dD, kM * kN * 4));
  CK(cudaMemcpy(dq, q.data(), q.size(), cudaMemcpyHostToDevice)); CK(cudaMemcpy(dt, t.data(), t.size(), cudaMemcpyHostToDevice));
  CK(cudaMemset(dD, 0x7f, kM * kN * 4));
  const int smemBytes = (kM + kN) * 256;
  CK(cudaFuncSetAttribute(k_tc_tile, cudaFuncAttributeMaxDynamicSharedMemorySize, smemBytes));
  k_tc_tile<<<1, 128, smemBytes>>>(dq, dt, dD);
  CK(cudaGetLastError());
  CK(cudaDeviceSynchronize());
  std::vector<int> D(kM * kN);
  CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
  int bad = 0;
  for (int i = 0; i < kM; i++)
    for (int j = 0; j < kN; j++) {
      int ham = 0;
      for (int b = 0; b < 32; b++) ham += __builtin_popcount(q[i * 32 + b] ^ t[j * 32 + b]);
      const int want = 256 - 2 * ham;
      if (D[i * kN + j] != want) { if (bad < 10) printf("mismatch at (%d,%d): got %d want %d\n", i, j, D[i * kN + j], want); bad++; }
    }
  printf("tile 128x256: %d mismatches; D[3][5] = %d (expect 256)\n", bad, D[3 * kN + 5]);
  return bad != 0;
}
