// Issue-rate probe for the Hamming inner loop (sm_100a): POPC alone, the xor+popc+add chain of one 256-bit pair, and the
// whole per-pair update of k_knn2 (8 LOP3 + 8 POPC + adds + key + 3 min/max).  SURVEY.md 10.3: calibrate the popc roof
// before quoting a fraction of it.   nvcc -arch=sm_100a -O3 -o popc_rate popc_rate.cu && ./popc_rate
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#define ITERS 2048
template <int OP>
__global__ void __launch_bounds__(256) k(uint32_t* out, uint32_t seed) {
  uint32_t a[8], q[8];
#pragma unroll
  for (int i = 0; i < 8; i++) { a[i] = seed * (threadIdx.x + 1) + i * 0x01010101u; q[i] = a[i] * 2654435761u; }
  uint32_t k0 = 0xffffffffu, k1 = 0xffffffffu, acc = 0;
  for (int it = 0; it < ITERS; it++) {
    if (OP == 0) {          // 8 independent POPC
#pragma unroll
      for (int i = 0; i < 8; i++) a[i] = __popc(a[i]) + 0x9e3779b9u * it;   // IMAD on the FMA pipe keeps the chain data-dependent
    } else if (OP == 1) {   // one pair: 8 x (xor, popc) + add tree
      int d = 0;
#pragma unroll
      for (int i = 0; i < 8; i++) d += __popc(q[i] ^ a[i]);
      acc += d;
#pragma unroll
      for (int i = 0; i < 8; i++) a[i] += 0x01000193u;
    } else {                // one pair of k_knn2: distance, packed key, running top-2
      int d = 0;
#pragma unroll
      for (int i = 0; i < 8; i++) d += __popc(q[i] ^ a[i]);
      const uint32_t key = ((uint32_t)d << 23) | (uint32_t)it;
      const uint32_t lo = min(key, k0), hi = max(key, k0);
      k0 = lo; k1 = min(k1, hi);
      a[it & 7] += 0x01000193u;
    }
  }
  uint32_t s = acc ^ k0 ^ k1;
#pragma unroll
  for (int i = 0; i < 8; i++) s ^= a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int OP>
void run(const char* name, uint32_t* d, int sms, double ghz, double popcPerIter, double pairsPerIter) {
  const int blocks = sms * 8;
  k<OP><<<blocks, 256>>>(d, 12345u);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  k<OP><<<blocks, 256>>>(d, 12345u);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double threads = (double)blocks * 256, sec = ms * 1e-3;
  const double popc = threads * ITERS * popcPerIter / sec;
  printf("%-34s %8.3f ms  %6.2f popc/clk/SM  %.3e popc/s", name, ms, popc / (ghz * 1e9) / sms, popc);
  if (pairsPerIter > 0) printf("  %.3e pairs/s", threads * ITERS * pairsPerIter / sec);
  printf("\n");
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  const double ghz = khz * 1e-6;
  uint32_t* d; cudaMalloc(&d, (size_t)p.multiProcessorCount * 8 * 256 * 4);
  printf("%s, %d SMs, %.3f GHz max\n", p.name, p.multiProcessorCount, ghz);
  run<0>("POPC (+IMAD)", d, p.multiProcessorCount, ghz, 8, 0);
  run<1>("pair: 8 x (LOP3 + POPC) + adds", d, p.multiProcessorCount, ghz, 8, 1);
  run<2>("pair + key + top-2 (k_knn2 loop)", d, p.multiProcessorCount, ghz, 8, 1);
  return 0;
}
