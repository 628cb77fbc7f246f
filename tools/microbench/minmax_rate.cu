// Issue-rate probe for the integer / half min-max instructions FAST could use (sm_100a).
// Prints warp-instructions per clock per SM for each op at full occupancy (8 independent chains per thread).
#include <cstdio>
#include <cstdint>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#define ITERS 4096
template <int OP>
__global__ void __launch_bounds__(256) k(uint32_t* out, uint32_t seed) {
  uint32_t a[8];
#pragma unroll
  for (int i = 0; i < 8; i++) a[i] = seed * (threadIdx.x + 1) + i * 0x01010101u;
  uint32_t b = seed ^ 0x00370019u, c = seed ^ 0x00210044u;
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      if (OP == 0) a[i] = __vimin3_u16x2(a[i], b, c) + 0u, a[i] = __vimax3_u16x2(a[i], c, b);
      if (OP == 1) a[i] = __vminu2(a[i], b), a[i] = __vmaxu2(a[i], c);
      if (OP == 2) a[i] = min(min(a[i], b), c), a[i] = max(max(a[i], c), b);
      if (OP == 3) a[i] = min(a[i], b), a[i] = max(a[i], c);
      if (OP == 4) { __half2 h = *reinterpret_cast<__half2*>(&a[i]); h = __hmin2(h, *reinterpret_cast<__half2*>(&b)); h = __hmax2(h, *reinterpret_cast<__half2*>(&c)); a[i] = *reinterpret_cast<uint32_t*>(&h); }
      if (OP == 5) a[i] = a[i] * b + c, a[i] = a[i] * c + b;                       // IMAD
      if (OP == 6) a[i] = (a[i] & b) ^ c, a[i] = (a[i] | c) ^ b;                   // LOP3
      if (OP == 7) {  // mix: VIMNMX3.U16x2 + HMNMX2 on separate chains
        if (i & 1) { __half2 h = *reinterpret_cast<__half2*>(&a[i]); h = __hmin2(h, *reinterpret_cast<__half2*>(&b)); h = __hmax2(h, *reinterpret_cast<__half2*>(&c)); a[i] = *reinterpret_cast<uint32_t*>(&h); }
        else a[i] = __vimin3_u16x2(a[i], b, c), a[i] = __vimax3_u16x2(a[i], c, b);
      }
      if (OP == 8) {  // mix: VIMNMX3.U16x2 + IMAD
        if (i & 1) a[i] = a[i] * b + c, a[i] = a[i] * c + b;
        else a[i] = __vimin3_u16x2(a[i], b, c), a[i] = __vimax3_u16x2(a[i], c, b);
      }
      if (OP == 9) a[i] = __vimin3_s16x2(a[i], b, c), a[i] = __vimax3_s16x2(a[i], c, b);
      if (OP == 10) a[i] = __vimin3_s32(a[i], b, c), a[i] = __vimax3_s32(a[i], c, b);
      if (OP == 11) a[i] = __vsubus2(a[i], b), a[i] = __vaddus2(a[i], c);          // saturating u16x2 add/sub
      if (OP == 12) a[i] = __vminu4(a[i], b), a[i] = __vmaxu4(a[i], c);
      if (OP == 13) a[i] = __byte_perm(a[i], b, 0x4140), a[i] = __byte_perm(a[i], c, 0x4342);
    }
    b += 0x00010001u; c ^= a[0] & 1;
  }
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) s ^= a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int OP>
void run(const char* name, uint32_t* d, int sms, double ghz) {
  const int blocks = sms * 8;
  k<OP><<<blocks, 256>>>(d, 12345u);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  k<OP><<<blocks, 256>>>(d, 12345u);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double winst = (double)blocks * 8 * ITERS * 16;  // warps * iterations * 16 ops
  printf("%-28s %8.3f ms  %6.2f warp-inst/clk/SM (at %.3f GHz)\n", name, ms, winst / (ms * 1e-3) / (ghz * 1e9) / sms, ghz);
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  const double ghz = khz * 1e-6;
  uint32_t* d; cudaMalloc(&d, (size_t)p.multiProcessorCount * 8 * 256 * 4);
  printf("%s, %d SMs\n", p.name, p.multiProcessorCount);
  run<0>("VIMNMX3.U16x2", d, p.multiProcessorCount, ghz);
  run<1>("VIMNMX.U16x2", d, p.multiProcessorCount, ghz);
  run<2>("VIMNMX3.U32", d, p.multiProcessorCount, ghz);
  run<3>("VIMNMX.U32", d, p.multiProcessorCount, ghz);
  run<4>("HMNMX2", d, p.multiProcessorCount, ghz);
  run<5>("IMAD", d, p.multiProcessorCount, ghz);
  run<6>("LOP3", d, p.multiProcessorCount, ghz);
  run<7>("VIMNMX3.U16x2 + HMNMX2", d, p.multiProcessorCount, ghz);
  run<8>("VIMNMX3.U16x2 + IMAD", d, p.multiProcessorCount, ghz);
  run<9>("VIMNMX3.S16x2", d, p.multiProcessorCount, ghz);
  run<10>("VIMNMX3.S32", d, p.multiProcessorCount, ghz);
  run<11>("vsubus2/vaddus2", d, p.multiProcessorCount, ghz);
  run<12>("vminu4/vmaxu4", d, p.multiProcessorCount, ghz);
  run<13>("PRMT", d, p.multiProcessorCount, ghz);
  return 0;
}
