// Exhaustive check for k_orient_desc2: is sincos((double)a) -> (float) identical to (float)cos((double)a), (float)sin((double)a) for EVERY float
// a in [0, 6.3] (the steering angle of rBRIEF is a float in [0, 2 pi))?  Prints the number of floats checked and of mismatches.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 --fmad=false -o sincos_check sincos_check.cu && ./sincos_check
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__global__ void k(uint32_t first, uint32_t count, unsigned long long* bad) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  unsigned long long mine = 0;
  for (uint32_t j = i; j < count; j += gridDim.x * blockDim.x) {
    const float a = __uint_as_float(first + j);
    const float c0 = (float)cos((double)a), s0 = (float)sin((double)a);
    double sd, cd;
    sincos((double)a, &sd, &cd);
    if (__float_as_uint((float)cd) != __float_as_uint(c0) || __float_as_uint((float)sd) != __float_as_uint(s0)) mine++;
  }
  if (mine) atomicAdd(bad, mine);
}
int main() {
  const float hi = 6.3f;
  const uint32_t last = *reinterpret_cast<const uint32_t*>(&hi);
  unsigned long long* bad; cudaMallocManaged(&bad, 8); *bad = 0;
  k<<<148 * 16, 256>>>(0u, last + 1, bad);
  cudaError_t e = cudaDeviceSynchronize();
  printf("%s: floats checked %u, mismatches %llu\n", cudaGetErrorString(e), last + 1, *bad);
  return 0;
}
