// Step 2 of the tensor-core all-pairs fold: the whole 2-NN fold of a query set against a train set on tcgen05 (development
// harness of csrc/knn2_tc.cuh: same kernels, checked here against a plain popc kernel and timed).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I../../3_orb_slam3_selfnote_b200/csrc -o tc_knn2 tc_knn2.cu && ./tc_knn2 [nq] [nt]
#include <climits>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#include "knn2_tc.cuh"

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

static const unsigned long long kNone = 0xffffffffffffffffull;

// reference: thread = query, plain popc, 64-bit keys (dist << 32 | global index), optional exclusion range
__global__ void k_ref(const uint8_t* q, int nq, const uint8_t* t, int nt, long long base, const int2* ex, unsigned long long* k0o, unsigned long long* k1o) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nq) return;
  uint32_t a[8];
  for (int w = 0; w < 8; w++) a[w] = reinterpret_cast<const uint32_t*>(q + 32ll * i)[w];
  unsigned long long k0 = kNone, k1 = kNone;
  for (int j = 0; j < nt; j++) {
    const long long gj = base + j;
    if (ex && gj >= ex[i].x && gj < ex[i].y) continue;
    int d = 0;
    for (int w = 0; w < 8; w++) d += __popc(a[w] ^ reinterpret_cast<const uint32_t*>(t + 32ll * j)[w]);
    const unsigned long long k = ((unsigned long long)d << 32) | (unsigned long long)gj;
    if (k < k0) { k1 = k0; k0 = k; } else if (k < k1) k1 = k;
  }
  k0o[i] = k0; k1o[i] = k1;
}

int main(int argc, char** argv) {
  const int nq = argc > 1 ? atoi(argv[1]) : 148 * 128 * 2;
  const int nt = argc > 2 ? atoi(argv[2]) : 256 * 400 + 77;
  const long long base = 1000000;
  std::vector<uint8_t> q((size_t)nq * 32), t((size_t)nt * 32);
  srand(11);
  for (auto& x : q) x = rand() & 255;
  for (auto& x : t) x = rand() & 255;
  const bool ties = !(argc > 3 && atoi(argv[3]) == 1);   // default: few distinct distances (many ties); 1 = fully random descriptors
  if (ties) {
    for (size_t i = 0; i < t.size(); i++) if ((i & 31) >= 6) t[i] = t[i & 31];
    for (size_t i = 0; i < q.size(); i++) if ((i & 31) >= 6) q[i] = t[i & 31];
  }
  for (int i = 0; i < 32; i++) { t[(size_t)12345 % nt * 32 + i] = q[7 * 32 + i]; t[(size_t)(nt - 1) * 32 + i] = q[9 * 32 + i]; }   // exact duplicates, one in the ragged last tile
  std::vector<int2> ex(nq);
  for (int i = 0; i < nq; i++) { const int kf = i / 1000; ex[i] = make_int2((int)base + kf * 1000, (int)base + kf * 1000 + 1000); }   // "own keyframe" ranges
  uint8_t *dq, *dt; int2* dex; unsigned long long *r0, *r1, *g0, *g1; int8_t* dexp;
  CK(cudaMalloc(&dq, q.size())); CK(cudaMalloc(&dt, t.size())); CK(cudaMalloc(&dex, nq * sizeof(int2)));
  CK(cudaMalloc(&r0, nq * 8)); CK(cudaMalloc(&r1, nq * 8)); CK(cudaMalloc(&g0, nq * 8)); CK(cudaMalloc(&g1, nq * 8));
  const size_t expBytes = orbx::knn2_tc_workspace_bytes(nt);
  CK(cudaMalloc(&dexp, expBytes));
  CK(cudaMemcpy(dq, q.data(), q.size(), cudaMemcpyHostToDevice)); CK(cudaMemcpy(dt, t.data(), t.size(), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dex, ex.data(), nq * sizeof(int2), cudaMemcpyHostToDevice));
  k_ref<<<(nq + 127) / 128, 128>>>(dq, nq, dt, nt, base, dex, r0, r1);
  CK(cudaDeviceSynchronize());
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e9f;
  for (int rep = 0; rep < 4; rep++) {
    cudaEventRecord(e0);
    CK(orbx::knn2_tc_launch(dq, nq, dt, nt, base, dex, g0, g1, /*init=*/1, dexp, 0));
    cudaEventRecord(e1);
    CK(cudaDeviceSynchronize());
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  std::vector<unsigned long long> R0(nq), R1(nq), G0(nq), G1(nq);
  CK(cudaMemcpy(R0.data(), r0, nq * 8, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(R1.data(), r1, nq * 8, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(G0.data(), g0, nq * 8, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(G1.data(), g1, nq * 8, cudaMemcpyDeviceToHost));
  int bad = 0;
  for (int i = 0; i < nq; i++)
    if (R0[i] != G0[i] || R1[i] != G1[i]) {
      if (bad < 8) printf("query %d: ref (%llu,%llu | %llu,%llu) got (%llu,%llu | %llu,%llu)\n", i, R0[i] >> 32, R0[i] & 0xffffffffull, R1[i] >> 32,
                          R1[i] & 0xffffffffull, G0[i] >> 32, G0[i] & 0xffffffffull, G1[i] >> 32, G1[i] & 0xffffffffull);
      bad++;
    }
  // second launch folding a second "chunk" into the running keys (init = 0), against the reference over both ranges
  printf("nq %d nt %d: %d mismatches; %.3f ms -> %.3e pairs/s\n", nq, nt, bad, best, (double)nq * nt / (best * 1e-3));
  return bad != 0;
}
