// All-pairs keyframe-descriptor 2-NN sweep (BASELINE.json config 5, SURVEY.md 8-e "the one exchange step") behind the C-ABI.
//
// Every rank owns `rows` descriptors (rows / desc_per_kf keyframes); global row = rank * rows + local row.  For each of its own
// descriptors a rank wants the two nearest descriptors among ALL descriptors of all OTHER keyframes.  The shards travel by
// ncclAllGather in `nchunks` pieces on a communication stream; while piece c + 1 is in flight over NVLink, the compute stream
// folds piece c (world x chunk rows) into the running (best, second) keys of the rank's own queries with the 2-NN tile kernel
// (orbx_knn2_chunk_device).  No reduction afterwards: each rank owns its queries.  Two gather buffers, ordered by events:
//   gather(c+1) waits for fold(c-1) (same buffer), fold(c) waits for gather(c).
//
// NCCL is bound at run time (dlopen of libnccl.so.2, preferring the copy already loaded in the process, e.g. the one PyTorch
// ships) so that liborbx.so has no link-time NCCL dependency for the single-GPU callers.  The few NCCL declarations used here
// follow nccl.h (2.27/2.28: ncclUniqueId = 128 bytes passed by value, ncclUint8 = 1).
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <mutex>
#include <vector>

#include "orbx_internal.h"

using namespace orbx;

namespace {

struct NcclId { char internal[128]; };
typedef void* NcclComm;
struct NcclApi {
  void* so = nullptr;
  int (*GetUniqueId)(NcclId*) = nullptr;
  int (*CommInitRank)(NcclComm*, int, NcclId, int) = nullptr;
  int (*CommDestroy)(NcclComm) = nullptr;
  int (*CommCount)(NcclComm, int*) = nullptr;
  int (*CommUserRank)(NcclComm, int*) = nullptr;
  int (*AllGather)(const void*, void*, size_t, int, NcclComm, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  int (*GetVersion)(int*) = nullptr;
};

NcclApi* nccl() {
  static NcclApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    void* so = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);   // the copy the process already uses, if any
    if (!so) so = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!so) so = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!so) return;
    api.so = so;
    api.GetUniqueId = (int (*)(NcclId*))dlsym(so, "ncclGetUniqueId");
    api.CommInitRank = (int (*)(NcclComm*, int, NcclId, int))dlsym(so, "ncclCommInitRank");
    api.CommDestroy = (int (*)(NcclComm))dlsym(so, "ncclCommDestroy");
    api.CommCount = (int (*)(NcclComm, int*))dlsym(so, "ncclCommCount");
    api.CommUserRank = (int (*)(NcclComm, int*))dlsym(so, "ncclCommUserRank");
    api.AllGather = (int (*)(const void*, void*, size_t, int, NcclComm, cudaStream_t))dlsym(so, "ncclAllGather");
    api.GetErrorString = (const char* (*)(int))dlsym(so, "ncclGetErrorString");
    api.GetVersion = (int (*)(int*))dlsym(so, "ncclGetVersion");
  });
  if (!api.so || !api.GetUniqueId || !api.CommInitRank || !api.CommDestroy || !api.CommCount || !api.CommUserRank || !api.AllGather) return nullptr;
  return &api;
}

#define NCK(call)                                                                                              \
  do {                                                                                                         \
    int r_ = (call);                                                                                           \
    if (r_ != 0) {                                                                                             \
      set_error("%s:%d %s -> NCCL error %d (%s)", __FILE__, __LINE__, #call, r_,                               \
                nccl()->GetErrorString ? nccl()->GetErrorString(r_) : "?");                                    \
      return ORBX_ERR_CUDA;                                                                                    \
    }                                                                                                          \
  } while (0)

struct Scope {   // releases the sweep's streams / events / buffers on every exit path
  std::vector<cudaEvent_t> ev;
  std::vector<void*> dev;
  cudaStream_t comm = nullptr, comp = nullptr;
  ~Scope() {
    if (comm) cudaStreamSynchronize(comm);
    if (comp) cudaStreamSynchronize(comp);
    for (cudaEvent_t e : ev) cudaEventDestroy(e);
    for (void* p : dev) cudaFree(p);
    if (comm) cudaStreamDestroy(comm);
    if (comp) cudaStreamDestroy(comp);
  }
};

__global__ void k_allpairs_exclude(int2* ex, long long rows, int descPerKf, long long rankBase) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= rows) return;
  const long long s = rankBase + (i / descPerKf) * descPerKf;
  ex[i] = make_int2((int)s, (int)(s + descPerKf));
}

}  // namespace

extern "C" {

int orbx_nccl_version(void) {
  NcclApi* n = nccl();
  int v = 0;
  if (!n || !n->GetVersion || n->GetVersion(&v) != 0) return 0;
  return v;
}

int orbx_nccl_unique_id(uint8_t id[128]) {
  NcclApi* n = nccl();
  if (!n) { set_error("libnccl.so.2 could not be loaded"); return ORBX_ERR_UNSUPPORTED; }
  if (!id) return ORBX_ERR_ARG;
  NcclId u;
  NCK(n->GetUniqueId(&u));
  memcpy(id, u.internal, 128);
  return ORBX_OK;
}

int orbx_nccl_comm_init(const uint8_t id[128], int nranks, int rank, int device, void** comm) {
  NcclApi* n = nccl();
  if (!n) { set_error("libnccl.so.2 could not be loaded"); return ORBX_ERR_UNSUPPORTED; }
  if (!id || !comm || nranks < 1 || rank < 0 || rank >= nranks) return ORBX_ERR_ARG;
  CK(cudaSetDevice(device));
  NcclId u;
  memcpy(u.internal, id, 128);
  NcclComm c = nullptr;
  NCK(n->CommInitRank(&c, nranks, u, rank));
  *comm = c;
  return ORBX_OK;
}

int orbx_nccl_comm_destroy(void* comm) {
  NcclApi* n = nccl();
  if (!n || !comm) return ORBX_ERR_ARG;
  NCK(n->CommDestroy((NcclComm)comm));
  return ORBX_OK;
}

int orbx_allpairs_knn2(orbx_matcher* m, void* nccl_comm, const uint8_t* d_desc, long long rows, int desc_per_kf, int nchunks,
                       long long* idx, int* dist, orbx_allpairs_stats* stats) {
  if (!m || !d_desc || rows < 1 || desc_per_kf < 1 || rows % desc_per_kf != 0 || !idx || !dist) { set_error("allpairs: bad argument"); return ORBX_ERR_ARG; }
  int world = 1, rank = 0;
  NcclApi* n = nullptr;
  if (nccl_comm) {
    n = nccl();
    if (!n) { set_error("libnccl.so.2 could not be loaded"); return ORBX_ERR_UNSUPPORTED; }
    NCK(n->CommCount((NcclComm)nccl_comm, &world));
    NCK(n->CommUserRank((NcclComm)nccl_comm, &rank));
  }
  if ((long long)world * rows >= (1ll << 31)) { set_error("allpairs: more than 2^31 descriptors in total"); return ORBX_ERR_UNSUPPORTED; }
  nchunks = (int)std::max(1ll, std::min<long long>(nchunks, rows));
  while ((rows + nchunks - 1) / nchunks > (1 << 23)) nchunks *= 2;   // the tile kernel indexes a chunk with 23 bits
  CK(cudaSetDevice(matcher_device(m)));
  Scope S;
  CK(cudaStreamCreateWithFlags(&S.comm, cudaStreamNonBlocking));
  CK(cudaStreamCreateWithFlags(&S.comp, cudaStreamNonBlocking));
  auto newEvent = [&](cudaEvent_t* e, bool timing) {
    cudaError_t r = cudaEventCreateWithFlags(e, timing ? cudaEventDefault : cudaEventDisableTiming);
    if (r == cudaSuccess) S.ev.push_back(*e);
    return r;
  };
  auto devAlloc = [&](void** p, size_t bytes) {
    cudaError_t r = cudaMalloc(p, bytes);
    if (r == cudaSuccess) S.dev.push_back(*p);
    return r;
  };
  std::vector<long long> bounds(nchunks + 1);
  long long maxr = 0;
  for (int c = 0; c <= nchunks; c++) bounds[c] = (long long)c * rows / nchunks;
  for (int c = 0; c < nchunks; c++) maxr = std::max(maxr, bounds[c + 1] - bounds[c]);
  int2* dEx = nullptr; unsigned long long *dK0 = nullptr, *dK1 = nullptr; uint8_t* buf[2] = {nullptr, nullptr};
  CK(devAlloc((void**)&dEx, (size_t)rows * sizeof(int2)));
  CK(devAlloc((void**)&dK0, (size_t)rows * 8));
  CK(devAlloc((void**)&dK1, (size_t)rows * 8));
  if (world > 1)
    for (int i = 0; i < 2; i++) CK(devAlloc((void**)&buf[i], (size_t)world * maxr * 32));
  k_allpairs_exclude<<<(unsigned)((rows + 255) / 256), 256, 0, S.comp>>>(dEx, rows, desc_per_kf, (long long)rank * rows);
  CK(cudaGetLastError());

  std::vector<cudaEvent_t> gBeg(nchunks), gEnd(nchunks), fBeg(nchunks), fEnd(nchunks);
  for (int c = 0; c < nchunks; c++) {
    CK(newEvent(&gBeg[c], true)); CK(newEvent(&gEnd[c], true)); CK(newEvent(&fBeg[c], true)); CK(newEvent(&fEnd[c], true));
  }
  cudaEvent_t t0, t1;
  CK(newEvent(&t0, true)); CK(newEvent(&t1, true));
  CK(cudaEventRecord(t0, S.comp));
  CK(cudaStreamWaitEvent(S.comm, t0, 0));

  auto gather = [&](int c) -> int {
    const long long a = bounds[c], b = bounds[c + 1];
    if (c >= 2) CK(cudaStreamWaitEvent(S.comm, fEnd[c - 2], 0));   // the buffer's previous contents have been folded
    CK(cudaEventRecord(gBeg[c], S.comm));
    NCK(n->AllGather(d_desc + a * 32, buf[c & 1], (size_t)(b - a) * 32, /*ncclUint8*/ 1, (NcclComm)nccl_comm, S.comm));
    CK(cudaEventRecord(gEnd[c], S.comm));
    return ORBX_OK;
  };
  int init = 1;
  if (world > 1) { const int rc = gather(0); if (rc != ORBX_OK) return rc; }
  for (int c = 0; c < nchunks; c++) {
    const long long a = bounds[c], b = bounds[c + 1];
    if (world > 1) {
      if (c + 1 < nchunks) { const int rc = gather(c + 1); if (rc != ORBX_OK) return rc; }   // travels while chunk c is folded
      CK(cudaStreamWaitEvent(S.comp, gEnd[c], 0));
    }
    CK(cudaEventRecord(fBeg[c], S.comp));
    for (int r = 0; r < world; r++) {
      const uint8_t* train = world > 1 ? buf[c & 1] + (size_t)r * (b - a) * 32 : d_desc + a * 32;
      const int rc = orbx_knn2_chunk_device(m, d_desc, (int)rows, train, (int)(b - a), (long long)r * rows + a, reinterpret_cast<const int*>(dEx), dK0,
                                            dK1, init, S.comp);
      if (rc != ORBX_OK) return rc;
      init = 0;
    }
    CK(cudaEventRecord(fEnd[c], S.comp));
  }
  CK(cudaEventRecord(t1, S.comp));
  std::vector<unsigned long long> h0(rows), h1(rows);
  CK(cudaMemcpyAsync(h0.data(), dK0, (size_t)rows * 8, cudaMemcpyDeviceToHost, S.comp));
  CK(cudaMemcpyAsync(h1.data(), dK1, (size_t)rows * 8, cudaMemcpyDeviceToHost, S.comp));
  CK(cudaStreamSynchronize(S.comp));
  CK(cudaStreamSynchronize(S.comm));
  const unsigned long long none = 0xffffffffffffffffull;
  for (long long i = 0; i < rows; i++) {
    const bool a = h0[i] != none, b = h1[i] != none;
    idx[2 * i] = a ? (long long)(h0[i] & 0xffffffffu) : -1; dist[2 * i] = a ? (int)(h0[i] >> 32) : -1;
    idx[2 * i + 1] = b ? (long long)(h1[i] & 0xffffffffu) : -1; dist[2 * i + 1] = b ? (int)(h1[i] >> 32) : -1;
  }
  if (stats) {
    memset(stats, 0, sizeof *stats);
    stats->world = world; stats->rank = rank; stats->nchunks = nchunks;
    stats->pairs = (double)rows * ((double)world * rows - desc_per_kf);
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, t0, t1)); stats->total_ms = ms;
    for (int c = 0; c < nchunks; c++) {
      CK(cudaEventElapsedTime(&ms, fBeg[c], fEnd[c])); stats->fold_ms += ms;
      if (world > 1) { CK(cudaEventElapsedTime(&ms, gBeg[c], gEnd[c])); stats->gather_ms += ms; }
    }
    stats->gather_bytes = world > 1 ? (double)world * rows * 32 : 0.0;
    const double hidden = stats->gather_ms + stats->fold_ms - stats->total_ms;
    stats->overlap_frac = (world > 1 && stats->gather_ms > 0) ? std::max(0.0, std::min(1.0, hidden / stats->gather_ms)) : 0.0;
  }
  return ORBX_OK;
}

}  // extern "C"
