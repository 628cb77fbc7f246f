// the Frame glue on either side of the path, so that keypoints go from the extractor to the searches
// (undistortion) and map points from the local map to SearchByProjection (frustum test, projection, scale prediction)
// without a host round trip.
//
//   Frame::UndistortKeyPoints / ComputeImageBounds   src/Frame.cc:837-870, 872-899  (cv::undistortPoints, 5 iterations, double)
//   Frame::isInFrustum, Nleft == -1 branch           src/Frame.cc:572-642
//   MapPoint::PredictScale                           src/MapPoint.cc:587-602
//   Pinhole::project                                 src/CameraModels/Pinhole.cpp:46-49
//
// Thread = one keypoint / one map point; the arithmetic mirrors the reference's evaluation order operation by operation
// (float where the reference's cv::Mat code is float, double where OpenCV accumulates in double), no contraction
// (--fmad=false), IEEE division and square root.  log(ratio) is (float)log((double)ratio) -- see DESIGN.md, "portable" modes.
#include <math.h>

#include "../../include/orbx.h"
#include "frustum.cuh"
#include "orbx_internal.h"

using namespace orbx;

// defined in orbx_preproc.cu
struct orbx_preproc;
namespace orbx {
int preproc_device(const orbx_preproc* p);
cudaStream_t preproc_stream(const orbx_preproc* p);
int preproc_scratch(orbx_preproc* p, size_t inBytes, size_t outBytes, uint8_t** dIn, uint8_t** dOut);
}  // namespace orbx

namespace {

struct UndistArgs {
  const orbx_keypoint* in; orbx_keypoint* out; int n;
  double fx, fy, cx, cy, ifx, ify, k1, k2, p1, p2, k3;
  int identity;  // mDistCoef.at<float>(0) == 0.0 -> mvKeysUn = mvKeys (Frame.cc:839-843)
};

__global__ void __launch_bounds__(256) k_undistort(const __grid_constant__ UndistArgs A) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= A.n) return;
  orbx_keypoint kp = A.in[i];
  if (!A.identity) {
    double x = kp.x, y = kp.y;
    const double u = x, v = y;
    x = (x - A.cx) * A.ifx; y = (y - A.cy) * A.ify;
    const double x0 = x, y0 = y;
    for (int j = 0; j < 5; ++j) {
      const double r2 = x * x + y * y;
      const double icdist = 1.0 / (1 + ((A.k3 * r2 + A.k2) * r2 + A.k1) * r2);
      if (icdist < 0) { x = (u - A.cx) * A.ifx; y = (v - A.cy) * A.ify; break; }
      const double deltaX = 2 * A.p1 * x * y + A.p2 * (r2 + 2 * x * x);
      const double deltaY = A.p1 * (r2 + 2 * y * y) + 2 * A.p2 * x * y;
      x = (x0 - deltaX) * icdist; y = (y0 - deltaY) * icdist;
    }
    kp.x = (float)(A.fx * x + A.cx); kp.y = (float)(A.fy * y + A.cy);
  }
  A.out[i] = kp;
}

struct FrustumArgs {
  orbx_frustum cam;
  int n;
  const float* P; const float* normal; const float* maxDist; const float* minDist;
  uint8_t* inView; float* projX; float* projY; float* projXR; float* depth; int* level; float* viewCos;
};

__global__ void __launch_bounds__(256) k_frustum(const __grid_constant__ FrustumArgs A) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= A.n) return;
  const FrustumOut o = frustum_point(A.cam, A.P[3 * i], A.P[3 * i + 1], A.P[3 * i + 2], A.normal[3 * i], A.normal[3 * i + 1], A.normal[3 * i + 2],
                                     A.maxDist[i], A.minDist[i]);
  A.inView[i] = o.inView; A.projX[i] = o.projX; A.projY[i] = o.projY; A.projXR[i] = o.projXR; A.depth[i] = o.depth; A.level[i] = o.level;
  A.viewCos[i] = o.viewCos;
}

inline size_t al16(size_t v) { return (v + 15) & ~(size_t)15; }

}  // namespace

extern "C" {

int orbx_undistort_keypoints(orbx_preproc* p, const orbx_keypoint* kps, int n, float fx, float fy, float cx, float cy, const float* dist5,
                             orbx_keypoint* kps_un, int mem, void* stream) {
  if (!p || !dist5 || n < 0 || (n > 0 && (!kps || !kps_un))) { set_error("undistort: bad argument"); return ORBX_ERR_ARG; }
  if (mem != ORBX_MEM_HOST && mem != ORBX_MEM_DEVICE) { set_error("mem must be ORBX_MEM_HOST or ORBX_MEM_DEVICE"); return ORBX_ERR_ARG; }
  if (n == 0) return ORBX_OK;
  CK(cudaSetDevice(preproc_device(p)));
  UndistArgs A{};
  A.n = n; A.fx = fx; A.fy = fy; A.cx = cx; A.cy = cy; A.ifx = 1. / A.fx; A.ify = 1. / A.fy;
  A.k1 = dist5[0]; A.k2 = dist5[1]; A.p1 = dist5[2]; A.p2 = dist5[3]; A.k3 = dist5[4];
  A.identity = dist5[0] == 0.0f;
  cudaStream_t st = (mem == ORBX_MEM_DEVICE && stream) ? (cudaStream_t)stream : preproc_stream(p);
  const size_t bytes = (size_t)n * sizeof(orbx_keypoint);
  if (mem == ORBX_MEM_HOST) {
    uint8_t *dIn, *dOut;
    int rc = preproc_scratch(p, bytes, bytes, &dIn, &dOut);
    if (rc) return rc;
    CK(cudaMemcpyAsync(dIn, kps, bytes, cudaMemcpyHostToDevice, st));
    A.in = (const orbx_keypoint*)dIn; A.out = (orbx_keypoint*)dOut;
  } else {
    A.in = kps; A.out = kps_un;
  }
  k_undistort<<<(n + 255) / 256, 256, 0, st>>>(A);
  CK(cudaGetLastError());
  if (mem == ORBX_MEM_HOST) {
    CK(cudaMemcpyAsync(kps_un, A.out, bytes, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
  }
  return ORBX_OK;
}

int orbx_is_in_frustum(orbx_preproc* p, const orbx_frustum* cam, int n, const float* world_pos, const float* normal, const float* max_dist,
                       const float* min_dist, uint8_t* in_view, float* proj_x, float* proj_y, float* proj_xr, float* depth, int* level,
                       float* view_cos, int mem, void* stream) {
  if (!p || !cam || n < 0) { set_error("is_in_frustum: bad argument"); return ORBX_ERR_ARG; }
  if (n > 0 && (!world_pos || !normal || !max_dist || !min_dist || !in_view || !proj_x || !proj_y || !proj_xr || !depth || !level || !view_cos)) {
    set_error("is_in_frustum: null array");
    return ORBX_ERR_ARG;
  }
  if (mem != ORBX_MEM_HOST && mem != ORBX_MEM_DEVICE) { set_error("mem must be ORBX_MEM_HOST or ORBX_MEM_DEVICE"); return ORBX_ERR_ARG; }
  if (n == 0) return ORBX_OK;
  CK(cudaSetDevice(preproc_device(p)));
  FrustumArgs A{};
  A.cam = *cam; A.n = n;
  cudaStream_t st = (mem == ORBX_MEM_DEVICE && stream) ? (cudaStream_t)stream : preproc_stream(p);
  const size_t b3 = al16((size_t)n * 12), b1 = al16((size_t)n * 4), bu = al16((size_t)n);
  uint8_t *dIn = nullptr, *dOut = nullptr;
  if (mem == ORBX_MEM_HOST) {
    int rc = preproc_scratch(p, 2 * b3 + 2 * b1, bu + 6 * b1, &dIn, &dOut);
    if (rc) return rc;
    CK(cudaMemcpyAsync(dIn, world_pos, (size_t)n * 12, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(dIn + b3, normal, (size_t)n * 12, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(dIn + 2 * b3, max_dist, (size_t)n * 4, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(dIn + 2 * b3 + b1, min_dist, (size_t)n * 4, cudaMemcpyHostToDevice, st));
    A.P = (const float*)dIn; A.normal = (const float*)(dIn + b3); A.maxDist = (const float*)(dIn + 2 * b3); A.minDist = (const float*)(dIn + 2 * b3 + b1);
    A.inView = dOut; A.projX = (float*)(dOut + bu); A.projY = (float*)(dOut + bu + b1); A.projXR = (float*)(dOut + bu + 2 * b1);
    A.depth = (float*)(dOut + bu + 3 * b1); A.level = (int*)(dOut + bu + 4 * b1); A.viewCos = (float*)(dOut + bu + 5 * b1);
  } else {
    A.P = world_pos; A.normal = normal; A.maxDist = max_dist; A.minDist = min_dist;
    A.inView = in_view; A.projX = proj_x; A.projY = proj_y; A.projXR = proj_xr; A.depth = depth; A.level = level; A.viewCos = view_cos;
  }
  k_frustum<<<(n + 255) / 256, 256, 0, st>>>(A);
  CK(cudaGetLastError());
  if (mem == ORBX_MEM_HOST) {
    CK(cudaMemcpyAsync(in_view, A.inView, (size_t)n, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(proj_x, A.projX, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(proj_y, A.projY, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(proj_xr, A.projXR, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(depth, A.depth, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(level, A.level, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(view_cos, A.viewCos, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
  }
  return ORBX_OK;
}

}  // extern "C"
