// Frame::isInFrustum for one map point (src/Frame.cc:572-642, Nleft == -1 branch) with MapPoint::PredictScale
// (src/MapPoint.cc:587-602) and Pinhole::project (src/CameraModels/Pinhole.cpp:46-49).  Shared by k_frustum (orbx_frame.cu) and
// the device-resident SearchLocalPoints path (orbx_match.cu).  Evaluation order = the reference's cv::Mat float code: each row of
// Rcw * P as ((a0*b0 + a1*b1) + a2*b2) in float, cv::norm / Mat::dot accumulated in double; no contraction.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/orbx.h"

namespace orbx {

struct FrustumOut { uint8_t inView; float projX, projY, projXR, depth, viewCos; int level; };

__device__ __forceinline__ FrustumOut frustum_point(const orbx_frustum& C, float p0, float p1, float p2, float n0, float n1, float n2,
                                                    float maxDistance, float minDistance) {
  // maxDistance / minDistance are MapPoint::mfMaxDistance / mfMinDistance as stored.  The range gate uses
  // GetMaxDistanceInvariance() = 1.2f * mfMaxDistance and GetMinDistanceInvariance() = 0.8f * mfMinDistance
  // (MapPoint.cc:555-565, Frame.cc:609-615); PredictScale divides the RAW mfMaxDistance (MapPoint.cc:587-602).
  const float maxDist = __fmul_rn(1.2f, maxDistance), minDist = __fmul_rn(0.8f, minDistance);
  FrustumOut o;
  o.inView = 0; o.projX = -1.f; o.projY = -1.f; o.projXR = 0.f; o.depth = 0.f; o.viewCos = 0.f; o.level = -1;
  float Pc[3];
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const float t = __fadd_rn(__fadd_rn(__fmul_rn(C.Rcw[3 * r], p0), __fmul_rn(C.Rcw[3 * r + 1], p1)), __fmul_rn(C.Rcw[3 * r + 2], p2));
    Pc[r] = __fadd_rn(t, C.tcw[r]);
  }
  const float PcDist = (float)sqrt((double)Pc[0] * Pc[0] + (double)Pc[1] * Pc[1] + (double)Pc[2] * Pc[2]);
  const float invz = __fdiv_rn(1.0f, Pc[2]);
  if (Pc[2] < 0.0f) return o;
  float ux, uy;
  if (C.camera_model == ORBX_CAMERA_KB8) {
    // KannalaBrandt8::project (KannalaBrandt8.cpp:29-45).  atan2f is evaluated as (float)atan2((double)..) (portable policy,
    // DESIGN.md section 2); cos / sin resolve to the double overloads in the reference (no using-directive in that file), so
    // the last line is double arithmetic rounded once to float, as written here.
    const float x2y2 = __fadd_rn(__fmul_rn(Pc[0], Pc[0]), __fmul_rn(Pc[1], Pc[1]));
    const float theta = (float)atan2((double)sqrtf(x2y2), (double)Pc[2]);
    const float psi = (float)atan2((double)Pc[1], (double)Pc[0]);
    const float t2 = __fmul_rn(theta, theta), t3 = __fmul_rn(theta, t2), t5 = __fmul_rn(t3, t2), t7 = __fmul_rn(t5, t2), t9 = __fmul_rn(t7, t2);
    const float r = __fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(theta, __fmul_rn(C.kb[0], t3)), __fmul_rn(C.kb[1], t5)), __fmul_rn(C.kb[2], t7)),
                              __fmul_rn(C.kb[3], t9));
    ux = (float)__dadd_rn(__dmul_rn((double)__fmul_rn(C.fx, r), cos((double)psi)), (double)C.cx);
    uy = (float)__dadd_rn(__dmul_rn((double)__fmul_rn(C.fy, r), sin((double)psi)), (double)C.cy);
  } else {
    ux = __fadd_rn(__fdiv_rn(__fmul_rn(C.fx, Pc[0]), Pc[2]), C.cx);
    uy = __fadd_rn(__fdiv_rn(__fmul_rn(C.fy, Pc[1]), Pc[2]), C.cy);
  }
  if (ux < C.min_x || ux > C.max_x) return o;
  if (uy < C.min_y || uy > C.max_y) return o;
  o.projX = ux; o.projY = uy;
  const float o0 = __fsub_rn(p0, C.Ow[0]), o1 = __fsub_rn(p1, C.Ow[1]), o2 = __fsub_rn(p2, C.Ow[2]);
  const float dist = (float)sqrt((double)o0 * o0 + (double)o1 * o1 + (double)o2 * o2);
  if (dist < minDist || dist > maxDist) return o;
  const double dot = (double)o0 * n0 + (double)o1 * n1 + (double)o2 * n2;
  const float viewCos = (float)(dot / (double)dist);
  if (viewCos < C.viewing_cos_limit) return o;
  const float ratio = __fdiv_rn(maxDistance, dist);
  const float lg = (float)log((double)ratio);
  int nScale = (int)ceilf(__fdiv_rn(lg, C.log_scale_factor));
  if (nScale < 0) nScale = 0;
  else if (nScale >= C.n_scale_levels) nScale = C.n_scale_levels - 1;
  o.inView = 1; o.projXR = __fsub_rn(ux, __fmul_rn(C.bf, invz)); o.depth = PcDist; o.level = nScale; o.viewCos = viewCos;
  return o;
}

}  // namespace orbx
