// CPU ORACLE (test infrastructure, not product code) -- SURVEY.md §8(f) rank 2: the Frame glue on either side of the path.
//
//   Frame::UndistortKeyPoints / ComputeImageBounds     src/Frame.cc:837-870, 872-899  -> cv::undistortPoints(mat, mat, K, D, Mat(), K)
//   Frame::isInFrustum (monocular / rectified branch)  src/Frame.cc:572-642
//   MapPoint::PredictScale                             src/MapPoint.cc:587-602
//   Pinhole::project                                   src/CameraModels/Pinhole.cpp:46-49
//
// cv::undistortPoints is OpenCV (external, un-vendored): restated from the published algorithm (cvUndistortPointsInternal:
// 5 fixed-point iterations in double, criteria COUNT only) and pinned against python cv2 in tests/test_oracle_frame.py.
// isInFrustum's arithmetic is the reference's cv::Mat float code: OpenCV's small-matrix gemm evaluates each row of Rcw * P
// as ((a0*b0 + a1*b1) + a2*b2) in float, cv::norm and Mat::dot accumulate in double.  Compiled with -ffp-contract=off.
#include <cmath>
#include <cstdint>

extern "C" {

// src/dst: n x 2 float (x, y).  K = fx, fy, cx, cy (float, widened to double as OpenCV does); D = k1 k2 p1 p2 k3.
// R = identity and P = K as in Frame.cc:856, so the final projection is ((fx*x + 0*y) + cx) * (1 / 1).
void oro_undistort_points(const float* src, int n, float fx_, float fy_, float cx_, float cy_, const float* D, float* dst) {
  const double fx = fx_, fy = fy_, cx = cx_, cy = cy_, ifx = 1. / fx, ify = 1. / fy;
  const double k1 = D[0], k2 = D[1], p1 = D[2], p2 = D[3], k3 = D[4];
  for (int i = 0; i < n; ++i) {
    double x = src[2 * i], y = src[2 * i + 1];
    const double u = x, v = y;
    x = (x - cx) * ifx; y = (y - cy) * ify;
    const double x0 = x, y0 = y;
    for (int j = 0; j < 5; ++j) {
      const double r2 = x * x + y * y;
      const double icdist = (1 + ((0 * r2 + 0) * r2 + 0) * r2) / (1 + ((k3 * r2 + k2) * r2 + k1) * r2);
      if (icdist < 0) { x = (u - cx) * ifx; y = (v - cy) * ify; break; }
      const double deltaX = 2 * p1 * x * y + p2 * (r2 + 2 * x * x) + 0 * r2 + 0 * r2 * r2;
      const double deltaY = p1 * (r2 + 2 * y * y) + 2 * p2 * x * y + 0 * r2 + 0 * r2 * r2;
      x = (x0 - deltaX) * icdist; y = (y0 - deltaY) * icdist;
    }
    const double xx = fx * x + 0 * y + cx, yy = 0 * x + fy * y + cy, ww = 1. / (0 * x + 0 * y + 1);
    dst[2 * i] = (float)(xx * ww); dst[2 * i + 1] = (float)(yy * ww);
  }
}

struct FrustumCam {
  float Rcw[9], tcw[3], Ow[3];
  float fx, fy, cx, cy;
  float minX, maxX, minY, maxY;
  float bf, logScaleFactor, viewingCosLimit;
  int nScaleLevels;
  int logMode;  // 0: logf / atan2f (reference-literal, host libm); 1: (float)log((double)x), (float)atan2((double)..) (portable, = CUDA path)
  int cameraModel;  // 0 Pinhole::project, 1 KannalaBrandt8::project (KannalaBrandt8.cpp:29-45)
  float kb[4];      // KB8 k1..k4
};

// Frame::isInFrustum, Nleft == -1 branch (Frame.cc:574-642).  Outputs per point: inView, mTrackProjX/Y/XR, mTrackDepth,
// mnTrackScaleLevel, mTrackViewCos.  Points that fail keep the values the reference leaves behind: projX/projY = -1 unless the
// image-bounds test passed (Frame.cc:577-578, 604-605); the other fields are written as 0 / -1 (the reference leaves them stale).
void oro_is_in_frustum(const FrustumCam* C, int n, const float* P /*n x 3*/, const float* normal /*n x 3*/, const float* maxDist /* mfMaxDistance */,
                       const float* minDist /* mfMinDistance */, uint8_t* inView, float* projX, float* projY, float* projXR, float* depth, int* level,
                       float* viewCosOut) {
  for (int i = 0; i < n; ++i) {
    inView[i] = 0; projX[i] = -1; projY[i] = -1; projXR[i] = 0; depth[i] = 0; level[i] = -1; viewCosOut[i] = 0;
    const float* p = P + 3 * i;
    float Pc[3];
    for (int r = 0; r < 3; ++r) {
      const float t = C->Rcw[3 * r] * p[0] + C->Rcw[3 * r + 1] * p[1] + C->Rcw[3 * r + 2] * p[2];
      Pc[r] = t + C->tcw[r];
    }
    const float PcDist = (float)std::sqrt((double)Pc[0] * Pc[0] + (double)Pc[1] * Pc[1] + (double)Pc[2] * Pc[2]);
    const float PcZ = Pc[2];
    const float invz = 1.0f / PcZ;
    if (PcZ < 0.0f) continue;
    float ux, uy;
    if (C->cameraModel == 1) {
      const float x2y2 = Pc[0] * Pc[0] + Pc[1] * Pc[1];
      const float theta = C->logMode == 0 ? atan2f(sqrtf(x2y2), Pc[2]) : (float)std::atan2((double)sqrtf(x2y2), (double)Pc[2]);
      const float psi = C->logMode == 0 ? atan2f(Pc[1], Pc[0]) : (float)std::atan2((double)Pc[1], (double)Pc[0]);
      const float theta2 = theta * theta, theta3 = theta * theta2, theta5 = theta3 * theta2, theta7 = theta5 * theta2, theta9 = theta7 * theta2;
      const float r = theta + C->kb[0] * theta3 + C->kb[1] * theta5 + C->kb[2] * theta7 + C->kb[3] * theta9;
      // cos(psi) / sin(psi): the double overloads (KannalaBrandt8.cpp has no using-directive) -> double product and sum
      ux = (float)((double)(C->fx * r) * std::cos((double)psi) + (double)C->cx);
      uy = (float)((double)(C->fy * r) * std::sin((double)psi) + (double)C->cy);
    } else {
      ux = C->fx * Pc[0] / Pc[2] + C->cx; uy = C->fy * Pc[1] / Pc[2] + C->cy;
    }
    if (ux < C->minX || ux > C->maxX) continue;
    if (uy < C->minY || uy > C->maxY) continue;
    projX[i] = ux; projY[i] = uy;
    const float PO[3] = {p[0] - C->Ow[0], p[1] - C->Ow[1], p[2] - C->Ow[2]};
    const float dist = (float)std::sqrt((double)PO[0] * PO[0] + (double)PO[1] * PO[1] + (double)PO[2] * PO[2]);
    // GetMinDistanceInvariance() = 0.8f * mfMinDistance, GetMaxDistanceInvariance() = 1.2f * mfMaxDistance (MapPoint.cc:555-565)
    if (dist < 0.8f * minDist[i] || dist > 1.2f * maxDist[i]) continue;
    const float* Pn = normal + 3 * i;
    const double dot = (double)PO[0] * Pn[0] + (double)PO[1] * Pn[1] + (double)PO[2] * Pn[2];
    const float viewCos = (float)(dot / dist);
    if (viewCos < C->viewingCosLimit) continue;
    const float ratio = maxDist[i] / dist;   // PredictScale: the RAW mfMaxDistance (MapPoint.cc:587-602)
    const float lg = C->logMode == 0 ? logf(ratio) : (float)std::log((double)ratio);
    int nScale = (int)std::ceil(lg / C->logScaleFactor);
    if (nScale < 0) nScale = 0;
    else if (nScale >= C->nScaleLevels) nScale = C->nScaleLevels - 1;
    inView[i] = 1; projXR[i] = ux - C->bf * invz; depth[i] = PcDist; level[i] = nScale; viewCosOut[i] = viewCos;
  }
}

}  // extern "C"
