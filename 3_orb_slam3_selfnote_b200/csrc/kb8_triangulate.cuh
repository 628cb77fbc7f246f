// KannalaBrandt8::epipolarConstrain on the device: the gate of ORBmatcher::SearchForTriangulation on fisheye keyframes
// (src/ORBmatcher.cc:1146 -> src/CameraModels/KannalaBrandt8.cpp:244-247 -> TriangulateMatches :343-418 -> Triangulate :430-444).
//
// Every cv::Mat expression of those lines is written out with the rounding OpenCV gives it (the rules tests/ hold bit for bit
// against the reference's own compiled file and against cv2): a small cv::gemm is float products summed
// left to right, Mat::dot / cv::norm accumulate in double, `s * row - row` is one rounding (addWeighted), `Mat / s` multiplies by
// (float)(1.0 / s), and cv::SVD::compute of a 4x4 is OpenCV's one-sided Jacobi (modules/core/src/lapack.cpp JacobiSVDImpl_<float>:
// matrices this small never reach LAPACK).  libm calls follow the portable policy of DESIGN.md section 2: tanf / atan2f are the double
// function rounded to float; cos / sin are the double overloads the reference's translation unit binds.  The file is compiled with
// --fmad=false (build.py), so a * b + c below is two roundings, as on the host.
#pragma once

namespace orbx {

// One (camera of KF1, camera of KF2) combination: R12 / t12 as ORBmatcher.cc:1000-1019 forms them, cam = fx fy cx cy k1..k4.
// Layout = orbx.h's orbx_kb8_pair.
struct Kb8PairDev { float R12[9]; float t12[3]; float cam1[8]; float cam2[8]; };

// KannalaBrandt8::unproject (KannalaBrandt8.cpp:112-137): Newton on theta + k1 theta^3 + .. = theta_d, ray = (x, y, 1).
__device__ __forceinline__ void kb8_unproject_dev(const float* cam, float u, float v, float* ray) {
  const float pwx = (u - cam[2]) / cam[0], pwy = (v - cam[3]) / cam[1];
  float scale = 1.f;
  float theta_d = sqrtf(pwx * pwx + pwy * pwy);
  theta_d = fminf(fmaxf(-1.57079637f, theta_d), 1.57079637f);   // (float)(-+CV_PI / 2.f)
  if ((double)theta_d > 1e-8) {
    float theta = theta_d;
    for (int j = 0; j < 10; j++) {
      const float theta2 = theta * theta, theta4 = theta2 * theta2, theta6 = theta4 * theta2, theta8 = theta4 * theta4;
      const float k0_theta2 = cam[4] * theta2, k1_theta4 = cam[5] * theta4;
      const float k2_theta6 = cam[6] * theta6, k3_theta8 = cam[7] * theta8;
      const float theta_fix = (theta * (1.f + k0_theta2 + k1_theta4 + k2_theta6 + k3_theta8) - theta_d) /
                              (1.f + 3.f * k0_theta2 + 5.f * k1_theta4 + 7.f * k2_theta6 + 9.f * k3_theta8);
      theta = theta - theta_fix;
      if (fabsf(theta_fix) < 1e-6f) break;
    }
    scale = (float)tan((double)theta) / theta_d;
  }
  ray[0] = pwx * scale; ray[1] = pwy * scale; ray[2] = 1.f;
}

// KannalaBrandt8::project (KannalaBrandt8.cpp:29-45), the arithmetic of frustum.cuh / k_project_points.
__device__ __forceinline__ void kb8_project_dev(const float* cam, const float* X, float* u, float* v) {
  const float x2y2 = X[0] * X[0] + X[1] * X[1];
  const float theta = (float)atan2((double)sqrtf(x2y2), (double)X[2]);
  const float psi = (float)atan2((double)X[1], (double)X[0]);
  const float t2 = theta * theta, t3 = theta * t2, t5 = t3 * t2, t7 = t5 * t2, t9 = t7 * t2;
  const float r = theta + cam[4] * t3 + cam[5] * t5 + cam[6] * t7 + cam[7] * t9;
  *u = (float)((double)(cam[0] * r) * cos((double)psi) + (double)cam[2]);
  *v = (float)((double)(cam[1] * r) * sin((double)psi) + (double)cam[3]);
}

// One Hestenes rotation of JacobiSVDImpl_<float> on rows I, J of At (and Vt); returns whether the pair was rotated.
template <int I, int J>
__device__ __forceinline__ bool jacobi_pair(float (&At)[4][4], float (&Vt)[4][4], double (&W)[4]) {
  double a = W[I], p = 0, b = W[J];
#pragma unroll
  for (int k = 0; k < 4; k++) p += (double)At[I][k] * (double)At[J][k];
  if (fabs(p) <= (double)(2 * 1.1920929e-07f) * sqrt(a * b)) return false;
  p *= 2;
  const double beta = a - b, gamma = hypot(p, beta);
  float c, s;
  if (beta < 0) {
    const double delta = (gamma - beta) * 0.5;
    s = (float)sqrt(delta / gamma);
    c = (float)(p / (gamma * (double)s * 2));
  } else {
    c = (float)sqrt((gamma + beta) / (gamma * 2));
    s = (float)(p / (gamma * (double)c * 2));
  }
  a = b = 0;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const float t0 = c * At[I][k] + s * At[J][k];
    const float t1 = -s * At[I][k] + c * At[J][k];
    At[I][k] = t0; At[J][k] = t1;
    a += (double)t0 * (double)t0; b += (double)t1 * (double)t1;
  }
  W[I] = a; W[J] = b;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const float t0 = c * Vt[I][k] + s * Vt[J][k];
    const float t1 = -s * Vt[I][k] + c * Vt[J][k];
    Vt[I][k] = t0; Vt[J][k] = t1;
  }
  return true;
}

// cv::SVD::compute(A, w, u, vt, MODIFY_A | FULL_UV) of a 4x4, reduced to what Triangulate reads: the row of vt that belongs to the
// singular value that the descending selection sort of JacobiSVDImpl_ leaves last (the sort swaps on a strictly larger later value
// only; it is replayed literally below on the row indices, so ties resolve as they do on the host).
__device__ __forceinline__ void jacobi_svd4_last_row(const float* A, float* v) {
  float At[4][4], Vt[4][4];
  double W[4];
#pragma unroll
  for (int i = 0; i < 4; i++) {
    double sd = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const float t = A[k * 4 + i];
      At[i][k] = t; Vt[i][k] = i == k ? 1.f : 0.f;
      sd += (double)t * (double)t;
    }
    W[i] = sd;
  }
  for (int iter = 0; iter < 30; iter++) {
    bool changed = false;
    changed |= jacobi_pair<0, 1>(At, Vt, W);
    changed |= jacobi_pair<0, 2>(At, Vt, W);
    changed |= jacobi_pair<0, 3>(At, Vt, W);
    changed |= jacobi_pair<1, 2>(At, Vt, W);
    changed |= jacobi_pair<1, 3>(At, Vt, W);
    changed |= jacobi_pair<2, 3>(At, Vt, W);
    if (!changed) break;
  }
#pragma unroll
  for (int i = 0; i < 4; i++) {
    double sd = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) sd += (double)At[i][k] * (double)At[i][k];
    W[i] = sqrt(sd);
  }
  // selection sort, descending, swapping on a strictly larger later value only: track which original row lands at position 3
  int pos[4] = {0, 1, 2, 3};
#pragma unroll
  for (int i = 0; i < 3; i++) {
    int j = i;
#pragma unroll
    for (int k = i + 1; k < 4; k++)
      if (W[j] < W[k]) j = k;
    if (i != j) {
      const double tw = W[i]; W[i] = W[j]; W[j] = tw;
      const int tp = pos[i]; pos[i] = pos[j]; pos[j] = tp;
    }
  }
  const int r = pos[3];
#pragma unroll
  for (int k = 0; k < 4; k++) v[k] = r == 0 ? Vt[0][k] : (r == 1 ? Vt[1][k] : (r == 2 ? Vt[2][k] : Vt[3][k]));
}

// KannalaBrandt8::TriangulateMatches with r1 = unproject(kp1) supplied (it does not depend on the candidate): z1 or -1.
__device__ __forceinline__ float kb8_triangulate_matches_dev(const Kb8PairDev& G, const float* r1, float x1, float y1, float x2, float y2,
                                                             float sigmaLevel, float unc, float* p3D = nullptr) {
  float r2[3], r21[3];
  kb8_unproject_dev(G.cam2, x2, y2, r2);
  const float* R = G.R12;
#pragma unroll
  for (int i = 0; i < 3; i++) r21[i] = R[3 * i] * r2[0] + R[3 * i + 1] * r2[1] + R[3 * i + 2] * r2[2];
  double dot = 0, n1 = 0, n2 = 0;
#pragma unroll
  for (int i = 0; i < 3; i++) {
    dot += (double)r1[i] * (double)r21[i]; n1 += (double)r1[i] * (double)r1[i]; n2 += (double)r21[i] * (double)r21[i];
  }
  const float cosParallaxRays = (float)(dot / (sqrt(n1) * sqrt(n2)));
  if ((double)cosParallaxRays > 0.9998) return -1.f;
  float T2[3][4];
#pragma unroll
  for (int i = 0; i < 3; i++) {
#pragma unroll
    for (int k = 0; k < 3; k++) T2[i][k] = R[3 * k + i];
    const float t = R[i] * G.t12[0] + R[3 + i] * G.t12[1] + R[6 + i] * G.t12[2];
    T2[i][3] = (float)((double)t * -1.0);
  }
  float A[16];
  const float ps[4] = {r1[0], r1[1], r2[0], r2[1]};
#pragma unroll
  for (int i = 0; i < 4; i++) {
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const float row2 = i < 2 ? (k == 2 ? 1.f : 0.f) : T2[2][k];
      const float row01 = i < 2 ? (k == (i & 1) ? 1.f : 0.f) : T2[i & 1][k];
      A[4 * i + k] = (float)((double)row2 * (double)ps[i] + (double)row01 * -1.0);
    }
  }
  float vt3[4];
  jacobi_svd4_last_row(A, vt3);
  const float inv = (float)(1.0 / (double)vt3[3]);
  const float x3D[3] = {vt3[0] * inv, vt3[1] * inv, vt3[2] * inv};
  const float z1 = x3D[2];
  if (!(z1 > 0)) return -1.f;   // `if (z1 <= 0) return -1` -- a NaN depth passes that test in the reference and fails `> 0.0001f` later
  const float z2 = (float)(((double)T2[2][0] * (double)x3D[0] + (double)T2[2][1] * (double)x3D[1] + (double)T2[2][2] * (double)x3D[2]) +
                           (double)T2[2][3]);
  if (z2 <= 0) return -1.f;
  float u, v;
  kb8_project_dev(G.cam1, x3D, &u, &v);
  const float errX1 = u - x1, errY1 = v - y1;
  if ((double)(errX1 * errX1 + errY1 * errY1) > 5.991 * (double)sigmaLevel) return -1.f;
  float x3D2[3];
#pragma unroll
  for (int i = 0; i < 3; i++) {
    const float t = T2[i][0] * x3D[0] + T2[i][1] * x3D[1] + T2[i][2] * x3D[2];
    x3D2[i] = (float)((double)t + (double)T2[i][3]);
  }
  kb8_project_dev(G.cam2, x3D2, &u, &v);
  const float errX2 = u - x2, errY2 = v - y2;
  if ((double)(errX2 * errX2 + errY2 * errY2) > 5.991 * (double)unc) return -1.f;
  if (p3D) { p3D[0] = x3D[0]; p3D[1] = x3D[1]; p3D[2] = x3D[2]; }
  return z1;
}

// n independent keypoint pairs through TriangulateMatches: what Frame::ComputeStereoFishEyeMatches does with every left-right match
// that survived the ratio test (src/Frame.cc:1255-1262, pair = (mRlr, mtlr, left camera, right camera)).  Thread = pair.
__global__ void __launch_bounds__(128) k_kb8_triangulate_pairs(const __grid_constant__ Kb8PairDev G, int n, const float* __restrict__ xy1,
                                                               const float* __restrict__ xy2, const float* __restrict__ sigma1,
                                                               const float* __restrict__ sigma2, float* __restrict__ z1,
                                                               float* __restrict__ p3D) {
  const int i = blockIdx.x * 128 + threadIdx.x;
  if (i >= n) return;
  float r1[3], X[3] = {0.f, 0.f, 0.f};
  kb8_unproject_dev(G.cam1, xy1[2 * i], xy1[2 * i + 1], r1);
  z1[i] = kb8_triangulate_matches_dev(G, r1, xy1[2 * i], xy1[2 * i + 1], xy2[2 * i], xy2[2 * i + 1], sigma1[i], sigma2[i], X);
  p3D[3 * i] = X[0]; p3D[3 * i + 1] = X[1]; p3D[3 * i + 2] = X[2];
}

struct Kb8TriArgs {
  Kb8PairDev pair[4];          // ll, lr, rl, rr (index 2 * bRight1 + bRight2); pair[0] alone without second cameras
  int nLeft1, nLeft2;          // -1: one camera per keyframe
  const float* levelSigma1;    // pKF1->mvLevelSigma2 (sigmaLevel of the gate); BowArgs::levelSigma2 is pKF2's (unc)
};

// SearchForTriangulation (ORBmatcher.cc:1046-1175) with the KannalaBrandt8 gate: same shape as k_bow_triangulate -- warp = one side-1
// feature of a common vocabulary node, lanes = the node's side-2 features, last minimum within TH_LOW among the gated candidates.
// The Hamming test runs first, so only the few candidates within TH_LOW pay for a triangulation.
__global__ void __launch_bounds__(256) k_bow_triangulate_kb8(const __grid_constant__ BowArgs A, const __grid_constant__ Kb8TriArgs T) {
  const int lane = threadIdx.x & 31;
  const int w = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (w >= A.nwork) return;
  const BowWork W = A.work[w];
  const int idx1 = (int)A.s1.indices[W.p1];
  if ((unsigned)idx1 >= (unsigned)A.s1.n || !A.s1.elig[idx1]) return;
  const bool two = T.nLeft1 != -1;
  const bool bStereo1 = !two && A.s1.stereo && A.s1.stereo[idx1];
  const bool bRight1 = two && idx1 >= T.nLeft1;
  const orbx_keypoint kp1 = A.s1.kps[idx1];
  uint32_t d1[8];
  load_desc(A.s1.desc + 32ll * idx1, d1);
  float r1[3];
  kb8_unproject_dev(T.pair[bRight1 ? 2 : 0].cam1, kp1.x, kp1.y, r1);
  const float sigmaLevel = T.levelSigma1[kp1.octave];
  unsigned long long best = kKeyNone;
  for (int p = W.b0 + lane; p < W.b1; p += 32) {
    const int idx2 = (int)A.s2.indices[p];
    if ((unsigned)idx2 >= (unsigned)A.s2.n || !A.s2.elig[idx2]) continue;
    uint32_t d2[8];
    load_desc(A.s2.desc + 32ll * idx2, d2);
    const int dist = hamming8(d1, d2);
    if (dist > 50) continue;  // TH_LOW
    const orbx_keypoint kp2 = A.s2.kps[idx2];
    const bool bStereo2 = T.nLeft2 == -1 && A.s2.stereo && A.s2.stereo[idx2];
    const bool bRight2 = T.nLeft2 != -1 && idx2 >= T.nLeft2;
    if (!bStereo1 && !bStereo2 && !two) {
      const float distex = A.epx - kp2.x, distey = A.epy - kp2.y;
      if (distex * distex + distey * distey < 100.0f * A.scaleFactors2[kp2.octave]) continue;
    }
    bool ok = A.coarse != 0;
    if (!ok) {
      const Kb8PairDev& G = T.pair[two ? 2 * (int)bRight1 + (int)bRight2 : 0];
      ok = kb8_triangulate_matches_dev(G, r1, kp1.x, kp1.y, kp2.x, kp2.y, sigmaLevel, A.levelSigma2[kp2.octave]) > 0.0001f;
    }
    if (!ok) continue;
    best = min(best, bow_key(dist, 0xfffff - (p - W.b0), idx2));   // later list position wins ties
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
  if (lane == 0 && best != kKeyNone) A.matches12[idx1] = (int)(best & 0xfffffu);
}

}  // namespace orbx
