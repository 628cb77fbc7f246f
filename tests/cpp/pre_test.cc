// C++ exercise of host/FramePreproc.h through the reference-side call shapes:
//   pre_test <bgr.raw> <mapx.f32> <mapy.f32> <W> <H> <out.bin>
// colour -> gray (Tracking.cc:998), remap (stereo_euroc.cc:166), CLAHE (mono_tum_vi.cc:109), UndistortKeyPoints (Frame.cc:837),
// isInFrustum over a small point set (Frame.cc:572).  Outputs are dumped for tests/test_cpp_pre_gpu.py to compare with the oracle.
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../../3_orb_slam3_selfnote_b200/host/FramePreproc.h"

template <class T> static std::vector<T> slurp(const char* fn, size_t n) {
  std::vector<T> v(n);
  FILE* f = fopen(fn, "rb");
  if (!f || fread(v.data(), sizeof(T), n, f) != n) { fprintf(stderr, "cannot read %s\n", fn); exit(2); }
  fclose(f);
  return v;
}
template <class T> static void dump(FILE* f, const T* p, size_t n) { fwrite(p, sizeof(T), n, f); }

int main(int argc, char** argv) {
  if (argc < 7) return 1;
  const int W = atoi(argv[4]), H = atoi(argv[5]);
  std::vector<uint8_t> bgr = slurp<uint8_t>(argv[1], (size_t)W * H * 3);
  std::vector<float> mx = slurp<float>(argv[2], (size_t)W * H), my = slurp<float>(argv[3], (size_t)W * H);
  ORB_SLAM3::FramePreproc pre;
  cvx::Mat gray, rect;
  pre.cvtColorToGray(bgr.data(), W, H, (size_t)W * 3, 3, /*bRGB=*/false, gray);
  const int id = pre.addRectifyMap(mx.data(), my.data(), W, H, W, W, H);
  pre.remap(gray, rect, id);
  cvx::Mat eq(H, W, ORBX_MAT_8U);
  for (int y = 0; y < H; ++y) memcpy(eq.ptr(y), rect.ptr(y), W);
  pre.clahe(eq);
  // keypoints on a lattice, EuRoC cam0 intrinsics / distortion (Examples/Monocular/EuRoC.yaml)
  std::vector<cvx::KeyPoint> keys, keysUn;
  for (int y = 5; y < H; y += 37)
    for (int x = 3; x < W; x += 41) { cvx::KeyPoint k; k.pt.x = x + 0.25f; k.pt.y = y + 0.5f; k.size = 31; k.octave = (x + y) % 8; keys.push_back(k); }
  const float K[4] = {458.654f, 457.296f, 367.215f, 248.375f}, D[4] = {-0.28340811f, 0.07395907f, 0.00019359f, 1.76187114e-05f};
  pre.UndistortKeyPoints(keys, K, D, 4, keysUn);
  // frustum: identity pose, points on a grid in front of / behind the camera
  orbx_frustum fr = {};
  fr.Rcw[0] = fr.Rcw[4] = fr.Rcw[8] = 1.f;
  fr.fx = K[0]; fr.fy = K[1]; fr.cx = K[2]; fr.cy = K[3];
  fr.min_x = 0; fr.max_x = (float)W; fr.min_y = 0; fr.max_y = (float)H;
  fr.bf = 47.9f; fr.log_scale_factor = 0.18232156f; fr.viewing_cos_limit = 0.5f; fr.n_scale_levels = 8;
  std::vector<float> P, N, maxD, minD;
  for (int i = 0; i < 500; ++i) {
    const float z = -1.f + 0.03f * i, x = -2.f + 0.013f * (i * 7 % 311), y = -1.5f + 0.011f * (i * 13 % 277);
    P.push_back(x); P.push_back(y); P.push_back(z);
    N.push_back(0.1f * (i % 5)); N.push_back(-0.1f * (i % 3)); N.push_back(0.9f);
    maxD.push_back(2.f + 0.05f * (i % 200)); minD.push_back(0.2f + 0.01f * (i % 50));
  }
  ORB_SLAM3::FramePreproc::FrustumResult r;
  const int visible = pre.isInFrustum(fr, P, N, maxD, minD, r);

  FILE* f = fopen(argv[6], "wb");
  for (int y = 0; y < H; ++y) dump(f, gray.ptr(y), W);
  for (int y = 0; y < H; ++y) dump(f, rect.ptr(y), W);
  for (int y = 0; y < H; ++y) dump(f, eq.ptr(y), W);
  const int nk = (int)keys.size();
  dump(f, &nk, 1); dump(f, keys.data(), nk); dump(f, keysUn.data(), nk);
  const int np = (int)maxD.size();
  dump(f, &np, 1); dump(f, &visible, 1);
  dump(f, P.data(), P.size()); dump(f, N.data(), N.size()); dump(f, maxD.data(), np); dump(f, minD.data(), np);
  dump(f, r.mbTrackInView.data(), np); dump(f, r.mTrackProjX.data(), np); dump(f, r.mTrackProjY.data(), np); dump(f, r.mTrackProjXR.data(), np);
  dump(f, r.mTrackDepth.data(), np); dump(f, r.mnTrackScaleLevel.data(), np); dump(f, r.mTrackViewCos.data(), np);
  fclose(f);
  return 0;
}
