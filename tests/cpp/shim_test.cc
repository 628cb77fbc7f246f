// C++ exercise of the reference-signature shims (host/ORBextractor.h, host/ORBmatcher.h) with small mock Frame / MapPoint
// types that carry the member names the reference's classes have.  Driven by tests/test_cpp_shim_gpu.py:
//   shim_test left.raw right.raw W H out.bin
// writes: [n, mono, kps(28n), desc(32n)] for the left image with lapping {0,1000}; then stereo (lapping {0,0}) N, mvuRight, mvDepth;
// then SearchByProjection of the left frame against map points made of the right frame's keypoints: nmatches, kp->mp index.
#include <cstdio>
#include <cstdlib>
#include <map>
#include <vector>

#include "../../3_orb_slam3_selfnote_b200/host/ORBmatcher.h"

using namespace ORB_SLAM3;

struct MockMapPoint {
  bool mbTrackInView = true, mbTrackInViewR = false;
  float mTrackDepth = 1.f, mTrackProjX = 0, mTrackProjY = 0, mTrackProjXR = 0, mTrackProjYR = 0, mTrackViewCos = 1.f, mTrackViewCosR = 1.f;
  int mnTrackScaleLevel = 0, mnTrackScaleLevelR = -1;
  cvx::Mat desc;
  int id = 0;
  bool isBad() const { return false; }
  int Observations() const { return 1; }
  cvx::Mat GetDescriptor() const { return desc; }
};

struct MockFrame {
  int N = 0, Nleft = -1;
  std::vector<cvx::KeyPoint> mvKeys, mvKeysUn, mvKeysRight;
  std::vector<int> mvLeftToRightMatch, mvRightToLeftMatch;
  cvx::Mat mDescriptors;
  std::vector<float> mvuRight, mvDepth, mvScaleFactors;
  float mnMinX = 0, mnMinY = 0, mnMaxX = 0, mnMaxY = 0, mb = 0, mbf = 0;
  std::vector<MockMapPoint*> mvpMapPoints;
};

// the members of KeyFrame the bag-of-words searches read (include/KeyFrame.h)
struct MockKeyFrame {
  int N = 0, NLeft = -1;
  void* mpCamera2 = nullptr;
  std::vector<cvx::KeyPoint> mvKeys, mvKeysUn, mvKeysRight;
  cvx::Mat mDescriptors;
  std::vector<float> mvuRight, mvScaleFactors, mvLevelSigma2, mvInvLevelSigma2;
  float mnMinX = 0, mnMinY = 0, mnMaxX = 0, mnMaxY = 0;
  std::map<unsigned, std::vector<unsigned> > mFeatVec;   // DBoW2::FeatureVector
  std::vector<MockMapPoint*> mps;
  std::vector<MockMapPoint*> GetMapPointMatches() const { return mps; }
  MockMapPoint* GetMapPoint(size_t i) const { return mps[i]; }
};
// test vocabulary: node = low 5 bits of the first descriptor byte (the Python side builds the same lists)
static void fill_featvec(std::map<unsigned, std::vector<unsigned> >& fv, const cvx::Mat& desc, int n) {
  for (int i = 0; i < n; i++) fv[(unsigned)(desc.ptr(i)[0] & 31)].push_back((unsigned)i);
}

static std::vector<uint8_t> read_raw(const char* fn, size_t n) {
  std::vector<uint8_t> v(n);
  FILE* f = fopen(fn, "rb");
  if (!f || fread(v.data(), 1, n, f) != n) { fprintf(stderr, "cannot read %s\n", fn); exit(2); }
  fclose(f);
  return v;
}

int main(int argc, char** argv) {
  if (argc < 6) return 2;
  const int W = atoi(argv[3]), H = atoi(argv[4]);
  std::vector<uint8_t> l = read_raw(argv[1], (size_t)W * H), r = read_raw(argv[2], (size_t)W * H);
  cvx::Mat imL(H, W, ORBX_MAT_8U, l.data(), W), imR(H, W, ORBX_MAT_8U, r.data(), W), mask;
  FILE* out = fopen(argv[5], "wb");
  try {
    // --- mono extraction, Frame::ExtractORB(0, im, 0, 1000) (src/Frame.cc:339)
    ORBextractor mono(1000, 1.2f, 8, 20, 7);
    std::vector<cvx::KeyPoint> keys;
    cvx::Mat desc;
    std::vector<int> lap = {0, 1000};
    int ret = mono(imL, mask, keys, desc, lap);
    int n = (int)keys.size();
    fwrite(&n, 4, 1, out); fwrite(&ret, 4, 1, out);
    fwrite(keys.data(), 28, n, out);
    for (int i = 0; i < n; i++) fwrite(desc.ptr(i), 1, 32, out);
    if (mono(cvx::Mat(), mask, keys, desc, lap) != -1) { fprintf(stderr, "empty image must return -1\n"); return 3; }
    // the public member is live: reading a level right after operator() fetches it (src/Frame.cc:908,998,1012 do exactly this)
    if (mono.mvImagePyramid.size() != 8 || mono.mvImagePyramid[1].cols != 627 || mono.mvImagePyramid[0].rows != H) { fprintf(stderr, "pyramid\n"); return 3; }
    { const cvx::Mat& l0 = mono.mvImagePyramid[0];
      for (int y = 0; y < H; y += 37) if (memcmp(l0.ptr(y), l.data() + (size_t)y * W, W) != 0) { fprintf(stderr, "pyramid level 0 pixels\n"); return 3; } }
    if (mono.GetLevels() != 8 || mono.GetScaleFactors().size() != 8) return 3;

    // --- stereo frame (src/Frame.cc:120-141)
    ORBextractor exL(1200, 1.2f, 8, 20, 7), exR(1200, 1.2f, 8, 20, 7);
    MockFrame F, FR;
    cvx::Mat descR;
    std::vector<int> lap0 = {0, 0};
    exL(imL, mask, F.mvKeys, F.mDescriptors, lap0);
    exR(imR, mask, FR.mvKeys, descR, lap0);
    F.N = (int)F.mvKeys.size(); F.mvKeysUn = F.mvKeys;
    F.mvScaleFactors = exL.GetScaleFactors();
    F.mnMaxX = (float)W; F.mnMaxY = (float)H;
    F.mbf = 47.90639384423901f; F.mb = F.mbf / 435.2046959714599f;
    ORBmatcher matcher(0.8f, true);
    matcher.ComputeStereoMatches(F, exL, exR);
    fwrite(&F.N, 4, 1, out);
    fwrite(F.mvuRight.data(), 4, F.N, out); fwrite(F.mvDepth.data(), 4, F.N, out);

    // --- SearchByProjection(F, vpMapPoints, th): map points = right-image keypoints
    std::vector<MockMapPoint> mps(FR.mvKeys.size());
    std::vector<MockMapPoint*> vp;
    for (size_t i = 0; i < mps.size(); i++) {
      mps[i].mTrackProjX = FR.mvKeys[i].pt.x + 5.f; mps[i].mTrackProjY = FR.mvKeys[i].pt.y;
      mps[i].mTrackProjXR = FR.mvKeys[i].pt.x; mps[i].mnTrackScaleLevel = FR.mvKeys[i].octave;
      mps[i].mTrackViewCos = (i % 2) ? 0.9f : 1.0f; mps[i].desc = descR.row((int)i); mps[i].id = (int)i;
      vp.push_back(&mps[i]);
    }
    F.mvpMapPoints.assign(F.N, nullptr);
    int nm = matcher.SearchByProjection(F, vp, 3.0f);
    fwrite(&nm, 4, 1, out);
    for (int k = 0; k < F.N; k++) { int id = F.mvpMapPoints[k] ? F.mvpMapPoints[k]->id : -1; fwrite(&id, 4, 1, out); }
    int dd = ORBmatcher::DescriptorDistance(F.mDescriptors.row(0), descR.row(0));
    fwrite(&dd, 4, 1, out);

    // --- SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches) and (KeyFrame*, KeyFrame*, vpMatches12): keyframes = the two images
    MockKeyFrame kfL, kfR;
    kfL.N = F.N; kfL.mvKeysUn = F.mvKeys; kfL.mDescriptors = F.mDescriptors;
    kfR.N = (int)FR.mvKeys.size(); kfR.mvKeysUn = FR.mvKeys; kfR.mDescriptors = descR;
    fill_featvec(kfL.mFeatVec, kfL.mDescriptors, kfL.N); fill_featvec(kfR.mFeatVec, kfR.mDescriptors, kfR.N);
    std::vector<MockMapPoint> mpsL(kfL.N);
    kfL.mps.assign(kfL.N, nullptr);
    for (int i = 0; i < kfL.N; i++) { mpsL[i].id = i; if (i % 5 != 0) kfL.mps[i] = &mpsL[i]; }
    kfR.mps.assign(kfR.N, nullptr);
    for (int i = 0; i < kfR.N; i++) if (i % 7 != 0) kfR.mps[i] = &mps[i];
    MockFrame FB;
    FB.N = kfR.N; FB.mvKeys = FR.mvKeys; FB.mvKeysUn = FR.mvKeys; FB.mDescriptors = descR;
    std::map<unsigned, std::vector<unsigned> >& fvB = kfR.mFeatVec;
    struct FrameWithBow : MockFrame { std::map<unsigned, std::vector<unsigned> > mFeatVec; } FBB;
    static_cast<MockFrame&>(FBB) = FB; FBB.mFeatVec = fvB;
    std::vector<MockMapPoint*> vpMatches;
    ORBmatcher bow(0.75f, true);
    int nb = bow.SearchByBoW(&kfL, FBB, vpMatches);
    fwrite(&nb, 4, 1, out);
    for (int k = 0; k < FBB.N; k++) { int id = vpMatches[k] ? vpMatches[k]->id : -1; fwrite(&id, 4, 1, out); }
    std::vector<MockMapPoint*> vpMatches12;
    int nb2 = bow.SearchByBoW(&kfL, &kfR, vpMatches12);
    fwrite(&nb2, 4, 1, out);
    for (int k = 0; k < kfL.N; k++) { int id = vpMatches12[k] ? vpMatches12[k]->id : -1; fwrite(&id, 4, 1, out); }
  } catch (const std::exception& e) {
    fprintf(stderr, "exception: %s\n", e.what());
    return 4;
  }
  fclose(out);
  return 0;
}
