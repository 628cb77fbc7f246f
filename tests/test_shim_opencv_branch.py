"""The ORBX_WITH_OPENCV branch of host/*.h (what a drop-in build of the reference compiles) must build: compiled here against the
stand-in opencv2/core/core.hpp of oracle/ref_shim (this image has no C++ OpenCV), with every template of the matcher shim
instantiated on a minimal Frame / KeyFrame / MapPoint model that has exactly the reference's member names."""
import os
import shutil
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SRC = r'''
#define ORBX_WITH_OPENCV 1
#include <list>
#include <map>
#include <mutex>
#include <set>
#include "slam_world.h"          // the reference's data model (member names of Frame.h / KeyFrame.h / MapPoint.h)
#include "host/ORBextractor.h"
#include "host/ORBmatcher.h"
#include "host/FramePreproc.h"
using namespace ORB_SLAM3;
int instantiate(Frame& F, Frame& G, KeyFrame* K, KeyFrame* K2, std::vector<MapPoint*>& mps, cv::Mat& img) {
  ORBextractor ex(1000, 1.2f, 8, 20, 7);
  std::vector<cv::KeyPoint> kps; cv::Mat desc; std::vector<int> lap{0, 0};
  int n = ex(img, cv::Mat(), kps, desc, lap);
  ex.SyncPyramid();
  ORBmatcher m(0.9f, true);
  std::set<MapPoint*> found; std::vector<KeyFrame*> kfs, mkfs; std::vector<std::pair<size_t, size_t> > pairs; std::vector<cv::Point2f> prev; std::vector<int> m12;
  cv::Mat Scw = cv::Mat::eye(4, 4, CV_32F), R = cv::Mat::eye(3, 3, CV_32F), t = cv::Mat::zeros(3, 1, CV_32F);
  n += m.SearchByProjection(F, mps, 3.f, false, 50.f);
  n += m.SearchByProjection(F, G, 7.f, false);
  n += m.SearchByProjection(F, K, found, 10.f, 100);
  n += m.SearchByProjection(K, Scw, mps, mps, 4, 1.0f);
  n += m.SearchByProjection(K, Scw, mps, kfs, mps, mkfs, 4, 1.0f);
  n += m.SearchForInitialization(F, G, prev, m12, 100);
  n += m.SearchByBoW(K, F, mps);
  n += m.SearchByBoW(K, K2, mps);
  n += m.SearchForTriangulation(K, K2, R, pairs, false, false);
  n += m.Fuse(K, mps, 3.f, false);
  n += m.Fuse(K, Scw, mps, 4.f, mps);
  n += m.SearchBySim3(K, K2, mps, 1.0f, R, t, 7.5f);
  n += ORBmatcher::DescriptorDistance(desc.row(0), desc.row(1));
  FramePreproc pre;
  cv::Mat gray;
  pre.cvtColorToGray(img, true, gray);
  pre.clahe(gray);
  return n;
}
'''


def test_opencv_branch_compiles(tmp_path):
    cxx = shutil.which("g++") or "/opt/gcc/bin/g++"
    src = tmp_path / "opencv_branch.cc"
    src.write_text(SRC)
    cmd = [cxx, "-std=c++11", "-fsyntax-only", "-Wall", "-Wno-sign-compare", "-Wno-unused",
           "-I" + os.path.join(ROOT, "oracle", "ref_shim"), "-I" + os.path.join(ROOT, "3_orb_slam3_selfnote_b200"), str(src)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-4000:]
