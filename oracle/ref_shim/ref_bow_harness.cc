// CPU ORACLE support (test infrastructure, not product code).
//
// C entry points around the reference's own, UNMODIFIED DBoW2 (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h, FORB.cpp, BowVector.cpp,
// FeatureVector.cpp, ScoringObject.cpp -- vendored in the reference's tree and compiled by oracle/Makefile by path against the stand-in
// cv:: header): the vocabulary as ORB-SLAM3 loads it (loadFromTextFile, System.cc:95) and the transform Frame::ComputeBoW runs
// (src/Frame.cc:828-835: transform(vCurrentDesc, mBowVec, mFeatVec, 4)).  Loaded by oracle/refbow.py for tests/ only.
#include <cstdint>
#include <cstring>
#include <vector>

#include "FORB.h"
#include "TemplatedVocabulary.h"

// include/ORBVocabulary.h:31-32
typedef DBoW2::TemplatedVocabulary<DBoW2::FORB::TDescriptor, DBoW2::FORB> ORBVocabulary;

extern "C" {

void* rb_vocab_load(const char* path) {
  ORBVocabulary* v = new ORBVocabulary();
  if (!v->loadFromTextFile(path)) { delete v; return nullptr; }
  return v;
}
void rb_vocab_destroy(void* h) { delete (ORBVocabulary*)h; }
int rb_vocab_words(void* h) { return (int)((ORBVocabulary*)h)->size(); }

// Frame::ComputeBoW: the descriptor matrix becomes one cv::Mat per row (src/Converter.cc:33-41, toDescriptorVector), then
// transform(features, BowVector, FeatureVector, levelsup).  Outputs in the containers' own (ascending key) order; capacities = n.
void rb_transform(void* h, const uint8_t* desc, int n, int levelsup, uint32_t* bowIds, double* bowVals, int* nbow, uint32_t* fvNodes,
                  int* fvStart, uint32_t* fvIdx, int* nfv) {
  cv::Mat D(n > 0 ? n : 1, 32, CV_8U);
  if (n > 0) memcpy(D.ptr(0), desc, (size_t)n * 32);
  std::vector<cv::Mat> vDesc;
  vDesc.reserve(n);
  for (int j = 0; j < n; j++) vDesc.push_back(D.row(j));
  DBoW2::BowVector bow;
  DBoW2::FeatureVector fv;
  ((ORBVocabulary*)h)->transform(vDesc, bow, fv, levelsup);
  int k = 0;
  for (DBoW2::BowVector::const_iterator it = bow.begin(); it != bow.end(); ++it, ++k) { bowIds[k] = it->first; bowVals[k] = it->second; }
  *nbow = k;
  int m = 0, o = 0;
  for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it, ++m) {
    fvNodes[m] = it->first;
    fvStart[m] = o;
    for (size_t q = 0; q < it->second.size(); q++) fvIdx[o++] = it->second[q];
  }
  fvStart[m] = o;
  *nfv = m;
}

// transform(feature) -> word id (TemplatedVocabulary.h:1200-1214), per feature
void rb_transform_words(void* h, const uint8_t* desc, int n, uint32_t* words) {
  for (int j = 0; j < n; j++) {
    cv::Mat d(1, 32, CV_8U);
    memcpy(d.ptr(0), desc + (size_t)j * 32, 32);
    words[j] = ((ORBVocabulary*)h)->transform(d);
  }
}

}  // extern "C"
