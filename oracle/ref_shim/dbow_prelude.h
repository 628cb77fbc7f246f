// CPU ORACLE support (test infrastructure, not product code).
// Force-included ahead of the reference's Thirdparty/DBoW2/DBoW2/*.{h,cpp} (oracle/Makefile, target _ref/libref_bow.so) so that those
// files compile UNMODIFIED against the stand-in cv:: header: TemplatedVocabulary.h uses std::stringstream without including <sstream>
// (the real opencv2/core/core.hpp drags it in) and declares YAML save / load through cv::FileStorage (TemplatedVocabulary.h:265-275,
// :1452-1618).  Those members are virtual, hence instantiated, but nothing here calls them: the stand-ins below abort when used.  The
// path under test is loadFromTextFile (:1338-1423, how ORB-SLAM3 loads ORBvoc.txt) and transform (:1130-1259).
#pragma once
#include <cstdio>
#include <cstdlib>
#include <sstream>
#include <string>

#include <opencv2/core/core.hpp>

namespace cv {
class FileNode {
 public:
  FileNode operator[](const char*) const { die(); return FileNode(); }
  FileNode operator[](const std::string&) const { die(); return FileNode(); }
  FileNode operator[](int) const { die(); return FileNode(); }
  size_t size() const { die(); return 0; }
  operator int() const { die(); return 0; }
  operator double() const { die(); return 0; }
  operator std::string() const { die(); return std::string(); }

 private:
  static void die() { fprintf(stderr, "oracle/_ref: cv::FileStorage is not available in the stand-in (YAML vocabularies are not on the path)\n"); abort(); }
};
class FileStorage {
 public:
  enum { READ = 0, WRITE = 1 };
  FileStorage(const char*, int) {}
  FileStorage(const std::string&, int) {}
  bool isOpened() const { return false; }
  FileNode operator[](const char*) const { return FileNode()[0]; }
  FileNode operator[](const std::string&) const { return FileNode()[0]; }
};
template <class T>
FileStorage& operator<<(FileStorage& fs, const T&) { return fs; }
}  // namespace cv
