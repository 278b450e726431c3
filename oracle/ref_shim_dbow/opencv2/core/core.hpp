// Header shim, TEST INFRASTRUCTURE ONLY: the few OpenCV core types the reference's Thirdparty/DBoW2 touches, so that
// its sources compile UNMODIFIED in an image without OpenCV headers (oracle/Makefile target `ref_dbow`).
//   cv::Mat                  a row-major byte / float matrix: create, zeros, clone, release, ptr<T>, rows, cols, empty
//   cv::FileStorage/FileNode declared so that the YAML save / load members of TemplatedVocabulary parse; never opened
//                            (the fork loads ORBvoc.txt through loadFromTextFile, System.cc:118)
#pragma once
// (the real core.hpp brings these in; TemplatedVocabulary.h relies on that for pow / log / stringstream)
#include <algorithm>
#include <cfloat>
#include <climits>
#include <cmath>
#include <cstdlib>
#include <iostream>
#include <sstream>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#define CV_8U 0
#define CV_32F 5

namespace cv {
typedef unsigned char uchar;

class Mat {
 public:
  int rows = 0, cols = 0;
  uchar* data = nullptr;
  Mat() {}
  void create(int r, int c, int type) {
    const size_t es = type == CV_32F ? 4 : 1;
    if (data && rows == r && cols == c && es == elem_) return;  // cv::Mat::create keeps a matching allocation
    buf_ = std::make_shared<std::vector<uchar> >((size_t)r * c * es);
    rows = r; cols = c; elem_ = es; data = buf_->data();
  }
  static Mat zeros(int r, int c, int type) { Mat m; m.create(r, c, type); memset(m.data, 0, (size_t)r * c * m.elem_); return m; }
  void release() { buf_.reset(); data = nullptr; rows = cols = 0; }
  Mat clone() const {
    Mat m;
    if (!data) return m;
    m.create(rows, cols, elem_ == 4 ? CV_32F : CV_8U);
    memcpy(m.data, data, (size_t)rows * cols * elem_);
    return m;
  }
  bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
  template <typename T> T* ptr(int y = 0) { return (T*)(data + (size_t)y * cols * elem_); }
  template <typename T> const T* ptr(int y = 0) const { return (const T*)(data + (size_t)y * cols * elem_); }

 private:
  size_t elem_ = 1;
  std::shared_ptr<std::vector<uchar> > buf_;
};

class FileNode {
 public:
  FileNode operator[](const std::string&) const { return FileNode(); }
  FileNode operator[](const char*) const { return FileNode(); }
  FileNode operator[](int) const { return FileNode(); }
  size_t size() const { return 0; }
  operator int() const { return 0; }
  operator float() const { return 0.f; }
  operator double() const { return 0.; }
  operator std::string() const { return std::string(); }
};

class FileStorage {
 public:
  enum { READ = 0, WRITE = 1 };
  FileStorage(const char*, int) {}
  FileStorage(const std::string&, int) {}
  bool isOpened() const { return false; }
  FileNode operator[](const std::string&) const { return FileNode(); }
  FileNode operator[](const char*) const { return FileNode(); }
};
template <class T> FileStorage& operator<<(FileStorage& fs, const T&) { return fs; }
}  // namespace cv
