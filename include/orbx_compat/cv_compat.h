// cv_compat.h — layout-identical stand-in for the handful of OpenCV types the ORBextractor /
// ORBmatcher signatures mention, used ONLY when OpenCV's headers are not installed
// (-DORBX_NO_OPENCV; this build image has no OpenCV C++ headers).  With real OpenCV the compat
// headers include <opencv2/core.hpp> instead and these definitions are not seen.
//
// sizeof(cv::KeyPoint) == 28 and the field order {pt.x, pt.y, size, angle, response, octave,
// class_id} match OpenCV, which is what lets std::vector<cv::KeyPoint> alias orbx_keypoint[].
#pragma once
#ifdef ORBX_NO_OPENCV
#include <stdint.h>
#include <string.h>

#include <memory>
#include <vector>

#define CV_8U 0
#define CV_8UC1 0

namespace cv {

struct Point2f {
  float x, y;
  Point2f() : x(0), y(0) {}
  Point2f(float x_, float y_) : x(x_), y(y_) {}
};

struct KeyPoint {
  Point2f pt;
  float size, angle, response;
  int octave, class_id;
  KeyPoint() : size(0), angle(-1), response(0), octave(0), class_id(-1) {}
};
static_assert(sizeof(KeyPoint) == 28, "cv::KeyPoint layout");

struct Rect {
  int x, y, width, height;
  Rect(int x_, int y_, int w_, int h_) : x(x_), y(y_), width(w_), height(h_) {}
};

// 8-bit single-channel matrix header with shared ownership, enough for ROI views.
class Mat {
 public:
  int rows = 0, cols = 0;
  unsigned char* data = nullptr;
  size_t step = 0;
  Mat() {}
  Mat(int r, int c, int /*type*/) { create(r, c, CV_8U); }
  Mat(int r, int c, int /*type*/, void* ext, size_t st) : rows(r), cols(c), data((unsigned char*)ext), step(st) {}
  void create(int r, int c, int /*type*/) {
    if (r == rows && c == cols && data && step == (size_t)c) return;
    buf_ = std::shared_ptr<unsigned char>(new unsigned char[(size_t)r * c + 1], std::default_delete<unsigned char[]>());
    rows = r; cols = c; step = (size_t)c; data = buf_.get();
  }
  void release() { buf_.reset(); rows = cols = 0; data = nullptr; step = 0; }
  bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
  int type() const { return CV_8UC1; }
  Mat operator()(const Rect& r) const {
    Mat m;
    m.rows = r.height; m.cols = r.width; m.step = step; m.data = data + (size_t)r.y * step + r.x; m.buf_ = buf_;
    return m;
  }
  template <typename T> T* ptr(int y = 0) { return (T*)(data + (size_t)y * step); }
  template <typename T> const T* ptr(int y = 0) const { return (const T*)(data + (size_t)y * step); }
  unsigned char* ptr(int y = 0) { return data + (size_t)y * step; }
  const unsigned char* ptr(int y = 0) const { return data + (size_t)y * step; }
  Mat row(int y) const { return (*this)(Rect(0, y, cols, 1)); }
  Mat getMat() const { return *this; }  // so that InputArray/OutputArray can be plain Mat references

 private:
  std::shared_ptr<unsigned char> buf_;
};

typedef const Mat& InputArray;
typedef Mat& OutputArray;

}  // namespace cv
#else
#include <opencv2/core/core.hpp>
#endif
