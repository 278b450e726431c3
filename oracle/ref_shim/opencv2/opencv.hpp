// opencv2/opencv.hpp — minimal OpenCV *header shim*, TEST INFRASTRUCTURE ONLY.
//
// Purpose: let /root/reference/src/ORBextractor.cc be compiled UNMODIFIED, where it lies, into
// oracle/_ref/libref_orbextractor.so (recipe: oracle/Makefile target `ref`).  The image has no
// OpenCV C++ headers (the cv2 wheel hides the C++ ABI), so this header provides only the ~25
// cv:: entities that file uses (ORBextractor.cc:55-58 includes; Mat / KeyPoint / Point / Size /
// Rect / InputArray / OutputArray / FAST / resize / copyMakeBorder / GaussianBlur / fastAtan2 /
// cvRound / cvFloor / cvCeil / KeyPointsFilter).
//
// The container types below reproduce cv::Mat's *sharing* semantics (refcounted buffer, ROI views,
// create() is a no-op on a same-size/type Mat, copyMakeBorder in place on a ROI of its own parent,
// non-ISOLATED borders taken from the parent of a ROI) because ComputePyramid relies on them
// (ORBextractor.cc:1170-1195).  The five image PRIMITIVES are not implemented here: they go through
// a function table (refshim::prims()) whose default entries are the oracle primitives that
// tests/test_oracle_vs_cv2.py pins bit-exact against the OpenCV 4.13 binary, and which a test can
// repoint at the real cv2 functions through ctypes callbacks (tests/test_ref_extractor.py), so the
// reference's own cell loop / octree / ordering / sampler run over real OpenCV results.
//
// Nothing in the product (orb_slam3_vio_fixes_b200/, include/) includes this file.
#ifndef REFSHIM_OPENCV_HPP
#define REFSHIM_OPENCV_HPP

#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>
#include <emmintrin.h>

typedef unsigned char uchar;

#define CV_PI 3.1415926535897932384626433832795
#define CV_8U 0
#define CV_8UC1 0

// core/fast_math.hpp: SSE2 builds use cvtsd2si / cvtss2si (current rounding mode = nearest even).
static inline int cvRound(double v) { return _mm_cvtsd_si32(_mm_set_sd(v)); }
static inline int cvRound(float v) { return _mm_cvtss_si32(_mm_set_ss(v)); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v) { int i = (int)v; return i - (i > v); }
static inline int cvFloor(float v) { int i = (int)v; return i - (i > v); }
static inline int cvFloor(int v) { return v; }
static inline int cvCeil(double v) { int i = (int)v; return i + (i < v); }
static inline int cvCeil(float v) { int i = (int)v; return i + (i < v); }
static inline int cvCeil(int v) { return v; }

// ---- primitive table (C layout, settable from ctypes) ------------------------------------------
extern "C" {
struct refshim_prims {
  // cv::resize(src, dst, Size(dw,dh), 0, 0, INTER_LINEAR), CV_8UC1, contiguous buffers
  void (*resize_linear)(const uint8_t* src, int sw, int sh, uint8_t* dst, int dw, int dh);
  // cv::copyMakeBorder(src, dst, b,b,b,b, BORDER_REFLECT_101 | ISOLATED); dst is (w+2b)x(h+2b)
  void (*border101)(const uint8_t* src, int w, int h, int b, uint8_t* dst);
  // cv::GaussianBlur(m, m, Size(7,7), 2, 2, BORDER_REFLECT_101) on a non-submatrix
  void (*gaussian_blur7)(const uint8_t* src, int w, int h, uint8_t* dst);
  // cv::fastAtan2
  float (*fast_atan2)(float y, float x);
  // cv::FAST(roi, kps, threshold, true) on the sub-rectangle [x0,x1)x[y0,y1) of a contiguous image;
  // out = (x, y, response) int triples in ROI coordinates, row-major order; returns the count
  int (*fast)(const uint8_t* img, int w, int h, int x0, int y0, int x1, int y1, int threshold, int* out,
              int cap);
};
refshim_prims* refshim_prims_table();
}

namespace cv {

template <typename T> struct Point_ {
  T x, y;
  Point_() : x(0), y(0) {}
  Point_(T _x, T _y) : x(_x), y(_y) {}
  template <typename U> Point_(const Point_<U>& o) : x((T)o.x), y((T)o.y) {}
  Point_& operator*=(float s) { x = (T)(x * s); y = (T)(y * s); return *this; }
  Point_& operator*=(double s) { x = (T)(x * s); y = (T)(y * s); return *this; }
  Point_& operator+=(const Point_& o) { x += o.x; y += o.y; return *this; }
};
typedef Point_<int> Point2i;
typedef Point_<float> Point2f;
typedef Point2i Point;

template <typename T> struct Size_ {
  T width, height;
  Size_() : width(0), height(0) {}
  Size_(T w, T h) : width(w), height(h) {}
};
typedef Size_<int> Size;

template <typename T> struct Rect_ {
  T x, y, width, height;
  Rect_() : x(0), y(0), width(0), height(0) {}
  Rect_(T _x, T _y, T w, T h) : x(_x), y(_y), width(w), height(h) {}
};
typedef Rect_<int> Rect;

struct KeyPoint {  // 28 bytes, the layout include/orbx.h's orbx_keypoint mirrors
  Point2f pt;
  float size, angle, response;
  int octave, class_id;
  KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
  KeyPoint(float x, float y, float _size, float _angle = -1, float _response = 0, int _octave = 0,
           int _class_id = -1)
      : pt(x, y), size(_size), angle(_angle), response(_response), octave(_octave), class_id(_class_id) {}
};
static_assert(sizeof(KeyPoint) == 28, "cv::KeyPoint layout");

enum { BORDER_CONSTANT = 0, BORDER_REPLICATE = 1, BORDER_REFLECT = 2, BORDER_WRAP = 3,
       BORDER_REFLECT_101 = 4, BORDER_REFLECT101 = 4, BORDER_DEFAULT = 4, BORDER_ISOLATED = 16 };
enum { INTER_NEAREST = 0, INTER_LINEAR = 1, INTER_CUBIC = 2, INTER_AREA = 3 };

struct MatStep {
  size_t v;
  MatStep(size_t s = 0) : v(s) {}
  operator size_t() const { return v; }
};

class _InputArray;
class _OutputArray;
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;

class Mat {
 public:
  int rows, cols;
  uchar* data;
  MatStep step;
  // parent allocation (for ROI bookkeeping, as cv::Mat::datastart / locateROI)
  std::shared_ptr<std::vector<uchar>> buf;
  int whole_rows, whole_cols, ofs_x, ofs_y;

  Mat() : rows(0), cols(0), data(nullptr), step(0), whole_rows(0), whole_cols(0), ofs_x(0), ofs_y(0) {}
  Mat(int r, int c, int type) : Mat() { create(r, c, type); }
  Mat(Size sz, int type) : Mat() { create(sz.height, sz.width, type); }
  // external data, not owned (cv::Mat(rows, cols, type, void* data, size_t step))
  Mat(int r, int c, int type, void* d, size_t st = 0) : Mat() {
    assert(type == CV_8UC1);
    rows = whole_rows = r; cols = whole_cols = c; data = (uchar*)d; step = st ? st : (size_t)c;
  }

  void create(int r, int c, int type) {
    assert(type == CV_8UC1);
    (void)type;
    if (data && rows == r && cols == c) return;  // cv::Mat::create: same size & type keeps the data
    buf = std::make_shared<std::vector<uchar>>((size_t)r * c);
    rows = whole_rows = r; cols = whole_cols = c; ofs_x = ofs_y = 0;
    data = buf->data(); step = (size_t)c;
  }
  void create(Size sz, int type) { create(sz.height, sz.width, type); }
  void release() { buf.reset(); data = nullptr; rows = cols = 0; step = 0; whole_rows = whole_cols = ofs_x = ofs_y = 0; }

  static Mat zeros(int r, int c, int type) { Mat m(r, c, type); if (m.data) memset(m.data, 0, (size_t)r * c); return m; }

  int type() const { return CV_8UC1; }
  bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
  size_t step1() const { return step.v; }
  size_t elemSize() const { return 1; }
  bool isSubmatrix() const { return rows != whole_rows || cols != whole_cols; }
  bool isContinuous() const { return step.v == (size_t)cols || rows == 1; }
  void locateROI(Size& whole, Point& ofs) const { whole = Size(whole_cols, whole_rows); ofs = Point(ofs_x, ofs_y); }

  Mat operator()(const Rect& r) const {
    assert(r.x >= 0 && r.y >= 0 && r.x + r.width <= cols && r.y + r.height <= rows);
    Mat m(*this);
    m.rows = r.height; m.cols = r.width;
    m.data = data + (size_t)r.y * step.v + r.x;
    m.ofs_x = ofs_x + r.x; m.ofs_y = ofs_y + r.y;
    return m;
  }
  Mat rowRange(int a, int b) const { return (*this)(Rect(0, a, cols, b - a)); }
  Mat colRange(int a, int b) const { return (*this)(Rect(a, 0, b - a, rows)); }
  Mat row(int y) const { return rowRange(y, y + 1); }

  uchar* ptr(int y = 0) { return data + (size_t)y * step.v; }
  const uchar* ptr(int y = 0) const { return data + (size_t)y * step.v; }
  template <typename T> T* ptr(int y = 0) { return (T*)(data + (size_t)y * step.v); }
  template <typename T> const T* ptr(int y = 0) const { return (const T*)(data + (size_t)y * step.v); }
  template <typename T> T& at(int y, int x) { return ((T*)(data + (size_t)y * step.v))[x]; }
  template <typename T> const T& at(int y, int x) const { return ((const T*)(data + (size_t)y * step.v))[x]; }

  Mat clone() const {
    Mat m(rows, cols, CV_8UC1);
    for (int y = 0; y < rows; y++) memcpy(m.ptr(y), ptr(y), (size_t)cols);
    return m;
  }
  inline void copyTo(OutputArray dst) const;

  // contiguous copy in / out (shim helpers for the primitive table)
  std::vector<uchar> packed() const {
    std::vector<uchar> v((size_t)rows * cols);
    for (int y = 0; y < rows; y++) memcpy(v.data() + (size_t)y * cols, ptr(y), (size_t)cols);
    return v;
  }
  void unpack(const std::vector<uchar>& v) {
    for (int y = 0; y < rows; y++) memcpy(ptr(y), v.data() + (size_t)y * cols, (size_t)cols);
  }
};

class _InputArray {
 public:
  const Mat* m;
  _InputArray() : m(nullptr) {}
  _InputArray(const Mat& mm) : m(&mm) {}
  bool empty() const { return !m || m->empty(); }
  Mat getMat() const { return m ? *m : Mat(); }
};
class _OutputArray : public _InputArray {
 public:
  _OutputArray() {}
  _OutputArray(Mat& mm) : _InputArray(mm) {}
  _OutputArray(const Mat& mm) : _InputArray(mm) {}  // fixed-size view (e.g. a row of another Mat)
  void create(int r, int c, int type) const { const_cast<Mat*>(m)->create(r, c, type); }
  void create(Size sz, int type) const { const_cast<Mat*>(m)->create(sz, type); }
  void release() const { const_cast<Mat*>(m)->release(); }
  Mat& getMatRef() const { return *const_cast<Mat*>(m); }
};
static inline InputArray noArray() { static _InputArray none; return none; }

inline void Mat::copyTo(OutputArray dst) const {
  dst.create(rows, cols, CV_8UC1);
  Mat d = dst.getMat();
  for (int y = 0; y < rows; y++) if (d.ptr(y) != ptr(y)) memcpy(d.ptr(y), ptr(y), (size_t)cols);
}

// ---- primitives routed through the table --------------------------------------------------------
static inline float fastAtan2(float y, float x) { return refshim_prims_table()->fast_atan2(y, x); }

static inline void resize(InputArray _src, OutputArray _dst, Size dsize, double fx = 0, double fy = 0,
                          int interpolation = INTER_LINEAR) {
  assert(interpolation == INTER_LINEAR && fx == 0 && fy == 0);
  (void)fx; (void)fy; (void)interpolation;
  Mat src = _src.getMat();
  _dst.create(dsize, src.type());
  Mat dst = _dst.getMat();
  std::vector<uchar> s = src.packed(), d((size_t)dsize.width * dsize.height);
  refshim_prims_table()->resize_linear(s.data(), src.cols, src.rows, d.data(), dsize.width, dsize.height);
  dst.unpack(d);
}

// imgproc/src/filter? no: core/src/copy.cpp cv::copyMakeBorder.  A ROI source without BORDER_ISOLATED
// takes as much of the border as exists from its parent Mat; the rest is extrapolated.
static inline void copyMakeBorder(InputArray _src, OutputArray _dst, int top, int bottom, int left, int right,
                                  int borderType) {
  Mat src = _src.getMat();
  assert((borderType & ~BORDER_ISOLATED) == BORDER_REFLECT_101);
  if (src.isSubmatrix() && (borderType & BORDER_ISOLATED) == 0) {
    Size wholeSize; Point ofs;
    src.locateROI(wholeSize, ofs);
    int dtop = std::min(ofs.y, top), dbottom = std::min(wholeSize.height - src.rows - ofs.y, bottom);
    int dleft = std::min(ofs.x, left), dright = std::min(wholeSize.width - src.cols - ofs.x, right);
    Mat wider(src);
    wider.data = src.data - (size_t)dtop * src.step.v - dleft;
    wider.rows = src.rows + dtop + dbottom; wider.cols = src.cols + dleft + dright;
    wider.ofs_x -= dleft; wider.ofs_y -= dtop;
    src = wider;
    top -= dtop; left -= dleft; bottom -= dbottom; right -= dright;
  }
  _dst.create(src.rows + top + bottom, src.cols + left + right, src.type());
  Mat dst = _dst.getMat();
  if (top == 0 && left == 0 && bottom == 0 && right == 0) { src.copyTo(dst); return; }
  // The table primitive makes a uniform border b on every side; pad asymmetric requests to the
  // largest side and crop (REFLECT_101 extrapolation of a pixel does not depend on the other sides).
  int b = std::max(std::max(top, bottom), std::max(left, right));
  std::vector<uchar> s = src.packed(), d((size_t)(src.cols + 2 * b) * (src.rows + 2 * b));
  refshim_prims_table()->border101(s.data(), src.cols, src.rows, b, d.data());
  const size_t dp = (size_t)src.cols + 2 * b;
  for (int y = 0; y < dst.rows; y++) {
    const uchar* from = d.data() + (size_t)(y + b - top) * dp + (b - left);
    uchar* to = dst.ptr(y);
    if (to != from) memmove(to, from, (size_t)dst.cols);
  }
}

static inline void GaussianBlur(InputArray _src, OutputArray _dst, Size ksize, double sigmaX, double sigmaY = 0,
                                int borderType = BORDER_DEFAULT) {
  Mat src = _src.getMat();
  assert(ksize.width == 7 && ksize.height == 7 && sigmaX == 2 && sigmaY == 2 && borderType == BORDER_REFLECT_101);
  assert(!src.isSubmatrix());  // the reference blurs a clone(); a ROI would read its parent's pixels
  (void)ksize; (void)sigmaX; (void)sigmaY; (void)borderType;
  std::vector<uchar> s = src.packed(), d(s.size());
  refshim_prims_table()->gaussian_blur7(s.data(), src.cols, src.rows, d.data());
  _dst.create(src.rows, src.cols, src.type());
  Mat dst = _dst.getMat();
  dst.unpack(d);
}

static inline void FAST(InputArray _img, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression = true) {
  assert(nonmaxSuppression);
  (void)nonmaxSuppression;
  Mat img = _img.getMat();
  keypoints.clear();
  std::vector<uchar> s = img.packed();
  std::vector<int> out((size_t)3 * (img.rows * img.cols / 4 + 16));  // NMS keeps at most 1 in 4 pixels
  int n = refshim_prims_table()->fast(s.data(), img.cols, img.rows, 0, 0, img.cols, img.rows, threshold,
                                      out.data(), (int)(out.size() / 3));
  keypoints.reserve(n);
  for (int i = 0; i < n; i++)  // features2d/src/fast.cpp: KeyPoint((float)j, (float)(i-1), 7.f, -1, (float)score)
    keypoints.push_back(KeyPoint((float)out[3 * i], (float)out[3 * i + 1], 7.f, -1.f, (float)out[3 * i + 2]));
}

struct KeyPointsFilter {
  // features2d/src/keypoint.cpp; only reachable from the reference's dead ComputeKeyPointsOld
  // (ORBextractor.cc:898-1075, call commented out at :1101).  Kept so the file links.
  static void retainBest(std::vector<KeyPoint>& kps, int n) {
    if (n >= 0 && kps.size() > (size_t)n) {
      if (n == 0) { kps.clear(); return; }
      std::nth_element(kps.begin(), kps.begin() + n - 1, kps.end(),
                       [](const KeyPoint& a, const KeyPoint& b) { return a.response > b.response; });
      float ambiguous = kps[n - 1].response;
      auto end = std::partition(kps.begin() + n, kps.end(), [=](const KeyPoint& k) { return k.response >= ambiguous; });
      kps.resize(end - kps.begin());
    }
  }
};

}  // namespace cv
#endif
