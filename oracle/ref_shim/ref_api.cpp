// ref_api.cpp — C entry points around the REFERENCE's own ORB_SLAM3::ORBextractor, compiled
// unmodified from /root/reference/src/ORBextractor.cc over the header shim in ./opencv2/.
// TEST INFRASTRUCTURE ONLY (oracle/_ref/libref_orbextractor.so): tests/ use it to pin
// oracle/orb_oracle.cpp and the committed goldens; bench.py's cpu_baseline may time it.
//
// Default primitive table = the oracle's exported primitives (liborb_oracle.so), each pinned
// bit-exact against the OpenCV 4.13 binary by tests/test_oracle_vs_cv2.py.  ref_set_prims()
// repoints the table, e.g. at ctypes callbacks into the real cv2.
#include <cstdint>
#include <cstring>
#include <vector>

#include "ORBextractor.h"  // /root/reference/include (which includes <opencv2/opencv.hpp> = the shim)

extern "C" {
// oracle/orb_oracle.cpp exports
void oracle_resize_linear(const uint8_t* src, int sw, int sh, uint8_t* dst, int dw, int dh);
void oracle_gaussian_blur7(const uint8_t* src, int w, int h, uint8_t* dst);
void oracle_border101(const uint8_t* src, int w, int h, int b, uint8_t* dst);
float oracle_fast_atan2(float y, float x);
int oracle_fast(const uint8_t* img, int w, int h, int x0, int y0, int x1, int y1, int threshold, int* out, int cap);

static refshim_prims g_prims = {oracle_resize_linear, oracle_border101, oracle_gaussian_blur7, oracle_fast_atan2,
                                oracle_fast};
refshim_prims* refshim_prims_table() { return &g_prims; }

// Any null entry keeps the current one.  Not thread-safe against running extractions.
void ref_set_prims(const refshim_prims* p) {
  if (!p) { g_prims = {oracle_resize_linear, oracle_border101, oracle_gaussian_blur7, oracle_fast_atan2, oracle_fast}; return; }
  if (p->resize_linear) g_prims.resize_linear = p->resize_linear;
  if (p->border101) g_prims.border101 = p->border101;
  if (p->gaussian_blur7) g_prims.gaussian_blur7 = p->gaussian_blur7;
  if (p->fast_atan2) g_prims.fast_atan2 = p->fast_atan2;
  if (p->fast) g_prims.fast = p->fast;
}

void* ref_create(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh) {
  return new ORB_SLAM3::ORBextractor(nfeatures, scaleFactor, nlevels, iniTh, minTh);
}
void ref_destroy(void* h) { delete (ORB_SLAM3::ORBextractor*)h; }

// ORBextractor::operator() (ORBextractor.h:57-59).  (px, py) = offset of the image inside a larger
// parent buffer of parent_w x parent_h (0,0,w,h for a plain image): level 0's border is not ISOLATED
// (ORBextractor.cc:1190), so a ROI input takes its border pixels from the parent.
// Returns monoIndex (-1 on an empty image); *n_out = number of keypoints.
int ref_extract_roi(void* h, const uint8_t* parent, int parent_w, int parent_h, size_t stride, int px, int py, int w,
                    int hgt, int lap0, int lap1, void* kps, uint8_t* desc, int cap, int* n_out) {
  auto* e = (ORB_SLAM3::ORBextractor*)h;
  std::vector<cv::KeyPoint> k;
  cv::Mat d;
  std::vector<int> lap = {lap0, lap1};
  cv::Mat img;
  if (parent && w > 0 && hgt > 0) {
    cv::Mat whole(parent_h, parent_w, CV_8UC1, (void*)parent, stride);
    img = whole(cv::Rect(px, py, w, hgt));
  }
  int mono = (*e)(img, cv::Mat(), k, d, lap);
  int n = (int)k.size();
  if (n_out) *n_out = n;
  if (mono < 0) return mono;
  if (n > cap) return -2;
  if (n) {
    memcpy(kps, k.data(), (size_t)n * sizeof(cv::KeyPoint));
    for (int i = 0; i < n; i++) memcpy(desc + (size_t)i * 32, d.ptr(i), 32);
  }
  return mono;
}
int ref_extract(void* h, const uint8_t* img, int w, int hgt, size_t stride, int lap0, int lap1, void* kps,
                uint8_t* desc, int cap, int* n_out) {
  return ref_extract_roi(h, img, w, hgt, stride, 0, 0, w, hgt, lap0, lap1, kps, desc, cap, n_out);
}

int ref_levels(void* h) { return ((ORB_SLAM3::ORBextractor*)h)->GetLevels(); }
void ref_tables(void* h, float* sf, float* inv, float* sig2, float* invsig2) {
  auto* e = (ORB_SLAM3::ORBextractor*)h;
  auto a = e->GetScaleFactors(), b = e->GetInverseScaleFactors(), c = e->GetScaleSigmaSquares(),
       d = e->GetInverseScaleSigmaSquares();
  for (size_t i = 0; i < a.size(); i++) { sf[i] = a[i]; inv[i] = b[i]; sig2[i] = c[i]; invsig2[i] = d[i]; }
}
int ref_level_size(void* h, int level, int* w, int* hgt) {
  auto* e = (ORB_SLAM3::ORBextractor*)h;
  if (level < 0 || level >= (int)e->mvImagePyramid.size() || e->mvImagePyramid[level].empty()) return -1;
  *w = e->mvImagePyramid[level].cols; *hgt = e->mvImagePyramid[level].rows;
  return 0;
}
// mvImagePyramid[level] (ORBextractor.h:83); with_border=1 returns the (w+38)x(h+38) parent buffer.
int ref_get_pyramid(void* h, int level, uint8_t* dst, size_t dst_stride, int with_border) {
  auto* e = (ORB_SLAM3::ORBextractor*)h;
  if (level < 0 || level >= (int)e->mvImagePyramid.size() || e->mvImagePyramid[level].empty()) return -1;
  const cv::Mat& m = e->mvImagePyramid[level];
  const int b = with_border ? 19 : 0;
  for (int y = -b; y < m.rows + b; y++)
    memcpy(dst + (size_t)(y + b) * dst_stride, m.data + (ptrdiff_t)y * (ptrdiff_t)m.step.v - b, (size_t)m.cols + 2 * b);
  return 0;
}

// Test access to the protected members the constructor fills (mnFeaturesPerLevel, umax).
struct RefPeek : ORB_SLAM3::ORBextractor {
  using ORB_SLAM3::ORBextractor::mnFeaturesPerLevel;
  using ORB_SLAM3::ORBextractor::umax;
};
void ref_quota_umax(void* h, int* nfeat, int* umax16) {
  auto* e = (RefPeek*)(ORB_SLAM3::ORBextractor*)h;
  for (size_t i = 0; i < e->mnFeaturesPerLevel.size(); i++) nfeat[i] = e->mnFeaturesPerLevel[i];
  for (size_t i = 0; i < e->umax.size() && i < 16; i++) umax16[i] = e->umax[i];
}
}  // extern "C"
