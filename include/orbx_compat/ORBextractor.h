// ORBextractor.h — drop-in for ORB-SLAM3's include/ORBextractor.h on top of liborbx.so.
//
// Same namespace, class name, constructor, operator(), getters and public mvImagePyramid as the
// reference (include/ORBextractor.h:43-112), so src/Frame.cc, src/Tracking.cc and the fork's
// RealSense / MAVLink drivers compile and link unchanged: replace the reference header with this
// one, drop src/ORBextractor.cc from the build and link -lorbx.
//
// Differences a caller can observe: none in results (keypoints, descriptors, pyramid bytes are
// bit-identical).  The CUDA plan for an image size is created on the first call
// with that size.  Errors of the C ABI surface as operator() returning -1 (the reference's only
// failure return, src/ORBextractor.cc:1090-1091) with orbx_last_error() set.
#ifndef ORBEXTRACTOR_H
#define ORBEXTRACTOR_H

#include <list>
#include <vector>

#include "../orbx.h"
#include "cv_compat.h"
#ifndef ORBX_NO_OPENCV
#include <opencv2/imgproc/imgproc.hpp>
#endif

namespace ORB_SLAM3 {

// Kept for source compatibility (include/ORBextractor.h:30-41); the quadtree itself runs on the GPU.
class ExtractorNode {
 public:
  ExtractorNode() : bNoMore(false) {}
  std::vector<cv::KeyPoint> vKeys;
  std::list<ExtractorNode>::iterator lit;
  bool bNoMore;
};

class ORBextractor {
 public:
  enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

  ORBextractor(int nfeatures_, float scaleFactor_, int nlevels_, int iniThFAST_, int minThFAST_)
      : nfeatures(nfeatures_), scaleFactor(scaleFactor_), nlevels(nlevels_), iniThFAST(iniThFAST_),
        minThFAST(minThFAST_), h_(nullptr), w_(0), hgt_(0) {
    // scale tables with the reference's float recurrences (src/ORBextractor.cc:414-430); they do
    // not depend on the image size, so they are available before the first frame like upstream
    mvScaleFactor.resize(nlevels); mvLevelSigma2.resize(nlevels);
    mvScaleFactor[0] = 1.0f; mvLevelSigma2[0] = 1.0f;
    for (int i = 1; i < nlevels; i++) {
      mvScaleFactor[i] = (float)(mvScaleFactor[i - 1] * scaleFactor);
      mvLevelSigma2[i] = mvScaleFactor[i] * mvScaleFactor[i];
    }
    mvInvScaleFactor.resize(nlevels); mvInvLevelSigma2.resize(nlevels);
    for (int i = 0; i < nlevels; i++) {
      mvInvScaleFactor[i] = 1.0f / mvScaleFactor[i];
      mvInvLevelSigma2[i] = 1.0f / mvLevelSigma2[i];
    }
    mvImagePyramid.resize(nlevels);
  }
  ~ORBextractor() { if (h_) orbx_destroy(h_); }
  ORBextractor(const ORBextractor&) = delete;
  ORBextractor& operator=(const ORBextractor&) = delete;

  // Compute the ORB features and descriptors on an image.  Mask is ignored, as upstream.
  int operator()(cv::InputArray _image, cv::InputArray /*_mask*/, std::vector<cv::KeyPoint>& _keypoints,
                 cv::OutputArray _descriptors, std::vector<int>& vLappingArea) {
    cv::Mat image = _image.getMat();
    if (image.empty()) return -1;
    if (image.type() != CV_8UC1) return -1;  // upstream asserts (src/ORBextractor.cc:1094)
    if (!ensure(image.cols, image.rows)) return -1;
    const int cap = orbx_max_keypoints(h_);
    kp_.resize(cap);
    desc_.resize((size_t)cap * 32);
    int n = 0, mono = 0;
    const int lap0 = vLappingArea.size() > 0 ? vLappingArea[0] : 0, lap1 = vLappingArea.size() > 1 ? vLappingArea[1] : 0;
    if (orbx_extract(h_, image.data, image.cols, image.rows, image.step, lap0, lap1,
                     reinterpret_cast<orbx_keypoint*>(kp_.data()), desc_.data(), cap, &n, &mono) != ORBX_OK)
      return -1;
    _keypoints.assign(kp_.begin(), kp_.begin() + n);
    if (n == 0) {
      _descriptors.release();
    } else {
      _descriptors.create(n, 32, CV_8U);
      cv::Mat d = _descriptors.getMat();
      for (int i = 0; i < n; i++) memcpy(d.ptr(i), desc_.data() + (size_t)i * 32, 32);
    }
    materialize_pyramid(image);
    return mono;
  }

  int inline GetLevels() { return nlevels; }
  float inline GetScaleFactor() { return (float)scaleFactor; }
  std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }
  std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }
  std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }
  std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

  // The C-ABI handle behind this instance (device-resident outputs and pyramid of the last call), for the
  // matcher-side entry points that run on them without a second upload (orbx_compat/StereoMatcher.h).
  orbx_handle handle() const { return h_; }

  // "GetPyramid": level image of the last call (same object mvImagePyramid[level] holds).
  const cv::Mat& GetPyramid(int level) const { return mvImagePyramid[level]; }

  // ROI views (step = w+38) into bordered buffers, refreshed on every call like upstream
  // (include/ORBextractor.h:83).  Set ORBX_LAZY_PYRAMID to skip the device->host copy when the
  // caller never reads it (monocular / RGB-D tracking does not).
  std::vector<cv::Mat> mvImagePyramid;

 protected:
  bool ensure(int w, int h) {
    if (h_ && w == w_ && h == hgt_) return true;
    if (h_) { orbx_destroy(h_); h_ = nullptr; }
    orbx_config cfg;
    cfg.nfeatures = nfeatures; cfg.scale_factor = (float)scaleFactor; cfg.nlevels = nlevels;
    cfg.ini_th_fast = iniThFAST; cfg.min_th_fast = minThFAST; cfg.width = w; cfg.height = h;
    cfg.max_batch = 1; cfg.device = 0; cfg.flags = 0;
    if (orbx_create(&cfg, &h_) != ORBX_OK) { h_ = nullptr; return false; }
    w_ = w; hgt_ = h;
    borders_.assign(nlevels, cv::Mat());
    return true;
  }
  void materialize_pyramid(const cv::Mat& image) {
    (void)image;
#ifndef ORBX_LAZY_PYRAMID
    // one transfer for all levels into the handle's pinned buffer, then one host copy per level into the Mats the
    // caller keeps (0.6 ms instead of 1.3 ms per 752x480 frame with eight separate orbx_get_pyramid calls)
    const uint8_t* base = nullptr;
    std::vector<size_t> ofs(nlevels), str(nlevels);
    const bool all = orbx_get_pyramid_all(h_, 0, 1, &base, ofs.data(), str.data()) == ORBX_OK;
    for (int l = 0; l < nlevels; l++) {
      int lw = 0, lh = 0;
      orbx_level_size(h_, l, &lw, &lh);
      borders_[l].create(lh + 38, lw + 38, CV_8U);
      if (all) {
        for (int y = 0; y < lh + 38; y++) memcpy(borders_[l].ptr(y), base + ofs[l] + (size_t)y * str[l], (size_t)lw + 38);
      } else {
        orbx_get_pyramid(h_, 0, l, borders_[l].data, borders_[l].step, 1);
      }
      mvImagePyramid[l] = borders_[l](cv::Rect(19, 19, lw, lh));
    }
    // Level 0 upstream uses a non-ISOLATED copyMakeBorder (src/ORBextractor.cc:1190): when the input
    // is a ROI of a larger Mat its border comes from the parent.  With real OpenCV available that
    // single border is refreshed here from the caller's Mat.
#ifndef ORBX_NO_OPENCV
    if (image.isSubmatrix()) cv::copyMakeBorder(image, borders_[0], 19, 19, 19, 19, cv::BORDER_REFLECT_101);
#endif
#endif
  }

  int nfeatures;
  double scaleFactor;
  int nlevels;
  int iniThFAST;
  int minThFAST;
  std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;

  orbx_handle h_;
  int w_, hgt_;
  std::vector<cv::KeyPoint> kp_;
  std::vector<unsigned char> desc_;
  std::vector<cv::Mat> borders_;
};

}  // namespace ORB_SLAM3

#endif
