// frame_ref_api.cpp — C entry points of oracle/_ref/libref_frame.so.  TEST INFRASTRUCTURE ONLY (tests/test_ref_frame.py
// compares these with the oracle restatements; nothing in the product links or loads it).
//
// /root/reference/src/Frame.cc cannot be compiled as a whole here (constructors, IMU and BoW members need Eigen, Sophus,
// g2o, boost, real OpenCV).  The four functions of the hot path can: `make ref_frame` cuts
//   void Frame::ComputeStereoMatches()            Frame.cc:811-981
//   void Frame::ComputeStereoFromRGBD(...)        Frame.cc:984-1005
//   bool Frame::isInFrustum(...)                  Frame.cc:512-586
//   bool Frame::isInFrustumChecks(...)            Frame.cc:1168-1242
//   void Frame::ComputeStereoFishEyeMatches()     Frame.cc:1126-1166 (TriangulateMatches answered from a table)
// verbatim from the reference file into a scratch file outside the repository (oracle/extract_ref_functions.py) and
// this translation unit includes it as REF_FRAME_FUNCTIONS inside namespace ORB_SLAM3, against the stand-in Frame /
// MapPoint / Eigen / Sophus / cv types of prelude.h and the reference's own include/ORBmatcher.h (TH_HIGH, TH_LOW,
// DescriptorDistance — linked from the reference's ORBmatcher.cc).  Every loop, gate, comparison, rounding call and
// store of those functions is the reference's text; the stand-ins hold data and do the 3-vector arithmetic in plain
// fp32 loops (this file is compiled with -ffp-contract=off, the fp contract DESIGN.md §2 states for the path).
#include <algorithm>
#include <climits>
#include <cmath>

#include "prelude.h"

namespace cv {
// cv::norm(a, b, NORM_L1) on 8-bit single-channel views (Frame.cc:932): sum of absolute differences
enum { NORM_L1 = 2 };
inline double norm(const Mat& a, const Mat& b, int /*normType*/) {
  double s = 0;
  for (int y = 0; y < a.rows; y++) {
    const uchar* pa = a.ptr(y);
    const uchar* pb = b.ptr(y);
    for (int x = 0; x < a.cols; x++) s += std::abs((int)pa[x] - (int)pb[x]);
  }
  return s;
}
}  // namespace cv

#include "ORBmatcher.h"

namespace ORB_SLAM3 {
#include REF_FRAME_FUNCTIONS
}  // namespace ORB_SLAM3

using namespace ORB_SLAM3;

cv::BFMatcher Frame::BFmatcher = cv::BFMatcher(cv::NORM_HAMMING);  // Frame.cc:43

namespace {
struct KP { float x, y, size, angle, response; int32_t octave, class_id; };  // cv::KeyPoint memory layout
std::vector<cv::KeyPoint> to_keys(const void* kps_v, int n) {
  const KP* k = (const KP*)kps_v;
  std::vector<cv::KeyPoint> v(n);
  for (int i = 0; i < n; i++) {
    v[i].pt.x = k[i].x; v[i].pt.y = k[i].y; v[i].size = k[i].size; v[i].angle = k[i].angle;
    v[i].response = k[i].response; v[i].octave = k[i].octave; v[i].class_id = k[i].class_id;
  }
  return v;
}
cv::Mat to_desc(const uint8_t* d, int n) {
  cv::Mat m(n > 0 ? n : 1, 32, CV_8UC1);
  if (n > 0) memcpy(m.data, d, (size_t)n * 32);
  return m;
}
Eigen::Matrix3f mat3(const float* m) { Eigen::Matrix3f r; for (int i = 0; i < 9; i++) r.m[i] = m[i]; return r; }
Eigen::Vector3f vec3(const float* v) { return Eigen::Vector3f(v[0], v[1], v[2]); }
}  // namespace

extern "C" {

// Frame::ComputeStereoMatches on one pair.  Pyramid level l of each eye: contiguous pw[l] x ph[l] bytes.
// Returns the number of keypoints with a depth; u_right / depth = mvuRight / mvDepth.
int ref_compute_stereo_matches(const void* kps_l, const uint8_t* desc_l, int n_l, const void* kps_r, const uint8_t* desc_r,
                               int n_r, const float* scale_factors, const float* inv_scale_factors, int nlevels,
                               const uint8_t* const* pyr_l, const uint8_t* const* pyr_r, const int* pw, const int* ph,
                               float mb, float mbf, float* u_right, float* depth) {
  Frame F;
  ORBextractor eL, eR;
  for (int l = 0; l < nlevels; l++) {
    eL.mvImagePyramid.push_back(cv::Mat(ph[l], pw[l], CV_8UC1, (void*)pyr_l[l]));
    eR.mvImagePyramid.push_back(cv::Mat(ph[l], pw[l], CV_8UC1, (void*)pyr_r[l]));
  }
  F.mpORBextractorLeft = &eL; F.mpORBextractorRight = &eR;
  F.N = n_l;
  F.mvKeys = to_keys(kps_l, n_l);
  F.mvKeysRight = to_keys(kps_r, n_r);
  F.mDescriptors = to_desc(desc_l, n_l);
  F.mDescriptorsRight = to_desc(desc_r, n_r);
  F.mvScaleFactors.assign(scale_factors, scale_factors + nlevels);
  F.mvInvScaleFactors.assign(inv_scale_factors, inv_scale_factors + nlevels);
  F.mb = mb; F.mbf = mbf;
  F.ComputeStereoMatches();
  int cnt = 0;
  for (int i = 0; i < n_l; i++) { u_right[i] = F.mvuRight[i]; depth[i] = F.mvDepth[i]; cnt += F.mvDepth[i] > 0; }
  return cnt;
}

// MapPoint::ComputeDistinctiveDescriptors (the reference's own text, mappoint_ref.cpp).  Observation o is keyframe o
// (keyframes live in one array, so the std::map<KeyFrame*, ...> of the MapPoint visits them in index order); it
// contributes descriptor row first[o] and, when second[o] >= 0, that row too (the right-camera index of a rig).
// kf_bad[o] != 0: that keyframe isBad().  Writes the chosen descriptor (32 bytes); returns 1 when one was chosen.
int ref_distinctive_descriptor(const uint8_t* desc, int n_rows, const int* first, const int* second, const uint8_t* kf_bad,
                               int n_obs, uint8_t* chosen) {
  std::vector<KeyFrame> kfs(n_obs);
  MapPoint p;
  for (int o = 0; o < n_obs; o++) {
    kfs[o].mDescriptors = to_desc(desc, n_rows);
    kfs[o].mbBad = kf_bad && kf_bad[o];
    p.mObservations[&kfs[o]] = std::tuple<int, int>(first[o], second[o]);
  }
  p.ComputeDistinctiveDescriptors();
  if (p.mDescriptor.empty()) return 0;
  memcpy(chosen, p.mDescriptor.data, 32);
  return 1;
}

// Frame::AssignFeaturesToGrid + GetFeaturesInArea (the reference's own text, grid_ref.cpp) on a single-camera frame:
// the candidates of query q, in the order the reference returns them, are idx[off[q] .. off[q+1]).  Returns the total
// number of candidates (idx holds the first `cap` of them).
int ref_grid_queries(const void* kps, int n, const float* bounds, const float* qx, const float* qy, const float* qr,
                     const int* qmin, const int* qmax, int nq, int* off, int* idx, int cap) {
  Frame F;
  F.N = n;
  F.mvKeysUn = to_keys(kps, n);
  F.mnMinX = bounds[0]; F.mnMaxX = bounds[1]; F.mnMinY = bounds[2]; F.mnMaxY = bounds[3];
  F.mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(F.mnMaxX - F.mnMinX);   // Frame.cc:118-119
  F.mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(F.mnMaxY - F.mnMinY);
  F.AssignFeaturesToGrid();
  int tot = 0;
  for (int q = 0; q < nq; q++) {
    off[q] = tot;
    const std::vector<std::size_t> v = F.GetFeaturesInArea(qx[q], qy[q], qr[q], qmin[q], qmax[q]);
    for (std::size_t j = 0; j < v.size(); j++, tot++)
      if (tot < cap) idx[tot] = (int)v[j];
  }
  off[nq] = tot;
  return tot;
}

// Frame::ComputeStereoFishEyeMatches.  tri_depth[l * n_r + r] = what TriangulateMatches returns for (left l, right r).
// Outputs: mvLeftToRightMatch [n_l], mvRightToLeftMatch [n_r], mvDepth [n_l].
void ref_stereo_fisheye_matches(const void* kps_l, const uint8_t* desc_l, int n_l, int mono_l, const void* kps_r,
                                const uint8_t* desc_r, int n_r, int mono_r, const float* level_sigma2, int nlevels,
                                const float* tri_depth, int* l2r, int* r2l, float* depth) {
  Frame F;
  KannalaBrandt8 c1, c2;
  c1.depth_table = tri_depth; c1.table_cols = n_r;
  F.mpCamera = &c1; F.mpCamera2 = &c2;
  F.mvKeys = to_keys(kps_l, n_l);
  F.mvKeysRight = to_keys(kps_r, n_r);
  for (int i = 0; i < n_l; i++) F.mvKeys[i].class_id = i;
  for (int i = 0; i < n_r; i++) F.mvKeysRight[i].class_id = i;
  F.mDescriptors = to_desc(desc_l, n_l);
  F.mDescriptorsRight = to_desc(desc_r, n_r);
  F.Nleft = n_l; F.Nright = n_r; F.N = n_l + n_r;
  F.monoLeft = mono_l; F.monoRight = mono_r;
  F.mvLevelSigma2.assign(level_sigma2, level_sigma2 + nlevels);
  F.ComputeStereoFishEyeMatches();
  for (int i = 0; i < n_l; i++) { l2r[i] = F.mvLeftToRightMatch[i]; depth[i] = F.mvDepth[i]; }
  for (int i = 0; i < n_r; i++) r2l[i] = F.mvRightToLeftMatch[i];
}

// Frame::ComputeStereoFromRGBD.  depth_img: w x h floats (the reference's imDepth after the depth-map factor).
void ref_rgbd_depth(const void* kps, const void* kps_un, int n, const float* depth_img, int w, int h, float mbf,
                    float* u_right, float* depth) {
  Frame F;
  F.N = n;
  F.mvKeys = to_keys(kps, n);
  F.mvKeysUn = to_keys(kps_un, n);
  F.mbf = mbf;
  // the shim's Mat is byte-typed: a float image is a view of 4*w bytes per row, at<float>(y, x) indexes it as floats
  cv::Mat im(h, 4 * w, CV_8UC1, (void*)depth_img);
  F.ComputeStereoFromRGBD(im);
  for (int i = 0; i < n; i++) { u_right[i] = F.mvuRight[i]; depth[i] = F.mvDepth[i]; }
}

// Frame::isInFrustum for n MapPoints (mf_min_dist / mf_max_dist = MapPoint::mfMinDistance / mfMaxDistance).  pose = Rcw[9] (row-major), tcw[3], Ow[3], Rwc[9]; cam / cam_r = fx, fy, cx, cy
// (pinhole).  rig != 0: Nleft != -1, trl = Rrl[9], trl[3] and tlr[3] = mTlr.translation().  Every tracking field starts
// at the value the oracle reports for "not written" (-1 / 0 / level -1).  Returns the number of points in view; for a
// rig also the composed right pose as the reference computes it (:1176-1180) in right_pose = R[9], t[3], twc[3].
int ref_is_in_frustum(const float* pos, const float* normal, const float* mf_min_dist, const float* mf_max_dist, int n,
                      const float* pose, const float* cam, const float* cam_r, float mbf, const float* bounds,
                      float cos_limit, float log_sf, int nlevels, int rig, const float* trl, const float* tlr,
                      uint8_t* in_view, float* proj_x, float* proj_y, float* proj_xr, float* depth, float* view_cos,
                      int8_t* level, uint8_t* in_view_r, float* proj_xr_r, float* proj_yr, float* depth_r,
                      float* view_cos_r, int8_t* level_r, float* right_pose) {
  Frame F;
  GeometricCamera c1, c2;
  c1.fx = cam[0]; c1.fy = cam[1]; c1.cx = cam[2]; c1.cy = cam[3];
  F.mpCamera = &c1;
  if (rig) {
    c2.fx = cam_r[0]; c2.fy = cam_r[1]; c2.cx = cam_r[2]; c2.cy = cam_r[3];
    F.mpCamera2 = &c2;
    F.Nleft = 1;
    F.mTrl = Sophus::SE3f(mat3(trl), vec3(trl + 9));
    F.mTlr = Sophus::SE3f(Eigen::Matrix3f(), vec3(tlr));
    const Eigen::Matrix3f Rr = F.mTrl.rotationMatrix() * mat3(pose);
    const Eigen::Vector3f tr = F.mTrl.rotationMatrix() * vec3(pose + 9) + F.mTrl.translation();
    const Eigen::Vector3f tw = mat3(pose + 15) * F.mTlr.translation() + vec3(pose + 12);
    for (int i = 0; i < 9; i++) right_pose[i] = Rr.m[i];
    for (int i = 0; i < 3; i++) { right_pose[9 + i] = tr.v[i]; right_pose[12 + i] = tw.v[i]; }
  }
  F.mRcw = mat3(pose); F.mtcw = vec3(pose + 9); F.mOw = vec3(pose + 12); F.mRwc = mat3(pose + 15);
  F.mbf = mbf;
  F.mnMinX = bounds[0]; F.mnMaxX = bounds[1]; F.mnMinY = bounds[2]; F.mnMaxY = bounds[3];
  F.mfLogScaleFactor = log_sf; F.mnScaleLevels = nlevels;
  int cnt = 0;
  for (int i = 0; i < n; i++) {
    MapPoint p;
    p.mWorldPos = vec3(pos + 3 * i); p.mNormalVector = vec3(normal + 3 * i);
    p.mfMinDistance = mf_min_dist[i]; p.mfMaxDistance = mf_max_dist[i];  // the gates use 0.8f * min, 1.2f * max
    p.mTrackProjX = -1; p.mTrackProjY = -1; p.mTrackProjXR = -1; p.mTrackDepth = 0; p.mTrackViewCos = 0;
    p.mnTrackScaleLevel = -1; p.mTrackProjYR = -1; p.mTrackDepthR = 0; p.mTrackViewCosR = 0; p.mnTrackScaleLevelR = -1;
    const bool r = F.isInFrustum(&p, cos_limit);
    cnt += r;
    in_view[i] = p.mbTrackInView; proj_x[i] = p.mTrackProjX; proj_y[i] = p.mTrackProjY; proj_xr[i] = p.mTrackProjXR;
    depth[i] = p.mTrackDepth; view_cos[i] = p.mTrackViewCos; level[i] = (int8_t)p.mnTrackScaleLevel;
    if (rig) {
      in_view_r[i] = p.mbTrackInViewR; proj_xr_r[i] = p.mTrackProjXR; proj_yr[i] = p.mTrackProjYR;
      depth_r[i] = p.mTrackDepthR; view_cos_r[i] = p.mTrackViewCosR; level_r[i] = (int8_t)p.mnTrackScaleLevelR;
    }
  }
  return cnt;
}

}  // extern "C"
