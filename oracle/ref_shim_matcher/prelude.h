// prelude.h — force-included (-include) in front of /root/reference/src/ORBmatcher.cc so that the reference's OWN
// matcher source compiles UNMODIFIED in an image without Eigen, Sophus, DBoW2, boost or OpenCV headers
// (oracle/Makefile target `ref_matcher` -> oracle/_ref/libref_orbmatcher.so).  TEST INFRASTRUCTURE ONLY.
//
// How: the reference's include/ORBmatcher.h is used as it is; the headers it pulls in are neutralised by defining
// their include guards here (FRAME_H, KEYFRAME_H, MAPPOINT_H, __D_T_FEATURE_VECTOR__) after declaring stand-ins that
// carry exactly the members ORBmatcher.cc touches:
//   Eigen::Vector2f / Vector3f / Matrix3f, Sophus::SE3f / Sim3f   the handful of operations the file uses (:427-646,
//                                                                 907-943, 1148-1674, 1686-1930), plain fp32 loops
//   DBoW2::FeatureVector                                          std::map<NodeId, std::vector<unsigned>> (FeatureVector.h:24)
//   ORB_SLAM3::GeometricCamera / Pinhole                          project + epipolarConstrain (Pinhole.cpp:49-55, 107-129)
//   ORB_SLAM3::MapPoint / KeyFrame / Frame                        data members under the reference's names; the grid
//                                                                 functions (AssignFeaturesToGrid / GetFeaturesInArea /
//                                                                 PosInGrid, Frame.cc:385-416, 657-735; KeyFrame.cc:704-748)
//                                                                 are the reference's own text (grid_ref.cpp)
// The OpenCV types come from the shim that already serves ORBextractor.cc (oracle/ref_shim/opencv2).
// What this pins: every loop, gate, tie rule, `continue`, histogram and index write of ORBmatcher.cc runs as the
// reference wrote and compiled it (-O3 -march=x86-64-v3 -std=c++14); the stand-ins only hold data.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <list>
#include <map>
#include <mutex>
#include <set>
#include <tuple>
#include <vector>

#include <opencv2/core/core.hpp>
#include <opencv2/features2d/features2d.hpp>

#define EIGEN_MAKE_ALIGNED_OPERATOR_NEW

namespace cv {
template <typename T> struct Point3_ {
  T x, y, z;
  Point3_() : x(0), y(0), z(0) {}
  Point3_(T _x, T _y, T _z) : x(_x), y(_y), z(_z) {}
};
typedef Point3_<float> Point3f;
}  // namespace cv

namespace cv {
// cv::DMatch and the brute-force Hamming matcher of Frame.h:337 (cv::BFMatcher(cv::NORM_HAMMING), crossCheck off).
// knnMatch: per query row the k nearest train rows by Hamming distance, nearest first; equal distances keep the lower
// train index first (OpenCV's batch distance + partial sort walks train rows in order and replaces on `<` only).
struct DMatch {
  int queryIdx = -1, trainIdx = -1, imgIdx = -1;
  float distance = 0;
};
enum { NORM_HAMMING = 6 };
class BFMatcher {
 public:
  explicit BFMatcher(int /*normType*/ = NORM_HAMMING, bool /*crossCheck*/ = false) {}
  void knnMatch(const Mat& query, const Mat& train, std::vector<std::vector<DMatch> >& matches, int k) const {
    matches.assign(query.rows, std::vector<DMatch>());
    for (int q = 0; q < query.rows; q++) {
      std::vector<DMatch>& out = matches[q];
      for (int t = 0; t < train.rows; t++) {
        int d = 0;
        const uchar *a = query.ptr(q), *b = train.ptr(t);
        for (int i = 0; i < query.cols; i++) d += __builtin_popcount((unsigned)(a[i] ^ b[i]));
        DMatch m; m.queryIdx = q; m.trainIdx = t; m.imgIdx = 0; m.distance = (float)d;
        std::size_t pos = out.size();
        while (pos > 0 && m.distance < out[pos - 1].distance) pos--;
        if ((int)pos < k) { out.insert(out.begin() + pos, m); if ((int)out.size() > k) out.pop_back(); }
      }
    }
  }
};
}  // namespace cv

namespace Eigen {
struct Vector2f {
  float v[2];
  Vector2f() : v{0, 0} {}
  Vector2f(float a, float b) : v{a, b} {}
  float operator()(int i) const { return v[i]; }
  float& operator()(int i) { return v[i]; }
  float operator[](int i) const { return v[i]; }
  float& operator[](int i) { return v[i]; }
};
struct Vector3f {
  float v[3];
  Vector3f() : v{0, 0, 0} {}
  Vector3f(float a, float b, float c) : v{a, b, c} {}
  float operator()(int i) const { return v[i]; }
  float& operator()(int i) { return v[i]; }
  float operator[](int i) const { return v[i]; }
  float& operator[](int i) { return v[i]; }
  Vector3f operator-(const Vector3f& o) const { return Vector3f(v[0] - o.v[0], v[1] - o.v[1], v[2] - o.v[2]); }
  Vector3f operator+(const Vector3f& o) const { return Vector3f(v[0] + o.v[0], v[1] + o.v[1], v[2] + o.v[2]); }
  Vector3f operator/(float s) const { return Vector3f(v[0] / s, v[1] / s, v[2] / s); }
  Vector3f operator*(float s) const { return Vector3f(v[0] * s, v[1] * s, v[2] * s); }
  Vector3f operator-() const { return Vector3f(-v[0], -v[1], -v[2]); }
  float dot(const Vector3f& o) const { return v[0] * o.v[0] + (v[1] * o.v[1] + v[2] * o.v[2]); }
  float norm() const { return std::sqrt(dot(*this)); }
};
struct Matrix3f {
  float m[9];  // row-major
  Matrix3f() : m{1, 0, 0, 0, 1, 0, 0, 0, 1} {}
  float operator()(int r, int c) const { return m[3 * r + c]; }
  float& operator()(int r, int c) { return m[3 * r + c]; }
  Vector3f operator*(const Vector3f& x) const {
    Vector3f o;
    for (int i = 0; i < 3; i++) o.v[i] = m[3 * i] * x.v[0] + (m[3 * i + 1] * x.v[1] + m[3 * i + 2] * x.v[2]);
    return o;
  }
  Matrix3f operator*(const Matrix3f& b) const {
    Matrix3f o;
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) o.m[3 * i + j] = m[3 * i] * b.m[j] + (m[3 * i + 1] * b.m[3 + j] + m[3 * i + 2] * b.m[6 + j]);
    return o;
  }
  Matrix3f transpose() const {
    Matrix3f o;
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) o.m[3 * i + j] = m[3 * j + i];
    return o;
  }
};
// Frame.cc:521,524 spell the 3-vector Eigen::Matrix<float,3,1>
template <class T, int R, int C> struct MatrixSel;
template <> struct MatrixSel<float, 3, 1> { typedef Vector3f type; };
template <> struct MatrixSel<float, 3, 3> { typedef Matrix3f type; };
template <class T, int R, int C> using Matrix = typename MatrixSel<T, R, C>::type;
}  // namespace Eigen

namespace Sophus {
template <class T> struct SE3 {
  Eigen::Matrix3f R;
  Eigen::Vector3f t;
  SE3() {}
  SE3(const Eigen::Matrix3f& r, const Eigen::Vector3f& tt) : R(r), t(tt) {}
  Eigen::Matrix3f rotationMatrix() const { return R; }
  Eigen::Vector3f translation() const { return t; }
  SE3 inverse() const { const Eigen::Matrix3f Rt = R.transpose(); return SE3(Rt, -(Rt * t)); }
  Eigen::Vector3f operator*(const Eigen::Vector3f& x) const { return R * x + t; }
  SE3 operator*(const SE3& o) const { return SE3(R * o.R, R * o.t + t); }
};
typedef SE3<float> SE3f;
template <class T> struct Sim3 {
  Eigen::Matrix3f R;
  Eigen::Vector3f t;
  float s = 1.f;
  Sim3() {}
  Sim3(float sc, const Eigen::Matrix3f& r, const Eigen::Vector3f& tt) : R(r), t(tt), s(sc) {}
  Eigen::Matrix3f rotationMatrix() const { return R; }
  Eigen::Vector3f translation() const { return t; }
  float scale() const { return s; }
  Sim3 inverse() const { const Eigen::Matrix3f Rt = R.transpose(); return Sim3(1.f / s, Rt, -((Rt * t) * (1.f / s))); }
  Eigen::Vector3f operator*(const Eigen::Vector3f& x) const { return (R * x) * s + t; }
};
typedef Sim3<float> Sim3f;
}  // namespace Sophus

namespace DBoW2 {
typedef unsigned int NodeId;
class FeatureVector : public std::map<NodeId, std::vector<unsigned int> > {};
}  // namespace DBoW2
#define __D_T_FEATURE_VECTOR__

namespace ORB_SLAM3 {

class KeyFrame;
class Frame;

// what Frame::ComputeStereoMatches reads of an extractor (ORBextractor.h:87): the pyramid it left behind
class ORBextractor {
 public:
  std::vector<cv::Mat> mvImagePyramid;
};

// GeometricCamera surface used by ORBmatcher.cc (:465, :920, :1072, :1210, :1713).  Defined out of line in
// matcher_ref_api.cpp (compiled with -ffp-contract=off: camera arithmetic is not part of ORBmatcher.cc).
class GeometricCamera {
 public:
  float fx = 1, fy = 1, cx = 0, cy = 0;
  virtual ~GeometricCamera() {}
  virtual Eigen::Vector2f project(const Eigen::Vector3f& v3D);
  virtual bool epipolarConstrain(GeometricCamera* pCamera2, const cv::KeyPoint& kp1, const cv::KeyPoint& kp2,
                                 const Eigen::Matrix3f& R12, const Eigen::Vector3f& t12, const float sigmaLevel,
                                 const float unc);
  // K1^-T [t12]x R12 K2^-1, row-major (Pinhole.cpp:109-112); exported so the tests hand the oracle the same matrix
  static void fundamental(const GeometricCamera& c1, const GeometricCamera& c2, const Eigen::Matrix3f& R12,
                          const Eigen::Vector3f& t12, float F12[9]);
};

// std::mutex that survives the copies the test harness makes of its stand-in objects (each copy gets a fresh mutex)
struct CopyableMutex : std::mutex {
  CopyableMutex() {}
  CopyableMutex(const CopyableMutex&) : std::mutex() {}
  CopyableMutex& operator=(const CopyableMutex&) { return *this; }
};

class MapPoint {
 public:
  int id = -1;
  // tracking fields written by Frame::isInFrustum (MapPoint.h:148-157)
  float mTrackProjX = 0, mTrackProjY = 0, mTrackDepth = 0, mTrackDepthR = 0, mTrackProjXR = 0, mTrackProjYR = 0;
  bool mbTrackInView = false, mbTrackInViewR = false;
  int mnTrackScaleLevel = 0, mnTrackScaleLevelR = -1;
  float mTrackViewCos = 0, mTrackViewCosR = 0;
  // state
  bool mbBad = false;
  int nObs = 0;
  cv::Mat mDescriptor;
  Eigen::Vector3f mWorldPos, mNormalVector;
  float mfMinDistance = 0, mfMaxDistance = 0;
  std::map<KeyFrame*, std::tuple<int, int> > mObservations;
  // operations log of the map-mutating calls (Fuse): (kind, keyframe feature index, other MapPoint id)
  struct Op { int kind, idx, other; };
  static std::vector<Op>* oplog;

  bool isBad() { return mbBad; }
  int Observations() { return nObs; }
  cv::Mat GetDescriptor() { return mDescriptor.clone(); }
  Eigen::Vector3f GetWorldPos() { return mWorldPos; }
  Eigen::Vector3f GetNormal() { return mNormalVector; }
  // the reference's own text, cut out of MapPoint.cc at build time (mappoint_ref.cpp)
  float GetMinDistanceInvariance();                                   // MapPoint.cc:498-502
  float GetMaxDistanceInvariance();                                   // MapPoint.cc:504-508
  int PredictScale(const float& currentDist, KeyFrame* pKF);          // MapPoint.cc:510-525
  int PredictScale(const float& currentDist, Frame* pF);              // MapPoint.cc:527-542
  void ComputeDistinctiveDescriptors();                               // MapPoint.cc:328-402
  CopyableMutex mMutexFeatures, mMutexPos;
  bool IsInKeyFrame(KeyFrame* pKF) { return mObservations.count(pKF) != 0; }
  std::tuple<int, int> GetIndexInKeyFrame(KeyFrame* pKF) {
    auto it = mObservations.find(pKF);
    return it == mObservations.end() ? std::tuple<int, int>(-1, -1) : it->second;
  }
  void AddObservation(KeyFrame* pKF, int idx);
  void Replace(MapPoint* pMP);
};

// What Frame::ComputeStereoFishEyeMatches (Frame.cc:1126-1166) uses of KannalaBrandt8: TriangulateMatches.  The real one
// needs Eigen's JacobiSVD (KannalaBrandt8.cpp:306-375; the product's version is on the tolerance contract),
// so this stand-in answers from a table the test supplies: depth[class_id of kp1][class_id of kp2], p3D = (0, 0, depth).
class KannalaBrandt8 : public GeometricCamera {
 public:
  // project / unproject / unprojectEig are the reference's own text (KannalaBrandt8.cpp:67-85, 111-146), camera_ref.cpp
  std::vector<float> mvParameters;   // fx, fy, cx, cy, k0..k3 (GeometricCamera.h:104)
  const float precision = 1e-6;      // KannalaBrandt8.h:42-47
  Eigen::Vector2f project(const Eigen::Vector3f& v3D);
  cv::Point3f unproject(const cv::Point2f& p2D);
  Eigen::Vector3f unprojectEig(const cv::Point2f& p2D);
  const float* depth_table = nullptr;
  int table_cols = 0;
  float TriangulateMatches(GeometricCamera* /*pCamera2*/, const cv::KeyPoint& kp1, const cv::KeyPoint& kp2,
                           const Eigen::Matrix3f& /*R12*/, const Eigen::Vector3f& /*t12*/, const float /*sigmaLevel*/,
                           const float /*unc*/, Eigen::Vector3f& p3D) {
    const float d = depth_table[(std::size_t)kp1.class_id * table_cols + kp2.class_id];
    p3D = Eigen::Vector3f(0.f, 0.f, d);
    return d;
  }
};

// Pinhole::project(const Eigen::Vector3f&) is the reference's own text (Pinhole.cpp:49-55), camera_ref.cpp
class Pinhole : public GeometricCamera {
 public:
  std::vector<float> mvParameters;   // fx, fy, cx, cy
  Eigen::Vector2f project(const Eigen::Vector3f& v3D);
};

// 64 x 48 feature grid (Frame.h:44-45, 250-252, 344; KeyFrame.h:320-321, 458, 518).  AssignFeaturesToGrid, PosInGrid and
// both GetFeaturesInArea are the REFERENCE's own text: grid_ref.cpp includes them, cut verbatim out of Frame.cc:385-416,
// 657-735 and KeyFrame.cc:704-748 at build time (oracle/extract_ref_functions.py).
#define FRAME_GRID_ROWS 48
#define FRAME_GRID_COLS 64

class Frame {
 public:
  int N = 0, Nleft = -1, Nright = -1;
  std::vector<cv::KeyPoint> mvKeys, mvKeysRight, mvKeysUn;
  cv::Mat mDescriptors;
  std::vector<float> mvuRight, mvScaleFactors;
  std::vector<MapPoint*> mvpMapPoints;
  std::vector<bool> mvbOutlier;
  std::vector<int> mvLeftToRightMatch, mvRightToLeftMatch;
  DBoW2::FeatureVector mFeatVec;
  float mnMinX = 0, mnMaxX = 0, mnMinY = 0, mnMaxY = 0, mb = 0, mbf = 0;
  float mfLogScaleFactor = 0;
  int mnScaleLevels = 8;
  GeometricCamera *mpCamera = nullptr, *mpCamera2 = nullptr;
  Sophus::SE3f mTcw, mTrl;
  float mfGridElementWidthInv = 1, mfGridElementHeightInv = 1;  // static in the reference (Frame.h:250-251)
  std::vector<std::size_t> mGrid[FRAME_GRID_COLS][FRAME_GRID_ROWS];
  std::vector<std::size_t> mGridRight[FRAME_GRID_COLS][FRAME_GRID_ROWS];
  bool PosInGrid(const cv::KeyPoint& kp, int& posX, int& posY);

  // members and functions of the Frame.cc functions compiled by `make ref_frame` (frame_ref_api.cpp): stereo matching
  // (Frame.cc:811-1005) and the frustum test (:512-586, :1168-1242)
  ORBextractor *mpORBextractorLeft = nullptr, *mpORBextractorRight = nullptr;
  cv::Mat mDescriptorsRight;
  std::vector<float> mvInvScaleFactors, mvDepth;
  Eigen::Matrix3f mRcw, mRwc;
  Eigen::Vector3f mtcw, mOw;
  Sophus::SE3f mTlr;
  int monoLeft = -1, monoRight = -1, mnCloseMPs = 0;
  std::vector<Eigen::Vector3f> mvStereo3Dpoints;
  std::vector<float> mvLevelSigma2;
  Eigen::Matrix3f mRlr;
  Eigen::Vector3f mtlr;
  static cv::BFMatcher BFmatcher;
  void ComputeStereoFishEyeMatches();
  void ComputeStereoMatches();
  void ComputeStereoFromRGBD(const cv::Mat& imDepth);
  bool isInFrustum(MapPoint* pMP, float viewingCosLimit);
  bool isInFrustumChecks(MapPoint* pMP, float viewingCosLimit, bool bRight = false);

  Sophus::SE3f GetPose() const { return mTcw; }
  Sophus::SE3f GetRelativePoseTrl() { return mTrl; }
  void AssignFeaturesToGrid();
  std::vector<std::size_t> GetFeaturesInArea(const float& x, const float& y, const float& r, const int minLevel = -1,
                                             const int maxLevel = -1, const bool bRight = false) const;
};

class KeyFrame {
 public:
  int N = 0, NLeft = -1;
  std::vector<cv::KeyPoint> mvKeys, mvKeysRight, mvKeysUn;
  cv::Mat mDescriptors;
  std::vector<float> mvuRight, mvScaleFactors, mvLevelSigma2, mvInvLevelSigma2;
  std::vector<MapPoint*> mvpMapPoints;
  DBoW2::FeatureVector mFeatVec;
  float mnMinX = 0, mnMaxX = 0, mnMinY = 0, mnMaxY = 0, fx = 1, fy = 1, cx = 0, cy = 0, mbf = 0;
  float mfLogScaleFactor = 0;
  int mnScaleLevels = 8;
  GeometricCamera *mpCamera = nullptr, *mpCamera2 = nullptr;
  Sophus::SE3f mTcw, mTrl;
  int mnGridCols = FRAME_GRID_COLS, mnGridRows = FRAME_GRID_ROWS;
  float mfGridElementWidthInv = 1, mfGridElementHeightInv = 1;
  std::vector<std::vector<std::vector<std::size_t> > > mGrid, mGridRight;

  bool mbBad = false;
  bool isBad() { return mbBad; }
  Sophus::SE3f GetPose() { return mTcw; }
  Sophus::SE3f GetPoseInverse() { return mTcw.inverse(); }
  Eigen::Vector3f GetCameraCenter() { return mTcw.inverse().translation(); }
  Sophus::SE3f GetRightPose() { return mTrl * mTcw; }
  Sophus::SE3f GetRightPoseInverse() { return (mTrl * mTcw).inverse(); }
  Eigen::Vector3f GetRightCameraCenter() { return (mTrl * mTcw).inverse().translation(); }
  std::vector<MapPoint*> GetMapPointMatches() { return mvpMapPoints; }
  MapPoint* GetMapPoint(const std::size_t& idx) { return mvpMapPoints[idx]; }
  std::set<MapPoint*> GetMapPoints() {
    std::set<MapPoint*> s;
    for (MapPoint* p : mvpMapPoints) if (p && !p->isBad()) s.insert(p);
    return s;
  }
  void AddMapPoint(MapPoint* pMP, const std::size_t& idx);
  bool IsInImage(const float& x, const float& y) const { return (x >= mnMinX && x < mnMaxX && y >= mnMinY && y < mnMaxY); }
  void AssignFeaturesToGrid();
  std::vector<std::size_t> GetFeaturesInArea(const float& x, const float& y, const float& r, const bool bRight = false) const;
};

}  // namespace ORB_SLAM3

// the reference's own headers of these three classes are skipped: their include guards are taken
#define MAPPOINT_H
#define KEYFRAME_H
#define FRAME_H

// include/ORBmatcher.h:76,84,87 write `pair` / `vector` unqualified: the reference gets the names from the global
// `using namespace std;` of Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:36, reached through KeyFrame.h -> ORBVocabulary.h
#ifndef ORBX_PRELUDE_NO_USING_STD
using namespace std;
#endif
