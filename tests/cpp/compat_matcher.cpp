// The ORBmatcher drop-in class (include/orbx_compat/ORBmatcher.h) driven through the reference's own method names
// and arguments, on stand-in Frame / KeyFrame / MapPoint structs that carry the members the reference bodies read
// (reference: include/Frame.h, include/KeyFrame.h, include/MapPoint.h) — the real classes need Eigen / Sophus / DBoW2,
// which this image does not have.
//   compat_matcher <in.bin> <out_prefix>
// in.bin is written by tests/test_gpu_compat_cpp.py::test_cpp_orbmatcher_class (arrays in the order read below).
#include <stdio.h>
#include <stdlib.h>

#include <math.h>

#include <map>
#include <set>
#include <tuple>
#include <vector>

#include "orbx_compat/ORBmatcher.h"

using namespace ORB_SLAM3;

// ---- minimal geometry stand-ins (Eigen::Vector3f / Sophus::SE3f surface used by ORBmatcher.cc:1686-1712) ----
struct Vec3 {
  float v[3];
  float operator()(int i) const { return v[i]; }
  Vec3 operator-(const Vec3& o) const { return Vec3{{v[0] - o.v[0], v[1] - o.v[1], v[2] - o.v[2]}}; }
  Vec3 operator/(float s) const { return Vec3{{v[0] / s, v[1] / s, v[2] / s}}; }
  float dot(const Vec3& o) const { return v[0] * o.v[0] + (v[1] * o.v[1] + v[2] * o.v[2]); }
  float norm() const { return sqrtf(dot(*this)); }
};
struct Vec2 {
  float v[2];
  float operator()(int i) const { return v[i]; }
};
struct Mat3 { float m[9]; };
struct SE3 {
  float R[9], t[3];
  SE3() : R{1, 0, 0, 0, 1, 0, 0, 0, 1}, t{0, 0, 0} {}
  SE3(const Mat3& r, const Vec3& tt) { memcpy(R, r.m, 36); memcpy(t, tt.v, 12); }
  Vec3 operator*(const Vec3& x) const {
    Vec3 o;
    for (int i = 0; i < 3; i++) o.v[i] = R[3 * i] * x.v[0] + R[3 * i + 1] * x.v[1] + R[3 * i + 2] * x.v[2] + t[i];
    return o;
  }
  SE3 inverse() const {
    SE3 o;
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) o.R[3 * i + j] = R[3 * j + i];
    for (int i = 0; i < 3; i++) o.t[i] = -(o.R[3 * i] * t[0] + o.R[3 * i + 1] * t[1] + o.R[3 * i + 2] * t[2]);
    return o;
  }
  Vec3 translation() const { return Vec3{{t[0], t[1], t[2]}}; }
};
struct Sim3 {  // Sophus::Sim3f surface used by ORBmatcher.cc:436, 1349, 1469, 1508, 1588 (identity here)
  Mat3 rotationMatrix() const { return Mat3{{1, 0, 0, 0, 1, 0, 0, 0, 1}}; }
  Vec3 translation() const { return Vec3{{0, 0, 0}}; }
  float scale() const { return 1.f; }
  Sim3 inverse() const { return Sim3(); }
  Vec3 operator*(const Vec3& x) const { return x; }
};
struct PinholeCam {  // Pinhole::project(Eigen::Vector3f), Pinhole.cpp:49-55
  float fx, fy, cx, cy;
  Vec2 project(const Vec3& p) const { return Vec2{{fx * p(0) / p(2) + cx, fy * p(1) / p(2) + cy}}; }
};

// ---- the three object types, reference member names ----
struct KeyFrame;
struct Frame;
struct MapPoint {
  int id = -1;
  bool mbTrackInView = false, mbTrackInViewR = false, bad = false;
  float mTrackProjX = 0, mTrackProjY = 0, mTrackProjXR = 0, mTrackProjYR = 0, mTrackViewCos = 0, mTrackViewCosR = 0,
        mTrackDepth = 0;
  int mnTrackScaleLevel = 0, mnTrackScaleLevelR = -1, nObs = 0;
  cv::Mat desc;
  Vec3 pos;
  Vec3 normal{{0, 0, 1}};
  float mfMinDistance = 0, mfMaxDistance = 0;
  std::map<KeyFrame*, int> obs;
  bool isBad() const { return bad; }
  int Observations() const { return nObs; }
  cv::Mat GetDescriptor() const { return desc; }
  Vec3 GetWorldPos() const { return pos; }
  Vec3 GetNormal() const { return normal; }
  float GetMinDistanceInvariance() const { return 0.8f * mfMinDistance; }
  float GetMaxDistanceInvariance() const { return 1.2f * mfMaxDistance; }
  template <class T> int PredictScale(const float& currentDist, T* pF) const {  // MapPoint.cc:514-546
    const float ratio = mfMaxDistance / currentDist;
    int nScale = (int)ceilf(logf(ratio) / pF->mfLogScaleFactor);
    if (nScale < 0) nScale = 0;
    else if (nScale >= pF->mnScaleLevels) nScale = pF->mnScaleLevels - 1;
    return nScale;
  }
  bool IsInKeyFrame(KeyFrame* pKF) const { return obs.count(pKF) != 0; }
  std::tuple<int, int> GetIndexInKeyFrame(KeyFrame* pKF) const {
    auto it = obs.find(pKF);
    return it == obs.end() ? std::tuple<int, int>(-1, -1) : std::tuple<int, int>(it->second, -1);
  }
  void AddObservation(KeyFrame* pKF, int idx) { obs[pKF] = idx; nObs++; }
  void Replace(MapPoint* pMP);
};

struct Frame {
  int N = 0, Nleft = -1;
  std::vector<cv::KeyPoint> mvKeys, mvKeysRight, mvKeysUn;
  cv::Mat mDescriptors;
  std::vector<float> mvuRight, mvScaleFactors;
  std::vector<MapPoint*> mvpMapPoints;
  std::vector<bool> mvbOutlier;
  std::vector<int> mvLeftToRightMatch, mvRightToLeftMatch;
  float mnMinX = 0, mnMaxX = 0, mnMinY = 0, mnMaxY = 0, mb = 0, mbf = 0;
  FeatureVectorMap mFeatVec;
  PinholeCam cam;
  PinholeCam* mpCamera = &cam;
  SE3 pose;
  float mfLogScaleFactor = logf(1.2f);
  int mnScaleLevels = 8;
  SE3 GetPose() const { return pose; }
};

struct KeyFrame {
  int N = 0;
  std::vector<cv::KeyPoint> mvKeysUn;
  cv::Mat mDescriptors;
  std::vector<float> mvuRight, mvScaleFactors, mvLevelSigma2;
  std::vector<MapPoint*> mps;
  FeatureVectorMap mFeatVec;
  float F12_to[9], ep_to[2];  // test geometry towards "the other keyframe"
  int NLeft = -1;
  std::vector<cv::KeyPoint> mvKeys, mvKeysRight;
  std::vector<float> mvInvLevelSigma2;
  float mnMinX = 0, mnMaxX = 0, mnMinY = 0, mnMaxY = 0, fx = 1, fy = 1, cx = 0, cy = 0, mbf = 0;
  float mfLogScaleFactor = logf(1.2f);
  int mnScaleLevels = 8;
  PinholeCam cam{1, 1, 0, 0};
  PinholeCam *mpCamera = &cam, *mpCamera2 = &cam;
  SE3 pose;
  std::vector<int> oplog;  // (kind, a, b): 1 AddObservation(idx, id), 2 Replace(self id, other id)
  SE3 GetPose() const { return pose; }
  SE3 GetRightPose() const { return pose; }
  Vec3 GetCameraCenter() const { return pose.inverse().translation(); }
  Vec3 GetRightCameraCenter() const { return pose.inverse().translation(); }
  std::vector<MapPoint*> GetMapPointMatches() const { return mps; }
  MapPoint* GetMapPoint(int i) const { return mps[i]; }
  std::set<MapPoint*> GetMapPoints() const {
    std::set<MapPoint*> s;
    for (MapPoint* p : mps) if (p && !p->isBad()) s.insert(p);
    return s;
  }
  void AddMapPoint(MapPoint* pMP, int idx) { mps[idx] = pMP; }
  bool IsInImage(const float& x, const float& y) const { return (x >= mnMinX && x < mnMaxX && y >= mnMinY && y < mnMaxY); }
};
inline void MapPoint::Replace(MapPoint* pMP) {  // what the matcher can observe of MapPoint.cc:226-280
  for (auto& ob : obs) {
    if (!pMP->IsInKeyFrame(ob.first)) { ob.first->mps[ob.second] = pMP; pMP->obs[ob.first] = ob.second; pMP->nObs++; }
    else ob.first->mps[ob.second] = nullptr;
  }
  bad = true;
}

// customisation point of ORBmatcherT::SearchForTriangulation (with the real classes: Eigen)
void orbx_epipolar_geometry(KeyFrame* pKF1, KeyFrame* /*pKF2*/, float F12[9], float ep[2]) {
  memcpy(F12, pKF1->F12_to, 36);
  memcpy(ep, pKF1->ep_to, 8);
}

typedef ORBmatcherT<Frame, KeyFrame, MapPoint> ORBmatcher;

// ---- I/O ----
static FILE* g_in;
template <class T> static std::vector<T> rd(size_t n) {
  std::vector<T> v(n);
  if (n && fread(v.data(), sizeof(T), n, g_in) != n) { fprintf(stderr, "short read\n"); exit(3); }
  return v;
}
static int rdi() { return rd<int>(1)[0]; }
static void dump(const char* prefix, const char* ext, const std::vector<int>& v) {
  char path[512];
  snprintf(path, sizeof path, "%s.%s", prefix, ext);
  FILE* f = fopen(path, "wb");
  if (!v.empty()) fwrite(v.data(), 4, v.size(), f);
  fclose(f);
}
static cv::Mat desc_mat(const std::vector<unsigned char>& d, int n) {
  cv::Mat m(n > 0 ? n : 1, 32, CV_8U);
  if (n) memcpy(m.data, d.data(), (size_t)n * 32);
  return m;
}
static FeatureVectorMap rd_fv(int n) {
  const int nn = rdi();
  std::vector<int> node = rd<int>(nn), off = rd<int>(nn + 1), idx = rd<int>(n);
  FeatureVectorMap fv;
  for (int a = 0; a < nn; a++)
    for (int i = off[a]; i < off[a + 1]; i++) fv[(unsigned)node[a]].push_back((unsigned)idx[i]);
  return fv;
}
// features of one camera set: n, keys, desc
static void rd_feats(int& n, std::vector<cv::KeyPoint>& k, std::vector<unsigned char>& d) {
  n = rdi();
  k = rd<cv::KeyPoint>(n);
  d = rd<unsigned char>((size_t)n * 32);
}
static std::vector<int> ids_of(const std::vector<MapPoint*>& v) {
  std::vector<int> o(v.size(), -1);
  for (size_t i = 0; i < v.size(); i++) o[i] = v[i] ? v[i]->id : -1;
  return o;
}
// MapPoints: M, then per-point arrays
static std::vector<MapPoint> rd_points() {
  const int M = rdi();
  std::vector<float> f = rd<float>((size_t)M * 10);  // projX, projY, projXR, projYR, viewCos, viewCosR, depth, pos[3]
  std::vector<int> iv = rd<int>((size_t)M * 6);      // inView, inViewR, level, levelR, bad, nObs
  std::vector<unsigned char> d = rd<unsigned char>((size_t)M * 32);
  std::vector<MapPoint> pts(M);
  for (int i = 0; i < M; i++) {
    MapPoint& p = pts[i];
    p.id = i;
    const float* a = &f[(size_t)i * 10];
    p.mTrackProjX = a[0]; p.mTrackProjY = a[1]; p.mTrackProjXR = a[2]; p.mTrackProjYR = a[3];
    p.mTrackViewCos = a[4]; p.mTrackViewCosR = a[5]; p.mTrackDepth = a[6];
    p.pos = Vec3{{a[7], a[8], a[9]}};
    const int* b = &iv[(size_t)i * 6];
    p.mbTrackInView = b[0] != 0; p.mbTrackInViewR = b[1] != 0; p.mnTrackScaleLevel = b[2]; p.mnTrackScaleLevelR = b[3];
    p.bad = b[4] != 0; p.nObs = b[5];
    p.desc = cv::Mat(1, 32, CV_8U);
    memcpy(p.desc.data, &d[(size_t)i * 32], 32);
  }
  return pts;
}
static void set_holders(std::vector<MapPoint*>& dst, const std::vector<int>& holder, std::vector<MapPoint>& pts) {
  dst.assign(holder.size(), nullptr);
  for (size_t j = 0; j < holder.size(); j++)
    if (holder[j] >= 0) dst[j] = &pts[holder[j]];
}

int main(int argc, char** argv) {
  if (argc < 3) return 2;
  g_in = fopen(argv[1], "rb");
  if (!g_in) return 3;
  const char* out = argv[2];
  std::vector<float> sf = rd<float>(8), sig = rd<float>(8), bounds = rd<float>(4);
  std::vector<int> rets;

  // ---- 1. SearchByProjection(Frame&, vector<MapPoint*>&, th) single camera ----
  {
    std::vector<MapPoint> pts = rd_points();
    Frame F;
    std::vector<unsigned char> d;
    rd_feats(F.N, F.mvKeysUn, d);
    F.mvKeys = F.mvKeysUn;
    F.mDescriptors = desc_mat(d, F.N);
    F.mvuRight = rd<float>(F.N);
    set_holders(F.mvpMapPoints, rd<int>(F.N), pts);
    F.mvScaleFactors = sf;
    F.mnMinX = bounds[0]; F.mnMaxX = bounds[1]; F.mnMinY = bounds[2]; F.mnMaxY = bounds[3];
    std::vector<MapPoint*> vp;
    for (auto& p : pts) vp.push_back(&p);
    ORBmatcher matcher(0.8f);
    rets.push_back(matcher.SearchByProjection(F, vp, 3.0f, true, 40.0f));
    dump(out, "proj_mono", ids_of(F.mvpMapPoints));
  }
  // ---- 2. the same on a two-camera (fisheye rig) frame ----
  {
    std::vector<MapPoint> pts = rd_points();
    Frame F;
    std::vector<unsigned char> dl, dr;
    int nl, nr;
    rd_feats(nl, F.mvKeys, dl);
    rd_feats(nr, F.mvKeysRight, dr);
    F.Nleft = nl; F.N = nl + nr;
    dl.insert(dl.end(), dr.begin(), dr.end());
    F.mDescriptors = desc_mat(dl, F.N);
    F.mvuRight.assign(F.N, -1.f);
    set_holders(F.mvpMapPoints, rd<int>(F.N), pts);
    F.mvLeftToRightMatch = rd<int>(nl);
    F.mvRightToLeftMatch = rd<int>(nr);
    F.mvScaleFactors = sf;
    F.mnMinX = bounds[0]; F.mnMaxX = bounds[1]; F.mnMinY = bounds[2]; F.mnMaxY = bounds[3];
    std::vector<MapPoint*> vp;
    for (auto& p : pts) vp.push_back(&p);
    ORBmatcher matcher(0.8f);
    rets.push_back(matcher.SearchByProjection(F, vp, 3.0f));
    dump(out, "proj_rig", ids_of(F.mvpMapPoints));
  }
  // ---- 3. SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, th, bMono) ----
  {
    std::vector<MapPoint> pts = rd_points();
    Frame C, Lf;
    std::vector<unsigned char> d;
    rd_feats(C.N, C.mvKeysUn, d);
    C.mvKeys = C.mvKeysUn;
    C.mDescriptors = desc_mat(d, C.N);
    C.mvuRight = rd<float>(C.N);
    set_holders(C.mvpMapPoints, rd<int>(C.N), pts);
    rd_feats(Lf.N, Lf.mvKeysUn, d);
    Lf.mvKeys = Lf.mvKeysUn;
    set_holders(Lf.mvpMapPoints, rd<int>(Lf.N), pts);
    std::vector<int> outl = rd<int>(Lf.N);
    Lf.mvbOutlier.assign(outl.begin(), outl.end());
    std::vector<float> cam = rd<float>(4), pc = rd<float>(12), pl = rd<float>(12), mbs = rd<float>(2);
    C.cam = PinholeCam{cam[0], cam[1], cam[2], cam[3]};
    memcpy(C.pose.R, pc.data(), 36); memcpy(C.pose.t, pc.data() + 9, 12);
    memcpy(Lf.pose.R, pl.data(), 36); memcpy(Lf.pose.t, pl.data() + 9, 12);
    C.mb = mbs[0]; C.mbf = mbs[1];
    C.mvScaleFactors = sf;
    C.mnMinX = bounds[0]; C.mnMaxX = bounds[1]; C.mnMinY = bounds[2]; C.mnMaxY = bounds[3];
    ORBmatcher matcher(0.9f, true);
    rets.push_back(matcher.SearchByProjection(C, Lf, 15.0f, true));
    dump(out, "proj_last", ids_of(C.mvpMapPoints));
  }
  // ---- 4./5./6. SearchByBoW (KF, Frame), on a rig frame, (KF, KF); SearchForTriangulation ----
  {
    std::vector<MapPoint> pts = rd_points();
    KeyFrame K1, K2;
    std::vector<unsigned char> d;
    for (KeyFrame* K : {&K1, &K2}) {
      rd_feats(K->N, K->mvKeysUn, d);
      K->mDescriptors = desc_mat(d, K->N);
      K->mFeatVec = rd_fv(K->N);
      set_holders(K->mps, rd<int>(K->N), pts);
      K->mvuRight = rd<float>(K->N);
      K->mvScaleFactors = sf; K->mvLevelSigma2 = sig;
    }
    std::vector<float> geo = rd<float>(11);
    memcpy(K1.F12_to, geo.data(), 36); memcpy(K1.ep_to, geo.data() + 9, 8);
    Frame F;  // the frame twin of K2
    F.N = K2.N; F.mvKeys = K2.mvKeysUn; F.mvKeysUn = K2.mvKeysUn; F.mDescriptors = K2.mDescriptors; F.mFeatVec = K2.mFeatVec;
    ORBmatcher matcher(0.7f, true);
    std::vector<MapPoint*> vm;
    rets.push_back(matcher.SearchByBoW(&K1, F, vm));
    dump(out, "bow_kf_f", ids_of(vm));
    rets.push_back(matcher.SearchByBoW(&K1, &K2, vm));
    dump(out, "bow_kf_kf", ids_of(vm));
    std::vector<std::pair<size_t, size_t> > pairs;
    rets.push_back(matcher.SearchForTriangulation(&K1, &K2, pairs, false, false));
    std::vector<int> flat;
    for (auto& pr : pairs) { flat.push_back((int)pr.first); flat.push_back((int)pr.second); }
    dump(out, "tri_pairs", flat);
    // rig frame: left|right features, FeatureVector over the concatenation
    Frame R;
    std::vector<unsigned char> dl, dr;
    int nl, nr;
    rd_feats(nl, R.mvKeys, dl);
    rd_feats(nr, R.mvKeysRight, dr);
    R.Nleft = nl; R.N = nl + nr;
    dl.insert(dl.end(), dr.begin(), dr.end());
    R.mDescriptors = desc_mat(dl, R.N);
    R.mFeatVec = rd_fv(R.N);
    rets.push_back(matcher.SearchByBoW(&K1, R, vm));
    dump(out, "bow_kf_rig", ids_of(vm));
  }
  // ---- 7. SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize) ----
  {
    Frame F1, F2;
    std::vector<unsigned char> d;
    rd_feats(F1.N, F1.mvKeysUn, d);
    F1.mDescriptors = desc_mat(d, F1.N);
    rd_feats(F2.N, F2.mvKeysUn, d);
    F2.mDescriptors = desc_mat(d, F2.N);
    for (Frame* F : {&F1, &F2}) { F->mnMinX = bounds[0]; F->mnMaxX = bounds[1]; F->mnMinY = bounds[2]; F->mnMaxY = bounds[3]; }
    std::vector<cv::Point2f> prev(F1.N);
    for (int i = 0; i < F1.N; i++) prev[i] = F1.mvKeysUn[i].pt;
    std::vector<int> m12;
    ORBmatcher matcher(0.9f, true);
    rets.push_back(matcher.SearchForInitialization(F1, F2, prev, m12, 100));
    dump(out, "init_m12", m12);
  }

  // ---- 8.-12. relocalisation / Sim3 projections x2 / Fuse / Fuse(Sim3) / SearchBySim3: identity poses, unit pinhole,
  //      MapPoint i at world (u_i, v_i, 1) -> the class's own projection code yields exactly (u_i, v_i) ----
  {
    auto make_kf = [&](KeyFrame& K, std::vector<MapPoint>& own) {
      std::vector<unsigned char> d;
      rd_feats(K.N, K.mvKeysUn, d);
      K.mvKeys = K.mvKeysUn;
      K.mDescriptors = desc_mat(d, K.N);
      std::vector<int> hf = rd<int>(K.N);  // bit0 holds a point, bit1 that point is bad
      K.mvuRight = rd<float>(K.N);
      K.mvScaleFactors = sf; K.mvLevelSigma2 = sig;
      K.mvInvLevelSigma2.resize(8);
      for (int l = 0; l < 8; l++) K.mvInvLevelSigma2[l] = 1.0f / sig[l];
      K.mnMinX = bounds[0]; K.mnMaxX = bounds[1]; K.mnMinY = bounds[2]; K.mnMaxY = bounds[3];
      own.resize(K.N);
      K.mps.assign(K.N, nullptr);
      for (int i = 0; i < K.N; i++) {
        own[i].id = -2 - i; own[i].nObs = 1; own[i].bad = (hf[i] & 2) != 0;
        own[i].obs[&K] = i;
        if (hf[i] & 1) K.mps[i] = &own[i];
      }
    };
    auto place = [&](std::vector<MapPoint>& pts, const std::vector<float>& uv, const std::vector<float>& mx,
                     const std::vector<float>& mn, const unsigned char* d) {
      for (size_t i = 0; i < pts.size(); i++) {
        const float u = uv[2 * i], v = uv[2 * i + 1];
        pts[i].pos = Vec3{{u, v, 1.f}};
        const float nrm = sqrtf(u * u + v * v + 1.f);
        pts[i].normal = Vec3{{u / nrm, v / nrm, 1.f / nrm}};
        pts[i].mfMaxDistance = mx[i]; pts[i].mfMinDistance = mn[i];
        if (d) { pts[i].desc = cv::Mat(1, 32, CV_8U); memcpy(pts[i].desc.data, d + i * 32, 32); }
      }
    };
    // queries: M points with flags (bit0 present, bit1 bad, bit2 already found), uv, ranges, descriptors, angles
    const int M = rdi();
    std::vector<int> qf = rd<int>(M);
    std::vector<float> uv = rd<float>((size_t)M * 2), mx = rd<float>(M), mn = rd<float>(M), ang = rd<float>(M);
    std::vector<unsigned char> qd = rd<unsigned char>((size_t)M * 32);
    auto fresh_points = [&](std::vector<MapPoint>& pts) {
      pts.assign(M, MapPoint());
      for (int i = 0; i < M; i++) { pts[i].id = i; pts[i].nObs = 1; pts[i].bad = (qf[i] & 2) != 0; }
      place(pts, uv, mx, mn, qd.data());
    };
    // 8. relocalisation: current frame = the keyframe data read below, keyframe = the query points
    {
      KeyFrame Kt; std::vector<MapPoint> own;
      make_kf(Kt, own);
      Frame C;
      C.N = Kt.N; C.mvKeysUn = Kt.mvKeysUn; C.mvKeys = Kt.mvKeysUn; C.mDescriptors = Kt.mDescriptors;
      C.mvuRight.assign(C.N, -1.f);
      C.mvpMapPoints = Kt.mps;       // non-null entries block their features
      C.mvScaleFactors = sf; C.cam = PinholeCam{1, 1, 0, 0};
      C.mnMinX = bounds[0]; C.mnMaxX = bounds[1]; C.mnMinY = bounds[2]; C.mnMaxY = bounds[3];
      std::vector<MapPoint> pts; fresh_points(pts);
      KeyFrame Kq;
      Kq.N = M; Kq.mvKeysUn.resize(M); Kq.mps.assign(M, nullptr);
      std::set<MapPoint*> found;
      for (int i = 0; i < M; i++) {
        Kq.mvKeysUn[i].angle = ang[i];
        if (qf[i] & 1) Kq.mps[i] = &pts[i];
        if (qf[i] & 4) found.insert(&pts[i]);
      }
      ORBmatcher matcher(0.9f, true);
      rets.push_back(matcher.SearchByProjection(C, &Kq, found, 10.0f, 100));
      dump(out, "reloc", ids_of(C.mvpMapPoints));
      // 9. Sim3 projections (both variants) on the same keyframe: vpMatched starts from the keyframe's own points
      for (int variant = 0; variant < 2; variant++) {
        std::vector<MapPoint> p2; fresh_points(p2);
        std::vector<MapPoint*> vp, vpMatched = Kt.mps;
        for (auto& p : p2) vp.push_back(&p);
        for (int i = 0; i < M; i++)
          if ((qf[i] & 4) && i < Kt.N) vpMatched[i] = &p2[i];   // "already found": held by feature i on entry
        std::vector<KeyFrame*> vpKFs(M, &Kt), vpMatchedKF(Kt.N, nullptr);
        Sim3 Scw;
        ORBmatcher m2(0.75f, true);
        rets.push_back(variant == 0 ? m2.SearchByProjection(&Kt, Scw, vp, vpMatched, 8, 1.0f)
                                    : m2.SearchByProjection(&Kt, Scw, vp, vpKFs, vpMatched, vpMatchedKF, 8, 1.0f));
        dump(out, variant == 0 ? "sim3_a" : "sim3_b", ids_of(vpMatched));
      }
      // 10. Fuse(pKF, vpMapPoints, th): which feature every point was fused into
      {
        KeyFrame Kf = Kt;
        std::vector<MapPoint> ownf(own);
        for (int i = 0; i < Kf.N; i++) { ownf[i].obs.clear(); ownf[i].obs[&Kf] = i; if (Kf.mps[i]) Kf.mps[i] = &ownf[i]; }
        Kf.mbf = 40.0f;
        std::vector<MapPoint> p3; fresh_points(p3);
        std::vector<MapPoint*> vp(M, nullptr);
        for (int i = 0; i < M; i++) if (qf[i] & 1) vp[i] = &p3[i];
        ORBmatcher m3(0.6f, true);
        rets.push_back(m3.Fuse(&Kf, vp, 3.0f));
        std::vector<int> where(M, -1);
        for (int j = 0; j < Kf.N; j++) if (Kf.mps[j] && Kf.mps[j]->id >= 0) where[Kf.mps[j]->id] = j;
        for (int i = 0; i < M; i++)   // a point that gave way to the held one (Replace) is found through its bad flag
          if (where[i] < 0 && p3[i].bad && !(qf[i] & 2)) where[i] = -3;
        dump(out, "fuse", where);
      }
      // 11. Fuse(pKF, Scw, vpPoints, th, vpReplacePoint)
      {
        KeyFrame Kf = Kt;
        std::vector<MapPoint> ownf(own);
        for (int i = 0; i < Kf.N; i++) { ownf[i].obs.clear(); ownf[i].obs[&Kf] = i; if (Kf.mps[i]) Kf.mps[i] = &ownf[i]; }
        std::vector<MapPoint> p4; fresh_points(p4);
        std::vector<MapPoint*> vp, repl(M, nullptr);
        for (auto& p : p4) vp.push_back(&p);
        Sim3 Scw;
        ORBmatcher m4(0.6f, true);
        rets.push_back(m4.Fuse(&Kf, Scw, vp, 4.0f, repl));
        std::vector<int> where(M, -1);
        for (int j = 0; j < Kf.N; j++) if (Kf.mps[j] && Kf.mps[j]->id >= 0) where[Kf.mps[j]->id] = j;
        for (int i = 0; i < M; i++) if (repl[i]) where[i] = -2 - repl[i]->id;   // held point's feature index
        dump(out, "fuse_sim3", where);
      }
    }
  }

  // ---- 12. SearchBySim3(pKF1, pKF2, vpMatches12, S12, th): every feature of both keyframes holds a point ----
  {
    KeyFrame K1, K2;
    std::vector<MapPoint> p1, p2;
    for (int side = 0; side < 2; side++) {
      KeyFrame& K = side == 0 ? K1 : K2;
      std::vector<MapPoint>& own = side == 0 ? p1 : p2;
      std::vector<unsigned char> d;
      rd_feats(K.N, K.mvKeysUn, d);
      K.mDescriptors = desc_mat(d, K.N);
      std::vector<int> hf = rd<int>(K.N);
      std::vector<float> uv = rd<float>((size_t)K.N * 2), mx = rd<float>(K.N), mn = rd<float>(K.N);
      K.mvScaleFactors = sf;
      K.mnMinX = bounds[0]; K.mnMaxX = bounds[1]; K.mnMinY = bounds[2]; K.mnMaxY = bounds[3];
      own.resize(K.N);
      K.mps.assign(K.N, nullptr);
      for (int i = 0; i < K.N; i++) {
        own[i].id = i; own[i].nObs = 1; own[i].bad = (hf[i] & 2) != 0;
        own[i].pos = Vec3{{uv[2 * i], uv[2 * i + 1], 1.f}};
        own[i].mfMaxDistance = mx[i]; own[i].mfMinDistance = mn[i];
        own[i].desc = cv::Mat(1, 32, CV_8U);
        memcpy(own[i].desc.data, &d[(size_t)i * 32], 32);
        if (hf[i] & 1) K.mps[i] = &own[i];
      }
    }
    for (int j = 0; j < K2.N; j++) p2[j].obs[&K2] = j;
    std::vector<int> init12 = rd<int>(K1.N);
    std::vector<MapPoint*> vpMatches12(K1.N, nullptr);
    for (int i = 0; i < K1.N; i++) if (init12[i] >= 0) vpMatches12[i] = &p2[init12[i]];
    Sim3 S12;
    ORBmatcher matcher(0.75f, true);
    rets.push_back(matcher.SearchBySim3(&K1, &K2, vpMatches12, S12, 7.5f));
    dump(out, "sim3_search", ids_of(vpMatches12));
  }
  fclose(g_in);
  for (size_t i = 0; i < rets.size(); i++) printf("%d ", rets[i]);
  printf("\n");
  for (int r : rets)
    if (r < 0) return 1;
  return 0;
}
