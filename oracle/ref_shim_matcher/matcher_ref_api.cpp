// matcher_ref_api.cpp — C entry points of oracle/_ref/libref_orbmatcher.so: build the stand-in Frame / KeyFrame /
// MapPoint objects (prelude.h) from flat arrays and call the REFERENCE's own ORBmatcher methods, compiled unmodified
// from /root/reference/src/ORBmatcher.cc.  TEST INFRASTRUCTURE ONLY (tests/test_ref_matcher.py compares these
// results with the oracle's restatements on the same inputs; that is what pins the matcher rows).
// Also holds the out-of-line members of the stand-ins that are NOT reference text (pinhole camera arithmetic, the
// map-mutation log); the grid functions and MapPoint::PredictScale / Get*DistanceInvariance /
// ComputeDistinctiveDescriptors are the reference's own text (grid_ref.cpp, mappoint_ref.cpp).  Compiled with
// -ffp-contract=off (camera arithmetic is not part of ORBmatcher.cc; the oracle evaluates it uncontracted too).
#include "prelude.h"

#include "ORBmatcher.h"

namespace ORB_SLAM3 {

std::vector<MapPoint::Op>* MapPoint::oplog = nullptr;

Eigen::Vector2f GeometricCamera::project(const Eigen::Vector3f& v3D) {  // Pinhole.cpp:49-55
  Eigen::Vector2f res;
  res(0) = fx * v3D(0) / v3D(2) + cx;
  res(1) = fy * v3D(1) / v3D(2) + cy;
  return res;
}

void GeometricCamera::fundamental(const GeometricCamera& c1, const GeometricCamera& c2, const Eigen::Matrix3f& R12,
                                  const Eigen::Vector3f& t12, float F12[9]) {
  // K1^-T * [t12]x * R12 * K2^-1 in double, rounded once: the matrix is an INPUT of the comparison (the oracle and
  // the device take F12 from the host), so only its use below has to follow Pinhole.cpp
  const double k1it[9] = {1.0 / c1.fx, 0, 0, 0, 1.0 / c1.fy, 0, -(double)c1.cx / c1.fx, -(double)c1.cy / c1.fy, 1};
  const double k2i[9] = {1.0 / c2.fx, 0, -(double)c2.cx / c2.fx, 0, 1.0 / c2.fy, -(double)c2.cy / c2.fy, 0, 0, 1};
  const double tx[9] = {0, -(double)t12(2), (double)t12(1), (double)t12(2), 0, -(double)t12(0), -(double)t12(1), (double)t12(0), 0};
  double a[9], b[9], c[9];
  auto mul = [](const double* x, const double* y, double* o) {
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) o[3 * i + j] = x[3 * i] * y[j] + x[3 * i + 1] * y[3 + j] + x[3 * i + 2] * y[6 + j];
  };
  double r[9];
  for (int i = 0; i < 9; i++) r[i] = R12.m[i];
  mul(k1it, tx, a); mul(a, r, b); mul(b, k2i, c);
  for (int i = 0; i < 9; i++) F12[i] = (float)c[i];
}

bool GeometricCamera::epipolarConstrain(GeometricCamera* pCamera2, const cv::KeyPoint& kp1, const cv::KeyPoint& kp2,
                                        const Eigen::Matrix3f& R12, const Eigen::Vector3f& t12, const float /*sigmaLevel*/,
                                        const float unc) {
  float F[9];
  fundamental(*this, *pCamera2, R12, t12, F);
  // Pinhole.cpp:114-129
  const float a = kp1.pt.x * F[0] + kp1.pt.y * F[3] + F[6];
  const float b = kp1.pt.x * F[1] + kp1.pt.y * F[4] + F[7];
  const float c = kp1.pt.x * F[2] + kp1.pt.y * F[5] + F[8];
  const float num = a * kp2.pt.x + b * kp2.pt.y + c;
  const float den = a * a + b * b;
  if (den == 0) return false;
  const float dsqr = num * num / den;
  return dsqr < 3.84 * unc;
}

void MapPoint::AddObservation(KeyFrame* pKF, int idx) {
  mObservations[pKF] = std::tuple<int, int>(idx, -1);
  nObs++;
  if (oplog) oplog->push_back(Op{1, idx, id});
}
void MapPoint::Replace(MapPoint* pMP) {  // MapPoint.cc:226-280, as far as the matcher can observe it
  if (oplog) oplog->push_back(Op{2, id, pMP ? pMP->id : -1});
  for (auto& ob : mObservations) {
    KeyFrame* pKF = ob.first;
    const int idx = std::get<0>(ob.second);
    if (!pMP->IsInKeyFrame(pKF)) { pKF->mvpMapPoints[idx] = pMP; pMP->mObservations[pKF] = ob.second; pMP->nObs++; }
    else pKF->mvpMapPoints[idx] = static_cast<MapPoint*>(NULL);
  }
  mbBad = true;
}
void KeyFrame::AddMapPoint(MapPoint* pMP, const std::size_t& idx) {
  mvpMapPoints[idx] = pMP;
  if (MapPoint::oplog) MapPoint::oplog->push_back(MapPoint::Op{3, (int)idx, pMP ? pMP->id : -1});
}

// The grid of a KeyFrame is the grid of the Frame it is made from: KeyFrame::KeyFrame(Frame&, ...) copies the bounds,
// the cell size and every cell (KeyFrame.cc:34-78).  The stand-in does the same through a temporary Frame whose
// AssignFeaturesToGrid is the reference's own text (grid_ref.cpp).
void KeyFrame::AssignFeaturesToGrid() {
  Frame F;
  F.N = N; F.Nleft = NLeft;
  F.mvKeys = mvKeys; F.mvKeysRight = mvKeysRight; F.mvKeysUn = mvKeysUn;
  F.mnMinX = mnMinX; F.mnMaxX = mnMaxX; F.mnMinY = mnMinY; F.mnMaxY = mnMaxY;
  F.mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(mnMaxX - mnMinX);   // Frame.cc:118-119
  F.mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(mnMaxY - mnMinY);
  F.AssignFeaturesToGrid();
  mfGridElementWidthInv = F.mfGridElementWidthInv; mfGridElementHeightInv = F.mfGridElementHeightInv;
  mGrid.assign(mnGridCols, std::vector<std::vector<std::size_t> >(mnGridRows));       // KeyFrame.cc:65-77
  mGridRight.assign(mnGridCols, std::vector<std::vector<std::size_t> >(mnGridRows));
  for (int i = 0; i < mnGridCols; i++)
    for (int j = 0; j < mnGridRows; j++) {
      mGrid[i][j] = F.mGrid[i][j];
      if (F.Nleft != -1) mGridRight[i][j] = F.mGridRight[i][j];
    }
}

}  // namespace ORB_SLAM3

using namespace ORB_SLAM3;

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
cv::Mat desc_row(const uint8_t* d) {
  cv::Mat m(1, 32, CV_8UC1);
  memcpy(m.data, d, 32);
  return m;
}
DBoW2::FeatureVector to_fv(const int* node, const int* off, const int* idx, int nn) {
  DBoW2::FeatureVector fv;
  for (int a = 0; a < nn; a++)
    for (int i = off[a]; i < off[a + 1]; i++) fv[(unsigned)node[a]].push_back((unsigned)idx[i]);
  return fv;
}

// A frame (single camera when n_r < 0, else a two-camera rig: left features then right features)
void fill_frame(Frame& F, const void* kps_l, const uint8_t* desc_l, int n_l, const void* kps_r, const uint8_t* desc_r,
                int n_r, const float* bounds, const float* scale_factors, int nlevels) {
  const bool rig = n_r >= 0;
  std::vector<cv::KeyPoint> kl = to_keys(kps_l, n_l);
  F.mvKeys = kl;
  F.mvKeysUn = kl;
  std::vector<uint8_t> d(desc_l, desc_l + (size_t)n_l * 32);
  if (rig) {
    F.mvKeysRight = to_keys(kps_r, n_r);
    d.insert(d.end(), desc_r, desc_r + (size_t)n_r * 32);
    F.Nleft = n_l; F.Nright = n_r; F.N = n_l + n_r;
    F.mvLeftToRightMatch.assign(n_l, -1);
    F.mvRightToLeftMatch.assign(n_r, -1);
  } else {
    F.N = n_l;
  }
  F.mDescriptors = to_desc(d.data(), F.N);
  F.mvuRight.assign(F.N, -1.f);
  F.mvpMapPoints.assign(F.N, static_cast<MapPoint*>(NULL));
  F.mvbOutlier.assign(F.N, false);
  F.mvScaleFactors.assign(scale_factors, scale_factors + nlevels);
  F.mnScaleLevels = nlevels;
  F.mfLogScaleFactor = logf(scale_factors[1]);
  F.mnMinX = bounds[0]; F.mnMaxX = bounds[1]; F.mnMinY = bounds[2]; F.mnMaxY = bounds[3];
  F.mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(F.mnMaxX - F.mnMinX);   // Frame.cc:118-119
  F.mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(F.mnMaxY - F.mnMinY);
  F.AssignFeaturesToGrid();
}

// features that hold an observed MapPoint on entry get a private dummy point
void lock_features(Frame& F, const uint8_t* locked, int offset, int n, std::list<MapPoint>& store) {
  if (!locked) return;
  for (int j = 0; j < n; j++)
    if (locked[j]) { store.emplace_back(); store.back().nObs = 1; store.back().id = -2; F.mvpMapPoints[offset + j] = &store.back(); }
}

}  // namespace

extern "C" {

// SearchByProjection(Frame &F, const vector<MapPoint*> &vpMapPoints, th, bFarPoints, thFarPoints)  ORBmatcher.cc:43-213.
// Per MapPoint: flags bit0 mbTrackInView, bit1 Observations()>0, bit2 mbTrackInViewR, bit3 isBad; proj = (mTrackProjX,
// mTrackProjY, mTrackProjXR, mTrackProjYR), view_cos / view_cos_r, level / level_r, depth.  n_r < 0: single camera.
// holder[j] = index of the MapPoint feature j holds after the call (-1 none, -2 a point it held on entry).
int ref_search_by_projection_mappoints(const void* kps_l, const uint8_t* desc_l, int n_l, const void* kps_r,
                                       const uint8_t* desc_r, int n_r, const float* bounds, const float* scale_factors,
                                       int nlevels, const float* u_right, const uint8_t* locked, const int* l2r,
                                       const int* r2l, int M, const uint8_t* flags, const float* proj,
                                       const float* view_cos, const float* view_cos_r, const int* level,
                                       const int* level_r, const float* depth, const uint8_t* qdesc, float nnratio,
                                       float th, int far_points, float th_far, int* holder) {
  Frame F;
  fill_frame(F, kps_l, desc_l, n_l, kps_r, desc_r, n_r, bounds, scale_factors, nlevels);
  if (u_right) F.mvuRight.assign(u_right, u_right + F.N);
  std::list<MapPoint> dummies;
  lock_features(F, locked, 0, F.N, dummies);
  if (n_r >= 0) {
    if (l2r) F.mvLeftToRightMatch.assign(l2r, l2r + n_l);
    if (r2l) F.mvRightToLeftMatch.assign(r2l, r2l + n_r);
  }
  std::vector<MapPoint> pts(M);
  std::vector<MapPoint*> vp(M);
  for (int i = 0; i < M; i++) {
    MapPoint& p = pts[i];
    p.id = i;
    p.mbTrackInView = flags[i] & 1; p.nObs = (flags[i] & 2) ? 1 : 0; p.mbTrackInViewR = (flags[i] & 4) != 0;
    p.mbBad = (flags[i] & 8) != 0;
    p.mTrackProjX = proj[4 * i]; p.mTrackProjY = proj[4 * i + 1]; p.mTrackProjXR = proj[4 * i + 2]; p.mTrackProjYR = proj[4 * i + 3];
    p.mTrackViewCos = view_cos[i]; p.mTrackViewCosR = view_cos_r ? view_cos_r[i] : 0.f;
    p.mnTrackScaleLevel = level[i]; p.mnTrackScaleLevelR = level_r ? level_r[i] : -1;
    p.mTrackDepth = depth ? depth[i] : 0.f;
    p.mDescriptor = desc_row(qdesc + (size_t)i * 32);
    vp[i] = &p;
  }
  ORBmatcher matcher(nnratio);
  const int nm = matcher.SearchByProjection(F, vp, th, far_points != 0, th_far);
  for (int j = 0; j < F.N; j++) holder[j] = F.mvpMapPoints[j] ? F.mvpMapPoints[j]->id : -1;
  return nm;
}

// SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, th, bMono)  ORBmatcher.cc:1676-1887.
// The last frame holds M MapPoints; point i sits at world (u_i, v_i, 1), the current pose is the identity and the
// camera is fx = fy = 1, cx = cy = 0, so the reference's own projection code yields exactly (u_i, v_i); the right
// camera of a rig is displaced by trl_x along x (GetRelativePoseTrl).  tlc_z = z of the last frame's translation
// (bForward / bBackward, :1692-1693).  last_flags bit0: mvpMapPoints[i] != NULL, bit1: Observations()>0, bit2: outlier.
int ref_search_by_projection_lastframe(const void* kps_l, const uint8_t* desc_l, int n_l, const void* kps_r,
                                       const uint8_t* desc_r, int n_r, const float* bounds, const float* scale_factors,
                                       int nlevels, const float* u_right, const uint8_t* locked, int M,
                                       const uint8_t* last_flags, const float* uv, const int* last_octave,
                                       const float* last_angle, const uint8_t* qdesc, float mb, float mbf, float trl_x,
                                       float tlc_z, float th, int mono, int check_ori, int* holder) {
  Frame C, L;
  fill_frame(C, kps_l, desc_l, n_l, kps_r, desc_r, n_r, bounds, scale_factors, nlevels);
  if (u_right) C.mvuRight.assign(u_right, u_right + C.N);
  std::list<MapPoint> dummies;
  lock_features(C, locked, 0, C.N, dummies);
  GeometricCamera cam;  // fx = fy = 1, cx = cy = 0
  C.mpCamera = &cam; C.mb = mb; C.mbf = mbf;
  C.mTrl.t = Eigen::Vector3f(trl_x, 0.f, 0.f);
  L.N = M; L.Nleft = -1;
  L.mvKeys.resize(M); L.mvKeysUn.resize(M);
  L.mvpMapPoints.assign(M, static_cast<MapPoint*>(NULL));
  L.mvbOutlier.assign(M, false);
  L.mTcw.t = Eigen::Vector3f(0.f, 0.f, tlc_z);
  std::vector<MapPoint> pts(M);
  for (int i = 0; i < M; i++) {
    pts[i].id = i;
    pts[i].nObs = (last_flags[i] & 2) ? 1 : 0;
    pts[i].mWorldPos = Eigen::Vector3f(uv[2 * i], uv[2 * i + 1], 1.f);
    pts[i].mDescriptor = desc_row(qdesc + (size_t)i * 32);
    L.mvKeys[i].octave = last_octave[i]; L.mvKeysUn[i].octave = last_octave[i];
    L.mvKeys[i].angle = last_angle[i]; L.mvKeysUn[i].angle = last_angle[i];
    if (last_flags[i] & 1) L.mvpMapPoints[i] = &pts[i];
    L.mvbOutlier[i] = (last_flags[i] & 4) != 0;
  }
  ORBmatcher matcher(0.9f, check_ori != 0);
  const int nm = matcher.SearchByProjection(C, L, th, mono != 0);
  for (int j = 0; j < C.N; j++) holder[j] = C.mvpMapPoints[j] ? C.mvpMapPoints[j]->id : -1;
  return nm;
}

// a keyframe from flat arrays; holder_flags[i]: bit0 GetMapPoint(i) != NULL, bit1 that point isBad()
static void fill_keyframe(KeyFrame& K, std::vector<MapPoint>& pts, const void* kps, const uint8_t* desc, int n, int nleft,
                          const uint8_t* holder_flags, const int* node, const int* off, const int* idx, int nn,
                          const float* u_right, const float* bounds, const float* scale_factors,
                          const float* level_sigma2, int nlevels) {
  std::vector<cv::KeyPoint> k = to_keys(kps, n);
  K.N = n; K.NLeft = nleft;
  if (nleft < 0) { K.mvKeysUn = k; K.mvKeys = k; }
  else {
    K.mvKeys.assign(k.begin(), k.begin() + nleft);
    K.mvKeysRight.assign(k.begin() + nleft, k.end());
    K.mvKeysUn = k;
  }
  K.mDescriptors = to_desc(desc, n);
  if (node) K.mFeatVec = to_fv(node, off, idx, nn);
  K.mvuRight.assign(n, -1.f);
  if (u_right) K.mvuRight.assign(u_right, u_right + n);
  K.mvScaleFactors.assign(scale_factors, scale_factors + nlevels);
  K.mvLevelSigma2.assign(level_sigma2, level_sigma2 + nlevels);
  K.mvInvLevelSigma2.resize(nlevels);
  for (int l = 0; l < nlevels; l++) K.mvInvLevelSigma2[l] = 1.0f / level_sigma2[l];
  K.mnScaleLevels = nlevels; K.mfLogScaleFactor = logf(scale_factors[1]);
  if (bounds) { K.mnMinX = bounds[0]; K.mnMaxX = bounds[1]; K.mnMinY = bounds[2]; K.mnMaxY = bounds[3]; K.AssignFeaturesToGrid(); }
  pts.resize(n);
  K.mvpMapPoints.assign(n, static_cast<MapPoint*>(NULL));
  for (int i = 0; i < n; i++) {
    pts[i].id = i;
    pts[i].mbBad = holder_flags && (holder_flags[i] & 2);
    pts[i].nObs = 1;
    if (!holder_flags || (holder_flags[i] & 1)) K.mvpMapPoints[i] = &pts[i];
  }
}

// SearchByBoW(KeyFrame* pKF, Frame &F, vector<MapPoint*> &vpMapPointMatches)  ORBmatcher.cc:223-425.
// nleft2 < 0: single-camera frame, else F.Nleft.  match21[j] = keyframe feature whose MapPoint frame feature j got.
int ref_search_by_bow_kf_frame(const void* kps1, const uint8_t* desc1, int n1, const uint8_t* holder_flags1,
                               const int* node1, const int* off1, const int* idx1, int nn1, const void* kps2,
                               const uint8_t* desc2, int n2, int nleft2, const int* node2, const int* off2,
                               const int* idx2, int nn2, const float* scale_factors, int nlevels, float nnratio,
                               int check_ori, int* match21) {
  KeyFrame K;
  std::vector<MapPoint> pts;
  std::vector<float> sig(nlevels, 1.f);
  fill_keyframe(K, pts, kps1, desc1, n1, -1, holder_flags1, node1, off1, idx1, nn1, nullptr, nullptr, scale_factors,
                sig.data(), nlevels);
  Frame F;
  std::vector<cv::KeyPoint> k2 = to_keys(kps2, n2);
  F.N = n2;
  if (nleft2 < 0) { F.mvKeys = k2; F.mvKeysUn = k2; }
  else {
    F.Nleft = nleft2; F.Nright = n2 - nleft2;
    F.mvKeys.assign(k2.begin(), k2.begin() + nleft2);
    F.mvKeysRight.assign(k2.begin() + nleft2, k2.end());
    static GeometricCamera cam2;
    F.mpCamera2 = &cam2;       // (!F.mpCamera2) selects F.mvKeys at :365-367
    K.mpCamera2 = &cam2;       // (!pKF->mpCamera2 || F.Nleft == -1) at :342-345; the keyframe itself stays NLeft = -1
    K.NLeft = n1;              // every keyframe feature is a left feature: realIdxKF >= NLeft never holds
    K.mvKeys = K.mvKeysUn;
  }
  F.mDescriptors = to_desc(desc2, n2);
  F.mFeatVec = to_fv(node2, off2, idx2, nn2);
  std::vector<MapPoint*> vm;
  ORBmatcher matcher(nnratio, check_ori != 0);
  const int nm = matcher.SearchByBoW(&K, F, vm);
  for (int j = 0; j < n2; j++) match21[j] = vm[j] ? vm[j]->id : -1;
  return nm;
}

// SearchByBoW(KeyFrame *pKF1, KeyFrame *pKF2, vector<MapPoint *> &vpMatches12)  ORBmatcher.cc:765-905
int ref_search_by_bow_kf_kf(const void* kps1, const uint8_t* desc1, int n1, const uint8_t* hf1, const int* node1,
                            const int* off1, const int* idx1, int nn1, const void* kps2, const uint8_t* desc2, int n2,
                            const uint8_t* hf2, const int* node2, const int* off2, const int* idx2, int nn2,
                            const float* scale_factors, int nlevels, float nnratio, int check_ori, int* match12) {
  KeyFrame K1, K2;
  std::vector<MapPoint> p1, p2;
  std::vector<float> sig(nlevels, 1.f);
  fill_keyframe(K1, p1, kps1, desc1, n1, -1, hf1, node1, off1, idx1, nn1, nullptr, nullptr, scale_factors, sig.data(), nlevels);
  fill_keyframe(K2, p2, kps2, desc2, n2, -1, hf2, node2, off2, idx2, nn2, nullptr, nullptr, scale_factors, sig.data(), nlevels);
  std::vector<MapPoint*> vm;
  ORBmatcher matcher(nnratio, check_ori != 0);
  const int nm = matcher.SearchByBoW(&K1, &K2, vm);
  for (int i = 0; i < n1; i++) match12[i] = vm[i] ? vm[i]->id : -1;
  return nm;
}

// SearchForTriangulation(KeyFrame *pKF1, KeyFrame *pKF2, vMatchedPairs, bOnlyStereo, bCoarse)  ORBmatcher.cc:907-1146,
// pinhole single-camera keyframes.  pose1 / pose2 = (R row-major, t) world->camera; cam = fx, fy, cx, cy (both).
// F12_out / ep_out = the fundamental matrix and epipole the reference code used (inputs of the oracle / the device).
int ref_search_for_triangulation(const void* kps1, const uint8_t* desc1, int n1, const uint8_t* hf1, const float* ur1,
                                 const int* node1, const int* off1, const int* idx1, int nn1, const void* kps2,
                                 const uint8_t* desc2, int n2, const uint8_t* hf2, const float* ur2, const int* node2,
                                 const int* off2, const int* idx2, int nn2, const float* scale_factors,
                                 const float* level_sigma2, int nlevels, const float* pose1, const float* pose2,
                                 const float* cam, int only_stereo, int coarse, int check_ori, int* match12,
                                 float* F12_out, float* ep_out) {
  KeyFrame K1, K2;
  std::vector<MapPoint> p1, p2;
  fill_keyframe(K1, p1, kps1, desc1, n1, -1, hf1, node1, off1, idx1, nn1, ur1, nullptr, scale_factors, level_sigma2, nlevels);
  fill_keyframe(K2, p2, kps2, desc2, n2, -1, hf2, node2, off2, idx2, nn2, ur2, nullptr, scale_factors, level_sigma2, nlevels);
  GeometricCamera c;
  c.fx = cam[0]; c.fy = cam[1]; c.cx = cam[2]; c.cy = cam[3];
  K1.mpCamera = &c; K2.mpCamera = &c;
  for (int i = 0; i < 9; i++) { K1.mTcw.R.m[i] = pose1[i]; K2.mTcw.R.m[i] = pose2[i]; }
  for (int i = 0; i < 3; i++) { K1.mTcw.t(i) = pose1[9 + i]; K2.mTcw.t(i) = pose2[9 + i]; }
  std::vector<std::pair<size_t, size_t> > pairs;
  ORBmatcher matcher(0.6f, check_ori != 0);
  const int nm = matcher.SearchForTriangulation(&K1, &K2, pairs, only_stereo != 0, coarse != 0);
  for (int i = 0; i < n1; i++) match12[i] = -1;
  for (auto& pr : pairs) match12[pr.first] = (int)pr.second;
  // what the reference computed at :914-932 (same stand-in arithmetic)
  const Sophus::SE3f T12 = K1.GetPose() * K2.GetPoseInverse();
  GeometricCamera::fundamental(c, c, T12.rotationMatrix(), T12.translation(), F12_out);
  const Eigen::Vector2f ep = c.project(K2.GetPose() * K1.GetCameraCenter());
  ep_out[0] = ep(0); ep_out[1] = ep(1);
  return nm;
}

// SearchForInitialization(Frame &F1, Frame &F2, vbPrevMatched, vnMatches12, windowSize)  ORBmatcher.cc:648-763
int ref_search_for_initialization(const void* kps1, const uint8_t* desc1, int n1, const void* kps2, const uint8_t* desc2,
                                  int n2, const float* bounds, const float* scale_factors, int nlevels, float* prev_matched,
                                  int window_size, float nnratio, int check_ori, int* match12) {
  Frame F1, F2;
  fill_frame(F1, kps1, desc1, n1, nullptr, nullptr, -1, bounds, scale_factors, nlevels);
  fill_frame(F2, kps2, desc2, n2, nullptr, nullptr, -1, bounds, scale_factors, nlevels);
  std::vector<cv::Point2f> prev(n1);
  for (int i = 0; i < n1; i++) prev[i] = cv::Point2f(prev_matched[2 * i], prev_matched[2 * i + 1]);
  std::vector<int> m12;
  ORBmatcher matcher(nnratio, check_ori != 0);
  const int nm = matcher.SearchForInitialization(F1, F2, prev, m12, window_size);
  for (int i = 0; i < n1; i++) { match12[i] = m12[i]; prev_matched[2 * i] = prev[i].x; prev_matched[2 * i + 1] = prev[i].y; }
  return nm;
}

// SearchByProjection(Frame &CurrentFrame, KeyFrame *pKF, const set<MapPoint*> &sAlreadyFound, th, ORBdist)
// ORBmatcher.cc:1889-2010 (relocalisation).  Keyframe point i sits at world (u_i, v_i, 1); identity pose, unit
// pinhole; its scale level is predicted by the reference from mfMaxDistance (the caller sets it so that
// PredictScale gives level_i); kf_flags bit0 pMP != NULL, bit1 isBad, bit2 in sAlreadyFound.
int ref_search_by_projection_reloc(const void* kps_c, const uint8_t* desc_c, int n_c, const float* bounds,
                                   const float* scale_factors, int nlevels, const uint8_t* holds_c, int M,
                                   const uint8_t* kf_flags, const float* uv, const float* max_dist,
                                   const float* min_dist, const float* kf_angle, const uint8_t* qdesc, float th,
                                   int orb_dist, int check_ori, int* holder) {
  Frame C;
  fill_frame(C, kps_c, desc_c, n_c, nullptr, nullptr, -1, bounds, scale_factors, nlevels);
  std::list<MapPoint> dummies;
  lock_features(C, holds_c, 0, n_c, dummies);   // any non-null entry blocks its feature here (:1951-1952)
  GeometricCamera cam;
  C.mpCamera = &cam;
  KeyFrame K;
  K.N = M;
  K.mvKeysUn.resize(M);
  K.mvpMapPoints.assign(M, static_cast<MapPoint*>(NULL));
  std::vector<MapPoint> pts(M);
  std::set<MapPoint*> found;
  for (int i = 0; i < M; i++) {
    pts[i].id = i; pts[i].nObs = 1;
    pts[i].mbBad = (kf_flags[i] & 2) != 0;
    pts[i].mWorldPos = Eigen::Vector3f(uv[2 * i], uv[2 * i + 1], 1.f);
    pts[i].mfMaxDistance = max_dist[i]; pts[i].mfMinDistance = min_dist[i];
    pts[i].mDescriptor = desc_row(qdesc + (size_t)i * 32);
    K.mvKeysUn[i].angle = kf_angle[i];
    if (kf_flags[i] & 1) K.mvpMapPoints[i] = &pts[i];
    if (kf_flags[i] & 4) found.insert(&pts[i]);
  }
  ORBmatcher matcher(0.9f, check_ori != 0);
  const int nm = matcher.SearchByProjection(C, &K, found, th, orb_dist);
  for (int j = 0; j < n_c; j++) holder[j] = C.mvpMapPoints[j] ? C.mvpMapPoints[j]->id : -1;
  return nm;
}

// Fuse(KeyFrame *pKF, const vector<MapPoint *> &vpMapPoints, th, bRight = false)  ORBmatcher.cc:1148-1338.
// MapPoint i sits at world (u_i, v_i, 1) (identity pose, unit pinhole -> the reference projects it to exactly
// (u_i, v_i)), its normal points at the camera, its distance range makes PredictScale return level_i.
// best_idx[i] = keyframe feature the reference fused MapPoint i into (Replace or AddObservation/AddMapPoint), -1 none.
int ref_fuse(const void* kps, const uint8_t* desc, int n, const uint8_t* holder_flags, const float* u_right,
             const float* bounds, const float* scale_factors, const float* level_sigma2, int nlevels, float bf, int M,
             const uint8_t* mp_flags, const float* uv, const float* max_dist, const float* min_dist,
             const uint8_t* qdesc, float th, int* best_idx) {
  KeyFrame K;
  std::vector<MapPoint> held;
  fill_keyframe(K, held, kps, desc, n, -1, holder_flags, nullptr, nullptr, nullptr, 0, u_right, bounds, scale_factors,
                level_sigma2, nlevels);
  GeometricCamera cam;
  K.mpCamera = &cam; K.fx = 1; K.fy = 1; K.cx = 0; K.cy = 0; K.mbf = bf;
  for (int j = 0; j < n; j++) { held[j].id = -2 - j; if (K.mvpMapPoints[j]) held[j].mObservations[&K] = std::tuple<int, int>(j, -1); }
  std::vector<MapPoint> pts(M);
  std::vector<MapPoint*> vp(M, static_cast<MapPoint*>(NULL));
  for (int i = 0; i < M; i++) {
    pts[i].id = i; pts[i].nObs = 1;
    pts[i].mbBad = (mp_flags[i] & 2) != 0;
    pts[i].mWorldPos = Eigen::Vector3f(uv[2 * i], uv[2 * i + 1], 1.f);
    const float nrm = std::sqrt(uv[2 * i] * uv[2 * i] + uv[2 * i + 1] * uv[2 * i + 1] + 1.f);
    pts[i].mNormalVector = Eigen::Vector3f(uv[2 * i] / nrm, uv[2 * i + 1] / nrm, 1.f / nrm);
    pts[i].mfMaxDistance = max_dist[i]; pts[i].mfMinDistance = min_dist[i];
    pts[i].mDescriptor = desc_row(qdesc + (size_t)i * 32);
    if (mp_flags[i] & 1) vp[i] = &pts[i];
    best_idx[i] = -1;
  }
  std::vector<MapPoint::Op> log;
  MapPoint::oplog = &log;
  ORBmatcher matcher(0.6f);
  const int nf = matcher.Fuse(&K, vp, th, false);
  MapPoint::oplog = nullptr;
  // kind 1 = AddObservation(pKF, bestIdx) logged by the fused point (idx, own id); kind 2 = Replace logged as
  // (own id, id of the point that takes over): either the fused point gives way to the held one (:1313-1314) or the
  // held one to the fused point (:1315-1316).  Held points carry id = -2 - feature, added points are looked up.
  std::map<int, int> feat_of;
  for (const MapPoint::Op& op : log) {
    if (op.kind == 1) { if (op.other >= 0 && best_idx[op.other] < 0) best_idx[op.other] = op.idx; feat_of[op.other] = op.idx; }
    else if (op.kind == 2) {
      const int self = op.idx, other = op.other;
      auto feature = [&](int id) { return id <= -2 ? -2 - id : (feat_of.count(id) ? feat_of[id] : -1); };
      if (self >= 0 && best_idx[self] < 0) { best_idx[self] = feature(other); feat_of[self] = best_idx[self]; }
      else if (other >= 0 && best_idx[other] < 0) { best_idx[other] = feature(self); feat_of[other] = best_idx[other]; }
    }
  }
  return nf;
}

// Fuse(KeyFrame *pKF, Sophus::Sim3f &Scw, vpPoints, th, vpReplacePoint)  ORBmatcher.cc:1340-1455 (loop closing).
// Scw = identity; points as in ref_fuse.  mp_flags bit1 = isBad, bit2 = the point is already one of the keyframe's
// own (it takes over feature own_feat[i] before the call, so spAlreadyFound holds it).  held_bad[j] = the point
// held by feature j isBad() (then vpReplacePoint stays empty but nFused still counts, :1440-1449).
// best_idx[i] = feature the point was fused into (-1 none); replaced[i] = 1 when that went through vpReplacePoint.
int ref_fuse_sim3(const void* kps, const uint8_t* desc, int n, const uint8_t* holder_flags, const uint8_t* held_bad,
                  const float* bounds, const float* scale_factors, int nlevels, int M, const uint8_t* mp_flags,
                  const int* own_feat, const float* uv, const float* max_dist, const float* min_dist,
                  const uint8_t* qdesc, float th, int* best_idx, uint8_t* replaced) {
  KeyFrame K;
  std::vector<MapPoint> held;
  std::vector<float> sig(nlevels, 1.f);
  fill_keyframe(K, held, kps, desc, n, -1, holder_flags, nullptr, nullptr, nullptr, 0, nullptr, bounds, scale_factors,
                sig.data(), nlevels);
  GeometricCamera cam;
  K.mpCamera = &cam; K.fx = 1; K.fy = 1; K.cx = 0; K.cy = 0;
  for (int j = 0; j < n; j++) { held[j].id = -2 - j; held[j].mbBad = held_bad && held_bad[j]; }
  std::vector<MapPoint> pts(M);
  std::vector<MapPoint*> vp(M), repl(M, static_cast<MapPoint*>(NULL));
  for (int i = 0; i < M; i++) {
    pts[i].id = i; pts[i].nObs = 1;
    pts[i].mbBad = (mp_flags[i] & 2) != 0;
    pts[i].mWorldPos = Eigen::Vector3f(uv[2 * i], uv[2 * i + 1], 1.f);
    const float nrm = std::sqrt(uv[2 * i] * uv[2 * i] + uv[2 * i + 1] * uv[2 * i + 1] + 1.f);
    pts[i].mNormalVector = Eigen::Vector3f(uv[2 * i] / nrm, uv[2 * i + 1] / nrm, 1.f / nrm);
    pts[i].mfMaxDistance = max_dist[i]; pts[i].mfMinDistance = min_dist[i];
    pts[i].mDescriptor = desc_row(qdesc + (size_t)i * 32);
    vp[i] = &pts[i];
    if ((mp_flags[i] & 4) && own_feat[i] >= 0 && own_feat[i] < n) K.mvpMapPoints[own_feat[i]] = &pts[i];
    best_idx[i] = -1; replaced[i] = 0;
  }
  std::vector<MapPoint::Op> log;
  MapPoint::oplog = &log;
  Sophus::Sim3f Scw;
  ORBmatcher matcher(0.8f);
  const int nf = matcher.Fuse(&K, Scw, vp, th, repl);
  MapPoint::oplog = nullptr;
  for (const MapPoint::Op& op : log)
    if (op.kind == 1 && op.other >= 0) best_idx[op.other] = op.idx;
  for (int i = 0; i < M; i++)
    if (repl[i]) { best_idx[i] = repl[i]->id <= -2 ? -2 - repl[i]->id : -3 - repl[i]->id; replaced[i] = 1; }
  return nf;
}

// SearchByProjection(KeyFrame* pKF, Sophus::Sim3f &Scw, vpPoints, vpMatched, th, ratioHamming)  ORBmatcher.cc:427-532
// (variant 0) and its twin with vpPointsKFs / vpMatchedKF :534-646 (variant 1).  Scw = identity with scale 1, unit
// pinhole, point i at world (u_i, v_i, 1) with its normal towards the camera.  matched_init[j] = query point held by
// keyframe feature j on entry (-1 none, -2 a foreign point); matched[j] = the same after the call.
int ref_search_by_projection_sim3(int variant, const void* kps, const uint8_t* desc, int n, const float* bounds,
                                  const float* scale_factors, int nlevels, const int* matched_init, int M,
                                  const uint8_t* mp_flags, const float* uv, const float* max_dist, const float* min_dist,
                                  const uint8_t* qdesc, int th, float ratio_hamming, int* matched) {
  KeyFrame K;
  std::vector<MapPoint> none;
  std::vector<float> sig(nlevels, 1.f);
  std::vector<uint8_t> nohold(n, 0);
  fill_keyframe(K, none, kps, desc, n, -1, nohold.data(), nullptr, nullptr, nullptr, 0, nullptr, bounds, scale_factors,
                sig.data(), nlevels);
  GeometricCamera cam;
  K.mpCamera = &cam;
  std::vector<MapPoint> pts(M);
  std::vector<MapPoint*> vp(M);
  std::vector<KeyFrame*> vpKFs(M, &K);
  for (int i = 0; i < M; i++) {
    pts[i].id = i; pts[i].nObs = 1;
    pts[i].mbBad = (mp_flags[i] & 2) != 0;
    pts[i].mWorldPos = Eigen::Vector3f(uv[2 * i], uv[2 * i + 1], 1.f);
    const float nrm = std::sqrt(uv[2 * i] * uv[2 * i] + uv[2 * i + 1] * uv[2 * i + 1] + 1.f);
    pts[i].mNormalVector = Eigen::Vector3f(uv[2 * i] / nrm, uv[2 * i + 1] / nrm, 1.f / nrm);
    pts[i].mfMaxDistance = max_dist[i]; pts[i].mfMinDistance = min_dist[i];
    pts[i].mDescriptor = desc_row(qdesc + (size_t)i * 32);
    vp[i] = &pts[i];
  }
  MapPoint foreign;
  foreign.id = -2;
  std::vector<MapPoint*> vpMatched(n, static_cast<MapPoint*>(NULL));
  std::vector<KeyFrame*> vpMatchedKF(n, static_cast<KeyFrame*>(NULL));
  for (int j = 0; j < n; j++) {
    if (matched_init[j] >= 0) vpMatched[j] = &pts[matched_init[j]];
    else if (matched_init[j] == -2) vpMatched[j] = &foreign;
  }
  Sophus::Sim3f Scw;
  ORBmatcher matcher(0.75f, true);
  const int nm = variant == 0 ? matcher.SearchByProjection(&K, Scw, vp, vpMatched, th, ratio_hamming)
                              : matcher.SearchByProjection(&K, Scw, vp, vpKFs, vpMatched, vpMatchedKF, th, ratio_hamming);
  for (int j = 0; j < n; j++) matched[j] = vpMatched[j] ? vpMatched[j]->id : -1;
  return nm;
}

// SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, vpMatches12, S12, th)  ORBmatcher.cc:1457-1674.  Identity poses,
// S12 = identity, fx = fy = 1, cx = cy = 0: MapPoint i of keyframe 1 sits at (uv1_i, 1) — where the reference's code
// projects it into keyframe 2 — and MapPoint j of keyframe 2 at (uv2_j, 1).  hf*: bit0 point present, bit1 isBad.
// init12[i] = feature of keyframe 2 whose point vpMatches12[i] holds on entry (-1 none).
// match12[i] = feature of keyframe 2 whose point vpMatches12[i] holds after the call.
int ref_search_by_sim3(const void* kps1, const uint8_t* desc1, int n1, const uint8_t* hf1, const float* uv1,
                       const float* maxd1, const float* mind1, const void* kps2, const uint8_t* desc2, int n2,
                       const uint8_t* hf2, const float* uv2, const float* maxd2, const float* mind2, const float* bounds,
                       const float* scale_factors, int nlevels, const int* init12, float th, int* match12) {
  KeyFrame K1, K2;
  std::vector<MapPoint> p1, p2;
  std::vector<float> sig(nlevels, 1.f);
  fill_keyframe(K1, p1, kps1, desc1, n1, -1, hf1, nullptr, nullptr, nullptr, 0, nullptr, bounds, scale_factors, sig.data(), nlevels);
  fill_keyframe(K2, p2, kps2, desc2, n2, -1, hf2, nullptr, nullptr, nullptr, 0, nullptr, bounds, scale_factors, sig.data(), nlevels);
  for (int i = 0; i < n1; i++) {
    p1[i].mWorldPos = Eigen::Vector3f(uv1[2 * i], uv1[2 * i + 1], 1.f);
    p1[i].mfMaxDistance = maxd1[i]; p1[i].mfMinDistance = mind1[i];
    p1[i].mDescriptor = desc_row(desc1 + (size_t)i * 32);
    p1[i].mObservations[&K1] = std::tuple<int, int>(i, -1);
  }
  for (int j = 0; j < n2; j++) {
    p2[j].mWorldPos = Eigen::Vector3f(uv2[2 * j], uv2[2 * j + 1], 1.f);
    p2[j].mfMaxDistance = maxd2[j]; p2[j].mfMinDistance = mind2[j];
    p2[j].mDescriptor = desc_row(desc2 + (size_t)j * 32);
    p2[j].mObservations[&K2] = std::tuple<int, int>(j, -1);
  }
  std::vector<MapPoint*> vpMatches12(n1, static_cast<MapPoint*>(NULL));
  for (int i = 0; i < n1; i++)
    if (init12[i] >= 0) vpMatches12[i] = &p2[init12[i]];
  Sophus::Sim3f S12;
  ORBmatcher matcher(0.75f, true);
  const int nf = matcher.SearchBySim3(&K1, &K2, vpMatches12, S12, th);
  for (int i = 0; i < n1; i++) match12[i] = vpMatches12[i] ? vpMatches12[i]->id : -1;
  return nf;
}

// SearchForTriangulation between keyframes of a two-camera rig (mpCamera2 != NULL), ORBmatcher.cc:907-1146 with the
// camera selection :1036-1062.  The geometric gate is KannalaBrandt8::TriangulateMatches (Eigen JacobiSVD): not
// compilable here, so the rig's cameras are table cameras whose epipolarConstrain looks the outcome up in
// gate[n1][n2] (keypoints carry their index in class_id) — and check that the reference handed them the camera pair and
// the (R12, t12) of the combination (left/right of KF1) x (left/right of KF2); *bad_calls counts violations.
namespace {
struct TableCamera : GeometricCamera {
  const uint8_t* gate; int n2, nleft1, nleft2; bool right;  // which camera of its keyframe this object is
  TableCamera* peer1[2]; TableCamera* peer2[2];
  Sophus::SE3f T[4]; int* bad;
  bool epipolarConstrain(GeometricCamera* pCamera2, const cv::KeyPoint& kp1, const cv::KeyPoint& kp2,
                         const Eigen::Matrix3f& R12, const Eigen::Vector3f& t12, const float, const float) override {
    const int r1 = kp1.class_id >= nleft1, r2 = kp2.class_id >= nleft2;
    const Sophus::SE3f& E = T[2 * r1 + r2];
    bool ok = this == peer1[r1] && pCamera2 == peer2[r2];
    // (the reference's translation unit may contract the pose products to FMAs, this file does not: compare to 1e-5;
    //  the four combinations differ by the 10 cm baseline)
    for (int i = 0; i < 9; i++) ok = ok && std::fabs(R12.m[i] - E.R.m[i]) < 1e-5f;
    for (int i = 0; i < 3; i++) ok = ok && std::fabs(t12(i) - E.t(i)) < 1e-5f;
    if (!ok) (*bad)++;
    return gate[(size_t)kp1.class_id * n2 + kp2.class_id] != 0;
  }
};
}  // namespace

int ref_search_for_triangulation_rig(const void* kps1, const uint8_t* desc1, int n1, int nleft1, const uint8_t* hf1,
                                     const int* node1, const int* off1, const int* idx1, int nn1, const void* kps2,
                                     const uint8_t* desc2, int n2, int nleft2, const uint8_t* hf2, const int* node2,
                                     const int* off2, const int* idx2, int nn2, const float* scale_factors, int nlevels,
                                     const float* pose1, const float* pose2, const float* pose_rl, const uint8_t* gate,
                                     int only_stereo, int coarse, int check_ori, int* match12, int* bad_calls) {
  KeyFrame K1, K2;
  std::vector<MapPoint> p1, p2;
  std::vector<float> sig(nlevels, 1.f);
  fill_keyframe(K1, p1, kps1, desc1, n1, nleft1, hf1, node1, off1, idx1, nn1, nullptr, nullptr, scale_factors, sig.data(), nlevels);
  fill_keyframe(K2, p2, kps2, desc2, n2, nleft2, hf2, node2, off2, idx2, nn2, nullptr, nullptr, scale_factors, sig.data(), nlevels);
  // keypoints carry their (global) feature index
  for (int i = 0; i < nleft1; i++) K1.mvKeys[i].class_id = i;
  for (int i = nleft1; i < n1; i++) K1.mvKeysRight[i - nleft1].class_id = i;
  for (int i = 0; i < nleft2; i++) K2.mvKeys[i].class_id = i;
  for (int i = nleft2; i < n2; i++) K2.mvKeysRight[i - nleft2].class_id = i;
  auto set_pose = [](Sophus::SE3f& T, const float* p) {
    for (int i = 0; i < 9; i++) T.R.m[i] = p[i];
    for (int i = 0; i < 3; i++) T.t(i) = p[9 + i];
  };
  set_pose(K1.mTcw, pose1); set_pose(K2.mTcw, pose2); set_pose(K1.mTrl, pose_rl); set_pose(K2.mTrl, pose_rl);
  TableCamera c1[2], c2[2];
  *bad_calls = 0;
  // the four combinations exactly as ORBmatcher.cc:934-939 forms them
  const Sophus::SE3f T1w = K1.GetPose(), Tw2 = K2.GetPoseInverse(), Tr1w = K1.GetRightPose(), Twr2 = K2.GetRightPoseInverse();
  const Sophus::SE3f Tc[4] = {T1w * Tw2, T1w * Twr2, Tr1w * Tw2, Tr1w * Twr2};
  for (TableCamera* c : {&c1[0], &c1[1], &c2[0], &c2[1]}) {
    c->gate = gate; c->n2 = n2; c->nleft1 = nleft1; c->nleft2 = nleft2; c->bad = bad_calls;
    c->peer1[0] = &c1[0]; c->peer1[1] = &c1[1]; c->peer2[0] = &c2[0]; c->peer2[1] = &c2[1];
    for (int i = 0; i < 4; i++) c->T[i] = Tc[i];
  }
  K1.mpCamera = &c1[0]; K1.mpCamera2 = &c1[1]; K2.mpCamera = &c2[0]; K2.mpCamera2 = &c2[1];
  std::vector<std::pair<size_t, size_t> > pairs;
  ORBmatcher matcher(0.6f, check_ori != 0);
  const int nm = matcher.SearchForTriangulation(&K1, &K2, pairs, only_stereo != 0, coarse != 0);
  for (int i = 0; i < n1; i++) match12[i] = -1;
  for (auto& pr : pairs) match12[pr.first] = (int)pr.second;
  return nm;
}

}  // extern "C"
