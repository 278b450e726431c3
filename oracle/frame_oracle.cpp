// frame_oracle.cpp — CPU ORACLE, part 2: per-point Frame geometry and the fisheye (two-camera rig) branches
// of the matcher.  TEST INFRASTRUCTURE ONLY (same rules as orb_oracle.cpp: only tests/, smoke() and bench.py's
// CPU legs may use it; nothing under orb_slam3_vio_fixes_b200/ links or calls it).
//
// Restates, in plain C++ with libm:
//   Frame::UndistortKeyPoints         src/Frame.cc:747-781   (cv::undistortPoints: double, source order of
//                                     opencv calib3d undistort.dispatch.cpp cvUndistortPointsInternal, 5 iterations)
//   Frame::isInFrustum                src/Frame.cc:512-586, isInFrustumChecks :1168-1242
//   MapPoint::PredictScale            src/MapPoint.cc:531-546 (libm logf)
//   GeometricCamera::project          src/CameraModels/Pinhole.cpp:49-55, KannalaBrandt8.cpp:67-85 (libm atan2f, double cos / sin)
//
// PINNING: undistortPoints is pinned bit for bit against the real OpenCV 4.13 binary (cv2.undistortPoints,
// tests/test_frame_geometry.py).  isInFrustum / isInFrustumChecks are pinned against the reference's own text:
// oracle/Makefile target `ref_frame` cuts the two functions verbatim out of /root/reference/src/Frame.cc at build time
// and compiles them over the stand-in Eigen / Sophus / MapPoint types (oracle/_ref/libref_frame.so);
// tests/test_ref_frame.py asserts array_equal of every tracking field (single camera and two-camera rig).  What that
// pins is the reference's logic (gate order, comparisons, the early mTrackProjX/Y write, untouched fields, the right
// pose composition); the 3-vector arithmetic itself is the stated contract — uncontracted fp32 in Eigen's fixed-size
// evaluation order (3-term dot = c0 + (c1 + c2)) — because Eigen is not in the image; it is cross-checked against an
// fp64 evaluation (the test reports how many decisions / levels differ).
// Must be compiled with -ffp-contract=off.
#include <cmath>
#include <cstdint>
#include <cstring>

namespace {

struct KeyPoint { float x, y, size, angle, response; int32_t octave, class_id; };

inline float dot3(float a0, float a1, float a2, float b0, float b1, float b2) {
  return a0 * b0 + (a1 * b1 + a2 * b2);
}

void project(int model, const float* cam /*fx,fy,cx,cy,k0..k3*/, float X, float Y, float Z, float* u, float* v) {
  if (model == 0) {  // Pinhole.cpp:49-55
    *u = cam[0] * X / Z + cam[2];
    *v = cam[1] * Y / Z + cam[3];
    return;
  }
  // KannalaBrandt8.cpp:67-85
  const float x2_plus_y2 = X * X + Y * Y;
  const float theta = atan2f(sqrtf(x2_plus_y2), Z);
  const float psi = atan2f(Y, X);
  const float theta2 = theta * theta;
  const float theta3 = theta * theta2;
  const float theta5 = theta3 * theta2;
  const float theta7 = theta5 * theta2;
  const float theta9 = theta7 * theta2;
  const float r = theta + cam[4] * theta3 + cam[5] * theta5 + cam[6] * theta7 + cam[7] * theta9;
  // `cos(psi)` / `sin(psi)` in KannalaBrandt8.cpp are the double functions (no using-directive for std reaches that
  // translation unit), so fx * r (float) is multiplied by the cosine and added to cx in double, then stored as float.
  // Pinned against the reference's own text (oracle/ref_shim_matcher/camera_ref.cpp, tests/test_ref_frame.py).
  *u = (float)((double)(cam[0] * r) * cos((double)psi) + (double)cam[2]);
  *v = (float)((double)(cam[1] * r) * sin((double)psi) + (double)cam[3]);
}

}  // namespace

extern "C" {

// cv::undistortPoints(src, dst, K, distCoeffs, noArray(), K) on keypoint positions.
void oracle_undistort_keypoints(const void* kps_v, int n, float fxf, float fyf, float cxf, float cyf,
                                const float* dist, int n_coef, void* out_v) {
  const KeyPoint* kps = (const KeyPoint*)kps_v;
  KeyPoint* out = (KeyPoint*)out_v;
  if (n_coef == 0 || dist[0] == 0.0f) {  // Frame.cc:749-753
    std::memcpy(out, kps, (size_t)n * sizeof(KeyPoint));
    return;
  }
  double k[14] = {0};
  for (int i = 0; i < n_coef && i < 12; i++) k[i] = (double)dist[i];
  const double fx = fxf, fy = fyf, cx = cxf, cy = cyf, ifx = 1. / fx, ify = 1. / fy;
  for (int i = 0; i < n; i++) {
    double x = (double)kps[i].x, y = (double)kps[i].y;
    const double u = x, v = y;
    x = (x - cx) * ifx;
    y = (y - cy) * ify;
    const double x0 = x, y0 = y;
    for (int j = 0; j < 5; j++) {
      const double r2 = x * x + y * y;
      const double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
      if (icdist < 0) { x = (u - cx) * ifx; y = (v - cy) * ify; break; }
      const double deltaX = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2;
      const double deltaY = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2;
      x = (x0 - deltaX) * icdist;
      y = (y0 - deltaY) * icdist;
    }
    const double xx = fx * x + 0.0 * y + cx, yy = 0.0 * x + fy * y + cy, ww = 1. / (0.0 * x + 0.0 * y + 1.0);
    out[i] = kps[i];
    out[i].x = (float)(xx * ww);
    out[i].y = (float)(yy * ww);
  }
}

// Frame::isInFrustum / isInFrustumChecks for n MapPoints against one camera pose.  pose = R[9], t[3], c[3].
// use_f64 != 0: the same gates evaluated in double (only for the "how many differ" report, not a reference form).
void oracle_is_in_frustum(const float* pos, const float* normal, const float* min_dist, const float* max_dist, int n,
                          const float* pose, int model, const float* cam, float mbf, float mnMinX, float mnMaxX,
                          float mnMinY, float mnMaxY, float viewingCosLimit, float mfLogScaleFactor, int mnScaleLevels,
                          int fisheye_rig, int use_f64, uint8_t* in_view, float* proj_x, float* proj_y, float* proj_xr,
                          float* depth, float* view_cos, int8_t* level) {
  const float* R = pose; const float* t = pose + 9; const float* c = pose + 12;
  for (int i = 0; i < n; i++) {
    in_view[i] = 0; proj_x[i] = -1; proj_y[i] = -1; proj_xr[i] = -1; depth[i] = 0; view_cos[i] = 0; level[i] = -1;
    const float* P = pos + 3 * i; const float* Pn = normal + 3 * i;
    if (use_f64) {
      const double X = (double)R[0] * P[0] + (double)R[1] * P[1] + (double)R[2] * P[2] + t[0];
      const double Y = (double)R[3] * P[0] + (double)R[4] * P[1] + (double)R[5] * P[2] + t[1];
      const double Z = (double)R[6] * P[0] + (double)R[7] * P[1] + (double)R[8] * P[2] + t[2];
      if (Z < 0) continue;
      double u, v;
      if (model == 0) { u = cam[0] * X / Z + cam[2]; v = cam[1] * Y / Z + cam[3]; }
      else {
        const double th = atan2(sqrt(X * X + Y * Y), Z), psi = atan2(Y, X), t2 = th * th;
        const double r = th * (1 + t2 * (cam[4] + t2 * (cam[5] + t2 * (cam[6] + t2 * cam[7]))));
        u = cam[0] * r * cos(psi) + cam[2]; v = cam[1] * r * sin(psi) + cam[3];
      }
      if (u < mnMinX || u > mnMaxX || v < mnMinY || v > mnMaxY) continue;
      if (!fisheye_rig) { proj_x[i] = (float)u; proj_y[i] = (float)v; }
      const double ox = (double)P[0] - c[0], oy = (double)P[1] - c[1], oz = (double)P[2] - c[2];
      const double dist = sqrt(ox * ox + oy * oy + oz * oz);
      if (dist < 0.8 * (double)min_dist[i] || dist > 1.2 * (double)max_dist[i]) continue;
      const double vc = (ox * Pn[0] + oy * Pn[1] + oz * Pn[2]) / dist;
      if (vc < viewingCosLimit) continue;
      int nScale = (int)ceil(log((double)max_dist[i] / dist) / (double)mfLogScaleFactor);
      if (nScale < 0) nScale = 0; else if (nScale >= mnScaleLevels) nScale = mnScaleLevels - 1;
      in_view[i] = 1; proj_x[i] = (float)u; proj_y[i] = (float)v;
      proj_xr[i] = fisheye_rig ? -1.f : (float)(u - mbf / Z);
      depth[i] = (float)sqrt(X * X + Y * Y + Z * Z); view_cos[i] = (float)vc; level[i] = (int8_t)nScale;
      continue;
    }
    // 3D in camera coordinates (:523, :1188)
    const float Pcx = dot3(R[0], R[1], R[2], P[0], P[1], P[2]) + t[0];
    const float Pcy = dot3(R[3], R[4], R[5], P[0], P[1], P[2]) + t[1];
    const float Pcz = dot3(R[6], R[7], R[8], P[0], P[1], P[2]) + t[2];
    const float Pc_dist = sqrtf(dot3(Pcx, Pcy, Pcz, Pcx, Pcy, Pcz));
    const float invz = 1.0f / Pcz;
    if (Pcz < 0.0f) continue;
    float u, v;
    project(model, cam, Pcx, Pcy, Pcz, &u, &v);
    if (u < mnMinX || u > mnMaxX) continue;
    if (v < mnMinY || v > mnMaxY) continue;
    if (!fisheye_rig) { proj_x[i] = u; proj_y[i] = v; }  // :539-540
    const float maxDistance = 1.2f * max_dist[i];  // MapPoint::GetMaxDistanceInvariance
    const float minDistance = 0.8f * min_dist[i];
    const float POx = P[0] - c[0], POy = P[1] - c[1], POz = P[2] - c[2];
    const float dist = sqrtf(dot3(POx, POy, POz, POx, POy, POz));
    if (dist < minDistance || dist > maxDistance) continue;
    const float viewCos = dot3(POx, POy, POz, Pn[0], Pn[1], Pn[2]) / dist;
    if (viewCos < viewingCosLimit) continue;
    // MapPoint::PredictScale
    const float ratio = max_dist[i] / dist;
    int nScale = (int)ceilf(logf(ratio) / mfLogScaleFactor);
    if (nScale < 0) nScale = 0;
    else if (nScale >= mnScaleLevels) nScale = mnScaleLevels - 1;
    in_view[i] = 1;
    proj_x[i] = u; proj_y[i] = v;
    proj_xr[i] = fisheye_rig ? -1.f : u - mbf * invz;
    depth[i] = Pc_dist; view_cos[i] = viewCos; level[i] = (int8_t)nScale;
  }
}

}  // extern "C"

// =====================================================================================================================
// Fisheye rig (F.Nleft != -1, KeyFrame::mpCamera2 != NULL) branches of the matcher, src/ORBmatcher.cc.
// A two-camera Frame holds its features concatenated: indices [0, Nleft) are mvKeys (left image), [Nleft, N) are
// mvKeysRight (right image); mDescriptors / mvpMapPoints follow the same numbering; mGrid indexes the left
// features, mGridRight the right ones by their index minus Nleft (Frame.cc:385-416).
// PINNED: SearchByProjection (both variants) and SearchByBoW on a rig frame are compared with the reference's own
// ORBmatcher.cc compiled unmodified (oracle/_ref/libref_orbmatcher.so, tests/test_ref_matcher.py) and with
// hand-derived known answers (tests/test_oracle_rig_known_answers.py).  The rig SearchForTriangulation needs
// KannalaBrandt8::TriangulateMatches (Eigen JacobiSVD) and stays on the tolerance contract.
// =====================================================================================================================
#include <algorithm>
#include <vector>

namespace {

int HammingDistance(const uint8_t* a, const uint8_t* b) {  // ORBmatcher::DescriptorDistance, ORBmatcher.cc:2058-2074
  int dist = 0;
  for (int i = 0; i < 8; i++) {
    uint32_t pa, pb;
    memcpy(&pa, a + 4 * i, 4); memcpy(&pb, b + 4 * i, 4);
    unsigned int v = pa ^ pb;
    v = v - ((v >> 1) & 0x55555555);
    v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
    dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
  }
  return dist;
}

int RotBin(float a1, float a2) {  // ORBmatcher.cc:332-338 and twins
  const float factor = 1.0f / 30;
  float rot = a1 - a2;
  if (rot < 0.0) rot += 360.0f;
  int bin = (int)std::round(rot * factor);
  if (bin == 30) bin = 0;
  return bin;
}

void ThreeMaxima(const std::vector<int>* histo, const int L, int& ind1, int& ind2, int& ind3) {  // :2012-2053
  int max1 = 0, max2 = 0, max3 = 0;
  for (int i = 0; i < L; i++) {
    const int s = (int)histo[i].size();
    if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
    else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
    else if (s > max3) { max3 = s; ind3 = i; }
  }
  if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
  else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
}

// Frame grid of one camera of the rig + GetFeaturesInArea (Frame.cc:385-416, 657-735)
struct Grid {
  static const int COLS = 64, ROWS = 48;
  std::vector<std::vector<int>> cell;
  float minX, minY, wInv, hInv;
  const KeyPoint* kps;  // that camera's keypoints, local numbering
  Grid(const KeyPoint* k, int n, float mnMinX, float mnMaxX, float mnMinY, float mnMaxY) : cell(COLS * ROWS), kps(k) {
    minX = mnMinX; minY = mnMinY;
    wInv = static_cast<float>(COLS) / static_cast<float>(mnMaxX - mnMinX);
    hInv = static_cast<float>(ROWS) / static_cast<float>(mnMaxY - mnMinY);
    for (int i = 0; i < n; i++) {
      const int px = (int)std::round((k[i].x - minX) * wInv), py = (int)std::round((k[i].y - minY) * hInv);
      if (px < 0 || px >= COLS || py < 0 || py >= ROWS) continue;
      cell[px * ROWS + py].push_back(i);
    }
  }
  std::vector<int> area(float x, float y, float r, int minLevel, int maxLevel) const {
    std::vector<int> v;
    const int nMinCellX = std::max(0, (int)std::floor((x - minX - r) * wInv));
    if (nMinCellX >= COLS) return v;
    const int nMaxCellX = std::min(COLS - 1, (int)std::ceil((x - minX + r) * wInv));
    if (nMaxCellX < 0) return v;
    const int nMinCellY = std::max(0, (int)std::floor((y - minY - r) * hInv));
    if (nMinCellY >= ROWS) return v;
    const int nMaxCellY = std::min(ROWS - 1, (int)std::ceil((y - minY + r) * hInv));
    if (nMaxCellY < 0) return v;
    const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
    for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
      for (int iy = nMinCellY; iy <= nMaxCellY; iy++)
        for (int j : cell[ix * ROWS + iy]) {
          const KeyPoint& kp = kps[j];
          if (bCheckLevels) {
            if (kp.octave < minLevel) continue;
            if (maxLevel >= 0 && kp.octave > maxLevel) continue;
          }
          if (std::fabs(kp.x - x) < r && std::fabs(kp.y - y) < r) v.push_back(j);
        }
    return v;
  }
};

}  // namespace

extern "C" {

// SearchByBoW(KeyFrame*, Frame&, vector<MapPoint*>&) with F.Nleft != -1 (ORBmatcher.cc:223-425, the `else` branch
// :295-323 and the right-side acceptance :356-380).  kps1 / kps2 = the concatenated left|right keypoints of the
// keyframe / the frame (what the mpCamera2 ternaries at :335-338, :342-345, :364-367, :371-374 select); nleft2 = F.Nleft.
// match21[j] = side-1 feature whose MapPoint vpMapPointMatches[j] holds; match12l / match12r = the left / right
// feature written for side-1 feature i.  Returns nmatches.
int oracle_search_by_bow_fisheye(const void* kps1_v, const uint8_t* desc1, int n1, const uint8_t* valid1,
                                 const int* fv1_node, const int* fv1_off, const int* fv1_idx, int nn1,
                                 const void* kps2_v, const uint8_t* desc2, int n2, int nleft2, const int* fv2_node,
                                 const int* fv2_off, const int* fv2_idx, int nn2, float mfNNratio, int check_orientation,
                                 int* match12l, int* match12r, int* match21) {
  const KeyPoint* kps1 = (const KeyPoint*)kps1_v;
  const KeyPoint* kps2 = (const KeyPoint*)kps2_v;
  const int TH_LOW = 50, HISTO_LENGTH = 30;
  for (int i = 0; i < n1; i++) { match12l[i] = -1; match12r[i] = -1; }
  for (int j = 0; j < n2; j++) match21[j] = -1;
  std::vector<int> rotHist[30];
  int nmatches = 0, a = 0, b = 0;
  while (a < nn1 && b < nn2) {
    if (fv1_node[a] == fv2_node[b]) {
      for (int iKF = fv1_off[a]; iKF < fv1_off[a + 1]; iKF++) {
        const int realIdxKF = fv1_idx[iKF];
        if (valid1 && !valid1[realIdxKF]) continue;  // !pMP || pMP->isBad()
        const uint8_t* dKF = desc1 + (size_t)realIdxKF * 32;
        int bestDist1 = 256, bestIdxF = -1, bestDist2 = 256;
        int bestDist1R = 256, bestIdxFR = -1, bestDist2R = 256;
        for (int iF = fv2_off[b]; iF < fv2_off[b + 1]; iF++) {
          const int realIdxF = fv2_idx[iF];
          if (match21[realIdxF] >= 0) continue;  // vpMapPointMatches[realIdxF]
          const int dist = HammingDistance(dKF, desc2 + (size_t)realIdxF * 32);
          if (realIdxF < nleft2 && dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdxF = realIdxF; }
          else if (realIdxF < nleft2 && dist < bestDist2) { bestDist2 = dist; }
          if (realIdxF >= nleft2 && dist < bestDist1R) { bestDist2R = bestDist1R; bestDist1R = dist; bestIdxFR = realIdxF; }
          else if (realIdxF >= nleft2 && dist < bestDist2R) { bestDist2R = dist; }
        }
        if (bestDist1 <= TH_LOW) {
          if (static_cast<float>(bestDist1) < mfNNratio * static_cast<float>(bestDist2)) {
            match21[bestIdxF] = realIdxKF; match12l[realIdxKF] = bestIdxF;
            if (check_orientation) rotHist[RotBin(kps1[realIdxKF].angle, kps2[bestIdxF].angle)].push_back(bestIdxF);
            nmatches++;
          }
          if (bestDist1R <= TH_LOW) {
            // `static_cast<float>(bestDist1R)<mfNNratio*static_cast<float>(bestDist2R) || true` (:358)
            (void)bestDist2R;
            match21[bestIdxFR] = realIdxKF; match12r[realIdxKF] = bestIdxFR;
            if (check_orientation) rotHist[RotBin(kps1[realIdxKF].angle, kps2[bestIdxFR].angle)].push_back(bestIdxFR);
            nmatches++;
          }
        }
      }
      a++; b++;
    } else if (fv1_node[a] < fv2_node[b]) {
      a = (int)(std::lower_bound(fv1_node, fv1_node + nn1, fv2_node[b]) - fv1_node);
    } else {
      b = (int)(std::lower_bound(fv2_node, fv2_node + nn2, fv1_node[a]) - fv2_node);
    }
  }
  if (check_orientation) {
    int ind1 = -1, ind2 = -1, ind3 = -1;
    ThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (int j : rotHist[i]) {  // vpMapPointMatches[rotHist[i][j]] = NULL
        const int i1 = match21[j];
        if (match12l[i1] == j) match12l[i1] = -1;
        if (match12r[i1] == j) match12r[i1] = -1;
        match21[j] = -1;
        nmatches--;
      }
    }
  }
  return nmatches;
}

// ORBmatcher::SearchByProjection on a two-camera Frame (F.Nleft != -1), one frame:
//   mode 0  (Frame&, const vector<MapPoint*>&, th, ...)  ORBmatcher.cc:43-213: left block :59-142 (the mvuRight gate
//           :92-97 is off when Nleft != -1), right block :144-210; a left match also writes the right feature paired
//           with it (mvLeftToRightMatch, :131-135) and vice versa (:197-201).  NOTE the two `continue`s of the left
//           block (:126-127 failed ratio test) skip the right block of the same MapPoint.
//   mode 1  (Frame& CurrentFrame, const Frame& LastFrame, th, bMono)  :1676-1887: left :1716-1792, right :1794-1862
//           (reached only when the left window was not empty, `continue` at :1737-1738); one rotation histogram.
// Left / right features, grids, locks and outputs use each camera's LOCAL numbering (right global index = local +
// Nleft).  q_flags: bit0 = left query takes part (mode 0: mbTrackInView; mode 1: the projection passed :1706-1714),
// bit1 = the MapPoint has Observations()>0 (a feature it is written to becomes "locked", :88-90), bit2 (mode 0) =
// mbTrackInViewR && mnTrackScaleLevelR != -1.  feat_query_*[j] = last query written to feature j (-1 none).
int oracle_search_by_projection_rig(int mode, const void* kps_l_v, const uint8_t* desc_l, int n_l, const void* kps_r_v,
                                    const uint8_t* desc_r, int n_r, float mnMinX, float mnMaxX, float mnMinY,
                                    float mnMaxY, const uint8_t* locked_l_in, const uint8_t* locked_r_in,
                                    const int* l2r, const int* r2l, const float* qx, const float* qy, const float* qr,
                                    const int8_t* qmin, const int8_t* qmax, const float* qxr, const float* qyr,
                                    const float* qrr, const int8_t* qminr, const int8_t* qmaxr, const uint8_t* qdesc,
                                    int nq, const uint8_t* q_flags, const float* q_angle, float mfNNratio,
                                    int check_orientation, int* match_l, int* match_r, int* feat_query_l,
                                    int* feat_query_r) {
  const KeyPoint* kl = (const KeyPoint*)kps_l_v;
  const KeyPoint* kr = (const KeyPoint*)kps_r_v;
  const int TH_HIGH = 100, HISTO_LENGTH = 30;
  const Grid gl(kl, n_l, mnMinX, mnMaxX, mnMinY, mnMaxY), gr(kr, n_r, mnMinX, mnMaxX, mnMinY, mnMaxY);
  std::vector<uint8_t> lockL(n_l, 0), lockR(n_r, 0);
  for (int j = 0; j < n_l; j++) { lockL[j] = locked_l_in ? locked_l_in[j] != 0 : 0; feat_query_l[j] = -1; }
  for (int j = 0; j < n_r; j++) { lockR[j] = locked_r_in ? locked_r_in[j] != 0 : 0; feat_query_r[j] = -1; }
  std::vector<std::pair<int, int>> rotHist[30];  // (side, local feature)
  int nmatches = 0;
  for (int i = 0; i < nq; i++) {
    match_l[i] = -1; match_r[i] = -1;
    const int fl = q_flags ? q_flags[i] : 7;
    const bool locker = (fl & 2) != 0;
    const uint8_t* dMP = qdesc + (size_t)i * 32;
    if (mode == 0) {
      if (!(fl & 1) && !(fl & 4)) continue;  // :52-53
      if (fl & 1) {
        const std::vector<int> vIndices = gl.area(qx[i], qy[i], qr[i], qmin[i], qmax[i]);
        if (!vIndices.empty()) {
          int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
          for (int idx : vIndices) {
            if (lockL[idx]) continue;
            const int dist = HammingDistance(dMP, desc_l + (size_t)idx * 32);
            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = kl[idx].octave; bestIdx = idx; }
            else if (dist < bestDist2) { bestLevel2 = kl[idx].octave; bestDist2 = dist; }
          }
          if (bestDist <= TH_HIGH) {
            if (bestLevel == bestLevel2 && bestDist > mfNNratio * bestDist2) continue;  // skips the right block too
            if (bestLevel != bestLevel2 || bestDist <= mfNNratio * bestDist2) {
              feat_query_l[bestIdx] = i; lockL[bestIdx] = locker; match_l[i] = bestIdx;
              if (l2r[bestIdx] != -1) { feat_query_r[l2r[bestIdx]] = i; lockR[l2r[bestIdx]] = locker; nmatches++; }
              nmatches++;
            }
          }
        }
      }
      if (fl & 4) {
        const std::vector<int> vIndices = gr.area(qxr[i], qyr[i], qrr[i], qminr[i], qmaxr[i]);
        if (vIndices.empty()) continue;
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (int idx : vIndices) {
          if (lockR[idx]) continue;
          const int dist = HammingDistance(dMP, desc_r + (size_t)idx * 32);
          if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = kr[idx].octave; bestIdx = idx; }
          else if (dist < bestDist2) { bestLevel2 = kr[idx].octave; bestDist2 = dist; }
        }
        if (bestDist <= TH_HIGH) {
          if (bestLevel == bestLevel2 && bestDist > mfNNratio * bestDist2) continue;
          if (r2l[bestIdx] != -1) { feat_query_l[r2l[bestIdx]] = i; lockL[r2l[bestIdx]] = locker; nmatches++; }
          feat_query_r[bestIdx] = i; lockR[bestIdx] = locker; match_r[i] = bestIdx;
          nmatches++;
        }
      }
    } else {
      if (!(fl & 1)) continue;
      const std::vector<int> vIndices2 = gl.area(qx[i], qy[i], qr[i], qmin[i], qmax[i]);
      if (vIndices2.empty()) continue;  // :1737-1738 (the right camera is not searched either)
      {
        int bestDist = 256, bestIdx2 = -1;
        for (int i2 : vIndices2) {
          if (lockL[i2]) continue;
          const int dist = HammingDistance(dMP, desc_l + (size_t)i2 * 32);
          if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
        }
        if (bestDist <= TH_HIGH) {
          feat_query_l[bestIdx2] = i; lockL[bestIdx2] = locker; match_l[i] = bestIdx2;
          nmatches++;
          if (check_orientation) rotHist[RotBin(q_angle[i], kl[bestIdx2].angle)].push_back({0, bestIdx2});
        }
      }
      {
        const std::vector<int> vR = gr.area(qxr[i], qyr[i], qrr[i], qminr[i], qmaxr[i]);
        int bestDist = 256, bestIdx2 = -1;
        for (int i2 : vR) {
          if (lockR[i2]) continue;
          const int dist = HammingDistance(dMP, desc_r + (size_t)i2 * 32);
          if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
        }
        if (bestDist <= TH_HIGH) {
          feat_query_r[bestIdx2] = i; lockR[bestIdx2] = locker; match_r[i] = bestIdx2;
          nmatches++;
          if (check_orientation) rotHist[RotBin(q_angle[i], kr[bestIdx2].angle)].push_back({1, bestIdx2});
        }
      }
    }
  }
  if (mode == 1 && check_orientation) {
    std::vector<int> sizes[30];
    for (int b = 0; b < 30; b++) sizes[b].resize(rotHist[b].size());
    int ind1 = -1, ind2 = -1, ind3 = -1;
    ThreeMaxima(sizes, HISTO_LENGTH, ind1, ind2, ind3);
    for (int b = 0; b < HISTO_LENGTH; b++) {
      if (b == ind1 || b == ind2 || b == ind3) continue;
      for (auto& e : rotHist[b]) {  // CurrentFrame.mvpMapPoints[rotHist[i][j]] = NULL, nmatches-- (:1870-1876)
        if (e.first == 0) feat_query_l[e.second] = -1; else feat_query_r[e.second] = -1;
        nmatches--;
      }
    }
    for (int i = 0; i < nq; i++) {  // a query's own record goes with its histogram entry
      if (match_l[i] >= 0) {
        const int b = RotBin(q_angle[i], kl[match_l[i]].angle);
        if (!(b == ind1 || b == ind2 || b == ind3)) match_l[i] = -1;
      }
      if (match_r[i] >= 0) {
        const int b = RotBin(q_angle[i], kr[match_r[i]].angle);
        if (!(b == ind1 || b == ind2 || b == ind3)) match_r[i] = -1;
      }
    }
  }
  return nmatches;
}

// ORBmatcher::SearchForTriangulation between two keyframes of a two-camera rig (mpCamera2 != NULL), ORBmatcher.cc:907-1146.
// bStereo1 / bStereo2 are false when mpCamera2 is set (:973, :998), so bOnlyStereo skips every feature and the epipole
// gate (:1022-1030, needs !pKF1->mpCamera2) never applies.  The geometric gate pCamera1->epipolarConstrain(...) of the
// camera combination chosen at :1036-1062 is floating-point triangulation (KannalaBrandt8.cpp:306-375, Eigen JacobiSVD)
// and is NOT restated here: the caller supplies its outcome per (idx1, idx2) in gate[n1][n2] (the tests evaluate it in
// fp64 numpy), so this function pins the loop logic only — skips, `dist>TH_LOW || dist>bestDist`, last minimum wins,
// rotation histogram.
int oracle_search_for_triangulation_rig(const void* kps1_v, const uint8_t* desc1, int n1, const uint8_t* has_mp1,
                                        const int* fv1_node, const int* fv1_off, const int* fv1_idx, int nn1,
                                        const void* kps2_v, const uint8_t* desc2, int n2, const uint8_t* has_mp2,
                                        const int* fv2_node, const int* fv2_off, const int* fv2_idx, int nn2,
                                        const uint8_t* gate, int only_stereo, int coarse, int check_orientation,
                                        int* match12) {
  const KeyPoint* kps1 = (const KeyPoint*)kps1_v;
  const KeyPoint* kps2 = (const KeyPoint*)kps2_v;
  const int TH_LOW = 50, HISTO_LENGTH = 30;
  for (int i = 0; i < n1; i++) match12[i] = -1;
  std::vector<int> rotHist[30];
  int nmatches = 0, a = 0, b = 0;
  while (a < nn1 && b < nn2) {
    if (fv1_node[a] == fv2_node[b]) {
      for (int i1 = fv1_off[a]; i1 < fv1_off[a + 1]; i1++) {
        const int idx1 = fv1_idx[i1];
        if (has_mp1 && has_mp1[idx1]) continue;
        const bool bStereo1 = false;
        if (only_stereo && !bStereo1) continue;
        int bestDist = TH_LOW, bestIdx2 = -1;
        for (int i2 = fv2_off[b]; i2 < fv2_off[b + 1]; i2++) {
          const int idx2 = fv2_idx[i2];
          if (has_mp2 && has_mp2[idx2]) continue;  // vbMatched2 is never set in the reference
          const int dist = HammingDistance(desc1 + (size_t)idx1 * 32, desc2 + (size_t)idx2 * 32);
          if (dist > TH_LOW || dist > bestDist) continue;
          if (coarse || gate[(size_t)idx1 * n2 + idx2]) { bestIdx2 = idx2; bestDist = dist; }
        }
        if (bestIdx2 >= 0) {
          match12[idx1] = bestIdx2;
          nmatches++;
          if (check_orientation) rotHist[RotBin(kps1[idx1].angle, kps2[bestIdx2].angle)].push_back(idx1);
        }
      }
      a++; b++;
    } else if (fv1_node[a] < fv2_node[b]) {
      a = (int)(std::lower_bound(fv1_node, fv1_node + nn1, fv2_node[b]) - fv1_node);
    } else {
      b = (int)(std::lower_bound(fv2_node, fv2_node + nn2, fv1_node[a]) - fv2_node);
    }
  }
  if (check_orientation) {
    int ind1 = -1, ind2 = -1, ind3 = -1;
    ThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (int idx1 : rotHist[i]) { match12[idx1] = -1; nmatches--; }
    }
  }
  return nmatches;
}

}  // extern "C"
