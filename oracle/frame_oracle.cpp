// frame_oracle.cpp — CPU ORACLE, part 2: per-point Frame geometry and the fisheye (two-camera rig) branches
// of the matcher.  TEST INFRASTRUCTURE ONLY (same rules as orb_oracle.cpp: only tests/, smoke() and bench.py's
// CPU legs may use it; nothing under orb_slam3_vio_fixes_b200/ links or calls it).
//
// Restates, in plain C++ with libm:
//   Frame::UndistortKeyPoints         src/Frame.cc:747-781   (cv::undistortPoints: double, source order of
//                                     opencv calib3d undistort.dispatch.cpp cvUndistortPointsInternal, 5 iterations)
//   Frame::isInFrustum                src/Frame.cc:512-586, isInFrustumChecks :1168-1242
//   MapPoint::PredictScale            src/MapPoint.cc:531-546 (libm logf)
//   GeometricCamera::project          src/CameraModels/Pinhole.cpp:49-55, KannalaBrandt8.cpp:67-85 (libm atan2f/cos/sin)
//
// PINNING: undistortPoints is pinned bit for bit against the real OpenCV 4.13 binary (cv2.undistortPoints,
// tests/test_frame_geometry.py).  isInFrustum is Eigen code; Eigen is absent from the image, so it is
// "parity unpinned" against an executable reference: the restatement follows the source with uncontracted fp32
// in Eigen's fixed-size evaluation order (3-term dot = c0 + (c1 + c2)) and is cross-checked against an fp64
// evaluation (the test reports how many decisions / levels differ).
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
  *u = cam[0] * r * cosf(psi) + cam[2];
  *v = cam[1] * r * sinf(psi) + cam[3];
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
