// frame_core.h — per-point Frame geometry shared by frame.cu and the CPU unit harness (tests/cpu_harness):
//   Frame::UndistortKeyPoints  src/Frame.cc:747-781  (cv::undistortPoints with R = I, P = K)
//   Frame::isInFrustum         src/Frame.cc:512-586  (Nleft == -1) and isInFrustumChecks :1168-1242 (fisheye rig)
//   MapPoint::PredictScale     src/MapPoint.cc:531-546
//
// Floating-point contract:
//   * cv::undistortPoints runs in double; OpenCV's baseline build does not contract: explicit uncontracted double
//     operations in cvUndistortPointsInternal's source order, 5 fixed-point iterations, result rounded to float once.
//     Pinned bit for bit against cv2.undistortPoints (tests/test_frame_geometry.py).
//   * isInFrustum is Eigen fp32 code in the reference.  Eigen is not in the image, so its compile cannot be pinned;
//     the contract here is uncontracted fp32 in Eigen's fixed-size evaluation order: a 3-term dot product is
//     c0 + (c1 + c2) (redux_novec_unroller halves 3 into 1 + 2), R*P + t adds t last, norm() = sqrtf(squaredNorm()).
//     logf is glibc's (orbx_logf, exhaustive sweep).  The KannalaBrandt8 projection uses atan2f/cos/sin of libm in the
//     reference; the device versions agree to a few ulp, so that branch carries a tolerance (1e-3 px), not bit parity.
#pragma once
#include "orbx_core.h"

#if defined(__CUDA_ARCH__)
ORBX_HD double d_div(double a, double b) { return __ddiv_rn(a, b); }
ORBX_HD float f_sqrt(float a) { return __fsqrt_rn(a); }
#else
ORBX_HD double d_div(double a, double b) { volatile double r = a / b; return r; }
ORBX_HD float f_sqrt(float a) { return sqrtf(a); }
#endif
ORBX_HD double d_sub(double a, double b) { return d_add(a, -b); }

// cv::undistortPoints for one point: K = (fx,fy,cx,cy) as floats widened to double, k[12] distortion coefficients
// (k1,k2,p1,p2,k3,k4,k5,k6,s1,s2,s3,s4; missing ones 0), R = I, P = K.  calib3d/src/undistort.dispatch.cpp,
// cvUndistortPointsInternal with the default criteria (5 iterations, no early exit for COUNT-only criteria).
ORBX_HD void orbx_undistort_point(float u, float v, float fxf, float fyf, float cxf, float cyf, const double* k,
                                  float* ou, float* ov) {
  const double fx = (double)fxf, fy = (double)fyf, cx = (double)cxf, cy = (double)cyf;
  const double ifx = d_div(1.0, fx), ify = d_div(1.0, fy);
  double x = d_mul(d_sub((double)u, cx), ifx), y = d_mul(d_sub((double)v, cy), ify);
  const double x0 = x, y0 = y;
  for (int j = 0; j < 5; j++) {
    const double r2 = d_add(d_mul(x, x), d_mul(y, y));
    const double num = d_add(1.0, d_mul(d_add(d_mul(d_add(d_mul(k[7], r2), k[6]), r2), k[5]), r2));
    const double den = d_add(1.0, d_mul(d_add(d_mul(d_add(d_mul(k[4], r2), k[1]), r2), k[0]), r2));
    double icdist = d_div(num, den);
    if (icdist < 0) { x = d_mul(d_sub((double)u, cx), ifx); y = d_mul(d_sub((double)v, cy), ify); break; }
    const double xx2 = d_mul(d_mul(2.0, x), x), yy2 = d_mul(d_mul(2.0, y), y);
    const double dX = d_add(d_add(d_add(d_mul(d_mul(d_mul(2.0, k[2]), x), y), d_mul(k[3], d_add(r2, xx2))),
                                  d_mul(k[8], r2)), d_mul(d_mul(k[9], r2), r2));
    const double dY = d_add(d_add(d_add(d_mul(k[2], d_add(r2, yy2)), d_mul(d_mul(d_mul(2.0, k[3]), x), y)),
                                  d_mul(k[10], r2)), d_mul(d_mul(k[11], r2), r2));
    x = d_mul(d_sub(x0, dX), icdist);
    y = d_mul(d_sub(y0, dY), icdist);
  }
  // RR = P * I: xx = fx*x + 0*y + cx, yy = 0*x + fy*y + cy, ww = 1/(0*x + 0*y + 1)
  const double xx = d_add(d_add(d_mul(fx, x), d_mul(0.0, y)), cx);
  const double yy = d_add(d_add(d_mul(0.0, x), d_mul(fy, y)), cy);
  const double ww = d_div(1.0, d_add(d_add(d_mul(0.0, x), d_mul(0.0, y)), 1.0));
  *ou = (float)d_mul(xx, ww);
  *ov = (float)d_mul(yy, ww);
}

// Eigen fixed-size 3-term dot product order
ORBX_HD float orbx_dot3(float a0, float a1, float a2, float b0, float b1, float b2) {
  return f_add(f_mul(a0, b0), f_add(f_mul(a1, b1), f_mul(a2, b2)));
}

struct OrbxFrustumCam {   // GeometricCamera parameters: model 0 = Pinhole, 1 = KannalaBrandt8
  int model;
  float fx, fy, cx, cy, k[4];
};

struct OrbxFrustumPose {  // R (row-major), t: world -> camera; c: camera centre in the world
  float R[9], t[3], c[3];
};

struct OrbxFrustumOut {
  int in_view;            // the return value (mbTrackInView / mbTrackInViewR)
  float proj_x, proj_y;   // mTrackProjX/Y (Nleft == -1: also set when a later gate fails, as the reference does)
  float proj_xr;          // mTrackProjXR = u - mbf/z (Nleft == -1 only)
  float depth;            // mTrackDepth = |Pc|
  float view_cos;         // mTrackViewCos
  int level;              // mnTrackScaleLevel
};

#if defined(__CUDA_ARCH__)
ORBX_HD float orbx_atan2f(float y, float x) { return atan2f(y, x); }
ORBX_HD float orbx_cosf(float a) { return cosf(a); }
ORBX_HD float orbx_sinf(float a) { return sinf(a); }
#else
ORBX_HD float orbx_atan2f(float y, float x) { return atan2f(y, x); }
ORBX_HD float orbx_cosf(float a) { return cosf(a); }
ORBX_HD float orbx_sinf(float a) { return sinf(a); }
#endif

// GeometricCamera::project(Eigen::Vector3f): Pinhole.cpp:49-55, KannalaBrandt8.cpp:67-85
ORBX_HD void orbx_project(const OrbxFrustumCam& C, float X, float Y, float Z, float* u, float* v) {
  if (C.model == 0) {
    *u = f_add(f_div(f_mul(C.fx, X), Z), C.cx);
    *v = f_add(f_div(f_mul(C.fy, Y), Z), C.cy);
    return;
  }
  const float x2y2 = f_add(f_mul(X, X), f_mul(Y, Y));
  const float theta = orbx_atan2f(f_sqrt(x2y2), Z);
  const float psi = orbx_atan2f(Y, X);
  const float t2 = f_mul(theta, theta), t3 = f_mul(theta, t2), t5 = f_mul(t3, t2), t7 = f_mul(t5, t2), t9 = f_mul(t7, t2);
  const float r = f_add(f_add(f_add(f_add(theta, f_mul(C.k[0], t3)), f_mul(C.k[1], t5)), f_mul(C.k[2], t7)), f_mul(C.k[3], t9));
  // KannalaBrandt8.cpp:81-82: `cos(psi)` / `sin(psi)` are the double functions there, so fx * r (float) times the cosine
  // plus cx is evaluated in double and stored as float (tolerance contract: the device's double cos / sin are not
  // glibc's; after the rounding to float the results agree except for rare double-rounding cases).
  *u = (float)d_add(d_mul((double)f_mul(C.fx, r), cos((double)psi)), (double)C.cx);
  *v = (float)d_add(d_mul((double)f_mul(C.fy, r), sin((double)psi)), (double)C.cy);
}

// One MapPoint against one camera of a frame.  fisheye_rig = 0: Frame::isInFrustum with Nleft == -1 (:514-571);
// 1: Frame::isInFrustumChecks (:1168-1242; the caller passes the left or the right camera's pose).
ORBX_HD OrbxFrustumOut orbx_in_frustum(const float* P, const float* Pn, float min_dist_raw, float max_dist_raw,
                                       const OrbxFrustumPose& T, const OrbxFrustumCam& C, float mbf, float min_x,
                                       float max_x, float min_y, float max_y, float cos_limit, float log_sf,
                                       int n_levels, int fisheye_rig) {
  OrbxFrustumOut o;
  o.in_view = 0; o.proj_x = -1.0f; o.proj_y = -1.0f; o.proj_xr = -1.0f; o.depth = 0.0f; o.view_cos = 0.0f; o.level = -1;
  const float Pcx = f_add(orbx_dot3(T.R[0], T.R[1], T.R[2], P[0], P[1], P[2]), T.t[0]);
  const float Pcy = f_add(orbx_dot3(T.R[3], T.R[4], T.R[5], P[0], P[1], P[2]), T.t[1]);
  const float Pcz = f_add(orbx_dot3(T.R[6], T.R[7], T.R[8], P[0], P[1], P[2]), T.t[2]);
  const float pc_dist = f_sqrt(orbx_dot3(Pcx, Pcy, Pcz, Pcx, Pcy, Pcz));
  const float invz = f_div(1.0f, Pcz);
  if (Pcz < 0.0f) return o;
  float u, v;
  orbx_project(C, Pcx, Pcy, Pcz, &u, &v);
  if (u < min_x || u > max_x) return o;
  if (v < min_y || v > max_y) return o;
  if (!fisheye_rig) { o.proj_x = u; o.proj_y = v; }
  const float max_d = f_mul(1.2f, max_dist_raw), min_d = f_mul(0.8f, min_dist_raw);  // MapPoint.cc:505,511
  const float POx = f_sub(P[0], T.c[0]), POy = f_sub(P[1], T.c[1]), POz = f_sub(P[2], T.c[2]);
  const float dist = f_sqrt(orbx_dot3(POx, POy, POz, POx, POy, POz));
  if (dist < min_d || dist > max_d) return o;
  const float view_cos = f_div(orbx_dot3(POx, POy, POz, Pn[0], Pn[1], Pn[2]), dist);
  if (view_cos < cos_limit) return o;
  // MapPoint::PredictScale: ratio = mfMaxDistance / dist (the raw member, not the 1.2x getter)
  const float ratio = f_div(max_dist_raw, dist);
  int n_scale = (int)ceilf(f_div(orbx_logf(ratio), log_sf));
  if (n_scale < 0) n_scale = 0;
  else if (n_scale >= n_levels) n_scale = n_levels - 1;
  o.in_view = 1;
  o.proj_x = u; o.proj_y = v;
  o.proj_xr = fisheye_rig ? -1.0f : f_sub(u, f_mul(mbf, invz));
  o.depth = pc_dist;
  o.view_cos = view_cos;
  o.level = n_scale;
  return o;
}
