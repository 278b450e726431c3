// tri_core.h — KannalaBrandt8::epipolarConstrain (src/CameraModels/KannalaBrandt8.cpp:216-220) =
// TriangulateMatches(...) > 0.0001f (:306-375) with unproject (:116-143), Triangulate (:394-406) and project (:67-85),
// shared by bow.cu (k_tri_nodes_rig) and the CPU unit harness.
//
// This is the per-candidate gate of ORBmatcher::SearchForTriangulation for a two-camera rig (ORBmatcher.cc:1036-1064).
// It is floating-point geometry: the reference runs it in fp32 through Eigen (JacobiSVD of a 4x4) and libm
// (tanf, atan2f, cos, sin).  Neither can be reproduced bit for bit here (no Eigen in the image; CUDA's libm differs
// from glibc by a few ulp), so this gate carries a TOLERANCE contract, stated in DESIGN.md §2: rays are unprojected in
// fp32 in source order, the homogeneous point is the null vector of A^T A by cyclic Jacobi in fp64, the gates are
// applied in source order with the reference's constants; decisions agree with an fp64 evaluation except for pairs
// that sit within 1e-3 (relative) of a threshold (tests/test_tri_rig.py counts them).
#pragma once
#include "frame_core.h"

struct OrbxRigPose { float R[9], t[3]; };  // R12 (row-major), t12 of one (camera of KF1, camera of KF2) combination

// KannalaBrandt8::unproject (:116-143), fp32 source order; precision = 1e-6
ORBX_HD void orbx_kb8_unproject(const OrbxFrustumCam& C, float u, float v, float* ray) {
  const float pwx = f_div(f_sub(u, C.cx), C.fx), pwy = f_div(f_sub(v, C.cy), C.fy);
  float scale = 1.f;
  float theta_d = f_sqrt(f_add(f_mul(pwx, pwx), f_mul(pwy, pwy)));
  const float hpi = 1.57079637f;  // (float)(CV_PI / 2.f): fminf(fmaxf(-CV_PI/2.f, theta_d), CV_PI/2.f) in double, then float
  theta_d = fminf(fmaxf(-hpi, theta_d), hpi);
  if (theta_d > 1e-8f) {
    float theta = theta_d;
    for (int j = 0; j < 10; j++) {
      const float theta2 = f_mul(theta, theta), theta4 = f_mul(theta2, theta2), theta6 = f_mul(theta4, theta2),
                  theta8 = f_mul(theta4, theta4);
      const float k0 = f_mul(C.k[0], theta2), k1 = f_mul(C.k[1], theta4), k2 = f_mul(C.k[2], theta6), k3 = f_mul(C.k[3], theta8);
      const float num = f_sub(f_mul(theta, f_add(f_add(f_add(f_add(1.f, k0), k1), k2), k3)), theta_d);
      const float den = f_add(f_add(f_add(f_add(1.f, f_mul(3.f, k0)), f_mul(5.f, k1)), f_mul(7.f, k2)), f_mul(9.f, k3));
      const float theta_fix = f_div(num, den);
      theta = f_sub(theta, theta_fix);
      if (fabsf(theta_fix) < 1e-6f) break;
    }
    scale = f_div(tanf(theta), theta_d);
  }
  ray[0] = f_mul(pwx, scale); ray[1] = f_mul(pwy, scale); ray[2] = 1.f;
}

// Null vector of the 4x4 system of KannalaBrandt8::Triangulate (:394-406): eigenvector of the smallest eigenvalue of
// A^T A (= last column of V in A = U S V^T), cyclic Jacobi in fp64.  Returns false when the homogeneous w is 0.
ORBX_HD bool orbx_triangulate_null(const double A[4][4], double* x3D) {
  double M[4][4], V[4][4];
  for (int i = 0; i < 4; i++)
    for (int j = 0; j < 4; j++) {
      double s = 0;
      for (int k = 0; k < 4; k++) s += A[k][i] * A[k][j];
      M[i][j] = s;
      V[i][j] = i == j ? 1.0 : 0.0;
    }
  for (int sweep = 0; sweep < 12; sweep++) {
    double off = 0;
    for (int p = 0; p < 3; p++)
      for (int q = p + 1; q < 4; q++) off += M[p][q] * M[p][q];
    if (off < 1e-300) break;
    for (int p = 0; p < 3; p++)
      for (int q = p + 1; q < 4; q++) {
        const double apq = M[p][q];
        if (apq == 0.0) continue;
        const double tau = (M[q][q] - M[p][p]) / (2.0 * apq);
        const double t = (tau >= 0 ? 1.0 : -1.0) / (fabs(tau) + sqrt(1.0 + tau * tau));
        const double c = 1.0 / sqrt(1.0 + t * t), s = t * c;
        for (int k = 0; k < 4; k++) {
          const double mkp = M[k][p], mkq = M[k][q];
          M[k][p] = c * mkp - s * mkq; M[k][q] = s * mkp + c * mkq;
        }
        for (int k = 0; k < 4; k++) {
          const double mpk = M[p][k], mqk = M[q][k];
          M[p][k] = c * mpk - s * mqk; M[q][k] = s * mpk + c * mqk;
        }
        for (int k = 0; k < 4; k++) {
          const double vkp = V[k][p], vkq = V[k][q];
          V[k][p] = c * vkp - s * vkq; V[k][q] = s * vkp + c * vkq;
        }
      }
  }
  int m = 0;
  for (int i = 1; i < 4; i++)
    if (M[i][i] < M[m][m]) m = i;
  const double w = V[3][m];
  if (w == 0.0) return false;
  x3D[0] = V[0][m] / w; x3D[1] = V[1][m] / w; x3D[2] = V[2][m] / w;
  return true;
}

// KannalaBrandt8::TriangulateMatches (:306-375): returns z1 of the triangulated point (in camera 1) or -1..-5 like the
// reference; p3D = the point in camera-1 coordinates.
ORBX_HD float orbx_kb8_triangulate_matches(const OrbxFrustumCam& C1, const OrbxFrustumCam& C2, float u1, float v1,
                                           float u2, float v2, const OrbxRigPose& T, float sigma1, float sigma2,
                                           float* p3D) {
  float r1[3], r2[3];
  orbx_kb8_unproject(C1, u1, v1, r1);
  orbx_kb8_unproject(C2, u2, v2, r2);
  const float* R = T.R;
  const float r21x = orbx_dot3(R[0], R[1], R[2], r2[0], r2[1], r2[2]);
  const float r21y = orbx_dot3(R[3], R[4], R[5], r2[0], r2[1], r2[2]);
  const float r21z = orbx_dot3(R[6], R[7], R[8], r2[0], r2[1], r2[2]);
  const float n1 = f_sqrt(orbx_dot3(r1[0], r1[1], r1[2], r1[0], r1[1], r1[2]));
  const float n21 = f_sqrt(orbx_dot3(r21x, r21y, r21z, r21x, r21y, r21z));
  const float cosParallax = f_div(orbx_dot3(r1[0], r1[1], r1[2], r21x, r21y, r21z), f_mul(n1, n21));
  if ((double)cosParallax > 0.9998) return -1.f;  // :316
  // Tcw1 = [I | 0], Tcw2 = [R21 | -R21 t12], R21 = R12^T (:330-336)
  double R21[3][3], t21[3];
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) R21[i][j] = (double)R[3 * j + i];
  for (int i = 0; i < 3; i++) t21[i] = -(R21[i][0] * T.t[0] + R21[i][1] * T.t[1] + R21[i][2] * T.t[2]);
  double A[4][4];
  const double p1x = r1[0], p1y = r1[1], p2x = r2[0], p2y = r2[1];
  A[0][0] = -1; A[0][1] = 0; A[0][2] = p1x; A[0][3] = 0;
  A[1][0] = 0; A[1][1] = -1; A[1][2] = p1y; A[1][3] = 0;
  for (int j = 0; j < 3; j++) { A[2][j] = p2x * R21[2][j] - R21[0][j]; A[3][j] = p2y * R21[2][j] - R21[1][j]; }
  A[2][3] = p2x * t21[2] - t21[0]; A[3][3] = p2y * t21[2] - t21[1];
  double X[3];
  if (!orbx_triangulate_null(A, X)) return -2.f;
  const float x = (float)X[0], y = (float)X[1], z1 = (float)X[2];
  if (z1 <= 0) return -2.f;                                                              // :342-345
  const float z2 = (float)(R21[2][0] * X[0] + R21[2][1] * X[1] + R21[2][2] * X[2] + t21[2]);
  if (z2 <= 0) return -3.f;                                                              // :347-350
  float pu, pv;
  orbx_project(C1, x, y, z1, &pu, &pv);
  const float ex1 = f_sub(pu, u1), ey1 = f_sub(pv, v1);
  if ((double)f_add(f_mul(ex1, ex1), f_mul(ey1, ey1)) > d_mul(5.991, (double)sigma1)) return -4.f;  // :358
  const float x2 = (float)(R21[0][0] * X[0] + R21[0][1] * X[1] + R21[0][2] * X[2] + t21[0]);
  const float y2 = (float)(R21[1][0] * X[0] + R21[1][1] * X[1] + R21[1][2] * X[2] + t21[1]);
  orbx_project(C2, x2, y2, z2, &pu, &pv);
  const float ex2 = f_sub(pu, u2), ey2 = f_sub(pv, v2);
  if ((double)f_add(f_mul(ex2, ex2), f_mul(ey2, ey2)) > d_mul(5.991, (double)sigma2)) return -5.f;  // :368
  p3D[0] = x; p3D[1] = y; p3D[2] = z1;
  return z1;
}

// KannalaBrandt8::epipolarConstrain(pCamera2, kp1, kp2, R12, t12, sigmaLevel, unc) (:216-220)
ORBX_HD bool orbx_kb8_epipolar_constrain(const OrbxFrustumCam& C1, const OrbxFrustumCam& C2, float u1, float v1, float u2,
                                         float v2, const OrbxRigPose& T, float sigma1, float sigma2) {
  float p3D[3];
  return orbx_kb8_triangulate_matches(C1, C2, u1, v1, u2, v2, T, sigma1, sigma2, p3D) > 0.0001f;
}
