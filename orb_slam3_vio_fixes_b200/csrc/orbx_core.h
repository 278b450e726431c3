// orbx_core.h — arithmetic shared by the kernels (and compiled for the host by the CPU unit
// harness in tests/cpu_harness, which is how the device functions are checked without a GPU).
//
// Everything here is the exact integer / fp32 arithmetic of the reference path
// (src/ORBextractor.cc) and of the OpenCV primitives it calls; see DESIGN.md for the map.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define ORBX_HD __host__ __device__ __forceinline__
#else
#define ORBX_HD inline
#endif

#define ORBX_EDGE 19          // EDGE_THRESHOLD, ORBextractor.cc:73
#define ORBX_HALF_PATCH 15    // HALF_PATCH_SIZE, :72
#define ORBX_MIN_BORDER 16    // EDGE_THRESHOLD-3, :789
#define ORBX_CELL_W 35.0f     // W, :785

// ---- exactly-rounded fp32 helpers (never contracted to FMA) --------------------------------
#if defined(__CUDA_ARCH__)
ORBX_HD float f_mul(float a, float b) { return __fmul_rn(a, b); }
ORBX_HD float f_add(float a, float b) { return __fadd_rn(a, b); }
ORBX_HD float f_sub(float a, float b) { return __fsub_rn(a, b); }
ORBX_HD float f_div(float a, float b) { return __fdiv_rn(a, b); }
ORBX_HD float f_fma(float a, float b, float c) { return __fmaf_rn(a, b, c); }
ORBX_HD double d_mul(double a, double b) { return __dmul_rn(a, b); }
ORBX_HD double d_add(double a, double b) { return __dadd_rn(a, b); }
ORBX_HD int f_round_int(float v) { return __float2int_rn(v); }  // cvRound: half to even
#else
#include <math.h>
// host build must use -ffp-contract=off
ORBX_HD float f_mul(float a, float b) { volatile float r = a * b; return r; }
ORBX_HD float f_add(float a, float b) { volatile float r = a + b; return r; }
ORBX_HD float f_sub(float a, float b) { volatile float r = a - b; return r; }
ORBX_HD float f_div(float a, float b) { volatile float r = a / b; return r; }
ORBX_HD float f_fma(float a, float b, float c) { return fmaf(a, b, c); }
ORBX_HD double d_mul(double a, double b) { volatile double r = a * b; return r; }
ORBX_HD double d_add(double a, double b) { volatile double r = a + b; return r; }
ORBX_HD int f_round_int(float v) { return (int)lrintf(v); }
#endif

// ---- cv::fastAtan2 (scalar path), fp32 Horner without FMA; used by IC_Angle (:102) ----------
ORBX_HD float orbx_fast_atan2(float y, float x) {
  const float k = 57.295780181884765625f;  // (float)(180/CV_PI)
  const float p1 = 0.9997878412794807f * k, p3 = -0.3258083974640975f * k;
  const float p5 = 0.1555786518463281f * k, p7 = -0.04432655554792128f * k;
  const float eps = 2.2204460492503131e-16f;  // (float)DBL_EPSILON
  float ax = x < 0 ? -x : x, ay = y < 0 ? -y : y;
  float a, c, c2;
  if (ax >= ay) {
    c = f_div(ay, f_add(ax, eps));
    c2 = f_mul(c, c);
    a = f_mul(f_add(f_mul(f_add(f_mul(f_add(f_mul(p7, c2), p5), c2), p3), c2), p1), c);
  } else {
    c = f_div(ax, f_add(ay, eps));
    c2 = f_mul(c, c);
    a = f_sub(90.f, f_mul(f_add(f_mul(f_add(f_mul(f_add(f_mul(p7, c2), p5), c2), p3), c2), p1), c));
  }
  if (x < 0) a = f_sub(180.f, a);
  if (y < 0) a = f_sub(360.f, a);
  return a;
}

// ---- sincosf, bit-identical to glibc 2.39's sincosf on [0, 120) ------------------------------
// The reference calls cos(float)/sin(float) (:112) = libm sincosf.  This is the same algorithm
// (double-precision range reduction by pi/2 and two degree-4/3 polynomials in x^2) written with
// explicit, uncontracted double operations; an exhaustive sweep of all 1.09e9 floats in
// [0, 6.3] against libm shows zero differences (tests/test_sincos.py).
ORBX_HD void orbx_sincosf(float y, float* sinp, float* cosp) {
  const double hpi_inv = 0x1.45F306DC9C883p+23, hpi = 0x1.921FB54442D18p0;
  const double c0 = 0x1p0, c1 = -0x1.ffffffd0c621cp-2, c2c = 0x1.55553e1068f19p-5;
  const double c3 = -0x1.6c087e89a359dp-10, c4 = 0x1.99343027bf8c3p-16;
  const double s1c = -0x1.555545995a603p-3, s2 = 0x1.1107605230bc4p-7, s3 = -0x1.994eb3774cf24p-13;
  union { float f; uint32_t u; } cvt;
  cvt.f = y;
  const uint32_t top12 = (cvt.u >> 20) & 0x7ff;
  double x = (double)y;
  int n = 0;
  double sgn_c = 1.0;  // sign of the cosine polynomial (second table of glibc = negated c's)
  if (top12 < 0x3f4) {             // |y| < pi/4  (abstop12(0x1.921FB6p-1f) == 0x3f4)
    if (top12 < 0x398) {           // |y| < 2^-12
      *sinp = y;
      *cosp = 1.0f;
      return;
    }
  } else {
    double r = d_mul(x, hpi_inv);
    n = ((int32_t)r + 0x800000) >> 24;
    x = d_add(x, -d_mul((double)n, hpi));
    const double sign_tab = (n & 3) == 1 || (n & 3) == 2 ? -1.0 : 1.0;  // {1,-1,-1,1}[n&3]
    if (n & 2) sgn_c = -1.0;
    const double x2r = d_mul(x, x);
    x = d_mul(x, sign_tab);
    // fallthrough with x2 = x*x of the unsigned x (same value)
    const double x2 = x2r;
    const double x4 = d_mul(x2, x2), x3 = d_mul(x2, x);
    const double pc2 = d_add(d_mul(sgn_c, c3), d_mul(x2, d_mul(sgn_c, c4)));
    const double ps1 = d_add(s2, d_mul(x2, s3));
    const double pc1 = d_add(d_mul(sgn_c, c0), d_mul(x2, d_mul(sgn_c, c1)));
    const double x5 = d_mul(x3, x2), x6 = d_mul(x4, x2);
    const double s = d_add(x, d_mul(x3, s1c));
    const double c = d_add(pc1, d_mul(x4, d_mul(sgn_c, c2c)));
    const float sv = (float)d_add(s, d_mul(x5, ps1));
    const float cv = (float)d_add(c, d_mul(x6, pc2));
    if (n & 1) { *sinp = cv; *cosp = sv; } else { *sinp = sv; *cosp = cv; }
    return;
  }
  {
    const double x2 = d_mul(x, x);
    const double x4 = d_mul(x2, x2), x3 = d_mul(x2, x);
    const double pc2 = d_add(c3, d_mul(x2, c4));
    const double ps1 = d_add(s2, d_mul(x2, s3));
    const double pc1 = d_add(c0, d_mul(x2, c1));
    const double x5 = d_mul(x3, x2), x6 = d_mul(x4, x2);
    const double s = d_add(x, d_mul(x3, s1c));
    const double c = d_add(pc1, d_mul(x4, c2c));
    *sinp = (float)d_add(s, d_mul(x5, ps1));
    *cosp = (float)d_add(c, d_mul(x6, pc2));
  }
}

// ---- FAST 9/16 arc strength ---------------------------------------------------------------------
// ring[k]: the 16 Bresenham-circle pixels in OpenCV's order; v: centre.  Returns
//   A = max over the 16 contiguous arcs of 9 of max( min(v-ring), min(ring-v) )   (>= 0 clamp)
// so that  corner(th) <=> A > th  and cv::FAST's cornerScore == A-1 (features2d/fast_score.cpp).
ORBX_HD int orbx_arc_strength(int v, const int* ring /*[16]*/) {
  int mn2[16], mx2[16];
#pragma unroll
  for (int k = 0; k < 16; k++) {
    int a = ring[k], b = ring[(k + 1) & 15];
    mn2[k] = a < b ? a : b;
    mx2[k] = a > b ? a : b;
  }
  int mn4[16], mx4[16];
#pragma unroll
  for (int k = 0; k < 16; k++) {
    int a = mn2[k], b = mn2[(k + 2) & 15];
    mn4[k] = a < b ? a : b;
    a = mx2[k]; b = mx2[(k + 2) & 15];
    mx4[k] = a > b ? a : b;
  }
  int best = 0;
#pragma unroll
  for (int k = 0; k < 16; k++) {
    int a = mn4[k], b = mn4[(k + 4) & 15];
    int mn8 = a < b ? a : b;
    a = mx4[k]; b = mx4[(k + 4) & 15];
    int mx8 = a > b ? a : b;
    int r8 = ring[(k + 8) & 15];
    int mn9 = mn8 < r8 ? mn8 : r8;  // darkest... smallest ring value on the arc
    int mx9 = mx8 > r8 ? mx8 : r8;
    int bright = v - mx9;  // min over arc of (v - ring)
    int dark = mn9 - v;    // min over arc of (ring - v)
    int s = bright > dark ? bright : dark;
    best = s > best ? s : best;
  }
  return best;
}

// BORDER_REFLECT_101 for |overshoot| < len
ORBX_HD int orbx_reflect101(int p, int len) {
  if (p < 0) p = -p;
  if (p >= len) p = 2 * (len - 1) - p;
  return p;
}

// ---- logf, bit-identical to glibc 2.39's logf for positive normal floats ------------------------
// MapPoint::PredictScale (src/MapPoint.cc:531-546) takes `log(ratio)` of a float with std::log(float)
// in scope (TemplatedVocabulary.h's global `using namespace std`), i.e. libm logf.  Same algorithm
// (16-entry table, degree-3 polynomial in double) in explicit uncontracted double operations; the
// sweep over all 2,130,706,432 positive normal floats against libm shows zero differences
// (tests/test_harness.py::test_logf_matches_libm_exhaustively).
ORBX_HD float orbx_logf(float xf) {
  const double T[16][2] = {
      {0x1.661ec79f8f3bep+0, -0x1.57bf7808caadep-2}, {0x1.571ed4aaf883dp+0, -0x1.2bef0a7c06ddbp-2},
      {0x1.49539f0f010bp+0, -0x1.01eae7f513a67p-2},  {0x1.3c995b0b80385p+0, -0x1.b31d8a68224e9p-3},
      {0x1.30d190c8864a5p+0, -0x1.6574f0ac07758p-3}, {0x1.25e227b0b8eap+0, -0x1.1aa2bc79c81p-3},
      {0x1.1bb4a4a1a343fp+0, -0x1.a4e76ce8c0e5ep-4}, {0x1.12358f08ae5bap+0, -0x1.1973c5a611cccp-4},
      {0x1.0953f419900a7p+0, -0x1.252f438e10c1ep-5}, {0x1p+0, 0x0p+0},
      {0x1.e608cfd9a47acp-1, 0x1.aa5aa5df25984p-5},  {0x1.ca4b31f026aap-1, 0x1.c5e53aa362eb4p-4},
      {0x1.b2036576afce6p-1, 0x1.526e57720db08p-3},  {0x1.9c2d163a1aa2dp-1, 0x1.bc2860d22477p-3},
      {0x1.886e6037841edp-1, 0x1.1058bc8a07ee1p-2},  {0x1.767dcf5534862p-1, 0x1.4043057b6ee09p-2}};
  const double Ln2 = 0x1.62e42fefa39efp-1;
  const double A0 = -0x1.00ea348b88334p-2, A1 = 0x1.5575b0be00b6ap-2, A2 = -0x1.ffffef20a4123p-2;
  union { float f; uint32_t u; } cvt;
  cvt.f = xf;
  const uint32_t ix = cvt.u;
  if (ix == 0x3f800000u) return 0.0f;
  const uint32_t tmp = ix - 0x3f330000u;
  const int i = (int)((tmp >> 19) & 15u);
  const int k = (int32_t)tmp >> 23;
  cvt.u = ix - (tmp & 0xff800000u);
  const double z = (double)cvt.f;
  const double r = d_add(d_mul(z, T[i][0]), -1.0);
  const double y0 = d_add(T[i][1], d_mul((double)k, Ln2));
  const double r2 = d_mul(r, r);
  double y = d_add(d_mul(A1, r), A2);
  y = d_add(d_mul(A0, r2), y);
  y = d_add(d_mul(y, r2), d_add(y0, r));
  return (float)y;
}
