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
