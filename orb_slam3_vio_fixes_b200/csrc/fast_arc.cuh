// fast_arc.cuh — arc strength of FAST-9/16 for two horizontally adjacent pixels at once (packed 16-bit SIMD),
// shared by the cell-per-warp kernel (fast_cell.cu) and the strip kernel (fast.cu).
// cv::FAST TYPE_9_16 / cornerScore: A = max over the 16 arcs of 9 of max(min(v-ring), min(ring-v));
// corner(th) <=> A > th, cornerScore == A-1.
#pragma once
#include <stdint.h>

// 16-bit lanes {w[A], w[A+1]} out of the 12-byte window w0|w1|w2 (A compile-time, 1..9)
template <int A>
__device__ __forceinline__ uint32_t pick2(uint32_t w0, uint32_t w1, uint32_t w2) {
  if (A <= 2) return __byte_perm(w0, 0, A | 0x40 | ((A + 1) << 8) | 0x4000);
  if (A == 3) return __byte_perm(w0, w1, 0x0403) & 0x00ff00ffu;
  if (A <= 6) return __byte_perm(w1, 0, (A - 4) | 0x40 | ((A - 3) << 8) | 0x4000);
  if (A == 7) return __byte_perm(w1, w2, 0x0403) & 0x00ff00ffu;
  return __byte_perm(w2, 0, (A - 8) | 0x40 | ((A - 7) << 8) | 0x4000);
}

// Arc strength of the two pixels at window bytes 4+J and 5+J.  rw[d] = the three words (q-1,q,q+1)
// of tile row (r-3+d).  Returns A0 | A1<<8.
template <int J>
__device__ __forceinline__ uint32_t arc_pair(const uint32_t (&rw)[7][3]) {
  uint32_t r[16];
#define RING(k, dx, dy) r[k] = pick2<4 + J + (dx)>(rw[(dy) + 3][0], rw[(dy) + 3][1], rw[(dy) + 3][2])
  // OpenCV ring order (features2d/fast_score.cpp makeOffsets, patternSize 16)
  RING(0, 0, 3);   RING(1, 1, 3);    RING(2, 2, 2);    RING(3, 3, 1);
  RING(4, 3, 0);   RING(5, 3, -1);   RING(6, 2, -2);   RING(7, 1, -3);
  RING(8, 0, -3);  RING(9, -1, -3);  RING(10, -2, -2); RING(11, -3, -1);
  RING(12, -3, 0); RING(13, -3, 1);  RING(14, -2, 2);  RING(15, -1, 3);
#undef RING
  const uint32_t vv = pick2<4 + J>(rw[3][0], rw[3][1], rw[3][2]);
  uint32_t mn3[16], mx3[16];
#pragma unroll
  for (int k = 0; k < 16; k++) {
    mn3[k] = __vimin3_u16x2(r[k], r[(k + 1) & 15], r[(k + 2) & 15]);
    mx3[k] = __vimax3_u16x2(r[k], r[(k + 1) & 15], r[(k + 2) & 15]);
  }
  // windows of 9 = three consecutive windows of 3; keep only the extreme over the 16 arcs:
  //   bright-centre strength = v - min_k max9[k],  dark-centre strength = max_k min9[k] - v
  uint32_t lo = 0xffffffffu, hi = 0;
#pragma unroll
  for (int k = 0; k < 16; k += 2) {
    const uint32_t mx9a = __vimax3_u16x2(mx3[k], mx3[(k + 3) & 15], mx3[(k + 6) & 15]);
    const uint32_t mx9b = __vimax3_u16x2(mx3[k + 1], mx3[(k + 4) & 15], mx3[(k + 7) & 15]);
    const uint32_t mn9a = __vimin3_u16x2(mn3[k], mn3[(k + 3) & 15], mn3[(k + 6) & 15]);
    const uint32_t mn9b = __vimin3_u16x2(mn3[k + 1], mn3[(k + 4) & 15], mn3[(k + 7) & 15]);
    lo = __vimin3_u16x2(lo, mx9a, mx9b);
    hi = __vimax3_u16x2(hi, mn9a, mn9b);
  }
  const int v0 = (int)(vv & 0xffffu), v1 = (int)(vv >> 16);
  const int a0 = max(max(v0 - (int)(lo & 0xffffu), (int)(hi & 0xffffu) - v0), 0);
  const int a1 = max(max(v1 - (int)(lo >> 16), (int)(hi >> 16) - v1), 0);
  return (uint32_t)a0 | ((uint32_t)a1 << 8);
}

