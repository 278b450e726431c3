// match_core.cuh — descriptor distance and best / second-best bookkeeping shared by match.cu and search.cu.
#pragma once
#include "../../include/orbx.h"
#include "orbx_kernels.cuh"

#define M_WARPS 8
#define KEY_NONE 0xFFFFFFFFu
#define FRAME_GRID_COLS 64  // Frame.h:44
#define FRAME_GRID_ROWS 48  // Frame.h:45

struct Desc256 {
  uint4 a, b;
};
__device__ __forceinline__ Desc256 load_desc(const uint8_t* p) {
  Desc256 d;
  d.a = __ldg(reinterpret_cast<const uint4*>(p));
  d.b = __ldg(reinterpret_cast<const uint4*>(p) + 1);
  return d;
}
// ORBmatcher::DescriptorDistance (ORBmatcher.cc:2058-2074): popcount of the XOR of 8 words
__device__ __forceinline__ int hamming(const Desc256& x, const Desc256& y) {
  return __popc(x.a.x ^ y.a.x) + __popc(x.a.y ^ y.a.y) + __popc(x.a.z ^ y.a.z) + __popc(x.a.w ^ y.a.w) +
         __popc(x.b.x ^ y.b.x) + __popc(x.b.y ^ y.b.y) + __popc(x.b.z ^ y.b.z) + __popc(x.b.w ^ y.b.w);
}

// Running (best, second) of packed keys in one lane, with the reference index of each.
struct Top2 {
  uint32_t k1 = KEY_NONE, k2 = KEY_NONE;
  int i1 = -1, i2 = -1;
  __device__ __forceinline__ void push(uint32_t k, int idx) {
    if (k < k1) { k2 = k1; i2 = i1; k1 = k; i1 = idx; }
    else if (k < k2) { k2 = k; i2 = idx; }
  }
};
// Warp-wide merge: after this every lane holds the global best and second.
__device__ __forceinline__ void warp_top2(Top2& t) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const uint32_t ok1 = __shfl_xor_sync(0xffffffffu, t.k1, o), ok2 = __shfl_xor_sync(0xffffffffu, t.k2, o);
    const int oi1 = __shfl_xor_sync(0xffffffffu, t.i1, o), oi2 = __shfl_xor_sync(0xffffffffu, t.i2, o);
    // keys are unique per query (distinct positions), so merging two sorted pairs is exact
    if (ok1 < t.k1) {
      if (t.k1 < ok2) { t.k2 = t.k1; t.i2 = t.i1; } else { t.k2 = ok2; t.i2 = oi2; }
      t.k1 = ok1; t.i1 = oi1;
    } else if (ok1 < t.k2) {
      t.k2 = ok1; t.i2 = oi1;
    }
  }
}

__device__ __forceinline__ uint32_t make_key(int dist, int pos, int tie_rule) {
  return ((uint32_t)dist << 20) | (uint32_t)(tie_rule == ORBM_TIE_LAST_MIN ? 0xFFFFF - pos : pos);
}

#define GRID_CELLS (FRAME_GRID_COLS * FRAME_GRID_ROWS)
__device__ __forceinline__ int grid_cell(const OrbxKeyPointDev& kp, float min_x, float min_y, float winv,
                                         float hinv) {
  // PosInGrid (Frame.cc:725-735): round() = half away from zero
  const int px = (int)roundf(f_mul(f_sub(kp.x, min_x), winv));
  const int py = (int)roundf(f_mul(f_sub(kp.y, min_y), hinv));
  if (px < 0 || px >= FRAME_GRID_COLS || py < 0 || py >= FRAME_GRID_ROWS) return -1;
  return px * FRAME_GRID_ROWS + py;
}

