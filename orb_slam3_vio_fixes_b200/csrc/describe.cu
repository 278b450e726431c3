// describe.cu — orientation (IC_Angle, src/ORBextractor.cc:76-103), the 7x7 sigma=2 Gaussian
// blur (:1131-1133) evaluated only where descriptors need it, the 256-bit rotated BRIEF
// descriptor (:107-146) and the final keypoint record (:880-890, :1143-1164), one warp per
// keypoint.
//
// The reference blurs the whole level and then samples 512 points within radius 19 of each
// keypoint.  Here each warp stages the 45x45 source neighbourhood of its keypoint (39x39 sampled
// window + 3-px blur halo, BORDER_REFLECT_101 at the level edge), runs the separable Q8 filter
// on it in shared memory — identical integer arithmetic, out = (sum + 32768) >> 16 — and samples
// from that.  No blurred level is ever written to HBM.
#include <math.h>

#include <vector>

#include "orbx_kernels.cuh"

#define DESC_WARPS 4
#define PATCH 45        // 2*22+1
#define PATCH_R 22
#define PW 13           // words per staged row: 52 bytes cover (alignment offset <= 3) + 45
#define BL 39           // blurred window 2*19+1
#define BW 11           // words per blurred row (44 bytes)

// bit_pattern_31_ as floats (x0,y0,x1,y1 per test), read through the L1 with 128-bit loads
__device__ const __align__(16) float c_patf[1024] = {
#include "orb_pattern.inc"
};

// Shared memory of one warp.  Column coordinates are "staged columns": staged byte c of row r is
// image pixel (ky-22+r, kx-22+c-off), off = alignment offset of the row start (0 on the border path).
//   patch : 45 rows x 13 words of source pixels; after the column pass it is dead and receives the
//           blurred 39x39 window (row pitch 11 words, same staged-column coordinates).
//   t     : column-pass sums for 39 rows x 13 words, two packed 16-bit lanes per register:
//           t[r][w].x = {T(4w), T(4w+2)}, .y = {T(4w+1), T(4w+3)}   (sums <= 255*256 fit 16 bits)
struct __align__(16) DescWarpSmem {
  uint32_t patch[PATCH * PW + 3];
  uint2 t[BL * PW];
};

// Row-pass tasks (row << 4 | word) that cover the blurred pixels BRIEF can sample: the pattern points lie
// within radius R = max |p| (18.4) of the keypoint, so after rotation and cvRound only pixels whose
// unit square meets that disc are read — 74 % of the 39x39 window.  One list per alignment offset.
#define ROW_TASKS_MAX 448
__device__ uint16_t c_row_tasks[4 * ROW_TASKS_MAX];
__device__ int c_row_ntasks[4];

// One output byte of the row pass: taps 18,34,48,56,48,34,18 over seven 16-bit sums, rounded
// like cv::GaussianBlur's fixed-point path: (acc + 2^15) >> 16.
__device__ __forceinline__ uint32_t blur_out(uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t a4,
                                             uint32_t a5, uint32_t a6) {
  return (18u * (a0 + a6) + 34u * (a1 + a5) + 48u * (a2 + a4) + 56u * a3 + 32768u) >> 16;
}

template <bool STRICT_FP>
__global__ void __launch_bounds__(DESC_WARPS * 32)
k_describe(const __grid_constant__ DevPlan P, Level0 l0, const uint8_t* __restrict__ arena,
           const uint32_t* __restrict__ sel, const int32_t* __restrict__ sel_cnt,
           const int32_t* __restrict__ dst, OrbxKeyPointDev* __restrict__ kps, uint8_t* __restrict__ desc,
           int capacity) {
  __shared__ DescWarpSmem sm_all[DESC_WARPS];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int frame = blockIdx.y;
  const int slot = blockIdx.x * DESC_WARPS + warp;
  if (slot >= P.kp_total_cap) return;
  int level = 0;
#pragma unroll 1
  for (int l = 1; l < P.nlevels; l++)
    if (slot >= P.lv[l].kp_base) level = l;
  const DevLevel& L = P.lv[level];
  const int idx = slot - L.kp_base;
  if (idx >= sel_cnt[frame * P.nlevels + level]) return;
  DescWarpSmem& sm = sm_all[warp];

  const uint32_t key = sel[(long long)frame * P.kp_total_cap + slot];
  const int kx = ORBX_KEY_X(key), ky = ORBX_KEY_Y(key);
  int pitch;
  const uint8_t* img;
  if (level == 0) { pitch = l0.pitch; img = l0.base + (long long)frame * l0.frame_stride; }
  else { pitch = L.pitch; img = arena + (long long)frame * P.arena_bytes + L.offset; }

  // ---- stage the 45x45 neighbourhood (13 words per row)
  int off;
  {
    const uint8_t* row0 = img + (long long)(ky - PATCH_R) * pitch + (kx - PATCH_R);
    const bool interior = (kx - PATCH_R >= 4) && (kx + PATCH_R + 8 < L.w) && (ky - PATCH_R >= 0) &&
                          (ky + PATCH_R < L.h) && ((pitch & 3) == 0);
    if (interior) {
      off = (int)((uintptr_t)row0 & 3);
      // word aligned; 13 words cover off+45 <= 48 bytes
      const uint32_t* a0 = reinterpret_cast<const uint32_t*>(row0 - off);
      const int wpitch = pitch >> 2;
      uint32_t v[19];
#pragma unroll
      for (int it = 0; it < 19; it++) {
        const int i = it * 32 + lane;
        const int r = (i * 5042) >> 16, c = i - r * PW;  // i / 13 for i < 608
        v[it] = (it < 18 || lane < PATCH * PW - 18 * 32) ? __ldg(a0 + (r * wpitch + c)) : 0u;
      }
#pragma unroll
      for (int it = 0; it < 19; it++)
        if (it < 18 || lane < PATCH * PW - 18 * 32) sm.patch[it * 32 + lane] = v[it];
    } else {
      // near the level edge (or unaligned rows): bytes, BORDER_REFLECT_101; lane = column(s)
      off = 0;
      uint8_t* pb = reinterpret_cast<uint8_t*>(sm.patch);
      const int c1 = min(lane + 32, PATCH - 1);
      const int sx0 = orbx_reflect101(kx - PATCH_R + lane, L.w), sx1 = orbx_reflect101(kx - PATCH_R + c1, L.w);
      for (int r = 0; r < PATCH; r++) {
        const uint8_t* rowp = img + (long long)orbx_reflect101(ky - PATCH_R + r, L.h) * pitch;
        const uint8_t p0 = rowp[sx0], p1 = rowp[sx1];
        pb[r * (PW * 4) + lane] = p0;
        pb[r * (PW * 4) + c1] = p1;
      }
    }
  }
  __syncwarp();
  const uint8_t* pc = reinterpret_cast<const uint8_t*>(sm.patch) + off;  // pc[r*52 + c]

  // ---- IC_Angle: integer moments over the circular patch of radius 15
  int m01 = 0, m10 = 0;
  if (lane < 31) {
    const int v = lane - ORBX_HALF_PATCH;
    const int d = P.umax[v < 0 ? -v : v];
    const uint8_t* rowp = pc + (PATCH_R + v) * (PW * 4) + PATCH_R;
    int s = 0, su = 0;
    for (int u = -d; u <= d; u++) {
      const int val = rowp[u];
      s += val;
      su += u * val;
    }
    m10 = su;
    m01 = v * s;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    m01 += __shfl_xor_sync(0xffffffffu, m01, o);
    m10 += __shfl_xor_sync(0xffffffffu, m10, o);
  }
  const float angle = orbx_fast_atan2((float)m01, (float)m10);

  // ---- separable 7-tap Q8 blur of the staged patch (exact integers, so the pass order is free).
  // Column pass: lane = (word column w, half), 4 pixels per word as two packed 16-bit pairs,
  // sliding window down 26 source rows -> 20 rows of sums.
  if (lane < 2 * PW) {
    const int w = lane < PW ? lane : lane - PW;
    const int r0 = lane < PW ? 0 : BL - 20;  // halves overlap on row 19
    const uint32_t* src = sm.patch + r0 * PW + w;
    uint2* tp = sm.t + r0 * PW + w;
    uint32_t e[7], o[7];
#pragma unroll
    for (int k = 0; k < 6; k++) {
      const uint32_t x = src[k * PW];
      e[k] = x & 0x00ff00ffu;
      o[k] = (x >> 8) & 0x00ff00ffu;
    }
#pragma unroll
    for (int r = 0; r < 20; r++) {
      const uint32_t x = src[(r + 6) * PW];
      e[6] = x & 0x00ff00ffu;
      o[6] = (x >> 8) & 0x00ff00ffu;
      uint2 s;
      s.x = 18u * (e[0] + e[6]) + 34u * (e[1] + e[5]) + 48u * (e[2] + e[4]) + 56u * e[3];
      s.y = 18u * (o[0] + o[6]) + 34u * (o[1] + o[5]) + 48u * (o[2] + o[4]) + 56u * o[3];
      tp[r * PW] = s;
#pragma unroll
      for (int k = 0; k < 6; k++) { e[k] = e[k + 1]; o[k] = o[k + 1]; }
    }
  }
  __syncwarp();
  // Row pass: one task = 4 adjacent output bytes (one word) of one row; the source patch is dead
  // now (orientation and column pass are done), so the blurred window overwrites it.
  {
    uint32_t* bw = sm.patch;
    const int ntask = c_row_ntasks[off];
    const uint16_t* tasks = c_row_tasks + off * ROW_TASKS_MAX;
    for (int i = lane; i < ntask; i += 32) {
      const int tk = tasks[i];
      const int r = tk >> 4, w = tk & 15, q = r * BW + w;
      const uint2* tp = sm.t + r * PW + w;
      const uint2 a = tp[0], b = tp[1], c = tp[2];
      // a.x = {T0,T2}, a.y = {T1,T3}, b.x = {T4,T6}, b.y = {T5,T7}, c.x = {T8,..}, c.y = {T9,..}: each output is a
      // 7-tap dot product over 16-bit sums = 4-5 dp2a (u16 x u8) on the packed pairs, no unpacking; the rounded
      // byte (acc + 2^15) >> 16 is byte 2 of the accumulator (acc < 2^24)
      uint32_t o0 = __dp2a_lo(a.x, 0x00003012u, 32768u);   // 18 T0 + 48 T2
      o0 = __dp2a_lo(a.y, 0x00003822u, o0);                // 34 T1 + 56 T3
      o0 = __dp2a_lo(b.x, 0x00001230u, o0);                // 48 T4 + 18 T6
      o0 = __dp2a_lo(b.y, 0x00000022u, o0);                // 34 T5
      uint32_t o1 = __dp2a_lo(a.y, 0x00003012u, 32768u);   // 18 T1 + 48 T3
      o1 = __dp2a_lo(a.x, 0x00002200u, o1);                // 34 T2
      o1 = __dp2a_lo(b.x, 0x00002238u, o1);                // 56 T4 + 34 T6
      o1 = __dp2a_lo(b.y, 0x00001230u, o1);                // 48 T5 + 18 T7
      uint32_t o2 = __dp2a_lo(a.x, 0x00001200u, 32768u);   // 18 T2
      o2 = __dp2a_lo(a.y, 0x00002200u, o2);                // 34 T3
      o2 = __dp2a_lo(b.x, 0x00003030u, o2);                // 48 T4 + 48 T6
      o2 = __dp2a_lo(b.y, 0x00002238u, o2);                // 56 T5 + 34 T7
      o2 = __dp2a_lo(c.x, 0x00000012u, o2);                // 18 T8
      uint32_t o3 = __dp2a_lo(a.y, 0x00001200u, 32768u);   // 18 T3
      o3 = __dp2a_lo(b.x, 0x00003822u, o3);                // 34 T4 + 56 T6
      o3 = __dp2a_lo(b.y, 0x00003030u, o3);                // 48 T5 + 48 T7
      o3 = __dp2a_lo(c.x, 0x00000022u, o3);                // 34 T8
      o3 = __dp2a_lo(c.y, 0x00000012u, o3);                // 18 T9
      bw[q] = __byte_perm(__byte_perm(o0, o1, 0x0062), __byte_perm(o2, o3, 0x0062), 0x5410);
    }
  }
  __syncwarp();

  // ---- rotated BRIEF: test t = 32*j + lane -> bit `lane` of descriptor word j
  const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
  float a, b;
  orbx_sincosf(f_mul(angle, factorPI), &b, &a);
  const uint8_t* centre = reinterpret_cast<const uint8_t*>(sm.patch) + 19 * (BW * 4) + 19 + off;
  // cvRound(v) for |v| < 2^22: v + 1.5*2^23 rounds to nearest-even at integer granularity, and the
  // integer sits in the low mantissa bits; row*44 + col is formed directly on the biased values.
  const float kMagic = 12582912.0f;
  const uint32_t kBias = 0x4B400000u * (uint32_t)(BW * 4 + 1);
  uint32_t word_mine = 0;
#pragma unroll
  for (int j = 0; j < 8; j++) {
    const float4 pt = __ldg(reinterpret_cast<const float4*>(c_patf) + 32 * j + lane);
    float fy0, fx0, fy1, fx1;
    if (STRICT_FP) {
      fy0 = f_add(f_mul(pt.x, b), f_mul(pt.y, a)); fx0 = f_sub(f_mul(pt.x, a), f_mul(pt.y, b));
      fy1 = f_add(f_mul(pt.z, b), f_mul(pt.w, a)); fx1 = f_sub(f_mul(pt.z, a), f_mul(pt.w, b));
    } else {
      fy0 = f_fma(pt.x, b, f_mul(pt.y, a)); fx0 = f_fma(pt.x, a, -f_mul(pt.y, b));
      fy1 = f_fma(pt.z, b, f_mul(pt.w, a)); fx1 = f_fma(pt.z, a, -f_mul(pt.w, b));
    }
    const uint32_t i0 = __float_as_uint(f_add(fy0, kMagic)) * (uint32_t)(BW * 4) + __float_as_uint(f_add(fx0, kMagic)) - kBias;
    const uint32_t i1 = __float_as_uint(f_add(fy1, kMagic)) * (uint32_t)(BW * 4) + __float_as_uint(f_add(fx1, kMagic)) - kBias;
    const int v0 = centre[(int)i0], v1 = centre[(int)i1];
    const uint32_t w = __ballot_sync(0xffffffffu, v0 < v1);
    if (lane == j) word_mine = w;
  }

  // ---- emit
  const int d = dst[(long long)frame * P.kp_total_cap + slot];
  if (d < 0 || d >= capacity) return;
  const long long o = (long long)frame * capacity + d;
  if (lane < 8) reinterpret_cast<uint32_t*>(desc + o * 32)[lane] = word_mine;
  if (lane == 8) {
    OrbxKeyPointDev kp;
    kp.x = (float)kx; kp.y = (float)ky;
    if (level != 0) { kp.x = f_mul(kp.x, L.scale); kp.y = f_mul(kp.y, L.scale); }
    kp.size = L.patch_size;
    kp.angle = angle;
    kp.response = (float)ORBX_KEY_S(key);
    kp.octave = level;
    kp.class_id = -1;
    kps[o] = kp;
  }
}

// ------------------------------------------------------------------------------------------------
// "Level blur" mode: the levels were blurred once by k_blur_levels (blur.cu), so a keypoint only
// needs its orientation (moments straight from the level image: lane = patch row, 9 aligned words
// per row, two dp4a per word against constant column weights) and 512 samples from the blurred level.
// No shared memory.
// ------------------------------------------------------------------------------------------------

__device__ __forceinline__ int dp4a_us(uint32_t a_u8x4, uint32_t b_s8x4, int c) {
  int d;
  asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a_u8x4), "r"(b_s8x4), "r"(c));
  return d;
}

// Moment weights for the coalesced patch read: the 31x36-byte window (9 aligned words per patch row) is
// word i = 9*r + j, lane-strided.  For alignment offset sh = (kx-15)&3, byte b of word i is patch column
// u = 4j + b - sh - 15 of row v = r - 15; g_mom_wu holds u and g_mom_wv holds v (signed bytes) where
// |u| <= umax[|v|] and 0 elsewhere, so m10 += dp4a(px, wu), m01 += dp4a(px, wv) (ORBextractor.cc:76-103).
#define MOM_WORDS 279
#define MOM_STRIDE 288
__device__ uint32_t g_mom_wu[4 * MOM_STRIDE];
__device__ uint32_t g_mom_wv[4 * MOM_STRIDE];
#define LITE_WORDS (BL * BW)  // 429 words: the 39 x 44-byte blurred window around the keypoint

template <bool STRICT_FP>
__global__ void __launch_bounds__(DESC_WARPS * 32)
k_describe_lite(const __grid_constant__ DevPlan P, Level0 l0, const uint8_t* __restrict__ arena,
                const uint8_t* __restrict__ blur, const uint32_t* __restrict__ sel,
                const int32_t* __restrict__ sel_cnt, const int32_t* __restrict__ dst,
                OrbxKeyPointDev* __restrict__ kps, uint8_t* __restrict__ desc, int capacity) {
  __shared__ __align__(16) uint32_t s_win[DESC_WARPS][LITE_WORDS + 3];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int frame = blockIdx.y;
  const int slot = blockIdx.x * DESC_WARPS + warp;
  if (slot >= P.kp_total_cap) return;
  int level = 0;
#pragma unroll 1
  for (int l = 1; l < P.nlevels; l++)
    if (slot >= P.lv[l].kp_base) level = l;
  const DevLevel& L = P.lv[level];
  const int idx = slot - L.kp_base;
  if (idx >= sel_cnt[frame * P.nlevels + level]) return;
  const uint32_t key = sel[(long long)frame * P.kp_total_cap + slot];
  const int kx = ORBX_KEY_X(key), ky = ORBX_KEY_Y(key);
  int pitch;
  const uint8_t* img;
  if (level == 0) { pitch = l0.pitch; img = l0.base + (long long)frame * l0.frame_stride; }
  else { pitch = L.pitch; img = arena + (long long)frame * P.arena_bytes + L.offset; }

  // ---- blurred 39x39 window -> shared memory, coalesced: word i = 11*r + j of the window that starts at
  // the aligned column (kx-19)&~3 (the blurred levels live in the library's own pitched buffer: always
  // 4-byte aligned, and the <= 5 bytes read past column w-1 stay inside the row pitch / the slack)
  const int bp = L.pitch;
  const int xb = kx - 19, shb = xb & 3;
  const uint8_t* bwin = blur + (long long)frame * P.arena_bytes + L.offset + (long long)(ky - 19) * bp + (xb - shb);
  uint32_t bw[14];
#pragma unroll
  for (int t = 0; t < 14; t++) {
    const int i = lane + 32 * t;
    const int r = (i * 745) >> 13, j = i - r * BW;  // i / 11 for i < 448
    bw[t] = 0;
    if (t < 13 || i < LITE_WORDS) bw[t] = __ldg(reinterpret_cast<const uint32_t*>(bwin + r * bp) + j);
  }

  // ---- IC_Angle: integer moments over the circular patch of radius 15
  int m01 = 0, m10 = 0;
  if ((pitch & 3) == 0 && ((uintptr_t)img & 3) == 0) {
    // rows ky-15..ky+15 never include the last image row, so the <= 1 byte read past column w-1 is in-bounds
    const int xa = kx - ORBX_HALF_PATCH, sh = xa & 3;
    const uint8_t* pwin = img + (long long)(ky - ORBX_HALF_PATCH) * pitch + (xa - sh);
    const uint32_t* wu = g_mom_wu + sh * MOM_STRIDE;
    const uint32_t* wv = g_mom_wv + sh * MOM_STRIDE;
#pragma unroll
    for (int t = 0; t < 9; t++) {
      const int i = lane + 32 * t;
      if (t < 8 || i < MOM_WORDS) {
        const int r = (i * 57) >> 9, j = i - r * 9;  // i / 9 for i < 288
        const uint32_t px = __ldg(reinterpret_cast<const uint32_t*>(pwin + r * pitch) + j);
        m10 = dp4a_us(px, __ldg(wu + i), m10);
        m01 = dp4a_us(px, __ldg(wv + i), m01);
      }
    }
  } else if (lane < 31) {
    const int v = lane - ORBX_HALF_PATCH, av = v < 0 ? -v : v;
    const uint8_t* rowp = img + (long long)(ky + v) * pitch + kx;
    const int d = P.umax[av];
    int s = 0, su = 0;
    for (int u = -d; u <= d; u++) {
      const int val = rowp[u];
      s += val;
      su += u * val;
    }
    m10 = su;
    m01 = v * s;
  }
#pragma unroll
  for (int t = 0; t < 14; t++) {
    const int i = lane + 32 * t;
    if (t < 13 || i < LITE_WORDS) s_win[warp][i] = bw[t];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    m01 += __shfl_xor_sync(0xffffffffu, m01, o);
    m10 += __shfl_xor_sync(0xffffffffu, m10, o);
  }
  const float angle = orbx_fast_atan2((float)m01, (float)m10);
  __syncwarp();

  // ---- rotated BRIEF on the staged window: test t = 32*j + lane -> bit `lane` of descriptor word j
  const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
  float a, b;
  orbx_sincosf(f_mul(angle, factorPI), &b, &a);
  const uint8_t* centre = reinterpret_cast<const uint8_t*>(s_win[warp]) + 19 * (BW * 4) + 19 + shb;
  const float kMagic = 12582912.0f;  // cvRound through the mantissa, see k_describe
  const uint32_t kBias = 0x4B400000u * (uint32_t)(BW * 4 + 1);
  uint32_t word_mine = 0;
#pragma unroll
  for (int j = 0; j < 8; j++) {
    const float4 pt = __ldg(reinterpret_cast<const float4*>(c_patf) + 32 * j + lane);
    float fy0, fx0, fy1, fx1;
    if (STRICT_FP) {
      fy0 = f_add(f_mul(pt.x, b), f_mul(pt.y, a)); fx0 = f_sub(f_mul(pt.x, a), f_mul(pt.y, b));
      fy1 = f_add(f_mul(pt.z, b), f_mul(pt.w, a)); fx1 = f_sub(f_mul(pt.z, a), f_mul(pt.w, b));
    } else {
      fy0 = f_fma(pt.x, b, f_mul(pt.y, a)); fx0 = f_fma(pt.x, a, -f_mul(pt.y, b));
      fy1 = f_fma(pt.z, b, f_mul(pt.w, a)); fx1 = f_fma(pt.z, a, -f_mul(pt.w, b));
    }
    const uint32_t i0 = __float_as_uint(f_add(fy0, kMagic)) * (uint32_t)(BW * 4) + __float_as_uint(f_add(fx0, kMagic)) - kBias;
    const uint32_t i1 = __float_as_uint(f_add(fy1, kMagic)) * (uint32_t)(BW * 4) + __float_as_uint(f_add(fx1, kMagic)) - kBias;
    const int v0 = centre[(int)i0], v1 = centre[(int)i1];
    const uint32_t wd = __ballot_sync(0xffffffffu, v0 < v1);
    if (lane == j) word_mine = wd;
  }

  // ---- emit
  const int d = dst[(long long)frame * P.kp_total_cap + slot];
  if (d < 0 || d >= capacity) return;
  const long long o = (long long)frame * capacity + d;
  if (lane < 8) reinterpret_cast<uint32_t*>(desc + o * 32)[lane] = word_mine;
  if (lane == 8) {
    OrbxKeyPointDev kp;
    kp.x = (float)kx; kp.y = (float)ky;
    if (level != 0) { kp.x = f_mul(kp.x, L.scale); kp.y = f_mul(kp.y, L.scale); }
    kp.size = L.patch_size;
    kp.angle = angle;
    kp.response = (float)ORBX_KEY_S(key);
    kp.octave = level;
    kp.class_id = -1;
    kps[o] = kp;
  }
}

// ------------------------------------------------------------------------------------------------
// k_describe_lite4: the same work with LITE_KPW keypoints per warp.  The orientation chain (fastAtan2 + the
// glibc-exact sincosf, ~225 warp instructions, a quarter of k_describe_lite) is a scalar computation per
// keypoint: here lane g evaluates it for keypoint g, so one pass of the chain serves LITE_KPW keypoints.  The
// slot -> level lookup and the result-slot fetch are lane-parallel too; window words are addressed
// incrementally (i -> i+32 is "three rows down, one word left" on an 11-word row) instead of by a division per
// load; the blurred window of keypoint g+1 is on its way while keypoint g is being sampled.
// ------------------------------------------------------------------------------------------------
#define LITE_KPW 4

template <bool STRICT_FP>
__global__ void __launch_bounds__(DESC_WARPS * 32)
k_describe_lite4(const __grid_constant__ DevPlan P, Level0 l0, const uint8_t* __restrict__ arena,
                 const uint8_t* __restrict__ blur, const uint32_t* __restrict__ sel,
                 const int32_t* __restrict__ sel_cnt, const int32_t* __restrict__ dst,
                 OrbxKeyPointDev* __restrict__ kps, uint8_t* __restrict__ desc, int capacity) {
  __shared__ __align__(16) uint32_t s_win[DESC_WARPS][LITE_WORDS + 3];
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int frame = blockIdx.y;
  const int slot0 = (blockIdx.x * DESC_WARPS + warp) * LITE_KPW;
  if (slot0 >= P.kp_total_cap) return;

  // ---- header, lane-parallel: lane handles keypoint gl = lane % LITE_KPW of the warp's group
  const int gl = lane & (LITE_KPW - 1);
  const int slot_l = slot0 + gl;
  int level_l = 0;
#pragma unroll
  for (int l = 1; l < ORBX_PLAN_MAX_LEVELS; l++)
    if (l < P.nlevels && slot_l >= P.lv[l].kp_base) level_l = l;
  uint32_t key_l = 0;
  int dst_l = -1;
  bool valid_l = slot_l < P.kp_total_cap;
  if (valid_l) valid_l = slot_l - P.lv[level_l].kp_base < sel_cnt[frame * P.nlevels + level_l];
  if (valid_l) {
    key_l = sel[(long long)frame * P.kp_total_cap + slot_l];
    dst_l = dst[(long long)frame * P.kp_total_cap + slot_l];
    valid_l = dst_l >= 0 && dst_l < capacity;  // a keypoint without a result slot is never written
  }
  const unsigned vmask = __ballot_sync(FULL, valid_l) & ((1u << LITE_KPW) - 1u);
  if (vmask == 0) return;

  // ---- IC_Angle moments of every keypoint of the group (ORBextractor.cc:76-103)
  int my01 = 0, my10 = 0;  // the sums of keypoint gl
#pragma unroll
  for (int g = 0; g < LITE_KPW; g++) {
    if (!((vmask >> g) & 1u)) continue;
    const uint32_t key = __shfl_sync(FULL, key_l, g);
    const int level = __shfl_sync(FULL, level_l, g);
    const DevLevel& L = P.lv[level];
    const int kx = ORBX_KEY_X(key), ky = ORBX_KEY_Y(key);
    int pitch;
    const uint8_t* img;
    if (level == 0) { pitch = l0.pitch; img = l0.base + (long long)frame * l0.frame_stride; }
    else { pitch = L.pitch; img = arena + (long long)frame * P.arena_bytes + L.offset; }
    int m01 = 0, m10 = 0;
    if ((pitch & 3) == 0 && ((uintptr_t)img & 3) == 0) {
      const int xa = kx - ORBX_HALF_PATCH, sh = xa & 3;
      // word i = 9 r + j of the 31 x 36-byte window; i -> i + 32: j + 5 (mod 9), three rows down (four on the wrap)
      const int r0 = (lane * 57) >> 9;
      int j = lane - 9 * r0;
      const uint8_t* q = img + (long long)(ky - ORBX_HALF_PATCH + r0) * pitch + (xa - sh) + 4 * j;
      const int step = 3 * pitch + 20, wrap = pitch - 36;
      const uint32_t* wu = g_mom_wu + sh * MOM_STRIDE + lane;
      const uint32_t* wv = g_mom_wv + sh * MOM_STRIDE + lane;
#pragma unroll
      for (int t = 0; t < 9; t++) {
        if (t < 8 || lane + 32 * 8 < MOM_WORDS) {
          const uint32_t px = __ldg(reinterpret_cast<const uint32_t*>(q));
          m10 = dp4a_us(px, __ldg(wu + 32 * t), m10);
          m01 = dp4a_us(px, __ldg(wv + 32 * t), m01);
        }
        q += step;
        j += 5;
        if (j >= 9) { j -= 9; q += wrap; }
      }
    } else if (lane < 31) {
      const int v = lane - ORBX_HALF_PATCH, av = v < 0 ? -v : v;
      const uint8_t* rowp = img + (long long)(ky + v) * pitch + kx;
      const int d = P.umax[av];
      int s = 0, su = 0;
      for (int u = -d; u <= d; u++) {
        const int val = rowp[u];
        s += val;
        su += u * val;
      }
      m10 = su;
      m01 = v * s;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      m01 += __shfl_xor_sync(FULL, m01, o);
      m10 += __shfl_xor_sync(FULL, m10, o);
    }
    if (gl == g) { my01 = m01; my10 = m10; }
  }

  // ---- orientation of keypoint gl on lane gl (lanes >= LITE_KPW repeat it), :103 and the rotation of :107-146
  const float angle_l = orbx_fast_atan2((float)my01, (float)my10);
  const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
  float a_l, b_l;
  orbx_sincosf(f_mul(angle_l, factorPI), &b_l, &a_l);

  // ---- blurred 39x39 windows -> shared memory -> rotated BRIEF, one keypoint after the other
  uint32_t bw[14];
  auto fetch_window = [&](int g) {
    const uint32_t key = __shfl_sync(FULL, key_l, g);
    const int level = __shfl_sync(FULL, level_l, g);
    const DevLevel& L = P.lv[level];
    const int kx = ORBX_KEY_X(key), ky = ORBX_KEY_Y(key);
    const int bp = L.pitch, xb = kx - 19, shb = xb & 3;
    // word i = 11 r + j of the 39 x 44-byte window; i -> i + 32: j - 1 (mod 11), three rows down (two on the wrap)
    const int r0 = (lane * 745) >> 13;
    int j = lane - BW * r0;
    const uint8_t* q = blur + (long long)frame * P.arena_bytes + L.offset + (long long)(ky - 19 + r0) * bp + (xb - shb) + 4 * j;
    const int step = 3 * bp - 4, wrap = 44 - bp;
#pragma unroll
    for (int t = 0; t < 14; t++) {
      bw[t] = 0;
      if (t < 13 || lane + 32 * 13 < LITE_WORDS) bw[t] = __ldg(reinterpret_cast<const uint32_t*>(q));
      q += step;
      if (--j < 0) { j = BW - 1; q += wrap; }
    }
  };
  const float kMagic = 12582912.0f;  // cvRound through the mantissa, see k_describe
  const uint32_t kBias = 0x4B400000u * (uint32_t)(BW * 4 + 1);
  int g = __ffs(vmask) - 1;
  fetch_window(g);
  while (g >= 0) {
    const unsigned rest = vmask & ~((2u << g) - 1u);
    const int gn = rest ? __ffs(rest) - 1 : -1;
#pragma unroll
    for (int t = 0; t < 14; t++)
      if (t < 13 || lane + 32 * 13 < LITE_WORDS) s_win[warp][lane + 32 * t] = bw[t];
    __syncwarp();
    if (gn >= 0) fetch_window(gn);  // in flight while this keypoint is sampled
    const uint32_t key = __shfl_sync(FULL, key_l, g);
    const int level = __shfl_sync(FULL, level_l, g);
    const int d = __shfl_sync(FULL, dst_l, g);
    const float angle = __shfl_sync(FULL, angle_l, g);
    const float a = __shfl_sync(FULL, a_l, g), b = __shfl_sync(FULL, b_l, g);
    const DevLevel& L = P.lv[level];
    const int kx = ORBX_KEY_X(key), ky = ORBX_KEY_Y(key);
    const int shb = (kx - 19) & 3;
    const uint8_t* centre = reinterpret_cast<const uint8_t*>(s_win[warp]) + 19 * (BW * 4) + 19 + shb;
    uint32_t word_mine = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) {
      const float4 pt = __ldg(reinterpret_cast<const float4*>(c_patf) + 32 * j + lane);
      float fy0, fx0, fy1, fx1;
      if (STRICT_FP) {
        fy0 = f_add(f_mul(pt.x, b), f_mul(pt.y, a)); fx0 = f_sub(f_mul(pt.x, a), f_mul(pt.y, b));
        fy1 = f_add(f_mul(pt.z, b), f_mul(pt.w, a)); fx1 = f_sub(f_mul(pt.z, a), f_mul(pt.w, b));
      } else {
        fy0 = f_fma(pt.x, b, f_mul(pt.y, a)); fx0 = f_fma(pt.x, a, -f_mul(pt.y, b));
        fy1 = f_fma(pt.z, b, f_mul(pt.w, a)); fx1 = f_fma(pt.z, a, -f_mul(pt.w, b));
      }
      const uint32_t i0 = __float_as_uint(f_add(fy0, kMagic)) * (uint32_t)(BW * 4) + __float_as_uint(f_add(fx0, kMagic)) - kBias;
      const uint32_t i1 = __float_as_uint(f_add(fy1, kMagic)) * (uint32_t)(BW * 4) + __float_as_uint(f_add(fx1, kMagic)) - kBias;
      const int v0 = centre[(int)i0], v1 = centre[(int)i1];
      const uint32_t wd = __ballot_sync(FULL, v0 < v1);
      if (lane == j) word_mine = wd;
    }
    const long long o = (long long)frame * capacity + d;
    if (lane < 8) reinterpret_cast<uint32_t*>(desc + o * 32)[lane] = word_mine;
    if (lane == 8) {
      OrbxKeyPointDev kp;
      kp.x = (float)kx; kp.y = (float)ky;
      if (level != 0) { kp.x = f_mul(kp.x, L.scale); kp.y = f_mul(kp.y, L.scale); }
      kp.size = L.patch_size;
      kp.angle = angle;
      kp.response = (float)ORBX_KEY_S(key);
      kp.octave = level;
      kp.class_id = -1;
      kps[o] = kp;
    }
    __syncwarp();  // every lane has sampled before the next window overwrites the buffer
    g = gn;
  }
}

// Fills the row-task and moment-weight tables from the plan's umax table (once per process; the table is a constant of the
// algorithm, ORBextractor.cc:453-468).
static const float h_patf[1024] = {
#include "orb_pattern.inc"
};

cudaError_t orbx_describe_init_tables(const int* umax) {
  {
    // blurred pixels (dx, dy) whose unit square meets the disc of radius max|p| (+ margin for fp32)
    double R = 0;
    for (int i = 0; i < 512; i++) R = fmax(R, hypot((double)h_patf[2 * i], (double)h_patf[2 * i + 1]));
    R += 0.05;
    std::vector<uint16_t> tab(4 * ROW_TASKS_MAX, 0);
    int cnt[4];
    for (int off = 0; off < 4; off++) {
      int n = 0;
      for (int r = 0; r < BL; r++) {
        const double ay = fmax(fabs((double)(r - 19)) - 0.5, 0.0);
        int m = -1;
        for (int dx = 0; dx <= 19; dx++)
          if (hypot(fmax(dx - 0.5, 0.0), ay) <= R) m = dx;
        if (m < 0) continue;
        for (int w = (19 - m + off) >> 2; w <= (19 + m + off) >> 2; w++) tab[off * ROW_TASKS_MAX + n++] = (uint16_t)(r << 4 | w);
      }
      cnt[off] = n;
    }
    cudaError_t e = cudaMemcpyToSymbol(c_row_tasks, tab.data(), tab.size() * sizeof(uint16_t));
    if (e != cudaSuccess) return e;
    e = cudaMemcpyToSymbol(c_row_ntasks, cnt, sizeof(cnt));
    if (e != cudaSuccess) return e;
  }
  std::vector<uint32_t> wu(4 * MOM_STRIDE, 0), wv(4 * MOM_STRIDE, 0);
  for (int sh = 0; sh < 4; sh++)
    for (int i = 0; i < MOM_WORDS; i++) {
      const int r = i / 9, j = i % 9, v = r - 15, av = v < 0 ? -v : v;
      for (int b = 0; b < 4; b++) {
        const int u = 4 * j + b - sh - 15;
        if ((u < 0 ? -u : u) > umax[av]) continue;
        wu[sh * MOM_STRIDE + i] |= (uint32_t)(uint8_t)(int8_t)u << (8 * b);
        wv[sh * MOM_STRIDE + i] |= (uint32_t)(uint8_t)(int8_t)v << (8 * b);
      }
    }
  cudaError_t e2 = cudaMemcpyToSymbol(g_mom_wu, wu.data(), wu.size() * sizeof(uint32_t));
  if (e2 != cudaSuccess) return e2;
  return cudaMemcpyToSymbol(g_mom_wv, wv.data(), wv.size() * sizeof(uint32_t));
}

cudaError_t orbx_launch_describe_lite(const DevPlan& P, Level0 l0, const uint8_t* arena, const uint8_t* blur,
                                      const uint32_t* sel, const int32_t* sel_cnt, const int32_t* dst,
                                      OrbxKeyPointDev* kps, uint8_t* desc, int capacity, int strict_fp, int batch,
                                      cudaStream_t st) {
  static const int kpw = [] { const char* e = getenv("ORBX_DESC_KPW"); return e ? atoi(e) : LITE_KPW; }();
  if (kpw == LITE_KPW) {
    dim3 grd((P.kp_total_cap + DESC_WARPS * LITE_KPW - 1) / (DESC_WARPS * LITE_KPW), batch);
    if (strict_fp) k_describe_lite4<true><<<grd, DESC_WARPS * 32, 0, st>>>(P, l0, arena, blur, sel, sel_cnt, dst, kps, desc, capacity);
    else k_describe_lite4<false><<<grd, DESC_WARPS * 32, 0, st>>>(P, l0, arena, blur, sel, sel_cnt, dst, kps, desc, capacity);
    return cudaGetLastError();
  }
  dim3 grd((P.kp_total_cap + DESC_WARPS - 1) / DESC_WARPS, batch);
  if (strict_fp) k_describe_lite<true><<<grd, DESC_WARPS * 32, 0, st>>>(P, l0, arena, blur, sel, sel_cnt, dst, kps, desc, capacity);
  else k_describe_lite<false><<<grd, DESC_WARPS * 32, 0, st>>>(P, l0, arena, blur, sel, sel_cnt, dst, kps, desc, capacity);
  return cudaGetLastError();
}

cudaError_t orbx_launch_describe(const DevPlan& P, Level0 l0, const uint8_t* arena, const uint32_t* sel,
                                 const int32_t* sel_cnt, const int32_t* dst, OrbxKeyPointDev* kps,
                                 uint8_t* desc, int capacity, int strict_fp, int batch, cudaStream_t st) {
  dim3 grd((P.kp_total_cap + DESC_WARPS - 1) / DESC_WARPS, batch);
  if (strict_fp) k_describe<true><<<grd, DESC_WARPS * 32, 0, st>>>(P, l0, arena, sel, sel_cnt, dst, kps, desc, capacity);
  else k_describe<false><<<grd, DESC_WARPS * 32, 0, st>>>(P, l0, arena, sel, sel_cnt, dst, kps, desc, capacity);
  return cudaGetLastError();
}
