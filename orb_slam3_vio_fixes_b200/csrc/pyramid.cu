// pyramid.cu — image pyramid (ORBextractor::ComputePyramid, src/ORBextractor.cc:1170-1195) for sm_100a.
#include "orbx_kernels.cuh"

__device__ __forceinline__ const uint8_t* level_ptr(const DevPlan& P, const Level0& l0, const uint8_t* arena,
                                                    int frame, int level, int* pitch) {
  if (level == 0) {
    *pitch = l0.pitch;
    return l0.base + (long long)frame * l0.frame_stride;
  }
  *pitch = P.lv[level].pitch;
  return arena + (long long)frame * P.arena_bytes + P.lv[level].offset;
}

// ------------------------------------------------------------------------------------------------
// cv::resize(level-1 -> level, INTER_LINEAR) on CV_8UC1: Q11 horizontal taps, then
//   out = ( ((b0*(H0>>4))>>16) + ((b1*(H1>>4))>>16) + 2 ) >> 2
// One thread produces 4 adjacent output pixels and stores them as one 32-bit word.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_resize(const uint8_t* __restrict__ src, int spitch, long long sframe, int sw, int sh,
         uint8_t* __restrict__ dst, int dpitch, long long dframe, int dw, int dh,
         const OrbxTap* __restrict__ xt, const OrbxTap* __restrict__ yt) {
  const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  const int dy = blockIdx.y * blockDim.y + threadIdx.y;
  if (x4 >= dw || dy >= dh) return;
  const uint8_t* S = src + (long long)blockIdx.z * sframe;
  uint8_t* D = dst + (long long)blockIdx.z * dframe;
  const OrbxTap ty = yt[dy];
  int sy0 = min(max(ty.ofs, 0), sh - 1), sy1 = min(max(ty.ofs + 1, 0), sh - 1);
  const uint8_t* r0 = S + (long long)sy0 * spitch;
  const uint8_t* r1 = S + (long long)sy1 * spitch;
  const int b0 = ty.c0, b1 = ty.c1;
  uint32_t out = 0;
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const int dx = x4 + i;
    if (dx < dw) {
      const OrbxTap tx = xt[dx];
      const int sx = tx.ofs, sx1 = min(sx + 1, sw - 1);
      const int h0 = r0[sx] * tx.c0 + r0[sx1] * tx.c1;
      const int h1 = r1[sx] * tx.c0 + r1[sx1] * tx.c1;
      int v = (((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2;
      v = min(max(v, 0), 255);
      out |= (uint32_t)v << (8 * i);
    }
  }
  // rows are 128-byte aligned and padded, so the full word store stays inside the row
  *reinterpret_cast<uint32_t*>(D + (long long)dy * dpitch + x4) = out;
}

// ------------------------------------------------------------------------------------------------
// The same arithmetic for down-scales below 4/3 (ORB-SLAM3 uses 1.2), word-aligned rows.
// A thread owns 4 adjacent output columns and walks down RS_ROWS output rows:
//   * its 4 horizontal taps read source bytes ofs_i, ofs_i+1 with ofs_i - ofs_0 in {0..4}; taps 0,1
//     lie in the 4-byte window at ofs_0 and taps 2,3 in the window at ofs_0+2 (checked on the host,
//     OrbxLevel::fast_resize), so a source row costs 4 aligned word loads, 2 funnel shifts and one
//     PRMT per byte (selectors are per-thread constants) instead of byte gathers;
//   * H = r0*c0 + r1*c1 is formed pre-shifted by 12 bits, (H>>4)<<16 = H4096 & 0xffff0000, so the
//     vertical term (b*(H>>4))>>16 is one multiply-high;
//   * consecutive output rows share a source row (the step is 1 or 2), which is kept in registers.
// ------------------------------------------------------------------------------------------------
#define RS_ROWS 16
#define RS_WARPS 4

struct RsTapRegs {
  uint32_t selA[4], selB[4];  // PRMT selectors: byte -> position 0 / position 1 (x256), rest zero
  uint32_t C0[4], C1[4];      // c0 << 12, c1 << 4
  int wa, wa1, wm, wm1;       // word indices of the two windows (clamped to the row)
  int sa, sm;                 // funnel shifts in bits
};

__device__ __forceinline__ void rs_row(const uint32_t* __restrict__ row, const RsTapRegs& T, uint32_t (&H)[4]) {
  const uint32_t lo = __funnelshift_r(__ldg(row + T.wa), __ldg(row + T.wa1), T.sa);
  const uint32_t mid = __funnelshift_r(__ldg(row + T.wm), __ldg(row + T.wm1), T.sm);
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const uint32_t win = i < 2 ? lo : mid;
    const uint32_t h = __byte_perm(win, 0, T.selA[i]) * T.C0[i] + __byte_perm(win, 0, T.selB[i]) * T.C1[i];
    H[i] = h & 0xffff0000u;  // ((r0*c0 + r1*c1) >> 4) << 16
  }
}

__global__ void __launch_bounds__(RS_WARPS * 32)
k_resize_walk(const uint8_t* __restrict__ src, int spitch, long long sframe, int sw, int sh,
              uint8_t* __restrict__ dst, int dpitch, long long dframe, int dw, int dh,
              const OrbxTap* __restrict__ xt, const OrbxTap* __restrict__ yt) {
  const int x4 = (blockIdx.x * 32 + (threadIdx.x & 31)) * 4;
  const int dy0 = (blockIdx.y * RS_WARPS + (threadIdx.x >> 5)) * RS_ROWS;
  if (x4 >= dw || dy0 >= dh) return;
  const uint8_t* S = src + (long long)blockIdx.z * sframe;
  uint8_t* D = dst + (long long)blockIdx.z * dframe;
  RsTapRegs T;
  {
    const int wlast = (spitch >> 2) - 1;
    const int o0 = xt[x4].ofs;
    T.wa = min(o0 >> 2, wlast); T.wa1 = min((o0 >> 2) + 1, wlast); T.sa = 8 * (o0 & 3);
    T.wm = min((o0 + 2) >> 2, wlast); T.wm1 = min(((o0 + 2) >> 2) + 1, wlast); T.sm = 8 * ((o0 + 2) & 3);
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const OrbxTap t = xt[min(x4 + i, dw - 1)];
      int d = t.ofs - o0 - (i < 2 ? 0 : 2);
      d = min(max(d, 0), 2);  // only the padding columns beyond dw can fall outside
      T.selA[i] = 0x4440u | (uint32_t)d;
      T.selB[i] = 0x4404u | ((uint32_t)(d + 1) << 4);
      T.C0[i] = (uint32_t)t.c0 << 12;
      T.C1[i] = (uint32_t)t.c1 << 4;
    }
  }
  uint32_t HA[4], HB[4];
  int ra = -1, rb = -1;  // source rows held in HA / HB
  const int dy1 = min(dy0 + RS_ROWS, dh);
  for (int dy = dy0; dy < dy1; dy++) {
    const OrbxTap ty = yt[dy];
    const int sy0 = min(max(ty.ofs, 0), sh - 1), sy1 = min(max(ty.ofs + 1, 0), sh - 1);
    if (sy0 == rb) {
#pragma unroll
      for (int i = 0; i < 4; i++) HA[i] = HB[i];
      ra = rb;
    } else if (sy0 != ra) {
      rs_row(reinterpret_cast<const uint32_t*>(S + (long long)sy0 * spitch), T, HA);
      ra = sy0;
    }
    if (sy1 == ra) {
#pragma unroll
      for (int i = 0; i < 4; i++) HB[i] = HA[i];
      rb = ra;
    } else if (sy1 != rb) {
      rs_row(reinterpret_cast<const uint32_t*>(S + (long long)sy1 * spitch), T, HB);
      rb = sy1;
    }
    const uint32_t b0 = (uint32_t)ty.c0, b1 = (uint32_t)ty.c1;
    uint32_t out = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const uint32_t v = (__umulhi(b0, HA[i]) + __umulhi(b1, HB[i]) + 2u) >> 2;  // <= 255, see DESIGN.md
      out |= v << (8 * i);
    }
    // rows are 128-byte aligned and padded, so the full word store stays inside the row
    *reinterpret_cast<uint32_t*>(D + (long long)dy * dpitch + x4) = out;
  }
}

cudaError_t orbx_launch_pyramid(const DevPlan& P, Level0 l0, uint8_t* arena, const OrbxTap* xtaps,
                                const OrbxTap* ytaps, int batch, cudaStream_t st) {
  for (int l = 1; l < P.nlevels; l++) {
    const DevLevel& D = P.lv[l];
    const DevLevel& S = P.lv[l - 1];
    const uint8_t* src;
    int spitch;
    long long sframe;
    if (l == 1) { src = l0.base; spitch = l0.pitch; sframe = l0.frame_stride; }
    else { src = arena + S.offset; spitch = S.pitch; sframe = P.arena_bytes; }
    const bool walk = D.fast_resize && (spitch & 3) == 0 && (sframe & 3) == 0 && ((uintptr_t)src & 3) == 0;
    if (walk) {
      dim3 grd((D.w + 127) / 128, (D.h + RS_WARPS * RS_ROWS - 1) / (RS_WARPS * RS_ROWS), batch);
      k_resize_walk<<<grd, RS_WARPS * 32, 0, st>>>(src, spitch, sframe, S.w, S.h, arena + D.offset, D.pitch,
                                                   P.arena_bytes, D.w, D.h, xtaps + D.xtab, ytaps + D.ytab);
    } else {
      dim3 blk(32, 8);
      dim3 grd((D.w + 127) / 128, (D.h + 7) / 8, batch);
      k_resize<<<grd, blk, 0, st>>>(src, spitch, sframe, S.w, S.h, arena + D.offset, D.pitch, P.arena_bytes, D.w,
                                    D.h, xtaps + D.xtab, ytaps + D.ytab);
    }
  }
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// copyMakeBorder(BORDER_REFLECT_101, 19 px) — only for orbx_get_pyramid(with_border=1); the
// compute kernels never read the border.
// ------------------------------------------------------------------------------------------------
__global__ void k_border(const uint8_t* __restrict__ src, int pitch, int w, int h, uint8_t* __restrict__ dst,
                         int dpitch) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= w + 2 * ORBX_EDGE || y >= h + 2 * ORBX_EDGE) return;
  const int sx = orbx_reflect101(x - ORBX_EDGE, w), sy = orbx_reflect101(y - ORBX_EDGE, h);
  dst[(long long)y * dpitch + x] = src[(long long)sy * pitch + sx];
}
cudaError_t orbx_launch_border(const uint8_t* src, int pitch, int w, int h, uint8_t* dst, int dst_pitch,
                               cudaStream_t st) {
  dim3 blk(32, 8), grd((w + 2 * ORBX_EDGE + 31) / 32, (h + 2 * ORBX_EDGE + 7) / 8);
  k_border<<<grd, blk, 0, st>>>(src, pitch, w, h, dst, dst_pitch);
  return cudaGetLastError();
}

