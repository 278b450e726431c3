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
    dim3 blk(32, 8);
    dim3 grd((D.w + 127) / 128, (D.h + 7) / 8, batch);
    k_resize<<<grd, blk, 0, st>>>(src, spitch, sframe, S.w, S.h, arena + D.offset, D.pitch, P.arena_bytes, D.w,
                                  D.h, xtaps + D.xtab, ytaps + D.ytab);
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

