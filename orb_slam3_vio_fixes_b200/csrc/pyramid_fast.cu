// pyramid_fast.cu — image pyramid (ORBextractor::ComputePyramid, src/ORBextractor.cc:1170-1195)
// and per-cell FAST detection (cell loop of ComputeKeyPointsOctTree, :781-872) for sm_100a.
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

// ------------------------------------------------------------------------------------------------
// FAST-9/16 + per-cell non-max suppression + per-cell threshold fallback.
//
// The reference runs cv::FAST on each ~35-px cell ROI (6 px overlap).  cv::FAST tests ROI pixels
// [3,W-3)x[3,H-3), so the tested zones of the cells tile [19,w-19)x[19,h-19) exactly on a
// (wCell,hCell) lattice; the NMS of a cell only sees that cell's own zone (outside = score 0).
// One CTA = one cell of one level of one frame:
//   1. stage the zone + 3-px halo in shared memory,
//   2. arc strength A per zone pixel (threshold independent; cornerScore = A-1),
//   3. NMS at iniThFAST; if the cell ends up empty, NMS again at minThFAST (:843-859),
//   4. ordered (row-major) compaction into the cell's slot array.
// ------------------------------------------------------------------------------------------------
#define FAST_THREADS 256
#define FAST_MAX_CELL 70  // wCell,hCell < 70 by construction (width/35 cells)

struct FastSmem {
  uint8_t px[(FAST_MAX_CELL + 6) * (FAST_MAX_CELL + 8)];  // zone + halo, pitch = zw+6 (<=76) rounded
  uint8_t amap[(FAST_MAX_CELL + 2) * (FAST_MAX_CELL + 2)];  // A with a zero ring
  int gcount[(FAST_MAX_CELL * FAST_MAX_CELL + 31) / 32 + 1];
  int any;
};

__device__ __forceinline__ bool nms_keep(const uint8_t* a, int ap, int th) {
  // a -> centre in amap (pitch ap).  score = A-1 if A>th else 0, strict > on all 8 neighbours.
  const int A = a[0];
  if (A <= th) return false;
  const int s = A - 1;
#define NB(o) { const int B = a[o]; const int sb = B > th ? B - 1 : 0; if (!(s > sb)) return false; }
  NB(-1) NB(1) NB(-ap - 1) NB(-ap) NB(-ap + 1) NB(ap - 1) NB(ap) NB(ap + 1)
#undef NB
  return true;
}

__global__ void __launch_bounds__(FAST_THREADS)
k_fast(const __grid_constant__ DevPlan P, Level0 l0, const uint8_t* __restrict__ arena,
       int32_t* __restrict__ cell_cnt, uint32_t* __restrict__ cand) {
  __shared__ FastSmem sm;
  const int frame = blockIdx.y;
  int cell = blockIdx.x;
  int level = 0;
#pragma unroll 1
  for (int l = 1; l < P.nlevels; l++)
    if (cell >= P.lv[l].cell_base) level = l;
  const DevLevel& L = P.lv[level];
  cell -= L.cell_base;
  const int ci = cell / L.nCols, cj = cell - ci * L.nCols;
  // tested zone of this cell, level coordinates
  const int zx0 = ORBX_EDGE + cj * L.wCell, zy0 = ORBX_EDGE + ci * L.hCell;
  const int zx1 = min(zx0 + L.wCell, L.w - ORBX_EDGE), zy1 = min(zy0 + L.hCell, L.h - ORBX_EDGE);
  const int zw = zx1 - zx0, zh = zy1 - zy0;
  int32_t* cnt_out = cell_cnt + (long long)frame * P.total_cells + L.cell_base + cell;
  if (zw <= 0 || zh <= 0) {  // cell skipped / empty in the reference (:810,:819)
    if (threadIdx.x == 0) *cnt_out = 0;
    return;
  }
  int pitch;
  const uint8_t* img = level_ptr(P, l0, arena, frame, level, &pitch);
  const int pp = zw + 6;  // px pitch
  const int ap = zw + 2;  // amap pitch
  // 1. stage pixels [zx0-3, zx1+3) x [zy0-3, zy1+3)
  for (int i = threadIdx.x; i < pp * (zh + 6); i += FAST_THREADS) {
    const int y = i / pp, x = i - y * pp;
    sm.px[i] = img[(long long)(zy0 - 3 + y) * pitch + (zx0 - 3 + x)];
  }
  for (int i = threadIdx.x; i < ap * (zh + 2); i += FAST_THREADS) sm.amap[i] = 0;
  if (threadIdx.x == 0) sm.any = 0;
  __syncthreads();
  // 2. arc strength
  const int tmin = min(P.ini_th, P.min_th);
  const int o1 = pp, o2 = 2 * pp, o3 = 3 * pp;
  for (int i = threadIdx.x; i < zw * zh; i += FAST_THREADS) {
    const int y = i / zw, x = i - y * zw;
    const uint8_t* c = sm.px + (y + 3) * pp + (x + 3);
    const int v = c[0];
    // OpenCV ring order: (0,3)(1,3)(2,2)(3,1)(3,0)(3,-1)(2,-2)(1,-3)(0,-3)(-1,-3)(-2,-2)(-3,-1)
    //                    (-3,0)(-3,1)(-2,2)(-1,3)
    int ring[16];
    ring[0] = c[o3]; ring[8] = c[-o3];
    const int lo = v - tmin, hi = v + tmin;
    // every arc of 9 holds one pixel of each opposite pair: quick reject on 4 pairs
    int m = ((ring[0] < lo) | ((ring[0] > hi) << 1)) | ((ring[8] < lo) | ((ring[8] > hi) << 1));
    if (m == 0) continue;
    ring[4] = c[3]; ring[12] = c[-3];
    m &= ((ring[4] < lo) | ((ring[4] > hi) << 1)) | ((ring[12] < lo) | ((ring[12] > hi) << 1));
    if (m == 0) continue;
    ring[2] = c[o2 + 2]; ring[10] = c[-o2 - 2];
    m &= ((ring[2] < lo) | ((ring[2] > hi) << 1)) | ((ring[10] < lo) | ((ring[10] > hi) << 1));
    if (m == 0) continue;
    ring[6] = c[-o2 + 2]; ring[14] = c[o2 - 2];
    m &= ((ring[6] < lo) | ((ring[6] > hi) << 1)) | ((ring[14] < lo) | ((ring[14] > hi) << 1));
    if (m == 0) continue;
    ring[1] = c[o3 + 1]; ring[3] = c[o1 + 3]; ring[5] = c[-o1 + 3]; ring[7] = c[-o3 + 1];
    ring[9] = c[-o3 - 1]; ring[11] = c[-o1 - 3]; ring[13] = c[o1 - 3]; ring[15] = c[o3 - 1];
    const int A = orbx_arc_strength(v, ring);
    if (A > tmin) sm.amap[(y + 1) * ap + (x + 1)] = (uint8_t)A;
  }
  __syncthreads();
  // 3. NMS at iniTh; fall back to minTh if the cell has no survivor
  int any = 0;
  for (int i = threadIdx.x; i < zw * zh; i += FAST_THREADS) {
    const int y = i / zw, x = i - y * zw;
    any |= nms_keep(sm.amap + (y + 1) * ap + (x + 1), ap, P.ini_th);
  }
  any = __syncthreads_or(any);
  const int th = any ? P.ini_th : P.min_th;
  // 4. ordered compaction: 32-pixel groups, ballot per group, scan of group counts
  const int npx = zw * zh, ngroups = (npx + 31) >> 5;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int g = warp; g < ngroups; g += FAST_THREADS / 32) {
    const int i = g * 32 + lane;
    bool k = false;
    if (i < npx) {
      const int y = i / zw, x = i - y * zw;
      k = nms_keep(sm.amap + (y + 1) * ap + (x + 1), ap, th);
    }
    const unsigned b = __ballot_sync(0xffffffffu, k);
    if (lane == 0) sm.gcount[g] = __popc(b);
  }
  __syncthreads();
  if (warp == 0) {  // exclusive scan of gcount (ngroups <= 154)
    int carry = 0;
    for (int base = 0; base < ngroups; base += 32) {
      const int g = base + lane;
      int v = g < ngroups ? sm.gcount[g] : 0;
      int inc = v;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
      }
      if (g < ngroups) sm.gcount[g] = carry + inc - v;
      carry += __shfl_sync(0xffffffffu, inc, 31);
    }
    if (lane == 0) {
      sm.gcount[ngroups] = carry;
      *cnt_out = carry;
    }
  }
  __syncthreads();
  uint32_t* out = cand + (long long)frame * P.cand_total + L.cand_base + (long long)cell * L.cell_cap;
  for (int g = warp; g < ngroups; g += FAST_THREADS / 32) {
    const int i = g * 32 + lane;
    bool k = false;
    int x = 0, y = 0;
    if (i < npx) {
      y = i / zw; x = i - y * zw;
      k = nms_keep(sm.amap + (y + 1) * ap + (x + 1), ap, th);
    }
    const unsigned b = __ballot_sync(0xffffffffu, k);
    if (k) {
      const int pos = sm.gcount[g] + __popc(b & ((1u << lane) - 1));
      // coordinates relative to (minBorderX,minBorderY)=(16,16), as DistributeOctTree sees them
      out[pos] = orbx_pack_key(zx0 + x - ORBX_MIN_BORDER, zy0 + y - ORBX_MIN_BORDER,
                               sm.amap[(y + 1) * ap + (x + 1)] - 1);
    }
  }
}

cudaError_t orbx_launch_fast(const DevPlan& P, Level0 l0, const uint8_t* arena, int32_t* cell_cnt,
                             uint32_t* cand, int batch, cudaStream_t st) {
  dim3 grd(P.total_cells, batch);
  k_fast<<<grd, FAST_THREADS, 0, st>>>(P, l0, arena, cell_cnt, cand);
  return cudaGetLastError();
}
