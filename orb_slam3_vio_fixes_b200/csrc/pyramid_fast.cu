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
// One CTA = one cell of one level of one frame.  The work is organised around compacted lists so
// that warps stay converged:
//   1. stage the zone + 3-px halo in shared memory (32-bit loads),
//   2. quick test on every zone pixel (4 opposite ring pairs) -> ordered list of survivors,
//   3. arc strength A for the listed pixels only (cornerScore = A-1, threshold independent),
//   4. NMS of the corners at iniThFAST; if none survives, NMS again at minThFAST (:843-859),
//   5. ordered (row-major) emission into the cell's slot array.
// ------------------------------------------------------------------------------------------------
#define FAST_THREADS 256
#define FAST_WARPS (FAST_THREADS / 32)
#define FAST_MAX_CELL 70                                 // wCell,hCell < 70 by construction
#define FAST_PXW ((FAST_MAX_CELL + 6 + 3 + 3) / 4 + 1)   // words per staged row
#define FAST_MAX_GROUPS ((FAST_MAX_CELL * FAST_MAX_CELL + 31) / 32)

struct FastSmem {
  uint32_t px[FAST_PXW * (FAST_MAX_CELL + 6)];              // zone + halo, word-aligned rows
  uint8_t amap[(FAST_MAX_CELL + 2) * (FAST_MAX_CELL + 2)];  // A with a zero ring
  uint32_t mask[FAST_MAX_GROUPS + 1];                       // ballots of a compaction pass
  int goff[FAST_MAX_GROUPS + 2];                            // their exclusive scan
  uint16_t list[FAST_MAX_CELL * FAST_MAX_CELL];             // pixels that pass the quick test
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

// exclusive scan of popc(mask[g]), g < ngroups, by warp 0; goff[ngroups] = total
__device__ __forceinline__ void scan_groups(FastSmem& sm, int ngroups, int lane) {
  int carry = 0;
  for (int base = 0; base < ngroups; base += 32) {
    const int g = base + lane;
    const int v = g < ngroups ? __popc(sm.mask[g]) : 0;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    if (g < ngroups) sm.goff[g] = carry + inc - v;
    carry += __shfl_sync(0xffffffffu, inc, 31);
  }
  if (lane == 0) sm.goff[ngroups] = carry;
}

__global__ void __launch_bounds__(FAST_THREADS)
k_fast(const __grid_constant__ DevPlan P, Level0 l0, const uint8_t* __restrict__ arena,
       int32_t* __restrict__ cell_cnt, uint32_t* __restrict__ cand) {
  __shared__ FastSmem sm;
  const int frame = blockIdx.y;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
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
    if (tid == 0) *cnt_out = 0;
    return;
  }
  int pitch;
  const uint8_t* img = level_ptr(P, l0, arena, frame, level, &pitch);
  const int ap = zw + 2;  // amap pitch
  const int npx = zw * zh;
  const uint32_t rcp = ((1u << 20) + zw - 1) / zw;  // i/zw == (i*rcp)>>20 for i*zw < 2^20

  // 1. stage pixels [zx0-3, zx1+3) x [zy0-3, zy1+3); row r, column c at pxb[r*pwb + c]
  const uint8_t* tile0 = img + (long long)(zy0 - 3) * pitch + (zx0 - 3);
  int sh, pwb;
  if ((pitch & 3) == 0) {
    sh = (int)((uintptr_t)tile0 & 3);
    const int nw = (sh + zw + 6 + 3) >> 2;  // words per row
    pwb = nw * 4;
    const uint8_t* a0 = tile0 - sh;
    const uint32_t rcw = ((1u << 20) + nw - 1) / nw;
    for (int i = tid; i < nw * (zh + 6); i += FAST_THREADS) {
      const int r = (int)(((uint32_t)i * rcw) >> 20), c = i - r * nw;
      sm.px[i] = __ldg(reinterpret_cast<const uint32_t*>(a0 + (long long)r * pitch) + c);
    }
  } else {
    sh = 0;
    pwb = (zw + 6 + 3) & ~3;
    uint8_t* pb = reinterpret_cast<uint8_t*>(sm.px);
    for (int i = tid; i < (zw + 6) * (zh + 6); i += FAST_THREADS) {
      const int r = i / (zw + 6), c = i - r * (zw + 6);
      pb[r * pwb + c] = tile0[(long long)r * pitch + c];
    }
  }
  for (int i = tid; i < (ap * (zh + 2) + 3) >> 2; i += FAST_THREADS) reinterpret_cast<uint32_t*>(sm.amap)[i] = 0;
  __syncthreads();
  const uint8_t* pxb = reinterpret_cast<const uint8_t*>(sm.px) + sh;

  // 2. quick test: every arc of 9 holds one pixel of each opposite ring pair, so a corner at
  //    threshold t needs, for all four pairs (0,8)(4,12)(2,10)(6,14), a member darker than v-t —
  //    or, for all four, a member brighter than v+t.
  const int tmin = min(P.ini_th, P.min_th);
  const int o1 = pwb, o2 = 2 * pwb, o3 = 3 * pwb;
  const int ngroups = (npx + 31) >> 5;
  for (int g = warp; g < ngroups; g += FAST_WARPS) {
    const int i = g * 32 + lane;
    bool pass = false;
    if (i < npx) {
      const int y = (int)(((uint32_t)i * rcp) >> 20), x = i - y * zw;
      const uint8_t* c = pxb + (y + 3) * pwb + (x + 3);
      const int v = c[0], lo = v - tmin, hi = v + tmin;
      const int r0 = c[o3], r8 = c[-o3], r4 = c[3], r12 = c[-3];
      const int r2 = c[o2 + 2], r10 = c[-o2 - 2], r6 = c[-o2 + 2], r14 = c[o2 - 2];
      const bool dark = ((r0 < lo) | (r8 < lo)) & ((r4 < lo) | (r12 < lo)) & ((r2 < lo) | (r10 < lo)) &
                        ((r6 < lo) | (r14 < lo));
      const bool bright = ((r0 > hi) | (r8 > hi)) & ((r4 > hi) | (r12 > hi)) & ((r2 > hi) | (r10 > hi)) &
                          ((r6 > hi) | (r14 > hi));
      pass = dark | bright;
    }
    const unsigned b = __ballot_sync(0xffffffffu, pass);
    if (lane == 0) sm.mask[g] = b;
  }
  __syncthreads();
  if (warp == 0) scan_groups(sm, ngroups, lane);
  __syncthreads();
  for (int g = warp; g < ngroups; g += FAST_WARPS) {
    const unsigned b = sm.mask[g];
    if ((b >> lane) & 1u) sm.list[sm.goff[g] + __popc(b & ((1u << lane) - 1))] = (uint16_t)(g * 32 + lane);
  }
  const int nlist = sm.goff[ngroups];
  __syncthreads();

  // 3. arc strength of the listed pixels (dense: warps stay converged)
  for (int e = tid; e < nlist; e += FAST_THREADS) {
    const int i = sm.list[e];
    const int y = (int)(((uint32_t)i * rcp) >> 20), x = i - y * zw;
    const uint8_t* c = pxb + (y + 3) * pwb + (x + 3);
    // OpenCV ring order: (0,3)(1,3)(2,2)(3,1)(3,0)(3,-1)(2,-2)(1,-3)(0,-3)(-1,-3)(-2,-2)(-3,-1)
    //                    (-3,0)(-3,1)(-2,2)(-1,3)
    int ring[16];
    ring[0] = c[o3]; ring[1] = c[o3 + 1]; ring[2] = c[o2 + 2]; ring[3] = c[o1 + 3];
    ring[4] = c[3]; ring[5] = c[-o1 + 3]; ring[6] = c[-o2 + 2]; ring[7] = c[-o3 + 1];
    ring[8] = c[-o3]; ring[9] = c[-o3 - 1]; ring[10] = c[-o2 - 2]; ring[11] = c[-o1 - 3];
    ring[12] = c[-3]; ring[13] = c[o1 - 3]; ring[14] = c[o2 - 2]; ring[15] = c[o3 - 1];
    const int A = orbx_arc_strength(c[0], ring);
    if (A > tmin) sm.amap[(y + 1) * ap + (x + 1)] = (uint8_t)A;
  }
  __syncthreads();

  // 4. NMS over the listed pixels at iniTh; fall back to minTh if the cell has no survivor.
  //    keep bit k of `mine` belongs to list entry tid + k*FAST_THREADS (k < 20).
  uint32_t mine = 0;
  for (int e = tid, k = 0; e < nlist; e += FAST_THREADS, k++) {
    const int i = sm.list[e];
    const int y = (int)(((uint32_t)i * rcp) >> 20), x = i - y * zw;
    if (nms_keep(sm.amap + (y + 1) * ap + (x + 1), ap, P.ini_th)) mine |= 1u << k;
  }
  const int any = __syncthreads_or(mine != 0);
  if (!any) {
    for (int e = tid, k = 0; e < nlist; e += FAST_THREADS, k++) {
      const int i = sm.list[e];
      const int y = (int)(((uint32_t)i * rcp) >> 20), x = i - y * zw;
      if (nms_keep(sm.amap + (y + 1) * ap + (x + 1), ap, P.min_th)) mine |= 1u << k;
    }
  }
  // 5. ordered emission: list entries e = k*256 + warp*32 + lane form group k*8 + warp
  const int ng2 = (nlist + 31) >> 5;
  for (int e = tid, k = 0; e - lane < nlist; e += FAST_THREADS, k++) {
    const unsigned b = __ballot_sync(0xffffffffu, (mine >> k) & 1u);
    if (lane == 0) sm.mask[k * FAST_WARPS + warp] = b;
  }
  __syncthreads();
  if (warp == 0) {
    scan_groups(sm, ng2, lane);
    if (lane == 0) *cnt_out = sm.goff[ng2];
  }
  __syncthreads();
  uint32_t* out = cand + (long long)frame * P.cand_total + L.cand_base + (long long)cell * L.cell_cap;
  for (int e = tid, k = 0; e < nlist; e += FAST_THREADS, k++) {
    if ((mine >> k) & 1u) {
      const int g = k * FAST_WARPS + warp;
      const int pos = sm.goff[g] + __popc(sm.mask[g] & ((1u << lane) - 1));
      const int i = sm.list[e];
      const int y = (int)(((uint32_t)i * rcp) >> 20), x = i - y * zw;
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
