// fast.cu — per-cell FAST-9/16 detection with non-max suppression and the per-cell threshold
// fallback (cell loop of ORBextractor::ComputeKeyPointsOctTree, src/ORBextractor.cc:781-872;
// cv::FAST TYPE_9_16 + cornerScore) for sm_100a.
//
// The reference runs cv::FAST on each ~35-px cell ROI (6 px overlap).  cv::FAST tests ROI pixels
// [3,W-3)x[3,H-3), so the tested zones of the cells tile [19,w-19)x[19,h-19) exactly on a
// (wCell,hCell) lattice, and the NMS of a cell only sees that cell's own zone (outside = 0).
//
// One WARP = one cell (no block-level synchronisation at all; a CTA is just 4 independent warps):
//   1. stage the zone + 3-px halo in the warp's shared-memory slice with 32-bit loads;
//   2. quick test, four pixels (one 32-bit word, a "quad") per lane with packed-byte arithmetic
//      (VABSDIFF4): a 9/16 corner at threshold t needs |v-ring| > t on at least one pixel of each
//      opposite ring pair (0,8)(4,12)(2,10)(6,14); surviving quads are compacted in order;
//   3. arc strength A of the surviving quads only, dense, with packed 16-bit SIMD
//      (VIMNMX3.U16x2: the 16 sliding windows of 9 ring pixels are two rounds of 3-input min/max);
//      corner(th) <=> A > th and cornerScore == A-1, so one pass serves both thresholds;
//   4. NMS at iniThFAST; if the cell ends up empty, NMS again at minThFAST (:843-859);
//   5. ordered (row-major) emission into the cell's slot array.
#include "orbx_kernels.cuh"

#define FAST_WARPS 4
#define FAST_THREADS (FAST_WARPS * 32)

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
  for (int k = 0; k < 16; k++) {
    const uint32_t mx9 = __vimax3_u16x2(mx3[k], mx3[(k + 3) & 15], mx3[(k + 6) & 15]);
    const uint32_t mn9 = __vimin3_u16x2(mn3[k], mn3[(k + 3) & 15], mn3[(k + 6) & 15]);
    lo = __vminu2(lo, mx9);
    hi = __vmaxu2(hi, mn9);
  }
  const int v0 = (int)(vv & 0xffffu), v1 = (int)(vv >> 16);
  const int a0 = max(max(v0 - (int)(lo & 0xffffu), (int)(hi & 0xffffu) - v0), 0);
  const int a1 = max(max(v1 - (int)(lo >> 16), (int)(hi >> 16) - v1), 0);
  return (uint32_t)a0 | ((uint32_t)a1 << 8);
}

// NMS decision for the pixel at byte j of the centre word: score = A-1 if A>th else 0, strict > on
// all 8 neighbours (cv::FAST_t, nonmax_suppression branch).  n[3][3] = the 3x3 words around it.
__device__ __forceinline__ bool nms_keep(const uint32_t (&n)[3][3], int j, int th) {
  auto at = [&](int row, int col) -> int {  // col in [-1, 4]
    const uint32_t w = col < 0 ? n[row][0] : (col > 3 ? n[row][2] : n[row][1]);
    return (int)((w >> (8 * (col & 3))) & 0xffu);
  };
  const int A = at(1, j);
  if (A <= th) return false;
  const int s = A - 1;
#pragma unroll
  for (int dr = 0; dr < 3; dr++)
#pragma unroll
    for (int dc = -1; dc <= 1; dc++) {
      if (dr == 1 && dc == 0) continue;
      const int B = at(dr, j + dc);
      const int sb = B > th ? B - 1 : 0;
      if (!(s > sb)) return false;
    }
  return true;
}

// per byte: 0x80 where x > t (t <= 126), via the carry into bit 7; c = (0x7f - t) * 0x01010101
__device__ __forceinline__ uint32_t gt_bytes(uint32_t x, uint32_t c) {
  return ((x & 0x7f7f7f7fu) + c) | x;  // bit 7 of each byte is the answer (mask later)
}

// Shared memory of one warp: px[pw*rows] | am[pw*rows] | list[list_cap] (u16), all in 32-bit words.
// px: pixel (tile row r, tile col c) at byte r*4*pw + 4 + sh + c;  am: arc strength A (u8), same
// geometry, 0 outside the zone;  list: quads that passed the quick test, in row-major order.
__global__ void __launch_bounds__(FAST_THREADS)
k_fast(const __grid_constant__ DevPlan P, Level0 l0, const uint8_t* __restrict__ arena,
       int32_t* __restrict__ cell_cnt, uint32_t* __restrict__ cand, int pw, int rows_max, int list_cap) {
  extern __shared__ uint32_t fast_smem[];
  const int frame = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int cell = blockIdx.x * FAST_WARPS + warp;
  if (cell >= P.total_cells) return;
  const int region = 2 * pw * rows_max + ((list_cap + 1) >> 1);
  uint32_t* px = fast_smem + warp * region;
  uint32_t* am = px + pw * rows_max;
  uint16_t* list = reinterpret_cast<uint16_t*>(am + pw * rows_max);
  uint8_t* flags = reinterpret_cast<uint8_t*>(px);  // px is dead once the arcs are done

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
    if (lane == 0) *cnt_out = 0;
    return;
  }
  int pitch;
  const uint8_t* img;
  if (level == 0) { pitch = l0.pitch; img = l0.base + (long long)frame * l0.frame_stride; }
  else { pitch = L.pitch; img = arena + (long long)frame * P.arena_bytes + L.offset; }

  // 1. stage pixels [zx0-3, zx1+3) x [zy0-3, zy1+3)
  const uint8_t* tile0 = img + (long long)(zy0 - 3) * pitch + (zx0 - 3);
  int sh;
  if ((pitch & 3) == 0) {
    sh = (int)((uintptr_t)tile0 & 3);
    const int nw = (sh + zw + 6 + 3) >> 2;  // data words per row
    const uint8_t* a0 = tile0 - sh;
    const uint32_t rcw = ((1u << 20) + nw - 1) / nw;
    for (int i = lane; i < nw * (zh + 6); i += 32) {
      const int r = (int)(((uint32_t)i * rcw) >> 20), c = i - r * nw;
      px[r * pw + 1 + c] = __ldg(reinterpret_cast<const uint32_t*>(a0 + (long long)r * pitch) + c);
    }
  } else {
    sh = 0;
    uint8_t* pb = reinterpret_cast<uint8_t*>(px);
    for (int i = lane; i < (zw + 6) * (zh + 6); i += 32) {
      const int r = i / (zw + 6), c = i - r * (zw + 6);
      pb[r * (pw * 4) + 4 + c] = tile0[(long long)r * pitch + c];
    }
  }
  // Quads: word q of tile row r holds smem columns 4q..4q+3 = zone x 4q-(7+sh)+j, j=0..3.
  const int colbase = 7 + sh;  // smem column of zone x = 0
  const int q_first = colbase >> 2, q_last = (colbase + zw - 1) >> 2;
  const int nqr = q_last - q_first + 1;  // quads per zone row
  const int nquads = nqr * zh;
  const uint32_t rcq = ((1u << 20) + nqr - 1) / nqr;
  const int tmin = min(P.ini_th, P.min_th);
  // zero ring of the A map: rows just above/below the zone and the words left/right of it
  for (int i = lane; i < nqr + 2; i += 32) {
    am[2 * pw + q_first - 1 + i] = 0;
    am[(zh + 3) * pw + q_first - 1 + i] = 0;
  }
  for (int i = lane; i < zh; i += 32) {
    am[(i + 3) * pw + q_first - 1] = 0;
    am[(i + 3) * pw + q_last + 1] = 0;
  }
  __syncwarp();

  // 2. quick test per quad, ordered compaction of the survivors
  const bool quick = tmin <= 126;
  const uint32_t cgt = (uint32_t)(0x7f - (quick ? tmin : 0)) * 0x01010101u;
  int nlist = 0;
  for (int base = 0; base < nquads; base += 32) {
    const int ql = base + lane;
    bool pass = false;
    if (ql < nquads) {
      const int yy = (int)(((uint32_t)ql * rcq) >> 20), q = q_first + ql - yy * nqr;
      const uint32_t* p = px + yy * pw + q;  // tile row yy = zone row yy - 3
      const uint32_t v = p[3 * pw];
      am[(yy + 3) * pw + q] = 0;
      if (quick) {
        // ring pixels 0 (0,+3) and 8 (0,-3): same columns
        uint32_t g = gt_bytes(__vabsdiffu4(v, p[6 * pw]), cgt) | gt_bytes(__vabsdiffu4(v, p[0]), cgt);
        // 4 (+3,0) and 12 (-3,0)
        const uint32_t c0 = p[3 * pw - 1], c2 = p[3 * pw + 1];
        g &= gt_bytes(__vabsdiffu4(v, __byte_perm(v, c2, 0x6543)), cgt) |
             gt_bytes(__vabsdiffu4(v, __byte_perm(c0, v, 0x4321)), cgt);
        // 2 (+2,+2) and 10 (-2,-2)
        const uint32_t d0 = p[5 * pw - 1], d1 = p[5 * pw], d2 = p[5 * pw + 1];
        const uint32_t u0 = p[pw - 1], u1 = p[pw], u2 = p[pw + 1];
        g &= gt_bytes(__vabsdiffu4(v, __byte_perm(d1, d2, 0x5432)), cgt) |
             gt_bytes(__vabsdiffu4(v, __byte_perm(u0, u1, 0x5432)), cgt);
        // 6 (+2,-2) and 14 (-2,+2)
        g &= gt_bytes(__vabsdiffu4(v, __byte_perm(u1, u2, 0x5432)), cgt) |
             gt_bytes(__vabsdiffu4(v, __byte_perm(d0, d1, 0x5432)), cgt);
        pass = (g & 0x80808080u) != 0;
      } else {
        pass = true;
      }
    }
    const unsigned b = __ballot_sync(0xffffffffu, pass);
    if (pass) list[nlist + __popc(b & ((1u << lane) - 1))] = (uint16_t)ql;
    nlist += __popc(b);
  }
  __syncwarp();

  // 3. arc strength of the listed quads (dense: the warp stays converged)
  for (int e = lane; e < nlist; e += 32) {
    const int ql = list[e];
    const int yy = (int)(((uint32_t)ql * rcq) >> 20), q = q_first + ql - yy * nqr;
    uint32_t rw[7][3];
#pragma unroll
    for (int d = 0; d < 7; d++) {
      const uint32_t* p = px + (yy + d) * pw + q - 1;
      rw[d][0] = p[0]; rw[d][1] = p[1]; rw[d][2] = p[2];
    }
    const uint32_t a = arc_pair<0>(rw) | (arc_pair<2>(rw) << 16);
    // keep A only where it can matter (A > tmin) and only inside the zone
    const int x0 = 4 * q - colbase;
    uint32_t keep = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int A = (int)((a >> (8 * j)) & 0xffu);
      if (A > tmin && (unsigned)(x0 + j) < (unsigned)zw) keep |= 0xffu << (8 * j);
    }
    am[(yy + 3) * pw + q] = a & keep;
  }
  __syncwarp();

  // 4. NMS at iniTh over the listed quads; fall back to minTh when the cell has no survivor
  int th = P.ini_th;
  for (int pass = 0; pass < 2; pass++) {
    unsigned any = 0;
    for (int base = 0; base < nlist; base += 32) {
      const int e = base + lane;
      uint32_t f = 0;
      if (e < nlist) {
        const int ql = list[e];
        const int yy = (int)(((uint32_t)ql * rcq) >> 20), q = q_first + ql - yy * nqr;
        if (am[(yy + 3) * pw + q] != 0) {
          uint32_t n[3][3];
#pragma unroll
          for (int d = 0; d < 3; d++) {
            const uint32_t* p = am + (yy + 2 + d) * pw + q - 1;
            n[d][0] = p[0]; n[d][1] = p[1]; n[d][2] = p[2];
          }
#pragma unroll
          for (int j = 0; j < 4; j++)
            if (nms_keep(n, j, th)) f |= 1u << j;
        }
        flags[e] = (uint8_t)f;
      }
      any |= __ballot_sync(0xffffffffu, f != 0);
    }
    if (any || P.min_th == th) break;
    th = P.min_th;
  }
  __syncwarp();

  // 5. ordered emission (list order == row-major quad order, bytes of a quad left to right)
  uint32_t* out = cand + (long long)frame * P.cand_total + L.cand_base + (long long)cell * L.cell_cap;
  int total = 0;
  for (int base = 0; base < nlist; base += 32) {
    const int e = base + lane;
    const unsigned f = e < nlist ? flags[e] : 0u;
    const unsigned b = __ballot_sync(0xffffffffu, f != 0);
    if (b == 0) continue;
    const int c = __popc(f);
    int inc = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    if (f) {
      const int ql = list[e];
      const int yy = (int)(((uint32_t)ql * rcq) >> 20), q = q_first + ql - yy * nqr;
      const uint32_t a = am[(yy + 3) * pw + q];
      int pos = total + inc - c;
#pragma unroll
      for (int j = 0; j < 4; j++)
        if ((f >> j) & 1u) {
          // coordinates relative to (minBorderX,minBorderY)=(16,16), as DistributeOctTree sees them
          out[pos++] = orbx_pack_key(zx0 + 4 * q - colbase + j - ORBX_MIN_BORDER, zy0 + yy - ORBX_MIN_BORDER,
                                     (int)((a >> (8 * j)) & 0xffu) - 1);
        }
    }
    total += __shfl_sync(0xffffffffu, inc, 31);
  }
  if (lane == 0) *cnt_out = total;
}

static int g_fast_smem_set = 0;

cudaError_t orbx_launch_fast(const DevPlan& P, Level0 l0, const uint8_t* arena, int32_t* cell_cnt,
                             uint32_t* cand, int batch, cudaStream_t st) {
  // shared-memory geometry from the largest cell of the plan
  int pw = 0, rows = 0, cap = 0;
  for (int l = 0; l < P.nlevels; l++) {
    const int nw = (3 + P.lv[l].wCell + 6 + 3) >> 2;
    pw = max(pw, nw + 2);
    rows = max(rows, P.lv[l].hCell + 6);
    cap = max(cap, ((P.lv[l].wCell + 3 + 3) / 4 + 2) * P.lv[l].hCell);
  }
  const size_t smem = (size_t)FAST_WARPS * (2 * pw * rows + ((cap + 1) >> 1)) * sizeof(uint32_t);
  if ((int)smem > g_fast_smem_set) {
    cudaError_t e = cudaFuncSetAttribute(k_fast, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    g_fast_smem_set = (int)smem;
  }
  dim3 grd((P.total_cells + FAST_WARPS - 1) / FAST_WARPS, batch);
  k_fast<<<grd, FAST_THREADS, smem, st>>>(P, l0, arena, cell_cnt, cand, pw, rows, cap);
  return cudaGetLastError();
}
