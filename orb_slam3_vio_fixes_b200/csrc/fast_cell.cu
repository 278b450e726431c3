// fast_cell.cu — (first-generation kernel, kept selectable with ORBX_FAST_IMPL=cell) per-cell FAST-9/16 detection with non-max suppression and the per-cell threshold
// fallback (cell loop of ORBextractor::ComputeKeyPointsOctTree, src/ORBextractor.cc:781-872;
// cv::FAST TYPE_9_16 + cornerScore) for sm_100a.
//
// The reference runs cv::FAST on each ~35-px cell ROI (6 px overlap).  cv::FAST tests ROI pixels
// [3,W-3)x[3,H-3), so the tested zones of the cells tile [19,w-19)x[19,h-19) exactly on a
// (wCell,hCell) lattice, and the NMS of a cell only sees that cell's own zone (outside = 0).
//
// One WARP = one cell (no block-level synchronisation at all; a CTA is just 4 independent warps):
//   1. stage the zone + 3-px halo in the warp's shared-memory slice with 32-bit loads, plus a
//      copy with every byte halved (7-bit pixels) for the quick test;
//   2. quick test at threshold t, four pixels (one 32-bit word, a "quad") per lane per step with
//      packed-byte arithmetic (VABSDIFF4): a 9/16 corner needs |v-ring| > t on at least one pixel
//      of each opposite ring pair (0,8)(4,12)(2,10)(6,14).  On 7-bit pixels the test
//      |v'-r'| >= ceil(t/2) is implied by |v-r| > t (never rejects a corner) and the compare is one
//      add whose carry lands in bit 7.  A lane walks down one quad column keeping the last 7 rows
//      in registers, so a step costs 3 shared loads; the pass masks go to a byte map;
//   3. the byte map is compacted, in row-major order, into a list of pixel PAIRS to score;
//   4. arc strength A of the listed pairs only, with packed 16-bit SIMD (VIMNMX3.U16x2: the 16
//      sliding windows of 9 ring pixels are two rounds of 3-input min/max); corner(t) <=> A > t and
//      cornerScore == A-1;
//   5. NMS over the listed pairs;
//   6. if nothing survives at iniThFAST the cell is redone at minThFAST (:843-859), else
//   7. ordered (row-major) emission into the cell's slot array.
#include <stdlib.h>

#include "orbx_kernels.cuh"

#define FAST_WARPS 4
#define FAST_THREADS (FAST_WARPS * 32)
// Opposite ring pairs the quick test checks: 4 = (0,8)(4,12)(2,10)(6,14); 2 = the axial pairs only.  On the
// SURVEY frames the diagonal pairs reject only 14 % of what the axial pairs let through (8.1 % -> 7.0 % of the
// pixels) but cost half of the quick test's ALU-pipe instructions, so 2 is faster (measured: DESIGN.md §4).
#ifndef FAST_QUICK_PAIRS
#define FAST_QUICK_PAIRS 2
#endif
#define FAST_LIST_CAP 512  // pair tasks scored per band of rows (a cell is one band unless it is corner-dense)

#include "fast_arc.cuh"

// NMS for the two pixels of a pair.  w0,w1,w2 = the rows above / at / below, each holding the A
// values (0 where A <= th) of columns c-1, c, c+1, c+2 with the pair at c, c+1.  cv::FAST_t keeps a
// corner when its score A-1 is strictly greater than all 8 neighbouring scores (0 for non-corners):
//   keep <=> A != 0 && A > max(1, max of the 8 neighbours' A).
// Returns bit0 / bit1 for the left / right pixel.
__device__ __forceinline__ uint32_t nms_pair(uint32_t w0, uint32_t w1, uint32_t w2) {
  const uint32_t e0 = w0 & 0x00ff00ffu, e1 = w1 & 0x00ff00ffu, e2 = w2 & 0x00ff00ffu;  // columns 0,2
  const uint32_t o0 = (w0 >> 8) & 0x00ff00ffu, o1 = (w1 >> 8) & 0x00ff00ffu, o2 = (w2 >> 8) & 0x00ff00ffu;  // 1,3
  const uint32_t ve = __vimax3_u16x2(e0, e1, e2), vo = __vimax3_u16x2(o0, o1, o2);  // column maxima
  const uint32_t te = __vmaxu2(e0, e2), to = __vmaxu2(o0, o2);                      // without the centre row
  const uint32_t ma = max(max(ve & 0xffffu, ve >> 16), max(to & 0xffffu, 1u));  // around column 1
  const uint32_t mb = max(max(vo & 0xffffu, vo >> 16), max(te >> 16, 1u));      // around column 2
  const uint32_t a = o1 & 0xffffu, b = e1 >> 16;
  return (a > ma ? 1u : 0u) | (b > mb ? 2u : 0u);
}

// One step of the quick test for the quad column a lane walks down.  C/L2/R2/L3/R3 hold, per
// staged row (slot = row mod 7), the centre word and the words shifted by -2, +2, -3, +3 pixels
// (7-bit pixels).  Slot NEW receives tile row `row`; the quad under test sits 3 rows above it.
template <int QP>
struct QuickWin {
  uint32_t C[7], L2[QP == 4 ? 7 : 1], R2[QP == 4 ? 7 : 1], L3[7], R3[7];
};

template <int NEW, int QP>
__device__ __forceinline__ void quick_load(QuickWin<QP>& W, const uint32_t* hrow) {
  const uint32_t n0 = hrow[-1], n1 = hrow[0], n2 = hrow[1];
  W.C[NEW] = n1;
  if (QP == 4) {
    W.L2[QP == 4 ? NEW : 0] = __byte_perm(n0, n1, 0x5432);
    W.R2[QP == 4 ? NEW : 0] = __byte_perm(n1, n2, 0x5432);
  }
  W.L3[NEW] = __byte_perm(n0, n1, 0x4321);
  W.R3[NEW] = __byte_perm(n1, n2, 0x6543);
}

// PH = slot of the oldest row (tile row yy); the centre row is slot PH+3, the newest PH+6.
// Returns the 4-bit mask of pixels that pass.
template <int PH, int QP>
__device__ __forceinline__ uint32_t quick_test(const QuickWin<QP>& W, uint32_t cadd) {
  constexpr int S0 = PH % 7, S1 = (PH + 1) % 7, S3 = (PH + 3) % 7, S5 = (PH + 5) % 7, S6 = (PH + 6) % 7;
  const uint32_t v = W.C[S3];
#define QT(x) (__vabsdiffu4(v, (x)) + cadd)
  uint32_t g = QT(W.C[S6]) | QT(W.C[S0]);     // ring 0 (0,+3) | ring 8 (0,-3)
  g &= QT(W.R3[S3]) | QT(W.L3[S3]);           // ring 4 (+3,0) | ring 12 (-3,0)
  if (QP == 4) {
    g &= QT(W.R2[QP == 4 ? S5 : 0]) | QT(W.L2[QP == 4 ? S1 : 0]);  // ring 2 (+2,+2) | ring 10 (-2,-2)
    g &= QT(W.R2[QP == 4 ? S1 : 0]) | QT(W.L2[QP == 4 ? S5 : 0]);  // ring 6 (+2,-2) | ring 14 (-2,+2)
  }
#undef QT
  return ((g & 0x80808080u) * 0x00204081u) >> 28;  // bits 7,15,23,31 -> 0..3
}

// Shared memory of one warp, in 32-bit words:
//   px[pw*rows]   staged pixels: tile row r (= zone row r-3), zone column x at byte r*4*pw + 7+sh + x
//   hm[pw*rows]   the same tile with 7-bit pixels during the quick test; afterwards the arc-strength
//                 map A (u8, 0 where A <= th or outside the zone), same geometry
//   list[FAST_LIST_CAP] u16 tasks (quad*2 + pair), row-major, for the current band of rows
//   qm[qcap]      u8 per quad: quick-test mask of its 4 pixels
struct FastCell {
  uint32_t *px, *hm;
  uint16_t* list;
  uint8_t* qm;
  int pw, nqr, q_first, colbase, zw, zh;
  uint32_t rcq;
};

// Arc strength of the listed pairs (dense: the warp stays converged) into the A map.
__device__ __forceinline__ void fast_score_list(const FastCell& c, int nlist, int th, int lane) {
  for (int e = lane; e < nlist; e += 32) {
    const int tk = c.list[e];
    const int ql = tk >> 1, pr = tk & 1;
    const int yy = (int)(((uint32_t)ql * c.rcq) >> 20), q = c.q_first + ql - yy * c.nqr;
    const int fs = 16 * pr;
    uint32_t rw[7][3];
#pragma unroll
    for (int d = 0; d < 7; d++) {
      const uint32_t* p = c.px + (yy + d) * c.pw + q - 1;
      const uint32_t w0 = p[0], w1 = p[1], w2 = p[2];
      rw[d][0] = __funnelshift_r(w0, w1, fs);
      rw[d][1] = __funnelshift_r(w1, w2, fs);
      rw[d][2] = w2 >> fs;
    }
    const uint32_t a = arc_pair<0>(rw);  // A of zone columns x0, x0+1
    const int x0 = 4 * q - c.colbase + 2 * pr;
    const int A0 = (int)(a & 0xffu), A1 = (int)(a >> 8);
    uint32_t keep = 0;
    if (A0 > th && (unsigned)x0 < (unsigned)c.zw) keep |= 0x00ffu;
    if (A1 > th && (unsigned)(x0 + 1) < (unsigned)c.zw) keep |= 0xff00u;
    reinterpret_cast<uint16_t*>(c.hm + (yy + 3) * c.pw + q)[pr] = (uint16_t)(a & keep);
  }
  __syncwarp();
}

// NMS over the listed pairs and ordered emission (list order == row-major, pairs and their pixels
// left to right).  Returns the new running total of the cell.
__device__ __forceinline__ int fast_nms_emit(const FastCell& c, int nlist, int total, uint32_t* out, int kx0,
                                             int ky0, int lane) {
  for (int base = 0; base < nlist; base += 32) {
    const int e = base + lane;
    uint32_t f = 0, a = 0;
    int yy = 0, q = 0, pr = 0;
    if (e < nlist) {
      const int tk = c.list[e];
      const int ql = tk >> 1;
      pr = tk & 1;
      yy = (int)(((uint32_t)ql * c.rcq) >> 20);
      q = c.q_first + ql - yy * c.nqr;
      const uint32_t* p = c.hm + (yy + 2) * c.pw + q - 1;
      a = reinterpret_cast<const uint16_t*>(p + c.pw + 1)[pr];
      if (a != 0) {
        // columns c-1..c+2 around the pair at c = 2*pr of word q
        const int fs = pr ? 8 : 24;
        uint32_t w[3];
#pragma unroll
        for (int d = 0; d < 3; d++) {
          const uint32_t lo = pr ? p[d * c.pw + 1] : p[d * c.pw], hi = pr ? p[d * c.pw + 2] : p[d * c.pw + 1];
          w[d] = __funnelshift_r(lo, hi, fs);
        }
        f = nms_pair(w[0], w[1], w[2]);
      }
    }
    const unsigned b = __ballot_sync(0xffffffffu, f != 0);
    if (b == 0) continue;
    const int cnt = __popc(f);
    int inc = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    if (f) {
      int pos = total + inc - cnt;
      // coordinates relative to (minBorderX,minBorderY)=(16,16), as DistributeOctTree sees them
      const int x = kx0 + 4 * q + 2 * pr, y = ky0 + yy;
      if (f & 1u) out[pos++] = orbx_pack_key(x, y, (int)(a & 0xffu) - 1);
      if (f & 2u) out[pos] = orbx_pack_key(x + 1, y, (int)(a >> 8) - 1);
    }
    total += __shfl_sync(0xffffffffu, inc, 31);
  }
  return total;
}

// Walks the quick-test byte map in row-major order, 256 quads (two words of qm per lane) at a time,
// compacting the pixel pairs that passed into the task list.  The list is handed to the scoring
// stage (MODE 0) or to NMS + emission (MODE 1) whenever the next 256 quads might not fit, and at the
// end; *bands counts the hand-overs.  Returns the running emission total (MODE 1).
template <int MODE>
__device__ __forceinline__ int fast_sweep(const FastCell& c, int nquads, int th, int lane, int* nlist_out,
                                          int* bands, int total, uint32_t* out, int kx0, int ky0) {
  int nlist = 0, nb = 0;
  for (int base = 0; base < nquads; base += 256) {
    const int i4 = (base >> 2) + 2 * lane;  // this lane's two words of qm = 8 quads
    uint32_t bits = 0;
    if (4 * i4 < nquads) {
      const uint2 mm = *reinterpret_cast<const uint2*>(c.qm + 4 * i4);
      uint32_t x = (mm.x | (mm.x >> 1)) & 0x05050505u;  // per quad: bit0 = left pair, bit2 = right pair
      uint32_t y = (mm.y | (mm.y >> 1)) & 0x05050505u;
      x = (x | (x >> 1)) & 0x03030303u;                 // -> bits 0,1
      y = (y | (y >> 1)) & 0x03030303u;
      bits = (x | (x >> 6) | (x >> 12) | (x >> 18)) & 0xffu;
      if (4 * i4 + 4 < nquads) bits |= ((y | (y >> 6) | (y >> 12) | (y >> 18)) & 0xffu) << 8;
    }
    const int cnt = __popc(bits);
    int inc = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    const int chunk = __shfl_sync(0xffffffffu, inc, 31);
    if (nlist + chunk > FAST_LIST_CAP) {
      __syncwarp();
      if (MODE == 0) fast_score_list(c, nlist, th, lane);
      else total = fast_nms_emit(c, nlist, total, out, kx0, ky0, lane);
      __syncwarp();
      nlist = 0;
      nb++;
    }
    int pos = nlist + inc - cnt;
    while (bits) {
      const int b = __ffs(bits) - 1;
      bits &= bits - 1;
      c.list[pos++] = (uint16_t)(8 * i4 + b);  // (4*i4 + b/2)*2 + (b&1), b < 16 spans both words
    }
    nlist += chunk;
  }
  __syncwarp();
  if (MODE == 0) fast_score_list(c, nlist, th, lane);
  else total = fast_nms_emit(c, nlist, total, out, kx0, ky0, lane);
  *nlist_out = nlist;
  *bands = nb;
  return total;
}

template <int QP>
__global__ void __launch_bounds__(FAST_THREADS, 8)
k_fast(const __grid_constant__ DevPlan P, Level0 l0, const uint8_t* __restrict__ arena,
       const uint32_t* __restrict__ cell_tab, int32_t* __restrict__ cell_cnt, uint32_t* __restrict__ cand,
       int pw, int rows_max, int qcap) {
  extern __shared__ uint32_t fast_smem[];
  const int frame = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int cell = blockIdx.x * FAST_WARPS + warp;
  if (cell >= P.total_cells) return;
  const int region = 2 * pw * rows_max + FAST_LIST_CAP / 2 + (qcap >> 2);
  FastCell c;
  c.px = fast_smem + warp * region;
  c.hm = c.px + pw * rows_max;
  c.list = reinterpret_cast<uint16_t*>(c.hm + pw * rows_max);
  c.qm = reinterpret_cast<uint8_t*>(c.hm + pw * rows_max + FAST_LIST_CAP / 2);
  c.pw = pw;

  // cell_tab: level << 24 | cell row << 12 | cell column (built with the plan)
  const uint32_t ct = __ldg(cell_tab + cell);
  const int level = (int)(ct >> 24), ci = (int)((ct >> 12) & 0xfff), cj = (int)(ct & 0xfff);
  const DevLevel& L = P.lv[level];
  cell -= L.cell_base;
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
  const bool aligned = (pitch & 3) == 0;
  if (aligned) {
    sh = (int)((uintptr_t)tile0 & 3);
    const int nw = (sh + zw + 6 + 3) >> 2;  // data words per row
    const uint32_t* a0 = reinterpret_cast<const uint32_t*>(tile0 - sh);
    const int wpitch = pitch >> 2;
    if (nw <= 16) {
      // half a warp per row: lane = (row parity, word)
      const int col = lane & 15;
      if (col < nw) {
        const uint32_t* g = a0 + (lane >> 4) * wpitch + col;
        int o = (lane >> 4) * pw + 1 + col;
#pragma unroll 4
        for (int r = lane >> 4; r < zh + 6; r += 2, g += 2 * wpitch, o += 2 * pw) {
          const uint32_t v = __ldg(g);
          c.px[o] = v;
          c.hm[o] = (v >> 1) & 0x7f7f7f7fu;
        }
      }
    } else {
      const uint32_t rcw = ((1u << 20) + nw - 1) / nw;
      const int nwords = nw * (zh + 6);
#pragma unroll 4
      for (int i = lane; i < nwords; i += 32) {
        const int r = (int)(((uint32_t)i * rcw) >> 20), col = i - r * nw;
        const uint32_t v = __ldg(a0 + r * wpitch + col);
        const int o = r * pw + 1 + col;
        c.px[o] = v;
        c.hm[o] = (v >> 1) & 0x7f7f7f7fu;
      }
    }
  } else {
    sh = 0;
    uint8_t* pb = reinterpret_cast<uint8_t*>(c.px);
    for (int i = lane; i < (zw + 6) * (zh + 6); i += 32) {
      const int r = i / (zw + 6), col = i - r * (zw + 6);
      pb[r * (pw * 4) + 4 + col] = tile0[(long long)r * pitch + col];
    }
  }
  // Quads: word q of tile row r holds smem columns 4q..4q+3 = zone x 4q-(7+sh)+j, j=0..3.
  c.colbase = 7 + sh;  // smem column of zone x = 0
  c.q_first = c.colbase >> 2;
  c.nqr = ((c.colbase + zw - 1) >> 2) - c.q_first + 1;  // quads per zone row
  c.zw = zw;
  c.zh = zh;
  c.rcq = ((1u << 20) + c.nqr - 1) / c.nqr;
  const int nqr = c.nqr, nquads = nqr * zh;
  uint32_t* out = cand + (long long)frame * P.cand_total + L.cand_base + (long long)cell * L.cell_cap;
  const int kx0 = zx0 - c.colbase - ORBX_MIN_BORDER, ky0 = zy0 - ORBX_MIN_BORDER;

  int th = P.ini_th, total = 0;
  bool restage = !aligned;  // hm does not hold the 7-bit tile yet
#pragma unroll 1
  for (int attempt = 0; attempt < 2; attempt++) {
    __syncwarp();
    if (restage) {
      for (int i = lane; i < pw * (zh + 6); i += 32) c.hm[i] = (c.px[i] >> 1) & 0x7f7f7f7fu;
      __syncwarp();
    }
    if (lane == 0) reinterpret_cast<uint32_t*>(c.qm)[(nquads - 1) >> 2] = 0;  // tail bytes of the last word
    __syncwarp();

    // 2. quick test: lane = (segment of rows, quad column)
    const uint32_t cadd = (uint32_t)(128 - ((th + 1) >> 1)) * 0x01010101u;
#pragma unroll 1
    for (int cg = 0; cg < nqr; cg += 32) {
      const int ncols = min(32, nqr - cg);
      const int nseg = ncols <= 10 ? 3 : (ncols <= 16 ? 2 : 1);
      const int rps = (zh + nseg - 1) / nseg;
      int seg = lane >= 2 * ncols ? 2 : (lane >= ncols ? 1 : 0);
      const int qi = lane - seg * ncols;
      const bool active = seg < nseg && qi < ncols;
      if (!active) seg = 0;
      const int y0 = seg * rps;
      const int nrow = active ? min(zh, y0 + rps) - y0 : 0;  // rows this lane reports
      const uint32_t* hcol = c.hm + c.q_first + cg + (active ? qi : 0);
      const int hlast = (zh + 5) * pw;  // last staged row
      int hp = y0 * pw;
      uint8_t* qp = c.qm + y0 * nqr + cg + (active ? qi : 0);
      QuickWin<QP> W;
      quick_load<0, QP>(W, hcol + hp); hp = min(hp + pw, hlast);
      quick_load<1, QP>(W, hcol + hp); hp = min(hp + pw, hlast);
      quick_load<2, QP>(W, hcol + hp); hp = min(hp + pw, hlast);
      quick_load<3, QP>(W, hcol + hp); hp = min(hp + pw, hlast);
      quick_load<4, QP>(W, hcol + hp); hp = min(hp + pw, hlast);
      quick_load<5, QP>(W, hcol + hp); hp = min(hp + pw, hlast);
#define QSTEP(PH)                                       \
  {                                                     \
    quick_load<(PH + 6) % 7, QP>(W, hcol + hp);                  \
    hp = min(hp + pw, hlast);                           \
    const uint32_t m = quick_test<PH, QP>(W, cadd);         \
    if (base + PH < nrow) qp[PH * nqr] = (uint8_t)m;    \
  }
#pragma unroll 1
      for (int base = 0; base < rps; base += 7) {
        QSTEP(0) QSTEP(1) QSTEP(2) QSTEP(3) QSTEP(4) QSTEP(5) QSTEP(6)
        qp += 7 * nqr;
      }
#undef QSTEP
    }
    __syncwarp();
    // hm becomes the A map: clear the rows NMS can read (tile rows 2 .. zh+3)
    for (int i = 2 * pw + lane; i < (zh + 4) * pw; i += 32) c.hm[i] = 0;
    __syncwarp();

    // 3.-5. list the pixel pairs that passed, score them, then NMS + ordered emission
    int nlist, bands;
    fast_sweep<0>(c, nquads, th, lane, &nlist, &bands, 0, out, kx0, ky0);
    if (bands == 0) total = fast_nms_emit(c, nlist, 0, out, kx0, ky0, lane);
    else total = fast_sweep<1>(c, nquads, th, lane, &nlist, &bands, 0, out, kx0, ky0);
    if (total > 0 || th == P.min_th) break;
    // 6. nothing at iniThFAST: the reference runs FAST again on the cell with minThFAST
    th = P.min_th;
    restage = true;
  }
  if (lane == 0) *cnt_out = total;
}

cudaError_t orbx_launch_fast_cell(const DevPlan& P, Level0 l0, const uint8_t* arena, const uint32_t* cell_tab,
                             int32_t* cell_cnt, uint32_t* cand, int batch, cudaStream_t st) {
  // shared-memory geometry from the largest cell of the plan
  int pw = 0, rows = 0, cap = 0;
  for (int l = 0; l < P.nlevels; l++) {
    const int nw = (3 + P.lv[l].wCell + 6 + 3) >> 2;
    pw = max(pw, nw + 2);
    rows = max(rows, P.lv[l].hCell + 6);
    cap = max(cap, ((P.lv[l].wCell + 3 + 3) / 4 + 2) * P.lv[l].hCell);
  }
  cap = (cap + 8 + 15) & ~15;  // quads per cell (bytes of qm) + one spare uint2, regions stay 8-byte aligned
  const size_t smem = (size_t)FAST_WARPS * (2 * pw * rows + FAST_LIST_CAP / 2 + (cap >> 2)) * sizeof(uint32_t);
  static const int qp = [] { const char* e = getenv("ORBX_FAST_PAIRS"); return e && atoi(e) == 4 ? 4 : FAST_QUICK_PAIRS; }();
  const void* fn = qp == 4 ? (const void*)k_fast<4> : (const void*)k_fast<2>;
  cudaError_t e = orbx_ensure_dyn_smem(qp == 4 ? ORBX_SMEM_FAST4 : ORBX_SMEM_FAST, fn, smem);
  if (e != cudaSuccess) return e;
  dim3 grd((P.total_cells + FAST_WARPS - 1) / FAST_WARPS, batch);
  if (qp == 4) k_fast<4><<<grd, FAST_THREADS, smem, st>>>(P, l0, arena, cell_tab, cell_cnt, cand, pw, rows, cap);
  else k_fast<2><<<grd, FAST_THREADS, smem, st>>>(P, l0, arena, cell_tab, cell_cnt, cand, pw, rows, cap);
  return cudaGetLastError();
}
