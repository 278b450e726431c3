// fast.cu — per-cell FAST-9/16 detection with non-max suppression and the per-cell threshold fallback
// (cell loop of ORBextractor::ComputeKeyPointsOctTree, src/ORBextractor.cc:781-872; cv::FAST TYPE_9_16 +
// cornerScore) for sm_100a: the strip kernel.
//
// The reference runs cv::FAST on each ~35-px cell ROI (6 px overlap).  cv::FAST tests ROI pixels
// [3,W-3)x[3,H-3), so the tested zones of the cells tile [19,w-19)x[19,h-19) exactly on a (wCell,hCell)
// lattice; the NMS of a cell only sees that cell's own zone (outside = 0), and a cell that is empty after NMS
// at iniThFAST is redone at minThFAST (:843-859).  The arc strength A of a pixel depends on neither the cell
// nor the threshold (corner(th) <=> A > th, cornerScore == A-1), so only NMS and emission are per cell.
//
// One WARP = one STRIP = strip_k consecutive cells of one cell row (<= 120 tested columns), no block-level
// synchronisation (a CTA is FS_WARPS independent warps; 1 by default so shared memory packs the SM tightly):
//   1. the strip's pixels (+3 px halo, 16-byte aligned columns) are fetched by the TMA engine: one 3-D
//      tensor-map box (144 columns x hCell+6 rows of one frame) per strip, completion on the warp's mbarrier,
//      while the A map is being cleared;
//   2. quick test, dense: lane = one 4-pixel word column (a "quad" per row), walking down the rows with the last
//      7 rows (7-bit pixels) in registers and the horizontal neighbours taken from the adjacent lanes (SHFL);
//      packed-byte arithmetic (VABSDIFF4): a 9/16 corner needs |v-ring| > t on at least one pixel of each of the
//      axial ring pairs (0,8) and (4,12).  Each row's passing quads are appended to a task list with one ballot,
//      so the list is in row-major order;
//   3. arc strength of the listed quads with packed 16-bit SIMD (fast_arc.cuh, two pixel pairs per quad) into the
//      A map (u8, 0 where A <= th or outside the zone); quads with a corner stay listed;
//   4. NMS of the kept quads against the A map (columns where a new cell starts masked out) and, in the same
//      sweep, ordered emission: a corner's slot in its cell is a running per-cell count + a warp prefix over the
//      batch of 32 quads + its rank inside the quad — the row-major order the reference's per-cell cv::FAST returns;
//   5. cells that stayed empty are redone from step 2 at minThFAST with the zone restricted to their columns.
// The kernel is bound by the integer ALU pipe (VABSDIFF4 / LOP3 / PRMT / VIMNMX3 issue at one warp instruction
// per two cycles per scheduler), not by memory: see DESIGN.md §4 for the instruction budget per pixel.
#include <cuda.h>
#include <stdlib.h>

#include "fast_arc.cuh"
#include "orbx_kernels.cuh"

#ifndef FS_WARPS
#define FS_WARPS 1
#endif
#define FS_THREADS (FS_WARPS * 32)
#define FS_TP 144            // tile row pitch in bytes: 9 vectors of 16 B cover lane words 0..31 at word offset <= 3
#define FS_TPW (FS_TP / 4)   // 36 words
#define FS_APW 32            // A-map row pitch in words: word 0 = pad, words 1..30 = strip columns 0..119, word 31 = pad
#define FS_MAXH 250          // tallest zone: a TMA box holds at most 256 rows

__device__ __forceinline__ uint32_t fs_halve(uint32_t v) { return (v >> 1) & 0x7f7f7f7fu; }

// Shared memory of one warp (32-bit words):
//   tile[FS_TPW * (hmax+6)]   staged pixels: tile row r = level row Y0-3+r, tile byte b = level column xt+b
//   amap[FS_APW * (hmax+2)]   arc strengths: row r+1 = zone row r, byte 4+c = strip column c (level column xs+c)
//   list[30 * (hmax+6)] u16   quad tasks, row-major: row << 5 | lane
//   misc[12]                  zbits[4] (zone columns), ledge[5] (columns where a cell starts / the zone ends)
//   rowbits[hmax]             per zone row: ballot of the quads that passed the quick test
struct FsStrip {
  uint32_t *tile, *amap, *misc, *rowbits;
  uint16_t* list;
  int wq0;  // tile word of lane 0
  int zh;
};

// Quick test over the whole strip: lane = one 4-pixel word column (a "quad" per row), walking down the rows with
// the last 7 rows (7-bit pixels) in registers.  A quad passes when one of its pixels has |v-ring| > t on at least
// one pixel of each axial ring pair.  Each row leaves ONE word (the ballot of its passing quads) in rowbits; the
// row-major task list is expanded from those words afterwards, 32 rows at a time (lane = row: a warp scan of the
// popcounts, then every lane writes its row's entries) — ~2 instructions per row instead of the ~9 a per-row
// compacting append costs every lane.  Returns the number of tasks.
#ifndef FS_QUICK_ROWBITS
#define FS_QUICK_ROWBITS 1
#endif
__device__ __forceinline__ int fs_quick(const FsStrip& s, int th, uint32_t zm, int lane) {
  const uint32_t cadd = (uint32_t)(128 - ((th + 1) >> 1)) * 0x01010101u;
  const int zh = s.zh, nblk = (zh + 6) / 7;
  const uint32_t* trow = s.tile + s.wq0 + lane;
  uint32_t C[7];
#pragma unroll
  for (int i = 0; i < 6; i++) C[i] = fs_halve(trow[i * FS_TPW]);
  trow += 6 * FS_TPW;
#define QT(x) (__vabsdiffu4(v, (x)) + cadd)
#define FS_QROW(ph)                                                                                                   \
  C[((ph) + 6) % 7] = fs_halve(trow[(ph) * FS_TPW]);                                                                   \
  const uint32_t v = C[((ph) + 3) % 7];                                                                                \
  const uint32_t lw = __shfl_up_sync(0xffffffffu, v, 1), rw = __shfl_down_sync(0xffffffffu, v, 1);                     \
  const uint32_t l3 = __byte_perm(lw, v, 0x4321), r3 = __byte_perm(v, rw, 0x6543);                                     \
  uint32_t g = (QT(C[((ph) + 6) % 7]) | QT(C[(ph) % 7])) & zm; /* ring 0 (0,+3) | ring 8 (0,-3); zm: bit 7 of zone bytes */ \
  g &= QT(r3) | QT(l3);                                         /* ring 4 (+3,0) | ring 12 (-3,0) */                    \
  const unsigned bal = __ballot_sync(0xffffffffu, g != 0);
#if FS_QUICK_ROWBITS
  uint32_t* rb = s.rowbits;
  int rows_left = zh;
#pragma unroll 1
  for (int blk = 0; blk < nblk; blk++, trow += 7 * FS_TPW, rows_left -= 7, rb += 7) {
    if (rows_left >= 7) {  // warp-uniform
#pragma unroll
      for (int ph = 0; ph < 7; ph++) {
        FS_QROW(ph)
        if (lane == 0) rb[ph] = bal;
      }
    } else {
#pragma unroll
      for (int ph = 0; ph < 7; ph++) {
        if (ph >= rows_left) break;  // the last block of rows may be short
        FS_QROW(ph)
        if (lane == 0) rb[ph] = bal;
      }
    }
  }
  __syncwarp();
  int base = 0;
#pragma unroll 1
  for (int r0 = 0; r0 < zh; r0 += 32) {
    const int r = r0 + lane;
    uint32_t bits = r < zh ? s.rowbits[r] : 0u;
    const int cnt = __popc(bits);
    int inc = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    uint16_t* lp = s.list + base + inc - cnt;
    const uint32_t code = (uint32_t)r << 5;
    while (bits) {
      const uint32_t b = (uint32_t)__ffs((int)bits) - 1u;
      bits &= bits - 1u;
      *lp++ = (uint16_t)(code | b);
    }
    base += __shfl_sync(0xffffffffu, inc, 31);
  }
  __syncwarp();
  return base;
#else
  const uint32_t ltm = (1u << lane) - 1u;
  int base = 0;
  uint32_t code = (uint32_t)lane;
  int rows_left = zh;
#pragma unroll 1
  for (int blk = 0; blk < nblk; blk++, trow += 7 * FS_TPW, rows_left -= 7) {
#pragma unroll
    for (int ph = 0; ph < 7; ph++) {
      if (ph >= rows_left) break;  // warp-uniform: the last block of rows may be short
      FS_QROW(ph)
      if (g != 0) s.list[base + __popc(bal & ltm)] = (uint16_t)(code + 32u * ph);
      base += __popc(bal);
    }
    code += 7u * 32u;
  }
  __syncwarp();
  return base;
#endif
#undef FS_QROW
#undef QT
}

// Arc strength of the n listed quads into the A map (0 where A <= th or outside the zone); the quads that
// hold a corner move to the front of the list (order kept); returns their number.
__device__ __forceinline__ int fs_score(const FsStrip& s, int n, int th, int lane) {
  const uint32_t th4 = (uint32_t)th * 0x01010101u;
  int nkeep = 0;
  for (int base = 0; base < n; base += 32) {
    const int e = base + lane;
    const bool act = e < n;
    const uint32_t tk = act ? (uint32_t)s.list[e] : 1u;  // idle lanes score the quad of lane 1, row 0 (discarded)
    const int r = (int)(tk >> 5), ql = (int)(tk & 31u);
    const uint32_t* tp = s.tile + r * FS_TPW + s.wq0 + ql - 1;
    uint32_t rw[7][3];
#pragma unroll
    for (int d = 0; d < 7; d++) {
      rw[d][0] = tp[d * FS_TPW];
      rw[d][1] = tp[d * FS_TPW + 1];
      rw[d][2] = tp[d * FS_TPW + 2];
    }
    const uint32_t a = arc_pair<0>(rw) | (arc_pair<2>(rw) << 16);
    const int c0 = 4 * (ql - 1);  // strip column of the quad's first pixel
    const uint32_t nib = (s.misc[c0 >> 5] >> (c0 & 31)) & 15u;
    const uint32_t zmask = ((nib * 0x00204081u) & 0x01010101u) * 0xffu;
    const uint32_t av = a & __vcmpgtu4(a, th4) & zmask;
    if (act) s.amap[(r + 1) * FS_APW + ql] = av;
    const bool sv = act && av != 0;
    const unsigned bal = __ballot_sync(0xffffffffu, sv);
    __syncwarp();  // every lane has read its entry of this batch before the compacted entries land
    if (sv) s.list[nkeep + __popc(bal & ((1u << lane) - 1u))] = (uint16_t)tk;
    nkeep += __popc(bal);
  }
  __syncwarp();
  return nkeep;
}

// NMS of a quad.  w[j][0..2] = A-map words ql-1, ql, ql+1 of the rows above / at / below; e5 bit i = a cell
// starts at the quad's column i (i = 4: the next quad's first column): the column to its left belongs to
// another cell, whose corners the reference's per-cell cv::FAST never sees.  cv::FAST keeps a corner when its
// score A-1 is strictly greater than all 8 neighbouring scores (0 for non-corners): keep <=> A > max(1,
// neighbours' A).  Returns one bit per pixel.
__device__ __forceinline__ uint32_t fs_nms_quad(const uint32_t (&w)[3][3], uint32_t e5) {
  uint32_t p02[3], p13[3], p24[3], p35[3];
#pragma unroll
  for (int j = 0; j < 3; j++) {
    const uint32_t X = __funnelshift_r(w[j][0], w[j][1], 24);  // columns -1..2 of the quad
    const uint32_t Y = __funnelshift_r(w[j][1], w[j][2], 8);   // columns 1..4
    p02[j] = X & 0x00ff00ffu; p13[j] = (X >> 8) & 0x00ff00ffu;
    p24[j] = Y & 0x00ff00ffu; p35[j] = (Y >> 8) & 0x00ff00ffu;
  }
  // window position p = column p-1; pixel i sits at position i+1, its neighbours at i and i+2
  uint32_t m02 = __vimax3_u16x2(p02[0], p02[1], p02[2]), m13 = __vimax3_u16x2(p13[0], p13[1], p13[2]);
  uint32_t m24 = __vimax3_u16x2(p24[0], p24[1], p24[2]), m35 = __vimax3_u16x2(p35[0], p35[1], p35[2]);
  const uint32_t c13 = __vmaxu2(p13[0], p13[2]), c24 = __vmaxu2(p24[0], p24[2]);
  uint32_t r24 = m24, l13 = m13;
  if (e5) {
    // pixel i: left column out when a cell starts at i, right column out when one starts at i+1
    const uint32_t k0 = (e5 & 1u) ? 0u : 0xffffu, k1 = (e5 & 2u) ? 0u : 0xffffu, k2 = (e5 & 4u) ? 0u : 0xffffu,
                   k3 = (e5 & 8u) ? 0u : 0xffffu, k4 = (e5 & 16u) ? 0u : 0xffffu;
    m02 &= k0 | (k2 << 16);  // left of pixels 0,2
    r24 &= k1 | (k3 << 16);  // right of pixels 0,2
    l13 &= k1 | (k3 << 16);  // left of pixels 1,3
    m35 &= k2 | (k4 << 16);  // right of pixels 1,3
  }
  const uint32_t n02 = __vmaxu2(__vimax3_u16x2(m02, r24, c13), 0x00010001u);
  const uint32_t n13 = __vmaxu2(__vimax3_u16x2(l13, m35, c24), 0x00010001u);
  // own > n  <=>  bit 8 of (256 + n - own) clear, per 16-bit lane (no borrow crosses lanes)
  const uint32_t k02 = ~((n02 | 0x01000100u) - p13[1]) & 0x01000100u;
  const uint32_t k13 = ~((n13 | 0x01000100u) - p24[1]) & 0x01000100u;
  return ((k02 >> 8) & 1u) | ((k13 >> 7) & 2u) | ((k02 >> 22) & 4u) | ((k13 >> 21) & 8u);
}

__device__ __forceinline__ uint32_t fs_below(int c, int w) {  // bits of bitmap word w that lie below column c
  const int n = c - 32 * w;
  return n <= 0 ? 0u : (n >= 32 ? 0xffffffffu : ((1u << n) - 1u));
}

__global__ void __launch_bounds__(FS_THREADS, 14 / FS_WARPS)
k_fast_strip(const __grid_constant__ DevPlan P, const __grid_constant__ OrbxFastMaps M, Level0 l0,
             const uint8_t* __restrict__ arena, const uint32_t* __restrict__ strip_tab, int32_t* __restrict__ cell_cnt,
             uint32_t* __restrict__ cand, int strip0, int strip1, int hmax, int region_words, int f0) {
  extern __shared__ __align__(128) uint32_t fs_smem[];
  const int frame = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int strip = strip0 + blockIdx.x * FS_WARPS + warp;
  if (strip >= strip1) return;
  FsStrip s;
  s.tile = fs_smem + warp * region_words;
  s.amap = s.tile + FS_TPW * (hmax + 6);
  s.misc = s.amap + FS_APW * (hmax + 2);
  s.rowbits = s.misc + 12;                              // misc[10..11] = the warp's mbarrier (8-byte aligned)
  s.list = reinterpret_cast<uint16_t*>(s.rowbits + hmax);

  // strip_tab: level << 24 | cell row << 12 | first cell column (built with the plan)
  const uint32_t stt = __ldg(strip_tab + strip);
  const int level = (int)(stt >> 24), ci = (int)((stt >> 12) & 0xfff), cj0 = (int)(stt & 0xfff);
  const DevLevel& L = P.lv[level];
  const int k = min(L.strip_k, L.nCols - cj0);
  const int Y0 = ORBX_EDGE + ci * L.hCell, Y1 = min(Y0 + L.hCell, L.h - ORBX_EDGE), zh = Y1 - Y0;
  const int X0 = ORBX_EDGE + cj0 * L.wCell, X1 = min(X0 + k * L.wCell, L.w - ORBX_EDGE);
  const int cell0 = L.cell_base + ci * L.nCols + cj0;
  int32_t* cnt_out = cell_cnt + (long long)frame * P.total_cells + cell0;
  if (zh <= 0 || X1 <= X0) {  // cells skipped / empty in the reference (:810,:819)
    if (lane < k) cnt_out[lane] = 0;
    return;
  }
  s.zh = zh;
  int pitch;
  const uint8_t* img;
  if (level == 0) { pitch = l0.pitch; img = l0.base + (long long)frame * l0.frame_stride; }
  else { pitch = L.pitch; img = arena + (long long)frame * P.arena_bytes + L.offset; }

  // 1. stage level rows [Y0-3, Y1+3), columns [xt, xt+144)
  const int xs = X0 & ~3;           // level column of lane 1's first pixel (strip column 0)
  const int xt = (xs - 4) & ~15;    // level column of tile byte 0
  s.wq0 = (xs - 4 - xt) >> 2;
  const int trows = zh + 6;
  const uint8_t* src0 = img + (long long)(Y0 - 3) * pitch + xt;
  // One 2-D box of the level's tensor map (144 columns x hCell+6 rows of frame f0+frame), fetched by the TMA
  // engine on behalf of the warp; columns / rows outside the level read as zeros (never used).  Levels whose
  // base or pitch is not 16-byte aligned (a caller's odd-width level 0) are staged with plain loads.
  const bool tma = (M.mask >> level) & 1u;
  const uint32_t bar_a = (uint32_t)__cvta_generic_to_shared(s.misc + 10);
  if (tma) {
    if (lane == 0) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_a));
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_a), "r"(FS_TP * (L.hCell + 6)) : "memory");
      asm volatile(
          "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
          ::"r"((uint32_t)__cvta_generic_to_shared(s.tile)), "l"(reinterpret_cast<const void*>(&M.m[level])), "r"(xt),
          "r"(Y0 - 3), "r"(f0 + frame), "r"(bar_a)
          : "memory");
    }
  } else {
    for (int i = lane; i < trows * FS_TPW; i += 32) {
      const int r = i / FS_TPW, wq = i - r * FS_TPW, col = xt + 4 * wq;
      const uint8_t* g = src0 + (long long)r * pitch + 4 * wq;
      uint32_t v = 0;
#pragma unroll
      for (int j = 0; j < 4; j++)
        if (col + j < L.w) v |= (uint32_t)g[j] << (8 * j);
      s.tile[i] = v;
    }
  }
  // clear the A map (rows 0..zh+1) while the copies are in flight
  {
    uint4* z = reinterpret_cast<uint4*>(s.amap);
    const int nz = (FS_APW * (zh + 2)) >> 2;
    for (int i = lane; i < nz; i += 32) z[i] = make_uint4(0, 0, 0, 0);
  }
  // zone columns / cell starts as bit masks over the strip's columns
  const int cx0 = X0 - xs, cx1 = X1 - xs;  // strip columns of the zone, cx1 <= 120
  const int wc = L.wCell;
  if (lane < 4) {
    s.misc[lane] = fs_below(cx1, lane) & ~fs_below(cx0, lane);
  } else if (lane < 9) {
    const int w = lane - 4;
    uint32_t m = 0;
    for (int j = 0; j <= k; j++) {
      const int c = j < k ? min(cx0 + j * wc, cx1) : cx1;
      if ((c >> 5) == w) m |= 1u << (c & 31);
    }
    s.misc[4 + w] = m;
  }
  // cells: [cs1, cs2) is cell 1, [cs2, ..) cell 2; cells that start past the zone are empty in the reference (:819)
  const int cs1 = k > 1 ? cx0 + wc : 128, cs2 = k > 2 ? cx0 + 2 * wc : 128;
  uint32_t valid = 1u;
  if (k > 1 && cs1 < cx1) valid |= 2u;
  if (k > 2 && cs2 < cx1) valid |= 4u;
  if (tma) {
    uint32_t done = 0;
    while (!done)
      asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }"
                   : "=r"(done) : "r"(bar_a) : "memory");
  }
  __syncwarp();

  uint32_t* out0 = cand + (long long)frame * P.cand_total + L.cand_base + (long long)(ci * L.nCols + cj0) * L.cell_cap;
  int th = P.ini_th;
  uint32_t todo = (1u << k) - 1u;  // cells this attempt covers
#pragma unroll 1
  for (int attempt = 0; attempt < 2; attempt++) {
    // 2./3. quick test -> task list
    uint32_t zm = 0;
    if (lane >= 1 && lane <= 30) {
      const int c = 4 * (lane - 1);
      const uint32_t nib = (s.misc[c >> 5] >> (c & 31)) & 15u;
      zm = ((nib * 0x00204081u) & 0x01010101u) << 7;
    }
    const int ntask = fs_quick(s, th, zm, lane);
    // 4. arc strength; the quads that hold a corner stay listed (row-major)
    const int nk = fs_score(s, ntask, th, lane);
    // 5./6. NMS and ordered emission in one sweep over the kept quads, 32 at a time: the position of a corner in
    // its cell's slot array is the number of corners the cell kept before it in row-major order, i.e. a running
    // per-cell count (11 | 11 | 10 bits) + a warp prefix over the batch + its rank inside the quad
    uint32_t run = 0;
    const uint32_t* ledge = s.misc + 4;
    const uint8_t* ab = reinterpret_cast<const uint8_t*>(s.amap);
    for (int base = 0; base < nk; base += 32) {
      const int e = base + lane;
      uint32_t f = 0;
      int r = 0, c0 = 0;
      if (e < nk) {
        const uint32_t tk = s.list[e];
        r = (int)(tk >> 5);
        const int ql = (int)(tk & 31u);
        c0 = 4 * (ql - 1);
        const uint32_t* wp = s.amap + r * FS_APW + ql - 1;  // rows r, r+1, r+2 = zone rows r-1, r, r+1
        uint32_t w[3][3];
#pragma unroll
        for (int j = 0; j < 3; j++) {
          w[j][0] = wp[j * FS_APW]; w[j][1] = wp[j * FS_APW + 1]; w[j][2] = wp[j * FS_APW + 2];
        }
        const uint32_t e5 = __funnelshift_r(ledge[c0 >> 5], ledge[(c0 >> 5) + 1], c0 & 31) & 31u;
        f = fs_nms_quad(w, e5);
      }
      // the quad's pixels by cell: columns below cs1 are cell 0, below cs2 cell 1, the rest cell 2
      const uint32_t m1 = (1u << min(max(cs1 - c0, 0), 4)) - 1u, m2 = (1u << min(max(cs2 - c0, 0), 4)) - 1u;
      const uint32_t f0 = f & m1, f1 = f & m2 & ~m1, f2 = f & ~m2;
      const uint32_t pk = (uint32_t)__popc(f0) | ((uint32_t)__popc(f1) << 11) | ((uint32_t)__popc(f2) << 22);
      uint32_t inc = pk;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
      }
      const uint32_t exc = run + inc - pk;
      run += __shfl_sync(0xffffffffu, inc, 31);
      while (f) {
        const int i = __ffs(f) - 1;
        const uint32_t below = (1u << i) - 1u;
        f &= f - 1;
        const int j = (int)(((m1 >> i) & 1u) ^ 1u) + (int)(((m2 >> i) & 1u) ^ 1u);
        if (!((todo >> j) & 1u)) continue;
        const uint32_t fj = j == 0 ? f0 : (j == 1 ? f1 : f2);
        const int pos = (int)((exc >> (11 * j)) & (j == 2 ? 0x3ffu : 0x7ffu)) + __popc(fj & below);
        const int c = c0 + i;
        const int a = ab[(r + 1) * (FS_APW * 4) + 4 + c];
        // coordinates relative to (minBorderX,minBorderY)=(16,16), as DistributeOctTree sees them
        out0[(long long)j * L.cell_cap + pos] = orbx_pack_key(xs + c - ORBX_MIN_BORDER, Y0 + r - ORBX_MIN_BORDER, a - 1);
      }
    }
    uint32_t redo = 0;
#pragma unroll
    for (int j = 0; j < 3; j++) {
      const int tj = (int)((run >> (11 * j)) & (j == 2 ? 0x3ffu : 0x7ffu));
      if ((todo >> j) & 1u) {
        if (lane == 0) cnt_out[j] = tj;
        if (tj == 0) redo |= 1u << j;
      }
    }
    redo &= valid;
    // 7. nothing at iniThFAST: the reference runs FAST again on the cell with minThFAST
    if (redo == 0 || th == P.min_th) break;
    th = P.min_th;
    todo = redo;
    __syncwarp();
    if (lane < 4) {
      uint32_t m = 0;
      for (int j = 0; j < k; j++) {
        if (!((redo >> j) & 1u)) continue;
        const int cs = cx0 + j * wc, ce = min(cs + wc, cx1);
        m |= fs_below(ce, lane) & ~fs_below(cs, lane);
      }
      s.misc[lane] = m;
    }
    __syncwarp();
  }
}

static size_t fs_region_words(int hmax) {
  const size_t w = (size_t)FS_TPW * (hmax + 6) + (size_t)FS_APW * (hmax + 2) + 12 + hmax + (30 * (size_t)(hmax + 6) + 1) / 2;
  return (w + 31) & ~(size_t)31;  // the TMA destination (the tile, first in the region) is 128-byte aligned
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link against libcuda)
typedef CUresult (*FsEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                               const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static FsEncodeFn fs_encode_fn() {
  static FsEncodeFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return (FsEncodeFn)p;
  }();
  return fn;
}

// Tensor map of one level: bytes [w] x rows [h] x frames [n], box = box_cols columns x box_rows rows x 1 frame.
bool orbx_encode_tmap3(void* m, const uint8_t* base, int w, int h, long long nframes, long long pitch,
                       long long frame_stride, int box_cols, int box_rows) {
  FsEncodeFn enc = fs_encode_fn();
  if (!enc || box_rows > 256 || box_rows < 1 || box_cols > 256 || (box_cols & 15)) return false;
  if ((reinterpret_cast<uintptr_t>(base) & 15u) || (pitch & 15) || (frame_stride & 15) || nframes < 1) return false;
  const cuuint64_t dims[3] = {(cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)nframes};
  const cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)frame_stride};
  const cuuint32_t box[3] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  return enc(reinterpret_cast<CUtensorMap*>(m), CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<uint8_t*>(base), dims, strides,
             box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
static bool fs_encode_level(CUtensorMap* m, const uint8_t* base, int w, int h, long long nframes, long long pitch,
                            long long frame_stride, int box_rows) {
  return orbx_encode_tmap3(m, base, w, h, nframes, pitch, frame_stride, FS_TP, box_rows);
}

void orbx_fast_maps_init(const DevPlan& P, const uint8_t* arena, int max_batch, OrbxFastMaps* M) {
  memset(M, 0, sizeof(*M));
  for (int l = 1; l < P.nlevels; l++) {
    const DevLevel& L = P.lv[l];
    if (fs_encode_level(reinterpret_cast<CUtensorMap*>(&M->m[l]), arena + L.offset, L.w, L.h, max_batch, L.pitch, P.arena_bytes, L.hCell + 6)) M->mask |= 1u << l;
  }
}

cudaError_t orbx_launch_fast_strip(const DevPlan& P, const OrbxFastMaps& maps, Level0 l0, const uint8_t* arena,
                                   const uint32_t* strip_tab, int32_t* cell_cnt, uint32_t* cand, int batch, int f0,
                                   cudaStream_t st) {
  // level 0 is the caller's buffer: its map is encoded per call (frame 0 of the map = l0.base)
  OrbxFastMaps M = maps;
  M.mask &= ~1u;
  static const bool no_tma = [] { const char* e = getenv("ORBX_FAST_TMA"); return e && e[0] == '0'; }();
  if (no_tma) M.mask = 0;
  else if (fs_encode_level(reinterpret_cast<CUtensorMap*>(&M.m[0]), l0.base - (long long)f0 * l0.frame_stride, P.lv[0].w, P.lv[0].h, (long long)f0 + batch, l0.pitch,
                           l0.frame_stride ? l0.frame_stride : 16, P.lv[0].hCell + 6))
    M.mask |= 1u;
  // two launches: the strips of the levels with ordinary cell heights, then those of the (small) levels whose
  // few cell rows are tall — shared memory per warp follows the tallest zone of the launch
  for (int g = 0; g < 2; g++) {
    const int s0 = g == 0 ? 0 : P.strips_a, s1 = g == 0 ? P.strips_a : P.total_strips;
    const int hmax = g == 0 ? P.hmax_a : P.hmax_b;
    if (s1 <= s0) continue;
    if (hmax > FS_MAXH) return cudaErrorInvalidValue;
    // (occupancy is not the lever here: padding the region by 2 / 4 / 8 KB, i.e. 13 / 11 / 9 instead of 15 strips per
    //  SM, costs 3.6 / 12.8 / 26.5 % — the ALU pipe is 65 % busy at 15 strips, so the 2.6 KB task list could buy ~3 %)
    const size_t region = fs_region_words(hmax);
    const size_t smem = FS_WARPS * region * sizeof(uint32_t);
    cudaError_t e = orbx_ensure_dyn_smem(ORBX_SMEM_FAST_STRIP, (const void*)k_fast_strip, smem);
    if (e != cudaSuccess) return e;
    dim3 grd((s1 - s0 + FS_WARPS - 1) / FS_WARPS, batch);
    k_fast_strip<<<grd, FS_THREADS, smem, st>>>(P, M, l0, arena, strip_tab, cell_cnt, cand, s0, s1, hmax, (int)region, f0);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}

cudaError_t orbx_launch_fast(const DevPlan& P, const OrbxFastMaps& maps, Level0 l0, const uint8_t* arena,
                             const uint32_t* cell_tab, const uint32_t* strip_tab, int32_t* cell_cnt, uint32_t* cand,
                             int batch, int f0, cudaStream_t st) {
  // ORBX_FAST_IMPL=cell selects the first-generation cell-per-warp kernel (A/B measurements)
  static const bool cell = [] { const char* e = getenv("ORBX_FAST_IMPL"); return e && e[0] == 'c'; }();
  if (cell) return orbx_launch_fast_cell(P, l0, arena, cell_tab, cell_cnt, cand, batch, st);
  return orbx_launch_fast_strip(P, maps, l0, arena, strip_tab, cell_cnt, cand, batch, f0, st);
}
