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
// synchronisation (a CTA is FS_WARPS independent warps):
//   1. the strip's pixels (+3 px halo, 16-byte aligned columns) go to shared memory with cp.async (16 B per
//      lane, three rows per step) while the A map and the keep bitmap are being cleared;
//   2. quick test, dense: lane = one 4-pixel word column, walking down the rows with the last 7 rows (7-bit
//      pixels) in registers and the horizontal neighbours taken from the adjacent lanes (SHFL); packed-byte
//      arithmetic (VABSDIFF4): a 9/16 corner needs |v-ring| > t on at least one pixel of each of the axial
//      ring pairs (0,8) and (4,12).  Two pass bits (one per pixel pair) per row accumulate in a register;
//   3. the pass bits are compacted into a task list (pixel pairs, any order);
//   4. arc strength of the listed pairs with packed 16-bit SIMD (fast_arc.cuh) into the A map (u8, 0 where
//      A <= th or outside the zone); pairs with a corner are kept in the list;
//   5. NMS of the kept pairs against the A map, with the columns where a new cell starts masked out, sets one
//      bit per surviving corner in the keep bitmap;
//   6. per cell: lanes own rows, count their bits, one warp scan, ordered (row-major) emission into the cell's
//      slot array — the order the reference's per-cell cv::FAST returns;
//   7. cells that stayed empty are redone from step 2 at minThFAST with the zone restricted to their columns.
#include <stdlib.h>

#include "fast_arc.cuh"
#include "orbx_kernels.cuh"

#define FS_WARPS 2
#define FS_THREADS (FS_WARPS * 32)
#define FS_TP 144            // tile row pitch in bytes: 9 vectors of 16 B cover lane words 0..31 at word offset <= 3
#define FS_TPW (FS_TP / 4)   // 36 words
#define FS_AP 132            // A-map row pitch in bytes: 4 (pad) + 120 (columns) + 8 (pad), odd word count
#define FS_APW (FS_AP / 4)   // 33 words
#define FS_KBP 5             // keep-bitmap words per row (4 used: 120 columns)
#define FS_CAP 512           // pair tasks per chunk of the list
#define FS_MAXB2 5           // pass-bit words per lane: 14 rows each (zh <= 69)

__device__ __forceinline__ uint32_t fs_halve(uint32_t v) { return (v >> 1) & 0x7f7f7f7fu; }

__device__ __forceinline__ void fs_cp_async16(uint32_t saddr, const void* g, uint32_t nbytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(saddr), "l"(g), "r"(nbytes) : "memory");
}

// NMS of the pixel pair in bytes 1,2 of w1.  w0,w1,w2 = rows above / at / below with the A values of columns
// c-1..c+2 (bytes 0..3); e3 bit i = a cell starts at column c+i (its left neighbour belongs to another cell,
// whose corners the reference's per-cell cv::FAST never sees).  cv::FAST keeps a corner when its score A-1 is
// strictly greater than all 8 neighbouring scores (0 for non-corners): keep <=> A > max(1, neighbours' A).
__device__ __forceinline__ uint32_t fs_nms_pair(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t e3) {
  uint32_t mA = 0x00ffffffu, mB = 0xffffff00u;
  if (e3 & 1u) mA &= 0xffffff00u;
  if (e3 & 2u) { mA &= 0xff00ffffu; mB &= 0xffff00ffu; }
  if (e3 & 4u) mB &= 0x00ffffffu;
  uint32_t keep = 0;
  {
    const uint32_t x0 = w0 & mA, x1 = w1 & mA & 0xffff00ffu, x2 = w2 & mA;
    const uint32_t ve = __vimax3_u16x2(x0 & 0x00ff00ffu, x1 & 0x00ff00ffu, x2 & 0x00ff00ffu);
    const uint32_t vo = __vimax3_u16x2((x0 >> 8) & 0x00ff00ffu, (x1 >> 8) & 0x00ff00ffu, (x2 >> 8) & 0x00ff00ffu);
    const uint32_t t = __vmaxu2(ve, vo);
    const uint32_t m = max(max(t & 0xffffu, t >> 16), 1u);
    if (((w1 >> 8) & 0xffu) > m) keep |= 1u;
  }
  {
    const uint32_t x0 = w0 & mB, x1 = w1 & mB & 0xff00ffffu, x2 = w2 & mB;
    const uint32_t ve = __vimax3_u16x2(x0 & 0x00ff00ffu, x1 & 0x00ff00ffu, x2 & 0x00ff00ffu);
    const uint32_t vo = __vimax3_u16x2((x0 >> 8) & 0x00ff00ffu, (x1 >> 8) & 0x00ff00ffu, (x2 >> 8) & 0x00ff00ffu);
    const uint32_t t = __vmaxu2(ve, vo);
    const uint32_t m = max(max(t & 0xffffu, t >> 16), 1u);
    if (((w1 >> 16) & 0xffu) > m) keep |= 2u;
  }
  return keep;
}

// Shared memory of one warp (32-bit words):
//   tile[FS_TPW * (hmax+6)]   staged pixels: tile row r = level row Y0-3+r, tile byte b = level column xt+b
//   amap[FS_APW * (hmax+2)]   arc strengths: row r+1 = zone row r, byte 4+c = strip column c (level column xs+c)
//   kb[FS_KBP * hmax]         keep bitmap: bit c of row r
//   pb[32 * FS_MAXB2]         pass bits: word b of lane l = rows 14b..14b+13 (two 7-row halves, see below)
//   list[FS_CAP] u16          tasks: row << 6 | lane << 1 | pair
//   misc[12]                  zbits[4] (zone columns), ledge[5] (columns where a cell starts / the zone ends)
struct FsStrip {
  uint32_t *tile, *amap, *kb, *pb, *misc;
  uint16_t* list;
  int wq0;  // tile word of lane 0
  int zh;
};

// Task list of chunk [cb, cb+FS_CAP) of the pass bits pw[] (lane-private), off = this lane's first index.
__device__ __forceinline__ void fs_fill(const FsStrip& s, const uint32_t (&pw)[FS_MAXB2], int off, int cb, int lane) {
  int pos = off - cb;
#pragma unroll
  for (int b = 0; b < FS_MAXB2; b++) {
    uint32_t w = pw[b];
    while (w) {
      const int bit = __ffs(w) - 1;
      w &= w - 1;
      // low half: rows 0..6 at bits 2+2j, high half: rows 7..13 at bits 18+2(j-7); even bit = pair 0
      const int row = 14 * b + 7 * (bit >> 4) + (((bit & 15) - 2) >> 1);
      if ((unsigned)pos < (unsigned)FS_CAP) s.list[pos] = (uint16_t)((row << 6) | (lane << 1) | (bit & 1));
      pos++;
    }
  }
  __syncwarp();
}

// Arc strength of the n listed pairs into the A map; with COMPACT the pairs that hold a corner move to the
// front of the list (returns their number), otherwise returns n.
template <bool COMPACT>
__device__ __forceinline__ int fs_score(const FsStrip& s, int n, int th, int lane) {
  int nkeep = 0;
  for (int base = 0; base < n; base += 32) {
    const int e = base + lane;
    const bool act = e < n;
    const uint32_t tk = act ? (uint32_t)s.list[e] : 2u;  // idle lanes score the pair of lane 1, row 0 (discarded)
    const int r = (int)(tk >> 6), ql = (int)((tk >> 1) & 31u), p = (int)(tk & 1u);
    const uint32_t* tp = s.tile + r * FS_TPW + s.wq0 + ql - 1;
    const int fs = 16 * p;
    uint32_t rw[7][3];
#pragma unroll
    for (int d = 0; d < 7; d++) {
      const uint32_t w0 = tp[d * FS_TPW], w1 = tp[d * FS_TPW + 1], w2 = tp[d * FS_TPW + 2];
      rw[d][0] = __funnelshift_r(w0, w1, fs);
      rw[d][1] = __funnelshift_r(w1, w2, fs);
      rw[d][2] = w2 >> fs;
    }
    const uint32_t a = arc_pair<0>(rw);
    const int c0 = 4 * (ql - 1) + 2 * p;  // strip column of the left pixel
    const uint32_t zz = (s.misc[c0 >> 5] >> (c0 & 31)) & 3u;
    uint32_t keep = 0;
    if ((int)(a & 0xffu) > th && (zz & 1u)) keep |= 0x00ffu;
    if ((int)(a >> 8) > th && (zz & 2u)) keep |= 0xff00u;
    const uint32_t av = a & keep;
    if (act) *reinterpret_cast<uint16_t*>(reinterpret_cast<uint8_t*>(s.amap) + (r + 1) * FS_AP + 4 + c0) = (uint16_t)av;
    if (COMPACT) {
      const bool sv = act && av != 0;
      const unsigned bal = __ballot_sync(0xffffffffu, sv);
      if (sv) s.list[nkeep + __popc(bal & ((1u << lane) - 1u))] = (uint16_t)tk;
      nkeep += __popc(bal);
    }
  }
  __syncwarp();
  return COMPACT ? nkeep : n;
}

// NMS of the n listed pairs (pairs without a corner are skipped) into the keep bitmap.
__device__ __forceinline__ void fs_nms(const FsStrip& s, int n, int lane) {
  const uint32_t* ledge = s.misc + 4;
  for (int e = lane; e < n; e += 32) {
    const uint32_t tk = s.list[e];
    const int r = (int)(tk >> 6), ql = (int)((tk >> 1) & 31u), p = (int)(tk & 1u);
    const int c0 = 4 * (ql - 1) + 2 * p;
    const uint32_t a = *reinterpret_cast<const uint16_t*>(reinterpret_cast<const uint8_t*>(s.amap) + (r + 1) * FS_AP + 4 + c0);
    if (a == 0) continue;
    // columns c0-1..c0+2 = bytes 3+c0..6+c0 of the A-map row; rows r, r+1, r+2 = zone rows r-1, r, r+1
    const uint32_t* wp = s.amap + r * FS_APW + ((3 + c0) >> 2);
    const int fs = p ? 8 : 24;
    const uint32_t w0 = __funnelshift_r(wp[0], wp[1], fs);
    const uint32_t w1 = __funnelshift_r(wp[FS_APW], wp[FS_APW + 1], fs);
    const uint32_t w2 = __funnelshift_r(wp[2 * FS_APW], wp[2 * FS_APW + 1], fs);
    const uint32_t e3 = __funnelshift_r(ledge[c0 >> 5], ledge[(c0 >> 5) + 1], c0 & 31) & 7u;
    const uint32_t f = fs_nms_pair(w0, w1, w2, e3);
    if (f) atomicOr(s.kb + r * FS_KBP + (c0 >> 5), f << (c0 & 31));
  }
  __syncwarp();
}

__global__ void __launch_bounds__(FS_THREADS, 7)
k_fast_strip(const __grid_constant__ DevPlan P, Level0 l0, const uint8_t* __restrict__ arena,
             const uint32_t* __restrict__ strip_tab, int32_t* __restrict__ cell_cnt, uint32_t* __restrict__ cand,
             int hmax, int region_words) {
  extern __shared__ __align__(16) uint32_t fs_smem[];
  const int frame = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int strip = blockIdx.x * FS_WARPS + warp;
  if (strip >= P.total_strips) return;
  const int amap_words = (FS_APW * (hmax + 2) + 3) & ~3, kb_words = (FS_KBP * hmax + 3) & ~3;
  FsStrip s;
  s.tile = fs_smem + warp * region_words;
  s.amap = s.tile + FS_TPW * (hmax + 6);
  s.kb = s.amap + amap_words;
  s.pb = s.kb + kb_words;
  s.list = reinterpret_cast<uint16_t*>(s.pb + 32 * FS_MAXB2);
  s.misc = s.pb + 32 * FS_MAXB2 + FS_CAP / 2;

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
  const bool al16 = ((reinterpret_cast<uintptr_t>(img) | (uintptr_t)(unsigned)pitch) & 15u) == 0;
  if (al16) {
    const int v = lane % 9, ro = lane / 9;
    if (ro < 3) {
      const bool inb = xt + 16 * v + 16 <= pitch;  // columns past the row's allocation are never needed: zero-fill
      const uint8_t* g = src0 + (inb ? 16 * v : 0) + (long long)ro * pitch;
      uint32_t sa = (uint32_t)__cvta_generic_to_shared(s.tile) + ro * FS_TP + 16 * v;
      const uint32_t nb = inb ? 16u : 0u;
      for (int r = ro; r < trows; r += 3, g += 3LL * pitch, sa += 3 * FS_TP) fs_cp_async16(sa, g, nb);
    }
    asm volatile("cp.async.commit_group;\n" ::: "memory");
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
  // clear the A map (rows 0..zh+1) and the keep bitmap while the copies are in flight
  {
    uint4* z = reinterpret_cast<uint4*>(s.amap);
    const int nz = (FS_APW * (zh + 2) + 3) >> 2;
    for (int i = lane; i < nz; i += 32) z[i] = make_uint4(0, 0, 0, 0);
    uint4* zk = reinterpret_cast<uint4*>(s.kb);
    const int nk = (FS_KBP * zh + 3) >> 2;
    for (int i = lane; i < nk; i += 32) zk[i] = make_uint4(0, 0, 0, 0);
  }
  // zone columns / cell starts as bit masks over the strip's columns
  const int cx0 = X0 - xs, cx1 = X1 - xs;  // strip columns of the zone, cx1 <= 120
  if (lane < 4) {
    const int lo = max(cx0 - 32 * lane, 0), hi = min(cx1 - 32 * lane, 32);
    s.misc[lane] = hi > lo ? ((hi - lo == 32) ? 0xffffffffu : (((1u << (hi - lo)) - 1u) << lo)) : 0u;
  } else if (lane < 9) {
    const int w = lane - 4;
    uint32_t m = 0;
    for (int j = 0; j <= k; j++) {
      const int c = j < k ? cx0 + j * L.wCell : cx1;
      if ((c >> 5) == w) m |= 1u << (c & 31);
    }
    if ((cx1 >> 5) == w) m |= 1u << (cx1 & 31);
    s.misc[4 + w] = m;
  }
  if (al16) asm volatile("cp.async.wait_all;\n" ::: "memory");
  __syncwarp();

  uint32_t* out0 = cand + (long long)frame * P.cand_total + L.cand_base + (long long)(ci * L.nCols + cj0) * L.cell_cap;
  int th = P.ini_th;
  uint32_t todo = (1u << k) - 1u;  // cells this attempt covers
#pragma unroll 1
  for (int attempt = 0; attempt < 2; attempt++) {
    // 2. quick test
    uint32_t zm = 0;
    if (lane >= 1 && lane <= 30) {
      const int c = 4 * (lane - 1);
      const uint32_t nib = (s.misc[c >> 5] >> (c & 31)) & 15u;
      zm = ((nib * 0x00204081u) & 0x01010101u) << 7;
    }
    const uint32_t cadd = (uint32_t)(128 - ((th + 1) >> 1)) * 0x01010101u;
    const int nblk = (zh + 6) / 7;
    {
      const uint32_t* trow = s.tile + s.wq0 + lane;
      uint32_t C[7];
#pragma unroll
      for (int i = 0; i < 6; i++) C[i] = fs_halve(trow[i * FS_TPW]);
      trow += 6 * FS_TPW;
      uint16_t* pb16 = reinterpret_cast<uint16_t*>(s.pb + lane);
#pragma unroll 1
      for (int blk = 0; blk < nblk; blk++, trow += 7 * FS_TPW) {
        uint32_t acc = 0;
#pragma unroll
        for (int ph = 0; ph < 7; ph++) {
          C[(ph + 6) % 7] = fs_halve(trow[ph * FS_TPW]);
          const uint32_t v = C[(ph + 3) % 7];
          const uint32_t lw = __shfl_up_sync(0xffffffffu, v, 1), rw = __shfl_down_sync(0xffffffffu, v, 1);
          const uint32_t l3 = __byte_perm(lw, v, 0x4321), r3 = __byte_perm(v, rw, 0x6543);
#define QT(x) (__vabsdiffu4(v, (x)) + cadd)
          uint32_t g = (QT(C[(ph + 6) % 7]) | QT(C[ph % 7])) & zm;  // ring 0 (0,+3) | ring 8 (0,-3)
          g &= QT(r3) | QT(l3);                                     // ring 4 (+3,0) | ring 12 (-3,0)
#undef QT
          // bit 7 / 23 of t: a pixel of pair 0 / 1 passed -> bits 30 / 31
          const uint32_t t = (g | (g >> 8)) & 0x00800080u;
          acc = (acc >> 2) | ((t * 0x00800100u) & 0xC0000000u);
        }
        // rows 7blk..7blk+6 sit at bits 18..31: keep them as the low/high half of the lane's word blk/2
        pb16[((blk >> 1) * 32) * 2 + (blk & 1)] = (uint16_t)(acc >> 16);
      }
    }
    __syncwarp();
    // 3. this lane's pass bits, rows past the zone masked off
    uint32_t pw[FS_MAXB2];
    int cnt = 0;
#pragma unroll
    for (int b = 0; b < FS_MAXB2; b++) {
      const int nv = min(max(zh - 14 * b, 0), 14);
      const int nlo = min(nv, 7), nhi = nv - nlo;
      const uint32_t m = (((1u << (2 * nlo)) - 1u) << 2) | (((1u << (2 * nhi)) - 1u) << 18);
      pw[b] = nv > 0 ? (s.pb[b * 32 + lane] & m) : 0u;
      cnt += __popc(pw[b]);
    }
    int inc = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    const int total_tasks = __shfl_sync(0xffffffffu, inc, 31), off = inc - cnt;
    // 4./5. score, then NMS
    if (total_tasks <= FS_CAP) {
      fs_fill(s, pw, off, 0, lane);
      const int nk = fs_score<true>(s, total_tasks, th, lane);
      fs_nms(s, nk, lane);
    } else {
      for (int cb = 0; cb < total_tasks; cb += FS_CAP) {
        fs_fill(s, pw, off, cb, lane);
        fs_score<false>(s, min(FS_CAP, total_tasks - cb), th, lane);
      }
      for (int cb = 0; cb < total_tasks; cb += FS_CAP) {
        fs_fill(s, pw, off, cb, lane);
        fs_nms(s, min(FS_CAP, total_tasks - cb), lane);
      }
    }
    // 6. per cell: ordered emission
    uint32_t redo = 0;
    const uint8_t* ab = reinterpret_cast<const uint8_t*>(s.amap);
#pragma unroll 1
    for (int j = 0; j < k; j++) {
      if (!((todo >> j) & 1u)) continue;
      const int cs = cx0 + j * L.wCell, ce = min(cs + L.wCell, cx1);
      int total = 0;
      if (ce > cs) {
        const int wlo = cs >> 5, whi = (ce - 1) >> 5;
        const uint32_t mlo = 0xffffffffu << (cs & 31), mhi = 0xffffffffu >> (31 - ((ce - 1) & 31));
        uint32_t* out = out0 + (long long)j * L.cell_cap;
        for (int rb = 0; rb < zh; rb += 64) {
          const int r0 = rb + 2 * lane;
          int c0n = 0, c1n = 0;
          for (int w = wlo; w <= whi; w++) {
            const uint32_t m = (w == wlo ? mlo : 0xffffffffu) & (w == whi ? mhi : 0xffffffffu);
            if (r0 < zh) c0n += __popc(s.kb[r0 * FS_KBP + w] & m);
            if (r0 + 1 < zh) c1n += __popc(s.kb[(r0 + 1) * FS_KBP + w] & m);
          }
          int inc2 = c0n + c1n;
#pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc2, o);
            if (lane >= o) inc2 += t;
          }
          int pos = total + inc2 - (c0n + c1n);
          total += __shfl_sync(0xffffffffu, inc2, 31);
          if (c0n + c1n) {
            for (int rr = r0; rr < r0 + 2; rr++) {
              if (rr >= zh) break;
              for (int w = wlo; w <= whi; w++) {
                uint32_t bits = s.kb[rr * FS_KBP + w] & (w == wlo ? mlo : 0xffffffffu) & (w == whi ? mhi : 0xffffffffu);
                while (bits) {
                  const int c = 32 * w + __ffs(bits) - 1;
                  bits &= bits - 1;
                  const int a = ab[(rr + 1) * FS_AP + 4 + c];
                  // coordinates relative to (minBorderX,minBorderY)=(16,16), as DistributeOctTree sees them
                  out[pos++] = orbx_pack_key(xs + c - ORBX_MIN_BORDER, Y0 + rr - ORBX_MIN_BORDER, a - 1);
                }
              }
            }
          }
        }
        if (total == 0) redo |= 1u << j;
      }
      if (lane == 0) cnt_out[j] = total;
    }
    // 7. nothing at iniThFAST: the reference runs FAST again on the cell with minThFAST
    if (redo == 0 || th == P.min_th) break;
    th = P.min_th;
    todo = redo;
    __syncwarp();
    if (lane < 4) {
      uint32_t m = 0;
      for (int j = 0; j < k; j++) {
        if (!((redo >> j) & 1u)) continue;
        const int cs = cx0 + j * L.wCell, ce = min(cs + L.wCell, cx1);
        const int lo = max(cs - 32 * lane, 0), hi = min(ce - 32 * lane, 32);
        if (hi > lo) m |= (hi - lo == 32) ? 0xffffffffu : (((1u << (hi - lo)) - 1u) << lo);
      }
      s.misc[lane] = m;
    }
    __syncwarp();
  }
}

static size_t fs_region_words(int hmax) {
  const int amap_words = (FS_APW * (hmax + 2) + 3) & ~3, kb_words = (FS_KBP * hmax + 3) & ~3;
  return (size_t)FS_TPW * (hmax + 6) + amap_words + kb_words + 32 * FS_MAXB2 + FS_CAP / 2 + 12;
}

cudaError_t orbx_launch_fast_strip(const DevPlan& P, Level0 l0, const uint8_t* arena, const uint32_t* strip_tab,
                                   int32_t* cell_cnt, uint32_t* cand, int batch, cudaStream_t st) {
  int hmax = 0;
  for (int l = 0; l < P.nlevels; l++) hmax = max(hmax, P.lv[l].hCell);
  if (hmax > 14 * FS_MAXB2) return cudaErrorInvalidValue;
  const size_t region = (fs_region_words(hmax) + 3) & ~(size_t)3;
  const size_t smem = FS_WARPS * region * sizeof(uint32_t);
  cudaError_t e = orbx_ensure_dyn_smem(ORBX_SMEM_FAST_STRIP, (const void*)k_fast_strip, smem);
  if (e != cudaSuccess) return e;
  dim3 grd((P.total_strips + FS_WARPS - 1) / FS_WARPS, batch);
  k_fast_strip<<<grd, FS_THREADS, smem, st>>>(P, l0, arena, strip_tab, cell_cnt, cand, hmax, (int)region);
  return cudaGetLastError();
}

cudaError_t orbx_launch_fast(const DevPlan& P, Level0 l0, const uint8_t* arena, const uint32_t* cell_tab,
                             const uint32_t* strip_tab, int32_t* cell_cnt, uint32_t* cand, int batch, cudaStream_t st) {
  // ORBX_FAST_IMPL=cell selects the first-generation cell-per-warp kernel (A/B measurements)
  static const bool cell = [] { const char* e = getenv("ORBX_FAST_IMPL"); return e && e[0] == 'c'; }();
  if (cell) return orbx_launch_fast_cell(P, l0, arena, cell_tab, cell_cnt, cand, batch, st);
  return orbx_launch_fast_strip(P, l0, arena, strip_tab, cell_cnt, cand, batch, st);
}
