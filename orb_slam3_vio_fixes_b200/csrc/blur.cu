// blur.cu — GaussianBlur(level, 7x7, sigma 2, BORDER_REFLECT_101) of every pyramid level
// (src/ORBextractor.cc:1131-1133) for sm_100a: the "level blur" mode of the descriptor stage.
//
// cv::GaussianBlur on CV_8U runs a separable fixed-point filter with taps 18,34,48,56,48,34,18 (/256)
// and a single rounding: out = (sum_ij k_i k_j I + 2^15) >> 16.  Everything is exact integer work, so
// the two passes may run in either order, and because every tap is even the filter is evaluated with
// the half taps t' = 9,17,24,28,24,17,9:  out = (sum_ij t'_i t'_j I + 2^13) >> 14  (the same integer).
//
// No shared memory, no barriers.  A thread owns one 32-bit word column (4 pixels) and walks BL_RB rows
// down it; a warp = 32 adjacent word columns, so every load and store is one coalesced 128-byte access:
//   row pass     the 12-byte window (x-4 .. x+7) of the new row is three L1-resident word loads; the four
//                7-tap row sums are 6 funnel shifts + 8 dp4a against the packed half taps;
//   column pass  the last 7 rows of row sums (4 x 7 registers) are combined with the symmetric form
//                36(a0+a6) + 68(a1+a5) + 96(a2+a4) + 112 a3 + 2^15 (half taps x4: the result byte lands in
//                bits 16..23), packed with PRMT and stored as one word.
// BORDER_REFLECT_101: rows are reflected through the row index; at the left/right level edge the window
// words of the (at most two) affected lanes are rebuilt from the loaded words with per-lane PRMT
// selectors computed once before the walk (warps that touch no edge run the plain path).
#include "orbx_kernels.cuh"

#define BL_RB ORBX_BLUR_ROWS  // 35 output rows per thread (a multiple of 7: the row-sum window is indexed statically)
#define BL_WARPS 4
#define BL_TA 0x1c181109u  // half taps 9,17,24,28 on window bytes x-3 .. x
#define BL_TB 0x00091118u  // half taps 24,17,9,0  on window bytes x+1 .. x+4

struct BlurEdge {  // per-lane fix-ups of the window words at the level's left / right edge
  uint32_t s0, s1, s2a, s2b;
  bool has_l, has_r, w2_from_01;
};

// selector nibble source: window byte si (0..11 = columns x-4 .. x+7) relative to the register pair
// (R0,R1) -> si, or (R1,R2) -> si-4
__device__ __forceinline__ BlurEdge blur_edge_setup(int x, int w, int pitch) {
  BlurEdge e;
  e.has_l = x >= 4;
  e.has_r = x + 8 <= pitch;
  int si[12];
#pragma unroll
  for (int i = 0; i < 12; i++) {
    int c = x - 4 + i;
    if (c < 0) c = -c;
    if (c >= w) c = 2 * (w - 1) - c;
    int s = c - (x - 4);
    si[i] = s < 0 ? 0 : (s > 11 ? 11 : s);
  }
  // when R0 / R2 cannot be loaded they are replaced by R1; bytes that would come from them are unused
  e.s0 = (uint32_t)(si[0] & 7) | ((uint32_t)(si[1] & 7) << 4) | ((uint32_t)(si[2] & 7) << 8) | ((uint32_t)(si[3] & 7) << 12);
  e.s1 = (uint32_t)(si[4] & 7) | ((uint32_t)(si[5] & 7) << 4) | ((uint32_t)(si[6] & 7) << 8) | ((uint32_t)(si[7] & 7) << 12);
  // W2: sources either all in (R1,R2) (si >= 4) or, when the level ends inside this lane's word, in (R0,R1)
  e.w2_from_01 = (w - 1) < x + 3;
  e.s2a = e.s2b = 0;
#pragma unroll
  for (int j = 0; j < 4; j++) {
    const int s = si[8 + j];
    e.s2a |= (uint32_t)(s & 7) << (4 * j);                        // from (R0,R1): valid when s < 8
    e.s2b |= (uint32_t)((s >= 4 ? s - 4 : 0) & 7) << (4 * j);      // from (R1,R2): valid when s >= 4
  }
  return e;
}

// The three raw window words of one image row (next row is fetched while the current one is being filtered).
struct BlurRaw { uint32_t r0, r1, r2; };

template <bool EDGE>
__device__ __forceinline__ BlurRaw blur_load(const uint8_t* __restrict__ rowp, bool active, const BlurEdge& e) {
  BlurRaw v;
  const uint32_t* p = reinterpret_cast<const uint32_t*>(rowp);
  if (!EDGE) {
    v.r0 = __ldg(p - 1); v.r1 = __ldg(p); v.r2 = __ldg(p + 1);
  } else {
    v.r0 = v.r1 = v.r2 = 0;
    if (active) {
      v.r1 = __ldg(p);
      v.r0 = e.has_l ? __ldg(p - 1) : v.r1;
      v.r2 = e.has_r ? __ldg(p + 1) : v.r1;
    }
  }
  return v;
}

// Row sums for this lane's 4 pixels.  EDGE: lanes at the level border rebuild their window first.
template <bool EDGE>
__device__ __forceinline__ void blur_row_sums(const BlurRaw& v, const BlurEdge& e, uint32_t (&H)[4]) {
  uint32_t w0 = v.r0, w1 = v.r1, w2 = v.r2;
  if (EDGE) {
    w0 = __byte_perm(v.r0, v.r1, e.s0);
    w1 = __byte_perm(v.r0, v.r1, e.s1);
    const uint32_t ua = __byte_perm(v.r0, v.r1, e.s2a), ub = __byte_perm(v.r1, v.r2, e.s2b);
    w2 = e.w2_from_01 ? ua : ub;
  }
  H[0] = __dp4a(__funnelshift_r(w0, w1, 8), BL_TA, __dp4a(__funnelshift_r(w1, w2, 8), BL_TB, 0u));
  H[1] = __dp4a(__funnelshift_r(w0, w1, 16), BL_TA, __dp4a(__funnelshift_r(w1, w2, 16), BL_TB, 0u));
  H[2] = __dp4a(__funnelshift_r(w0, w1, 24), BL_TA, __dp4a(__funnelshift_r(w1, w2, 24), BL_TB, 0u));
  H[3] = __dp4a(w1, BL_TA, __dp4a(w2, BL_TB, 0u));
}

// Column pass for one output row: H rows oldest..newest = a0..a6.  The column taps are the half taps x4, so
// the rounded result byte sits in bits 16..23 of the accumulator and PRMT packs four of them.
__device__ __forceinline__ uint32_t blur_col(const uint32_t (&a0)[4], const uint32_t (&a1)[4], const uint32_t (&a2)[4],
                                             const uint32_t (&a3)[4], const uint32_t (&a4)[4], const uint32_t (&a5)[4],
                                             const uint32_t (&a6)[4]) {
  uint32_t acc[4];
#pragma unroll
  for (int k = 0; k < 4; k++)
    acc[k] = 36u * (a0[k] + a6[k]) + 68u * (a1[k] + a5[k]) + 96u * (a2[k] + a4[k]) + 112u * a3[k] + 32768u;
  return __byte_perm(__byte_perm(acc[0], acc[1], 0x0062), __byte_perm(acc[2], acc[3], 0x0062), 0x5410);
}

// ROWS: VEDGE = the tile touches the first / last rows of the level (row index reflected per row);
// FULL = the tile has all BL_RB output rows (static trip counts).
template <bool EDGE, bool VEDGE, bool FULL>
__device__ __forceinline__ void blur_walk(const uint8_t* __restrict__ img, int pitch, int w, int h, int x, int y0,
                                          bool active, uint8_t* __restrict__ out, int opitch) {
  BlurEdge e = {};
  if (EDGE) e = blur_edge_setup(x, w, pitch);
  else active = true;
  const int nout = FULL ? BL_RB : min(BL_RB, h - y0);
  uint32_t H[7][4];
  const uint8_t* colp = img + x;
  const uint8_t* rowp = colp + (long long)(y0 - 3) * pitch;  // !VEDGE: rows y0-3 .. y0+nout+3 are all inside
  uint8_t* op = out + (long long)y0 * opitch + x;
  int yy = y0 - 3;
  const int ylast = y0 + nout + 2;  // last source row (unreflected index) this tile needs
  if (VEDGE) {
    int sy = yy < 0 ? -yy : yy;
    sy = sy >= h ? 2 * (h - 1) - sy : sy;
    rowp = colp + (long long)sy * pitch;
  }
  BlurRaw nxt = blur_load<EDGE>(rowp, active, e);
  // filter the prefetched row into H[slot] after issuing the loads of the row below it
#define BLUR_ROW(slot)                                                            \
  do {                                                                            \
    const BlurRaw cur = nxt;                                                      \
    yy++;                                                                         \
    if (VEDGE) {                                                                  \
      int sy = yy < 0 ? -yy : yy;                                                 \
      sy = sy >= h ? 2 * (h - 1) - sy : sy;                                       \
      rowp = colp + (long long)sy * pitch;                                        \
    } else {                                                                      \
      rowp += pitch;                                                              \
    }                                                                             \
    if (FULL || yy <= ylast) nxt = blur_load<EDGE>(rowp, active, e);              \
    blur_row_sums<EDGE>(cur, e, H[slot]);                                         \
  } while (0)
#define BLUR_OUT(s)                                                                                           \
  do {                                                                                                        \
    const uint32_t v = blur_col(H[(s + 1) % 7], H[(s + 2) % 7], H[(s + 3) % 7], H[(s + 4) % 7], H[(s + 5) % 7], \
                                H[(s + 6) % 7], H[s]);                                                        \
    if (active) *reinterpret_cast<uint32_t*>(op) = v;                                                         \
    op += opitch;                                                                                             \
  } while (0)
  BLUR_ROW(0); BLUR_ROW(1); BLUR_ROW(2); BLUR_ROW(3); BLUR_ROW(4); BLUR_ROW(5);
  if (FULL) {
#pragma unroll 1
    for (int g = 0; g < BL_RB / 7; g++) {
      BLUR_ROW(6); BLUR_OUT(6);
      BLUR_ROW(0); BLUR_OUT(0);
      BLUR_ROW(1); BLUR_OUT(1);
      BLUR_ROW(2); BLUR_OUT(2);
      BLUR_ROW(3); BLUR_OUT(3);
      BLUR_ROW(4); BLUR_OUT(4);
      BLUR_ROW(5); BLUR_OUT(5);
    }
  } else {
    int left = nout;
#pragma unroll 1
    while (left > 0) {
      BLUR_ROW(6); BLUR_OUT(6); if (--left == 0) break;
      BLUR_ROW(0); BLUR_OUT(0); if (--left == 0) break;
      BLUR_ROW(1); BLUR_OUT(1); if (--left == 0) break;
      BLUR_ROW(2); BLUR_OUT(2); if (--left == 0) break;
      BLUR_ROW(3); BLUR_OUT(3); if (--left == 0) break;
      BLUR_ROW(4); BLUR_OUT(4); if (--left == 0) break;
      BLUR_ROW(5); BLUR_OUT(5); --left;
    }
  }
#undef BLUR_ROW
#undef BLUR_OUT
}

// Unaligned level-0 input (caller's buffer with an odd pitch or base): same arithmetic from byte loads.
__device__ __forceinline__ void blur_walk_bytes(const uint8_t* __restrict__ img, int pitch, int w, int h, int x,
                                                int y0, bool active, uint8_t* __restrict__ out, int opitch) {
  if (!active) return;
  const int nout = min(BL_RB, h - y0);
  const int taps[7] = {9, 17, 24, 28, 24, 17, 9};
  for (int r = 0; r < nout; r++) {
    uint32_t o[4];
    for (int k = 0; k < 4; k++) {
      uint32_t acc = 8192u;
      for (int dy = -3; dy <= 3; dy++) {
        const uint8_t* rowp = img + (long long)orbx_reflect101(y0 + r + dy, h) * pitch;
        uint32_t hs = 0;
        for (int dx = -3; dx <= 3; dx++) {
          int c = x + k + dx;
          c = c < 0 ? -c : c;
          c = c >= w ? 2 * (w - 1) - c : c;
          c = c < 0 ? 0 : c;  // columns past the level edge (never stored as valid pixels)
          hs += (uint32_t)taps[dx + 3] * rowp[c];
        }
        acc += (uint32_t)taps[dy + 3] * hs;
      }
      o[k] = acc >> 14;
    }
    *reinterpret_cast<uint32_t*>(out + (long long)(y0 + r) * opitch + x) = o[0] | (o[1] << 8) | (o[2] << 16) | (o[3] << 24);
  }
}

__global__ void __launch_bounds__(BL_WARPS * 32)
k_blur_levels(const __grid_constant__ DevPlan P, Level0 l0, const uint8_t* __restrict__ arena,
              uint8_t* __restrict__ blur) {
  const int lane = threadIdx.x & 31, frame = blockIdx.y;
  int tile = blockIdx.x * BL_WARPS + (threadIdx.x >> 5);
  if (tile >= P.total_btiles) return;
  int level = 0;
#pragma unroll 1
  for (int l = 1; l < P.nlevels; l++)
    if (tile >= P.lv[l].btile_base) level = l;
  const DevLevel& L = P.lv[level];
  tile -= L.btile_base;
  const int ty = tile / L.btiles_x, tx = tile - ty * L.btiles_x;
  const int w = L.w, h = L.h;
  const int x = 4 * (tx * 32 + lane), y0 = ty * BL_RB;
  int pitch;
  const uint8_t* img;
  if (level == 0) { pitch = l0.pitch; img = l0.base + (long long)frame * l0.frame_stride; }
  else { pitch = L.pitch; img = arena + (long long)frame * P.arena_bytes + L.offset; }
  uint8_t* out = blur + (long long)frame * P.arena_bytes + L.offset;
  const bool active = x < w;
  if ((pitch & 3) != 0 || ((uintptr_t)img & 3) != 0) {
    blur_walk_bytes(img, pitch, w, h, x, y0, active, out, L.pitch);
    return;
  }
  // warp-uniform: does any lane's 12-byte window leave [0, w)?
  const bool edge = tx == 0 || 4 * (tx * 32 + 33) > w;
  const bool full = y0 + BL_RB <= h;
  const bool vedge = y0 < 3 || y0 + BL_RB + 4 > h;  // FULL tiles also fetch one row past their last
  if (full && !vedge) {
    if (edge) blur_walk<true, false, true>(img, pitch, w, h, x, y0, active, out, L.pitch);
    else blur_walk<false, false, true>(img, pitch, w, h, x, y0, active, out, L.pitch);
  } else {  // first / last tile rows of a level: one generic variant (keeps the code footprint small)
    blur_walk<true, true, false>(img, pitch, w, h, x, y0, active, out, L.pitch);
  }
}

cudaError_t orbx_launch_blur(const DevPlan& P, Level0 l0, const uint8_t* arena, uint8_t* blur, int batch,
                             cudaStream_t st) {
  dim3 grd((P.total_btiles + BL_WARPS - 1) / BL_WARPS, batch);
  k_blur_levels<<<grd, BL_WARPS * 32, 0, st>>>(P, l0, arena, blur);
  return cudaGetLastError();
}
