// blur.cu — GaussianBlur(level, 7x7, sigma 2, BORDER_REFLECT_101) of every pyramid level
// (src/ORBextractor.cc:1131-1133) for sm_100a: the "level blur" mode of the descriptor stage.
//
// cv::GaussianBlur on CV_8U runs a separable fixed-point filter with taps 18,34,48,56,48,34,18 (/256)
// and a single rounding: out = (sum_ij k_i k_j I + 2^15) >> 16.  Everything is exact integer work, so
// the two passes may run in either order, and because every tap is even the filter is evaluated with
// the half taps t' = 9,17,24,28,24,17,9:  out = (sum_ij t'_i t'_j I + 2^13) >> 14  (the same integer).
//
// No shared memory, no barriers.  A thread owns one 32-bit word column (4 pixels) and walks BL_RB rows
// down it; a warp = 32 adjacent word columns (lanes 1..30 produce output, lanes 0 and 31 carry the
// horizontal halo), so every load and store is one coalesced access and each pixel is loaded ONCE:
//   column pass  first, on the raw bytes: the word is split into its even / odd bytes (two pixels per
//                register in 16-bit lanes); with the half taps the 7-row sum of a pixel is <= 255*128 =
//                32640, so 9(a0+a6) + 17(a1+a5) + 24(a2+a4) + 28 a3 runs on two pixels per IADD / IMAD;
//   row pass     the neighbours' column sums come from the adjacent lanes (4 SHFL); the 7-tap row sums of
//                the four pixels are 18 dp2a (u16x2 . u8x2) against the half taps x4 (36,68,96,112: the
//                rounded result byte then lands in bits 16..23), packed with PRMT and stored as one word.
// Loads run four rows ahead of the arithmetic (the kernel was latency-bound on its row loads
// with a one-row prefetch: profiles/r2a_*).
// BORDER_REFLECT_101: rows are reflected through the row index; a lane whose word is not entirely inside [0,w)
// rebuilds it from the two aligned words that hold its four reflected columns (one PRMT, selector per lane).
#include "orbx_kernels.cuh"

#ifndef BL_PTR
#define BL_PTR 1   // 1: row-pointer walk in the interior variant, 2: in the edge variant too, 0: off
#endif
#ifndef BL_MINB
#define BL_MINB 8  // 64 registers: unconstrained ptxas takes 94 and the walk loses its latency cover (0.96 -> 0.91 ms describe stage
                   // at C1 with the cap, the row pointer and the packed tap constants together; profiles/r2ag_blur_variants.txt)
#endif

#define BL_RB ORBX_BLUR_ROWS  // 35 output rows per thread (a multiple of 7: the row window is indexed statically)
#define BL_WARPS 4
#define BL_TW ORBX_BLUR_TILE_WORDS  // 30 output words per warp row

// even / odd bytes of a word as two 16-bit lanes each
#define BL_EVEN(w) __byte_perm((w), 0u, 0x4240)
#define BL_ODD(w) __byte_perm((w), 0u, 0x4341)

__device__ __forceinline__ uint32_t blur_vsum(uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t a4,
                                              uint32_t a5, uint32_t a6) {
  return 9u * (a0 + a6) + 17u * (a1 + a5) + 24u * (a2 + a4) + 28u * a3;
}
// (Measured and dropped: the same sum from adds and rotates on the ALU pipe only — 8 ALU instructions instead of
//  4 IMAD + 3 adds — for one or both byte parities: describe stage 0.915 / 0.924 ms against 0.905; the fmaheavy pipe
//  is the busiest unit at 71 % but the kernel also issues at 71 % of its slots, so extra instructions cost as much
//  as the multiplies they remove.)

// Row pass + pack for this lane's 4 pixels from the column sums of pixels -4..7 around it:
// Le=(V-4,V-2) Lo=(V-3,V-1) Ce=(V0,V2) Co=(V1,V3) Re=(V4,V6) Ro=(V5,V7), 16-bit lanes (low, high).
#define BL_K(lo, hi) ((uint32_t)(lo) | ((uint32_t)(hi) << 8))
// Two tap pairs share one 32-bit constant (dp2a.lo reads bytes 0-1 of the multiplier, dp2a.hi bytes 2-3): five
// constants instead of nine, so they stay in registers instead of being re-materialised row after row.
#ifndef BL_PACKK
#define BL_PACKK 1
#endif
#if !BL_PACKK
__device__ __forceinline__ uint32_t blur_hsum(uint32_t Le, uint32_t Lo, uint32_t Ce, uint32_t Co, uint32_t Re,
                                              uint32_t Ro) {
  uint32_t a0 = __dp2a_lo(Lo, BL_K(36, 96), 32768u);   // 36 V-3 + 96 V-1
  a0 = __dp2a_lo(Le, BL_K(0, 68), a0);                 // 68 V-2
  a0 = __dp2a_lo(Ce, BL_K(112, 68), a0);               // 112 V0 + 68 V2
  a0 = __dp2a_lo(Co, BL_K(96, 36), a0);                // 96 V1 + 36 V3
  uint32_t a1 = __dp2a_lo(Le, BL_K(0, 36), 32768u);    // 36 V-2
  a1 = __dp2a_lo(Lo, BL_K(0, 68), a1);                 // 68 V-1
  a1 = __dp2a_lo(Ce, BL_K(96, 96), a1);                // 96 V0 + 96 V2
  a1 = __dp2a_lo(Co, BL_K(112, 68), a1);               // 112 V1 + 68 V3
  a1 = __dp2a_lo(Re, BL_K(36, 0), a1);                 // 36 V4
  uint32_t a2 = __dp2a_lo(Lo, BL_K(0, 36), 32768u);    // 36 V-1
  a2 = __dp2a_lo(Ce, BL_K(68, 112), a2);               // 68 V0 + 112 V2
  a2 = __dp2a_lo(Co, BL_K(96, 96), a2);                // 96 V1 + 96 V3
  a2 = __dp2a_lo(Re, BL_K(68, 0), a2);                 // 68 V4
  a2 = __dp2a_lo(Ro, BL_K(36, 0), a2);                 // 36 V5
  uint32_t a3 = __dp2a_lo(Ce, BL_K(36, 96), 32768u);   // 36 V0 + 96 V2
  a3 = __dp2a_lo(Co, BL_K(68, 112), a3);               // 68 V1 + 112 V3
  a3 = __dp2a_lo(Re, BL_K(96, 36), a3);                // 96 V4 + 36 V6
  a3 = __dp2a_lo(Ro, BL_K(68, 0), a3);                 // 68 V5
  return __byte_perm(__byte_perm(a0, a1, 0x0062), __byte_perm(a2, a3, 0x0062), 0x5410);
}
#else
#define BL_KK(p, q) ((uint32_t)(p) | ((uint32_t)(q) << 16))
__device__ __forceinline__ uint32_t blur_hsum(uint32_t Le, uint32_t Lo, uint32_t Ce, uint32_t Co, uint32_t Re,
                                              uint32_t Ro) {
  const uint32_t KA = BL_KK(BL_K(36, 96), BL_K(0, 68));     // lo: 36,96   hi: 0,68
  const uint32_t KB = BL_KK(BL_K(112, 68), BL_K(96, 36));   // lo: 112,68  hi: 96,36
  const uint32_t KC = BL_KK(BL_K(0, 36), BL_K(96, 96));     // lo: 0,36    hi: 96,96
  const uint32_t KD = BL_KK(BL_K(36, 0), BL_K(68, 112));    // lo: 36,0    hi: 68,112
  const uint32_t KE = BL_KK(BL_K(68, 0), 0);                // lo: 68,0
  uint32_t a0 = __dp2a_lo(Lo, KA, 32768u);   // 36 V-3 + 96 V-1
  a0 = __dp2a_hi(Le, KA, a0);                // 68 V-2
  a0 = __dp2a_lo(Ce, KB, a0);                // 112 V0 + 68 V2
  a0 = __dp2a_hi(Co, KB, a0);                // 96 V1 + 36 V3
  uint32_t a1 = __dp2a_lo(Le, KC, 32768u);   // 36 V-2
  a1 = __dp2a_hi(Lo, KA, a1);                // 68 V-1
  a1 = __dp2a_hi(Ce, KC, a1);                // 96 V0 + 96 V2
  a1 = __dp2a_lo(Co, KB, a1);                // 112 V1 + 68 V3
  a1 = __dp2a_lo(Re, KD, a1);                // 36 V4
  uint32_t a2 = __dp2a_lo(Lo, KC, 32768u);   // 36 V-1
  a2 = __dp2a_hi(Ce, KD, a2);                // 68 V0 + 112 V2
  a2 = __dp2a_hi(Co, KC, a2);                // 96 V1 + 96 V3
  a2 = __dp2a_lo(Re, KE, a2);                // 68 V4
  a2 = __dp2a_lo(Ro, KD, a2);                // 36 V5
  uint32_t a3 = __dp2a_lo(Ce, KA, 32768u);   // 36 V0 + 96 V2
  a3 = __dp2a_hi(Co, KD, a3);                // 68 V1 + 112 V3
  a3 = __dp2a_hi(Re, KB, a3);                // 96 V4 + 36 V6
  a3 = __dp2a_lo(Ro, KE, a3);                // 68 V5
  return __byte_perm(__byte_perm(a0, a1, 0x0062), __byte_perm(a2, a3, 0x0062), 0x5410);
}
#endif

// EDGE  = some lane of the warp has a word that is not entirely inside [0,w) (warp-uniform);
// VEDGE = the tile touches the first / last rows of the level or is shorter than BL_RB rows.
template <bool EDGE, bool VEDGE>
__device__ __forceinline__ void blur_walk(const uint8_t* __restrict__ img, int pitch, int w, int h, int wq, int y0,
                                          uint8_t* __restrict__ out, int opitch, int lane) {
  const int c = 4 * wq;
  // EDGE: the lane's four columns reflected into [0,w) are 4 consecutive-ish columns, so they always lie in two
  // adjacent aligned words: two word loads (byte offsets qa, qb) and one PRMT with a per-lane selector rebuild the
  // lane's word; interior lanes of the same warp get the identity selector.  No divergence, no byte loads.
  int qa = c, qb = c;
  uint32_t sel = 0x3210u;
  if (EDGE) {
    int xs[4], lo = 1 << 30;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      int x = c + j;
      x = x < 0 ? -x : x;
      x = x >= w ? 2 * (w - 1) - x : x;
      xs[j] = x < 0 ? 0 : (x >= w ? w - 1 : x);  // lanes further out are never used: any valid column will do
      lo = min(lo, xs[j]);
    }
    qa = lo & ~3;
    qb = qa;
    sel = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      if (xs[j] >= qa + 4) qb = qa + 4;
      sel |= (uint32_t)(xs[j] - qa) << (4 * j);
    }
  }
  const int nout = VEDGE ? min(BL_RB, h - y0) : BL_RB;
  // !VEDGE: the source rows are consecutive, so the walk keeps a row pointer instead of multiplying out every
  // row's address (5 instructions per row, 3 of them on the fmaheavy pipe)
  const uint8_t* rpa = img + (long long)(y0 - 3) * pitch + (EDGE ? qa : c);
  const int qd = qb - qa;
  auto fetch = [&](int yy) -> uint32_t {
    if (!VEDGE && BL_PTR && (!EDGE || BL_PTR > 1)) {
      const uint8_t* rp = rpa;
      rpa += pitch;
      if (!EDGE) return __ldg(reinterpret_cast<const uint32_t*>(rp));
      return __byte_perm(__ldg(reinterpret_cast<const uint32_t*>(rp)), __ldg(reinterpret_cast<const uint32_t*>(rp + qd)), sel);
    }
    int sy = yy;
    if (VEDGE) {
      sy = sy < 0 ? -sy : sy;
      sy = sy >= h ? 2 * (h - 1) - sy : sy;
      sy = sy < 0 ? 0 : (sy >= h ? h - 1 : sy);  // rows past the reflected band are never used
    }
    const uint8_t* rp = img + (long long)sy * pitch;
    if (!EDGE) return __ldg(reinterpret_cast<const uint32_t*>(rp + c));
    return __byte_perm(__ldg(reinterpret_cast<const uint32_t*>(rp + qa)), __ldg(reinterpret_cast<const uint32_t*>(rp + qb)),
                       sel);
  };
  const bool writer = lane >= 1 && lane <= BL_TW && c < w;  // c >= 0 for lanes >= 1
  uint8_t* op = out + (long long)y0 * opitch + c;
  // E/O: the last 7 source rows (even / odd bytes), slot = source row mod 7; q0..q3: raw words of the next four
  // source rows, already on their way (the loads run 4 rows ahead of the arithmetic).  Source row r of the walk
  // is level row y0 - 3 + r; output row o needs source rows o .. o+6.
  uint32_t E[7], O[7];
  int yy = y0 - 3;
  uint32_t q0 = fetch(yy), q1 = fetch(yy + 1), q2 = fetch(yy + 2), q3 = fetch(yy + 3);
  yy += 4;
#define BL_ROW(s)                                    \
  do {                                               \
    const uint32_t raw = q0;                         \
    q0 = q1; q1 = q2; q2 = q3; q3 = fetch(yy); yy++; \
    E[s] = BL_EVEN(raw); O[s] = BL_ODD(raw);         \
  } while (0)
#define BL_OUT(s)                                                                                                  \
  do {                                                                                                             \
    const uint32_t Ce = blur_vsum(E[(s + 1) % 7], E[(s + 2) % 7], E[(s + 3) % 7], E[(s + 4) % 7], E[(s + 5) % 7],   \
                                  E[(s + 6) % 7], E[s]);                                                           \
    const uint32_t Co = blur_vsum(O[(s + 1) % 7], O[(s + 2) % 7], O[(s + 3) % 7], O[(s + 4) % 7], O[(s + 5) % 7],   \
                                  O[(s + 6) % 7], O[s]);                                                           \
    const uint32_t Le = __shfl_up_sync(0xffffffffu, Ce, 1), Lo = __shfl_up_sync(0xffffffffu, Co, 1);               \
    const uint32_t Re = __shfl_down_sync(0xffffffffu, Ce, 1), Ro = __shfl_down_sync(0xffffffffu, Co, 1);           \
    const uint32_t v = blur_hsum(Le, Lo, Ce, Co, Re, Ro);                                                          \
    if (writer && (!VEDGE || left > 0)) *reinterpret_cast<uint32_t*>(op) = v;                                      \
    op += opitch;                                                                                                  \
    if (VEDGE) left--;                                                                                             \
  } while (0)
  int left = nout;
  BL_ROW(0); BL_ROW(1); BL_ROW(2); BL_ROW(3); BL_ROW(4); BL_ROW(5);
  // (the queue runs up to 4 source rows past the last one used: in range for !VEDGE tiles, clamped otherwise)
#pragma unroll 1
  for (int g = 0; g < BL_RB / 7; g++) {
    BL_ROW(6); BL_OUT(6);
    BL_ROW(0); BL_OUT(0);
    BL_ROW(1); BL_OUT(1);
    BL_ROW(2); BL_OUT(2);
    BL_ROW(3); BL_OUT(3);
    BL_ROW(4); BL_OUT(4);
    BL_ROW(5); BL_OUT(5);
  }
#undef BL_ROW
#undef BL_OUT
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

__global__ void __launch_bounds__(BL_WARPS * 32, BL_MINB)
k_blur_levels(const __grid_constant__ DevPlan P, Level0 l0, const uint8_t* __restrict__ arena,
              uint8_t* __restrict__ blur, int l0_bytes) {
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
  const int y0 = ty * BL_RB;
  int pitch;
  const uint8_t* img;
  if (level == 0) { pitch = l0.pitch; img = l0.base + (long long)frame * l0.frame_stride; }
  else { pitch = L.pitch; img = arena + (long long)frame * P.arena_bytes + L.offset; }
  uint8_t* out = blur + (long long)frame * P.arena_bytes + L.offset;
  if (level == 0 && l0_bytes) return;  // unaligned caller buffer: k_blur_level0_bytes does level 0
  const int wq = tx * BL_TW + lane - 1;  // word column of this lane; lanes 0 and 31 are halo
  // warp-uniform: is any lane's word (word columns tx*BL_TW-1 .. tx*BL_TW+30) not entirely inside [0,w)?
  const bool edge = tx == 0 || 4 * (tx * BL_TW + BL_TW + 1) > w;
  // the walk fetches source rows y0-3 .. y0+BL_RB+2
  const bool vedge = y0 < 3 || y0 + BL_RB + 7 > h;  // (the load queue runs 4 rows past the last row it uses)
  if (!vedge) {
    if (edge) blur_walk<true, false>(img, pitch, w, h, wq, y0, out, L.pitch, lane);
    else blur_walk<false, false>(img, pitch, w, h, wq, y0, out, L.pitch, lane);
  } else {  // first / last tile rows of a level: one generic variant (keeps the code footprint small)
    blur_walk<true, true>(img, pitch, w, h, wq, y0, out, L.pitch, lane);
  }
}

// Level 0 of a caller-owned buffer whose pitch or base is not a multiple of 4: the same arithmetic from byte loads.
__global__ void __launch_bounds__(BL_WARPS * 32)
k_blur_level0_bytes(const __grid_constant__ DevPlan P, Level0 l0, uint8_t* __restrict__ blur) {
  const int lane = threadIdx.x & 31, frame = blockIdx.y;
  const int tile = blockIdx.x * BL_WARPS + (threadIdx.x >> 5);
  const DevLevel& L = P.lv[0];
  const int ntiles = P.nlevels > 1 ? P.lv[1].btile_base : P.total_btiles;
  if (tile >= ntiles) return;
  const int ty = tile / L.btiles_x, tx = tile - ty * L.btiles_x;
  const int x = 4 * (tx * BL_TW + lane);
  if (lane < BL_TW)
    blur_walk_bytes(l0.base + (long long)frame * l0.frame_stride, l0.pitch, L.w, L.h, x, ty * BL_RB, x < L.w,
                    blur + (long long)frame * P.arena_bytes + L.offset, L.pitch);
}

cudaError_t orbx_launch_blur(const DevPlan& P, Level0 l0, const uint8_t* arena, uint8_t* blur, int batch,
                             cudaStream_t st) {
  const int l0_bytes = ((l0.pitch & 3) != 0 || ((uintptr_t)l0.base & 3) != 0 || (l0.frame_stride & 3) != 0) ? 1 : 0;
  dim3 grd((P.total_btiles + BL_WARPS - 1) / BL_WARPS, batch);
  k_blur_levels<<<grd, BL_WARPS * 32, 0, st>>>(P, l0, arena, blur, l0_bytes);
  if (l0_bytes) {
    const int nt0 = P.nlevels > 1 ? P.lv[1].btile_base : P.total_btiles;
    dim3 g0((nt0 + BL_WARPS - 1) / BL_WARPS, batch);
    k_blur_level0_bytes<<<g0, BL_WARPS * 32, 0, st>>>(P, l0, blur);
  }
  return cudaGetLastError();
}
