// pyramid.cu — image pyramid (ORBextractor::ComputePyramid, src/ORBextractor.cc:1170-1195) for sm_100a.
#include <stdlib.h>
#include <string.h>

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
  int sa;                     // funnel shift in bits: 8 * (ofs_0 & 3)
  int o1, o2;                 // byte offsets of the 2nd and 3rd source word (4, 8 unless clamped)
};

// H[i] = ((r0*c0 + r1*c1) >> 4) << 16 for the thread's 4 columns on one source row.
// p = the row's word that holds byte ofs_0.
template <bool CLAMP>
__device__ __forceinline__ void rs_row(const uint8_t* __restrict__ p, const RsTapRegs& T, uint32_t (&H)[4]) {
  const uint32_t w0 = __ldg(reinterpret_cast<const uint32_t*>(p));
  const uint32_t w1 = __ldg(reinterpret_cast<const uint32_t*>(p + (CLAMP ? T.o1 : 4)));
  const uint32_t w2 = __ldg(reinterpret_cast<const uint32_t*>(p + (CLAMP ? T.o2 : 8)));
  const uint32_t lo = __funnelshift_r(w0, w1, T.sa);  // source bytes ofs_0 .. ofs_0+3
  const uint32_t hi = __funnelshift_r(w1, w2, T.sa);  // ofs_0+4 .. ofs_0+7
  const uint32_t mid = __funnelshift_r(lo, hi, 16);   // ofs_0+2 .. ofs_0+5
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const uint32_t win = i < 2 ? lo : mid;
    const uint32_t h = __byte_perm(win, 0, T.selA[i]) * T.C0[i] + __byte_perm(win, 0, T.selB[i]) * T.C1[i];
    H[i] = h & 0xffff0000u;
  }
}

// One output row: source rows sy0 / sy1 live in slots P / P^1 of the two-row register cache
// (slot = parity of the source row, so a row shared by consecutive output rows never moves).
template <int P, bool CLAMP>
__device__ __forceinline__ uint32_t rs_out_row(const uint8_t* __restrict__ base, int spitch, int sy0, int sy1,
                                               uint32_t b0, uint32_t b1, const RsTapRegs& T,
                                               uint32_t (&H)[2][4], int (&held)[2]) {
  if (held[P] != sy0) {
    rs_row<CLAMP>(base + sy0 * spitch, T, H[P]);
    held[P] = sy0;
  }
  if (held[P ^ 1] != sy1) {
    rs_row<CLAMP>(base + sy1 * spitch, T, H[P ^ 1]);
    held[P ^ 1] = sy1;
  }
  uint32_t out = 0;
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const uint32_t v = (__umulhi(b0, H[P][i]) + __umulhi(b1, H[P ^ 1][i]) + 2u) >> 2;  // <= 255
    out |= v << (8 * i);
  }
  return out;
}

template <bool CLAMP>
__global__ void __launch_bounds__(RS_WARPS * 32)
k_resize_walk(const uint8_t* __restrict__ src, int spitch, long long sframe, int sw, int sh,
              uint8_t* __restrict__ dst, int dpitch, long long dframe, int dw, int dh,
              const OrbxTap* __restrict__ xt, const OrbxTap* __restrict__ yt) {
  const int x4 = (blockIdx.x * 32 + (threadIdx.x & 31)) * 4;
  const int dy0 = (blockIdx.y * RS_WARPS + (threadIdx.x >> 5)) * RS_ROWS;
  if (x4 >= dw || dy0 >= dh) return;
  RsTapRegs T;
  const int o0 = xt[x4].ofs;
  {
    T.sa = 8 * (o0 & 3);
    const int wlast = (spitch >> 2) - 1, wa = o0 >> 2;  // CLAMP: never read past the row's last word
    T.o1 = 4 * (min(wa + 1, wlast) - wa);
    T.o2 = 4 * (min(wa + 2, wlast) - wa);
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
  const uint8_t* base = src + (long long)blockIdx.z * sframe + (o0 & ~3);
  uint8_t* out = dst + (long long)blockIdx.z * dframe + (long long)dy0 * dpitch + x4;
  uint32_t H[2][4];
  int held[2] = {-1, -1};
  const int dy1 = min(dy0 + RS_ROWS, dh);
#pragma unroll 1
  for (int dy = dy0; dy < dy1; dy++, out += dpitch) {
    const OrbxTap ty = yt[dy];
    const int sy0 = min(max(ty.ofs, 0), sh - 1), sy1 = min(max(ty.ofs + 1, 0), sh - 1);
    const uint32_t b0 = (uint32_t)ty.c0, b1 = (uint32_t)ty.c1;
    uint32_t v;
    if (sy0 & 1) v = rs_out_row<1, CLAMP>(base, spitch, sy0, sy1, b0, b1, T, H, held);
    else v = rs_out_row<0, CLAMP>(base, spitch, sy0, sy1, b0, b1, T, H, held);
    // rows are 128-byte aligned and padded, so the full word store stays inside the row
    *reinterpret_cast<uint32_t*>(out) = v;
  }
}

// ------------------------------------------------------------------------------------------------
// Tiled variant: the CTA's 128 x 64 output tile needs <= RT_SROWS source rows of <= RT_SPITCH bytes.
// They are brought into shared memory by the TMA engine (cp.async.bulk, one 16-byte-aligned row
// segment per copy, completion counted on an mbarrier), so the row walk below never waits on HBM:
// the only global traffic is one coalesced bulk read of the source tile and the word stores.
// ------------------------------------------------------------------------------------------------
#define RT_SROWS 88
#define RT_SPITCH 208

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void rs_row_smem(const uint8_t* p, const RsTapRegs& T, uint32_t (&H)[4]) {
  const uint32_t w0 = *reinterpret_cast<const uint32_t*>(p);
  const uint32_t w1 = *reinterpret_cast<const uint32_t*>(p + 4);
  const uint32_t w2 = *reinterpret_cast<const uint32_t*>(p + 8);
  const uint32_t lo = __funnelshift_r(w0, w1, T.sa);
  const uint32_t hi = __funnelshift_r(w1, w2, T.sa);
  const uint32_t mid = __funnelshift_r(lo, hi, 16);
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const uint32_t win = i < 2 ? lo : mid;
    const uint32_t h = __byte_perm(win, 0, T.selA[i]) * T.C0[i] + __byte_perm(win, 0, T.selB[i]) * T.C1[i];
    H[i] = h & 0xffff0000u;
  }
}

// Vertical taps + pack of one output word from the H rows of source rows sy0 (A) and sy0+1 (B).
__device__ __forceinline__ uint32_t rs_vert(const uint32_t (&A)[4], const uint32_t (&B)[4], uint32_t b0, uint32_t b1) {
  uint32_t v[4];
#pragma unroll
  for (int i = 0; i < 4; i++) v[i] = (__umulhi(b0, A[i]) + __umulhi(b1, B[i]) + 2u) >> 2;  // <= 255
  return __byte_perm(__byte_perm(v[0], v[1], 0x0040), __byte_perm(v[2], v[3], 0x0040), 0x5410);
}

// dp2a form of the horizontal taps (RS_DP2A, default): one PRMT puts the two source bytes of a column into bytes 0 / 1,
// IDP.2A against the packed coefficients (c0 | c1 << 16) gives h = r0*c0 + r1*c1 (< 2^19); the row keeps h & ~15 and the
// vertical blend uses umulhi(b << 12, h & ~15) = (b * (h >> 4)) >> 16 — the same integers as cv::resize.
#ifndef RS_DP2A
#define RS_DP2A 2   // 2: rows keep h >> 4 and the vertical blend uses 32-bit products (below); 1: umulhi form
#endif
struct RsTapRegs2 {
  uint32_t sel[4];   // PRMT selector: bytes (d, d+1) of the window -> bytes 0, 1; rest zero
  uint32_t coef[4];  // c0 | c1 << 16
  int sa;
};
__device__ __forceinline__ void rs_row_smem2(const uint8_t* p, const RsTapRegs2& T, uint32_t (&H)[4]) {
  const uint32_t w0 = *reinterpret_cast<const uint32_t*>(p);
  const uint32_t w1 = *reinterpret_cast<const uint32_t*>(p + 4);
  const uint32_t w2 = *reinterpret_cast<const uint32_t*>(p + 8);
  const uint32_t lo = __funnelshift_r(w0, w1, T.sa);
  const uint32_t hi = __funnelshift_r(w1, w2, T.sa);
  const uint32_t mid = __funnelshift_r(lo, hi, 16);
#pragma unroll
#if RS_DP2A == 2
  for (int i = 0; i < 4; i++) H[i] = __dp2a_lo(T.coef[i], __byte_perm(i < 2 ? lo : mid, 0, T.sel[i]), 0u) >> 4;
#else
  for (int i = 0; i < 4; i++) H[i] = __dp2a_lo(T.coef[i], __byte_perm(i < 2 ? lo : mid, 0, T.sel[i]), 0u) & ~15u;
#endif
}
// Vertical blend without the high-half multiplies (IMAD.HI takes two passes of the half-rate fmaheavy pipe, the
// kernel's busiest, and needs a zeroed addend register): with A = h >> 4 (< 2^15) and b <= 2048 the products fit
// 32 bits, so (b * A) >> 16 is the upper half of a plain IMAD; the upper halves of two pixels are gathered by one
// PRMT into 16-bit lanes, summed and rounded there ((x0 + x1 + 2) >> 2 <= 255), and packed by a last PRMT.
__device__ __forceinline__ uint32_t rs_vert3(const uint32_t (&A)[4], const uint32_t (&B)[4], uint32_t b0, uint32_t b1) {
  const uint32_t a01 = __byte_perm(b0 * A[0], b0 * A[1], 0x7632), a23 = __byte_perm(b0 * A[2], b0 * A[3], 0x7632);
  const uint32_t c01 = __byte_perm(b1 * B[0], b1 * B[1], 0x7632), c23 = __byte_perm(b1 * B[2], b1 * B[3], 0x7632);
  const uint32_t s01 = (a01 + c01 + 0x00020002u) >> 2, s23 = (a23 + c23 + 0x00020002u) >> 2;
  return __byte_perm(s01, s23, 0x6420);
}
__device__ __forceinline__ uint32_t rs_vert2(const uint32_t (&A)[4], const uint32_t (&B)[4], uint32_t b0s, uint32_t b1s) {
  uint32_t v[4];
#pragma unroll
  for (int i = 0; i < 4; i++) v[i] = (__umulhi(b0s, A[i]) + __umulhi(b1s, B[i]) + 2u) >> 2;  // <= 255
  return __byte_perm(__byte_perm(v[0], v[1], 0x0040), __byte_perm(v[2], v[3], 0x0040), 0x5410);
}

// TMA = the source window arrives as ONE 3-D tensor-map box (208 columns x box_rows rows of frame zf0 + blockIdx.z,
// zero-filled outside the level) issued by one thread; otherwise one cp.async.bulk per row (each lane's copy is
// serialised through the uniform datapath: ~10 % of the kernel's instructions, all before the first output row).
template <bool TMA>
__global__ void __launch_bounds__(RS_WARPS * 32)
k_resize_tile(const __grid_constant__ OrbxTmap map, int box_rows, int zf0,
              const uint8_t* __restrict__ src, int spitch, long long sframe, int sw, int sh,
              uint8_t* __restrict__ dst, int dpitch, long long dframe, int dw, int dh,
              const OrbxTap* __restrict__ xt, const OrbxTap* __restrict__ yt) {
  __shared__ __align__(128) uint8_t tile[RT_SROWS * RT_SPITCH];
  __shared__ __align__(8) uint64_t bar;
  const int tid = threadIdx.x;
  const int x0c = blockIdx.x * 128, dy0c = blockIdx.y * (RS_WARPS * RS_ROWS);
  // source window of the CTA (uniform)
  const int sy_first = min(max(yt[dy0c].ofs, 0), sh - 1);
  const int sy_last = min(max(yt[min(dy0c + RS_WARPS * RS_ROWS - 1, dh - 1)].ofs + 1, 0), sh - 1);
  const int nrows = sy_last - sy_first + 1;
  const int xs0 = xt[x0c].ofs & ~15;
  const int seg = min((xt[min(x0c + 127, dw - 1)].ofs + 12 - xs0 + 15) & ~15, spitch - xs0);
  const uint32_t bar_a = smem_u32(&bar);
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_a));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_a), "r"(TMA ? box_rows * RT_SPITCH : nrows * seg) : "memory");
    if (TMA)
      asm volatile(
          "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
          ::"r"(smem_u32(tile)), "l"(reinterpret_cast<const void*>(&map)), "r"(xs0), "r"(sy_first), "r"(zf0 + (int)blockIdx.z),
          "r"(bar_a)
          : "memory");
  }
  __syncthreads();
  if (!TMA) {
    const uint8_t* g = src + (long long)blockIdx.z * sframe + (long long)sy_first * spitch + xs0;
    for (int r = tid; r < nrows; r += RS_WARPS * 32)
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(smem_u32(tile + r * RT_SPITCH)), "l"(g + (long long)r * spitch), "r"(seg), "r"(bar_a)
                   : "memory");
  }
  const int x4 = x0c + (tid & 31) * 4;
  const int dy0 = dy0c + (tid >> 5) * RS_ROWS;
  const bool live = x4 < dw && dy0 < dh;
#if RS_DP2A
  RsTapRegs2 T;
#else
  RsTapRegs T;
#endif
  int o0 = 0;
  if (live) {
    o0 = xt[x4].ofs;
    T.sa = 8 * (o0 & 3);
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const OrbxTap t = xt[min(x4 + i, dw - 1)];
      int d = t.ofs - o0 - (i < 2 ? 0 : 2);
      d = min(max(d, 0), 2);  // only the padding columns beyond dw can fall outside
#if RS_DP2A
      T.sel[i] = 0x4400u | ((uint32_t)(d + 1) << 4) | (uint32_t)d;
      T.coef[i] = (uint32_t)(uint16_t)t.c0 | ((uint32_t)(uint16_t)t.c1 << 16);
#else
      T.selA[i] = 0x4440u | (uint32_t)d;
      T.selB[i] = 0x4404u | ((uint32_t)(d + 1) << 4);
      T.C0[i] = (uint32_t)t.c0 << 12;
      T.C1[i] = (uint32_t)t.c1 << 4;
#endif
    }
  }
  {  // wait for the tile
    uint32_t done = 0;
    while (!done)
      asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }"
                   : "=r"(done) : "r"(bar_a) : "memory");
  }
  if (!live) return;
  const uint8_t* base = tile + ((o0 & ~3) - xs0) - sy_first * RT_SPITCH;
  uint8_t* out = dst + (long long)blockIdx.z * dframe + (long long)dy0 * dpitch + x4;
  // Row schedule (warp-uniform): consecutive output rows use source rows (sy0, sy0+1) with sy0 advancing by 1
  // or 2 (the scale is between 1 and 2), so the H row of sy0 is either the previous output row's second row
  // or a new one; the second row is always new.  HA/HB swap roles instead of moving.
  uint32_t HA[4], HB[4];
  int sy_b = -1 << 20;  // source row held in the buffer that was the second row of the previous output row
  const int dy1 = min(dy0 + RS_ROWS, dh);
#if RS_DP2A == 2
#define RS_VERT(P, Q, ty) rs_vert3(P, Q, (uint32_t)ty.c0, (uint32_t)ty.c1)
#else
#define RS_VERT(P, Q, ty) rs_vert2(P, Q, (uint32_t)ty.c0 << 12, (uint32_t)ty.c1 << 12)
#endif
#if RS_DP2A
  // two output rows per trip; the buffers swap roles (P = holds the previous second row, Q = receives the new one)
#define RS_STEP(P, Q, DY)                                                                         \
  do {                                                                                            \
    const OrbxTap ty = yt[DY];                                                                    \
    const int sy0 = min(max(ty.ofs, 0), sh - 1), sy1 = min(max(ty.ofs + 1, 0), sh - 1);           \
    if (sy0 != sy_b) rs_row_smem2(base + sy0 * RT_SPITCH, T, P);                                  \
    rs_row_smem2(base + sy1 * RT_SPITCH, T, Q);                                                   \
    sy_b = sy1;                                                                                   \
    *reinterpret_cast<uint32_t*>(out) = RS_VERT(P, Q, ty);                                        \
    out += dpitch;                                                                                \
  } while (0)
#pragma unroll 1
  for (int dy = dy0; dy < dy1; dy += 2) {
    RS_STEP(HA, HB, dy);
    if (dy + 1 < dy1) RS_STEP(HB, HA, dy + 1);
  }
#undef RS_STEP
#else
#pragma unroll 1
  for (int dy = dy0; dy < dy1; dy++, out += dpitch) {
    const OrbxTap ty = yt[dy];
    const int sy0 = min(max(ty.ofs, 0), sh - 1), sy1 = min(max(ty.ofs + 1, 0), sh - 1);
    if (sy0 == sy_b) {
#pragma unroll
      for (int i = 0; i < 4; i++) HA[i] = HB[i];
    } else {
      rs_row_smem(base + sy0 * RT_SPITCH, T, HA);
    }
    rs_row_smem(base + sy1 * RT_SPITCH, T, HB);
    sy_b = sy1;
    // rows are 128-byte aligned and padded, so the full word store stays inside the row
    *reinterpret_cast<uint32_t*>(out) = rs_vert(HA, HB, (uint32_t)ty.c0, (uint32_t)ty.c1);
  }
#endif
}

// Maps of the arena levels that serve as resize sources (levels 1 .. nlevels-2), over all max_batch frame slots.
void orbx_resize_maps_init(const DevPlan& P, const uint8_t* arena, int max_batch, OrbxResizeMaps* M) {
  memset(M, 0, sizeof(*M));
  for (int l = 1; l + 1 < P.nlevels; l++) {
    const DevLevel& S = P.lv[l];
    const DevLevel& D = P.lv[l + 1];
    if (!D.tile_resize || D.tile_rows < 1 || D.tile_rows > RT_SROWS) continue;
    if (orbx_encode_tmap3(&M->m[l], arena + S.offset, S.w, S.h, max_batch, S.pitch, P.arena_bytes, RT_SPITCH, D.tile_rows))
      M->mask |= 1u << l;
  }
}

cudaError_t orbx_launch_pyramid(const DevPlan& P, const OrbxResizeMaps& maps, Level0 l0, uint8_t* arena,
                                const OrbxTap* xtaps, const OrbxTap* ytaps, int batch, int f0, cudaStream_t st) {
  for (int l = 1; l < P.nlevels; l++) {
    const DevLevel& D = P.lv[l];
    const DevLevel& S = P.lv[l - 1];
    const uint8_t* src;
    int spitch;
    long long sframe;
    if (l == 1) { src = l0.base; spitch = l0.pitch; sframe = l0.frame_stride; }
    else { src = arena + S.offset; spitch = S.pitch; sframe = P.arena_bytes; }
    const bool walk = D.fast_resize && (spitch & 3) == 0 && (sframe & 3) == 0 && ((uintptr_t)src & 3) == 0;
    const bool tiled = walk && D.tile_resize && (spitch & 15) == 0 && (sframe & 15) == 0 && ((uintptr_t)src & 15) == 0;
    if (tiled) {
      dim3 grd((D.w + 127) / 128, (D.h + RS_WARPS * RS_ROWS - 1) / (RS_WARPS * RS_ROWS), batch);
      static const bool no_tma = [] { const char* e = getenv("ORBX_RESIZE_TMA"); return e && e[0] == '0'; }();
      OrbxTmap map = maps.m[l - 1];
      bool tma = !no_tma && D.tile_rows >= 1 && D.tile_rows <= RT_SROWS;
      int zf0 = f0;
      if (tma && l == 1) {  // level 0 is the caller's buffer: its map is encoded per call (frame 0 of the map = l0.base)
        tma = orbx_encode_tmap3(&map, src, S.w, S.h, batch, spitch, sframe ? sframe : 16, RT_SPITCH, D.tile_rows);
        zf0 = 0;
      } else if (tma) {
        tma = (maps.mask >> (l - 1)) & 1u;
      }
      if (tma)
        k_resize_tile<true><<<grd, RS_WARPS * 32, 0, st>>>(map, D.tile_rows, zf0, src, spitch, sframe, S.w, S.h, arena + D.offset,
                                                           D.pitch, P.arena_bytes, D.w, D.h, xtaps + D.xtab, ytaps + D.ytab);
      else
        k_resize_tile<false><<<grd, RS_WARPS * 32, 0, st>>>(map, 0, 0, src, spitch, sframe, S.w, S.h, arena + D.offset,
                                                            D.pitch, P.arena_bytes, D.w, D.h, xtaps + D.xtab, ytaps + D.ytab);
    } else if (walk) {
      dim3 grd((D.w + 127) / 128, (D.h + RS_WARPS * RS_ROWS - 1) / (RS_WARPS * RS_ROWS), batch);
      // Arena rows are padded (and the arena has slack after its last row), so only a caller-owned
      // level 0 whose pitch leaves < 12 spare bytes needs the clamped word offsets.
      if (l == 1 && spitch - S.w < 12)
        k_resize_walk<true><<<grd, RS_WARPS * 32, 0, st>>>(src, spitch, sframe, S.w, S.h, arena + D.offset, D.pitch,
                                                           P.arena_bytes, D.w, D.h, xtaps + D.xtab, ytaps + D.ytab);
      else
        k_resize_walk<false><<<grd, RS_WARPS * 32, 0, st>>>(src, spitch, sframe, S.w, S.h, arena + D.offset,
                                                            D.pitch, P.arena_bytes, D.w, D.h, xtaps + D.xtab,
                                                            ytaps + D.ytab);
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
// A caller's level 0 whose base, pitch or frame stride is not 16-byte aligned (e.g. tightly packed 1241-px
// KITTI rows) is copied once into the arena's level-0 slot (pitch a multiple of 128): every later kernel
// then takes its aligned path (TMA boxes in k_fast_strip, bulk copies in k_resize_tile).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_repack_level0(const uint8_t* __restrict__ src, long long sframe, int spitch, uint8_t* __restrict__ dst,
                long long dframe, int dpitch, int w, int h) {
  const int y = blockIdx.y, frame = blockIdx.z;
  const int x4 = (blockIdx.x * 256 + threadIdx.x) * 4;
  if (x4 >= w) return;
  const uint8_t* g = src + (long long)frame * sframe + (long long)y * spitch + x4;
  uint32_t v = 0;
#pragma unroll
  for (int j = 0; j < 4; j++)
    if (x4 + j < w) v |= (uint32_t)g[j] << (8 * j);
  *reinterpret_cast<uint32_t*>(dst + (long long)frame * dframe + (long long)y * dpitch + x4) = v;
}

cudaError_t orbx_launch_repack_level0(const DevPlan& P, Level0 l0, uint8_t* arena, int batch, cudaStream_t st) {
  const DevLevel& L = P.lv[0];
  dim3 grd((L.w + 1023) / 1024, L.h, batch);
  k_repack_level0<<<grd, 256, 0, st>>>(l0.base, l0.frame_stride, l0.pitch, arena + L.offset, P.arena_bytes, L.pitch, L.w, L.h);
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

