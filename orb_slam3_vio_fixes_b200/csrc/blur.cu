// blur.cu — GaussianBlur(level, 7x7, sigma 2, BORDER_REFLECT_101) of every pyramid level
// (src/ORBextractor.cc:1131-1133) for sm_100a: the "level blur" mode of the descriptor stage.
//
// cv::GaussianBlur on CV_8U runs a separable fixed-point filter with taps 18,34,48,56,48,34,18 (/256)
// and a single rounding: out = (sum_ij k_i k_j I + 2^15) >> 16.  Everything is exact integer work, so
// the two passes may run in either order.
//
// One CTA = one 128 x 32 output tile of one level of one frame:
//   1. the 38 source rows (3-px halo) are staged in shared memory with 32-bit coalesced loads; rows
//      outside the level are the reflected rows, columns outside are patched up in place;
//   2. row pass on bytes, 4 pixels per task with two packed 16-bit lanes per register (row sums
//      <= 255*256): the 8 byte pairs (s_k, s_k+2) a quad needs come from funnel shifts + PRMT;
//   3. column pass: a thread owns two pixel columns and walks 16 rows down with the last 7 row sums
//      in registers (24-bit accumulators), then pairs up with its neighbour to store 32-bit words.
#include "orbx_kernels.cuh"

#define BL_TW 128
#define BL_TH 32
#define BL_ROWS (BL_TH + 6)
#define BL_SW 36   // staged words per row: bytes 0..143, image column x at byte (x - x0 + 4)
#define BL_QUADS (BL_TW / 4)
#define BL_THREADS 128

__device__ __forceinline__ uint32_t even_bytes(uint32_t x) { return __byte_perm(x, 0, 0x4240); }  // (b0, b2)
__device__ __forceinline__ uint32_t odd_bytes(uint32_t x) { return __byte_perm(x, 0, 0x4341); }   // (b1, b3)

__global__ void __launch_bounds__(BL_THREADS)
k_blur_levels(const __grid_constant__ DevPlan P, Level0 l0, const uint8_t* __restrict__ arena,
              uint8_t* __restrict__ blur) {
  __shared__ __align__(16) uint32_t S[BL_ROWS * BL_SW];
  __shared__ __align__(16) uint2 T[BL_ROWS * BL_QUADS];
  const int tid = threadIdx.x, frame = blockIdx.y;
  int tile = blockIdx.x, level = 0;
#pragma unroll 1
  for (int l = 1; l < P.nlevels; l++)
    if (tile >= P.lv[l].btile_base) level = l;
  const DevLevel& L = P.lv[level];
  tile -= L.btile_base;
  const int ty = tile / L.btiles_x, tx = tile - ty * L.btiles_x;
  const int x0 = tx * BL_TW, y0 = ty * BL_TH;
  const int w = L.w, h = L.h;
  int pitch;
  const uint8_t* img;
  if (level == 0) { pitch = l0.pitch; img = l0.base + (long long)frame * l0.frame_stride; }
  else { pitch = L.pitch; img = arena + (long long)frame * P.arena_bytes + L.offset; }
  const int nrows = min(BL_TH, h - y0) + 6;  // staged rows actually needed

  // 1. stage: smem row r = level row reflect(y0 - 3 + r); smem byte c = level column x0 - 4 + c
  if ((pitch & 3) == 0 && ((uintptr_t)img & 3) == 0) {
    const int wlo = x0 == 0 ? 1 : 0;                       // word 0 would start at column -4
    const int whi = min(BL_SW, (pitch - (x0 - 4)) >> 2);   // stay inside the row's pitch
    const int nw = whi - wlo;
    const uint32_t rc = ((1u << 20) + nw - 1) / nw;
    for (int i = tid; i < nw * nrows; i += BL_THREADS) {
      const int r = (int)(((uint32_t)i * rc) >> 20), c = wlo + i - r * nw;
      const int sy = orbx_reflect101(y0 - 3 + r, h);
      S[r * BL_SW + c] = __ldg(reinterpret_cast<const uint32_t*>(img + (long long)sy * pitch + (x0 - 4)) + c);
    }
  } else {
    uint8_t* sb = reinterpret_cast<uint8_t*>(S);
    for (int i = tid; i < nrows * (BL_SW * 4); i += BL_THREADS) {
      const int r = i / (BL_SW * 4), c = i - r * (BL_SW * 4);
      const int x = x0 - 4 + c;
      if (x >= 0 && x < w) sb[i] = img[(long long)orbx_reflect101(y0 - 3 + r, h) * pitch + x];
    }
  }
  __syncthreads();
  // columns outside the level: BORDER_REFLECT_101, x = -k -> k and x = w-1+k -> w-1-k (k = 1..3)
  {
    uint8_t* sb = reinterpret_cast<uint8_t*>(S);
    const int cw = w - x0 + 4;  // smem byte of column w
    for (int i = tid; i < nrows * 6; i += BL_THREADS) {
      const int r = i / 6, k = i - r * 6;
      if (k < 3) {
        if (x0 == 0) sb[r * (BL_SW * 4) + 3 - k] = sb[r * (BL_SW * 4) + 5 + k];
      } else if (cw + (k - 3) < BL_SW * 4) {
        sb[r * (BL_SW * 4) + cw + (k - 3)] = sb[r * (BL_SW * 4) + cw - 2 - (k - 3)];
      }
    }
  }
  __syncthreads();

  // 2. row pass: T[r][q] = packed sums of output columns x0+4q .. x0+4q+3:
  //      .x = {T0, T2}, .y = {T1, T3},  T_i = sum_k tap_k * s_{i+k},  s_k = level column x0+4q-3+k
  const int nq = min(BL_QUADS, (w - x0 + 3) >> 2);
  {
    const uint32_t rq = ((1u << 20) + nq - 1) / nq;
    for (int i = tid; i < nq * nrows; i += BL_THREADS) {
      const int r = (int)(((uint32_t)i * rq) >> 20), q = i - r * nq;
      const uint32_t* sp = S + r * BL_SW + q;  // s_0 is byte 1 of sp[0]
      const uint32_t w0 = sp[0], w1 = sp[1], w2 = sp[2];
      const uint32_t a = __funnelshift_r(w0, w1, 8);    // s0..s3
      const uint32_t b = __funnelshift_r(w1, w2, 8);    // s4..s7
      const uint32_t c = w2 >> 8;                       // s8, s9
      const uint32_t f1 = __funnelshift_r(a, b, 16);    // s2..s5
      const uint32_t f2 = __funnelshift_r(b, c, 16);    // s6..s9
      const uint32_t ea = even_bytes(a), oa = odd_bytes(a), eb = even_bytes(b), ob = odd_bytes(b);
      const uint32_t e1 = even_bytes(f1), o1 = odd_bytes(f1), e2 = even_bytes(f2), o2 = odd_bytes(f2);
      uint2 t;
      t.x = 18u * (ea + e2) + 34u * (oa + ob) + 48u * (e1 + eb) + 56u * o1;  // {T0, T2}
      t.y = 18u * (oa + o2) + 34u * (e1 + e2) + 48u * (o1 + ob) + 56u * eb;  // {T1, T3}
      T[r * BL_QUADS + q] = t;
    }
  }
  __syncthreads();

  // 3. column pass: thread = (half of the tile rows, quad, even/odd pixel pair)
  {
    const int half = tid >> 6, q = (tid >> 1) & 31, odd = tid & 1;
    const int r0 = half * (BL_TH / 2);
    const int ny = min(BL_TH / 2, h - (y0 + r0));  // output rows of this thread (may be <= 0)
    const uint32_t* tp = reinterpret_cast<const uint32_t*>(T + r0 * BL_QUADS + q) + odd;
    uint8_t* out = blur + (long long)frame * P.arena_bytes + L.offset + (long long)(y0 + r0) * L.pitch + x0 + 4 * q;
    const bool live = q < nq;
    uint32_t lo[7], hi[7];
#pragma unroll
    for (int k = 0; k < 6; k++) {
      const uint32_t v = tp[k * (BL_QUADS * 2)];
      lo[k] = v & 0xffffu; hi[k] = v >> 16;
    }
#pragma unroll
    for (int r = 0; r < BL_TH / 2; r++) {
      const uint32_t v = tp[(r + 6) * (BL_QUADS * 2)];
      lo[6] = v & 0xffffu; hi[6] = v >> 16;
      const uint32_t p0 = (18u * (lo[0] + lo[6]) + 34u * (lo[1] + lo[5]) + 48u * (lo[2] + lo[4]) + 56u * lo[3] + 32768u) >> 16;
      const uint32_t p1 = (18u * (hi[0] + hi[6]) + 34u * (hi[1] + hi[5]) + 48u * (hi[2] + hi[4]) + 56u * hi[3] + 32768u) >> 16;
      const uint32_t mine = p0 | (p1 << 16);            // even lane: pixels 0,2; odd lane: pixels 1,3
      const uint32_t other = __shfl_down_sync(0xffffffffu, mine, 1);
      if (!odd && live && r < ny) *reinterpret_cast<uint32_t*>(out + (long long)r * L.pitch) = mine | (other << 8);
#pragma unroll
      for (int k = 0; k < 6; k++) { lo[k] = lo[k + 1]; hi[k] = hi[k + 1]; }
    }
  }
}

cudaError_t orbx_launch_blur(const DevPlan& P, Level0 l0, const uint8_t* arena, uint8_t* blur, int batch,
                             cudaStream_t st) {
  dim3 grd(P.total_btiles, batch);
  k_blur_levels<<<grd, BL_THREADS, 0, st>>>(P, l0, arena, blur);
  return cudaGetLastError();
}
