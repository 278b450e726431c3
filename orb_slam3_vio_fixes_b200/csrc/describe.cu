// describe.cu — orientation (IC_Angle, src/ORBextractor.cc:76-103), the 7x7 sigma=2 Gaussian
// blur (:1131-1133) evaluated only where descriptors need it, the 256-bit rotated BRIEF
// descriptor (:107-146) and the final keypoint record (:880-890, :1143-1164), one warp per
// keypoint.
//
// The reference blurs the whole level and then samples 512 points within radius 19 of each
// keypoint.  Here each warp stages the 45x45 source neighbourhood of its keypoint (39x39 sampled
// window + 3-px blur halo, BORDER_REFLECT_101 at the level edge), runs the separable Q8 filter
// on it in shared memory — identical integer arithmetic, out = (sum + 32768) >> 16 — and samples
// from that.  No blurred level is ever written to HBM.
#include "orbx_kernels.cuh"

#define DESC_WARPS 4
#define PATCH 45        // 2*22+1
#define PATCH_R 22
#define PPITCH 52       // 13 words: rows start word aligned, lanes at stride 13 words never conflict
#define TPITCH 41       // u16 row-pass pitch (odd: spreads banks)
#define BL 39           // blurred window 2*19+1
#define BPITCH 40

__constant__ int8_t c_pattern[1024] = {
#include "orb_pattern.inc"
};

struct __align__(16) DescWarpSmem {
  uint8_t patch[PATCH * PPITCH + 12];  // source neighbourhood; reused for the blurred window
  uint16_t tmp[PATCH * TPITCH + 3];    // row-pass sums (<= 255*256)
};

__global__ void __launch_bounds__(DESC_WARPS * 32)
k_describe(const __grid_constant__ DevPlan P, Level0 l0, const uint8_t* __restrict__ arena,
           const uint32_t* __restrict__ sel, const int32_t* __restrict__ sel_cnt,
           const int32_t* __restrict__ dst, OrbxKeyPointDev* __restrict__ kps, uint8_t* __restrict__ desc,
           int capacity, int strict_fp) {
  __shared__ DescWarpSmem sm_all[DESC_WARPS];
  __shared__ __align__(4) int8_t s_pattern[1024];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 256; i += DESC_WARPS * 32)
    reinterpret_cast<uint32_t*>(s_pattern)[i] = reinterpret_cast<const uint32_t*>(c_pattern)[i];
  __syncthreads();
  const int frame = blockIdx.y;
  const int slot = blockIdx.x * DESC_WARPS + warp;
  if (slot >= P.kp_total_cap) return;
  int level = 0;
#pragma unroll 1
  for (int l = 1; l < P.nlevels; l++)
    if (slot >= P.lv[l].kp_base) level = l;
  const DevLevel& L = P.lv[level];
  const int idx = slot - L.kp_base;
  if (idx >= sel_cnt[frame * P.nlevels + level]) return;
  DescWarpSmem& sm = sm_all[warp];

  const uint32_t key = sel[(long long)frame * P.kp_total_cap + slot];
  const int kx = ORBX_KEY_X(key), ky = ORBX_KEY_Y(key);
  int pitch;
  const uint8_t* img;
  if (level == 0) { pitch = l0.pitch; img = l0.base + (long long)frame * l0.frame_stride; }
  else { pitch = L.pitch; img = arena + (long long)frame * P.arena_bytes + L.offset; }

  // ---- stage the 45x45 neighbourhood; patch(r,c) = I(ky-22+r, kx-22+c) lives at
  //      sm.patch[r*PPITCH + c + off]
  int off;
  {
    const uint8_t* row0 = img + (long long)(ky - PATCH_R) * pitch + (kx - PATCH_R);
    const bool interior = (kx - PATCH_R >= 4) && (kx + PATCH_R + 8 < L.w) && (ky - PATCH_R >= 0) &&
                          (ky + PATCH_R < L.h) && ((pitch & 3) == 0);
    if (interior) {
      off = (int)((uintptr_t)row0 & 3);
      const uint8_t* a0 = row0 - off;  // word aligned; 13 words cover off+45 <= 48 bytes
      for (int r = 0; r < PATCH; r++) {
        if (lane < 13)
          *reinterpret_cast<uint32_t*>(&sm.patch[r * PPITCH + 4 * lane]) =
              *reinterpret_cast<const uint32_t*>(a0 + (long long)r * pitch + 4 * lane);
      }
    } else {
      off = 0;
      for (int i = lane; i < PATCH * PATCH; i += 32) {
        const int r = i / PATCH, c = i - r * PATCH;
        const int sy = orbx_reflect101(ky - PATCH_R + r, L.h), sx = orbx_reflect101(kx - PATCH_R + c, L.w);
        sm.patch[r * PPITCH + c] = img[(long long)sy * pitch + sx];
      }
    }
  }
  __syncwarp();
  const uint8_t* pc = sm.patch + off;  // pc[r*PPITCH + c]

  // ---- IC_Angle: integer moments over the circular patch of radius 15
  int m01 = 0, m10 = 0;
  if (lane < 31) {
    const int v = lane - ORBX_HALF_PATCH;
    const int d = P.umax[v < 0 ? -v : v];
    const uint8_t* rowp = pc + (PATCH_R + v) * PPITCH + PATCH_R;
    int s = 0, su = 0;
    for (int u = -d; u <= d; u++) {
      const int val = rowp[u];
      s += val;
      su += u * val;
    }
    m10 = su;
    m01 = v * s;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    m01 += __shfl_xor_sync(0xffffffffu, m01, o);
    m10 += __shfl_xor_sync(0xffffffffu, m10, o);
  }
  const float angle = orbx_fast_atan2((float)m01, (float)m10);

  // ---- separable 7-tap Q8 blur on the staged patch.  Row pass: lane = patch row, sliding window.
  for (int r = lane; r < PATCH; r += 32) {
    const uint8_t* rp = pc + r * PPITCH;
    int w0 = rp[0], w1 = rp[1], w2 = rp[2], w3 = rp[3], w4 = rp[4], w5 = rp[5];
    uint16_t* tp = sm.tmp + r * TPITCH;
#pragma unroll 3
    for (int c = 0; c < BL; c++) {
      const int w6 = rp[c + 6];
      tp[c] = (uint16_t)(18 * (w0 + w6) + 34 * (w1 + w5) + 48 * (w2 + w4) + 56 * w3);
      w0 = w1; w1 = w2; w2 = w3; w3 = w4; w4 = w5; w5 = w6;
    }
  }
  __syncwarp();
  // Column pass: lane = column, sliding window down the 45 rows.  The source patch is dead now
  // (orientation and row pass are done), so the blurred 39x39 window overwrites it.
  uint8_t* blur = sm.patch;
  for (int c = lane; c < BL; c += 32) {
    const uint16_t* tp = sm.tmp + c;
    int w0 = tp[0], w1 = tp[TPITCH], w2 = tp[2 * TPITCH], w3 = tp[3 * TPITCH], w4 = tp[4 * TPITCH],
        w5 = tp[5 * TPITCH];
#pragma unroll 3
    for (int r = 0; r < BL; r++) {
      const int w6 = tp[(r + 6) * TPITCH];
      const uint32_t acc = 18u * (w0 + w6) + 34u * (w1 + w5) + 48u * (w2 + w4) + 56u * w3;
      blur[r * BPITCH + c] = (uint8_t)((acc + 32768u) >> 16);
      w0 = w1; w1 = w2; w2 = w3; w3 = w4; w4 = w5; w5 = w6;
    }
  }
  __syncwarp();

  // ---- rotated BRIEF: test t = 32*j + lane -> bit `lane` of descriptor word j
  const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
  float a, b;
  orbx_sincosf(f_mul(angle, factorPI), &b, &a);
  const uint8_t* centre = blur + 19 * BPITCH + 19;
  uint32_t word_mine = 0;
#pragma unroll
  for (int j = 0; j < 8; j++) {
    const int t = 32 * j + lane;
    const char4 pt = reinterpret_cast<const char4*>(s_pattern)[t];
    int v0, v1;
    {
      const float x = (float)pt.x, y = (float)pt.y;
      int iy, ix;
      if (strict_fp) { iy = f_round_int(f_add(f_mul(x, b), f_mul(y, a))); ix = f_round_int(f_sub(f_mul(x, a), f_mul(y, b))); }
      else { iy = f_round_int(f_fma(x, b, f_mul(y, a))); ix = f_round_int(f_fma(x, a, -f_mul(y, b))); }
      v0 = centre[iy * BPITCH + ix];
    }
    {
      const float x = (float)pt.z, y = (float)pt.w;
      int iy, ix;
      if (strict_fp) { iy = f_round_int(f_add(f_mul(x, b), f_mul(y, a))); ix = f_round_int(f_sub(f_mul(x, a), f_mul(y, b))); }
      else { iy = f_round_int(f_fma(x, b, f_mul(y, a))); ix = f_round_int(f_fma(x, a, -f_mul(y, b))); }
      v1 = centre[iy * BPITCH + ix];
    }
    const uint32_t w = __ballot_sync(0xffffffffu, v0 < v1);
    if (lane == j) word_mine = w;
  }

  // ---- emit
  const int d = dst[(long long)frame * P.kp_total_cap + slot];
  if (d < 0 || d >= capacity) return;
  const long long o = (long long)frame * capacity + d;
  if (lane < 8) reinterpret_cast<uint32_t*>(desc + o * 32)[lane] = word_mine;
  if (lane == 8) {
    OrbxKeyPointDev kp;
    kp.x = (float)kx; kp.y = (float)ky;
    if (level != 0) { kp.x = f_mul(kp.x, L.scale); kp.y = f_mul(kp.y, L.scale); }
    kp.size = L.patch_size;
    kp.angle = angle;
    kp.response = (float)ORBX_KEY_S(key);
    kp.octave = level;
    kp.class_id = -1;
    kps[o] = kp;
  }
}

cudaError_t orbx_launch_describe(const DevPlan& P, Level0 l0, const uint8_t* arena, const uint32_t* sel,
                                 const int32_t* sel_cnt, const int32_t* dst, OrbxKeyPointDev* kps,
                                 uint8_t* desc, int capacity, int strict_fp, int batch, cudaStream_t st) {
  dim3 grd((P.kp_total_cap + DESC_WARPS - 1) / DESC_WARPS, batch);
  k_describe<<<grd, DESC_WARPS * 32, 0, st>>>(P, l0, arena, sel, sel_cnt, dst, kps, desc, capacity, strict_fp);
  return cudaGetLastError();
}
