// match.cu — 256-bit Hamming candidate scoring behind ORBmatcher's Search* family and
// Frame::ComputeStereoMatches (src/ORBmatcher.cc, src/Frame.cc), one warp per query.
//
// All the reference loops have the shape
//     for cand in <ordered candidate list>: d = DescriptorDistance(q, cand)
//         if d < best: second=best; best=d ... else if d < second: second=d      (first-min)
//     or  if d <= best: best=d                                                   (last-min)
// which is order dependent only through ties.  Tagging each candidate with its position in
// the reference's visiting order turns the loop into two min-reductions over the packed key
// (dist << 20 | pos): best = smallest key, second = second smallest key.
#include "../../include/orbx.h"
#include "orbx_kernels.cuh"

#include <stdlib.h>

#include "match_core.cuh"

__device__ __forceinline__ void write_result(orbm_result* out, const Top2& t, const int8_t* ref_level,
                                             int tie_rule) {
  orbm_result r;
  r.best_idx = t.i1;
  r.best_dist = t.i1 >= 0 ? (int16_t)(t.k1 >> 20) : (int16_t)256;
  r.best_level = t.i1 >= 0 ? (ref_level ? ref_level[t.i1] : (int8_t)0) : (int8_t)-1;
  if (tie_rule == ORBM_TIE_FIRST_MIN && t.i2 >= 0) {
    r.second_dist = (int16_t)(t.k2 >> 20);
    r.second_level = ref_level ? ref_level[t.i2] : (int8_t)0;
  } else {
    r.second_dist = 256;
    r.second_level = -1;
  }
  r.pad = 0;
  *out = r;
}

// ---- explicit candidate lists -----------------------------------------------------------------
__global__ void __launch_bounds__(M_WARPS * 32)
k_argmin2_csr(const uint8_t* __restrict__ q, int nq, const uint8_t* __restrict__ r,
              const int32_t* __restrict__ cand_off, const int32_t* __restrict__ cand_idx,
              const int8_t* __restrict__ ref_level, int tie_rule, orbm_result* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int qi = blockIdx.x * M_WARPS + (threadIdx.x >> 5);
  if (qi >= nq) return;
  const Desc256 dq = load_desc(q + (size_t)qi * 32);
  const int c0 = cand_off[qi], c1 = cand_off[qi + 1];
  Top2 t;
  for (int c = c0 + lane; c < c1; c += 32) {
    const int idx = cand_idx[c];
    const int d = hamming(dq, load_desc(r + (size_t)idx * 32));
    t.push(make_key(d, (c - c0) & 0xFFFFF, tie_rule), idx);
  }
  warp_top2(t);
  if (lane == 0) write_result(out + qi, t, ref_level, tie_rule);
}

// ---- Frame::AssignFeaturesToGrid (Frame.cc:385-416): CSR over 64x48 cells, cell = ix*48+iy ------
__global__ void __launch_bounds__(128)
k_grid_build(const OrbxKeyPointDev* __restrict__ kps, const int32_t* __restrict__ n_arr, int kp_stride,
             float min_x, float min_y, float winv, float hinv, int32_t* __restrict__ cell_off,
             int32_t* __restrict__ cell_idx) {
  __shared__ int cnt[GRID_CELLS + 1];
  __shared__ int wsum[4];
  __shared__ int carry_s;
  const int f = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const OrbxKeyPointDev* k = kps + (size_t)f * kp_stride;
  const int n = n_arr[f];
  int32_t* off = cell_off + (size_t)f * (GRID_CELLS + 1);
  int32_t* idx = cell_idx + (size_t)f * kp_stride;
  for (int i = tid; i <= GRID_CELLS; i += 128) cnt[i] = 0;
  if (tid == 0) carry_s = 0;
  __syncthreads();
  for (int i = tid; i < n; i += 128) {
    const int c = grid_cell(k[i], min_x, min_y, winv, hinv);
    if (c >= 0) atomicAdd(&cnt[c], 1);
  }
  __syncthreads();
  for (int base = 0; base < GRID_CELLS; base += 128) {  // exclusive scan, 24 chunks
    const int v = cnt[base + tid];
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    int woff = 0;
    for (int w = 0; w < warp; w++) woff += wsum[w];
    const int carry = carry_s;
    const int ex = carry + woff + inc - v;
    cnt[base + tid] = ex;  // becomes the write cursor
    off[base + tid] = ex;
    __syncthreads();
    if (tid == 127) carry_s = carry + woff + inc;
    __syncthreads();
  }
  if (tid == 0) off[GRID_CELLS] = carry_s;
  __syncthreads();
  // ordered fill: one warp walks the keypoints in index order so that every cell lists its
  // members in insertion order, exactly like the push_back loop of the reference
  if (warp == 0) {
    for (int base = 0; base < n; base += 32) {
      const int i = base + lane;
      const int c = i < n ? grid_cell(k[i], min_x, min_y, winv, hinv) : -1;
      const unsigned peers = __match_any_sync(0xffffffffu, c);
      const int rank = __popc(peers & ((1u << lane) - 1));
      int basepos = 0;
      if (c >= 0 && rank == 0) basepos = atomicAdd(&cnt[c], __popc(peers));
      basepos = __shfl_sync(0xffffffffu, basepos, __ffs(peers) - 1);
      if (c >= 0) idx[basepos + rank] = i;
      __syncwarp();
    }
  }
}

// ---- fused GetFeaturesInArea + scoring --------------------------------------------------------
__global__ void __launch_bounds__(M_WARPS * 32)
k_search_area(const OrbxKeyPointDev* __restrict__ kps, const uint8_t* __restrict__ desc,
              const int32_t* __restrict__ n_arr, int kp_stride, const int32_t* __restrict__ cell_off,
              const int32_t* __restrict__ cell_idx, float min_x, float min_y, float winv, float hinv,
              const float* __restrict__ qx, const float* __restrict__ qy, const float* __restrict__ qr,
              const int8_t* __restrict__ qmin, const int8_t* __restrict__ qmax,
              const uint8_t* __restrict__ qdesc, const int32_t* __restrict__ nq_arr, int q_stride, int tie_rule,
              orbm_result* __restrict__ out) {
  const int lane = threadIdx.x & 31, f = blockIdx.y;
  const int qi = blockIdx.x * M_WARPS + (threadIdx.x >> 5);
  if (qi >= nq_arr[f]) return;
  const size_t qo = (size_t)f * q_stride + qi;
  const float x = qx[qo], y = qy[qo], r = qr[qo];
  const int minLevel = qmin[qo], maxLevel = qmax[qo];
  const OrbxKeyPointDev* k = kps + (size_t)f * kp_stride;
  const uint8_t* dsc = desc + (size_t)f * kp_stride * 32;
  const int32_t* off = cell_off + (size_t)f * (GRID_CELLS + 1);
  const int32_t* idx = cell_idx + (size_t)f * kp_stride;
  Top2 t;
  // Frame.cc:665-687
  const int nMinCellX = max(0, (int)floorf(f_mul(f_sub(f_sub(x, min_x), r), winv)));
  const int nMaxCellX = min(FRAME_GRID_COLS - 1, (int)ceilf(f_mul(f_add(f_sub(x, min_x), r), winv)));
  const int nMinCellY = max(0, (int)floorf(f_mul(f_sub(f_sub(y, min_y), r), hinv)));
  const int nMaxCellY = min(FRAME_GRID_ROWS - 1, (int)ceilf(f_mul(f_add(f_sub(y, min_y), r), hinv)));
  const bool ok = nMinCellX < FRAME_GRID_COLS && nMaxCellX >= 0 && nMinCellY < FRAME_GRID_ROWS && nMaxCellY >= 0;
  const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
  if (ok) {
    const Desc256 dq = load_desc(qdesc + qo * 32);
    int seq = 0;
    for (int ix = nMinCellX; ix <= nMaxCellX; ix++) {
      // cells (ix, nMinCellY..nMaxCellY) are adjacent in the CSR: one contiguous run
      const int e0 = off[ix * FRAME_GRID_ROWS + nMinCellY], e1 = off[ix * FRAME_GRID_ROWS + nMaxCellY + 1];
      for (int e = e0 + lane; e < e1; e += 32) {
        const int j = idx[e];
        const OrbxKeyPointDev kp = k[j];
        bool pass = true;
        if (bCheckLevels) {
          if (kp.octave < minLevel) pass = false;
          if (maxLevel >= 0 && kp.octave > maxLevel) pass = false;
        }
        const float distx = f_sub(kp.x, x), disty = f_sub(kp.y, y);
        if (!(fabsf(distx) < r && fabsf(disty) < r)) pass = false;
        if (pass) {
          const int d = hamming(dq, load_desc(dsc + (size_t)j * 32));
          t.push(make_key(d, (seq + e - e0) & 0xFFFFF, tie_rule), j);
        }
      }
      seq += e1 - e0;
    }
  }
  warp_top2(t);
  if (lane == 0) {
    orbm_result res;
    res.best_idx = t.i1;
    res.best_dist = t.i1 >= 0 ? (int16_t)(t.k1 >> 20) : (int16_t)256;
    res.best_level = t.i1 >= 0 ? (int8_t)k[t.i1].octave : (int8_t)-1;
    const bool has2 = tie_rule == ORBM_TIE_FIRST_MIN && t.i2 >= 0;
    res.second_dist = has2 ? (int16_t)(t.k2 >> 20) : (int16_t)256;
    res.second_level = has2 ? (int8_t)k[t.i2].octave : (int8_t)-1;
    res.pad = 0;
    out[qo] = res;
  }
}

// Same result, one THREAD per query: a window holds only a few dozen candidates spread over several
// short CSR runs, which leaves a warp-per-query mostly idle.  Here a lane walks its own query's runs
// (best/second are plain sequential updates on the packed keys) and the 32 lanes of a warp work on
// 32 neighbouring queries; queries are level-major, so the lanes of a warp have similar trip counts.
#define SA_THREADS 128
__global__ void __launch_bounds__(SA_THREADS)
k_search_area_tq(const OrbxKeyPointDev* __restrict__ kps, const uint8_t* __restrict__ desc,
                 const int32_t* __restrict__ n_arr, int kp_stride, const int32_t* __restrict__ cell_off,
                 const int32_t* __restrict__ cell_idx, float min_x, float min_y, float winv, float hinv,
                 const float* __restrict__ qx, const float* __restrict__ qy, const float* __restrict__ qr,
                 const int8_t* __restrict__ qmin, const int8_t* __restrict__ qmax,
                 const uint8_t* __restrict__ qdesc, const int32_t* __restrict__ nq_arr, int q_stride, int tie_rule,
                 orbm_result* __restrict__ out) {
  const int f = blockIdx.y;
  const int qi = blockIdx.x * SA_THREADS + threadIdx.x;
  if (qi >= nq_arr[f]) return;
  const size_t qo = (size_t)f * q_stride + qi;
  const float x = qx[qo], y = qy[qo], r = qr[qo];
  const int minLevel = qmin[qo], maxLevel = qmax[qo];
  const OrbxKeyPointDev* k = kps + (size_t)f * kp_stride;
  const uint8_t* dsc = desc + (size_t)f * kp_stride * 32;
  const int32_t* off = cell_off + (size_t)f * (GRID_CELLS + 1);
  const int32_t* idx = cell_idx + (size_t)f * kp_stride;
  Top2 t;
  // Frame.cc:665-687
  const int nMinCellX = max(0, (int)floorf(f_mul(f_sub(f_sub(x, min_x), r), winv)));
  const int nMaxCellX = min(FRAME_GRID_COLS - 1, (int)ceilf(f_mul(f_add(f_sub(x, min_x), r), winv)));
  const int nMinCellY = max(0, (int)floorf(f_mul(f_sub(f_sub(y, min_y), r), hinv)));
  const int nMaxCellY = min(FRAME_GRID_ROWS - 1, (int)ceilf(f_mul(f_add(f_sub(y, min_y), r), hinv)));
  const bool ok = nMinCellX < FRAME_GRID_COLS && nMaxCellX >= 0 && nMinCellY < FRAME_GRID_ROWS && nMaxCellY >= 0;
  const int lo = minLevel > 0 ? minLevel : -128;     // bCheckLevels folded into a closed band
  const int hi = maxLevel >= 0 ? maxLevel : 127;
  if (ok) {
    const Desc256 dq = load_desc(qdesc + qo * 32);
    int seq = 0, ix = nMinCellX, e = 0, e1 = 0;
    while (true) {
      if (e >= e1) {  // next column of cells: (ix, nMinCellY..nMaxCellY) is one contiguous CSR run
        if (ix > nMaxCellX) break;
        e = off[ix * FRAME_GRID_ROWS + nMinCellY];
        e1 = off[ix * FRAME_GRID_ROWS + nMaxCellY + 1];
        ix++;
        continue;
      }
      if (e + 1 < e1) {
        // two candidates per trip: both index / keypoint / descriptor fetches are in flight before the first distance is
        // needed (the walk is a chain of dependent loads, idx -> keypoint -> descriptor; the kernel is latency-bound):
        // match stage 0.235 -> 0.217 ms at C1.  Three / four per trip cost registers and lose (0.247 / 0.300 ms).
        const int j0 = idx[e], j1 = idx[e + 1];
        const float kx0 = k[j0].x, ky0 = k[j0].y, kx1 = k[j1].x, ky1 = k[j1].y;
        const int oct0 = k[j0].octave, oct1 = k[j1].octave;
        const bool p0 = oct0 >= lo && oct0 <= hi && fabsf(f_sub(kx0, x)) < r && fabsf(f_sub(ky0, y)) < r;
        const bool p1 = oct1 >= lo && oct1 <= hi && fabsf(f_sub(kx1, x)) < r && fabsf(f_sub(ky1, y)) < r;
        Desc256 d0 = dq, d1 = dq;
        if (p0) d0 = load_desc(dsc + (size_t)j0 * 32);
        if (p1) d1 = load_desc(dsc + (size_t)j1 * 32);
        if (p0) t.push(make_key(hamming(dq, d0), seq & 0xFFFFF, tie_rule), j0);
        if (p1) t.push(make_key(hamming(dq, d1), (seq + 1) & 0xFFFFF, tie_rule), j1);
        seq += 2;
        e += 2;
        continue;
      }
      const int j = idx[e];
      const float kx = k[j].x, ky = k[j].y;
      const int oct = k[j].octave;
      if (oct >= lo && oct <= hi && fabsf(f_sub(kx, x)) < r && fabsf(f_sub(ky, y)) < r) {
        const int d = hamming(dq, load_desc(dsc + (size_t)j * 32));
        t.push(make_key(d, seq & 0xFFFFF, tie_rule), j);
      }
      seq++;
      e++;
    }
  }
  orbm_result res;
  res.best_idx = t.i1;
  res.best_dist = t.i1 >= 0 ? (int16_t)(t.k1 >> 20) : (int16_t)256;
  res.best_level = t.i1 >= 0 ? (int8_t)k[t.i1].octave : (int8_t)-1;
  const bool has2 = tie_rule == ORBM_TIE_FIRST_MIN && t.i2 >= 0;
  res.second_dist = has2 ? (int16_t)(t.k2 >> 20) : (int16_t)256;
  res.second_level = has2 ? (int8_t)k[t.i2].octave : (int8_t)-1;
  res.pad = 0;
  out[qo] = res;
}

// (Four lanes per query — cell columns dealt to a quad, shuffle merge — was measured at 0.44 ms against 0.23 ms for
// the thread-per-query kernel on 512 C1 frames: the extra threads only replicate the column bookkeeping.)
// ---- candidate scoring of ORBmatcher::Fuse (ORBmatcher.cc:1148-1338, :1340-1455) ---------------------
// KeyFrame::GetFeaturesInArea (KeyFrame.cc:704-748, no octave band) + the in-loop gates: octave in
// [l-1, l], chi-square reprojection gate (5.99 mono, 7.8 with a stereo coordinate), first minimum.
struct InvSigmaTab { float s[ORBX_MAX_LEVELS]; };

__global__ void __launch_bounds__(SA_THREADS)
k_fuse_tq(const OrbxKeyPointDev* __restrict__ kps, const uint8_t* __restrict__ desc,
          const int32_t* __restrict__ n_arr, int kp_stride, const int32_t* __restrict__ cell_off,
          const int32_t* __restrict__ cell_idx, float min_x, float min_y, float winv, float hinv,
          const float* __restrict__ u_right, InvSigmaTab inv_sigma2, const float* __restrict__ qx,
          const float* __restrict__ qy, const float* __restrict__ qr, const float* __restrict__ q_ur,
          const int8_t* __restrict__ qlevel, const uint8_t* __restrict__ qdesc,
          const int32_t* __restrict__ nq_arr, int q_stride, orbm_result* __restrict__ out) {
  const int f = blockIdx.y;
  const int qi = blockIdx.x * SA_THREADS + threadIdx.x;
  if (qi >= nq_arr[f]) return;
  const size_t qo = (size_t)f * q_stride + qi;
  const float x = qx[qo], y = qy[qo], r = qr[qo];
  const float ur = q_ur ? q_ur[qo] : 0.0f;
  const int level = qlevel[qo];
  const OrbxKeyPointDev* k = kps + (size_t)f * kp_stride;
  const uint8_t* dsc = desc + (size_t)f * kp_stride * 32;
  const float* urk = u_right ? u_right + (size_t)f * kp_stride : nullptr;
  const int32_t* off = cell_off + (size_t)f * (GRID_CELLS + 1);
  const int32_t* idx = cell_idx + (size_t)f * kp_stride;
  const int nMinCellX = max(0, (int)floorf(f_mul(f_sub(f_sub(x, min_x), r), winv)));
  const int nMaxCellX = min(FRAME_GRID_COLS - 1, (int)ceilf(f_mul(f_add(f_sub(x, min_x), r), winv)));
  const int nMinCellY = max(0, (int)floorf(f_mul(f_sub(f_sub(y, min_y), r), hinv)));
  const int nMaxCellY = min(FRAME_GRID_ROWS - 1, (int)ceilf(f_mul(f_add(f_sub(y, min_y), r), hinv)));
  const bool ok = nMinCellX < FRAME_GRID_COLS && nMaxCellX >= 0 && nMinCellY < FRAME_GRID_ROWS && nMaxCellY >= 0;
  uint32_t best = KEY_NONE;
  int best_i = -1;
  if (ok) {
    const Desc256 dq = load_desc(qdesc + qo * 32);
    int seq = 0, ix = nMinCellX, e = 0, e1 = 0;
    while (true) {
      if (e >= e1) {
        if (ix > nMaxCellX) break;
        e = off[ix * FRAME_GRID_ROWS + nMinCellY];
        e1 = off[ix * FRAME_GRID_ROWS + nMaxCellY + 1];
        ix++;
        continue;
      }
      const int j = idx[e];
      e++;
      const float kx = k[j].x, ky = k[j].y;
      const int oct = k[j].octave;
      if (!(fabsf(f_sub(kx, x)) < r && fabsf(f_sub(ky, y)) < r)) continue;  // GetFeaturesInArea
      const int pos = seq++;
      if (oct < level - 1 || oct > level) continue;                            // :1269
      const float ex = f_sub(x, kx), ey = f_sub(y, ky);
      float e2 = f_add(f_mul(ex, ex), f_mul(ey, ey));
      const float kr = urk ? urk[j] : -1.0f;
      double lim = 5.99;
      if (kr >= 0.0f) {                                                        // :1272-1285
        const float er = f_sub(ur, kr);
        e2 = f_add(e2, f_mul(er, er));
        lim = 7.8;
      }
      if ((double)f_mul(e2, inv_sigma2.s[oct]) > lim) continue;
      const int d = hamming(dq, load_desc(dsc + (size_t)j * 32));
      const uint32_t key = make_key(d, pos & 0xFFFFF, ORBM_TIE_FIRST_MIN);
      if (key < best) { best = key; best_i = j; }
    }
  }
  orbm_result res;
  res.best_idx = best_i;
  res.best_dist = best_i >= 0 ? (int16_t)(best >> 20) : (int16_t)256;
  res.best_level = best_i >= 0 ? (int8_t)k[best_i].octave : (int8_t)-1;
  res.second_dist = 256; res.second_level = -1; res.pad = 0;
  out[qo] = res;
}

// ---- ORBmatcher::DescriptorDistance (ORBmatcher.cc:2058-2074) over explicit index pairs ----------------
__global__ void __launch_bounds__(256)
k_distance_pairs(const uint8_t* __restrict__ a, const uint8_t* __restrict__ b, const int32_t* __restrict__ ia,
                 const int32_t* __restrict__ ib, int n, int32_t* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const size_t ra = ia ? (size_t)ia[i] : (size_t)i, rb = ib ? (size_t)ib[i] : (size_t)i;
  out[i] = hamming(load_desc(a + ra * 32), load_desc(b + rb * 32));
}

// ---- Hamming stage of Frame::ComputeStereoMatches (Frame.cc:811-894) -------------------------------
struct ScaleTab { float s[ORBX_MAX_LEVELS]; };

__global__ void __launch_bounds__(M_WARPS * 32)
k_stereo_rowband(const OrbxKeyPointDev* __restrict__ kl, const uint8_t* __restrict__ dl,
                 const int32_t* __restrict__ nl_arr, const OrbxKeyPointDev* __restrict__ kr,
                 const uint8_t* __restrict__ dr, const int32_t* __restrict__ nr_arr, int kp_stride,
                 ScaleTab scales, float max_d, int32_t* __restrict__ best_idx, int32_t* __restrict__ best_dist) {
  const int lane = threadIdx.x & 31, f = blockIdx.y;
  const int iL = blockIdx.x * M_WARPS + (threadIdx.x >> 5);
  if (iL >= nl_arr[f]) return;
  const size_t base = (size_t)f * kp_stride;
  const OrbxKeyPointDev kpL = kl[base + iL];
  const int nR = nr_arr[f];
  const int rowL = (int)kpL.y;  // vRowIndices[vL] (:852)
  const float minU = f_sub(kpL.x, max_d), maxU = kpL.x;  // minD = 0
  const Desc256 dq = load_desc(dl + (base + iL) * 32);
  uint32_t bestk = KEY_NONE;
  if (!(maxU < 0)) {
    for (int iR = lane; iR < nR; iR += 32) {
      const OrbxKeyPointDev kpR = kr[base + iR];
      // membership of iR in vRowIndices[rowL] (:828-838)
      const float rr = f_mul(2.0f, scales.s[kpR.octave]);
      const int maxr = (int)ceilf(f_add(kpR.y, rr)), minr = (int)floorf(f_sub(kpR.y, rr));
      if (rowL < minr || rowL > maxr) continue;
      if (kpR.octave < kpL.octave - 1 || kpR.octave > kpL.octave + 1) continue;
      if (!(kpR.x >= minU && kpR.x <= maxU)) continue;
      const int d = hamming(dq, load_desc(dr + (base + iR) * 32));
      const uint32_t k = ((uint32_t)d << 20) | (uint32_t)iR;  // ascending iR = visiting order
      bestk = min(bestk, k);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) bestk = min(bestk, __shfl_xor_sync(0xffffffffu, bestk, o));
  if (lane == 0) {
    const int d = bestk == KEY_NONE ? 256 : (int)(bestk >> 20);
    // bestDist starts at TH_HIGH=100 and only strictly smaller distances replace it (:862,:888)
    if (d < 100) { best_idx[base + iL] = (int)(bestk & 0xFFFFF); best_dist[base + iL] = d; }
    else { best_idx[base + iL] = -1; best_dist[base + iL] = 100; }
  }
}


// Same result with the right keypoints bucketed by image row first (the reference's vRowIndices, Frame.cc:820-838,
// at 16-row granularity): one CTA = one slice of the left keypoints of one pair; it builds the buckets in shared
// memory (count, scan, fill), then each warp takes left keypoints and scans only the bucket of their row.  The
// visiting order inside a bucket is irrelevant: ties are decided by the packed key (dist << 20 | iR).
#define RB_SHIFT 4
#define RB_BINS 256          // rows < 4096
#define RB_THREADS 256
#define RB_SPAN 3            // a right keypoint's band (<= 2*2*scale+2 rows) touches at most 3 buckets up to scale 7.5
__global__ void __launch_bounds__(RB_THREADS)
k_stereo_rowband_binned(const OrbxKeyPointDev* __restrict__ kl, const uint8_t* __restrict__ dl,
                        const int32_t* __restrict__ nl_arr, const OrbxKeyPointDev* __restrict__ kr,
                        const uint8_t* __restrict__ dr, const int32_t* __restrict__ nr_arr, int kp_stride,
                        ScaleTab scales, float max_d, int32_t* __restrict__ best_idx, int32_t* __restrict__ best_dist) {
  extern __shared__ uint16_t rb_list[];  // RB_SPAN * kp_stride entries
  __shared__ int cnt[RB_BINS + 1];
  __shared__ int cur[RB_BINS];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, f = blockIdx.y;
  const size_t base = (size_t)f * kp_stride;
  const int nL = nl_arr[f], nR = nr_arr[f];
  for (int i = tid; i <= RB_BINS; i += RB_THREADS) cnt[i] = 0;
  __syncthreads();
  for (int iR = tid; iR < nR; iR += RB_THREADS) {
    const OrbxKeyPointDev kpR = kr[base + iR];
    const float rr = f_mul(2.0f, scales.s[kpR.octave]);
    const int maxr = (int)ceilf(f_add(kpR.y, rr)), minr = (int)floorf(f_sub(kpR.y, rr));
    const int b0 = max(minr, 0) >> RB_SHIFT, b1 = min(maxr >> RB_SHIFT, RB_BINS - 1);
    for (int b = b0; b <= b1 && b < b0 + RB_SPAN; b++) atomicAdd(&cnt[b], 1);
  }
  __syncthreads();
  if (warp == 0) {  // exclusive scan of 256 counts
    int carry = 0;
    for (int c = 0; c < RB_BINS; c += 32) {
      const int v = cnt[c + lane];
      int inc = v;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
      }
      cnt[c + lane] = carry + inc - v;
      cur[c + lane] = carry + inc - v;
      carry += __shfl_sync(0xffffffffu, inc, 31);
    }
    if (lane == 0) cnt[RB_BINS] = carry;
  }
  __syncthreads();
  for (int iR = tid; iR < nR; iR += RB_THREADS) {
    const OrbxKeyPointDev kpR = kr[base + iR];
    const float rr = f_mul(2.0f, scales.s[kpR.octave]);
    const int maxr = (int)ceilf(f_add(kpR.y, rr)), minr = (int)floorf(f_sub(kpR.y, rr));
    const int b0 = max(minr, 0) >> RB_SHIFT, b1 = min(maxr >> RB_SHIFT, RB_BINS - 1);
    for (int b = b0; b <= b1 && b < b0 + RB_SPAN; b++) rb_list[atomicAdd(&cur[b], 1)] = (uint16_t)iR;
  }
  __syncthreads();
  // left keypoints of this slice, one per warp at a time
  const int per = (nL + gridDim.x - 1) / gridDim.x;
  const int l0 = blockIdx.x * per, l1 = min(nL, l0 + per);
  for (int iL = l0 + warp; iL < l1; iL += RB_THREADS / 32) {
    const OrbxKeyPointDev kpL = kl[base + iL];
    const int rowL = (int)kpL.y;  // vRowIndices[vL] (:852)
    const float minU = f_sub(kpL.x, max_d), maxU = kpL.x;  // minD = 0
    uint32_t bestk = KEY_NONE;
    if (!(maxU < 0) && rowL >= 0 && (rowL >> RB_SHIFT) < RB_BINS) {
      const Desc256 dq = load_desc(dl + (base + iL) * 32);
      const int b = rowL >> RB_SHIFT;
      for (int e = cnt[b] + lane; e < cnt[b + 1]; e += 32) {
        const int iR = rb_list[e];
        const OrbxKeyPointDev kpR = kr[base + iR];
        const float rr = f_mul(2.0f, scales.s[kpR.octave]);
        const int maxr = (int)ceilf(f_add(kpR.y, rr)), minr = (int)floorf(f_sub(kpR.y, rr));
        if (rowL < minr || rowL > maxr) continue;
        if (kpR.octave < kpL.octave - 1 || kpR.octave > kpL.octave + 1) continue;
        if (!(kpR.x >= minU && kpR.x <= maxU)) continue;
        const int d = hamming(dq, load_desc(dr + (base + iR) * 32));
        bestk = min(bestk, ((uint32_t)d << 20) | (uint32_t)iR);
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) bestk = min(bestk, __shfl_xor_sync(0xffffffffu, bestk, o));
    if (lane == 0) {
      const int d = bestk == KEY_NONE ? 256 : (int)(bestk >> 20);
      if (d < 100) { best_idx[base + iL] = (int)(bestk & 0xFFFFF); best_dist[base + iL] = d; }
      else { best_idx[base + iL] = -1; best_dist[base + iL] = 100; }
    }
  }
}

// ---- window queries for SearchByProjection(Frame&, const Frame&) under a pure image shift ------------
// The reference projects last-frame map points with the current pose (host geometry, out of scope)
// and then searches radius th*scale[octave] in the octave band [o+lo, o+hi] (ORBmatcher.cc:1725-1733).
// For device-resident pipelines the projection of frame `src` into frame f is a per-frame 2-D shift.
__global__ void __launch_bounds__(256)
k_queries_shift(const OrbxKeyPointDev* __restrict__ kps, const uint8_t* __restrict__ desc,
                const int32_t* __restrict__ n_arr, int kp_stride, const int32_t* __restrict__ src_frame,
                const float* __restrict__ shift_xy, float th, ScaleTab scales, int band_lo, int band_hi,
                float* __restrict__ qx, float* __restrict__ qy, float* __restrict__ qr, int8_t* __restrict__ qmin,
                int8_t* __restrict__ qmax, uint8_t* __restrict__ qdesc, int32_t* __restrict__ nq) {
  const int f = blockIdx.y, src = src_frame[f];
  const int n = n_arr[src];
  if (blockIdx.x == 0 && threadIdx.x == 0) nq[f] = n;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const OrbxKeyPointDev kp = kps[(size_t)src * kp_stride + i];
  const size_t o = (size_t)f * kp_stride + i;
  qx[o] = f_add(kp.x, shift_xy[2 * f]);
  qy[o] = f_add(kp.y, shift_xy[2 * f + 1]);
  qr[o] = f_mul(th, scales.s[kp.octave]);
  qmin[o] = (int8_t)(kp.octave + band_lo);
  qmax[o] = (int8_t)(kp.octave + band_hi);
  const uint4* s4 = reinterpret_cast<const uint4*>(desc + ((size_t)src * kp_stride + i) * 32);
  uint4* d4 = reinterpret_cast<uint4*>(qdesc + o * 32);
  d4[0] = s4[0];
  d4[1] = s4[1];
}

// ---- cv::BFMatcher(NORM_HAMMING).knnMatch(k=2) (Frame.cc:1144) ------------------------------------
__global__ void __launch_bounds__(M_WARPS * 32)
k_knn2_dense(const uint8_t* __restrict__ q, int nq, const uint8_t* __restrict__ r, int nr,
             int32_t* __restrict__ idx2, int32_t* __restrict__ dist2) {
  const int lane = threadIdx.x & 31;
  const int qi = blockIdx.x * M_WARPS + (threadIdx.x >> 5);
  if (qi >= nq) return;
  const Desc256 dq = load_desc(q + (size_t)qi * 32);
  Top2 t;
  for (int j = lane; j < nr; j += 32) {
    const int d = hamming(dq, load_desc(r + (size_t)j * 32));
    t.push(((uint32_t)d << 20) | (uint32_t)j, j);
  }
  warp_top2(t);
  if (lane == 0) {
    idx2[2 * qi] = t.i1; idx2[2 * qi + 1] = t.i2;
    dist2[2 * qi] = t.i1 >= 0 ? (int)(t.k1 >> 20) : -1;
    dist2[2 * qi + 1] = t.i2 >= 0 ? (int)(t.k2 >> 20) : -1;
  }
}

#define CUM(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return ORBX_E_CUDA; } while (0)

extern "C" {

int orbm_argmin2_csr(const uint8_t* d_q, int nq, const uint8_t* d_r, int nr, const int32_t* d_cand_off,
                     const int32_t* d_cand_idx, const int8_t* d_ref_level, int tie_rule, orbm_result* d_out,
                     void* stream) {
  if (nq < 0 || nr < 0 || (nq > 0 && (!d_q || !d_cand_off || !d_out))) return ORBX_E_INVALID;
  if (nq == 0) return ORBX_OK;
  k_argmin2_csr<<<(nq + M_WARPS - 1) / M_WARPS, M_WARPS * 32, 0, (cudaStream_t)stream>>>(
      d_q, nq, d_r, d_cand_off, d_cand_idx, d_ref_level, tie_rule, d_out);
  CUM(cudaGetLastError());
  return ORBX_OK;
}

int orbm_grid_build(const orbx_keypoint* d_kps, const int32_t* d_n, int kp_stride, int batch, float min_x,
                    float max_x, float min_y, float max_y, int32_t* d_cell_off, int32_t* d_cell_idx, void* stream) {
  if (!d_kps || !d_n || !d_cell_off || !d_cell_idx || batch < 1 || !(max_x > min_x) || !(max_y > min_y))
    return ORBX_E_INVALID;
  // mfGridElementWidthInv / HeightInv (Frame.cc:171-172)
  const float winv = (float)FRAME_GRID_COLS / (max_x - min_x), hinv = (float)FRAME_GRID_ROWS / (max_y - min_y);
  k_grid_build<<<batch, 128, 0, (cudaStream_t)stream>>>((const OrbxKeyPointDev*)d_kps, d_n, kp_stride, min_x, min_y,
                                                        winv, hinv, d_cell_off, d_cell_idx);
  CUM(cudaGetLastError());
  return ORBX_OK;
}

int orbm_search_area(const orbx_keypoint* d_kps, const uint8_t* d_desc, const int32_t* d_n, int kp_stride,
                     const int32_t* d_cell_off, const int32_t* d_cell_idx, float min_x, float max_x, float min_y,
                     float max_y, const float* d_qx, const float* d_qy, const float* d_qr,
                     const int8_t* d_qmin_level, const int8_t* d_qmax_level, const uint8_t* d_qdesc,
                     const int32_t* d_nq, int q_stride, int batch, int tie_rule, orbm_result* d_out, void* stream) {
  if (!d_kps || !d_desc || !d_n || !d_cell_off || !d_cell_idx || !d_qx || !d_qy || !d_qr || !d_qmin_level ||
      !d_qmax_level || !d_qdesc || !d_nq || !d_out || batch < 1 || q_stride < 1 || !(max_x > min_x) || !(max_y > min_y))
    return ORBX_E_INVALID;
  const float winv = (float)FRAME_GRID_COLS / (max_x - min_x), hinv = (float)FRAME_GRID_ROWS / (max_y - min_y);
  dim3 grd((q_stride + SA_THREADS - 1) / SA_THREADS, batch);
  k_search_area_tq<<<grd, SA_THREADS, 0, (cudaStream_t)stream>>>(
      (const OrbxKeyPointDev*)d_kps, d_desc, d_n, kp_stride, d_cell_off, d_cell_idx, min_x, min_y, winv, hinv, d_qx,
      d_qy, d_qr, d_qmin_level, d_qmax_level, d_qdesc, d_nq, q_stride, tie_rule, d_out);
  CUM(cudaGetLastError());
  return ORBX_OK;
}

int orbm_descriptor_distance_pairs(const uint8_t* d_a, const uint8_t* d_b, const int32_t* d_idx_a,
                                   const int32_t* d_idx_b, int n_pairs, int32_t* d_dist, void* stream) {
  if (!d_a || !d_b || !d_dist || n_pairs < 0) return ORBX_E_INVALID;
  if (n_pairs == 0) return ORBX_OK;
  k_distance_pairs<<<(n_pairs + 255) / 256, 256, 0, (cudaStream_t)stream>>>(d_a, d_b, d_idx_a, d_idx_b, n_pairs, d_dist);
  CUM(cudaGetLastError());
  return ORBX_OK;
}

int orbm_fuse_candidates(const orbx_keypoint* d_kps, const uint8_t* d_desc, const int32_t* d_n, int kp_stride,
                         const int32_t* d_cell_off, const int32_t* d_cell_idx, float min_x, float max_x, float min_y,
                         float max_y, const float* d_u_right, const float* inv_level_sigma2_host, int nlevels,
                         const float* d_qx, const float* d_qy, const float* d_qr, const float* d_q_ur,
                         const int8_t* d_qlevel, const uint8_t* d_qdesc, const int32_t* d_nq, int q_stride,
                         int batch, orbm_result* d_out, void* stream) {
  if (!d_kps || !d_desc || !d_n || !d_cell_off || !d_cell_idx || !inv_level_sigma2_host || !d_qx || !d_qy ||
      !d_qr || !d_qlevel || !d_qdesc || !d_nq || !d_out || batch < 1 || kp_stride < 1 || q_stride < 1 ||
      nlevels < 1 || nlevels > ORBX_MAX_LEVELS || !(max_x > min_x) || !(max_y > min_y) || (d_u_right && !d_q_ur))
    return ORBX_E_INVALID;
  InvSigmaTab tab;
  for (int i = 0; i < ORBX_MAX_LEVELS; i++) tab.s[i] = i < nlevels ? inv_level_sigma2_host[i] : 0.0f;
  const float winv = (float)FRAME_GRID_COLS / (max_x - min_x), hinv = (float)FRAME_GRID_ROWS / (max_y - min_y);
  dim3 grd((q_stride + SA_THREADS - 1) / SA_THREADS, batch);
  k_fuse_tq<<<grd, SA_THREADS, 0, (cudaStream_t)stream>>>(
      (const OrbxKeyPointDev*)d_kps, d_desc, d_n, kp_stride, d_cell_off, d_cell_idx, min_x, min_y, winv, hinv,
      d_u_right, tab, d_qx, d_qy, d_qr, d_q_ur, d_qlevel, d_qdesc, d_nq, q_stride, d_out);
  CUM(cudaGetLastError());
  return ORBX_OK;
}

int orbm_stereo_rowband(const orbx_keypoint* d_kps_l, const uint8_t* d_desc_l, const int32_t* d_n_l,
                        const orbx_keypoint* d_kps_r, const uint8_t* d_desc_r, const int32_t* d_n_r, int kp_stride,
                        int batch, const float* scale_factors_host, int nlevels, float max_d, int32_t* d_best_idx_r,
                        int32_t* d_best_dist, void* stream) {
  if (!d_kps_l || !d_desc_l || !d_n_l || !d_kps_r || !d_desc_r || !d_n_r || !scale_factors_host ||
      !d_best_idx_r || !d_best_dist || batch < 1 || nlevels < 1 || nlevels > ORBX_MAX_LEVELS)
    return ORBX_E_INVALID;
  ScaleTab st;
  for (int i = 0; i < ORBX_MAX_LEVELS; i++) st.s[i] = i < nlevels ? scale_factors_host[i] : 1.0f;
  float smax = 0;
  for (int i = 0; i < nlevels; i++) smax = scale_factors_host[i] > smax ? scale_factors_host[i] : smax;
  const size_t smem = (size_t)RB_SPAN * kp_stride * sizeof(uint16_t);
  if (kp_stride < 65536 && smem <= 160 * 1024 && 4.0f * smax + 3.0f <= (float)((RB_SPAN - 1) << RB_SHIFT)) {
    // bucketed by row: a band of <= 4*scale+3 rows touches at most RB_SPAN buckets
    CUM(orbx_ensure_dyn_smem(ORBX_SMEM_ROWBAND, (const void*)k_stereo_rowband_binned, smem));
    dim3 grd(4, batch);
    k_stereo_rowband_binned<<<grd, RB_THREADS, smem, (cudaStream_t)stream>>>(
        (const OrbxKeyPointDev*)d_kps_l, d_desc_l, d_n_l, (const OrbxKeyPointDev*)d_kps_r, d_desc_r, d_n_r, kp_stride,
        st, max_d, d_best_idx_r, d_best_dist);
  } else {
    dim3 grd((kp_stride + M_WARPS - 1) / M_WARPS, batch);
    k_stereo_rowband<<<grd, M_WARPS * 32, 0, (cudaStream_t)stream>>>(
        (const OrbxKeyPointDev*)d_kps_l, d_desc_l, d_n_l, (const OrbxKeyPointDev*)d_kps_r, d_desc_r, d_n_r, kp_stride,
        st, max_d, d_best_idx_r, d_best_dist);
  }
  CUM(cudaGetLastError());
  return ORBX_OK;
}

int orbm_queries_shift(const orbx_keypoint* d_kps, const uint8_t* d_desc, const int32_t* d_n, int kp_stride,
                       int batch, const int32_t* d_src_frame, const float* d_shift_xy, float th,
                       const float* scale_factors_host, int nlevels, int band_lo, int band_hi, float* d_qx,
                       float* d_qy, float* d_qr, int8_t* d_qmin_level, int8_t* d_qmax_level, uint8_t* d_qdesc,
                       int32_t* d_nq, void* stream) {
  if (!d_kps || !d_desc || !d_n || !d_src_frame || !d_shift_xy || !scale_factors_host || !d_qx || !d_qy ||
      !d_qr || !d_qmin_level || !d_qmax_level || !d_qdesc || !d_nq || batch < 1 || nlevels < 1 ||
      nlevels > ORBX_MAX_LEVELS)
    return ORBX_E_INVALID;
  ScaleTab st;
  for (int i = 0; i < ORBX_MAX_LEVELS; i++) st.s[i] = i < nlevels ? scale_factors_host[i] : 1.0f;
  dim3 grd((kp_stride + 255) / 256, batch);
  k_queries_shift<<<grd, 256, 0, (cudaStream_t)stream>>>((const OrbxKeyPointDev*)d_kps, d_desc, d_n, kp_stride,
                                                         d_src_frame, d_shift_xy, th, st, band_lo, band_hi, d_qx, d_qy,
                                                         d_qr, d_qmin_level, d_qmax_level, d_qdesc, d_nq);
  CUM(cudaGetLastError());
  return ORBX_OK;
}

int orbm_knn2_dense(const uint8_t* d_q, int nq, const uint8_t* d_r, int nr, int32_t* d_idx2, int32_t* d_dist2,
                    void* stream) {
  if (nq < 0 || nr < 0 || (nq > 0 && (!d_q || !d_idx2 || !d_dist2))) return ORBX_E_INVALID;
  if (nq == 0) return ORBX_OK;
  k_knn2_dense<<<(nq + M_WARPS - 1) / M_WARPS, M_WARPS * 32, 0, (cudaStream_t)stream>>>(d_q, nq, d_r, nr, d_idx2,
                                                                                       d_dist2);
  CUM(cudaGetLastError());
  return ORBX_OK;
}

}  // extern "C"
