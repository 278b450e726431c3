// search.cu — ORBmatcher::SearchByProjection as a whole function (single-camera case, Nleft == -1):
//   mode ORBM_PROJ_MAPPOINTS  (Frame&, const vector<MapPoint*>&, th, ...)   src/ORBmatcher.cc:43-142
//   mode ORBM_PROJ_LASTFRAME  (Frame&, const Frame&, th, bMono)             src/ORBmatcher.cc:1676-1887
// The pose/camera projection is host geometry and arrives as per-query (u, v, radius, octave band, ...).
//
// The reference loop is sequential through one thing only: a match writes F.mvpMapPoints[best], and a
// feature that holds a MapPoint with Observations()>0 is skipped by every later query (:88-90, :1741-1743).
// Two kernels:
//   k_proj_score    one thread per query: GetFeaturesInArea fused with scoring under all the gates that
//                   are known on entry (octave band, window, features already holding a point, stereo
//                   consistency) -> the 4 best candidates (index + distance), sorted.
//   k_proj_resolve  one warp per frame commits the queries in order.  A query's result is the first (two)
//                   entries of its list that no earlier query has locked — removing any other candidate
//                   cannot change it; only when a full list runs out the warp rescans that query's window
//                   with the current lock bitmap.  The rotation histogram, ComputeThreeMaxima (:2012-2053)
//                   and the erasure pass follow in the same warp.
#include <stdio.h>

#include <mutex>

#include "match_core.cuh"

struct ProjArgs {
  const OrbxKeyPointDev* kps; const uint8_t* desc; const int32_t* n; int kp_stride;
  const int32_t* cell_off; const int32_t* cell_idx;
  float min_x, min_y, winv, hinv;
  const float* u_right; const uint8_t* feat_locked;
  const float *qx, *qy, *qr; const int8_t *qmin, *qmax; const uint8_t* qdesc; const int32_t* nq; int q_stride;
  const float* q_ur; const uint8_t* q_flags; const float* q_angle;
  int mode; float nnratio; int check_orientation; int th_dist;
  int flag_bit;  // bit of q_flags that enables the query in k_proj_score (1; 4 for the right camera of a rig)
};

struct ProjQuery {
  float x, y, r, ur;
  int lo, hi;
  int cx0, cx1, cy0, cy1;
  bool ok;
};

__device__ __forceinline__ ProjQuery proj_query(const ProjArgs& A, size_t qo) {
  ProjQuery q;
  q.x = A.qx[qo]; q.y = A.qy[qo]; q.r = A.qr[qo];
  q.ur = A.q_ur ? A.q_ur[qo] : 0.0f;
  const int minLevel = A.qmin[qo], maxLevel = A.qmax[qo];
  q.lo = minLevel > 0 ? minLevel : -128;  // bCheckLevels folded into a closed band (Frame.cc:689-705)
  q.hi = maxLevel >= 0 ? maxLevel : 127;
  // Frame.cc:665-687
  q.cx0 = max(0, (int)floorf(f_mul(f_sub(f_sub(q.x, A.min_x), q.r), A.winv)));
  q.cx1 = min(FRAME_GRID_COLS - 1, (int)ceilf(f_mul(f_add(f_sub(q.x, A.min_x), q.r), A.winv)));
  q.cy0 = max(0, (int)floorf(f_mul(f_sub(f_sub(q.y, A.min_y), q.r), A.hinv)));
  q.cy1 = min(FRAME_GRID_ROWS - 1, (int)ceilf(f_mul(f_add(f_sub(q.y, A.min_y), q.r), A.hinv)));
  q.ok = q.cx0 < FRAME_GRID_COLS && q.cx1 >= 0 && q.cy0 < FRAME_GRID_ROWS && q.cy1 >= 0;
  return q;
}

// GetFeaturesInArea membership (Frame.cc:689-712): octave band and window
__device__ __forceinline__ bool proj_gate_area(const ProjQuery& q, const OrbxKeyPointDev* k, int j) {
  const float kx = k[j].x, ky = k[j].y;
  const int oct = k[j].octave;
  return oct >= q.lo && oct <= q.hi && fabsf(f_sub(kx, q.x)) < q.r && fabsf(f_sub(ky, q.y)) < q.r;
}

// every gate of the candidate loops that does not depend on earlier queries
__device__ __forceinline__ bool proj_gate(const ProjArgs& A, const ProjQuery& q, const OrbxKeyPointDev* k,
                                          size_t fbase, int j) {
  if (!proj_gate_area(q, k, j)) return false;
  if (A.feat_locked && A.feat_locked[fbase + j]) return false;       // :88-90 / :1741-1743 (state on entry)
  if (A.u_right && A.q_ur) {                                         // :92-97 / :1745-1751
    const float ur = A.u_right[fbase + j];
    if (ur > 0 && fabsf(f_sub(q.ur, ur)) > q.r) return false;
  }
  return true;
}

// The PROJ_K best candidates of a query under the entry gates, sorted by (distance, visiting position).
// If fewer than PROJ_K passed, the list is complete (idx = -1 beyond its end).
#define PROJ_K 4
struct ProjCand { uint32_t key[PROJ_K]; int32_t idx[PROJ_K]; };

struct TopK {
  uint32_t k[PROJ_K];
  int i[PROJ_K];
  __device__ __forceinline__ TopK() {
#pragma unroll
    for (int j = 0; j < PROJ_K; j++) { k[j] = KEY_NONE; i[j] = -1; }
  }
  __device__ __forceinline__ void push(uint32_t key, int idx) {
    if (key >= k[PROJ_K - 1]) return;
    k[PROJ_K - 1] = key; i[PROJ_K - 1] = idx;
#pragma unroll
    for (int j = PROJ_K - 1; j > 0; j--)
      if (k[j] < k[j - 1]) {
        const uint32_t tk = k[j]; k[j] = k[j - 1]; k[j - 1] = tk;
        const int ti = i[j]; i[j] = i[j - 1]; i[j - 1] = ti;
      }
  }
};

#define PS_THREADS 128
__global__ void __launch_bounds__(PS_THREADS)
k_proj_score(const ProjArgs A, ProjCand* __restrict__ cand, uint8_t* __restrict__ area_any = nullptr) {
  const int f = blockIdx.y;
  const int qi = blockIdx.x * PS_THREADS + threadIdx.x;
  if (qi >= A.nq[f]) return;
  const size_t qo = (size_t)f * A.q_stride + qi;
  TopK t;
  bool any = false;  // vIndices non-empty (before the lock / stereo gates)
  if (!A.q_flags || (A.q_flags[qo] & A.flag_bit)) {
    const size_t fbase = (size_t)f * A.kp_stride;
    const OrbxKeyPointDev* k = A.kps + fbase;
    const uint8_t* dsc = A.desc + fbase * 32;
    const int32_t* off = A.cell_off + (size_t)f * (GRID_CELLS + 1);
    const int32_t* idx = A.cell_idx + fbase;
    const ProjQuery q = proj_query(A, qo);
    if (q.ok) {
      const Desc256 dq = load_desc(A.qdesc + qo * 32);
      int seq = 0, ix = q.cx0, e = 0, e1 = 0;
      while (true) {
        if (e >= e1) {  // next column of cells: (ix, cy0..cy1) is one contiguous CSR run
          if (ix > q.cx1) break;
          e = off[ix * FRAME_GRID_ROWS + q.cy0];
          e1 = off[ix * FRAME_GRID_ROWS + q.cy1 + 1];
          ix++;
          continue;
        }
        // (two candidates per trip, which helps k_search_area_tq, loses here: 0.467 vs 0.434 ms whole function — the
        //  top-4 list and the extra gates leave no registers for the second descriptor)
        const int j = idx[e];
        if (area_any && !any && proj_gate_area(q, k, j)) any = true;
        if (proj_gate(A, q, k, fbase, j)) {
          const int d = hamming(dq, load_desc(dsc + (size_t)j * 32));
          t.push(make_key(d, seq & 0xFFFFF, ORBM_TIE_FIRST_MIN), j);
        }
        seq++;
        e++;
      }
    }
  }
  if (area_any) area_any[qo] = any ? 1 : 0;
  ProjCand c;
#pragma unroll
  for (int j = 0; j < PROJ_K; j++) { c.key[j] = t.k[j]; c.idx[j] = t.i[j]; }
  cand[qo] = c;
}

// rotation-histogram bin of a match (ORBmatcher.cc:1778-1785): factor = 1.0f/HISTO_LENGTH
__device__ __forceinline__ int proj_rot_bin(float angle_last, float angle_cur) {
  float rot = f_sub(angle_last, angle_cur);
  if (rot < 0.0f) rot = f_add(rot, 360.0f);
  int bin = (int)roundf(f_mul(rot, 1.0f / 30.0f));
  if (bin == 30) bin = 0;
  return bin;
}

// One warp per frame; lane = query inside a chunk of 32 consecutive queries.  A chunk is committed in rounds:
// every pending lane checks whether its scored result is still valid (best / second best not locked so far,
// and no EARLIER pending lane is about to lock them); the clean prefix of lanes commits in parallel, the first
// unclean lane is resolved alone (window rescan with the current locks if needed), and so on.  Commit order
// matters only through the lock bitmap and through "last writer wins" on feat_query (atomicMax on the query
// index).  Dynamic shared memory: two bitmaps of (kp_stride + 31) / 32 words (locks, chunk-pending).
__global__ void __launch_bounds__(32)
k_proj_resolve(const ProjArgs A, const ProjCand* __restrict__ cand, int32_t* __restrict__ match,
               int32_t* __restrict__ feat_query, int32_t* __restrict__ nmatches) {
  extern __shared__ uint32_t s_bits[];
  __shared__ int hist[30];
  __shared__ int s_ind[3];
  const int lane = threadIdx.x, f = blockIdx.x;
  const unsigned FULL = 0xffffffffu;
  const int nwords = (A.kp_stride + 31) / 32;
  uint32_t* lockbits = s_bits;
  uint32_t* pend = s_bits + nwords;
  const size_t fbase = (size_t)f * A.kp_stride, qbase = (size_t)f * A.q_stride;
  const OrbxKeyPointDev* k = A.kps + fbase;
  const uint8_t* dsc = A.desc + fbase * 32;
  const int32_t* off = A.cell_off + (size_t)f * (GRID_CELLS + 1);
  const int32_t* idx = A.cell_idx + fbase;
  const int n = A.n[f], nq = A.nq[f];
  for (int i = lane; i < 2 * nwords; i += 32) s_bits[i] = 0;
  if (lane < 30) hist[lane] = 0;
  for (int j = lane; j < n; j += 32) feat_query[fbase + j] = -1;
  __syncwarp();
  const bool ratio_mode = A.mode == ORBM_PROJ_MAPPOINTS;
  const bool rot_check = A.mode == ORBM_PROJ_LASTFRAME && A.check_orientation;
  int accepted = 0;  // per lane, summed at the end
#define BIT(bm, j) (((bm)[(j) >> 5] >> ((j) & 31)) & 1u)
  for (int base = 0; base < nq; base += 32) {
    const int qi = base + lane;
    const size_t qo = qbase + qi;
    int i1 = -1, i2 = -1, d1 = 256, d2 = 256, fl = 0;
    ProjCand c;
#pragma unroll
    for (int j = 0; j < PROJ_K; j++) { c.key[j] = KEY_NONE; c.idx[j] = -1; }
    if (qi < nq) {
      fl = A.q_flags ? A.q_flags[qo] : 3;
      if (fl & 1) c = cand[qo];
    }
    // best / second best among this lane's list entries that are not locked yet; returns false when the
    // list cannot answer (it was full and ran out), i.e. the window has to be rescanned
    auto pick = [&]() {
      i1 = i2 = -1; d1 = d2 = 256;
#pragma unroll
      for (int j = 0; j < PROJ_K; j++) {
        const int x = c.idx[j];
        if (x < 0 || BIT(lockbits, x)) continue;
        if (i1 < 0) { i1 = x; d1 = (int)(c.key[j] >> 20); }
        else if (i2 < 0) { i2 = x; d2 = (int)(c.key[j] >> 20); }
      }
      const bool full = c.idx[PROJ_K - 1] >= 0;
      return !full || (i1 >= 0 && (!ratio_mode || i2 >= 0));
    };
    // acceptance of the scored result (re-evaluated after a rescan)
    auto accept_of = [&](int a1, int a2, int e1, int e2) {
      bool acc = a1 >= 0 && e1 <= A.th_dist;  // TH_HIGH (or ORBdist, ORBmatcher.cc:1961)
      if (acc && ratio_mode) {
        // bestLevel==bestLevel2 && bestDist > mfNNratio*bestDist2 -> rejected (:123-127); no second: levels differ
        const int lv1 = k[a1].octave, lv2 = a2 >= 0 ? k[a2].octave : -1;
        if (lv1 == lv2 && (float)e1 > f_mul(A.nnratio, (float)e2)) acc = false;
      }
      return acc;
    };
    bool acc = false;
    int result = -1;  // feature this lane's query is written to
    unsigned pending = __ballot_sync(FULL, c.idx[0] >= 0);
    while (pending) {
      const bool mine = (pending >> lane) & 1u;
      bool answered = true;
      if (mine) { answered = pick(); acc = answered && accept_of(i1, i2, d1, d2); }
      // lanes that would lock their best when committed
      const bool locker = mine && acc && (fl & 2);
      if (locker) atomicOr(&pend[i1 >> 5], 1u << (i1 & 31));
      __syncwarp();
      bool unclean = false;
      if (mine) {
        unclean = !answered;
        // an earlier pending locker on the same best (exact), or anybody's best == my second (conservative)
        const unsigned lockers = __ballot_sync(pending, locker);
        const unsigned same = __match_any_sync(pending, i1);
        if (i1 >= 0 && (same & lockers & ((1u << lane) - 1))) unclean = true;
        if (ratio_mode && i2 >= 0 && BIT(pend, i2)) unclean = true;
      }
      __syncwarp();
      if (locker) atomicAnd(&pend[i1 >> 5], ~(1u << (i1 & 31)));
      const unsigned bad = __ballot_sync(FULL, unclean);
      const int first = bad ? __ffs(bad) - 1 : 32;
      // commit the clean prefix in parallel
      if (mine && lane < first && acc) {
        result = i1;
        if (fl & 2) atomicOr(&lockbits[i1 >> 5], 1u << (i1 & 31));
      }
      pending &= first >= 32 ? 0u : ~((2u << first) - 1u);  // lanes <= first are done after this round
      __syncwarp();
      if (first < 32) {  // resolve lane `first` alone, with every earlier query committed
        bool ok = true;
        if (lane == first) { ok = pick(); acc = ok && accept_of(i1, i2, d1, d2); }
        ok = __shfl_sync(FULL, ok, first);
        if (!ok) {  // warp-cooperative rescan of that query's window under the current locks
          const size_t qf = qbase + base + first;
          const ProjQuery q = proj_query(A, qf);
          const Desc256 dq = load_desc(A.qdesc + qf * 32);
          Top2 t;
          int seq = 0;
          for (int ix = q.cx0; ix <= q.cx1; ix++) {
            const int e0 = off[ix * FRAME_GRID_ROWS + q.cy0], e1 = off[ix * FRAME_GRID_ROWS + q.cy1 + 1];
            for (int e = e0 + lane; e < e1; e += 32) {
              const int j = idx[e];
              if (proj_gate(A, q, k, fbase, j) && !BIT(lockbits, j)) {
                const int d = hamming(dq, load_desc(dsc + (size_t)j * 32));
                t.push(make_key(d, (seq + e - e0) & 0xFFFFF, ORBM_TIE_FIRST_MIN), j);
              }
            }
            seq += e1 - e0;
          }
          warp_top2(t);
          if (lane == first) {
            i1 = t.i1; i2 = t.i2;
            d1 = t.i1 >= 0 ? (int)(t.k1 >> 20) : 256;
            d2 = t.i2 >= 0 ? (int)(t.k2 >> 20) : 256;
            acc = accept_of(i1, i2, d1, d2);
          }
        }
        if (lane == first && acc) {
          result = i1;
          if (fl & 2) atomicOr(&lockbits[i1 >> 5], 1u << (i1 & 31));
        }
        __syncwarp();
      }
    }
    // outputs of the chunk
    if (qi < nq) {
      match[qo] = result;
      if (result >= 0) {
        atomicMax(&feat_query[fbase + result], qi);  // the last query (in order) to write the feature holds it
        if (rot_check) atomicAdd(&hist[proj_rot_bin(A.q_angle[qo], k[result].angle)], 1);
        accepted++;
      }
    }
  }
#undef BIT
  __syncwarp();
  int erased = 0;
  if (rot_check) {
    if (lane == 0) {  // ComputeThreeMaxima (:2012-2053)
      int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
      for (int i = 0; i < 30; i++) {
        const int s = hist[i];
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
        else if (s > max3) { max3 = s; ind3 = i; }
      }
      if ((float)max2 < f_mul(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
      else if ((float)max3 < f_mul(0.1f, (float)max1)) { ind3 = -1; }
      s_ind[0] = ind1; s_ind[1] = ind2; s_ind[2] = ind3;
    }
    __syncwarp();
    const int ind1 = s_ind[0], ind2 = s_ind[1], ind3 = s_ind[2];
    for (int qi = lane; qi < nq; qi += 32) {
      const int m = match[qbase + qi];
      if (m < 0) continue;
      const int bin = proj_rot_bin(A.q_angle[qbase + qi], k[m].angle);
      if (bin != ind1 && bin != ind2 && bin != ind3) {
        match[qbase + qi] = -1;
        feat_query[fbase + m] = -1;  // the pointer is nulled whoever wrote it last (:1870-1876)
        erased++;
      }
    }
  }
  int total = accepted - erased;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) total += __shfl_xor_sync(FULL, total, o);
  if (lane == 0) nmatches[f] = total;
}

extern "C" int orbm_search_by_projection_th(
    const orbx_keypoint* d_kps, const uint8_t* d_desc, const int32_t* d_n, int kp_stride, int batch,
    const int32_t* d_cell_off, const int32_t* d_cell_idx, float min_x, float max_x, float min_y, float max_y,
    const float* d_u_right, const uint8_t* d_feat_locked, const float* d_qx, const float* d_qy, const float* d_qr,
    const int8_t* d_qmin_level, const int8_t* d_qmax_level, const uint8_t* d_qdesc, const int32_t* d_nq,
    int q_stride, const float* d_q_ur, const uint8_t* d_q_flags, const float* d_q_angle, int mode, float nn_ratio,
    int check_orientation, int th_dist, void* d_scratch, int32_t* d_match, int32_t* d_feat_query, int32_t* d_nmatches,
    void* stream) {
  if (!d_kps || !d_desc || !d_n || !d_cell_off || !d_cell_idx || !d_qx || !d_qy || !d_qr || !d_qmin_level ||
      !d_qmax_level || !d_qdesc || !d_nq || !d_scratch || !d_match || !d_feat_query || !d_nmatches || batch < 1 ||
      kp_stride < 1 || q_stride < 1 || !(max_x > min_x) || !(max_y > min_y))
    return ORBX_E_INVALID;
  if (mode != ORBM_PROJ_MAPPOINTS && mode != ORBM_PROJ_LASTFRAME) return ORBX_E_INVALID;
  if (th_dist < 0 || th_dist > 256) return ORBX_E_INVALID;
  if (mode == ORBM_PROJ_LASTFRAME && check_orientation && !d_q_angle) return ORBX_E_INVALID;
  const size_t smem = 2 * (size_t)((kp_stride + 31) / 32) * sizeof(uint32_t);
  if (smem > 40 * 1024) return ORBX_E_INVALID;  // > 160k features per frame
  ProjArgs A;
  A.kps = (const OrbxKeyPointDev*)d_kps; A.desc = d_desc; A.n = d_n; A.kp_stride = kp_stride;
  A.cell_off = d_cell_off; A.cell_idx = d_cell_idx;
  A.min_x = min_x; A.min_y = min_y;
  A.winv = (float)FRAME_GRID_COLS / (max_x - min_x); A.hinv = (float)FRAME_GRID_ROWS / (max_y - min_y);
  A.u_right = d_u_right; A.feat_locked = d_feat_locked;
  A.qx = d_qx; A.qy = d_qy; A.qr = d_qr; A.qmin = d_qmin_level; A.qmax = d_qmax_level; A.qdesc = d_qdesc;
  A.nq = d_nq; A.q_stride = q_stride; A.q_ur = d_q_ur; A.q_flags = d_q_flags; A.q_angle = d_q_angle;
  A.mode = mode; A.nnratio = nn_ratio; A.check_orientation = check_orientation; A.th_dist = th_dist;
  A.flag_bit = 1;
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grd((q_stride + PS_THREADS - 1) / PS_THREADS, batch);
  k_proj_score<<<grd, PS_THREADS, 0, st>>>(A, (ProjCand*)d_scratch);
  k_proj_resolve<<<batch, 32, smem, st>>>(A, (const ProjCand*)d_scratch, d_match, d_feat_query, d_nmatches);
  return cudaGetLastError() == cudaSuccess ? ORBX_OK : ORBX_E_CUDA;
}

extern "C" int orbm_search_by_projection(
    const orbx_keypoint* d_kps, const uint8_t* d_desc, const int32_t* d_n, int kp_stride, int batch,
    const int32_t* d_cell_off, const int32_t* d_cell_idx, float min_x, float max_x, float min_y, float max_y,
    const float* d_u_right, const uint8_t* d_feat_locked, const float* d_qx, const float* d_qy, const float* d_qr,
    const int8_t* d_qmin_level, const int8_t* d_qmax_level, const uint8_t* d_qdesc, const int32_t* d_nq,
    int q_stride, const float* d_q_ur, const uint8_t* d_q_flags, const float* d_q_angle, int mode, float nn_ratio,
    int check_orientation, void* d_scratch, int32_t* d_match, int32_t* d_feat_query, int32_t* d_nmatches,
    void* stream) {
  return orbm_search_by_projection_th(d_kps, d_desc, d_n, kp_stride, batch, d_cell_off, d_cell_idx, min_x, max_x,
                                      min_y, max_y, d_u_right, d_feat_locked, d_qx, d_qy, d_qr, d_qmin_level,
                                      d_qmax_level, d_qdesc, d_nq, q_stride, d_q_ur, d_q_flags, d_q_angle, mode,
                                      nn_ratio, check_orientation, 100 /* TH_HIGH */, d_scratch, d_match,
                                      d_feat_query, d_nmatches, stream);
}

// ---------------------------------------------------------------------------------------------------------
// orbm_search_by_projection_static: the same whole function for callers that cannot allocate (code under CUDA
// graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The candidate lists that orbm_search_by_projection takes from the
// caller as d_scratch live in a file-scope device array of ORBM_STATIC_SCRATCH_BYTES (4 MiB = 131 072 queries of
// 32 bytes; every GPU that loads the library holds it until the process exits).
//  * g_proj_static_scratch is, in host code, only a stand-in for the symbol: the device address comes from
//    cudaGetSymbolAddress, looked up once per device with that device current, and is what the kernels get.
//  * The entry point calls no cudaMalloc / cudaFree / allocating library routine.
//  * Every call shares the array and it keeps the last call's contents.  Nothing in it has to start at zero:
//    k_proj_score writes entry (frame, query) for every query < nq[frame] before k_proj_resolve reads it, and no
//    other entry is read.  Two calls that overlap on one GPU (different streams, or two host threads) race on it;
//    calls on one stream, or on different GPUs, do not.
// ---------------------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_proj_static_scratch[ORBM_STATIC_SCRATCH_BYTES];

#define ORBM_STATIC_MAX_DEVICES 64
static std::mutex g_static_scratch_mu;
static void* g_static_scratch_addr[ORBM_STATIC_MAX_DEVICES];  // device address of the array, per device (nullptr = not looked up)

// Device address of g_proj_static_scratch on the current device (cached per device).
static int proj_static_scratch(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= ORBM_STATIC_MAX_DEVICES) {
    orbx_set_error_text("orbm_search_by_projection_static: no current CUDA device (or device ordinal >= 64)");
    return e != cudaSuccess ? ORBX_E_CUDA : ORBX_E_INVALID;
  }
  std::lock_guard<std::mutex> lk(g_static_scratch_mu);
  if (!g_static_scratch_addr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_proj_static_scratch);
    if (e != cudaSuccess || !p) {
      char msg[200];
      snprintf(msg, sizeof(msg), "orbm_search_by_projection_static: cudaGetSymbolAddress(scratch) on device %d: %s", dev,
               cudaGetErrorString(e));
      orbx_set_error_text(msg);
      return ORBX_E_CUDA;
    }
    g_static_scratch_addr[dev] = p;
  }
  *out = g_static_scratch_addr[dev];
  return ORBX_OK;
}

extern "C" int orbm_search_by_projection_static(
    const orbx_keypoint* d_kps, const uint8_t* d_desc, const int32_t* d_n, int kp_stride, int batch,
    const int32_t* d_cell_off, const int32_t* d_cell_idx, float min_x, float max_x, float min_y, float max_y,
    const float* d_u_right, const uint8_t* d_feat_locked, const float* d_qx, const float* d_qy, const float* d_qr,
    const int8_t* d_qmin_level, const int8_t* d_qmax_level, const uint8_t* d_qdesc, const int32_t* d_nq,
    int q_stride, const float* d_q_ur, const uint8_t* d_q_flags, const float* d_q_angle, int mode, float nn_ratio,
    int check_orientation, int32_t* d_match, int32_t* d_feat_query, int32_t* d_nmatches, void* stream) {
  if (batch < 1 || q_stride < 1) return ORBX_E_INVALID;
  const unsigned long long need = (unsigned long long)batch * (unsigned long long)q_stride * sizeof(ProjCand);
  if (need > (unsigned long long)ORBM_STATIC_SCRATCH_BYTES) {
    char msg[256];
    snprintf(msg, sizeof(msg),
             "orbm_search_by_projection_static: %llu queries need %llu bytes of scratch, the static array holds %u; use "
             "orbm_search_by_projection",
             (unsigned long long)batch * q_stride, need, (unsigned)ORBM_STATIC_SCRATCH_BYTES);
    orbx_set_error_text(msg);
    return ORBX_E_CAPACITY;
  }
  void* scratch = nullptr;
  const int rc = proj_static_scratch(&scratch);
  if (rc != ORBX_OK) return rc;
  return orbm_search_by_projection_th(d_kps, d_desc, d_n, kp_stride, batch, d_cell_off, d_cell_idx, min_x, max_x,
                                      min_y, max_y, d_u_right, d_feat_locked, d_qx, d_qy, d_qr, d_qmin_level,
                                      d_qmax_level, d_qdesc, d_nq, q_stride, d_q_ur, d_q_flags, d_q_angle, mode,
                                      nn_ratio, check_orientation, 100 /* TH_HIGH */, scratch, d_match, d_feat_query,
                                      d_nmatches, stream);
}

// ---------------------------------------------------------------------------------------------------------
// Two-camera rig (F.Nleft != -1): the fisheye twins of the two functions above.
//   mode ORBM_PROJ_MAPPOINTS  src/ORBmatcher.cc:43-213  (right block :144-210, pair propagation :131-135, :197-201)
//   mode ORBM_PROJ_LASTFRAME  src/ORBmatcher.cc:1676-1887 (right block :1794-1862, one rotation histogram)
// Each camera keeps its own features, grid, lock bitmap and top-4 lists (k_proj_score, once per camera).  The commit
// is one warp per frame walking the MapPoints IN ORDER (left query, then right query): the two cameras are coupled
// through mvLeftToRightMatch / mvRightToLeftMatch (a match also writes — and re-locks or UNLOCKS — the paired feature
// of the other camera) and through the reference's `continue`s (a failed left ratio test, or an empty left window in
// the last-frame variant, skips the right search of that point), so the parallel prefix commit of k_proj_resolve
// does not apply; a query costs a broadcast list read and a few ballots.  Lists were built under the locks known on
// entry; if a paired write ever unlocks a feature that was locked on entry, that camera's lists are no longer
// supersets and every later query of it rescans its window (flag `dirty`).
// ---------------------------------------------------------------------------------------------------------
struct RigSide {
  ProjArgs A;                 // that camera's features, grid, window queries
  const ProjCand* cand;       // [batch][q_stride]
  const int32_t* partner;     // mvLeftToRightMatch / mvRightToLeftMatch [batch][kp_stride] (local index of the other camera, -1)
  int32_t* match;             // [batch][q_stride]
  int32_t* feat_query;        // [batch][kp_stride]
};

struct RigPick { int i1, i2, d1, d2; };

// best / second best of query qo on one camera under the CURRENT lock bitmap (warp-uniform result)
__device__ __forceinline__ RigPick rig_pick(const RigSide& S, const uint32_t* lockbits, bool dirty, bool want_second,
                                            int f, size_t qo, int lane) {
#define RBIT(bm, j) (((bm)[(j) >> 5] >> ((j) & 31)) & 1u)
  RigPick r;
  r.i1 = r.i2 = -1; r.d1 = r.d2 = 256;
  bool ok = false;
  if (!dirty) {
    const ProjCand c = S.cand[qo];
#pragma unroll
    for (int j = 0; j < PROJ_K; j++) {
      const int x = c.idx[j];
      if (x < 0 || RBIT(lockbits, x)) continue;
      if (r.i1 < 0) { r.i1 = x; r.d1 = (int)(c.key[j] >> 20); }
      else if (r.i2 < 0) { r.i2 = x; r.d2 = (int)(c.key[j] >> 20); }
    }
    const bool full = c.idx[PROJ_K - 1] >= 0;
    ok = !full || (r.i1 >= 0 && (!want_second || r.i2 >= 0));
  }
  if (!ok) {  // warp-cooperative rescan of the window under the current locks (entry locks are in the bitmap)
    const ProjArgs& A = S.A;
    const size_t fbase = (size_t)f * A.kp_stride;
    const OrbxKeyPointDev* k = A.kps + fbase;
    const uint8_t* dsc = A.desc + fbase * 32;
    const int32_t* off = A.cell_off + (size_t)f * (GRID_CELLS + 1);
    const int32_t* idx = A.cell_idx + fbase;
    const ProjQuery q = proj_query(A, qo);
    Top2 t;
    if (q.ok) {
      const Desc256 dq = load_desc(A.qdesc + qo * 32);
      int seq = 0;
      for (int ix = q.cx0; ix <= q.cx1; ix++) {
        const int e0 = off[ix * FRAME_GRID_ROWS + q.cy0], e1 = off[ix * FRAME_GRID_ROWS + q.cy1 + 1];
        for (int e = e0 + lane; e < e1; e += 32) {
          const int j = idx[e];
          if (proj_gate_area(q, k, j) && !RBIT(lockbits, j)) {
            const int d = hamming(dq, load_desc(dsc + (size_t)j * 32));
            t.push(make_key(d, (seq + e - e0) & 0xFFFFF, ORBM_TIE_FIRST_MIN), j);
          }
        }
        seq += e1 - e0;
      }
    }
    warp_top2(t);
    r.i1 = t.i1; r.i2 = t.i2;
    r.d1 = t.i1 >= 0 ? (int)(t.k1 >> 20) : 256;
    r.d2 = t.i2 >= 0 ? (int)(t.k2 >> 20) : 256;
  }
  return r;
#undef RBIT
}

__global__ void __launch_bounds__(32)
k_proj_resolve_rig(const RigSide L, const RigSide R, const uint8_t* __restrict__ area_any_l, int mode, float nnratio,
                   int check_orientation, int th_dist, int32_t* __restrict__ nmatches) {
  extern __shared__ uint32_t s_bits[];
  __shared__ int hist[30];
  __shared__ int s_ind[3];
  const int lane = threadIdx.x, f = blockIdx.x;
  const int nwl = (L.A.kp_stride + 31) / 32, nwr = (R.A.kp_stride + 31) / 32;
  uint32_t* lockL = s_bits;
  uint32_t* lockR = s_bits + nwl;
  const size_t fbl = (size_t)f * L.A.kp_stride, fbr = (size_t)f * R.A.kp_stride, qbase = (size_t)f * L.A.q_stride;
  const int nl = L.A.n[f], nr = R.A.n[f], nq = L.A.nq[f];
  // lock bitmaps = the state on entry; feat_query = -1
  for (int w = lane; w < nwl; w += 32) {
    uint32_t m = 0;
    if (L.A.feat_locked)
      for (int b = 0; b < 32; b++) { const int j = 32 * w + b; if (j < nl && L.A.feat_locked[fbl + j]) m |= 1u << b; }
    lockL[w] = m;
  }
  for (int w = lane; w < nwr; w += 32) {
    uint32_t m = 0;
    if (R.A.feat_locked)
      for (int b = 0; b < 32; b++) { const int j = 32 * w + b; if (j < nr && R.A.feat_locked[fbr + j]) m |= 1u << b; }
    lockR[w] = m;
  }
  for (int j = lane; j < nl; j += 32) L.feat_query[fbl + j] = -1;
  for (int j = lane; j < nr; j += 32) R.feat_query[fbr + j] = -1;
  if (lane < 30) hist[lane] = 0;
  __syncwarp();
  const bool ratio_mode = mode == ORBM_PROJ_MAPPOINTS;
  const bool rot_check = mode == ORBM_PROJ_LASTFRAME && check_orientation;
  const OrbxKeyPointDev* kl = L.A.kps + fbl;
  const OrbxKeyPointDev* kr = R.A.kps + fbr;
  bool dirtyL = false, dirtyR = false;
  int total = 0;
  // writes a MapPoint into a feature of one camera: holder, lock bit (set or cleared), "dirty" on an entry lock lost
  auto write_feat = [&](const RigSide& S, uint32_t* bits, bool& dirty, size_t fb, int j, int qi, bool locker) {
    if (lane == 0) {
      S.feat_query[fb + j] = qi;
      if (locker) bits[j >> 5] |= 1u << (j & 31); else bits[j >> 5] &= ~(1u << (j & 31));
    }
    if (!locker && S.A.feat_locked && S.A.feat_locked[fb + j]) dirty = true;
    __syncwarp();
  };
  auto accept = [&](const RigPick& p, const OrbxKeyPointDev* k, bool& ratio_failed) {
    ratio_failed = false;
    if (!(p.i1 >= 0 && p.d1 <= th_dist)) return false;
    if (ratio_mode) {
      const int lv1 = k[p.i1].octave, lv2 = p.i2 >= 0 ? k[p.i2].octave : -1;
      if (lv1 == lv2 && (float)p.d1 > f_mul(nnratio, (float)p.d2)) { ratio_failed = true; return false; }
    }
    return true;
  };
  for (int qi = 0; qi < nq; qi++) {
    const size_t qo = qbase + qi;
    const int fl = L.A.q_flags ? L.A.q_flags[qo] : 7;
    const bool locker = (fl & 2) != 0;
    int ml = -1, mr = -1;
    bool skip_right = false;
    if (fl & 1) {
      const RigPick p = rig_pick(L, lockL, dirtyL, ratio_mode, f, qo, lane);
      bool rf;
      if (accept(p, kl, rf)) {
        ml = p.i1;
        write_feat(L, lockL, dirtyL, fbl, p.i1, qi, locker);
        total++;
        if (ratio_mode) {
          const int pr = L.partner ? L.partner[fbl + p.i1] : -1;   // :131-135
          if (pr >= 0) { write_feat(R, lockR, dirtyR, fbr, pr, qi, locker); total++; }
        } else if (rot_check && lane == 0) {
          hist[proj_rot_bin(L.A.q_angle[qo], kl[p.i1].angle)]++;
        }
      } else if (rf) {
        skip_right = true;                                          // `continue` at :126-127
      }
      if (!ratio_mode && !area_any_l[qo]) skip_right = true;        // `continue` at :1737-1738
    } else if (!ratio_mode) {
      skip_right = true;                                            // the projection gates failed (:1706-1714)
    }
    const bool right_on = ratio_mode ? (fl & 4) != 0 : true;
    if (!skip_right && right_on) {
      const RigPick p = rig_pick(R, lockR, dirtyR, ratio_mode, f, qo, lane);
      bool rf;
      if (accept(p, kr, rf)) {
        mr = p.i1;
        if (ratio_mode) {
          const int pl = R.partner ? R.partner[fbr + p.i1] : -1;    // :197-201
          if (pl >= 0) { write_feat(L, lockL, dirtyL, fbl, pl, qi, locker); total++; }
        }
        write_feat(R, lockR, dirtyR, fbr, p.i1, qi, locker);
        total++;
        if (rot_check && lane == 0) hist[proj_rot_bin(L.A.q_angle[qo], kr[p.i1].angle)]++;
      }
    }
    if (lane == 0) { L.match[qo] = ml; R.match[qo] = mr; }
  }
  __syncwarp();
  int erased = 0;
  if (rot_check) {
    if (lane == 0) {  // ComputeThreeMaxima (:2012-2053)
      int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
      for (int i = 0; i < 30; i++) {
        const int s = hist[i];
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
        else if (s > max3) { max3 = s; ind3 = i; }
      }
      if ((float)max2 < f_mul(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
      else if ((float)max3 < f_mul(0.1f, (float)max1)) { ind3 = -1; }
      s_ind[0] = ind1; s_ind[1] = ind2; s_ind[2] = ind3;
    }
    __syncwarp();
    const int ind1 = s_ind[0], ind2 = s_ind[1], ind3 = s_ind[2];
    for (int qi = lane; qi < nq; qi += 32) {
      const size_t qo = qbase + qi;
      const int a = L.match[qo], b = R.match[qo];
      if (a >= 0) {
        const int bin = proj_rot_bin(L.A.q_angle[qo], kl[a].angle);
        if (bin != ind1 && bin != ind2 && bin != ind3) { L.match[qo] = -1; L.feat_query[fbl + a] = -1; erased++; }
      }
      if (b >= 0) {
        const int bin = proj_rot_bin(L.A.q_angle[qo], kr[b].angle);
        if (bin != ind1 && bin != ind2 && bin != ind3) { R.match[qo] = -1; R.feat_query[fbr + b] = -1; erased++; }
      }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) erased += __shfl_xor_sync(0xffffffffu, erased, o);
  if (lane == 0) nmatches[f] = total - erased;
}

static ProjArgs rig_args(const orbm_rig_camera* c, float min_x, float max_x, float min_y, float max_y,
                         const float* qx, const float* qy, const float* qr, const int8_t* qmin, const int8_t* qmax,
                         const uint8_t* qdesc, const int32_t* nq, int q_stride, const uint8_t* q_flags,
                         const float* q_angle, int mode, float nn_ratio, int check_orientation, int flag_bit) {
  ProjArgs A;
  A.kps = (const OrbxKeyPointDev*)c->d_kps; A.desc = c->d_desc; A.n = c->d_n; A.kp_stride = c->kp_stride;
  A.cell_off = c->d_cell_off; A.cell_idx = c->d_cell_idx;
  A.min_x = min_x; A.min_y = min_y;
  A.winv = (float)FRAME_GRID_COLS / (max_x - min_x); A.hinv = (float)FRAME_GRID_ROWS / (max_y - min_y);
  A.u_right = nullptr; A.feat_locked = c->d_feat_locked;
  A.qx = qx; A.qy = qy; A.qr = qr; A.qmin = qmin; A.qmax = qmax; A.qdesc = qdesc; A.nq = nq; A.q_stride = q_stride;
  A.q_ur = nullptr; A.q_flags = q_flags; A.q_angle = q_angle;
  A.mode = mode; A.nnratio = nn_ratio; A.check_orientation = check_orientation; A.th_dist = 100; A.flag_bit = flag_bit;
  return A;
}

extern "C" int orbm_search_by_projection_rig(
    int mode, const orbm_rig_camera* left, const orbm_rig_camera* right, int batch, float min_x, float max_x,
    float min_y, float max_y, const float* d_qx, const float* d_qy, const float* d_qr, const int8_t* d_qmin_level,
    const int8_t* d_qmax_level, const float* d_qx_r, const float* d_qy_r, const float* d_qr_r,
    const int8_t* d_qmin_level_r, const int8_t* d_qmax_level_r, const uint8_t* d_qdesc, const int32_t* d_nq,
    int q_stride, const uint8_t* d_q_flags, const float* d_q_angle, float nn_ratio, int check_orientation,
    void* d_scratch, int32_t* d_match_l, int32_t* d_match_r, int32_t* d_feat_query_l, int32_t* d_feat_query_r,
    int32_t* d_nmatches, void* stream) {
  if ((mode != ORBM_PROJ_MAPPOINTS && mode != ORBM_PROJ_LASTFRAME) || !left || !right || batch < 1 || q_stride < 1 ||
      !(max_x > min_x) || !(max_y > min_y) || !d_qx || !d_qy || !d_qr || !d_qmin_level || !d_qmax_level || !d_qx_r ||
      !d_qy_r || !d_qr_r || !d_qmin_level_r || !d_qmax_level_r || !d_qdesc || !d_nq || !d_scratch || !d_match_l ||
      !d_match_r || !d_feat_query_l || !d_feat_query_r || !d_nmatches)
    return ORBX_E_INVALID;
  for (const orbm_rig_camera* c : {left, right})
    if (!c->d_kps || !c->d_desc || !c->d_n || !c->d_cell_off || !c->d_cell_idx || c->kp_stride < 1) return ORBX_E_INVALID;
  if (mode == ORBM_PROJ_LASTFRAME && check_orientation && !d_q_angle) return ORBX_E_INVALID;
  const size_t smem = ((size_t)(left->kp_stride + 31) / 32 + (size_t)(right->kp_stride + 31) / 32) * sizeof(uint32_t);
  if (smem > 40 * 1024) return ORBX_E_INVALID;
  RigSide L, R;
  L.A = rig_args(left, min_x, max_x, min_y, max_y, d_qx, d_qy, d_qr, d_qmin_level, d_qmax_level, d_qdesc, d_nq, q_stride,
                 d_q_flags, d_q_angle, mode, nn_ratio, check_orientation, 1);
  R.A = rig_args(right, min_x, max_x, min_y, max_y, d_qx_r, d_qy_r, d_qr_r, d_qmin_level_r, d_qmax_level_r, d_qdesc,
                 d_nq, q_stride, d_q_flags, d_q_angle, mode, nn_ratio, check_orientation,
                 mode == ORBM_PROJ_MAPPOINTS ? 4 : 1);
  // scratch: two candidate arrays [batch][q_stride] + the left camera's "window not empty" bytes
  ProjCand* cand_l = (ProjCand*)d_scratch;
  ProjCand* cand_r = cand_l + (size_t)batch * q_stride;
  uint8_t* any_l = (uint8_t*)(cand_r + (size_t)batch * q_stride);
  L.cand = cand_l; R.cand = cand_r;
  L.partner = left->d_partner; R.partner = right->d_partner;
  L.match = d_match_l; R.match = d_match_r; L.feat_query = d_feat_query_l; R.feat_query = d_feat_query_r;
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grd((q_stride + PS_THREADS - 1) / PS_THREADS, batch);
  k_proj_score<<<grd, PS_THREADS, 0, st>>>(L.A, cand_l, any_l);
  k_proj_score<<<grd, PS_THREADS, 0, st>>>(R.A, cand_r, nullptr);
  k_proj_resolve_rig<<<batch, 32, smem, st>>>(L, R, any_l, mode, nn_ratio, check_orientation, 100, d_nmatches);
  return cudaGetLastError() == cudaSuccess ? ORBX_OK : ORBX_E_CUDA;
}
