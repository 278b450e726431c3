// octree.cu — ORBextractor::DistributeOctTree (src/ORBextractor.cc:555-779) on device.
// One CTA per (level, frame).  The node list is an array in list order that is rebuilt every round
// from block-wide prefix sums (octree_core.h, "scan" form — the CPU harness runs the same
// formulation with plain loops against the oracle); per-key work (root/quadrant labelling,
// counting, relabelling, best-response selection) is spread over the CTA.  The only serial step
// left is the libstdc++ std::sort replica that orders the careful phase.
#include <stdlib.h>

#include "octree_core.h"
#include "orbx_kernels.cuh"

#define OCT_THREADS 256      // ordinary frames; levels with many keypoints (4K) run OCT_THREADS_BIG threads per CTA:
#define OCT_THREADS_BIG 1024 // the per-key loops and block scans of a round shrink with the CTA width
#define OCT_THREADS_SMALL 128
#define OCT_SMALL_CAP 300     // node lists below this (<= ~290 keypoints on the largest level) use the narrow CTA
#define OCT_MAX_WARPS 32
#define OCT_MAX_ROOTS 64

struct OctShared {
  int root_cnt[OCT_MAX_ROOTS];
  int root_pos[OCT_MAX_ROOTS];
  int wsum[2][OCT_MAX_WARPS];  // block-scan scratch
  int bcast[4];            // block-wide results (totals, first hit)
};

// Dynamic shared memory carve-up (see orbx_octree_smem_bytes):
//   OctShared | nodes A[cap] | nodes B[cap] | qc[4*cap] | vec[cap] | vec2[cap] | best[cap]
//   | newpos[cap] (i16) | preK[cap] (i16) | preE[cap] (i16) | proc[cap] (u8)
//   | cell_off[max_cells+1] | keys[smem_keys] | labels[smem_keys]
size_t orbx_octree_smem_bytes(int smem_keys, int node_cap, int max_cells) {
  const size_t cap = (size_t)((node_cap + 7) & ~7);
  size_t b = (sizeof(OctShared) + 15) & ~(size_t)15;
  b += 2 * cap * sizeof(OctPNode);
  b += 4 * cap * sizeof(int);
  b += 2 * cap * sizeof(uint64_t);
  b += cap * sizeof(uint32_t);
  b += 3 * cap * sizeof(int16_t);
  b += cap;
  b += (((size_t)max_cells + 1) * sizeof(int) + 15) & ~(size_t)15;
  b += (size_t)smem_keys * sizeof(uint32_t);
  b += (size_t)smem_keys * sizeof(int16_t);
  return (b + 15) & ~(size_t)15;
}

// Exclusive block scan of two values per thread (a, b), with the block totals.  Two barriers.
template <int NT>
__device__ __forceinline__ void block_scan2(OctShared& sh, int a, int b, int& ea, int& eb, int& ta, int& tb) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int ia = a, ib = b;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int xa = __shfl_up_sync(0xffffffffu, ia, o), xb = __shfl_up_sync(0xffffffffu, ib, o);
    if (lane >= o) { ia += xa; ib += xb; }
  }
  if (lane == 31) { sh.wsum[0][warp] = ia; sh.wsum[1][warp] = ib; }
  __syncthreads();
  int oa = 0, ob = 0;
  ta = 0; tb = 0;
#pragma unroll
  for (int w = 0; w < NT / 32; w++) {
    const int va = sh.wsum[0][w], vb = sh.wsum[1][w];
    if (w < warp) { oa += va; ob += vb; }
    ta += va; tb += vb;
  }
  ea = oa + ia - a;
  eb = ob + ib - b;
  __syncthreads();
}

__device__ __forceinline__ int quad_K(const int* c) { return (c[0] > 0) + (c[1] > 0) + (c[2] > 0) + (c[3] > 0); }
__device__ __forceinline__ int quad_E(const int* c) { return (c[0] > 1) + (c[1] > 1) + (c[2] > 1) + (c[3] > 1); }

template <int NT>
__global__ void __launch_bounds__(NT)
k_octree(const __grid_constant__ DevPlan P, const int32_t* __restrict__ cell_cnt,
         const uint32_t* __restrict__ cand, uint32_t* __restrict__ gkeys, int16_t* __restrict__ glabels,
         uint32_t* __restrict__ sel, int32_t* __restrict__ sel_cnt, int32_t* __restrict__ status,
         int smem_keys, int node_cap, int max_cells, int level_major) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  // all CTAs of level 0 come first in launch order, then level 1, ...: the longest-running CTAs (most candidates)
  // start first and the short ones fill in behind them (was (level, frame): 0.265 -> see DESIGN.md §4)
  const int level = level_major ? blockIdx.y : blockIdx.x, frame = level_major ? blockIdx.x : blockIdx.y;
  const DevLevel& L = P.lv[level];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int cap = (node_cap + 7) & ~7;

  unsigned char* p = smem_raw;
  OctShared& sh = *reinterpret_cast<OctShared*>(p);
  p += (sizeof(OctShared) + 15) & ~(size_t)15;
  OctPNode* cur = reinterpret_cast<OctPNode*>(p);
  p += (size_t)cap * sizeof(OctPNode);
  OctPNode* nxt = reinterpret_cast<OctPNode*>(p);
  p += (size_t)cap * sizeof(OctPNode);
  int* qc = reinterpret_cast<int*>(p);
  p += 4 * (size_t)cap * sizeof(int);
  uint64_t* vec = reinterpret_cast<uint64_t*>(p);
  p += (size_t)cap * sizeof(uint64_t);
  uint64_t* vec2 = reinterpret_cast<uint64_t*>(p);
  p += (size_t)cap * sizeof(uint64_t);
  uint32_t* best = reinterpret_cast<uint32_t*>(p);
  p += (size_t)cap * sizeof(uint32_t);
  int16_t* newpos = reinterpret_cast<int16_t*>(p);
  p += (size_t)cap * sizeof(int16_t);
  int16_t* preKs = reinterpret_cast<int16_t*>(p);
  p += (size_t)cap * sizeof(int16_t);
  int16_t* preEs = reinterpret_cast<int16_t*>(p);
  p += (size_t)cap * sizeof(int16_t);
  uint8_t* proc = reinterpret_cast<uint8_t*>(p);
  p += (size_t)cap;
  int* cell_off = reinterpret_cast<int*>(p);
  p += (((size_t)max_cells + 1) * sizeof(int) + 15) & ~(size_t)15;
  uint32_t* skeys = reinterpret_cast<uint32_t*>(p);
  p += (size_t)smem_keys * sizeof(uint32_t);
  int16_t* slabels = reinterpret_cast<int16_t*>(p);

  // ---- 1. exclusive scan of the per-cell counts (cell-major order = reference's input order)
  const int ncell = L.nCols * L.nRows;
  const int32_t* cc = cell_cnt + (long long)frame * P.total_cells + L.cell_base;
  int n = 0;
  for (int base = 0; base < ncell; base += NT) {
    const int c = base + tid;
    const int v = c < ncell ? cc[c] : 0;
    int ev, e2, tv, t2;
    block_scan2<NT>(sh, v, 0, ev, e2, tv, t2);
    if (c < ncell) cell_off[c] = n + ev;
    n += tv;
  }
  if (tid == 0) cell_off[ncell] = n;
  for (int i = tid; i < OCT_MAX_ROOTS; i += NT) sh.root_cnt[i] = 0;
  __syncthreads();
  int32_t* cnt_out = sel_cnt + frame * P.nlevels + level;
  if (n == 0) {
    if (tid == 0) *cnt_out = 0;
    return;
  }

  // ---- 2. gather keys (warp per cell), label with the root node, count keys per root
  const long long lvl_off = (long long)frame * P.cand_total + L.cand_base;
  uint32_t* keys = n <= smem_keys ? skeys : gkeys + lvl_off;
  int16_t* labels = n <= smem_keys ? slabels : glabels + lvl_off;
  const uint32_t* lc = cand + lvl_off;
  for (int c = warp; c < ncell; c += (NT / 32)) {
    const int o = cell_off[c], m = cell_off[c + 1] - o;
    for (int i = lane; i < m; i += 32) {
      const uint32_t k = lc[(long long)c * L.cell_cap + i];
      keys[o + i] = k;
      const int r = (int)f_div((float)ORBX_KEY_X(k), L.hX);  // vpIniNodes[kp.pt.x/hX] (:585)
      labels[o + i] = (int16_t)r;
      atomicAdd(&sh.root_cnt[r], 1);
    }
  }
  __syncthreads();
  // roots that received keys, in index order, are the initial list (:589-601)
  const int N = L.nfeat;
  int S;
  {
    const int nIni = L.nIni;  // <= OCT_MAX_ROOTS (checked at plan time)
    const int c = tid < nIni ? sh.root_cnt[tid] : 0;
    int e, e2, tot, t2;
    block_scan2<NT>(sh, c > 0 ? 1 : 0, 0, e, e2, tot, t2);
    if (tid < nIni) {
      sh.root_pos[tid] = c > 0 ? e : -1;
      if (c > 0) cur[e] = octp_root(tid, L.hX, L.h - 2 * ORBX_MIN_BORDER, c);
    }
    S = tot;
  }
  __syncthreads();
  for (int i = tid; i < n; i += NT) labels[i] = (int16_t)sh.root_pos[labels[i]];
  __syncthreads();

  // ---- 3. rounds (octree_core.h, "scan" form)
  int nvec = 0;
  bool careful = false, overflow = false;
  while (true) {
    const int prevS = S;
    for (int q = tid; q < S; q += NT) {
      proc[q] = 0;
      *reinterpret_cast<int4*>(qc + 4 * q) = make_int4(0, 0, 0, 0);
    }
    __syncthreads();
    // (a) keys of nodes that may be divided vote for their quadrant
    for (int i = tid; i < n; i += NT) {
      const int pl = labels[i];
      const OctPNode nd = cur[pl];
      if (nd.cnt > 1) {
        const uint32_t k = keys[i];
        atomicAdd(&qc[4 * pl + octp_quadrant(nd, ORBX_KEY_X(k), ORBX_KEY_Y(k))], 1);
      }
    }
    __syncthreads();
    int totK = 0, totE = 0, nproc = 0;
    if (!careful) {
      // (b) every node with more than one key is divided, in list order (:620-676)
      {
        int sk = 0, sd = 0;
        for (int q = tid; q < S; q += NT)
          if (cur[q].cnt > 1) { sk += quad_K(qc + 4 * q); sd++; }
        int e1, e2;
        block_scan2<NT>(sh, sk, sd, e1, e2, totK, nproc);
      }
      if (totK + S - nproc > cap) { overflow = true; break; }
      int carryK = 0, carryE = 0;
      for (int base = 0; base < S; base += NT) {
        const int q = base + tid;
        const bool dv = q < S && cur[q].cnt > 1;
        int c[4] = {0, 0, 0, 0};
        if (dv) { const int4 v = *reinterpret_cast<const int4*>(qc + 4 * q); c[0] = v.x; c[1] = v.y; c[2] = v.z; c[3] = v.w; }
        const int K = quad_K(c), E = quad_E(c);
        int eK, eE, tK, tE;
        block_scan2<NT>(sh, K, E, eK, eE, tK, tE);
        if (dv) {
          proc[q] = 1;
          octp_emit(cur[q], c, totK - (carryK + eK) - K, carryE + eE, nxt, vec2);
          *reinterpret_cast<int4*>(qc + 4 * q) = make_int4(c[0], c[1], c[2], c[3]);
        }
        carryK += tK; carryE += tE;
      }
      totE = carryE;
    } else {
      // (c) careful loop (:691-754): candidates sorted by (count, UL.x) with libstdc++'s std::sort,
      // divided from the back until the list reaches N
      if (tid == 0) { oct_std_sort(vec, nvec); sh.bcast[0] = 0x7fffffff; }
      __syncthreads();
      int carryK = 0, carryE = 0;
      bool found = false;
      for (int base = 0; base < nvec; base += NT) {
        const int s = base + tid;
        const bool valid = s < nvec;
        int K = 0, E = 0;
        if (valid) { const int* c = qc + 4 * (int)(vec[nvec - 1 - s] & 0xffffu); K = quad_K(c); E = quad_E(c); }
        int eK, eE, tK, tE;
        block_scan2<NT>(sh, K, E, eK, eE, tK, tE);
        if (valid) { preKs[s] = (int16_t)(carryK + eK); preEs[s] = (int16_t)(carryE + eE); }
        const bool hit = valid && (S + carryK + eK + K - (s + 1) >= N);
        if (hit) atomicMin(&sh.bcast[0], s);
        __syncthreads();
        const int f = sh.bcast[0];
        if (f != 0x7fffffff) {
          if (s == f) { sh.bcast[1] = carryK + eK + K; sh.bcast[2] = carryE + eE + E; }
          __syncthreads();
          totK = sh.bcast[1]; totE = sh.bcast[2]; nproc = f + 1;
          found = true;
          break;
        }
        carryK += tK; carryE += tE;
      }
      if (!found) { totK = carryK; totE = carryE; nproc = nvec; }
      if (totK + S - nproc > cap) { overflow = true; break; }
      for (int s = tid; s < nproc; s += NT) {
        const int q = (int)(vec[nvec - 1 - s] & 0xffffu);
        const int4 v = *reinterpret_cast<const int4*>(qc + 4 * q);
        int c[4] = {v.x, v.y, v.z, v.w};
        proc[q] = 1;
        octp_emit(cur[q], c, totK - preKs[s] - quad_K(c), preEs[s], nxt, vec2);
        *reinterpret_cast<int4*>(qc + 4 * q) = make_int4(c[0], c[1], c[2], c[3]);
      }
    }
    __syncthreads();
    // (d) nodes that were not divided keep their relative order behind all the new children
    int stay = 0;
    for (int base = 0; base < S; base += NT) {
      const int q = base + tid;
      const bool st = q < S && !proc[q];
      int e, e2, t, t2;
      block_scan2<NT>(sh, st ? 1 : 0, 0, e, e2, t, t2);
      if (st) {
        const int pos = totK + stay + e;
        nxt[pos] = cur[q];
        newpos[q] = (int16_t)pos;
      }
      stay += t;
    }
    __syncthreads();
    // (e) keys follow their node
    for (int i = tid; i < n; i += NT) {
      const int pl = labels[i];
      if (proc[pl]) {
        const uint32_t k = keys[i];
        labels[i] = (int16_t)qc[4 * pl + octp_quadrant(cur[pl], ORBX_KEY_X(k), ORBX_KEY_Y(k))];
      } else {
        labels[i] = newpos[pl];
      }
    }
    __syncthreads();
    { OctPNode* t = cur; cur = nxt; nxt = t; }
    { uint64_t* t = vec; vec = vec2; vec2 = t; }
    nvec = totE;
    S = totK + stay;
    if (S >= N || S == prevS) break;                       // :678-683, :749-750
    if (!careful && S + 3 * totE > N) careful = true;      // :684
  }

  // ---- 4. best response per node, first key wins ties (:757-778); list order = array order
  if (overflow) {
    if (tid == 0) { *cnt_out = 0; atomicOr(status, 1); }
    return;
  }
  for (int i = tid; i < S; i += NT) best[i] = 0;
  __syncthreads();
  for (int i = tid; i < n; i += NT)
    atomicMax(&best[labels[i]], ((uint32_t)ORBX_KEY_S(keys[i]) << 24) | (0xFFFFFFu - (uint32_t)i));
  __syncthreads();
  if (tid == 0) *cnt_out = S;
  uint32_t* out = sel + (long long)frame * P.kp_total_cap + L.kp_base;
  for (int r = tid; r < S; r += NT) {
    const int i = (int)(0xFFFFFFu - (best[r] & 0xFFFFFFu));
    const uint32_t k = keys[i];
    // back to level coordinates (:883-884)
    out[r] = orbx_pack_key(ORBX_KEY_X(k) + ORBX_MIN_BORDER, ORBX_KEY_Y(k) + ORBX_MIN_BORDER, ORBX_KEY_S(k));
  }
}

#include <mutex>
cudaError_t orbx_ensure_dyn_smem(OrbxSmemSlot slot, const void* func, size_t bytes) {
  static std::mutex mu;
  static int have[64][ORBX_SMEM_SLOTS];  // zero-initialised; 48 KB needs no opt-in
  const bool fast = slot == ORBX_SMEM_FAST || slot == ORBX_SMEM_FAST4 || slot == ORBX_SMEM_FAST_STRIP;
  if (bytes <= 48 * 1024 && !fast) return cudaSuccess;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  std::lock_guard<std::mutex> lk(mu);
  if (dev < 0 || dev >= 64) return cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if ((int)bytes <= have[dev][slot]) return cudaSuccess;
  // the FAST kernels are limited by shared memory per SM: ask for the largest carveout
  if (fast)
    cudaFuncSetAttribute(func, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (e == cudaSuccess) have[dev][slot] = (int)bytes;
  return e;
}

cudaError_t orbx_launch_octree(const DevPlan& P, const int32_t* cell_cnt, const uint32_t* cand,
                               uint32_t* gkeys, int16_t* glabels, uint32_t* sel, int32_t* sel_cnt,
                               int32_t* status, int batch, size_t smem_bytes, int smem_keys, int node_cap,
                               int max_cells, cudaStream_t st) {
  // the opt-in is per (function, device) and is only ever raised (a smaller handle created later must not lower it)
  // CTA width by the size of the largest level's node list: the rounds are barrier-bound, so narrow CTAs win while the
  // per-key loops are short (C1: 0.30 -> 0.26 ms / 512 frames at 128 threads) and wide ones once they are long (4K)
  static const int force = [] { const char* e = getenv("ORBX_OCTREE_THREADS"); return e ? atoi(e) : 0; }();
  const int nt = force ? force : (node_cap >= 768 ? OCT_THREADS_BIG : (node_cap >= OCT_SMALL_CAP ? OCT_THREADS : OCT_THREADS_SMALL));
  static const int level_major = [] { const char* e = getenv("ORBX_OCT_LEVEL_MAJOR"); return e ? atoi(e) : 1; }();
  dim3 grd(P.nlevels, batch);
  if (level_major) grd = dim3(batch, P.nlevels);
  cudaError_t e;
#define OCT_LAUNCH(NT, SLOT)                                                                                       \
  do {                                                                                                             \
    e = orbx_ensure_dyn_smem(SLOT, (const void*)k_octree<NT>, smem_bytes);                                         \
    if (e != cudaSuccess) return e;                                                                                \
    k_octree<NT><<<grd, NT, smem_bytes, st>>>(P, cell_cnt, cand, gkeys, glabels, sel, sel_cnt, status, smem_keys,   \
                                              node_cap, max_cells, level_major);                                                \
  } while (0)
  if (nt >= OCT_THREADS_BIG) OCT_LAUNCH(OCT_THREADS_BIG, ORBX_SMEM_OCTREE_BIG);
  else if (nt >= OCT_THREADS) OCT_LAUNCH(OCT_THREADS, ORBX_SMEM_OCTREE);
  else OCT_LAUNCH(OCT_THREADS_SMALL, ORBX_SMEM_OCTREE_X);  // (64 threads: 0.35 ms, slower again)
#undef OCT_LAUNCH
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// Output ordering (src/ORBextractor.cc:1120-1166): level-major traversal; keypoints whose scaled x
// lies in the lapping area fill the output from the back, the others from the front.
// One CTA per frame computes every keypoint's destination row.
// ------------------------------------------------------------------------------------------------
#define ORD_THREADS 256
__global__ void __launch_bounds__(ORD_THREADS)
k_order(const __grid_constant__ DevPlan P, const uint32_t* __restrict__ sel, const int32_t* __restrict__ sel_cnt,
        int lap0, int lap1, int32_t* __restrict__ dst, int32_t* __restrict__ n_out,
        int32_t* __restrict__ mono_index) {
  const int frame = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  __shared__ int ws[ORD_THREADS / 32];
  __shared__ int s_carry;
  int total = 0;
  for (int l = 0; l < P.nlevels; l++) total += sel_cnt[frame * P.nlevels + l];
  if (tid == 0) s_carry = 0;
  __syncthreads();
  const float flap0 = (float)lap0, flap1 = (float)lap1;
  for (int l = 0; l < P.nlevels; l++) {
    const int m = sel_cnt[frame * P.nlevels + l];
    const uint32_t* s = sel + (long long)frame * P.kp_total_cap + P.lv[l].kp_base;
    int32_t* d = dst + (long long)frame * P.kp_total_cap + P.lv[l].kp_base;
    for (int base = 0; base < m; base += ORD_THREADS) {
      const int i = base + tid;
      int flag = 0;
      if (i < m) {
        float x = (float)ORBX_KEY_X(s[i]);
        if (l != 0) x = f_mul(x, P.lv[l].scale);
        flag = (x >= flap0 && x <= flap1) ? 1 : 0;
      }
      int inc = flag;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
      }
      if (lane == 31) ws[warp] = inc;
      __syncthreads();
      int woff = 0;
#pragma unroll
      for (int w = 0; w < ORD_THREADS / 32; w++)
        if (w < warp) woff += ws[w];
      const int carry = s_carry;       // lapping keypoints before this chunk
      const int before = carry + woff + inc - flag;  // lapping keypoints before this one
      if (i < m) {
        // position in traversal order = (sum of previous levels) + i
        int trav = i;
        for (int q = 0; q < l; q++) trav += sel_cnt[frame * P.nlevels + q];
        d[i] = flag ? (total - 1 - before) : (trav - before);
      }
      __syncthreads();
      if (tid == ORD_THREADS - 1) s_carry = carry + woff + inc;
      __syncthreads();
    }
  }
  if (tid == 0) {
    n_out[frame] = total;
    mono_index[frame] = total - s_carry;
  }
}

cudaError_t orbx_launch_order(const DevPlan& P, const uint32_t* sel, const int32_t* sel_cnt, int lap0,
                              int lap1, int32_t* dst, int32_t* n_out, int32_t* mono_index, int batch,
                              cudaStream_t st) {
  k_order<<<batch, ORD_THREADS, 0, st>>>(P, sel, sel_cnt, lap0, lap1, dst, n_out, mono_index);
  return cudaGetLastError();
}
