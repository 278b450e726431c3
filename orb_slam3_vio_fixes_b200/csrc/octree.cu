// octree.cu — ORBextractor::DistributeOctTree (src/ORBextractor.cc:555-779) on device.
// One CTA per (level, frame).  Per-key work (root/quadrant labelling, counting, best-response
// selection) is spread over the CTA; the list surgery of each round runs in thread 0 through the
// shared octree_core.h (the same code the CPU harness checks against the oracle).
#include "octree_core.h"
#include "orbx_kernels.cuh"

#define OCT_THREADS 256
#define OCT_MAX_ROOTS 64

struct OctShared {
  OctState S;
  int root_cnt[OCT_MAX_ROOTS];
  int n;         // candidates of this level
  int finished;  // broadcast copy
};

// Dynamic shared memory carve-up (see orbx_octree_smem_bytes):
//   OctShared | nodes[node_cap] | vec[node_cap] | vec2[node_cap] | divlist[node_cap]
//   | best[node_cap] | cell_off[max_cells+1] | keys[smem_keys] | labels[smem_keys]
size_t orbx_octree_smem_bytes(int smem_keys, int node_cap, int max_cells) {
  size_t b = (sizeof(OctShared) + 15) & ~(size_t)15;
  b += ((size_t)node_cap * sizeof(OctNode) + 15) & ~(size_t)15;
  b += 2 * (size_t)node_cap * sizeof(uint64_t);
  b += (size_t)node_cap * sizeof(int);
  b += (size_t)node_cap * sizeof(uint32_t);
  b += (((size_t)max_cells + 1) * sizeof(int) + 15) & ~(size_t)15;
  b += (size_t)smem_keys * sizeof(uint32_t);
  b += (size_t)smem_keys * sizeof(int16_t);
  return (b + 15) & ~(size_t)15;
}

__global__ void __launch_bounds__(OCT_THREADS)
k_octree(const __grid_constant__ DevPlan P, const int32_t* __restrict__ cell_cnt,
         const uint32_t* __restrict__ cand, uint32_t* __restrict__ gkeys, int16_t* __restrict__ glabels,
         uint32_t* __restrict__ sel, int32_t* __restrict__ sel_cnt, int32_t* __restrict__ status,
         int smem_keys, int node_cap, int max_cells) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int level = blockIdx.x, frame = blockIdx.y;
  const DevLevel& L = P.lv[level];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  unsigned char* p = smem_raw;
  OctShared& sh = *reinterpret_cast<OctShared*>(p);
  p += (sizeof(OctShared) + 15) & ~(size_t)15;
  OctNode* nodes = reinterpret_cast<OctNode*>(p);
  p += ((size_t)node_cap * sizeof(OctNode) + 15) & ~(size_t)15;
  uint64_t* vec = reinterpret_cast<uint64_t*>(p);
  p += (size_t)node_cap * sizeof(uint64_t);
  uint64_t* vec2 = reinterpret_cast<uint64_t*>(p);
  p += (size_t)node_cap * sizeof(uint64_t);
  int* divlist = reinterpret_cast<int*>(p);
  p += (size_t)node_cap * sizeof(int);
  uint32_t* best = reinterpret_cast<uint32_t*>(p);
  p += (size_t)node_cap * sizeof(uint32_t);
  int* cell_off = reinterpret_cast<int*>(p);
  p += (((size_t)max_cells + 1) * sizeof(int) + 15) & ~(size_t)15;
  uint32_t* skeys = reinterpret_cast<uint32_t*>(p);
  p += (size_t)smem_keys * sizeof(uint32_t);
  int16_t* slabels = reinterpret_cast<int16_t*>(p);

  // ---- 1. exclusive scan of the per-cell counts (cell-major order = reference's input order)
  const int ncell = L.nCols * L.nRows;
  const int32_t* cc = cell_cnt + (long long)frame * P.total_cells + L.cell_base;
  __shared__ int warp_sums[OCT_THREADS / 32];
  __shared__ int carry_s;
  if (tid == 0) carry_s = 0;
  __syncthreads();
  for (int base = 0; base < ncell; base += OCT_THREADS) {
    const int c = base + tid;
    const int v = c < ncell ? cc[c] : 0;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    int woff = 0;
#pragma unroll
    for (int w = 0; w < OCT_THREADS / 32; w++)
      if (w < warp) woff += warp_sums[w];
    const int carry = carry_s;
    if (c < ncell) cell_off[c] = carry + woff + inc - v;
    __syncthreads();
    if (tid == OCT_THREADS - 1) carry_s = carry + woff + inc;
    __syncthreads();
  }
  const int n = carry_s;
  if (tid == 0) { cell_off[ncell] = n; sh.n = n; }
  for (int i = tid; i < OCT_MAX_ROOTS; i += OCT_THREADS) sh.root_cnt[i] = 0;
  __syncthreads();

  // ---- 2. gather keys (warp per cell), label with the root node, count keys per root
  const long long lvl_off = (long long)frame * P.cand_total + L.cand_base;
  uint32_t* keys = n <= smem_keys ? skeys : gkeys + lvl_off;
  int16_t* labels = n <= smem_keys ? slabels : glabels + lvl_off;
  const uint32_t* lc = cand + lvl_off;
  for (int c = warp; c < ncell; c += OCT_THREADS / 32) {
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

  // ---- 3. rounds
  if (tid == 0) {
    OctState& S = sh.S;
    S.nodes = nodes; S.node_cap = node_cap; S.vec = vec; S.vec2 = vec2; S.divlist = divlist; S.N = L.nfeat;
    oct_init(S, L.nIni, L.hX, L.h - 2 * ORBX_MIN_BORDER, sh.root_cnt);
    sh.finished = 0;
  }
  __syncthreads();
  while (true) {
    // (1) quadrant counts of candidate nodes
    for (int i = tid; i < n; i += OCT_THREADS) {
      OctNode& nd = nodes[labels[i]];
      if (nd.cand) {
        const uint32_t k = keys[i];
        atomicAdd(&nd.c[oct_quadrant(nd, ORBX_KEY_X(k), ORBX_KEY_Y(k))], 1);
      }
    }
    __syncthreads();
    // (2) list surgery
    if (tid == 0) {
      oct_round(sh.S);
      sh.finished = sh.S.finished;
    }
    __syncthreads();
    // (3) relabel keys of divided nodes
    for (int i = tid; i < n; i += OCT_THREADS) {
      const OctNode& nd = nodes[labels[i]];
      if (nd.divided) {
        const uint32_t k = keys[i];
        labels[i] = (int16_t)nd.c[oct_quadrant(nd, ORBX_KEY_X(k), ORBX_KEY_Y(k))];
      }
    }
    const int fin = sh.finished;
    __syncthreads();
    if (fin) break;
  }

  // ---- 4. best response per node, first key wins ties (:757-778)
  for (int i = tid; i < node_cap; i += OCT_THREADS) best[i] = 0;
  __syncthreads();
  for (int i = tid; i < n; i += OCT_THREADS)
    atomicMax(&best[labels[i]], ((uint32_t)ORBX_KEY_S(keys[i]) << 24) | (0xFFFFFFu - (uint32_t)i));
  __syncthreads();
  // list order -> rank (serial walk), then parallel emit
  int* order = divlist;
  if (tid == 0) {
    int r = 0;
    for (int it = sh.S.head; it >= 0; it = nodes[it].next) order[r++] = it;
    sel_cnt[frame * P.nlevels + level] = sh.S.overflow ? 0 : r;
    if (sh.S.overflow) atomicOr(status, 1);
    sh.n = sh.S.overflow ? 0 : r;
  }
  __syncthreads();
  const int nsel = sh.n;
  uint32_t* out = sel + (long long)frame * P.kp_total_cap + L.kp_base;
  for (int r = tid; r < nsel; r += OCT_THREADS) {
    const int i = (int)(0xFFFFFFu - (best[order[r]] & 0xFFFFFFu));
    const uint32_t k = keys[i];
    // back to level coordinates (:883-884)
    out[r] = orbx_pack_key(ORBX_KEY_X(k) + ORBX_MIN_BORDER, ORBX_KEY_Y(k) + ORBX_MIN_BORDER, ORBX_KEY_S(k));
  }
}

cudaError_t orbx_octree_set_smem(size_t bytes) {
  return cudaFuncSetAttribute(k_octree, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
}

cudaError_t orbx_launch_octree(const DevPlan& P, const int32_t* cell_cnt, const uint32_t* cand,
                               uint32_t* gkeys, int16_t* glabels, uint32_t* sel, int32_t* sel_cnt,
                               int32_t* status, int batch, size_t smem_bytes, int smem_keys, int node_cap,
                               int max_cells, cudaStream_t st) {
  dim3 grd(P.nlevels, batch);
  k_octree<<<grd, OCT_THREADS, smem_bytes, st>>>(P, cell_cnt, cand, gkeys, glabels, sel, sel_cnt, status,
                                                 smem_keys, node_cap, max_cells);
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
