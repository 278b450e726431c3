// bow.cu — vocabulary-node matching for sm_100a: ORBmatcher::SearchByBoW (src/ORBmatcher.cc:223-425, 765-905),
// ORBmatcher::SearchForTriangulation (:907-1146, Pinhole::epipolarConstrain CameraModels/Pinhole.cpp:107-129),
// DBoW2 transform (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1127-1259) and the all-pairs median of
// MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:367-401).
//
//   k_bow_nodes     one warp per vocabulary node of side 1.  The merge walk of the two FeatureVector maps visits
//                   the node ids present in both, so the warp binary-searches its node id in side 2's sorted node
//                   list.  Features of a node only ever meet features of the same node, hence the reference's
//                   sequential "already matched" skip (:278-279, :823) is a dependence INSIDE one warp: side-1
//                   features of the node are taken in order, lanes score the node's side-2 features.
//   k_tri_nodes     same partition for SearchForTriangulation (one CTA per node, its warps share the node's
//                   queries); vbMatched2 is never set there, so the queries are independent: lanes score candidates, `dist<=bestDist` (last minimum wins, :1017) is a min over
//                   (dist, ~position) among the candidates that pass the epipole and epipolar-line gates.
//   k_bow_rot       one CTA per pair: rotation histogram, ComputeThreeMaxima (:2012-2053), erasure, nmatches.
//   k_bow_descend   one thread per feature: 10-ary tree descent with first-minimum child (FORB::distance).
//   k_bow_vectors   one CTA per frame: bitonic sort of (node, feature) / (word, feature) keys in shared memory,
//                   FeatureVector CSR, BowVector sums in feature order and the L1 norm in word order (fp64,
//                   sequential where the reference's summation order is).
//   k_distinctive   one warp per MapPoint: per row a 257-bin histogram of distances -> order statistic.
#include "../../include/orbx.h"
#include "orbx_kernels.cuh"
#include "orbx_core.h"
#include "match_core.cuh"
#include "tri_core.h"

#define BOW_WARPS 4
#define FULLMASK 0xffffffffu

struct BowSide {
  const OrbxKeyPointDev* kps;
  const uint8_t* desc;
  const int32_t* n;        // [batch]
  const uint8_t* flag;     // valid / has-map-point, may be NULL
  const float* u_right;    // triangulation only, may be NULL
  const int32_t* fv_node;  // [batch][node_stride]
  const int32_t* fv_off;   // [batch][node_stride+1]
  const int32_t* fv_idx;   // [batch][kp_stride]
  const int32_t* fv_nn;    // [batch]
};

__device__ __forceinline__ int bow_find_node(const int32_t* nodes, int nn, int id) {
  int lo = 0, hi = nn;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(nodes + mid) < id) lo = mid + 1; else hi = mid;
  }
  return (lo < nn && __ldg(nodes + lo) == id) ? lo : -1;
}

__device__ __forceinline__ int bow_rot_bin(float a1, float a2) {
  float rot = f_sub(a1, a2);
  if (rot < 0.0f) rot = f_add(rot, 360.0f);
  int bin = (int)roundf(f_mul(rot, 1.0f / 30.0f));
  if (bin == 30) bin = 0;
  return bin;
}

// mode 0: KeyFrame x Frame (best <= TH_LOW); mode 1: KeyFrame x KeyFrame (best < TH_LOW, side-2 validity).
__global__ void __launch_bounds__(BOW_WARPS * 32)
k_bow_nodes(BowSide S1, BowSide S2, int kp_stride, int node_stride, int mode, float nnratio,
            int32_t* __restrict__ match12, int32_t* __restrict__ match21) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, p = blockIdx.y;
  const int a = blockIdx.x * BOW_WARPS + warp;
  if (a >= S1.fv_nn[p]) return;
  const size_t nb = (size_t)p * node_stride, fb = (size_t)p * kp_stride;
  const int node = S1.fv_node[nb + a];
  const int b = bow_find_node(S2.fv_node + nb, S2.fv_nn[p], node);
  if (b < 0) return;
  const int o1 = S1.fv_off[nb + p + a], e1 = S1.fv_off[nb + p + a + 1];
  const int o2 = S2.fv_off[nb + p + b], e2 = S2.fv_off[nb + p + b + 1];
  const int32_t* idx1s = S1.fv_idx + fb;
  const int32_t* idx2s = S2.fv_idx + fb;
  volatile int32_t* m21 = match21 + fb;
  const int TH_LOW = 50;
  for (int i1 = o1; i1 < e1; i1++) {
    const int idx1 = idx1s[i1];
    if (S1.flag && !S1.flag[fb + idx1]) continue;
    const Desc256 d1 = load_desc(S1.desc + (fb + idx1) * 32);
    uint32_t k1 = KEY_NONE; int d2 = 256;  // lane-local best key (dist<<16|pos) and second distance
    for (int c = o2 + lane; c < e2; c += 32) {
      const int idx2 = idx2s[c];
      if (m21[idx2] >= 0) continue;
      if (mode == 1 && S2.flag && !S2.flag[fb + idx2]) continue;
      const int dist = hamming(d1, load_desc(S2.desc + (fb + idx2) * 32));
      const uint32_t key = ((uint32_t)dist << 16) | (uint32_t)((c - o2) & 0xffff);
      if (key < k1) { d2 = min(d2, (int)(k1 >> 16)); k1 = key; }
      else d2 = min(d2, dist);
    }
    if (k1 == KEY_NONE) d2 = 256; else d2 = min(d2, 256);
    uint32_t best = k1;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(FULLMASK, best, o));
    if (best == KEY_NONE) continue;
    // second smallest of the multiset: the winner lane offers its own runner-up, the others their best
    int sec = (k1 == best) ? d2 : (k1 == KEY_NONE ? 256 : (int)(k1 >> 16));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sec = min(sec, __shfl_xor_sync(FULLMASK, sec, o));
    const int bestDist1 = (int)(best >> 16), bestDist2 = sec;
    const bool pass = mode == 0 ? bestDist1 <= TH_LOW : bestDist1 < TH_LOW;
    if (pass && (float)bestDist1 < f_mul(nnratio, (float)bestDist2)) {
      const int idx2 = idx2s[o2 + (int)(best & 0xffff)];
      if (lane == 0) { match12[fb + idx1] = idx2; m21[idx2] = idx1; }
    }
    __syncwarp();
  }
}

// SearchByBoW(KeyFrame*, Frame&) on a two-camera Frame (F.Nleft != -1; ORBmatcher.cc:295-323, 356-380): side 2 holds
// the left features [0, Nleft) and the right features [Nleft, N) in one numbering; every side-1 feature keeps a
// separate (best, second) pair per camera.  The left match needs bestDist1 <= TH_LOW and the ratio test; the right
// match is taken whenever bestDist1 <= TH_LOW and bestDist1R <= TH_LOW (`|| true` at :358).
__global__ void __launch_bounds__(BOW_WARPS * 32)
k_bow_nodes_rig(BowSide S1, BowSide S2, const int32_t* __restrict__ nleft2, int kp_stride, int node_stride,
                float nnratio, int32_t* __restrict__ match12l, int32_t* __restrict__ match12r,
                int32_t* __restrict__ match21) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, p = blockIdx.y;
  const int a = blockIdx.x * BOW_WARPS + warp;
  if (a >= S1.fv_nn[p]) return;
  const size_t nb = (size_t)p * node_stride, fb = (size_t)p * kp_stride;
  const int node = S1.fv_node[nb + a];
  const int b = bow_find_node(S2.fv_node + nb, S2.fv_nn[p], node);
  if (b < 0) return;
  const int o1 = S1.fv_off[nb + p + a], e1 = S1.fv_off[nb + p + a + 1];
  const int o2 = S2.fv_off[nb + p + b], e2 = S2.fv_off[nb + p + b + 1];
  const int32_t* idx1s = S1.fv_idx + fb;
  const int32_t* idx2s = S2.fv_idx + fb;
  volatile int32_t* m21 = match21 + fb;
  const int nleft = nleft2[p];
  const int TH_LOW = 50;
  for (int i1 = o1; i1 < e1; i1++) {
    const int idx1 = idx1s[i1];
    if (S1.flag && !S1.flag[fb + idx1]) continue;
    const Desc256 d1 = load_desc(S1.desc + (fb + idx1) * 32);
    uint32_t kL = KEY_NONE, kR = KEY_NONE; int sL = 256, sR = 256;  // lane-local best key / second distance per camera
    for (int c = o2 + lane; c < e2; c += 32) {
      const int idx2 = idx2s[c];
      if (m21[idx2] >= 0) continue;
      const int dist = hamming(d1, load_desc(S2.desc + (fb + idx2) * 32));
      const uint32_t key = ((uint32_t)dist << 16) | (uint32_t)((c - o2) & 0xffff);
      if (idx2 < nleft) {
        if (key < kL) { sL = min(sL, (int)(kL >> 16)); kL = key; } else sL = min(sL, dist);
      } else {
        if (key < kR) { sR = min(sR, (int)(kR >> 16)); kR = key; } else sR = min(sR, dist);
      }
    }
    if (kL == KEY_NONE) sL = 256; else sL = min(sL, 256);
    if (kR == KEY_NONE) sR = 256; else sR = min(sR, 256);
    uint32_t bestL = kL, bestR = kR;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      bestL = min(bestL, __shfl_xor_sync(FULLMASK, bestL, o));
      bestR = min(bestR, __shfl_xor_sync(FULLMASK, bestR, o));
    }
    if (bestL == KEY_NONE) { __syncwarp(); continue; }  // bestDist1 = 256 > TH_LOW: neither camera is written (:327)
    int secL = (kL == bestL) ? sL : (kL == KEY_NONE ? 256 : (int)(kL >> 16));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) secL = min(secL, __shfl_xor_sync(FULLMASK, secL, o));
    const int bestDist1 = (int)(bestL >> 16), bestDist2 = secL;
    if (bestDist1 <= TH_LOW) {
      if ((float)bestDist1 < f_mul(nnratio, (float)bestDist2)) {
        const int idx2 = idx2s[o2 + (int)(bestL & 0xffff)];
        if (lane == 0) { match12l[fb + idx1] = idx2; m21[idx2] = idx1; }
      }
      if (bestR != KEY_NONE && (int)(bestR >> 16) <= TH_LOW) {
        const int idx2 = idx2s[o2 + (int)(bestR & 0xffff)];
        if (lane == 0) { match12r[fb + idx1] = idx2; m21[idx2] = idx1; }
      }
    }
    __syncwarp();
  }
}

// Rotation histogram of the rig variant: the entries are the written side-2 features (rotHist[bin].push_back(
// bestIdxF / bestIdxFR), :352, :376), so the sweep runs over match21.
__global__ void __launch_bounds__(256)
k_bow_rot_rig(const OrbxKeyPointDev* __restrict__ k1, const OrbxKeyPointDev* __restrict__ k2,
              const int32_t* __restrict__ n2_arr, int kp_stride, int check_orientation, int32_t* __restrict__ match12l,
              int32_t* __restrict__ match12r, int32_t* __restrict__ match21, int32_t* __restrict__ nmatches) {
  __shared__ int hist[30];
  __shared__ int s_ind[3];
  __shared__ int s_cnt;
  const int p = blockIdx.x, n2 = n2_arr[p];
  const size_t fb = (size_t)p * kp_stride;
  if (threadIdx.x < 30) hist[threadIdx.x] = 0;
  if (threadIdx.x == 0) s_cnt = 0;
  __syncthreads();
  int cnt = 0;
  for (int j = threadIdx.x; j < n2; j += blockDim.x) {
    const int m = match21[fb + j];
    if (m < 0) continue;
    cnt++;
    if (check_orientation) atomicAdd(&hist[bow_rot_bin(k1[fb + m].angle, k2[fb + j].angle)], 1);
  }
  __syncthreads();
  if (check_orientation) {
    if (threadIdx.x == 0) {
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
    __syncthreads();
    const int ind1 = s_ind[0], ind2 = s_ind[1], ind3 = s_ind[2];
    for (int j = threadIdx.x; j < n2; j += blockDim.x) {
      const int m = match21[fb + j];
      if (m < 0) continue;
      const int bin = bow_rot_bin(k1[fb + m].angle, k2[fb + j].angle);
      if (bin != ind1 && bin != ind2 && bin != ind3) {
        match21[fb + j] = -1;
        if (match12l[fb + m] == j) match12l[fb + m] = -1;  // the left and the right entry of m are distinct features
        if (match12r[fb + m] == j) match12r[fb + m] = -1;
        cnt--;
      }
    }
  }
  atomicAdd(&s_cnt, cnt);
  __syncthreads();
  if (threadIdx.x == 0) nmatches[p] = s_cnt;
}

struct TriGeom {       // per pair, host geometry
  const float* F12;    // [batch][9] row-major
  const float* ep;     // [batch][2]
  float scale2[ORBX_MAX_LEVELS], sigma2_2[ORBX_MAX_LEVELS];
  int only_stereo, coarse;
};

__global__ void __launch_bounds__(BOW_WARPS * 32)
k_tri_nodes(BowSide S1, BowSide S2, TriGeom G, int kp_stride, int node_stride, int32_t* __restrict__ match12) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, p = blockIdx.y;
  const int a = blockIdx.x;  // one CTA per node: the queries are independent, its warps take them round-robin
  if (a >= S1.fv_nn[p]) return;
  const size_t nb = (size_t)p * node_stride, fb = (size_t)p * kp_stride;
  const int node = S1.fv_node[nb + a];
  const int b = bow_find_node(S2.fv_node + nb, S2.fv_nn[p], node);
  if (b < 0) return;
  const int o1 = S1.fv_off[nb + p + a], e1 = S1.fv_off[nb + p + a + 1];
  const int o2 = S2.fv_off[nb + p + b], e2 = S2.fv_off[nb + p + b + 1];
  const int32_t* idx1s = S1.fv_idx + fb;
  const int32_t* idx2s = S2.fv_idx + fb;
  float F[9];
#pragma unroll
  for (int i = 0; i < 9; i++) F[i] = G.coarse ? 0.f : __ldg(G.F12 + (size_t)p * 9 + i);
  const float epx = __ldg(G.ep + 2 * p), epy = __ldg(G.ep + 2 * p + 1);
  const int TH_LOW = 50;
  for (int i1 = o1 + warp; i1 < e1; i1 += BOW_WARPS) {
    const int idx1 = idx1s[i1];
    if (S1.flag && S1.flag[fb + idx1]) continue;
    const bool st1 = S1.u_right && S1.u_right[fb + idx1] >= 0.0f;
    if (G.only_stereo && !st1) continue;
    const OrbxKeyPointDev kp1 = S1.kps[fb + idx1];
    const Desc256 d1 = load_desc(S1.desc + (fb + idx1) * 32);
    // epipolar line of kp1 in image 2 (Pinhole.cpp:115-117), source-order fp32
    const float la = f_add(f_add(f_mul(kp1.x, F[0]), f_mul(kp1.y, F[3])), F[6]);
    const float lb = f_add(f_add(f_mul(kp1.x, F[1]), f_mul(kp1.y, F[4])), F[7]);
    const float lc = f_add(f_add(f_mul(kp1.x, F[2]), f_mul(kp1.y, F[5])), F[8]);
    const float den = f_add(f_mul(la, la), f_mul(lb, lb));
    uint32_t k1 = KEY_NONE;
    for (int c = o2 + lane; c < e2; c += 32) {
      const int idx2 = idx2s[c];
      if (S2.flag && S2.flag[fb + idx2]) continue;
      const bool st2 = S2.u_right && S2.u_right[fb + idx2] >= 0.0f;
      if (G.only_stereo && !st2) continue;
      const int dist = hamming(d1, load_desc(S2.desc + (fb + idx2) * 32));
      if (dist > TH_LOW) continue;
      const OrbxKeyPointDev kp2 = S2.kps[fb + idx2];
      if (!st1 && !st2) {
        const float dx = f_sub(epx, kp2.x), dy = f_sub(epy, kp2.y);
        if (f_add(f_mul(dx, dx), f_mul(dy, dy)) < f_mul(100.0f, G.scale2[kp2.octave])) continue;
      }
      if (!G.coarse) {
        const float num = f_add(f_add(f_mul(la, kp2.x), f_mul(lb, kp2.y)), lc);
        if (den == 0.0f) continue;
        const float dsqr = f_div(f_mul(num, num), den);
        if (!((double)dsqr < d_mul(3.84, (double)G.sigma2_2[kp2.octave]))) continue;
      }
      const uint32_t key = ((uint32_t)dist << 16) | (uint32_t)(0xffff - ((c - o2) & 0xffff));
      k1 = min(k1, key);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) k1 = min(k1, __shfl_xor_sync(FULLMASK, k1, o));
    if (k1 != KEY_NONE && lane == 0) match12[fb + idx1] = idx2s[o2 + (0xffff - (int)(k1 & 0xffff))];
  }
}

// SearchForTriangulation between two keyframes of a two-camera rig (pKF->mpCamera2 != NULL; ORBmatcher.cc:907-1146 with
// the camera selection :1036-1062): features [0, NLeft) belong to the left camera, the rest to the right one; there is
// no mvuRight stereo (bStereo1/2 = false, :973,:998: with bOnlyStereo every feature is skipped) and no epipole gate
// (:1022 needs !mpCamera2); each candidate pair picks (R12, t12, camera of KF1, camera of KF2) from the four
// left/right combinations and passes KannalaBrandt8::epipolarConstrain (tri_core.h) unless bCoarse.
struct TriRig {
  const OrbxRigPose* pose;   // [batch][4]: index 2*bRight1 + bRight2 = (Tll, Tlr, Trl, Trr) as (R12, t12)
  const int32_t *nleft1, *nleft2;
  OrbxFrustumCam cam1[2], cam2[2];  // [0] = mpCamera, [1] = mpCamera2 of KF1 / KF2
  float sigma2_1[ORBX_MAX_LEVELS], sigma2_2[ORBX_MAX_LEVELS];
  int only_stereo, coarse;
};

__global__ void __launch_bounds__(BOW_WARPS * 32)
k_tri_nodes_rig(BowSide S1, BowSide S2, const __grid_constant__ TriRig G, int kp_stride, int node_stride,
                int32_t* __restrict__ match12) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, p = blockIdx.y;
  const int a = blockIdx.x;
  if (a >= S1.fv_nn[p] || G.only_stereo) return;
  const size_t nb = (size_t)p * node_stride, fb = (size_t)p * kp_stride;
  const int node = S1.fv_node[nb + a];
  const int b = bow_find_node(S2.fv_node + nb, S2.fv_nn[p], node);
  if (b < 0) return;
  const int o1 = S1.fv_off[nb + p + a], e1 = S1.fv_off[nb + p + a + 1];
  const int o2 = S2.fv_off[nb + p + b], e2 = S2.fv_off[nb + p + b + 1];
  const int32_t* idx1s = S1.fv_idx + fb;
  const int32_t* idx2s = S2.fv_idx + fb;
  const int nl1 = G.nleft1[p], nl2 = G.nleft2[p];
  const int TH_LOW = 50;
  for (int i1 = o1 + warp; i1 < e1; i1 += BOW_WARPS) {
    const int idx1 = idx1s[i1];
    if (S1.flag && S1.flag[fb + idx1]) continue;
    const OrbxKeyPointDev kp1 = S1.kps[fb + idx1];
    const int right1 = idx1 >= nl1 ? 1 : 0;
    const Desc256 d1 = load_desc(S1.desc + (fb + idx1) * 32);
    uint32_t k1 = KEY_NONE;
    for (int c = o2 + lane; c < e2; c += 32) {
      const int idx2 = idx2s[c];
      if (S2.flag && S2.flag[fb + idx2]) continue;
      const int dist = hamming(d1, load_desc(S2.desc + (fb + idx2) * 32));
      if (dist > TH_LOW) continue;
      if (!G.coarse) {
        const OrbxKeyPointDev kp2 = S2.kps[fb + idx2];
        const int right2 = idx2 >= nl2 ? 1 : 0;
        if (!orbx_kb8_epipolar_constrain(G.cam1[right1], G.cam2[right2], kp1.x, kp1.y, kp2.x, kp2.y,
                                         G.pose[(size_t)p * 4 + 2 * right1 + right2], G.sigma2_1[kp1.octave],
                                         G.sigma2_2[kp2.octave]))
          continue;
      }
      const uint32_t key = ((uint32_t)dist << 16) | (uint32_t)(0xffff - ((c - o2) & 0xffff));
      k1 = min(k1, key);  // dist <= bestDist: the last minimum wins (:1017)
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) k1 = min(k1, __shfl_xor_sync(FULLMASK, k1, o));
    if (k1 != KEY_NONE && lane == 0) match12[fb + idx1] = idx2s[o2 + (0xffff - (int)(k1 & 0xffff))];
  }
}

// Rotation histogram + ComputeThreeMaxima + erasure for one pair; also counts the matches.
__global__ void __launch_bounds__(256)
k_bow_rot(const OrbxKeyPointDev* __restrict__ k1, const OrbxKeyPointDev* __restrict__ k2,
          const int32_t* __restrict__ n1_arr, int kp_stride, int check_orientation,
          int32_t* __restrict__ match12, int32_t* __restrict__ match21, int32_t* __restrict__ nmatches) {
  __shared__ int hist[30];
  __shared__ int s_ind[3];
  __shared__ int s_cnt;
  const int p = blockIdx.x, n1 = n1_arr[p];
  const size_t fb = (size_t)p * kp_stride;
  if (threadIdx.x < 30) hist[threadIdx.x] = 0;
  if (threadIdx.x == 0) s_cnt = 0;
  __syncthreads();
  int cnt = 0;
  for (int i = threadIdx.x; i < n1; i += blockDim.x) {
    const int m = match12[fb + i];
    if (m < 0) continue;
    cnt++;
    if (check_orientation) atomicAdd(&hist[bow_rot_bin(k1[fb + i].angle, k2[fb + m].angle)], 1);
  }
  __syncthreads();
  if (check_orientation) {
    if (threadIdx.x == 0) {
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
    __syncthreads();
    const int ind1 = s_ind[0], ind2 = s_ind[1], ind3 = s_ind[2];
    for (int i = threadIdx.x; i < n1; i += blockDim.x) {
      const int m = match12[fb + i];
      if (m < 0) continue;
      const int bin = bow_rot_bin(k1[fb + i].angle, k2[fb + m].angle);
      if (bin != ind1 && bin != ind2 && bin != ind3) {
        match12[fb + i] = -1;
        if (match21) match21[fb + m] = -1;
        cnt--;
      }
    }
  }
  atomicAdd(&s_cnt, cnt);
  __syncthreads();
  if (threadIdx.x == 0) nmatches[p] = s_cnt;
}

// ---------------------------------------------------------------------------------------------------------
// DBoW2 transform
// ---------------------------------------------------------------------------------------------------------
struct BowVoc {
  const int32_t* child_off;   // [nnodes+1]
  const int32_t* child_idx;
  const uint8_t* node_desc;   // [nnodes][32]
  const int32_t* node_word;   // [nnodes]
  const double* node_weight;  // [nnodes]
  int L, levelsup;
};

__global__ void __launch_bounds__(128)
k_bow_descend(BowVoc V, const uint8_t* __restrict__ desc, const int32_t* __restrict__ n_arr, int kp_stride,
              int32_t* __restrict__ feat_word, int32_t* __restrict__ feat_node, double* __restrict__ feat_weight) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x, p = blockIdx.y;
  if (f >= n_arr[p]) return;
  const size_t o = (size_t)p * kp_stride + f;
  const Desc256 d = load_desc(desc + o * 32);
  const int nid_level = V.L - V.levelsup;
  int nid = 0, cur = 0, level = 0;
  int c0 = __ldg(V.child_off), c1 = __ldg(V.child_off + 1);
  do {  // TemplatedVocabulary.h:1224-1247: first child, then `d < best_d`
    ++level;
    int best = __ldg(V.child_idx + c0);
    int best_d = hamming(d, load_desc(V.node_desc + (size_t)best * 32));
    for (int c = c0 + 1; c < c1; c++) {
      const int id = __ldg(V.child_idx + c);
      const int dd = hamming(d, load_desc(V.node_desc + (size_t)id * 32));
      if (dd < best_d) { best_d = dd; best = id; }
    }
    cur = best;
    if (level == nid_level) nid = cur;
    c0 = __ldg(V.child_off + cur); c1 = __ldg(V.child_off + cur + 1);
  } while (c0 != c1);
  feat_word[o] = __ldg(V.node_word + cur);
  feat_node[o] = nid;
  feat_weight[o] = __ldg(V.node_weight + cur);
}

#define BV_THREADS 512
__device__ void bv_sort(unsigned long long* keys, int np2) {  // ascending bitonic sort in shared memory
  for (int k = 2; k <= np2; k <<= 1)
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < np2; i += BV_THREADS) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const unsigned long long x = keys[i], y = keys[ixj];
          const bool up = (i & k) == 0;
          if ((x > y) == up) { keys[i] = y; keys[ixj] = x; }
        }
      }
      __syncthreads();
    }
}

// Exclusive block scan of one int per thread (BV_THREADS threads); returns the prefix, *total = sum.
__device__ int bv_scan(int v, int* s_warp, int* total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(FULLMASK, inc, o); if (lane >= o) inc += t; }
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    int w = lane < BV_THREADS / 32 ? s_warp[lane] : 0, wi = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(FULLMASK, wi, o); if (lane >= o) wi += t; }
    if (lane < BV_THREADS / 32) s_warp[lane] = wi - w;
    if (lane == 31) s_warp[32] = wi;
  }
  __syncthreads();
  const int r = s_warp[warp] + inc - v;
  *total = s_warp[32];
  __syncthreads();
  return r;
}

__global__ void __launch_bounds__(BV_THREADS)
k_bow_vectors(const int32_t* __restrict__ n_arr, int kp_stride, int node_stride, int np2,
              const int32_t* __restrict__ feat_word, const int32_t* __restrict__ feat_node,
              const double* __restrict__ feat_weight, int32_t* __restrict__ fv_node, int32_t* __restrict__ fv_off,
              int32_t* __restrict__ fv_idx, int32_t* __restrict__ fv_nn, int32_t* __restrict__ bow_word,
              double* __restrict__ bow_val, int32_t* __restrict__ bow_n) {
  extern __shared__ unsigned long long keys[];
  __shared__ int s_warp[33];
  __shared__ double s_norm;
  const int p = blockIdx.x, n = n_arr[p];
  const size_t fb = (size_t)p * kp_stride, nb = (size_t)p * node_stride;
  const int ch = np2 / BV_THREADS > 0 ? np2 / BV_THREADS : 1;  // np2 >= BV_THREADS by construction
  const int lo = threadIdx.x * ch, hi = lo + ch;
  const unsigned long long NONE = ~0ull;
  // ---- FeatureVector: (node, feature) ascending; stopped words (weight <= 0) take no part (:1159-1163)
  for (int i = threadIdx.x; i < np2; i += BV_THREADS)
    keys[i] = (i < n && feat_weight[fb + i] > 0.0) ? ((unsigned long long)(uint32_t)feat_node[fb + i] << 32) | (uint32_t)i : NONE;
  __syncthreads();
  bv_sort(keys, np2);
  int heads = 0, valid = 0;
  for (int i = lo; i < hi; i++) {
    const unsigned long long k = keys[i];
    if (k == NONE) break;
    valid++;
    if (i == 0 || (keys[i - 1] >> 32) != (k >> 32)) heads++;
  }
  int total;
  int rank = bv_scan(heads, s_warp, &total);
  for (int i = lo; i < lo + valid; i++) {
    const unsigned long long k = keys[i];
    fv_idx[fb + i] = (int)(uint32_t)k;
    if (i == 0 || (keys[i - 1] >> 32) != (k >> 32)) {
      if (rank < node_stride) { fv_node[nb + rank] = (int)(k >> 32); fv_off[nb + p + rank] = i; }
      rank++;
    }
  }
  int nvalid;
  const int vpre = bv_scan(valid, s_warp, &nvalid);
  (void)vpre;
  if (threadIdx.x == 0) {
    fv_nn[p] = total <= node_stride ? total : -total;
    fv_off[nb + p + (total <= node_stride ? total : node_stride)] = nvalid;
  }
  __syncthreads();
  // ---- BowVector: (word, feature) ascending; v[word] += w in feature order (BowVector.cpp:33-45)
  for (int i = threadIdx.x; i < np2; i += BV_THREADS)
    keys[i] = (i < n && feat_weight[fb + i] > 0.0) ? ((unsigned long long)(uint32_t)feat_word[fb + i] << 32) | (uint32_t)i : NONE;
  __syncthreads();
  bv_sort(keys, np2);
  heads = 0;
  for (int i = lo; i < hi; i++) {
    const unsigned long long k = keys[i];
    if (k == NONE) break;
    if (i == 0 || (keys[i - 1] >> 32) != (k >> 32)) heads++;
  }
  rank = bv_scan(heads, s_warp, &total);
  for (int i = lo; i < hi; i++) {
    const unsigned long long k = keys[i];
    if (k == NONE) break;
    if (i == 0 || (keys[i - 1] >> 32) != (k >> 32)) {
      double s = feat_weight[fb + (uint32_t)k];
      for (int j = i + 1; j < np2 && (keys[j] >> 32) == (k >> 32); j++) s = d_add(s, feat_weight[fb + (uint32_t)keys[j]]);
      bow_word[fb + rank] = (int)(k >> 32);
      bow_val[fb + rank] = s;
      rank++;
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {  // L1 norm in map (= word) order, BowVector.cpp:66-70
    double norm = 0.0;
    for (int i = 0; i < total; i++) norm = d_add(norm, fabs(bow_val[fb + i]));
    s_norm = norm;
    bow_n[p] = total;
  }
  __syncthreads();
  const double norm = s_norm;
  if (norm > 0.0)
    for (int i = threadIdx.x; i < total; i += BV_THREADS) bow_val[fb + i] = __ddiv_rn(bow_val[fb + i], norm);
}

// ---------------------------------------------------------------------------------------------------------
// MapPoint::ComputeDistinctiveDescriptors: least median row of the all-pairs distance matrix
// ---------------------------------------------------------------------------------------------------------
#define DD_BINS 288
__global__ void __launch_bounds__(BOW_WARPS * 32)
k_distinctive(const uint8_t* __restrict__ pool, const int32_t* __restrict__ obs_off,
              const int32_t* __restrict__ obs_idx, int m_points, int32_t* __restrict__ best_out,
              int32_t* __restrict__ median_out) {
  __shared__ int hist[BOW_WARPS][DD_BINS];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int mp = blockIdx.x * BOW_WARPS + warp;
  if (mp >= m_points) return;
  const int o = obs_off[mp], N = obs_off[mp + 1] - o;
  int* h = hist[warp];
  for (int i = lane; i < DD_BINS; i += 32) h[i] = 0;
  __syncwarp();
  int bestMedian = 0x7fffffff, bestIdx = 0;
  const int kth = (N - 1) / 2;  // vDists[0.5*(N-1)]
  for (int i = 0; i < N; i++) {
    const Desc256 di = load_desc(pool + (size_t)obs_idx[o + i] * 32);
    for (int j = lane; j < N; j += 32)
      atomicAdd(&h[j == i ? 0 : hamming(di, load_desc(pool + (size_t)obs_idx[o + j] * 32))], 1);
    __syncwarp();
    int c[9], s = 0;
#pragma unroll
    for (int b = 0; b < 9; b++) { c[b] = h[lane * 9 + b]; s += c[b]; h[lane * 9 + b] = 0; }
    int inc = s;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(FULLMASK, inc, d); if (lane >= d) inc += t; }
    int med = -1, run = inc - s;
    if (run <= kth && kth < inc) {
#pragma unroll
      for (int b = 0; b < 9; b++) { if (med < 0 && kth < run + c[b]) med = lane * 9 + b; run += c[b]; }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) med = max(med, __shfl_xor_sync(FULLMASK, med, d));
    if (med < bestMedian) { bestMedian = med; bestIdx = i; }
    __syncwarp();
  }
  if (lane == 0) { best_out[mp] = N > 0 ? bestIdx : -1; if (median_out) median_out[mp] = N > 0 ? bestMedian : -1; }
}

// ---------------------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------------------
static BowSide make_side(const orbx_keypoint* kps, const uint8_t* desc, const int32_t* n, const uint8_t* flag,
                         const float* u_right, const orbm_feature_vector* fv) {
  BowSide s;
  s.kps = (const OrbxKeyPointDev*)kps; s.desc = desc; s.n = n; s.flag = flag; s.u_right = u_right;
  s.fv_node = fv->d_node; s.fv_off = fv->d_off; s.fv_idx = fv->d_idx; s.fv_nn = fv->d_nn;
  return s;
}
static bool fv_ok(const orbm_feature_vector* fv) { return fv && fv->d_node && fv->d_off && fv->d_idx && fv->d_nn; }

extern "C" int orbm_search_by_bow(int mode, const orbx_keypoint* d_kps1, const uint8_t* d_desc1,
                                  const int32_t* d_n1, const uint8_t* d_valid1, const orbm_feature_vector* fv1,
                                  const orbx_keypoint* d_kps2, const uint8_t* d_desc2, const int32_t* d_n2,
                                  const uint8_t* d_valid2, const orbm_feature_vector* fv2, int kp_stride,
                                  int node_stride, int batch, float nn_ratio, int check_orientation,
                                  int32_t* d_match12, int32_t* d_match21, int32_t* d_nmatches, void* stream) {
  if ((mode != ORBM_BOW_KF_FRAME && mode != ORBM_BOW_KF_KF) || !d_kps1 || !d_desc1 || !d_n1 || !fv_ok(fv1) ||
      !d_kps2 || !d_desc2 || !d_n2 || !fv_ok(fv2) || kp_stride < 1 || node_stride < 1 || batch < 1 || !d_match12 ||
      !d_match21 || !d_nmatches)
    return ORBX_E_INVALID;
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(d_match12, 0xff, (size_t)batch * kp_stride * sizeof(int32_t), st);
  cudaMemsetAsync(d_match21, 0xff, (size_t)batch * kp_stride * sizeof(int32_t), st);
  const BowSide S1 = make_side(d_kps1, d_desc1, d_n1, d_valid1, nullptr, fv1);
  const BowSide S2 = make_side(d_kps2, d_desc2, d_n2, d_valid2, nullptr, fv2);
  dim3 grd((node_stride + BOW_WARPS - 1) / BOW_WARPS, batch);
  k_bow_nodes<<<grd, BOW_WARPS * 32, 0, st>>>(S1, S2, kp_stride, node_stride, mode, nn_ratio, d_match12, d_match21);
  k_bow_rot<<<batch, 256, 0, st>>>(S1.kps, S2.kps, d_n1, kp_stride, check_orientation, d_match12, d_match21,
                                   d_nmatches);
  return cudaGetLastError() == cudaSuccess ? ORBX_OK : ORBX_E_CUDA;
}

extern "C" int orbm_search_by_bow_rig(const orbx_keypoint* d_kps1, const uint8_t* d_desc1, const int32_t* d_n1,
                                      const uint8_t* d_valid1, const orbm_feature_vector* fv1,
                                      const orbx_keypoint* d_kps2, const uint8_t* d_desc2, const int32_t* d_n2,
                                      const int32_t* d_nleft2, const orbm_feature_vector* fv2, int kp_stride,
                                      int node_stride, int batch, float nn_ratio, int check_orientation,
                                      int32_t* d_match12_left, int32_t* d_match12_right, int32_t* d_match21,
                                      int32_t* d_nmatches, void* stream) {
  if (!d_kps1 || !d_desc1 || !d_n1 || !fv_ok(fv1) || !d_kps2 || !d_desc2 || !d_n2 || !d_nleft2 || !fv_ok(fv2) ||
      kp_stride < 1 || node_stride < 1 || batch < 1 || !d_match12_left || !d_match12_right || !d_match21 || !d_nmatches)
    return ORBX_E_INVALID;
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(d_match12_left, 0xff, (size_t)batch * kp_stride * sizeof(int32_t), st);
  cudaMemsetAsync(d_match12_right, 0xff, (size_t)batch * kp_stride * sizeof(int32_t), st);
  cudaMemsetAsync(d_match21, 0xff, (size_t)batch * kp_stride * sizeof(int32_t), st);
  const BowSide S1 = make_side(d_kps1, d_desc1, d_n1, d_valid1, nullptr, fv1);
  const BowSide S2 = make_side(d_kps2, d_desc2, d_n2, nullptr, nullptr, fv2);
  dim3 grd((node_stride + BOW_WARPS - 1) / BOW_WARPS, batch);
  k_bow_nodes_rig<<<grd, BOW_WARPS * 32, 0, st>>>(S1, S2, d_nleft2, kp_stride, node_stride, nn_ratio, d_match12_left,
                                                 d_match12_right, d_match21);
  k_bow_rot_rig<<<batch, 256, 0, st>>>(S1.kps, S2.kps, d_n2, kp_stride, check_orientation, d_match12_left,
                                       d_match12_right, d_match21, d_nmatches);
  return cudaGetLastError() == cudaSuccess ? ORBX_OK : ORBX_E_CUDA;
}

extern "C" int orbm_search_for_triangulation(
    const orbx_keypoint* d_kps1, const uint8_t* d_desc1, const int32_t* d_n1, const uint8_t* d_has_mp1,
    const float* d_u_right1, const orbm_feature_vector* fv1, const orbx_keypoint* d_kps2, const uint8_t* d_desc2,
    const int32_t* d_n2, const uint8_t* d_has_mp2, const float* d_u_right2, const orbm_feature_vector* fv2,
    int kp_stride, int node_stride, int batch, const float* d_F12, const float* d_ep,
    const float* scale_factors2_host, const float* level_sigma2_2_host, int nlevels, int only_stereo, int coarse,
    int check_orientation, int32_t* d_match12, int32_t* d_nmatches, void* stream) {
  if (!d_kps1 || !d_desc1 || !d_n1 || !fv_ok(fv1) || !d_kps2 || !d_desc2 || !d_n2 || !fv_ok(fv2) || kp_stride < 1 ||
      node_stride < 1 || batch < 1 || !d_ep || (!coarse && !d_F12) || !scale_factors2_host ||
      !level_sigma2_2_host || nlevels < 1 || nlevels > ORBX_MAX_LEVELS || !d_match12 || !d_nmatches)
    return ORBX_E_INVALID;
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(d_match12, 0xff, (size_t)batch * kp_stride * sizeof(int32_t), st);
  const BowSide S1 = make_side(d_kps1, d_desc1, d_n1, d_has_mp1, d_u_right1, fv1);
  const BowSide S2 = make_side(d_kps2, d_desc2, d_n2, d_has_mp2, d_u_right2, fv2);
  TriGeom G;
  G.F12 = d_F12; G.ep = d_ep; G.only_stereo = only_stereo; G.coarse = coarse;
  for (int l = 0; l < ORBX_MAX_LEVELS; l++) {
    G.scale2[l] = l < nlevels ? scale_factors2_host[l] : 0.f;
    G.sigma2_2[l] = l < nlevels ? level_sigma2_2_host[l] : 0.f;
  }
  dim3 grd(node_stride, batch);
  k_tri_nodes<<<grd, BOW_WARPS * 32, 0, st>>>(S1, S2, G, kp_stride, node_stride, d_match12);
  k_bow_rot<<<batch, 256, 0, st>>>(S1.kps, S2.kps, d_n1, kp_stride, check_orientation, d_match12, nullptr,
                                   d_nmatches);
  return cudaGetLastError() == cudaSuccess ? ORBX_OK : ORBX_E_CUDA;
}

static OrbxFrustumCam to_cam(const orbm_camera& c) {
  OrbxFrustumCam o;
  o.model = c.model; o.fx = c.fx; o.fy = c.fy; o.cx = c.cx; o.cy = c.cy;
  for (int i = 0; i < 4; i++) o.k[i] = c.k[i];
  return o;
}

extern "C" int orbm_search_for_triangulation_rig(
    const orbx_keypoint* d_kps1, const uint8_t* d_desc1, const int32_t* d_n1, const int32_t* d_nleft1,
    const uint8_t* d_has_mp1, const orbm_feature_vector* fv1, const orbx_keypoint* d_kps2, const uint8_t* d_desc2,
    const int32_t* d_n2, const int32_t* d_nleft2, const uint8_t* d_has_mp2, const orbm_feature_vector* fv2,
    int kp_stride, int node_stride, int batch, const float* d_rig_pose, const orbm_camera* cams1,
    const orbm_camera* cams2, const float* level_sigma2_1_host, const float* level_sigma2_2_host, int nlevels,
    int only_stereo, int coarse, int check_orientation, int32_t* d_match12, int32_t* d_nmatches, void* stream) {
  if (!d_kps1 || !d_desc1 || !d_n1 || !d_nleft1 || !fv_ok(fv1) || !d_kps2 || !d_desc2 || !d_n2 || !d_nleft2 ||
      !fv_ok(fv2) || kp_stride < 1 || node_stride < 1 || batch < 1 || (!coarse && (!d_rig_pose || !cams1 || !cams2)) ||
      !level_sigma2_1_host || !level_sigma2_2_host || nlevels < 1 || nlevels > ORBX_MAX_LEVELS || !d_match12 ||
      !d_nmatches)
    return ORBX_E_INVALID;
  TriRig G;
  memset(&G, 0, sizeof(G));
  if (!coarse)
    for (int i = 0; i < 2; i++) {
      if (cams1[i].model != 1 || cams2[i].model != 1) return ORBX_E_INVALID;  // rigs are KannalaBrandt8 pairs (Settings.cc)
      G.cam1[i] = to_cam(cams1[i]); G.cam2[i] = to_cam(cams2[i]);
    }
  G.pose = (const OrbxRigPose*)d_rig_pose; G.nleft1 = d_nleft1; G.nleft2 = d_nleft2;
  G.only_stereo = only_stereo; G.coarse = coarse;
  for (int l = 0; l < ORBX_MAX_LEVELS; l++) {
    G.sigma2_1[l] = l < nlevels ? level_sigma2_1_host[l] : 0.f;
    G.sigma2_2[l] = l < nlevels ? level_sigma2_2_host[l] : 0.f;
  }
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(d_match12, 0xff, (size_t)batch * kp_stride * sizeof(int32_t), st);
  const BowSide S1 = make_side(d_kps1, d_desc1, d_n1, d_has_mp1, nullptr, fv1);
  const BowSide S2 = make_side(d_kps2, d_desc2, d_n2, d_has_mp2, nullptr, fv2);
  dim3 grd(node_stride, batch);
  k_tri_nodes_rig<<<grd, BOW_WARPS * 32, 0, st>>>(S1, S2, G, kp_stride, node_stride, d_match12);
  k_bow_rot<<<batch, 256, 0, st>>>(S1.kps, S2.kps, d_n1, kp_stride, check_orientation, d_match12, nullptr, d_nmatches);
  return cudaGetLastError() == cudaSuccess ? ORBX_OK : ORBX_E_CUDA;
}

extern "C" int orbm_bow_transform(const orbm_vocabulary* voc, const uint8_t* d_desc, const int32_t* d_n,
                                  int kp_stride, int batch, int levelsup, int32_t* d_feat_word,
                                  int32_t* d_feat_node, double* d_feat_weight, const orbm_feature_vector* fv_out,
                                  int node_stride, int32_t* d_bow_word, double* d_bow_val, int32_t* d_bow_n,
                                  void* stream) {
  if (!voc || !voc->d_child_off || !voc->d_child_idx || !voc->d_node_desc || !voc->d_node_word ||
      !voc->d_node_weight || voc->L < 1 || !d_desc || !d_n || kp_stride < 1 || kp_stride > 8192 || batch < 1 ||
      !d_feat_word || !d_feat_node || !d_feat_weight || !fv_ok(fv_out) || node_stride < 1 || !d_bow_word ||
      !d_bow_val || !d_bow_n)
    return ORBX_E_INVALID;
  cudaStream_t st = (cudaStream_t)stream;
  BowVoc V;
  V.child_off = voc->d_child_off; V.child_idx = voc->d_child_idx; V.node_desc = voc->d_node_desc;
  V.node_word = voc->d_node_word; V.node_weight = voc->d_node_weight; V.L = voc->L; V.levelsup = levelsup;
  dim3 grd((kp_stride + 127) / 128, batch);
  k_bow_descend<<<grd, 128, 0, st>>>(V, d_desc, d_n, kp_stride, d_feat_word, d_feat_node, d_feat_weight);
  int np2 = BV_THREADS;
  while (np2 < kp_stride) np2 <<= 1;
  const size_t smem = (size_t)np2 * sizeof(unsigned long long);
  if (orbx_ensure_dyn_smem(ORBX_SMEM_BOW_VECTORS, (const void*)k_bow_vectors, smem) != cudaSuccess) return ORBX_E_CUDA;
  k_bow_vectors<<<batch, BV_THREADS, smem, st>>>(d_n, kp_stride, node_stride, np2, d_feat_word, d_feat_node,
                                                 d_feat_weight, fv_out->d_node, fv_out->d_off, fv_out->d_idx,
                                                 fv_out->d_nn, d_bow_word, d_bow_val, d_bow_n);
  return cudaGetLastError() == cudaSuccess ? ORBX_OK : ORBX_E_CUDA;
}

extern "C" int orbm_distinctive_descriptors(const uint8_t* d_desc_pool, const int32_t* d_obs_off,
                                            const int32_t* d_obs_idx, int n_points, int32_t* d_best,
                                            int32_t* d_median, void* stream) {
  if (!d_desc_pool || !d_obs_off || !d_obs_idx || n_points < 1 || !d_best) return ORBX_E_INVALID;
  k_distinctive<<<(n_points + BOW_WARPS - 1) / BOW_WARPS, BOW_WARPS * 32, 0, (cudaStream_t)stream>>>(
      d_desc_pool, d_obs_off, d_obs_idx, n_points, d_best, d_median);
  return cudaGetLastError() == cudaSuccess ? ORBX_OK : ORBX_E_CUDA;
}

// ---------------------------------------------------------------------------------------------------------
// ORBmatcher::SearchForInitialization (ORBmatcher.cc:648-763).  The loop is sequential through
// vMatchedDistance / vnMatches21 (a later query may steal a feature from an earlier one), and it only runs
// while the monocular map is being initialised, so one warp walks the level-0 keypoints of F1 in order and
// spreads each query's window candidates over its lanes.
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(32)
k_search_init(const OrbxKeyPointDev* __restrict__ k1, const uint8_t* __restrict__ d1,
              const int32_t* __restrict__ n1_arr, const OrbxKeyPointDev* __restrict__ k2,
              const uint8_t* __restrict__ d2, const int32_t* __restrict__ n2_arr, int kp_stride,
              const int32_t* __restrict__ cell_off, const int32_t* __restrict__ cell_idx, float min_x, float min_y,
              float winv, float hinv, float* __restrict__ prev_matched, float window, float nnratio,
              int check_orientation, int32_t* __restrict__ match12, int32_t* __restrict__ match21,
              int32_t* __restrict__ matched_dist, int32_t* __restrict__ nmatches_out) {
  __shared__ int hist[30];
  const int lane = threadIdx.x, p = blockIdx.x;
  const size_t fb = (size_t)p * kp_stride;
  const int n1 = n1_arr[p], n2 = n2_arr[p];
  const int32_t* off = cell_off + (size_t)p * (GRID_CELLS + 1);
  const int32_t* idx = cell_idx + fb;
  volatile int32_t* m12 = match12 + fb;
  volatile int32_t* m21 = match21 + fb;
  volatile int32_t* md = matched_dist + fb;
  if (lane < 30) hist[lane] = 0;
  for (int i = lane; i < n1; i += 32) m12[i] = -1;
  for (int i = lane; i < n2; i += 32) { m21[i] = -1; md[i] = 0x7fffffff; }
  __syncwarp();
  const float r = window;
  for (int i1 = 0; i1 < n1; i1++) {
    const OrbxKeyPointDev kp1 = k1[fb + i1];
    if (kp1.octave > 0) continue;
    const float x = prev_matched[(fb + i1) * 2], y = prev_matched[(fb + i1) * 2 + 1];
    const int nMinCellX = max(0, (int)floorf(f_mul(f_sub(f_sub(x, min_x), r), winv)));
    const int nMaxCellX = min(FRAME_GRID_COLS - 1, (int)ceilf(f_mul(f_add(f_sub(x, min_x), r), winv)));
    const int nMinCellY = max(0, (int)floorf(f_mul(f_sub(f_sub(y, min_y), r), hinv)));
    const int nMaxCellY = min(FRAME_GRID_ROWS - 1, (int)ceilf(f_mul(f_add(f_sub(y, min_y), r), hinv)));
    if (!(nMinCellX < FRAME_GRID_COLS && nMaxCellX >= 0 && nMinCellY < FRAME_GRID_ROWS && nMaxCellY >= 0)) continue;
    const Desc256 dq = load_desc(d1 + (fb + i1) * 32);
    Top2 t;
    int seq = 0;
    for (int ix = nMinCellX; ix <= nMaxCellX; ix++) {
      const int e0 = off[ix * FRAME_GRID_ROWS + nMinCellY], e1 = off[ix * FRAME_GRID_ROWS + nMaxCellY + 1];
      for (int e = e0 + lane; e < e1; e += 32) {
        const int j = idx[e];
        const OrbxKeyPointDev kp2 = k2[fb + j];
        if (kp2.octave != 0) continue;  // GetFeaturesInArea(..., level1, level1) with level1 == 0
        if (!(fabsf(f_sub(kp2.x, x)) < r && fabsf(f_sub(kp2.y, y)) < r)) continue;
        const int dist = hamming(dq, load_desc(d2 + (fb + j) * 32));
        if (md[j] <= dist) continue;    // :689
        t.push(make_key(dist, (seq + e - e0) & 0xFFFFF, ORBM_TIE_FIRST_MIN), j);
      }
      seq += e1 - e0;
    }
    warp_top2(t);
    if (t.i1 < 0) continue;
    const int bestDist = (int)(t.k1 >> 20);
    const float second = t.i2 >= 0 ? (float)(int)(t.k2 >> 20) : 2147483648.0f;  // (float)INT_MAX
    if (bestDist <= 50 && (float)bestDist < f_mul(second, nnratio)) {
      const int j = t.i1;
      if (lane == 0) {
        const int old = m21[j];
        if (old >= 0) m12[old] = -1;  // a stolen feature: the earlier match is dropped (:709-713)
        m12[i1] = j; m21[j] = i1; md[j] = bestDist;
        if (check_orientation) hist[bow_rot_bin(kp1.angle, k2[fb + j].angle)]++;
      }
    }
    __syncwarp();
  }
  // nmatches bookkeeping: count what is live now (every steal did nmatches-- and nmatches++)
  __syncwarp();
  int live = 0;
  for (int i = lane; i < n1; i += 32) live += m12[i] >= 0;
  if (check_orientation) {
    int ind1 = -1, ind2 = -1, ind3 = -1;
    {
      int max1 = 0, max2 = 0, max3 = 0;
      for (int i = 0; i < 30; i++) {
        const int s = hist[i];
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
        else if (s > max3) { max3 = s; ind3 = i; }
      }
      if ((float)max2 < f_mul(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
      else if ((float)max3 < f_mul(0.1f, (float)max1)) { ind3 = -1; }
    }
    for (int i = lane; i < n1; i += 32) {
      const int m = m12[i];
      if (m < 0) continue;
      const int bin = bow_rot_bin(k1[fb + i].angle, k2[fb + m].angle);
      if (bin != ind1 && bin != ind2 && bin != ind3) { m12[i] = -1; live--; }
    }
  }
  __syncwarp();
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) live += __shfl_xor_sync(FULLMASK, live, o);
  if (lane == 0) nmatches_out[p] = live;
  // Update prev matched (:757-759)
  for (int i = lane; i < n1; i += 32) {
    const int m = m12[i];
    if (m >= 0) { prev_matched[(fb + i) * 2] = k2[fb + m].x; prev_matched[(fb + i) * 2 + 1] = k2[fb + m].y; }
  }
}

extern "C" int orbm_search_for_initialization(
    const orbx_keypoint* d_kps1, const uint8_t* d_desc1, const int32_t* d_n1, const orbx_keypoint* d_kps2,
    const uint8_t* d_desc2, const int32_t* d_n2, int kp_stride, int batch, const int32_t* d_cell_off2,
    const int32_t* d_cell_idx2, float min_x, float max_x, float min_y, float max_y, float* d_prev_matched,
    int window_size, float nn_ratio, int check_orientation, int32_t* d_match12, int32_t* d_scratch,
    int32_t* d_nmatches, void* stream) {
  if (!d_kps1 || !d_desc1 || !d_n1 || !d_kps2 || !d_desc2 || !d_n2 || kp_stride < 1 || batch < 1 || !d_cell_off2 ||
      !d_cell_idx2 || !(max_x > min_x) || !(max_y > min_y) || !d_prev_matched || !d_match12 || !d_scratch ||
      !d_nmatches)
    return ORBX_E_INVALID;
  const float winv = (float)FRAME_GRID_COLS / (max_x - min_x), hinv = (float)FRAME_GRID_ROWS / (max_y - min_y);
  k_search_init<<<batch, 32, 0, (cudaStream_t)stream>>>(
      (const OrbxKeyPointDev*)d_kps1, d_desc1, d_n1, (const OrbxKeyPointDev*)d_kps2, d_desc2, d_n2, kp_stride,
      d_cell_off2, d_cell_idx2, min_x, min_y, winv, hinv, d_prev_matched, (float)window_size, nn_ratio,
      check_orientation, d_match12, d_scratch, d_scratch + (size_t)batch * kp_stride, d_nmatches);
  return cudaGetLastError() == cudaSuccess ? ORBX_OK : ORBX_E_CUDA;
}
