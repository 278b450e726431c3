// CPU unit harness: compiles the product's host/device-shared headers for the HOST so the
// arithmetic and the octree round logic can be checked against the oracle without a GPU.
// Test infrastructure (built by tests/conftest.py with g++ -ffp-contract=off).
#include <math.h>
#include <stdint.h>
#include <string.h>

#include <vector>

#include "octree_core.h"
#include "orbx_core.h"
#include "orbx_plan.h"
#include "frame_core.h"
#include "tri_core.h"

extern "C" {

// DistributeOctTree through oct_init/oct_round with the parallel steps emulated serially.
// keys: (x, y, score) int triples relative to (16,16).  Returns kept count, out = triples.
int h_distribute(const int* xys, int n, int width, int height, int N, int* out) {
  int nIni = (int)roundf((float)width / (float)height);
  float hX = (float)width / (float)nIni;
  int cap = 2 * ((N + 3 > 4 * nIni ? N + 3 : 4 * nIni)) + 8;
  std::vector<OctNode> nodes(cap);
  std::vector<uint64_t> v1(cap), v2(cap);
  std::vector<int> divl(cap), label(n), rootc(nIni, 0);
  OctState S;
  S.nodes = nodes.data(); S.node_cap = cap; S.vec = v1.data(); S.vec2 = v2.data(); S.divlist = divl.data(); S.N = N;
  for (int i = 0; i < n; i++) { label[i] = (int)f_div((float)xys[3 * i], hX); rootc[label[i]]++; }
  oct_init(S, nIni, hX, height, rootc.data());
  for (int i = 0; i < nIni; i++) if (S.nodes[i].cand) S.vec[S.nvec++] = 0;  // only the count matters
  int rounds = 0;
  while (!S.finished) {
    for (int i = 0; i < n; i++) {
      OctNode& nd = S.nodes[label[i]];
      if (nd.cand) nd.c[oct_quadrant(nd, xys[3 * i], xys[3 * i + 1])]++;
    }
    oct_round(S);
    for (int i = 0; i < n; i++) {
      OctNode& nd = S.nodes[label[i]];
      if (nd.divided) label[i] = nd.c[oct_quadrant(nd, xys[3 * i], xys[3 * i + 1])];
    }
    if (++rounds > 100000) return -1;
  }
  if (S.overflow) return -2;
  std::vector<uint32_t> best(cap, 0);
  for (int i = 0; i < n; i++) {
    uint32_t key = ((uint32_t)xys[3 * i + 2] << 24) | (0xFFFFFFu - (uint32_t)i);
    if (key > best[label[i]]) best[label[i]] = key;
  }
  int k = 0;
  for (int it = S.head; it >= 0; it = S.nodes[it].next) {
    int i = (int)(0xFFFFFFu - (best[it] & 0xFFFFFFu));
    out[3 * k] = xys[3 * i]; out[3 * k + 1] = xys[3 * i + 1]; out[3 * k + 2] = xys[3 * i + 2];
    k++;
  }
  return k;
}

// The array ("scan") form of DistributeOctTree that the CUDA kernel runs (octree_core.h, second
// half), with the block scans written as plain loops.
int h_distribute_scan(const int* xys, int n, int width, int height, int N, int* out) {
  const int nIni = (int)roundf((float)width / (float)height);
  const float hX = (float)width / (float)nIni;
  const int cap = (N + 3 > 4 * nIni ? N + 3 : 4 * nIni) + 4;
  std::vector<OctPNode> bufA(cap), bufB(cap);
  OctPNode *cur = bufA.data(), *next = bufB.data();
  std::vector<uint64_t> v1(cap), v2(cap);
  uint64_t *vec = v1.data(), *vec2 = v2.data();
  std::vector<int> qc(4 * cap), newpos(cap), label(n), rootc(nIni, 0), rootpos(nIni, -1);
  std::vector<uint8_t> proc(cap);
  for (int i = 0; i < n; i++) { label[i] = (int)f_div((float)xys[3 * i], hX); rootc[label[i]]++; }
  int S = 0;
  for (int i = 0; i < nIni; i++)
    if (rootc[i] > 0) { rootpos[i] = S; cur[S++] = octp_root(i, hX, height, rootc[i]); }
  for (int i = 0; i < n; i++) label[i] = rootpos[label[i]];
  int nvec = 0, careful = 0, rounds = 0;
  bool finished = n == 0;
  while (!finished) {
    if (++rounds > 100000) return -1;
    const int prevS = S;
    for (int p = 0; p < S; p++) { proc[p] = 0; qc[4 * p] = qc[4 * p + 1] = qc[4 * p + 2] = qc[4 * p + 3] = 0; }
    for (int i = 0; i < n; i++) {
      const OctPNode& nd = cur[label[i]];
      if (nd.cnt > 1) qc[4 * label[i] + octp_quadrant(nd, xys[3 * i], xys[3 * i + 1])]++;
    }
    auto K = [&](int p) { return (qc[4 * p] > 0) + (qc[4 * p + 1] > 0) + (qc[4 * p + 2] > 0) + (qc[4 * p + 3] > 0); };
    auto E = [&](int p) { return (qc[4 * p] > 1) + (qc[4 * p + 1] > 1) + (qc[4 * p + 2] > 1) + (qc[4 * p + 3] > 1); };
    int totK = 0, totE = 0, m = 0;
    if (!careful) {
      for (int p = 0; p < S; p++) if (cur[p].cnt > 1) totK += K(p);
      int preK = 0;
      for (int p = 0; p < S; p++) {
        if (cur[p].cnt <= 1) continue;
        const int k = K(p), e = E(p);
        proc[p] = 1; m++;
        octp_emit(cur[p], &qc[4 * p], totK - preK - k, totE, next, vec2);
        preK += k; totE += e;
      }
    } else {
      oct_std_sort(vec, nvec);
      int size = S;
      for (int s = 0; s < nvec; s++) {  // which candidates are processed
        const int p = (int)(vec[nvec - 1 - s] & 0xffff);
        totK += K(p); size += K(p) - 1; m = s + 1;
        if (size >= N) break;
      }
      int preK = 0;
      for (int s = 0; s < m; s++) {
        const int p = (int)(vec[nvec - 1 - s] & 0xffff);
        const int k = K(p), e = E(p);
        proc[p] = 1;
        octp_emit(cur[p], &qc[4 * p], totK - preK - k, totE, next, vec2);
        preK += k; totE += e;
      }
    }
    int stay = 0;
    for (int p = 0; p < S; p++)
      if (!proc[p]) { newpos[p] = totK + stay; next[totK + stay] = cur[p]; stay++; }
    for (int i = 0; i < n; i++) {
      const int p = label[i];
      label[i] = proc[p] ? qc[4 * p + octp_quadrant(cur[p], xys[3 * i], xys[3 * i + 1])] : newpos[p];
    }
    S = totK + stay;
    if (S > cap) return -2;
    std::swap(cur, next); std::swap(vec, vec2); nvec = totE;
    if (S >= N || S == prevS) finished = true;
    else if (!careful && S + 3 * totE > N) careful = 1;
  }
  std::vector<uint32_t> best(cap, 0);
  for (int i = 0; i < n; i++) {
    uint32_t key = ((uint32_t)xys[3 * i + 2] << 24) | (0xFFFFFFu - (uint32_t)i);
    if (key > best[label[i]]) best[label[i]] = key;
  }
  for (int r = 0; r < S; r++) {
    int i = (int)(0xFFFFFFu - (best[r] & 0xFFFFFFu));
    out[3 * r] = xys[3 * i]; out[3 * r + 1] = xys[3 * i + 1]; out[3 * r + 2] = xys[3 * i + 2];
  }
  return S;
}

// counts float bit patterns in [lo,hi] where orbx_sincosf differs from libm sincosf
long h_sincos_mismatches(uint32_t lo, uint32_t hi, uint32_t stride) {
  long bad = 0;
  for (uint64_t u = lo; u <= hi; u += stride) {
    float y, s, c, s2, c2; uint32_t uu = (uint32_t)u; memcpy(&y, &uu, 4);
    sincosf(y, &s, &c);
    orbx_sincosf(y, &s2, &c2);
    if (memcmp(&s, &s2, 4) || memcmp(&c, &c2, 4)) bad++;
  }
  return bad;
}

// counts positive normal float bit patterns in [lo,hi] where orbx_logf differs from libm logf
long h_logf_mismatches(uint32_t lo, uint32_t hi, uint32_t stride) {
  long bad = 0;
  for (uint64_t u = lo; u <= hi; u += stride) {
    float x, a, b; uint32_t uu = (uint32_t)u; memcpy(&x, &uu, 4);
    a = logf(x); b = orbx_logf(x);
    if (memcmp(&a, &b, 4)) bad++;
  }
  return bad;
}

// frame_core.h compiled for the host: cv::undistortPoints on (x,y) pairs, isInFrustum on n MapPoints
void h_undistort(const float* xy, int n, float fx, float fy, float cx, float cy, const float* dist, int n_coef,
                 float* out_xy) {
  double k[12];
  for (int i = 0; i < 12; i++) k[i] = i < n_coef ? (double)dist[i] : 0.0;
  for (int i = 0; i < n; i++) orbx_undistort_point(xy[2 * i], xy[2 * i + 1], fx, fy, cx, cy, k, out_xy + 2 * i, out_xy + 2 * i + 1);
}

void h_in_frustum(const float* pos, const float* normal, const float* min_dist, const float* max_dist, int n,
                  const float* pose, int model, const float* cam, float mbf, float min_x, float max_x, float min_y,
                  float max_y, float cos_limit, float log_sf, int n_levels, int fisheye_rig, uint8_t* in_view,
                  float* proj_x, float* proj_y, float* proj_xr, float* depth, float* view_cos, int8_t* level) {
  OrbxFrustumPose T; memcpy(&T, pose, sizeof(T));
  OrbxFrustumCam C; C.model = model; C.fx = cam[0]; C.fy = cam[1]; C.cx = cam[2]; C.cy = cam[3];
  for (int i = 0; i < 4; i++) C.k[i] = cam[4 + i];
  for (int i = 0; i < n; i++) {
    const OrbxFrustumOut o = orbx_in_frustum(pos + 3 * i, normal + 3 * i, min_dist[i], max_dist[i], T, C, mbf, min_x,
                                             max_x, min_y, max_y, cos_limit, log_sf, n_levels, fisheye_rig);
    in_view[i] = (uint8_t)o.in_view; proj_x[i] = o.proj_x; proj_y[i] = o.proj_y; proj_xr[i] = o.proj_xr;
    depth[i] = o.depth; view_cos[i] = o.view_cos; level[i] = (int8_t)o.level;
  }
}

// tri_core.h for the host: KannalaBrandt8::epipolarConstrain on n keypoint pairs (cam = fx,fy,cx,cy,k0..k3)
void h_kb8_epipolar(const float* cam1, const float* cam2, const float* uv1, const float* uv2, int n, const float* pose12,
                    const float* sigma1, const float* sigma2, uint8_t* out) {
  OrbxFrustumCam C1, C2;
  C1.model = C2.model = 1;
  C1.fx = cam1[0]; C1.fy = cam1[1]; C1.cx = cam1[2]; C1.cy = cam1[3];
  C2.fx = cam2[0]; C2.fy = cam2[1]; C2.cx = cam2[2]; C2.cy = cam2[3];
  for (int i = 0; i < 4; i++) { C1.k[i] = cam1[4 + i]; C2.k[i] = cam2[4 + i]; }
  OrbxRigPose T; memcpy(&T, pose12, sizeof(T));
  for (int i = 0; i < n; i++)
    out[i] = orbx_kb8_epipolar_constrain(C1, C2, uv1[2 * i], uv1[2 * i + 1], uv2[2 * i], uv2[2 * i + 1], T, sigma1[i],
                                         sigma2[i]) ? 1 : 0;
}

// tri_core.h for the host: KannalaBrandt8::unprojectEig (cam = fx,fy,cx,cy,k0..k3), uv [n][2] -> ray [n][3]
void h_kb8_unproject(const float* cam, const float* uv, int n, float* ray) {
  OrbxFrustumCam C;
  C.model = 1;
  C.fx = cam[0]; C.fy = cam[1]; C.cx = cam[2]; C.cy = cam[3];
  for (int i = 0; i < 4; i++) C.k[i] = cam[4 + i];
  for (int i = 0; i < n; i++) orbx_kb8_unproject(C, uv[2 * i], uv[2 * i + 1], ray + 3 * i);
}

void h_fast_atan2(const float* y, const float* x, float* out, int n) {
  for (int i = 0; i < n; i++) out[i] = orbx_fast_atan2(y[i], x[i]);
}

int h_arc_strength(int v, const int* ring) { return orbx_arc_strength(v, ring); }

// plan summary: per level {w,h,nCols,nRows,wCell,hCell,nfeat,nIni,kp_max}, returns 0/-1
int h_plan(int nfeatures, float sf, int nlevels, int ini, int mn, int w, int h, int* out9, float* scales,
           int* umax, int* totals) {
  OrbxPlan P; std::vector<OrbxTap> xt, yt;
  int rc = orbx_build_plan(nfeatures, sf, nlevels, ini, mn, w, h, &P, &xt, &yt);
  if (rc) return rc;
  for (int l = 0; l < nlevels; l++) {
    const OrbxLevel& L = P.lv[l];
    int* o = out9 + 9 * l;
    o[0] = L.w; o[1] = L.h; o[2] = L.nCols; o[3] = L.nRows; o[4] = L.wCell; o[5] = L.hCell; o[6] = L.nfeat;
    o[7] = L.nIni; o[8] = L.kp_max;
    scales[4 * l] = P.scale[l]; scales[4 * l + 1] = P.inv_scale[l]; scales[4 * l + 2] = P.sigma2[l];
    scales[4 * l + 3] = P.inv_sigma2[l];
  }
  for (int i = 0; i < 16; i++) umax[i] = P.umax[i];
  totals[0] = P.total_cells; totals[1] = (int)P.cand_total; totals[2] = (int)P.arena_bytes; totals[3] = P.max_keypoints;
  return 0;
}
}
