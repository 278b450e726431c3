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
