// orbx_plan.h — host-side plan for one (frame size, ORB parameters) pair: everything the
// reference derives in ORBextractor::ORBextractor (src/ORBextractor.cc:409-469),
// ComputePyramid (:1170-1195) and the cell-grid prologue of ComputeKeyPointsOctTree (:781-803),
// plus the device memory layout.  Plain C++ (no CUDA) so the CPU harness can check it.
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#include <vector>

#include "orbx_core.h"

#define ORBX_PLAN_MAX_LEVELS 12
#ifndef ORBX_BLUR_ROWS
#define ORBX_BLUR_ROWS 35
#endif  // rows per warp tile of k_blur_levels (blur.cu BL_RB); 56 / 70 / 105 measured: describe stage 0.914 / 0.911 / 0.958 ms vs 0.904
#define ORBX_FAST_STRIP_PX 120  // columns a FAST strip can test per row (lanes 1..30 x 4 px)
#define ORBX_BLUR_TILE_WORDS 30  // output words (4 px) per warp tile row: lanes 1..30, lanes 0/31 are halo

struct OrbxLevel {
  int w, h;          // level size (cvRound(cols*inv), cvRound(rows*inv))
  int pitch;         // row pitch in bytes in the pyramid arena (multiple of 128)
  long long offset;  // byte offset inside a frame's arena (level 0 included)
  // FAST cell grid: the tested zone [19,w-19)x[19,h-19) split on a (wCell,hCell) lattice
  int nCols, nRows, wCell, hCell;
  int cell_base;        // first cell of this level in the per-frame flattened cell array
  int cell_cap;         // candidate slots per cell (worst case after NMS)
  long long cand_base;  // offset (uint32 units) of this level's cell slots in the frame's arena
  long long cand_cap;   // nCols*nRows*cell_cap
  int nfeat;            // mnFeaturesPerLevel[level]
  int kp_max;           // most keypoints the level can return: max(nfeat+3, 4*nIni)
  int kp_base, kp_cap;  // selected-keypoint slots (kp_max rounded up)
  float scale;          // mvScaleFactor[level]
  float patch_size;     // (float)(int)(31*scale)
  int nIni;             // DistributeOctTree root count
  float hX;
  int xtab, ytab;  // offsets into the resize tables (entries), level>=1
  int fast_resize;  // the 4-column tap windows of k_resize_walk hold for every column group
  int tile_resize;  // every 128x64 output tile fits k_resize_tile's shared-memory source window
  int tile_rows;    // most source rows any of those tiles needs (rows of the TMA box)
  int btile_base, btiles_x;  // warp tiles (ORBX_BLUR_TILE_WORDS words x ORBX_BLUR_ROWS rows) of k_blur_levels: first tile of this level, tiles per row
  // FAST strips (k_fast_strip): a strip = strip_k consecutive cells of one cell row, one warp each
  int strip_k, strips_per_row, strip_base;
};

struct OrbxPlan {
  int nlevels, width, height, nfeatures, ini_th, min_th;
  OrbxLevel lv[ORBX_PLAN_MAX_LEVELS];
  int total_cells;           // per frame
  long long cand_total;      // uint32 per frame
  long long arena_bytes;     // pyramid bytes per frame (all levels)
  int kp_total_cap;          // Σ kp_cap
  int total_btiles;          // Σ blur tiles per frame
  int total_strips;          // Σ FAST strips per frame
  int strips_a, hmax_a, hmax_b;  // strips [0, strips_a): levels with ordinary cell heights (<= hmax_a), the rest: tall cells (<= hmax_b)
  int max_keypoints;         // Σ kp_max (what one frame can return; nfeatures+3*nlevels normally)
  float scale[ORBX_PLAN_MAX_LEVELS], inv_scale[ORBX_PLAN_MAX_LEVELS];
  float sigma2[ORBX_PLAN_MAX_LEVELS], inv_sigma2[ORBX_PLAN_MAX_LEVELS];
  int umax[16];
};

// Resize tap tables: per destination index {source offset, c0 | c1<<16}.
struct OrbxTap {
  int32_t ofs;
  int16_t c0, c1;
};

inline int orbx_cv_round(float v) { return (int)lrintf(v); }

// Returns 0 on success, -1 when the geometry is unsupported (a level too small for one
// 35-px FAST cell, portrait aspect ratios that make nIni 0, scale chains that trigger
// OpenCV's INTER_AREA fast path).
inline int orbx_build_plan(int nfeatures, float scaleFactorF, int nlevels, int iniTh, int minTh, int width,
                           int height, OrbxPlan* P, std::vector<OrbxTap>* xtaps, std::vector<OrbxTap>* ytaps) {
  if (nlevels < 1 || nlevels > ORBX_PLAN_MAX_LEVELS || nfeatures < 1 || width < 1 || height < 1) return -1;
  if (!(scaleFactorF > 1.0f) && nlevels > 1) return -1;
  if (iniTh < 0 || iniTh > 255 || minTh < 0 || minTh > 255) return -1;
  memset(P, 0, sizeof(*P));
  P->nlevels = nlevels; P->width = width; P->height = height; P->nfeatures = nfeatures;
  P->ini_th = iniTh; P->min_th = minTh;
  const double scaleFactor = (double)scaleFactorF;  // member is a double (ORBextractor.h:97)
  P->scale[0] = 1.0f; P->sigma2[0] = 1.0f;
  for (int i = 1; i < nlevels; i++) {
    P->scale[i] = (float)(P->scale[i - 1] * scaleFactor);
    P->sigma2[i] = P->scale[i] * P->scale[i];
  }
  for (int i = 0; i < nlevels; i++) {
    P->inv_scale[i] = 1.0f / P->scale[i];
    P->inv_sigma2[i] = 1.0f / P->sigma2[i];
  }
  // features per level (:434-445)
  int nfeat[ORBX_PLAN_MAX_LEVELS];
  {
    float factor = (float)(1.0f / scaleFactor);
    float nDesired = nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nlevels));
    int sum = 0;
    for (int l = 0; l < nlevels - 1; l++) {
      nfeat[l] = orbx_cv_round(nDesired);
      sum += nfeat[l];
      nDesired *= factor;
    }
    nfeat[nlevels - 1] = nfeatures - sum > 0 ? nfeatures - sum : 0;
  }
  // umax (:453-468)
  {
    int v, v0, vmax = (int)floorf(ORBX_HALF_PATCH * sqrtf(2.f) / 2 + 1);
    int vmin = (int)ceilf(ORBX_HALF_PATCH * sqrtf(2.f) / 2);
    const double hp2 = ORBX_HALF_PATCH * ORBX_HALF_PATCH;
    for (v = 0; v <= vmax; ++v) P->umax[v] = (int)lrint(sqrt(hp2 - v * v));
    for (v = ORBX_HALF_PATCH, v0 = 0; v >= vmin; --v) {
      while (P->umax[v0] == P->umax[v0 + 1]) ++v0;
      P->umax[v] = v0;
      ++v0;
    }
  }
  long long offset = 0, cand = 0;
  int cells = 0, kpb = 0, maxk = 0, btiles = 0, strips = 0;
  xtaps->clear(); ytaps->clear();
  for (int l = 0; l < nlevels; l++) {
    OrbxLevel& L = P->lv[l];
    L.w = orbx_cv_round((float)width * P->inv_scale[l]);
    L.h = orbx_cv_round((float)height * P->inv_scale[l]);
    L.pitch = (L.w + 127) & ~127;
    L.offset = offset;
    offset += (long long)L.pitch * L.h;
    L.btile_base = btiles;
    L.btiles_x = ((L.w + 3) / 4 + ORBX_BLUR_TILE_WORDS - 1) / ORBX_BLUR_TILE_WORDS;
    btiles += L.btiles_x * ((L.h + ORBX_BLUR_ROWS - 1) / ORBX_BLUR_ROWS);
    L.scale = P->scale[l];
    L.patch_size = (float)(int)(31 * P->scale[l]);
    L.nfeat = nfeat[l];
    // cell grid (:789-803)
    const int minB = ORBX_MIN_BORDER, maxBX = L.w - ORBX_EDGE + 3, maxBY = L.h - ORBX_EDGE + 3;
    const float fw = (float)(maxBX - minB), fh = (float)(maxBY - minB);
    L.nCols = (int)(fw / ORBX_CELL_W);
    L.nRows = (int)(fh / ORBX_CELL_W);
    if (L.nCols < 1 || L.nRows < 1) return -1;
    L.wCell = (int)ceilf(fw / L.nCols);
    L.hCell = (int)ceilf(fh / L.nRows);
    L.cell_base = cells;
    cells += L.nCols * L.nRows;
    // strips: as many whole cells as fit the 120 tested columns after the <= 3 px of word alignment slack
    L.strip_k = (ORBX_FAST_STRIP_PX - 3) / L.wCell;
    if (L.strip_k < 1) return -1;
    if (L.strip_k > L.nCols) L.strip_k = L.nCols;
    L.strips_per_row = (L.nCols + L.strip_k - 1) / L.strip_k;
    L.cell_cap = ((L.wCell + 1) / 2) * ((L.hCell + 1) / 2);
    L.cand_base = cand;
    L.cand_cap = (long long)L.nCols * L.nRows * L.cell_cap;
    cand += L.cand_cap;
    // octree roots (:559-561)
    L.nIni = (int)roundf((float)(maxBX - minB) / (float)(maxBY - minB));
    if (L.nIni < 1) return -1;
    L.hX = (float)(maxBX - minB) / (float)L.nIni;
    // the list never exceeds max(N+3, 4*nIni) nodes (careful phase stops at the first size>=N,
    // a split adds at most 3; only the very first pass can overshoot, by 4 children per root)
    L.kp_max = L.nfeat + 3 > 4 * L.nIni ? L.nfeat + 3 : 4 * L.nIni;
    L.kp_base = kpb;
    L.kp_cap = (L.kp_max + 3) & ~3;
    kpb += L.kp_cap;
    maxk += L.kp_max;
    // resize taps from level l-1 (cv::resize INTER_LINEAR 8U)
    if (l > 0) {
      const OrbxLevel& S = P->lv[l - 1];
      if (S.w == 2 * L.w && S.h == 2 * L.h) return -1;  // OpenCV switches to INTER_AREA here
      L.xtab = (int)xtaps->size();
      L.ytab = (int)ytaps->size();
      double sx_scale = 1. / ((double)L.w / S.w), sy_scale = 1. / ((double)L.h / S.h);
      for (int dx = 0; dx < L.w; dx++) {
        float fx = (float)((dx + 0.5) * sx_scale - 0.5);
        int sx = (int)floorf(fx);
        fx -= sx;
        if (sx < 0) { fx = 0; sx = 0; }
        if (sx >= S.w - 1) { fx = 0; sx = S.w - 1; }
        OrbxTap t;
        t.ofs = sx;
        t.c0 = (int16_t)orbx_cv_round((1.f - fx) * 2048.f);
        t.c1 = (int16_t)orbx_cv_round(fx * 2048.f);
        xtaps->push_back(t);
      }
      // k_resize_walk: taps 0,1 of a 4-column group read inside [ofs0, ofs0+4), taps 2,3 inside
      // [ofs0+2, ofs0+6); true for every down-scale below 4/3
      L.fast_resize = 1;
      for (int dx = 0; dx < L.w; dx++) {
        const int o0 = (*xtaps)[L.xtab + (dx & ~3)].ofs;
        const int d = (*xtaps)[L.xtab + dx].ofs - o0 - ((dx & 3) < 2 ? 0 : 2);
        if (d < 0 || d > 2) L.fast_resize = 0;
      }
      for (int dy = 0; dy < L.h; dy++) {
        float fy = (float)((dy + 0.5) * sy_scale - 0.5);
        int sy = (int)floorf(fy);
        fy -= sy;
        OrbxTap t;
        t.ofs = sy;
        t.c0 = (int16_t)orbx_cv_round((1.f - fy) * 2048.f);
        t.c1 = (int16_t)orbx_cv_round(fy * 2048.f);
        ytaps->push_back(t);
      }
      // k_resize_tile: source window of a 128 x 64 output tile (88 rows x 208 bytes of shared memory)
      L.tile_resize = L.fast_resize;
      L.tile_rows = 0;
      for (int x0 = 0; x0 < L.w; x0 += 128) {
        const int xs0 = (*xtaps)[L.xtab + x0].ofs & ~15;
        const int xl = (*xtaps)[L.xtab + (x0 + 127 < L.w ? x0 + 127 : L.w - 1)].ofs;
        if (((xl + 12 - xs0 + 15) & ~15) > 208) L.tile_resize = 0;
      }
      for (int y0 = 0; y0 < L.h; y0 += 64) {
        int a = (*ytaps)[L.ytab + y0].ofs, b = (*ytaps)[L.ytab + (y0 + 63 < L.h ? y0 + 63 : L.h - 1)].ofs + 1;
        a = a < 0 ? 0 : (a > S.h - 1 ? S.h - 1 : a);
        b = b < 0 ? 0 : (b > S.h - 1 ? S.h - 1 : b);
        if (b - a + 1 > 88) L.tile_resize = 0;
        if (b - a + 1 > L.tile_rows) L.tile_rows = b - a + 1;
      }
    }
  }
  P->total_cells = cells;
  P->cand_total = cand;
  P->arena_bytes = (offset + 255) & ~255LL;
  P->kp_total_cap = kpb;
  P->total_btiles = btiles;
  // strip order: the levels whose cells are about as tall as level 0's first, then the levels with tall cells
  // (the small levels with few cell rows) — k_fast_strip sizes a warp's shared memory by the tallest zone of a launch
  P->hmax_a = P->hmax_b = 0;
  for (int g = 0; g < 2; g++) {
    for (int l = 0; l < nlevels; l++) {
      OrbxLevel& L = P->lv[l];
      if ((L.hCell <= P->lv[0].hCell + 2) != (g == 0)) continue;
      L.strip_base = strips;
      strips += L.strips_per_row * L.nRows;
      if (g == 0) { if (L.hCell > P->hmax_a) P->hmax_a = L.hCell; }
      else if (L.hCell > P->hmax_b) P->hmax_b = L.hCell;
    }
    if (g == 0) P->strips_a = strips;
  }
  P->total_strips = strips;
  P->max_keypoints = maxk;
  return 0;
}
