// api.cu — the C ABI of include/orbx.h: handle management and the per-frame launch sequence
// (ORBextractor::operator(), src/ORBextractor.cc:1086-1168).
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <new>
#include <vector>

#include "../../include/orbx.h"
#include "orbx_kernels.cuh"
#include "orbx_core.h"

static thread_local char g_err[256] = "";
static void set_err(const char* what, cudaError_t e) {
  snprintf(g_err, sizeof(g_err), "%s: %s", what, e == cudaSuccess ? "" : cudaGetErrorString(e));
}
void orbx_set_error_text(const char* msg) { snprintf(g_err, sizeof(g_err), "%s", msg ? msg : ""); }
#define CU(call)                                  \
  do {                                            \
    cudaError_t e_ = (call);                      \
    if (e_ != cudaSuccess) {                      \
      set_err(#call, e_);                         \
      return e_ == cudaErrorMemoryAllocation ? ORBX_E_NOMEM : ORBX_E_CUDA; \
    }                                             \
  } while (0)

struct orbx_extractor_s {
  orbx_config cfg;
  OrbxPlan plan;
  DevPlan dplan;
  int device;
  cudaStream_t stream;
  // device buffers
  uint8_t* arena = nullptr;
  OrbxTap *xtaps = nullptr, *ytaps = nullptr;
  int32_t* cell_cnt = nullptr;
  uint32_t* cell_tab = nullptr;  // per cell: level << 24 | row << 12 | column
  OrbxResizeMaps resize_maps;     // tensor maps of the source levels for k_resize_tile (level 0 is encoded per call)
  OrbxFastMaps fast_maps;         // tensor maps of arena levels 1.. for k_fast_strip (level 0 is encoded per call)
  uint32_t* strip_tab = nullptr;  // per FAST strip: level << 24 | cell row << 12 | first cell column
  uint32_t* cand = nullptr;
  uint32_t* gkeys = nullptr;
  int16_t* glabels = nullptr;
  uint32_t* sel = nullptr;
  int32_t* sel_cnt = nullptr;
  int32_t* dst = nullptr;
  int32_t* status = nullptr;
  // outputs of the host-pointer entry points
  OrbxKeyPointDev* o_kps = nullptr;
  uint8_t* o_desc = nullptr;
  int32_t *o_n = nullptr, *o_mono = nullptr;
  uint8_t* border_tmp = nullptr;
  // orbx_get_pyramid_all: every level of one frame, packed, on the device and in pinned host memory
  uint8_t* pyr_stage_d = nullptr; uint8_t* pyr_stage_h = nullptr; size_t pyr_stage_bytes = 0;
  uint8_t* stage = nullptr;  // tightly packed level-0 frames of the host batch API (lazy)
  uint8_t* blur = nullptr;   // blurred levels (arena layout) in level-blur mode
  bool level_blur = false;
  // octree launch geometry
  size_t oct_smem;
  int oct_smem_keys, oct_node_cap, oct_max_cells;
  // state of the last call
  Level0 last_l0;
  int last_batch = 0;
  cudaEvent_t ev[ORBX_NUM_STAGES + 1] = {};
  bool timed = false;
  // orbx_extract_batch_host pipeline: copy-in / two compute / copy-out streams, one event pair per chunk
  cudaStream_t s_in = nullptr, s_out = nullptr, s_cmp[4] = {nullptr, nullptr, nullptr, nullptr};
  int host_streams = 2;  // compute streams the host batch pipeline rotates its chunks over (ORBX_HOST_STREAMS, 1..4)
  std::vector<cudaEvent_t> ev_in, ev_done;
  cudaEvent_t ev_fork = nullptr, ev_join[4] = {nullptr, nullptr, nullptr, nullptr};
  // level-blur mode: k_blur_levels only needs the pyramid, so it can run beside FAST / octree of the same
  // sub-batch on a side stream (ORBX_BLUR_SIDE=1): fork after the pyramid, join before the sampler
  cudaStream_t s_side[4] = {nullptr, nullptr, nullptr, nullptr};
  cudaStream_t s_hi[4] = {nullptr, nullptr, nullptr, nullptr};  // high-priority twins of s_cmp for the octree
  cudaEvent_t ev_fast[4] = {nullptr, nullptr, nullptr, nullptr}, ev_oct[4] = {nullptr, nullptr, nullptr, nullptr};
  int oct_prio = 0;  // ORBX_OCT_PRIO
  cudaEvent_t ev_pyr[4] = {nullptr, nullptr, nullptr, nullptr}, ev_blur[4] = {nullptr, nullptr, nullptr, nullptr};
  int blur_side = 0;
  bool streams_ready = false;
  int host_chunk = 64;
  int host_edge = 0;   // length of the first and last chunk of a long host batch (ORBX_HOST_EDGE, 0 = off)
  int dev_split = 2;  // sub-batches of orbx_extract_batch_device (ORBX_DEVICE_SPLIT)
  int dev_streams = 2;  // internal streams the sub-batches rotate over (ORBX_DEVICE_STREAMS, 1..4)
};

static void fill_devplan(const OrbxPlan& P, DevPlan* D) {
  memset(D, 0, sizeof(*D));
  D->nlevels = P.nlevels; D->total_cells = P.total_cells; D->kp_total_cap = P.kp_total_cap;
  D->ini_th = P.ini_th; D->min_th = P.min_th; D->total_btiles = P.total_btiles; D->total_strips = P.total_strips; D->strips_a = P.strips_a; D->hmax_a = P.hmax_a; D->hmax_b = P.hmax_b; D->cand_total = P.cand_total; D->arena_bytes = P.arena_bytes;
  for (int i = 0; i < 16; i++) D->umax[i] = P.umax[i];
  for (int l = 0; l < P.nlevels; l++) {
    const OrbxLevel& s = P.lv[l];
    DevLevel& d = D->lv[l];
    d.w = s.w; d.h = s.h; d.pitch = s.pitch; d.nCols = s.nCols; d.nRows = s.nRows; d.wCell = s.wCell;
    d.hCell = s.hCell; d.cell_base = s.cell_base; d.cell_cap = s.cell_cap; d.nfeat = s.nfeat;
    d.kp_base = s.kp_base; d.kp_cap = s.kp_cap; d.kp_max = s.kp_max; d.nIni = s.nIni; d.hX = s.hX;
    d.scale = s.scale; d.patch_size = s.patch_size; d.xtab = s.xtab; d.ytab = s.ytab; d.fast_resize = s.fast_resize; d.tile_resize = s.tile_resize; d.tile_rows = s.tile_rows; d.btile_base = s.btile_base; d.btiles_x = s.btiles_x; d.offset = s.offset;
    d.cand_base = s.cand_base; d.strip_k = s.strip_k;
  }
}

// Test hook kernel: order-independent 64-bit checksum per chunk of the device build of orbx_sincosf over the
// float bit patterns lo, lo+1, ... (the nvcc compile of the glibc-exact sincosf the descriptor stage uses).
__device__ __forceinline__ unsigned long long sincos_mix(uint32_t bits, float s, float c) {
  unsigned long long h = ((unsigned long long)__float_as_uint(c) << 32) | __float_as_uint(s);
  h = h * 0x9E3779B97F4A7C15ull + bits;
  return h ^ (h >> 29);
}
__global__ void k_sincos_checksum(uint32_t lo, unsigned long long count, uint32_t chunk, unsigned long long* out) {
  const unsigned long long c0 = (unsigned long long)blockIdx.x * chunk;
  unsigned long long acc = 0;
  for (unsigned long long i = c0 + threadIdx.x; i < c0 + chunk && i < count; i += blockDim.x) {
    const uint32_t bits = lo + (uint32_t)i;
    float s, c;
    orbx_sincosf(__uint_as_float(bits), &s, &c);
    acc += sincos_mix(bits, s, c);
  }
  atomicAdd(out + blockIdx.x, acc);
}

// Waits for everything the handle's internal streams still have in flight (error paths: no D2H copy
// into the caller's buffers may outlive the call, and forked streams must not be left unjoined).
static void drain_streams(orbx_extractor_s* h) {
  if (h->s_in) cudaStreamSynchronize(h->s_in);
  for (int i = 0; i < 4; i++)
    if (h->s_cmp[i]) cudaStreamSynchronize(h->s_cmp[i]);
  if (h->s_out) cudaStreamSynchronize(h->s_out);
  if (h->stream) cudaStreamSynchronize(h->stream);
}

// Reads and clears the device-side status word (bit 0: k_octree ran out of node pool and returned an
// empty level).  The work that could set it must have completed.
static int consume_status(orbx_extractor_s* h) {
  int32_t st = 0;
  cudaError_t e = cudaMemcpy(&st, h->status, sizeof(int32_t), cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) { set_err("status readback", e); return ORBX_E_CUDA; }
  if (st == 0) return ORBX_OK;
  cudaMemset(h->status, 0, sizeof(int32_t));
  set_err("internal bound violated (octree node pool overflow): this call's keypoints are incomplete", cudaSuccess);
  return ORBX_E_INTERNAL;
}

extern "C" {

const char* orbx_version(void) { return "orbx 0.1 (sm_100a)"; }
const char* orbx_last_error(void) { return g_err; }

void orbx_destroy(orbx_handle h) {
  if (!h) return;
  cudaSetDevice(h->device);
  cudaFree(h->arena); cudaFree(h->xtaps); cudaFree(h->ytaps); cudaFree(h->cell_cnt); cudaFree(h->cell_tab); cudaFree(h->strip_tab); cudaFree(h->cand);
  cudaFree(h->gkeys); cudaFree(h->glabels); cudaFree(h->sel); cudaFree(h->sel_cnt); cudaFree(h->dst);
  cudaFree(h->status); cudaFree(h->o_kps); cudaFree(h->o_desc); cudaFree(h->o_n); cudaFree(h->o_mono);
  cudaFree(h->border_tmp); cudaFree(h->stage); cudaFree(h->blur);
  cudaFree(h->pyr_stage_d); if (h->pyr_stage_h) cudaFreeHost(h->pyr_stage_h);
  for (int i = 0; i <= ORBX_NUM_STAGES; i++)
    if (h->ev[i]) cudaEventDestroy(h->ev[i]);
  for (cudaEvent_t e : h->ev_in) cudaEventDestroy(e);
  for (cudaEvent_t e : h->ev_done) cudaEventDestroy(e);
  if (h->ev_fork) cudaEventDestroy(h->ev_fork);
  for (int i = 0; i < 4; i++)
    if (h->ev_join[i]) cudaEventDestroy(h->ev_join[i]);
  for (int i = 0; i < 4; i++) {
    if (h->s_side[i]) { cudaStreamSynchronize(h->s_side[i]); cudaStreamDestroy(h->s_side[i]); }
    if (h->s_hi[i]) { cudaStreamSynchronize(h->s_hi[i]); cudaStreamDestroy(h->s_hi[i]); }
    if (h->ev_fast[i]) cudaEventDestroy(h->ev_fast[i]);
    if (h->ev_oct[i]) cudaEventDestroy(h->ev_oct[i]);
    if (h->ev_pyr[i]) cudaEventDestroy(h->ev_pyr[i]);
    if (h->ev_blur[i]) cudaEventDestroy(h->ev_blur[i]);
  }
  if (h->s_in) cudaStreamDestroy(h->s_in);
  if (h->s_out) cudaStreamDestroy(h->s_out);
  for (int i = 0; i < 4; i++)
    if (h->s_cmp[i]) cudaStreamDestroy(h->s_cmp[i]);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
}

int orbx_create(const orbx_config* cfg, orbx_handle* out) {
  if (!cfg || !out) return ORBX_E_INVALID;
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    set_err("no CUDA device (this library has no CPU fallback)", e);
    return ORBX_E_NODEVICE;
  }
  if (cfg->device < 0 || cfg->device >= ndev || cfg->max_batch < 1) return ORBX_E_INVALID;
  if (cfg->width > 4095 || cfg->height > 4095) { set_err("frame larger than 4095 px", cudaSuccess); return ORBX_E_INVALID; }
  orbx_extractor_s* h = new (std::nothrow) orbx_extractor_s();
  if (!h) return ORBX_E_NOMEM;
  h->cfg = *cfg;
  h->device = cfg->device;
  h->stream = nullptr;
  std::vector<OrbxTap> xt, yt;
  if (orbx_build_plan(cfg->nfeatures, cfg->scale_factor, cfg->nlevels, cfg->ini_th_fast, cfg->min_th_fast,
                      cfg->width, cfg->height, &h->plan, &xt, &yt) != 0) {
    set_err("unsupported geometry (level smaller than one 35-px cell, portrait aspect, or exact 2x scale)",
            cudaSuccess);
    delete h;
    return ORBX_E_INVALID;
  }
  const OrbxPlan& P = h->plan;
  int node_cap = 0, max_cells = 0;
  for (int l = 0; l < P.nlevels; l++) {
    if (P.lv[l].nIni > 64) { delete h; set_err("aspect ratio > 64", cudaSuccess); return ORBX_E_INVALID; }
    node_cap = node_cap > P.lv[l].kp_max + 4 ? node_cap : P.lv[l].kp_max + 4;  // list never exceeds kp_max
    const int nc = P.lv[l].nCols * P.lv[l].nRows;
    max_cells = max_cells > nc ? max_cells : nc;
  }
  if (node_cap > 8000) { delete h; set_err("nfeatures too large for 16-bit node positions", cudaSuccess); return ORBX_E_INVALID; }
  fill_devplan(P, &h->dplan);
  h->oct_node_cap = node_cap;
  h->oct_max_cells = max_cells;
  // keys that fit next to the node pool in <= ~96 KB of shared memory; larger levels spill to HBM
  {
    const size_t fixed = orbx_octree_smem_bytes(0, node_cap, max_cells);
    // The kernel is latency bound (one thread replays the list surgery), so residency matters more
    // than keeping every key on chip: 2048 keys cover the usual level (<~1000 candidates).
    const size_t budget = 96 * 1024;
    long long keys = fixed < budget ? (long long)(budget - fixed) / 6 : 0;
    if (keys > 2048) keys = 2048;
    h->oct_smem_keys = (int)(keys & ~7LL);
    h->oct_smem = orbx_octree_smem_bytes(h->oct_smem_keys, node_cap, max_cells);
    if (h->oct_smem > 220 * 1024) { delete h; set_err("octree shared memory", cudaSuccess); return ORBX_E_INVALID; }
  }
  const long long B = cfg->max_batch;
#define ALLOC(ptr, bytes)                                                        \
  do {                                                                           \
    cudaError_t e_ = cudaMalloc((void**)&(ptr), (size_t)(bytes));                \
    if (e_ != cudaSuccess) { set_err("cudaMalloc " #ptr, e_); orbx_destroy(h); return e_ == cudaErrorMemoryAllocation ? ORBX_E_NOMEM : ORBX_E_CUDA; } \
  } while (0)
  if ((e = cudaSetDevice(h->device)) != cudaSuccess) { set_err("cudaSetDevice", e); delete h; return ORBX_E_CUDA; }
  if ((e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking)) != cudaSuccess) {
    set_err("cudaStreamCreate", e); delete h; return ORBX_E_CUDA;
  }
  ALLOC(h->arena, B * P.arena_bytes + 512);
  ALLOC(h->xtaps, (xt.size() + 1) * sizeof(OrbxTap));
  ALLOC(h->ytaps, (yt.size() + 1) * sizeof(OrbxTap));
  ALLOC(h->cell_cnt, B * P.total_cells * sizeof(int32_t));
  ALLOC(h->cell_tab, (size_t)P.total_cells * sizeof(uint32_t));
  ALLOC(h->strip_tab, (size_t)P.total_strips * sizeof(uint32_t));
  ALLOC(h->cand, B * P.cand_total * sizeof(uint32_t));
  ALLOC(h->gkeys, B * P.cand_total * sizeof(uint32_t));
  ALLOC(h->glabels, B * P.cand_total * sizeof(int16_t));
  ALLOC(h->sel, B * P.kp_total_cap * sizeof(uint32_t));
  ALLOC(h->sel_cnt, B * P.nlevels * sizeof(int32_t));
  ALLOC(h->dst, B * P.kp_total_cap * sizeof(int32_t));
  ALLOC(h->status, sizeof(int32_t));
  ALLOC(h->o_kps, B * P.max_keypoints * sizeof(OrbxKeyPointDev));
  ALLOC(h->o_desc, B * P.max_keypoints * 32);
  ALLOC(h->o_n, B * sizeof(int32_t));
  ALLOC(h->o_mono, B * sizeof(int32_t));
  {
    // Descriptor stage mode.  Blurring whole levels (k_blur_levels, then a light per-keypoint sampler) costs per
    // level pixel; blurring each keypoint's own 45x45 neighbourhood in shared memory (fused k_describe) costs per
    // keypoint and writes no blurred level.  Measured on B200 at 752x480 / 1000 kp (1078 level pixels per keypoint)
    // the level mode is 3 % faster per step; at 4K / 5000 kp (5000 px per keypoint) the fused mode wins clearly.
    long long px = 0;
    for (int l = 0; l < P.nlevels; l++) px += (long long)P.lv[l].w * P.lv[l].h;
    h->level_blur = px < 1200LL * P.max_keypoints;
    if (cfg->flags & ORBX_FLAG_LEVEL_BLUR) h->level_blur = true;
    if (cfg->flags & ORBX_FLAG_FUSED_BLUR) h->level_blur = false;
    if (h->level_blur) ALLOC(h->blur, B * P.arena_bytes + 512);
  }
  ALLOC(h->border_tmp, (size_t)(P.lv[0].w + 2 * ORBX_EDGE) * (P.lv[0].h + 2 * ORBX_EDGE));
#undef ALLOC
  // from here on a failure must release the handle and its device buffers
#define CUH(call)                                                     \
  do {                                                                \
    cudaError_t e_ = (call);                                          \
    if (e_ != cudaSuccess) { set_err(#call, e_); orbx_destroy(h); return ORBX_E_CUDA; } \
  } while (0)
  if (!xt.empty()) {
    CUH(cudaMemcpy(h->xtaps, xt.data(), xt.size() * sizeof(OrbxTap), cudaMemcpyHostToDevice));
    CUH(cudaMemcpy(h->ytaps, yt.data(), yt.size() * sizeof(OrbxTap), cudaMemcpyHostToDevice));
  }
  {
    std::vector<uint32_t> tab(P.total_cells);
    for (int l = 0; l < P.nlevels; l++)
      for (int c = 0; c < P.lv[l].nCols * P.lv[l].nRows; c++)
        tab[P.lv[l].cell_base + c] = ((uint32_t)l << 24) | ((uint32_t)(c / P.lv[l].nCols) << 12) | (uint32_t)(c % P.lv[l].nCols);
    CUH(cudaMemcpy(h->cell_tab, tab.data(), tab.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
    std::vector<uint32_t> stab(P.total_strips);
    for (int l = 0; l < P.nlevels; l++)
      for (int i = 0; i < P.lv[l].nRows; i++)
        for (int sj = 0; sj < P.lv[l].strips_per_row; sj++)
          stab[P.lv[l].strip_base + i * P.lv[l].strips_per_row + sj] =
              ((uint32_t)l << 24) | ((uint32_t)i << 12) | (uint32_t)(sj * P.lv[l].strip_k);
    CUH(cudaMemcpy(h->strip_tab, stab.data(), stab.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
  }
  CUH(cudaMemset(h->status, 0, sizeof(int32_t)));
  CUH(cudaMemset(h->arena, 0, B * P.arena_bytes + 512));
  orbx_fast_maps_init(h->dplan, h->arena, cfg->max_batch, &h->fast_maps);
  orbx_resize_maps_init(h->dplan, h->arena, cfg->max_batch, &h->resize_maps);
  CUH(orbx_describe_init_tables(P.umax));
  if (cfg->flags & ORBX_FLAG_STAGE_TIMING)
    for (int i = 0; i <= ORBX_NUM_STAGES; i++) CUH(cudaEventCreate(&h->ev[i]));
#undef CUH
  h->last_l0.base = h->arena; h->last_l0.frame_stride = P.arena_bytes; h->last_l0.pitch = P.lv[0].pitch;
  *out = h;
  return ORBX_OK;
}

int orbx_max_keypoints(orbx_handle h) { return h ? h->plan.max_keypoints : ORBX_E_INVALID; }

// Runs the launch sequence for `batch` frames whose per-frame scratch (pyramid arena, candidate
// slots, selections) starts at frame slot f0 of the handle; l0 and the outputs point at the first
// of those frames.
static int ensure_streams(orbx_handle h) {
  if (h->streams_ready) return ORBX_OK;
  // all or nothing: a partial failure releases what was created, so a later call starts over
  cudaError_t e = cudaSuccess;
  cudaStream_t* ss[10] = {&h->s_in, &h->s_out, &h->s_cmp[0], &h->s_cmp[1], &h->s_cmp[2], &h->s_cmp[3],
                          &h->s_side[0], &h->s_side[1], &h->s_side[2], &h->s_side[3]};
  cudaEvent_t* es[13] = {&h->ev_fork, &h->ev_join[0], &h->ev_join[1], &h->ev_join[2], &h->ev_join[3],
                         &h->ev_pyr[0], &h->ev_pyr[1], &h->ev_pyr[2], &h->ev_pyr[3],
                         &h->ev_blur[0], &h->ev_blur[1], &h->ev_blur[2], &h->ev_blur[3]};
  for (int i = 0; i < 10 && e == cudaSuccess; i++) e = cudaStreamCreateWithFlags(ss[i], cudaStreamNonBlocking);
  for (int i = 0; i < 13 && e == cudaSuccess; i++) e = cudaEventCreateWithFlags(es[i], cudaEventDisableTiming);
  if (e != cudaSuccess) {
    set_err("creating the handle's internal streams", e);
    for (int i = 0; i < 10; i++) if (*ss[i]) { cudaStreamDestroy(*ss[i]); *ss[i] = nullptr; }
    for (int i = 0; i < 13; i++) if (*es[i]) { cudaEventDestroy(*es[i]); *es[i] = nullptr; }
    return ORBX_E_CUDA;
  }
  h->streams_ready = true;
  if (const char* e = getenv("ORBX_HOST_CHUNK")) { const int v = atoi(e); if (v > 0) h->host_chunk = v; }
  if (const char* e = getenv("ORBX_HOST_STREAMS")) { const int v = atoi(e); if (v >= 1 && v <= 4) h->host_streams = v; }
  if (const char* e = getenv("ORBX_HOST_EDGE")) { const int v = atoi(e); if (v >= 0) h->host_edge = v; }
  if (const char* e = getenv("ORBX_BLUR_SIDE")) h->blur_side = atoi(e);
  // Octree on a high-priority twin of the sub-batch stream: its CTAs then take SM slots as soon as FAST is done
  // instead of queueing behind the other sub-batch's kernels.  Measured (profiles/r2ac_octprio.txt): +2.3 % at
  // 4K (1024-thread octree CTAs, 32 frames), -8 % at 752x480 x 512, so on only for the wide-CTA octree.
  h->oct_prio = h->oct_node_cap >= 768 ? 1 : 0;
  if (const char* e = getenv("ORBX_OCT_PRIO")) h->oct_prio = atoi(e);
  if (h->oct_prio) {
    int lo = 0, hi = 0;
    cudaDeviceGetStreamPriorityRange(&lo, &hi);
    for (int i = 0; i < 4; i++) {
      if (cudaStreamCreateWithPriority(&h->s_hi[i], cudaStreamNonBlocking, hi) != cudaSuccess ||
          cudaEventCreateWithFlags(&h->ev_fast[i], cudaEventDisableTiming) != cudaSuccess ||
          cudaEventCreateWithFlags(&h->ev_oct[i], cudaEventDisableTiming) != cudaSuccess) { h->oct_prio = 0; break; }
    }
  }
  if (const char* e = getenv("ORBX_DEVICE_SPLIT")) { const int v = atoi(e); if (v > 0) h->dev_split = v; }
  if (const char* e = getenv("ORBX_DEVICE_STREAMS")) { const int v = atoi(e); if (v >= 1 && v <= 4) h->dev_streams = v; }
  return ORBX_OK;
}

static int run_pipeline(orbx_handle h, Level0 l0, int batch, int lap0, int lap1, OrbxKeyPointDev* d_kps,
                        uint8_t* d_desc, int capacity, int32_t* d_n, int32_t* d_mono, cudaStream_t st,
                        int f0 = 0, bool allow_timing = true, int side = -1) {
  const DevPlan& D = h->dplan;
  const bool tm = allow_timing && (h->cfg.flags & ORBX_FLAG_STAGE_TIMING) != 0;
  uint8_t* arena = h->arena + (long long)f0 * D.arena_bytes;
  int32_t* cell_cnt = h->cell_cnt + (long long)f0 * D.total_cells;
  uint32_t* cand = h->cand + (long long)f0 * D.cand_total;
  uint32_t* gkeys = h->gkeys + (long long)f0 * D.cand_total;
  int16_t* glabels = h->glabels + (long long)f0 * D.cand_total;
  uint32_t* sel = h->sel + (long long)f0 * D.kp_total_cap;
  int32_t* sel_cnt = h->sel_cnt + (long long)f0 * D.nlevels;
  int32_t* dst = h->dst + (long long)f0 * D.kp_total_cap;
#define STAGE_MARK(i) do { if (tm) CU(cudaEventRecord(h->ev[i], st)); } while (0)
  STAGE_MARK(0);
  if (((uintptr_t)l0.base | (uintptr_t)(unsigned)l0.pitch | (uintptr_t)l0.frame_stride) & 15u) {
    // unaligned caller buffer: one pass into the arena's (unused in batch mode) level-0 slot, see pyramid.cu
    CU(orbx_launch_repack_level0(D, l0, arena, batch, st));
    l0.base = arena + D.lv[0].offset; l0.pitch = D.lv[0].pitch; l0.frame_stride = D.arena_bytes;
  }
  CU(orbx_launch_pyramid(D, h->resize_maps, l0, arena, h->xtaps, h->ytaps, batch, f0, st));
  STAGE_MARK(1);
  const bool blur_aside = h->level_blur && side >= 0 && h->blur_side && !tm;
  if (blur_aside) {
    CU(cudaEventRecord(h->ev_pyr[side], st));
    CU(cudaStreamWaitEvent(h->s_side[side], h->ev_pyr[side], 0));
    CU(orbx_launch_blur(D, l0, arena, h->blur + (long long)f0 * D.arena_bytes, batch, h->s_side[side]));
    CU(cudaEventRecord(h->ev_blur[side], h->s_side[side]));
  }
  CU(orbx_launch_fast(D, h->fast_maps, l0, arena, h->cell_tab, h->strip_tab, cell_cnt, cand, batch, f0, st));
  STAGE_MARK(2);
  const bool oct_hi = side >= 0 && h->oct_prio && !tm;
  cudaStream_t so = oct_hi ? h->s_hi[side] : st;
  if (oct_hi) {
    CU(cudaEventRecord(h->ev_fast[side], st));
    CU(cudaStreamWaitEvent(so, h->ev_fast[side], 0));
  }
  CU(orbx_launch_octree(D, cell_cnt, cand, gkeys, glabels, sel, sel_cnt, h->status, batch,
                        h->oct_smem, h->oct_smem_keys, h->oct_node_cap, h->oct_max_cells, so));
  STAGE_MARK(3);
  CU(orbx_launch_order(D, sel, sel_cnt, lap0, lap1, dst, d_n, d_mono, batch, so));
  STAGE_MARK(4);
  if (oct_hi) {
    if (h->oct_prio == 2 && h->level_blur) {
      // the blur does not depend on the selection: let it start beside the octree
      CU(orbx_launch_blur(D, l0, arena, h->blur + (long long)f0 * D.arena_bytes, batch, st));
    }
    CU(cudaEventRecord(h->ev_oct[side], so));
    CU(cudaStreamWaitEvent(st, h->ev_oct[side], 0));
  }
  if (h->level_blur) {
    uint8_t* blur = h->blur + (long long)f0 * D.arena_bytes;
    if (blur_aside) CU(cudaStreamWaitEvent(st, h->ev_blur[side], 0));
    else if (!(oct_hi && h->oct_prio == 2)) CU(orbx_launch_blur(D, l0, arena, blur, batch, st));
    CU(orbx_launch_describe_lite(D, l0, arena, blur, sel, sel_cnt, dst, d_kps, d_desc, capacity,
                                 (h->cfg.flags & ORBX_FLAG_STRICT_FP) ? 1 : 0, batch, st));
  } else {
    CU(orbx_launch_describe(D, l0, arena, sel, sel_cnt, dst, d_kps, d_desc, capacity,
                            (h->cfg.flags & ORBX_FLAG_STRICT_FP) ? 1 : 0, batch, st));
  }
  STAGE_MARK(5);
#undef STAGE_MARK
  if (allow_timing) h->timed = tm;
  if (f0 == 0) {
    h->last_l0 = l0;
    h->last_batch = batch;
  }
  return ORBX_OK;
}

int orbx_extract_batch_device(orbx_handle h, const uint8_t* d_imgs, size_t pitch, size_t frame_stride,
                              int batch, int lap0, int lap1, orbx_keypoint* d_kps, uint8_t* d_desc,
                              int capacity, int32_t* d_n_out, int32_t* d_mono_index, void* stream) {
  if (!h || !d_imgs || !d_kps || !d_desc || !d_n_out || !d_mono_index) return ORBX_E_INVALID;
  if (batch < 1 || batch > h->cfg.max_batch || (int)pitch < h->plan.width) return ORBX_E_INVALID;
  if (capacity < h->plan.max_keypoints) { set_err("capacity < orbx_max_keypoints()", cudaSuccess); return ORBX_E_CAPACITY; }
  CU(cudaSetDevice(h->device));
  Level0 l0;
  l0.base = d_imgs; l0.frame_stride = (long long)frame_stride; l0.pitch = (int)pitch;
  cudaStream_t st = (cudaStream_t)stream;
  // Large batches are cut into sub-batches that alternate between two internal streams (forked from and
  // joined back into the caller's stream): the latency-bound octree of one sub-batch then shares the SMs
  // with the issue-bound FAST / describe kernels of its neighbour instead of running alone.  Sub-batches
  // own disjoint frame slots of the scratch buffers.  Off while per-stage events are being recorded.
  int rc = ensure_streams(h);
  if (rc) return rc;
  // a sub-batch should still fill the GPU: >= 64 frames of 752x480, fewer of larger frames (3 at 4K)
  int nsplit = h->dev_split;
  const long long px = (long long)h->plan.width * h->plan.height;
  long long min_frames = (64LL * 752 * 480 + px - 1) / px;
  if (min_frames < 1) min_frames = 1;
  while (nsplit > 1 && batch < min_frames * nsplit) nsplit >>= 1;
  if (h->cfg.flags & ORBX_FLAG_STAGE_TIMING) nsplit = 1;
  const int nstreams = h->dev_streams < nsplit ? h->dev_streams : nsplit;
  if (nsplit <= 1)
    return run_pipeline(h, l0, batch, lap0, lap1, (OrbxKeyPointDev*)d_kps, d_desc, capacity, d_n_out, d_mono_index, st);
  CU(cudaEventRecord(h->ev_fork, st));
  for (int i = 0; i < nstreams; i++) CU(cudaStreamWaitEvent(h->s_cmp[i], h->ev_fork, 0));
  const int chunk = (batch + nsplit - 1) / nsplit;
  for (int c = 0, f0 = 0; f0 < batch; c++, f0 += chunk) {
    const int nb = (batch - f0) < chunk ? (batch - f0) : chunk;
    Level0 l = l0;
    l.base = l0.base + (long long)f0 * l0.frame_stride;
    rc = run_pipeline(h, l, nb, lap0, lap1, (OrbxKeyPointDev*)d_kps + (long long)f0 * capacity,
                      d_desc + (long long)f0 * capacity * 32, capacity, d_n_out + f0, d_mono_index + f0,
                      h->s_cmp[c % nstreams], f0, false, c % nstreams);
    if (rc) break;  // still join the forked streams back into the caller's stream below
  }
  for (int i = 0; i < nstreams; i++) {
    CU(cudaEventRecord(h->ev_join[i], h->s_cmp[i]));
    CU(cudaStreamWaitEvent(st, h->ev_join[i], 0));
  }
  h->last_l0 = l0;
  h->last_batch = batch;
  h->timed = false;
  return rc;
}

int orbx_check_status(orbx_handle h, void* stream) {
  if (!h) return ORBX_E_INVALID;
  CU(cudaSetDevice(h->device));
  CU(cudaStreamSynchronize((cudaStream_t)stream));
  return consume_status(h);
}

int orbx_extract_batch_host(orbx_handle h, const uint8_t* imgs, int batch, int lap0, int lap1,
                            orbx_keypoint* kps, uint8_t* desc, int capacity, int32_t* n_out,
                            int32_t* mono_index) {
  if (!h || !imgs || !kps || !desc || !n_out || !mono_index) return ORBX_E_INVALID;
  if (batch < 1 || batch > h->cfg.max_batch) return ORBX_E_INVALID;
  const OrbxPlan& P = h->plan;
  if (capacity < P.max_keypoints) { set_err("capacity < orbx_max_keypoints()", cudaSuccess); return ORBX_E_CAPACITY; }
  CU(cudaSetDevice(h->device));
  const int W = P.width, H = P.height, mk = P.max_keypoints;
  // Chunked pipeline: while chunk c is being extracted, chunk c+1 is on its way in and chunk c-1 on its
  // way out (PCIe is full duplex).  Chunks own disjoint frame slots of the handle's scratch, so two
  // compute streams may run neighbouring chunks concurrently and fill each other's kernel tails.
  { int rc_ = ensure_streams(h); if (rc_) return rc_; }
  if (!h->stage) {
    cudaError_t e_ = cudaMalloc((void**)&h->stage, (size_t)h->cfg.max_batch * W * H + 512);
    if (e_ != cudaSuccess) { set_err("cudaMalloc stage", e_); return e_ == cudaErrorMemoryAllocation ? ORBX_E_NOMEM : ORBX_E_CUDA; }
  }
  const int chunk = h->host_chunk < batch ? h->host_chunk : batch;
  // Chunk schedule: full chunks in the middle, a shorter first chunk (the kernels start sooner) and a shorter
  // last one (less work left once the last frame has arrived), when the batch is long enough to afford it.
  std::vector<int> sched;
  {
    const int edge = h->host_edge > 0 && h->host_edge < chunk && batch >= 4 * chunk ? h->host_edge : 0;
    int left = batch;
    if (edge) { sched.push_back(edge); left -= 2 * edge; }
    while (left > 0) { const int c_ = left < chunk ? left : chunk; sched.push_back(c_); left -= c_; }
    if (edge) sched.push_back(edge);
  }
  const int nchunks = (int)sched.size();
  while ((int)h->ev_in.size() < nchunks) {
    cudaEvent_t a, b;
    CU(cudaEventCreateWithFlags(&a, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&b, cudaEventDisableTiming));
    h->ev_in.push_back(a); h->ev_done.push_back(b);
  }
  // anything still running on the handle's own stream (single-frame API) must be done first
  CU(cudaStreamSynchronize(h->stream));
  int f_next = 0;
  for (int c = 0; c < nchunks; c++) {
    const int f0 = f_next, nb = sched[c];
    f_next += nb;
    cudaStream_t cs = h->s_cmp[c % h->host_streams];
    // One contiguous H2D per chunk into the tightly packed staging frames, which then serve as
    // level 0 directly (row pitch = W): strided 2-D copies of 752-byte rows run far below PCIe rate.
    CU(cudaMemcpyAsync(h->stage + (size_t)f0 * W * H, imgs + (size_t)f0 * W * H, (size_t)nb * W * H,
                       cudaMemcpyHostToDevice, h->s_in));
    CU(cudaEventRecord(h->ev_in[c], h->s_in));
    CU(cudaStreamWaitEvent(cs, h->ev_in[c], 0));
    Level0 l0;
    l0.base = h->stage + (size_t)f0 * W * H; l0.frame_stride = (long long)W * H; l0.pitch = W;
    int rc = run_pipeline(h, l0, nb, lap0, lap1, h->o_kps + (long long)f0 * mk, h->o_desc + (long long)f0 * mk * 32, mk,
                          h->o_n + f0, h->o_mono + f0, cs, f0, false);
    if (rc) { drain_streams(h); return rc; }
    CU(cudaEventRecord(h->ev_done[c], cs));
    CU(cudaStreamWaitEvent(h->s_out, h->ev_done[c], 0));
    CU(cudaMemcpyAsync(n_out + f0, h->o_n + f0, nb * sizeof(int32_t), cudaMemcpyDeviceToHost, h->s_out));
    CU(cudaMemcpyAsync(mono_index + f0, h->o_mono + f0, nb * sizeof(int32_t), cudaMemcpyDeviceToHost, h->s_out));
    CU(cudaMemcpy2DAsync(kps + (size_t)f0 * capacity, (size_t)capacity * sizeof(orbx_keypoint),
                         h->o_kps + (long long)f0 * mk, (size_t)mk * sizeof(orbx_keypoint),
                         (size_t)mk * sizeof(orbx_keypoint), nb, cudaMemcpyDeviceToHost, h->s_out));
    CU(cudaMemcpy2DAsync(desc + (size_t)f0 * capacity * 32, (size_t)capacity * 32, h->o_desc + (long long)f0 * mk * 32,
                         (size_t)mk * 32, (size_t)mk * 32, nb, cudaMemcpyDeviceToHost, h->s_out));
  }
  h->last_l0.base = h->stage; h->last_l0.frame_stride = (long long)W * H; h->last_l0.pitch = W;
  h->last_batch = batch;
  CU(cudaStreamSynchronize(h->s_out));
  for (int i = 0; i < h->host_streams; i++) CU(cudaStreamSynchronize(h->s_cmp[i]));
  return consume_status(h);
}

/* Device copies of what the last orbx_extract / orbx_extract_batch_host call returned (so that the
 * matcher entry points can run on them without a second upload). */
int orbx_last_outputs_device(orbx_handle h, const orbx_keypoint** d_kps, const uint8_t** d_desc,
                             const int32_t** d_n, const int32_t** d_mono, int* capacity) {
  if (!h) return ORBX_E_INVALID;
  if (d_kps) *d_kps = (const orbx_keypoint*)h->o_kps;
  if (d_desc) *d_desc = h->o_desc;
  if (d_n) *d_n = h->o_n;
  if (d_mono) *d_mono = h->o_mono;
  if (capacity) *capacity = h->plan.max_keypoints;
  return ORBX_OK;
}

int orbx_extract(orbx_handle h, const uint8_t* img, int width, int height, size_t stride, int lap0, int lap1,
                 orbx_keypoint* kps, uint8_t* desc, int capacity, int* n_out, int* mono_index) {
  if (!h || !kps || !desc || !n_out || !mono_index) return ORBX_E_INVALID;
  if (!img || width <= 0 || height <= 0) return ORBX_E_EMPTY;  // operator() returns -1 (:1090-1091)
  const OrbxPlan& P = h->plan;
  if (width != P.width || height != P.height || stride < (size_t)width) return ORBX_E_INVALID;
  CU(cudaSetDevice(h->device));
  cudaStream_t st = h->stream;
  CU(cudaMemcpy2DAsync(h->arena, P.lv[0].pitch, img, stride, width, height, cudaMemcpyHostToDevice, st));
  Level0 l0;
  l0.base = h->arena; l0.frame_stride = P.arena_bytes; l0.pitch = P.lv[0].pitch;
  int rc = run_pipeline(h, l0, 1, lap0, lap1, h->o_kps, h->o_desc, P.max_keypoints, h->o_n, h->o_mono, st);
  if (rc) return rc;
  int32_t nm[2];
  CU(cudaMemcpyAsync(&nm[0], h->o_n, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(&nm[1], h->o_mono, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  *n_out = nm[0];
  *mono_index = nm[1];
  { const int rs = consume_status(h); if (rs) return rs; }
  if (nm[0] > capacity) { set_err("keypoint buffer too small", cudaSuccess); return ORBX_E_CAPACITY; }
  if (nm[0] > 0) {
    CU(cudaMemcpyAsync(kps, h->o_kps, (size_t)nm[0] * sizeof(orbx_keypoint), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(desc, h->o_desc, (size_t)nm[0] * 32, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
  }
  return ORBX_OK;
}

int orbx_level_size(orbx_handle h, int level, int* width, int* height) {
  if (!h || level < 0 || level >= h->plan.nlevels) return ORBX_E_INVALID;
  if (width) *width = h->plan.lv[level].w;
  if (height) *height = h->plan.lv[level].h;
  return ORBX_OK;
}

int orbx_pyramid_device(orbx_handle h, int frame, int level, const uint8_t** d_ptr, size_t* pitch, int* width,
                        int* height) {
  if (!h || level < 0 || level >= h->plan.nlevels || frame < 0 || frame >= h->cfg.max_batch) return ORBX_E_INVALID;
  const OrbxLevel& L = h->plan.lv[level];
  if (level == 0) {
    if (d_ptr) *d_ptr = h->last_l0.base + (long long)frame * h->last_l0.frame_stride;
    if (pitch) *pitch = (size_t)h->last_l0.pitch;
  } else {
    if (d_ptr) *d_ptr = h->arena + (long long)frame * h->plan.arena_bytes + L.offset;
    if (pitch) *pitch = (size_t)L.pitch;
  }
  if (width) *width = L.w;
  if (height) *height = L.h;
  return ORBX_OK;
}

int orbx_get_pyramid(orbx_handle h, int frame, int level, uint8_t* dst, size_t dst_stride, int with_border) {
  const uint8_t* p; size_t pitch; int w, hh;
  int rc = orbx_pyramid_device(h, frame, level, &p, &pitch, &w, &hh);
  if (rc) return rc;
  if (!dst) return ORBX_E_INVALID;
  CU(cudaSetDevice(h->device));
  CU(cudaDeviceSynchronize());  // the last call may have run on a caller stream
  if (!with_border) {
    if (dst_stride < (size_t)w) return ORBX_E_INVALID;
    CU(cudaMemcpy2DAsync(dst, dst_stride, p, pitch, w, hh, cudaMemcpyDeviceToHost, h->stream));
  } else {
    const int bw = w + 2 * ORBX_EDGE, bh = hh + 2 * ORBX_EDGE;
    if (dst_stride < (size_t)bw) return ORBX_E_INVALID;
    CU(orbx_launch_border(p, (int)pitch, w, hh, h->border_tmp, bw, h->stream));
    CU(cudaMemcpy2DAsync(dst, dst_stride, h->border_tmp, bw, bw, bh, cudaMemcpyDeviceToHost, h->stream));
  }
  CU(cudaStreamSynchronize(h->stream));
  return ORBX_OK;
}

// All levels of one frame in ONE device->host transfer: the levels (with their 19-px REFLECT_101 ring when
// with_border) are packed into a device staging buffer by L small kernels / copies and land in a pinned host buffer
// owned by the handle.  This is what the compat operator() needs after every call (mvImagePyramid, ORBextractor.h:83);
// eight orbx_get_pyramid calls cost eight synchronisations.
int orbx_get_pyramid_all(orbx_handle h, int frame, int with_border, const uint8_t** host_base, size_t* offsets,
                         size_t* strides) {
  if (!h || !host_base || !offsets || !strides || frame < 0 || frame >= h->cfg.max_batch) return ORBX_E_INVALID;
  const OrbxPlan& P = h->plan;
  const int e = with_border ? ORBX_EDGE : 0;
  size_t total = 0;
  for (int l = 0; l < P.nlevels; l++) {
    offsets[l] = total;
    strides[l] = (size_t)(P.lv[l].w + 2 * e);
    total += ((strides[l] * (size_t)(P.lv[l].h + 2 * e)) + 63) & ~(size_t)63;
  }
  CU(cudaSetDevice(h->device));
  if (h->pyr_stage_bytes < total) {
    cudaFree(h->pyr_stage_d); h->pyr_stage_d = nullptr;
    if (h->pyr_stage_h) { cudaFreeHost(h->pyr_stage_h); h->pyr_stage_h = nullptr; }
    h->pyr_stage_bytes = 0;
    // both ring modes fit in the bordered size
    size_t cap = 0;
    for (int l = 0; l < P.nlevels; l++)
      cap += (((size_t)(P.lv[l].w + 2 * ORBX_EDGE) * (size_t)(P.lv[l].h + 2 * ORBX_EDGE)) + 63) & ~(size_t)63;
    if (cudaMalloc(&h->pyr_stage_d, cap) != cudaSuccess) { cudaGetLastError(); return ORBX_E_NOMEM; }
    if (cudaHostAlloc((void**)&h->pyr_stage_h, cap, cudaHostAllocDefault) != cudaSuccess) {
      cudaGetLastError(); cudaFree(h->pyr_stage_d); h->pyr_stage_d = nullptr; return ORBX_E_NOMEM;
    }
    h->pyr_stage_bytes = cap;
  }
  CU(cudaDeviceSynchronize());  // the last call may have run on a caller stream
  for (int l = 0; l < P.nlevels; l++) {
    const uint8_t* p; size_t pitch; int w, hh;
    const int rc = orbx_pyramid_device(h, frame, l, &p, &pitch, &w, &hh);
    if (rc) return rc;
    if (with_border)
      CU(orbx_launch_border(p, (int)pitch, w, hh, h->pyr_stage_d + offsets[l], (int)strides[l], h->stream));
    else
      CU(cudaMemcpy2DAsync(h->pyr_stage_d + offsets[l], strides[l], p, pitch, w, hh, cudaMemcpyDeviceToDevice, h->stream));
  }
  CU(cudaMemcpyAsync(h->pyr_stage_h, h->pyr_stage_d, total, cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  *host_base = h->pyr_stage_h;
  return ORBX_OK;
}

int orbx_get_scale_tables(orbx_handle h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2,
                          int32_t* features_per_level) {
  if (!h) return ORBX_E_INVALID;
  for (int l = 0; l < h->plan.nlevels; l++) {
    if (scale) scale[l] = h->plan.scale[l];
    if (inv_scale) inv_scale[l] = h->plan.inv_scale[l];
    if (sigma2) sigma2[l] = h->plan.sigma2[l];
    if (inv_sigma2) inv_sigma2[l] = h->plan.inv_sigma2[l];
    if (features_per_level) features_per_level[l] = h->plan.lv[l].nfeat;
  }
  return ORBX_OK;
}

// Sub-pixel stage + median cut of Frame::ComputeStereoMatches (Frame.cc:896-981), reading both eyes'
// pyramids where the two extractors left them (frame f of each handle's last batch).
int orbm_stereo_refine(orbx_handle hl, orbx_handle hr, const orbx_keypoint* d_kps_l, const int32_t* d_n_l,
                       const orbx_keypoint* d_kps_r, int kp_stride, int batch, const int32_t* d_best_idx_r,
                       const int32_t* d_best_dist, float mbf, float max_d, float* d_u_right, float* d_depth,
                       int32_t* d_sad, int32_t* d_kept, void* stream) {
  if (!hl || !hr || !d_kps_l || !d_n_l || !d_kps_r || !d_best_idx_r || !d_best_dist || !d_u_right || !d_depth ||
      !d_sad || !d_kept || batch < 1 || kp_stride < 1)
    return ORBX_E_INVALID;
  if (hl->device != hr->device || hl->plan.nlevels != hr->plan.nlevels || batch > hl->last_batch ||
      batch > hr->last_batch) {
    set_err("orbm_stereo_refine: handles disagree or no extraction of this batch yet", cudaSuccess);
    return ORBX_E_INVALID;
  }
  OrbxStereoPyr pyr;
  memset(&pyr, 0, sizeof(pyr));
  for (int l = 0; l < hl->plan.nlevels; l++) {
    const OrbxLevel &A = hl->plan.lv[l], &B = hr->plan.lv[l];
    if (A.w != B.w || A.h != B.h) { set_err("orbm_stereo_refine: level sizes differ", cudaSuccess); return ORBX_E_INVALID; }
    if (l == 0) {
      pyr.base_l[0] = hl->last_l0.base; pyr.stride_l[0] = hl->last_l0.frame_stride; pyr.pitch_l[0] = hl->last_l0.pitch;
      pyr.base_r[0] = hr->last_l0.base; pyr.stride_r[0] = hr->last_l0.frame_stride; pyr.pitch_r[0] = hr->last_l0.pitch;
    } else {
      pyr.base_l[l] = hl->arena + A.offset; pyr.stride_l[l] = hl->plan.arena_bytes; pyr.pitch_l[l] = A.pitch;
      pyr.base_r[l] = hr->arena + B.offset; pyr.stride_r[l] = hr->plan.arena_bytes; pyr.pitch_r[l] = B.pitch;
    }
    pyr.w[l] = A.w; pyr.h[l] = A.h;
    pyr.scale[l] = hl->plan.scale[l]; pyr.inv_scale[l] = hl->plan.inv_scale[l];  // Frame copies the left extractor's tables (Frame.cc:110-116)
  }
  CU(cudaSetDevice(hl->device));
  CU(orbx_launch_stereo_refine((const OrbxKeyPointDev*)d_kps_l, d_n_l, (const OrbxKeyPointDev*)d_kps_r, kp_stride,
                               batch, d_best_idx_r, d_best_dist, pyr, mbf, max_d, d_u_right, d_depth, d_sad, d_kept,
                               (cudaStream_t)stream));
  return ORBX_OK;
}

// Frame::ComputeStereoFromRGBD (Frame.cc:984-1005)
int orbm_rgbd_depth(const orbx_keypoint* d_kps, const orbx_keypoint* d_kps_un, const int32_t* d_n, int kp_stride,
                    int batch, const float* d_depth_img, size_t pitch_bytes, size_t frame_stride_bytes, int width,
                    int height, float mbf, float* d_u_right, float* d_depth, void* stream) {
  if (!d_kps || !d_n || !d_depth_img || !d_u_right || !d_depth || batch < 1 || kp_stride < 1 || width < 1 ||
      height < 1 || (pitch_bytes & 3) || (frame_stride_bytes & 3) || pitch_bytes < (size_t)width * 4)
    return ORBX_E_INVALID;
  if (!d_kps_un) d_kps_un = d_kps;  // no distortion: mvKeysUn = mvKeys (Frame.cc:747-752)
  CU(orbx_launch_rgbd_depth((const OrbxKeyPointDev*)d_kps, (const OrbxKeyPointDev*)d_kps_un, d_n, kp_stride, batch,
                            d_depth_img, (long long)(pitch_bytes / 4), (long long)(frame_stride_bytes / 4), width,
                            height, mbf, d_u_right, d_depth, (cudaStream_t)stream));
  return ORBX_OK;
}

int orbx_stage_times(orbx_handle h, float* ms) {
  if (!h || !ms || !h->timed) return ORBX_E_INVALID;
  CU(cudaSetDevice(h->device));
  CU(cudaEventSynchronize(h->ev[ORBX_NUM_STAGES]));
  for (int i = 0; i < ORBX_NUM_STAGES; i++) CU(cudaEventElapsedTime(&ms[i], h->ev[i], h->ev[i + 1]));
  return ORBX_OK;
}

/* Test hook: per-chunk checksums of the DEVICE orbx_sincosf over `count` consecutive float bit patterns from `lo`
 * (chunk patterns per output word); tests compare them with the same checksum of glibc's sincosf. */
int orbx_debug_sincos_checksum(int device, uint32_t lo, uint64_t count, uint32_t chunk, uint64_t* out, int n_out) {
  if (!out || chunk == 0 || count == 0) return ORBX_E_INVALID;
  const int nchunks = (int)((count + chunk - 1) / chunk);
  if (n_out < nchunks) return ORBX_E_CAPACITY;
  CU(cudaSetDevice(device));
  unsigned long long* d = nullptr;
  CU(cudaMalloc((void**)&d, (size_t)nchunks * 8));
  cudaError_t e = cudaMemset(d, 0, (size_t)nchunks * 8);
  if (e == cudaSuccess) {
    k_sincos_checksum<<<nchunks, 256>>>(lo, count, chunk, d);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpy(out, d, (size_t)nchunks * 8, cudaMemcpyDeviceToHost);
  cudaFree(d);
  if (e != cudaSuccess) { set_err("orbx_debug_sincos_checksum", e); return ORBX_E_CUDA; }
  return ORBX_OK;
}

/* Test hook: the blurred level the descriptor stage sampled (level-blur mode only), border-free. */
int orbx_debug_blur_level(orbx_handle h, int frame, int level, uint8_t* dst, size_t dst_stride) {
  if (!h || !dst || level < 0 || level >= h->plan.nlevels || frame < 0 || frame >= h->cfg.max_batch) return ORBX_E_INVALID;
  if (!h->level_blur || !h->blur) { set_err("handle is not in level-blur mode", cudaSuccess); return ORBX_E_INVALID; }
  const OrbxLevel& L = h->plan.lv[level];
  if (dst_stride < (size_t)L.w) return ORBX_E_INVALID;
  CU(cudaSetDevice(h->device));
  CU(cudaDeviceSynchronize());
  CU(cudaMemcpy2D(dst, dst_stride, h->blur + (long long)frame * h->plan.arena_bytes + L.offset, L.pitch, L.w, L.h,
                  cudaMemcpyDeviceToHost));
  return ORBX_OK;
}

int orbx_debug_level_keypoints(orbx_handle h, int frame, int level, int stage, int32_t* xyr, int cap, int* n) {
  if (!h || !n || level < 0 || level >= h->plan.nlevels || frame < 0 || frame >= h->cfg.max_batch)
    return ORBX_E_INVALID;
  const OrbxPlan& P = h->plan;
  const OrbxLevel& L = P.lv[level];
  CU(cudaSetDevice(h->device));
  CU(cudaDeviceSynchronize());
  std::vector<uint32_t> keys;
  if (stage == 0) {
    const int ncell = L.nCols * L.nRows;
    std::vector<int32_t> cnt(ncell);
    CU(cudaMemcpy(cnt.data(), h->cell_cnt + (long long)frame * P.total_cells + L.cell_base, ncell * sizeof(int32_t),
                  cudaMemcpyDeviceToHost));
    std::vector<uint32_t> cellbuf(L.cell_cap);
    for (int c = 0; c < ncell; c++) {
      if (cnt[c] <= 0) continue;
      CU(cudaMemcpy(cellbuf.data(), h->cand + (long long)frame * P.cand_total + L.cand_base + (long long)c * L.cell_cap,
                    (size_t)cnt[c] * sizeof(uint32_t), cudaMemcpyDeviceToHost));
      keys.insert(keys.end(), cellbuf.begin(), cellbuf.begin() + cnt[c]);
    }
  } else {
    int32_t m = 0;
    CU(cudaMemcpy(&m, h->sel_cnt + frame * P.nlevels + level, sizeof(int32_t), cudaMemcpyDeviceToHost));
    keys.resize(m);
    if (m) CU(cudaMemcpy(keys.data(), h->sel + (long long)frame * P.kp_total_cap + L.kp_base, (size_t)m * sizeof(uint32_t),
                         cudaMemcpyDeviceToHost));
  }
  *n = (int)keys.size();
  for (int i = 0; i < (int)keys.size() && i < cap; i++) {
    xyr[3 * i] = ORBX_KEY_X(keys[i]); xyr[3 * i + 1] = ORBX_KEY_Y(keys[i]); xyr[3 * i + 2] = ORBX_KEY_S(keys[i]);
  }
  return ORBX_OK;
}

/* Test hook: runs DistributeOctTree of `level` on caller-supplied keys (x, y relative to the
 * (16,16) border like the reference's vToDistributeKeys, response) and returns the selected keys in
 * list order.  The keys are laid into the level's cell slots in order, so the kernel gathers them
 * back in the given order. */
int orbx_debug_octree(orbx_handle h, int level, const int32_t* xys, int n, int32_t* out_xys, int cap, int* n_out) {
  if (!h || !n_out || level < 0 || level >= h->plan.nlevels || n < 0 || (n > 0 && !xys)) return ORBX_E_INVALID;
  const OrbxPlan& P = h->plan;
  const OrbxLevel& L = P.lv[level];
  const int ncell = L.nCols * L.nRows;
  if ((long long)n > (long long)ncell * L.cell_cap) return ORBX_E_CAPACITY;
  CU(cudaSetDevice(h->device));
  CU(cudaDeviceSynchronize());
  std::vector<int32_t> cnt(P.total_cells, 0);
  std::vector<uint32_t> slots((size_t)L.cand_cap, 0);
  for (int i = 0; i < n; i++) {
    const int c = i / L.cell_cap, j = i % L.cell_cap;
    if (xys[3 * i] < 0 || xys[3 * i] > 4095 || xys[3 * i + 1] < 0 || xys[3 * i + 1] > 4095) return ORBX_E_INVALID;
    slots[(size_t)c * L.cell_cap + j] = orbx_pack_key(xys[3 * i], xys[3 * i + 1], xys[3 * i + 2] & 0xff);
    cnt[L.cell_base + c]++;
  }
  CU(cudaMemcpy(h->cell_cnt, cnt.data(), cnt.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
  if (n) CU(cudaMemcpy(h->cand + L.cand_base, slots.data(), slots.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
  CU(orbx_launch_octree(h->dplan, h->cell_cnt, h->cand, h->gkeys, h->glabels, h->sel, h->sel_cnt, h->status, 1,
                        h->oct_smem, h->oct_smem_keys, h->oct_node_cap, h->oct_max_cells, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  int32_t m = 0, st = 0;
  CU(cudaMemcpy(&m, h->sel_cnt + level, sizeof(int32_t), cudaMemcpyDeviceToHost));
  CU(cudaMemcpy(&st, h->status, sizeof(int32_t), cudaMemcpyDeviceToHost));
  if (st) { CU(cudaMemset(h->status, 0, sizeof(int32_t))); set_err("octree node pool overflow", cudaSuccess); return ORBX_E_CUDA; }
  *n_out = m;
  std::vector<uint32_t> keys(m > 0 ? m : 1);
  if (m) CU(cudaMemcpy(keys.data(), h->sel + L.kp_base, (size_t)m * sizeof(uint32_t), cudaMemcpyDeviceToHost));
  for (int i = 0; i < m && i < cap; i++) {
    out_xys[3 * i] = ORBX_KEY_X(keys[i]) - ORBX_MIN_BORDER;
    out_xys[3 * i + 1] = ORBX_KEY_Y(keys[i]) - ORBX_MIN_BORDER;
    out_xys[3 * i + 2] = ORBX_KEY_S(keys[i]);
  }
  return ORBX_OK;
}

}  // extern "C"
