// orbx_kernels.cuh — device-side plan mirror and launch declarations shared by the .cu files.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "orbx_plan.h"

// Per-level constants as the kernels need them (subset of OrbxLevel, passed by value as a
// __grid_constant__ kernel parameter: 12 levels * 80 B < 1 KB).
struct DevLevel {
  int w, h, pitch;
  int nCols, nRows, wCell, hCell;
  int cell_base, cell_cap;
  int nfeat, kp_base, kp_cap, kp_max;
  int nIni;
  float hX, scale, patch_size;
  int xtab, ytab, fast_resize, tile_resize, tile_rows;
  int btile_base, btiles_x;
  int strip_k;
  long long offset;     // bytes, inside a frame's pyramid arena
  long long cand_base;  // uint32 units, inside a frame's candidate arena
};

struct DevPlan {
  DevLevel lv[ORBX_PLAN_MAX_LEVELS];
  int nlevels, total_cells, kp_total_cap, ini_th, min_th, total_btiles, total_strips, strips_a, hmax_a, hmax_b;
  long long cand_total;   // uint32 per frame
  long long arena_bytes;  // per frame
  int umax[16];
};

// Where each frame's level-0 image lives (user buffer for the batched device API, or the
// arena's own level-0 slot for the host API).
struct Level0 {
  const uint8_t* base;
  long long frame_stride;  // bytes
  int pitch;
};

// Packed candidate / keypoint key: x<<20 | y<<8 | score  (x,y < 4096, score < 256)
__host__ __device__ __forceinline__ uint32_t orbx_pack_key(int x, int y, int s) {
  return ((uint32_t)x << 20) | ((uint32_t)y << 8) | (uint32_t)s;
}
#define ORBX_KEY_X(k) ((int)((k) >> 20))
#define ORBX_KEY_Y(k) ((int)(((k) >> 8) & 0xfff))
#define ORBX_KEY_S(k) ((int)((k) & 0xff))

struct OrbxKeyPointDev {  // = orbx_keypoint = cv::KeyPoint
  float x, y, size, angle, response;
  int32_t octave, class_id;
};

// Both eyes' pyramids as the stereo refinement reads them (levels of frame f of each handle's last batch).
struct OrbxStereoPyr {
  const uint8_t* base_l[ORBX_PLAN_MAX_LEVELS];
  const uint8_t* base_r[ORBX_PLAN_MAX_LEVELS];
  long long stride_l[ORBX_PLAN_MAX_LEVELS], stride_r[ORBX_PLAN_MAX_LEVELS];  // bytes between frames
  int pitch_l[ORBX_PLAN_MAX_LEVELS], pitch_r[ORBX_PLAN_MAX_LEVELS];
  int w[ORBX_PLAN_MAX_LEVELS], h[ORBX_PLAN_MAX_LEVELS];
  float scale[ORBX_PLAN_MAX_LEVELS], inv_scale[ORBX_PLAN_MAX_LEVELS];
};

// --- launches (each returns cudaGetLastError()) -------------------------------------------------
cudaError_t orbx_launch_repack_level0(const DevPlan& P, Level0 l0, uint8_t* arena, int batch, cudaStream_t st);
// Tensor maps of the SOURCE levels of the tiled resize (m[l] = level l over (columns, rows, frames), box =
// 208 columns x lv[l+1].tile_rows rows x 1 frame); mask bit l = level l has a map.  Encoded by fast.cu's helper.
struct alignas(64) OrbxTmap { unsigned long long v[16]; };
struct alignas(64) OrbxResizeMaps {
  OrbxTmap m[ORBX_PLAN_MAX_LEVELS];
  unsigned mask;
};
bool orbx_encode_tmap3(void* m, const uint8_t* base, int w, int h, long long nframes, long long pitch,
                       long long frame_stride, int box_cols, int box_rows);
void orbx_resize_maps_init(const DevPlan& P, const uint8_t* arena, int max_batch, OrbxResizeMaps* M);
cudaError_t orbx_launch_pyramid(const DevPlan& P, const OrbxResizeMaps& maps, Level0 l0, uint8_t* arena,
                                const OrbxTap* xtaps, const OrbxTap* ytaps, int batch, int f0, cudaStream_t st);
// Tensor maps of the pyramid levels for k_fast_strip's TMA staging: one CUtensorMap (128 bytes, 64-byte aligned)
// per level over (columns, rows, frames); mask bit l = level l has a map.  Kept opaque here so that only fast.cu
// needs <cuda.h>.
struct alignas(64) OrbxFastMaps {
  unsigned long long m[ORBX_PLAN_MAX_LEVELS][16];
  unsigned mask;
};
void orbx_fast_maps_init(const DevPlan& P, const uint8_t* arena, int max_batch, OrbxFastMaps* M);
cudaError_t orbx_launch_fast(const DevPlan& P, const OrbxFastMaps& maps, Level0 l0, const uint8_t* arena,
                             const uint32_t* cell_tab, const uint32_t* strip_tab, int32_t* cell_cnt, uint32_t* cand,
                             int batch, int f0, cudaStream_t st);
cudaError_t orbx_launch_fast_cell(const DevPlan& P, Level0 l0, const uint8_t* arena, const uint32_t* cell_tab,
                                  int32_t* cell_cnt, uint32_t* cand, int batch, cudaStream_t st);
cudaError_t orbx_launch_fast_strip(const DevPlan& P, const OrbxFastMaps& maps, Level0 l0, const uint8_t* arena,
                                   const uint32_t* strip_tab, int32_t* cell_cnt, uint32_t* cand, int batch, int f0,
                                   cudaStream_t st);
cudaError_t orbx_launch_octree(const DevPlan& P, const int32_t* cell_cnt, const uint32_t* cand,
                               uint32_t* gkeys, int16_t* glabels, uint32_t* sel, int32_t* sel_cnt,
                               int32_t* status, int batch, size_t smem_bytes, int smem_keys, int node_cap,
                               int max_cells, cudaStream_t st);
cudaError_t orbx_launch_order(const DevPlan& P, const uint32_t* sel, const int32_t* sel_cnt, int lap0,
                              int lap1, int32_t* dst, int32_t* n_out, int32_t* mono_index, int batch,
                              cudaStream_t st);
cudaError_t orbx_launch_describe(const DevPlan& P, Level0 l0, const uint8_t* arena, const uint32_t* sel,
                                 const int32_t* sel_cnt, const int32_t* dst, OrbxKeyPointDev* kps,
                                 uint8_t* desc, int capacity, int strict_fp, int batch, cudaStream_t st);
cudaError_t orbx_launch_blur(const DevPlan& P, Level0 l0, const uint8_t* arena, uint8_t* blur, int batch,
                             cudaStream_t st);
cudaError_t orbx_launch_describe_lite(const DevPlan& P, Level0 l0, const uint8_t* arena, const uint8_t* blur,
                                      const uint32_t* sel, const int32_t* sel_cnt, const int32_t* dst,
                                      OrbxKeyPointDev* kps, uint8_t* desc, int capacity, int strict_fp, int batch,
                                      cudaStream_t st);
cudaError_t orbx_launch_stereo_refine(const OrbxKeyPointDev* kl, const int32_t* nl, const OrbxKeyPointDev* kr,
                                      int kp_stride, int batch, const int32_t* best_idx, const int32_t* best_dist,
                                      const OrbxStereoPyr& pyr, float mbf, float max_d, float* u_right, float* depth,
                                      int32_t* sad, int32_t* kept, cudaStream_t st);
cudaError_t orbx_launch_rgbd_depth(const OrbxKeyPointDev* kps, const OrbxKeyPointDev* kps_un, const int32_t* n,
                                   int kp_stride, int batch, const float* depth_img, long long pitch_f,
                                   long long frame_stride_f, int w, int h, float mbf, float* u_right, float* depth,
                                   cudaStream_t st);
cudaError_t orbx_describe_init_tables(const int* umax);
cudaError_t orbx_launch_border(const uint8_t* src, int pitch, int w, int h, uint8_t* dst, int dst_pitch,
                               cudaStream_t st);
size_t orbx_octree_smem_bytes(int smem_keys, int node_cap, int max_cells);

// cudaFuncAttributeMaxDynamicSharedMemorySize belongs to (function, device), not to a handle: this keeps a
// per-device, monotone (only ever raised) cache per kernel slot, under a mutex, so handles of different sizes,
// on different devices, created from different threads never lower or miss each other's opt-in.
enum OrbxSmemSlot { ORBX_SMEM_FAST = 0, ORBX_SMEM_FAST4, ORBX_SMEM_FAST_STRIP, ORBX_SMEM_OCTREE, ORBX_SMEM_OCTREE_BIG, ORBX_SMEM_OCTREE_X, ORBX_SMEM_BOW_VECTORS, ORBX_SMEM_ROWBAND, ORBX_SMEM_SLOTS };
cudaError_t orbx_ensure_dyn_smem(OrbxSmemSlot slot, const void* func, size_t bytes);
// Sets the calling thread's orbx_last_error() text (for entry points outside api.cu).
void orbx_set_error_text(const char* msg);
