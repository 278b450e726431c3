/*
 * orbx.h — C ABI of the B200-native ORB front-end (liborbx.so).
 *
 * This is the drop-in boundary for ORB-SLAM3's per-frame hot path
 * (vdoom/ORB_SLAM3_VIO_FIXES).  The reference has no FFI layer: the boundary there is two C++
 * classes.  Each entry point below names the reference interface it replaces (file:line in
 * the reference tree); include/orbx_compat/ORBextractor.h and ORBmatcher_core.h put the
 * reference's own C++ signatures back on top of these calls.
 *
 * Conventions: plain pointers and sizes only; every function returns 0 (ORBX_OK) or a
 * negative ORBX_E_* code; nothing throws, nothing aborts.  "d_" pointers are CUDA device
 * pointers, `stream` is a cudaStream_t passed as void* and used as given (NULL = CUDA's
 * default stream); the host-buffer entry points run on a private stream of the handle.
 * A handle is single-caller (like an ORBextractor instance, ORBextractor.h:83 mutable
 * mvImagePyramid); distinct handles may be used concurrently from different threads, as
 * Frame.cc:122-125 does for the left/right extractors.
 */
#ifndef ORBX_H_
#define ORBX_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORBX_OK 0
#define ORBX_E_INVALID (-1)     /* bad argument / unsupported geometry                     */
#define ORBX_E_CUDA (-2)        /* CUDA runtime error (sticky: destroy and recreate)        */
#define ORBX_E_CAPACITY (-3)    /* caller's output buffer too small                         */
#define ORBX_E_NOMEM (-4)       /* device allocation failed                                 */
#define ORBX_E_EMPTY (-5)       /* empty image: operator() returns -1 (ORBextractor.cc:1090) */
#define ORBX_E_NODEVICE (-6)    /* no CUDA device / driver: there is no CPU fallback        */
#define ORBX_E_INTERNAL (-7)    /* an internal capacity bound was violated on the device (octree
                                   node pool): the call's output is incomplete, never silently    */

#define ORBX_MAX_LEVELS 12

/* flags */
#define ORBX_FLAG_STRICT_FP 1u  /* evaluate GET_VALUE (ORBextractor.cc:117-119) uncontracted
                                   (mul,mul,add) instead of the FMA form g++ -O3 -march=native
                                   emits for the reference; default = as compiled (FMA).     */

#define ORBX_FLAG_STAGE_TIMING 2u /* record CUDA events between the stages of every call so that
                                   orbx_stage_times() can report per-kernel device time        */
#define ORBX_FLAG_FUSED_BLUR 4u   /* descriptor stage: blur each keypoint's neighbourhood in shared memory
                                   * (no blurred level is written); default is chosen per plan */
#define ORBX_FLAG_LEVEL_BLUR 8u   /* descriptor stage: blur whole levels once, then sample them */
#define ORBX_NUM_STAGES 5       /* pyramid, fast, octree, order, describe                     */

/* cv::KeyPoint memory layout, 28 bytes (what std::vector<cv::KeyPoint> holds). */
typedef struct orbx_keypoint {
  float x, y;      /* pt, level-0 pixel coordinates                                  */
  float size;      /* (int)(31*scale[octave]), ORBextractor.cc:880                   */
  float angle;     /* degrees in [0,360), IC_Angle + cv::fastAtan2, :76-103          */
  float response;  /* FAST score                                                     */
  int32_t octave;  /* pyramid level                                                  */
  int32_t class_id;/* -1                                                             */
} orbx_keypoint;

/* Constructor arguments of ORBextractor (ORBextractor.h:49-50) plus the plan geometry. */
typedef struct orbx_config {
  int32_t nfeatures;     /* ORBextractor.nFeatures                                       */
  float scale_factor;    /* ORBextractor.scaleFactor                                     */
  int32_t nlevels;       /* ORBextractor.nLevels (<= ORBX_MAX_LEVELS)                    */
  int32_t ini_th_fast;   /* ORBextractor.iniThFAST                                       */
  int32_t min_th_fast;   /* ORBextractor.minThFAST                                       */
  int32_t width, height; /* frame size the handle is planned for                         */
  int32_t max_batch;     /* frames per orbx_extract_batch_* call (>=1)                   */
  int32_t device;        /* CUDA device ordinal                                          */
  uint32_t flags;        /* ORBX_FLAG_*                                                  */
} orbx_config;

typedef struct orbx_extractor_s* orbx_handle;

/* Replaces ORBextractor::ORBextractor (ORBextractor.cc:409-469): builds the scale tables,
 * per-level feature quotas, umax, resize coefficient tables and FAST cell grid for one frame
 * size, and allocates every device buffer once (no allocation on the per-frame path). */
int orbx_create(const orbx_config* cfg, orbx_handle* out);
void orbx_destroy(orbx_handle h);

/* Upper bound of keypoints one frame can return: nfeatures + 3*nlevels. */
int orbx_max_keypoints(orbx_handle h);

/* Replaces ORBextractor::operator() (ORBextractor.h:57-59, ORBextractor.cc:1086-1168) for one
 * frame in HOST memory: H2D copy, all kernels, D2H copy, synchronous.  `lap0/lap1` is
 * vLappingArea; *mono_index receives operator()'s return value.  `kps`/`desc` (n x 32 bytes,
 * row-major like the CV_8U descriptor Mat) must hold `capacity` entries. */
int orbx_extract(orbx_handle h, const uint8_t* img, int width, int height, size_t stride,
                 int lap0, int lap1, orbx_keypoint* kps, uint8_t* desc, int capacity,
                 int* n_out, int* mono_index);

/* Batched operator() on frames resident in device memory (throughput path).  Frame f starts
 * at d_imgs + f*frame_stride, rows are `pitch` bytes apart.  Outputs are device arrays:
 * d_kps[batch][capacity], d_desc[batch][capacity][32], d_n_out[batch], d_mono_index[batch].
 * Asynchronous on `stream`.  The input must stay alive until the stream has drained (level 0
 * of the pyramid aliases it). */
int orbx_extract_batch_device(orbx_handle h, const uint8_t* d_imgs, size_t pitch,
                              size_t frame_stride, int batch, int lap0, int lap1,
                              orbx_keypoint* d_kps, uint8_t* d_desc, int capacity,
                              int32_t* d_n_out, int32_t* d_mono_index, void* stream);

/* For the asynchronous device path: waits for `stream`, then reads and clears the handle's device-side
 * status word.  ORBX_OK, or ORBX_E_INTERNAL if a kernel of an earlier orbx_extract_batch_device call hit
 * an internal bound (the synchronous entry points check this themselves before returning). */
int orbx_check_status(orbx_handle h, void* stream);

/* Same through HOST buffers (pinned for full speed): copies the batch in, runs, copies the
 * results out, synchronises.  Frames are tightly packed width*height. */
int orbx_extract_batch_host(orbx_handle h, const uint8_t* imgs, int batch, int lap0, int lap1,
                            orbx_keypoint* kps, uint8_t* desc, int capacity, int32_t* n_out,
                            int32_t* mono_index);

/* Device copies of what the last orbx_extract / orbx_extract_batch_host call returned, [batch][capacity]
 * records with capacity = orbx_max_keypoints(): lets the orbm_* entry points score them without a
 * second upload.  Valid until the next call on the handle. */
int orbx_last_outputs_device(orbx_handle h, const orbx_keypoint** d_kps, const uint8_t** d_desc,
                             const int32_t** d_n, const int32_t** d_mono_index, int* capacity);

/* The "GetPyramid" surface: replaces reads of the public field mvImagePyramid
 * (ORBextractor.h:83; Frame.cc:818,908,918,923).  Copies level `level` of frame `frame` of
 * the last call to host memory; with_border=1 reproduces the (w+38)x(h+38) parent buffer of
 * the reference's ROI (19-px BORDER_REFLECT_101 ring, ORBextractor.cc:1185-1191). */
int orbx_get_pyramid(orbx_handle h, int frame, int level, uint8_t* dst, size_t dst_stride,
                     int with_border);

/* Every level of frame `frame` in ONE device->host transfer (what the compat operator() needs to
 * refill mvImagePyramid, ORBextractor.h:83 / ORBextractor.cc:1170-1195): *host_base = a pinned host
 * buffer owned by the handle and valid until the next call on it; level l starts at offsets[l] with
 * row stride strides[l] (= w_l + 38 with the 19-px REFLECT_101 ring when with_border, else w_l).
 * offsets / strides must hold nlevels entries. */
int orbx_get_pyramid_all(orbx_handle h, int frame, int with_border, const uint8_t** host_base,
                         size_t* offsets, size_t* strides);
/* Device view of the same level (border-free). */
int orbx_pyramid_device(orbx_handle h, int frame, int level, const uint8_t** d_ptr,
                        size_t* pitch, int* width, int* height);
int orbx_level_size(orbx_handle h, int level, int* width, int* height);

/* Replaces GetScaleFactors/GetInverseScaleFactors/GetScaleSigmaSquares/
 * GetInverseScaleSigmaSquares (ORBextractor.h:61-81); arrays of nlevels floats.  Any pointer
 * may be NULL.  features_per_level = mnFeaturesPerLevel (ORBextractor.cc:434-445). */
int orbx_get_scale_tables(orbx_handle h, float* scale, float* inv_scale, float* sigma2,
                          float* inv_sigma2, int32_t* features_per_level);

/* Device time of each stage of the last extract call, in milliseconds, measured with CUDA events
 * on the stream the kernels ran on (the reference's REGISTER_TIMES counterpart,
 * include/Settings.h:24, Frame.cc:119-146).  Needs ORBX_FLAG_STAGE_TIMING; waits for the call to
 * finish.  ms[ORBX_NUM_STAGES] = {pyramid (all level launches), fast, octree, order, describe}. */
int orbx_stage_times(orbx_handle h, float* ms);

/* Introspection for the parity tests: order-independent 64-bit checksums (one per `chunk` patterns) of the device
 * build of the glibc-exact sincosf (ORBextractor.cc:112) over `count` consecutive float bit patterns from `lo`. */
int orbx_debug_sincos_checksum(int device, uint32_t lo, uint64_t count, uint32_t chunk, uint64_t* out, int n_out);
/* Introspection for the parity tests: the GaussianBlur'ed level (ORBextractor.cc:1131-1133) the descriptor
 * stage sampled for frame `frame` of the last call (level-blur mode only), border-free, row stride dst_stride. */
int orbx_debug_blur_level(orbx_handle h, int frame, int level, uint8_t* dst, size_t dst_stride);
/* Introspection for the parity tests: per-level intermediate keypoint lists of frame `frame`
 * of the last call.  stage 0 = FAST candidates in DistributeOctTree input order (x,y relative
 * to (16,16), response); stage 1 = after the octree, in list order (level coordinates).
 * Fills up to cap entries of (x, y, response) int32 triples; returns the count in *n. */
int orbx_debug_level_keypoints(orbx_handle h, int frame, int level, int stage, int32_t* xyr,
                               int cap, int* n);
/* Test hook: DistributeOctTree (ORBextractor.cc:555-779) of one level on caller-supplied keys
 * (x, y relative to the 16-px border, response); returns the kept keys in list order. */
int orbx_debug_octree(orbx_handle h, int level, const int32_t* xys, int n, int32_t* out_xys, int cap,
                      int* n_out);


/* ----------------------------------------------------------------------------------------
 * Matcher cores.  ORBmatcher::DescriptorDistance (ORBmatcher.h:43, ORBmatcher.cc:2058-2074)
 * for a single pair stays a host inline (include/orbx_compat/ORBmatcher_core.h); the
 * library exports the batched candidate-scoring loops that sit behind the Search* family.
 * Descriptors are rows of 32 bytes (the reference's N x 32 CV_8U Mat).
 * -------------------------------------------------------------------------------------- */
typedef struct orbm_result {
  int32_t best_idx;     /* reference index of the best candidate, -1 if none            */
  int16_t best_dist;    /* its Hamming distance (256 if none)                           */
  int16_t second_dist;  /* second best distance (256 if none)                           */
  int8_t best_level;    /* octave of best (-1 if none)                                  */
  int8_t second_level;  /* octave of second best (-1 if none)                           */
  int16_t pad;
} orbm_result;

#define ORBM_TIE_FIRST_MIN 0 /* `dist<bestDist` loops: ORBmatcher.cc:84-120, 273-325, 1743-1768 */
#define ORBM_TIE_LAST_MIN 1  /* `dist<=bestDist`: SearchForTriangulation, ORBmatcher.cc:1017     */

/* ORBmatcher::DescriptorDistance (ORBmatcher.cc:2058-2074) for n_pairs explicit pairs in one launch:
 * d_dist[i] = Hamming(d_a[d_idx_a[i]], d_b[d_idx_b[i]]) (index arrays may be NULL = row i).  For callers
 * that hold many pairs at once (e.g. re-scoring accepted matches); a single pair stays the host inline. */
int orbm_descriptor_distance_pairs(const uint8_t* d_a, const uint8_t* d_b, const int32_t* d_idx_a,
                                   const int32_t* d_idx_b, int n_pairs, int32_t* d_dist, void* stream);

/* Scores explicit candidate lists (CSR: d_cand_off[nq+1], d_cand_idx in the order the
 * reference loop visits them).  d_ref_level[nr] = octave of each reference keypoint (may be
 * NULL).  Result per query as the reference's bestDist/bestIdx/bestLevel/bestDist2/
 * bestLevel2 after the loop. */
int orbm_argmin2_csr(const uint8_t* d_q, int nq, const uint8_t* d_r, int nr,
                     const int32_t* d_cand_off, const int32_t* d_cand_idx,
                     const int8_t* d_ref_level, int tie_rule, orbm_result* d_out, void* stream);

/* Frame::AssignFeaturesToGrid (Frame.cc:385-416) on device: bins n keypoints into the 64x48
 * grid over [min_x,max_x) x [min_y,max_y); d_cell_off[64*48+1] / d_cell_idx[n] form a CSR in
 * (column-major cell, insertion) order — the order GetFeaturesInArea walks.  Batched: frame f
 * uses d_kps + f*kp_stride, d_n[f] keypoints. */
int orbm_grid_build(const orbx_keypoint* d_kps, const int32_t* d_n, int kp_stride, int batch,
                    float min_x, float max_x, float min_y, float max_y, int32_t* d_cell_off,
                    int32_t* d_cell_idx, void* stream);

/* Fused Frame::GetFeaturesInArea (Frame.cc:657-723) + best/second scoring loop: for each
 * query (x, y, radius r, octave band [min_level,max_level], descriptor) walks the grid cells
 * in the reference's order, gates candidates and scores them without materialising the list.
 * Queries of frame f: d_q* + f*q_stride, d_nq[f] of them; references: frame f's keypoints. */
int orbm_search_area(const orbx_keypoint* d_kps, const uint8_t* d_desc, const int32_t* d_n,
                     int kp_stride, const int32_t* d_cell_off, const int32_t* d_cell_idx,
                     float min_x, float max_x, float min_y, float max_y, const float* d_qx,
                     const float* d_qy, const float* d_qr, const int8_t* d_qmin_level,
                     const int8_t* d_qmax_level, const uint8_t* d_qdesc, const int32_t* d_nq,
                     int q_stride, int batch, int tie_rule, orbm_result* d_out, void* stream);

/* Candidate scoring of ORBmatcher::Fuse (ORBmatcher.cc:1148-1338 and the Sim3 twin :1340-1455): per
 * projected MapPoint (d_qx,d_qy = uv, d_q_ur = ur, d_qr = th*scale[l], d_qlevel = nPredictedLevel,
 * descriptor) KeyFrame::GetFeaturesInArea (KeyFrame.cc:704-748) on the orbm_grid_build CSR, octave in
 * [l-1, l], chi-square gate e2*invLevelSigma2[octave] <= 5.99 (7.8 with d_u_right[j] >= 0), first
 * minimum.  d_out[q].best_idx/best_dist feed the host's `bestDist<=TH_LOW` replace-or-add bookkeeping,
 * which mutates the map and stays on the host.  Queries are independent. */
int orbm_fuse_candidates(const orbx_keypoint* d_kps, const uint8_t* d_desc, const int32_t* d_n,
                         int kp_stride, const int32_t* d_cell_off, const int32_t* d_cell_idx,
                         float min_x, float max_x, float min_y, float max_y, const float* d_u_right,
                         const float* inv_level_sigma2_host, int nlevels, const float* d_qx,
                         const float* d_qy, const float* d_qr, const float* d_q_ur,
                         const int8_t* d_qlevel, const uint8_t* d_qdesc, const int32_t* d_nq,
                         int q_stride, int batch, orbm_result* d_out, void* stream);

/* Window queries of SearchByProjection(Frame&, const Frame&, th, bMono) (ORBmatcher.cc:1676-1768)
 * for a pure image-plane shift: query i of frame f is keypoint i of frame d_src_frame[f] moved by
 * d_shift_xy[f] = (dx,dy), radius th*scale[octave], octave band [octave+band_lo, octave+band_hi],
 * descriptor copied.  (The pose-based projection itself is host geometry and stays there.) */
int orbm_queries_shift(const orbx_keypoint* d_kps, const uint8_t* d_desc, const int32_t* d_n,
                       int kp_stride, int batch, const int32_t* d_src_frame,
                       const float* d_shift_xy, float th, const float* scale_factors_host,
                       int nlevels, int band_lo, int band_hi, float* d_qx, float* d_qy, float* d_qr,
                       int8_t* d_qmin_level, int8_t* d_qmax_level, uint8_t* d_qdesc, int32_t* d_nq,
                       void* stream);

/* ORBmatcher::SearchByProjection as a whole function, single-camera case (Nleft == -1), batched over
 * `batch` current frames:
 *   ORBM_PROJ_MAPPOINTS  (Frame&, const vector<MapPoint*>&, th, ...)  ORBmatcher.cc:43-142: best and second
 *                        best with their levels, TH_HIGH, ratio test nn_ratio when both are on one level;
 *   ORBM_PROJ_LASTFRAME  (Frame& Current, const Frame& Last, th, bMono) ORBmatcher.cc:1676-1887: best only,
 *                        TH_HIGH, rotation histogram + ComputeThreeMaxima when check_orientation.
 * The pose / camera projection stays on the host and arrives per query: (d_qx,d_qy) projected pixel,
 * d_qr search radius, [d_qmin_level, d_qmax_level] the octave band passed to GetFeaturesInArea (max < 0 =
 * open), d_q_ur projected right coordinate (NULL with d_u_right NULL: no stereo check), d_q_flags bit 0 =
 * query takes part (mbTrackInView / a non-outlier MapPoint), bit 1 = its MapPoint has Observations()>0, so
 * the feature it is written to is skipped by all later queries (NULL = 3 for all), d_q_angle = kpLF.angle.
 * d_feat_locked[j] != 0: F.mvpMapPoints[j] holds a point with Observations()>0 on entry (NULL = none).
 * The grid is orbm_grid_build's.  The sequential "already matched" dependence of the reference loop is
 * reproduced exactly (queries are committed in index order).  Outputs: d_match[i] = feature that query i
 * was written to (-1 none / erased by the rotation check), d_feat_query[j] = query whose MapPoint
 * F.mvpMapPoints[j] holds on return (-1 = untouched or erased), d_nmatches[f] = the return value.
 * d_scratch: batch * q_stride * 32 bytes. */
#define ORBM_PROJ_MAPPOINTS 0
#define ORBM_PROJ_LASTFRAME 1
int orbm_search_by_projection(const orbx_keypoint* d_kps, const uint8_t* d_desc, const int32_t* d_n,
                              int kp_stride, int batch, const int32_t* d_cell_off,
                              const int32_t* d_cell_idx, float min_x, float max_x, float min_y,
                              float max_y, const float* d_u_right, const uint8_t* d_feat_locked,
                              const float* d_qx, const float* d_qy, const float* d_qr,
                              const int8_t* d_qmin_level, const int8_t* d_qmax_level,
                              const uint8_t* d_qdesc, const int32_t* d_nq, int q_stride,
                              const float* d_q_ur, const uint8_t* d_q_flags, const float* d_q_angle,
                              int mode, float nn_ratio, int check_orientation, void* d_scratch,
                              int32_t* d_match, int32_t* d_feat_query, int32_t* d_nmatches,
                              void* stream);

/* The same with the acceptance threshold as a parameter (th_dist replaces TH_HIGH = 100).  With
 * ORBM_PROJ_LASTFRAME, every query flagged 3, d_feat_locked = "CurrentFrame.mvpMapPoints[j] != NULL" and
 * no stereo inputs this is SearchByProjection(Frame&, KeyFrame*, const set<MapPoint*>&, th, ORBdist)
 * (ORBmatcher.cc:1889-2010, relocalisation: octave band [l-1, l+1], best <= ORBdist = 100 or 64); queries
 * the host drops (bad point, in sAlreadyFound, outside the image or the distance range) clear bit 0. */
int orbm_search_by_projection_th(const orbx_keypoint* d_kps, const uint8_t* d_desc, const int32_t* d_n,
                                 int kp_stride, int batch, const int32_t* d_cell_off,
                                 const int32_t* d_cell_idx, float min_x, float max_x, float min_y,
                                 float max_y, const float* d_u_right, const uint8_t* d_feat_locked,
                                 const float* d_qx, const float* d_qy, const float* d_qr,
                                 const int8_t* d_qmin_level, const int8_t* d_qmax_level,
                                 const uint8_t* d_qdesc, const int32_t* d_nq, int q_stride,
                                 const float* d_q_ur, const uint8_t* d_q_flags, const float* d_q_angle,
                                 int mode, float nn_ratio, int check_orientation, int th_dist,
                                 void* d_scratch, int32_t* d_match, int32_t* d_feat_query,
                                 int32_t* d_nmatches, void* stream);

/* orbm_search_by_projection without any allocation and without a caller scratch buffer.  NO COUNTERPART IN THE
 * REFERENCE: added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arguments, outputs and results as
 * orbm_search_by_projection minus d_scratch: the candidate lists live in a static device array of
 * ORBM_STATIC_SCRATCH_BYTES inside the library (every GPU that loads the library holds it until exit), so
 * batch * q_stride * 32 must not exceed it (131 072 queries); larger calls return ORBX_E_CAPACITY with the
 * sizes in orbx_last_error().  The call itself performs no cudaMalloc / cudaFree.  All calls on a GPU share the
 * array (it keeps the last call's contents; nothing in it needs clearing, every entry read is written first by
 * the same call): calls that overlap on one GPU — different streams or host threads — RACE; order them on one
 * stream.  Calls on different GPUs are independent.  The device must be current, as for every orbm_* call. */
#define ORBM_STATIC_SCRATCH_BYTES (4u << 20)
int orbm_search_by_projection_static(const orbx_keypoint* d_kps, const uint8_t* d_desc, const int32_t* d_n,
                                     int kp_stride, int batch, const int32_t* d_cell_off,
                                     const int32_t* d_cell_idx, float min_x, float max_x, float min_y,
                                     float max_y, const float* d_u_right, const uint8_t* d_feat_locked,
                                     const float* d_qx, const float* d_qy, const float* d_qr,
                                     const int8_t* d_qmin_level, const int8_t* d_qmax_level,
                                     const uint8_t* d_qdesc, const int32_t* d_nq, int q_stride,
                                     const float* d_q_ur, const uint8_t* d_q_flags, const float* d_q_angle,
                                     int mode, float nn_ratio, int check_orientation, int32_t* d_match,
                                     int32_t* d_feat_query, int32_t* d_nmatches, void* stream);

/* Hamming stage of Frame::ComputeStereoMatches (Frame.cc:811-894): per left keypoint the best
 * right keypoint in its row band (|dy| within 2*scale[octR], |octave diff|<=1,
 * uL-max_d <= uR <= uL), first minimum below TH_HIGH=100 wins, candidates visited in
 * ascending right index (the order vRowIndices is filled).  Batched over `batch` pairs. */
int orbm_stereo_rowband(const orbx_keypoint* d_kps_l, const uint8_t* d_desc_l,
                        const int32_t* d_n_l, const orbx_keypoint* d_kps_r,
                        const uint8_t* d_desc_r, const int32_t* d_n_r, int kp_stride, int batch,
                        const float* scale_factors_host, int nlevels, float max_d,
                        int32_t* d_best_idx_r, int32_t* d_best_dist, void* stream);

/* Sub-pixel stage and median cut of Frame::ComputeStereoMatches (Frame.cc:896-981), after
 * orbm_stereo_rowband: for every left keypoint whose best Hamming distance is < (TH_HIGH+TH_LOW)/2
 * the 11x11 L1-SAD slide over +-5 px on its pyramid level, the parabola fit, the disparity gate
 * 0 <= d < max_d (= mbf/mb) and depth = mbf/d; then matches whose SAD >= 1.5f*1.4f*median are
 * erased.  The pyramids are read on the device from the two extractor handles (frame f of the last
 * orbx_extract_batch_device call of each).  d_u_right / d_depth = mvuRight / mvDepth (-1 = none),
 * d_sad = best SAD per kept-or-cut match (-1 where no match was formed), d_kept[f] = matches left. */
int orbm_stereo_refine(orbx_handle left, orbx_handle right, const orbx_keypoint* d_kps_l,
                       const int32_t* d_n_l, const orbx_keypoint* d_kps_r, int kp_stride, int batch,
                       const int32_t* d_best_idx_r, const int32_t* d_best_dist, float mbf,
                       float max_d, float* d_u_right, float* d_depth, int32_t* d_sad,
                       int32_t* d_kept, void* stream);

/* Frame::ComputeStereoFromRGBD (Frame.cc:984-1005): d = imDepth.at<float>(v,u) at the (truncated)
 * keypoint position; d > 0 -> mvDepth = d, mvuRight = kpUn.x - mbf/d, else -1.  d_kps_un = the
 * undistorted keypoints (NULL = same as d_kps).  The depth image is float32, pitched. */
int orbm_rgbd_depth(const orbx_keypoint* d_kps, const orbx_keypoint* d_kps_un, const int32_t* d_n,
                    int kp_stride, int batch, const float* d_depth_img, size_t pitch_bytes,
                    size_t frame_stride_bytes, int width, int height, float mbf, float* d_u_right,
                    float* d_depth, void* stream);

/* cv::BFMatcher(NORM_HAMMING).knnMatch(k=2) (Frame.cc:1144): all pairs, two best per query,
 * ties to the lower reference index.  d_idx2/d_dist2 are [nq][2]. */
int orbm_knn2_dense(const uint8_t* d_q, int nq, const uint8_t* d_r, int nr, int32_t* d_idx2,
                    int32_t* d_dist2, void* stream);

/* --------------------------------------------------------------------------------------
 * Vocabulary-node matching (§8(f).2-4).
 * A DBoW2::FeatureVector (std::map<NodeId, std::vector<unsigned>>, Thirdparty/DBoW2/DBoW2/
 * FeatureVector.h) is a CSR per frame: d_node[f][node_stride] ascending node ids,
 * d_off[f][node_stride+1], d_idx[f][kp_stride] feature indices (ascending inside a node),
 * d_nn[f] nodes in use.
 * -------------------------------------------------------------------------------------- */
typedef struct orbm_feature_vector {
  int32_t* d_node;
  int32_t* d_off;
  int32_t* d_idx;
  int32_t* d_nn;
} orbm_feature_vector;

/* ORBmatcher::SearchByBoW, whole function, single-camera case, batched over `batch` pairs:
 *   ORBM_BOW_KF_FRAME  (KeyFrame*, Frame&, vector<MapPoint*>&)      ORBmatcher.cc:223-425   best <= TH_LOW
 *   ORBM_BOW_KF_KF     (KeyFrame*, KeyFrame*, vector<MapPoint*>&)   ORBmatcher.cc:765-905   best <  TH_LOW
 * side 1 = the (first) keyframe, side 2 = the frame / second keyframe.  d_valid1[i] != 0: feature i
 * of side 1 holds a MapPoint that is not bad (NULL = all); d_valid2 likewise (KF_KF only).
 * d_match12[i] = side-2 feature matched to side-1 feature i (vpMatches12 of KF_KF), d_match21 the
 * inverse (vpMapPointMatches of KF_FRAME); -1 = none or erased by the rotation check.
 * The reference's "already matched" skip is reproduced exactly. */
#define ORBM_BOW_KF_FRAME 0
#define ORBM_BOW_KF_KF 1
int orbm_search_by_bow(int mode, const orbx_keypoint* d_kps1, const uint8_t* d_desc1, const int32_t* d_n1,
                       const uint8_t* d_valid1, const orbm_feature_vector* fv1,
                       const orbx_keypoint* d_kps2, const uint8_t* d_desc2, const int32_t* d_n2,
                       const uint8_t* d_valid2, const orbm_feature_vector* fv2, int kp_stride,
                       int node_stride, int batch, float nn_ratio, int check_orientation,
                       int32_t* d_match12, int32_t* d_match21, int32_t* d_nmatches, void* stream);

/* ORBmatcher::SearchForTriangulation (ORBmatcher.cc:907-1146), pinhole single-camera case, with
 * Pinhole::epipolarConstrain (CameraModels/Pinhole.cpp:107-129), batched over keyframe pairs.
 * Host geometry per pair: d_F12[p][9] = K1^-T [t12]x R12 K2^-1 (row-major), d_ep[p][2] = epipole.
 * d_has_mp*[i] != 0: GetMapPoint(i) != NULL (skipped); d_u_right* = mvuRight (NULL = monocular).
 * scale_factors2 / level_sigma2_2 = pKF2->mvScaleFactors / mvLevelSigma2 (host arrays).
 * d_match12[i] = vMatches12[i] after the rotation check; d_nmatches[p] = the return value. */
int orbm_search_for_triangulation(
    const orbx_keypoint* d_kps1, const uint8_t* d_desc1, const int32_t* d_n1, const uint8_t* d_has_mp1,
    const float* d_u_right1, const orbm_feature_vector* fv1, const orbx_keypoint* d_kps2,
    const uint8_t* d_desc2, const int32_t* d_n2, const uint8_t* d_has_mp2, const float* d_u_right2,
    const orbm_feature_vector* fv2, int kp_stride, int node_stride, int batch, const float* d_F12,
    const float* d_ep, const float* scale_factors2_host, const float* level_sigma2_2_host, int nlevels,
    int only_stereo, int coarse, int check_orientation, int32_t* d_match12, int32_t* d_nmatches,
    void* stream);

/* DBoW2 vocabulary tree, flat, device-resident (TemplatedVocabulary.h m_nodes): node i has children
 * d_child_idx[d_child_off[i] .. d_child_off[i+1]) (none = leaf), descriptor d_node_desc[i*32],
 * word id d_node_word[i] and weight d_node_weight[i] (leaves).  Node 0 is the root. */
typedef struct orbm_vocabulary {
  const int32_t* d_child_off;
  const int32_t* d_child_idx;
  const uint8_t* d_node_desc;
  const int32_t* d_node_word;
  const double* d_node_weight;
  int n_nodes, L;
} orbm_vocabulary;

/* The same tree in host memory, as TemplatedVocabulary::loadFromTextFile builds it
 * (TemplatedVocabulary.h:1338-1423): node ids in file order, children in push_back order, word ids in
 * file order of the leaves (-1 for inner nodes); k, L, scoring (ScoringType), weighting (WeightingType)
 * from the first line. */
typedef struct orbm_vocabulary_host {
  int32_t k, L, scoring, weighting, n_nodes, n_words;
  int32_t* child_off;   /* [n_nodes+1] */
  int32_t* child_idx;   /* [n_nodes-1] */
  uint8_t* node_desc;   /* [n_nodes][32] */
  int32_t* node_word;   /* [n_nodes] */
  double* node_weight;  /* [n_nodes] */
  int32_t* node_parent; /* [n_nodes] */
} orbm_vocabulary_host;

/* ORBVocabulary::loadFromTextFile (System.cc:118 loads ORBvoc.txt with it): parses the reference's text
 * format ("k L scoring weighting", then "parent isLeaf d0..d31 weight" per node) on the host; needs no
 * GPU.  ORBX_E_INVALID for an unreadable / malformed file (the reference returns false).  Blank lines
 * are skipped (the reference turns the trailing newline into a node of uninitialised memory). */
int orbm_vocabulary_parse_text(const char* path, orbm_vocabulary_host* out);
/* TemplatedVocabulary::saveToTextFile (TemplatedVocabulary.h:1427-1464); weights written losslessly. */
int orbm_vocabulary_save_text(const char* path, const orbm_vocabulary_host* voc);
void orbm_vocabulary_host_free(orbm_vocabulary_host* voc);
/* Host tree -> device arrays for orbm_bow_transform; orbm_vocabulary_release frees them. */
int orbm_vocabulary_upload(const orbm_vocabulary_host* voc, orbm_vocabulary* out);
void orbm_vocabulary_release(orbm_vocabulary* voc);
/* parse + upload in one call; k / n_words may be NULL. */
int orbm_vocabulary_load_text(const char* path, orbm_vocabulary* out, int32_t* k, int32_t* n_words);

/* TemplatedVocabulary::transform(features, BowVector&, FeatureVector&, levelsup)
 * (TemplatedVocabulary.h:1127-1259; Frame.cc:738-745 calls it with levelsup = 4), TF_IDF + L1.
 * Per feature: word id, node id at level L-levelsup, weight.  Per frame: the FeatureVector CSR
 * (fv_out; d_nn < 0 = node_stride too small, -needed) and the BowVector (d_bow_word ascending,
 * d_bow_val normalised, d_bow_n entries; rows of kp_stride).  kp_stride <= 8192. */
int orbm_bow_transform(const orbm_vocabulary* voc, const uint8_t* d_desc, const int32_t* d_n,
                       int kp_stride, int batch, int levelsup, int32_t* d_feat_word,
                       int32_t* d_feat_node, double* d_feat_weight, const orbm_feature_vector* fv_out,
                       int node_stride, int32_t* d_bow_word, double* d_bow_val, int32_t* d_bow_n,
                       void* stream);

/* MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:367-401), batched over map points: point m
 * observes descriptors d_desc_pool[d_obs_idx[d_obs_off[m] .. d_obs_off[m+1])]; d_best[m] = position
 * (in that list) of the descriptor with the least median distance to the rest, d_median[m] its
 * median (may be NULL). */
int orbm_distinctive_descriptors(const uint8_t* d_desc_pool, const int32_t* d_obs_off,
                                 const int32_t* d_obs_idx, int n_points, int32_t* d_best,
                                 int32_t* d_median, void* stream);

/* ORBmatcher::SearchForInitialization (ORBmatcher.cc:648-763), batched over frame pairs (F1, F2):
 * level-0 keypoints of F1, window `window_size` around d_prev_matched[i] (vbPrevMatched, [batch]
 * [kp_stride][2], updated in place like the reference does at :757-759) on F2's orbm_grid_build CSR,
 * vMatchedDistance / vnMatches21 stealing rule, TH_LOW, ratio, rotation histogram (stale entries of
 * stolen matches keep counting, as in the reference).  d_match12 = vnMatches12; d_scratch: 2 * batch *
 * kp_stride int32. */
int orbm_search_for_initialization(
    const orbx_keypoint* d_kps1, const uint8_t* d_desc1, const int32_t* d_n1, const orbx_keypoint* d_kps2,
    const uint8_t* d_desc2, const int32_t* d_n2, int kp_stride, int batch, const int32_t* d_cell_off2,
    const int32_t* d_cell_idx2, float min_x, float max_x, float min_y, float max_y, float* d_prev_matched,
    int window_size, float nn_ratio, int check_orientation, int32_t* d_match12, int32_t* d_scratch,
    int32_t* d_nmatches, void* stream);

/* --------------------------------------------------------------------------------------
 * Per-point Frame geometry.
 * -------------------------------------------------------------------------------------- */

/* Frame::UndistortKeyPoints (Frame.cc:747-781): mvKeysUn = cv::undistortPoints(mvKeys, K, mDistCoef,
 * R = I, P = K) — 5 fixed-point iterations in double, bit-exact against OpenCV 4.13; every other
 * cv::KeyPoint field is copied.  dist_coef_host = mDistCoef (k1,k2,p1,p2[,k3[,k4,k5,k6[,s1..s4]]],
 * n_coef <= 12; dist_coef_host[0] == 0 -> mvKeysUn = mvKeys as the reference does).  The same call on
 * the four image corners is Frame::ComputeImageBounds (Frame.cc:783-809). */
int orbm_undistort_keypoints(const orbx_keypoint* d_kps, const int32_t* d_n, int kp_stride, int batch,
                             float fx, float fy, float cx, float cy, const float* dist_coef_host,
                             int n_coef, orbx_keypoint* d_kps_un, void* stream);

/* Frame::ComputeImageBounds (Frame.cc:782-808): bounds_host = {mnMinX, mnMaxX, mnMinY, mnMaxY} — the undistorted
 * image corners (same cv::undistortPoints arithmetic as above, min / max paired as the reference pairs them), or
 * {0, width, 0, height} when dist_coef_host[0] == 0.  Host only, no stream: these are the `min_x .. max_y` every
 * grid / search entry point takes. */
int orbm_compute_image_bounds(int width, int height, float fx, float fy, float cx, float cy,
                              const float* dist_coef_host, int n_coef, float* bounds_host);

/* GeometricCamera parameters (CameraModels/Pinhole.h, KannalaBrandt8.h: mvParameters). */
typedef struct orbm_camera {
  int32_t model;          /* 0 = Pinhole, 1 = KannalaBrandt8 */
  float fx, fy, cx, cy;
  float k[4];             /* KannalaBrandt8 k0..k3 */
} orbm_camera;

/* Frame::isInFrustum (Frame.cc:512-586, Nleft == -1; fisheye_rig = 0) and Frame::isInFrustumChecks
 * (Frame.cc:1168-1242; fisheye_rig = 1, one call per camera of the rig) with MapPoint::PredictScale
 * (MapPoint.cc:531-546), batched: frame f tests its d_nmp[f] MapPoints (rows of mp_stride).
 * d_pos / d_normal = GetWorldPos / GetNormal ([..][3]); d_min_dist / d_max_dist = the raw
 * mfMinDistance / mfMaxDistance members (the 0.8 / 1.2 factors of the getters are applied here).
 * d_pose[f][15] = R (row-major 3x3), t, camera centre: (mRcw, mtcw, mOw) for the left camera,
 * (Rrl*mRcw, Rrl*mtcw+trl, mRwc*tlr+mOw) for the right one (:1174-1184, host geometry).
 * Outputs: d_in_view = the return value; d_proj_x/y = mTrackProjX/Y (fisheye_rig = 0: also set when
 * only the distance / viewing-angle gates fail, like the reference; else -1), d_proj_xr =
 * mTrackProjXR (u - mbf/z; fisheye_rig = 0), d_depth = mTrackDepth, d_view_cos, d_level =
 * mnTrackScaleLevel (-1 when not in view).  d_proj_xr, d_depth, d_view_cos may be NULL.
 * Optional (d_qr != NULL): the search window of SearchByProjection(Frame&, vector<MapPoint*>&, th)
 * (ORBmatcher.cc:62-71): d_qr = RadiusByViewingCos(viewCos) [* th if th != 1] * scale[level],
 * d_qmin = level-1, d_qmax = level — the (qr, qmin, qmax) inputs of orbm_search_by_projection.
 * fp32 contract: uncontracted, Eigen's fixed-size evaluation order. */
int orbm_is_in_frustum(const float* d_pos, const float* d_normal, const float* d_min_dist,
                       const float* d_max_dist, const int32_t* d_nmp, int mp_stride, int batch,
                       const float* d_pose, const orbm_camera* cam, float mbf, float min_x, float max_x,
                       float min_y, float max_y, float viewing_cos_limit, float log_scale_factor,
                       int n_levels, int fisheye_rig, uint8_t* d_in_view, float* d_proj_x,
                       float* d_proj_y, float* d_proj_xr, float* d_depth, float* d_view_cos,
                       int8_t* d_level, float th, const float* scale_factors_host, float* d_qr,
                       int8_t* d_qmin, int8_t* d_qmax, void* stream);

/* --------------------------------------------------------------------------------------
 * Two-camera rig (fisheye stereo: F.Nleft != -1, KeyFrame::mpCamera2 != NULL) twins of the matcher
 * functions.  The reference numbers the rig's features left camera
 * first, right camera from Nleft on; mGrid / mGridRight are per camera (Frame.cc:385-416).
 * -------------------------------------------------------------------------------------- */

/* One camera of a rig as the projection search sees it: that camera's own keypoints / descriptors
 * (LOCAL numbering — the right camera's global index is local + Nleft), its orbm_grid_build CSR,
 * d_feat_locked[j] != 0: mvpMapPoints[j] holds a point with Observations()>0 on entry (NULL = none),
 * d_partner[j] = mvLeftToRightMatch[j] (left camera) / mvRightToLeftMatch[j] (right camera): local
 * index of the paired feature in the other camera or -1 (NULL = no pairs). */
typedef struct orbm_rig_camera {
  const orbx_keypoint* d_kps;
  const uint8_t* d_desc;
  const int32_t* d_n;
  int32_t kp_stride;
  const int32_t* d_cell_off;
  const int32_t* d_cell_idx;
  const uint8_t* d_feat_locked;
  const int32_t* d_partner;
} orbm_rig_camera;

/* ORBmatcher::SearchByProjection on a two-camera Frame, whole function, batched over frames:
 *   ORBM_PROJ_MAPPOINTS  ORBmatcher.cc:43-213  left block :59-142 + right block :144-210, with the
 *                        pair propagation (:131-135, :197-201) and the reference's `continue`s (a
 *                        failed left ratio test skips the right search of that MapPoint)
 *   ORBM_PROJ_LASTFRAME  ORBmatcher.cc:1676-1887  left :1716-1792 + right :1794-1862 (searched only
 *                        when the left window was not empty), one rotation histogram for both
 * Per query (MapPoint): the left window (d_qx, d_qy, d_qr, level band) and the right window (*_r),
 * one descriptor, d_q_flags bit0 = left query takes part, bit1 = Observations()>0, bit2 = right
 * query takes part (MAPPOINTS: mbTrackInViewR && mnTrackScaleLevelR != -1), d_q_angle (LASTFRAME).
 * d_scratch: batch * q_stride * 65 bytes.  Outputs in local numbering: d_match_l/r[q] = feature the
 * query itself was matched to, d_feat_query_l/r[j] = last query written into feature j (including
 * pair propagation; -1 none / erased), d_nmatches[f] = the return value. */
int orbm_search_by_projection_rig(
    int mode, const orbm_rig_camera* left, const orbm_rig_camera* right, int batch, float min_x,
    float max_x, float min_y, float max_y, const float* d_qx, const float* d_qy, const float* d_qr,
    const int8_t* d_qmin_level, const int8_t* d_qmax_level, const float* d_qx_r, const float* d_qy_r,
    const float* d_qr_r, const int8_t* d_qmin_level_r, const int8_t* d_qmax_level_r,
    const uint8_t* d_qdesc, const int32_t* d_nq, int q_stride, const uint8_t* d_q_flags,
    const float* d_q_angle, float nn_ratio, int check_orientation, void* d_scratch,
    int32_t* d_match_l, int32_t* d_match_r, int32_t* d_feat_query_l, int32_t* d_feat_query_r,
    int32_t* d_nmatches, void* stream);

/* ORBmatcher::SearchByBoW(KeyFrame*, Frame&, ...) with F.Nleft != -1 (ORBmatcher.cc:295-323,
 * 356-380): side 2 = the frame's left|right features in the reference's concatenated numbering,
 * d_nleft2[p] = F.Nleft; side 1 = the keyframe (its left|right keypoints concatenated likewise when
 * it has mpCamera2).  Separate best pairs per camera; d_match12_left / d_match12_right = the left /
 * right feature written for keyframe feature i, d_match21 = vpMapPointMatches (keyframe feature
 * whose MapPoint frame feature j received); the rotation histogram runs over the written features. */
int orbm_search_by_bow_rig(const orbx_keypoint* d_kps1, const uint8_t* d_desc1, const int32_t* d_n1,
                           const uint8_t* d_valid1, const orbm_feature_vector* fv1,
                           const orbx_keypoint* d_kps2, const uint8_t* d_desc2, const int32_t* d_n2,
                           const int32_t* d_nleft2, const orbm_feature_vector* fv2, int kp_stride,
                           int node_stride, int batch, float nn_ratio, int check_orientation,
                           int32_t* d_match12_left, int32_t* d_match12_right, int32_t* d_match21,
                           int32_t* d_nmatches, void* stream);

/* ORBmatcher::SearchForTriangulation between two keyframes of a two-camera rig (mpCamera2 != NULL;
 * ORBmatcher.cc:907-1146 with the camera selection :1036-1062), batched over keyframe pairs.  Features
 * are numbered left camera first (d_nleft* = NLeft).  d_rig_pose[p][4][12] = (R12 row-major, t12) of
 * (Tll, Tlr, Trl, Trr) (:929-937, host geometry); cams1 / cams2 = {mpCamera, mpCamera2} of KF1 / KF2
 * (KannalaBrandt8).  No mvuRight stereo and no epipole gate exist in this case (:973, :1022), so
 * only_stereo != 0 returns no matches; every candidate passes KannalaBrandt8::epipolarConstrain
 * (KannalaBrandt8.cpp:216-220, 306-375) with level_sigma2_1[kp1.octave] / level_sigma2_2[kp2.octave]
 * unless `coarse`.  That gate is floating-point triangulation: tolerance contract of DESIGN.md §2
 * (fp32 rays in source order, fp64 Jacobi null vector), not bit parity. */
int orbm_search_for_triangulation_rig(
    const orbx_keypoint* d_kps1, const uint8_t* d_desc1, const int32_t* d_n1, const int32_t* d_nleft1,
    const uint8_t* d_has_mp1, const orbm_feature_vector* fv1, const orbx_keypoint* d_kps2,
    const uint8_t* d_desc2, const int32_t* d_n2, const int32_t* d_nleft2, const uint8_t* d_has_mp2,
    const orbm_feature_vector* fv2, int kp_stride, int node_stride, int batch, const float* d_rig_pose,
    const orbm_camera* cams1, const orbm_camera* cams2, const float* level_sigma2_1_host,
    const float* level_sigma2_2_host, int nlevels, int only_stereo, int coarse, int check_orientation,
    int32_t* d_match12, int32_t* d_nmatches, void* stream);

/* Frame::ComputeStereoFishEyeMatches (Frame.cc:1126-1166), batched over rig frames: knnMatch(k=2) of the
 * lapping-area keypoints (indices >= d_mono_l / d_mono_r = the monoIndex the two extractors returned,
 * Frame.cc:1063-1067), Lowe ratio 0.7, KannalaBrandt8::TriangulateMatches with (mRlr, mtlr) and
 * mvLevelSigma2.  Outputs in each camera's numbering: d_left_to_right = mvLeftToRightMatch,
 * d_right_to_left = mvRightToLeftMatch (rows of kp_stride, -1 = none), d_depth = mvDepth (-1),
 * d_p3d[..][3] = mvStereo3Dpoints.  The triangulation carries the tolerance contract of
 * orbm_search_for_triangulation_rig. */
int orbm_stereo_fisheye_matches(const orbx_keypoint* d_kps_l, const uint8_t* d_desc_l, const int32_t* d_n_l,
                                const int32_t* d_mono_l, const orbx_keypoint* d_kps_r,
                                const uint8_t* d_desc_r, const int32_t* d_n_r, const int32_t* d_mono_r,
                                int kp_stride, int batch, const orbm_camera* cam_l,
                                const orbm_camera* cam_r, const float* Rlr_host, const float* tlr_host,
                                const float* level_sigma2_host, int nlevels, int32_t* d_left_to_right,
                                int32_t* d_right_to_left, float* d_depth, float* d_p3d, void* stream);

/* Library / build identification. */
const char* orbx_version(void);
const char* orbx_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* ORBX_H_ */
