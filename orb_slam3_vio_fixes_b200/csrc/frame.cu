// frame.cu — per-point Frame geometry on the device, batched over frames:
//   k_undistort    Frame::UndistortKeyPoints (src/Frame.cc:747-781): cv::undistortPoints, one thread per keypoint,
//                  double arithmetic in OpenCV's source order (frame_core.h) -> mvKeysUn.  Also serves
//                  Frame::ComputeImageBounds (:783-809), which undistorts the four image corners the same way.
//   k_in_frustum   Frame::isInFrustum (:512-586) / isInFrustumChecks (:1168-1242) + MapPoint::PredictScale
//                  (src/MapPoint.cc:531-546), one thread per MapPoint, and — optionally — the search window of
//                  ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&) (src/ORBmatcher.cc:62-71) so that the
//                  outputs feed orbm_search_by_projection without a host pass.
// Both are pure streaming kernels (28-byte records in, 28-byte records out; 32 B in / 24 B out per MapPoint).
#include "../../include/orbx.h"
#include "frame_core.h"
#include "orbx_kernels.cuh"

struct UndistortArgs {
  double k[12];
  float fx, fy, cx, cy;
  int identity;  // mDistCoef.at<float>(0) == 0.0: mvKeysUn = mvKeys (:749-753)
};

__global__ void __launch_bounds__(256)
k_undistort(const OrbxKeyPointDev* __restrict__ kps, const int32_t* __restrict__ n, int kp_stride,
            const __grid_constant__ UndistortArgs A, OrbxKeyPointDev* __restrict__ out) {
  const int f = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n[f]) return;
  const size_t o = (size_t)f * kp_stride + i;
  OrbxKeyPointDev kp = kps[o];
  if (!A.identity) {
    float u, v;
    orbx_undistort_point(kp.x, kp.y, A.fx, A.fy, A.cx, A.cy, A.k, &u, &v);
    kp.x = u; kp.y = v;
  }
  out[o] = kp;
}

extern "C" int orbm_undistort_keypoints(const orbx_keypoint* d_kps, const int32_t* d_n, int kp_stride, int batch,
                                        float fx, float fy, float cx, float cy, const float* dist_coef_host,
                                        int n_coef, orbx_keypoint* d_kps_un, void* stream) {
  if (!d_kps || !d_n || !d_kps_un || kp_stride < 1 || batch < 1 || n_coef < 0 || n_coef > 12 ||
      (n_coef > 0 && !dist_coef_host) || !(fx != 0.0f) || !(fy != 0.0f))
    return ORBX_E_INVALID;
  UndistortArgs A;
  for (int i = 0; i < 12; i++) A.k[i] = i < n_coef ? (double)dist_coef_host[i] : 0.0;
  A.fx = fx; A.fy = fy; A.cx = cx; A.cy = cy;
  A.identity = (n_coef == 0 || dist_coef_host[0] == 0.0f) ? 1 : 0;
  dim3 grd((kp_stride + 255) / 256, batch);
  k_undistort<<<grd, 256, 0, (cudaStream_t)stream>>>((const OrbxKeyPointDev*)d_kps, d_n, kp_stride, A,
                                                      (OrbxKeyPointDev*)d_kps_un);
  return cudaGetLastError() == cudaSuccess ? ORBX_OK : ORBX_E_CUDA;
}

// Frame::ComputeImageBounds (Frame.cc:782-808): the four image corners through cv::undistortPoints, min / max taken
// the way the reference pairs them (left corners for mnMinX, right corners for mnMaxX, ...).  Host only: four points.
extern "C" int orbm_compute_image_bounds(int width, int height, float fx, float fy, float cx, float cy,
                                         const float* dist_coef_host, int n_coef, float* bounds_host) {
  if (!bounds_host || width < 1 || height < 1 || n_coef < 0 || n_coef > 12 || (n_coef > 0 && !dist_coef_host) ||
      !(fx != 0.0f) || !(fy != 0.0f))
    return ORBX_E_INVALID;
  if (n_coef == 0 || dist_coef_host[0] == 0.0f) {  // :802-808
    bounds_host[0] = 0.0f; bounds_host[1] = (float)width; bounds_host[2] = 0.0f; bounds_host[3] = (float)height;
    return ORBX_OK;
  }
  double k[12];
  for (int i = 0; i < 12; i++) k[i] = i < n_coef ? (double)dist_coef_host[i] : 0.0;
  const float cu[4] = {0.0f, (float)width, 0.0f, (float)width}, cv[4] = {0.0f, 0.0f, (float)height, (float)height};
  float x[4], y[4];
  for (int i = 0; i < 4; i++) orbx_undistort_point(cu[i], cv[i], fx, fy, cx, cy, k, &x[i], &y[i]);
  bounds_host[0] = x[0] < x[2] ? x[0] : x[2];  // mnMinX = min(mat(0,0), mat(2,0))
  bounds_host[1] = x[1] > x[3] ? x[1] : x[3];  // mnMaxX = max(mat(1,0), mat(3,0))
  bounds_host[2] = y[0] < y[1] ? y[0] : y[1];  // mnMinY = min(mat(0,1), mat(1,1))
  bounds_host[3] = y[2] > y[3] ? y[2] : y[3];  // mnMaxY = max(mat(2,1), mat(3,1))
  return ORBX_OK;
}

struct FrustumArgs {
  const float *pos, *normal, *min_dist, *max_dist;
  const int32_t* nmp;
  int mp_stride;
  const OrbxFrustumPose* pose;  // [batch]
  OrbxFrustumCam cam;
  float mbf, min_x, max_x, min_y, max_y, cos_limit, log_sf;
  int n_levels, fisheye_rig;
  float th;                               // window factor of SearchByProjection (<= 0: no window outputs)
  float scale[ORBX_MAX_LEVELS];
  uint8_t* in_view;
  float *proj_x, *proj_y, *proj_xr, *depth, *view_cos;
  int8_t* level;
  float* qr; int8_t *qmin, *qmax;
};

__global__ void __launch_bounds__(256)
k_in_frustum(const __grid_constant__ FrustumArgs A) {
  const int f = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= A.nmp[f]) return;
  const size_t o = (size_t)f * A.mp_stride + i;
  const float P[3] = {A.pos[3 * o], A.pos[3 * o + 1], A.pos[3 * o + 2]};
  const float N[3] = {A.normal[3 * o], A.normal[3 * o + 1], A.normal[3 * o + 2]};
  const OrbxFrustumOut r = orbx_in_frustum(P, N, A.min_dist[o], A.max_dist[o], A.pose[f], A.cam, A.mbf, A.min_x, A.max_x,
                                           A.min_y, A.max_y, A.cos_limit, A.log_sf, A.n_levels, A.fisheye_rig);
  A.in_view[o] = (uint8_t)r.in_view;
  A.proj_x[o] = r.proj_x; A.proj_y[o] = r.proj_y;
  if (A.proj_xr) A.proj_xr[o] = r.proj_xr;
  if (A.depth) A.depth[o] = r.depth;
  if (A.view_cos) A.view_cos[o] = r.view_cos;
  A.level[o] = (int8_t)r.level;
  if (A.qr) {
    float q = 0.0f;
    int lo = 0, hi = 0;
    if (r.in_view) {
      // RadiusByViewingCos (ORBmatcher.cc:215-221): `viewCos>0.998` compares the float promoted to double
      float rad = (double)r.view_cos > 0.998 ? 2.5f : 4.0f;
      if (A.th != 1.0f) rad = f_mul(rad, A.th);     // bFactor, :47,69-70
      q = f_mul(rad, A.scale[r.level]);
      lo = r.level - 1; hi = r.level;
    }
    A.qr[o] = q; A.qmin[o] = (int8_t)lo; A.qmax[o] = (int8_t)hi;
  }
}

extern "C" int orbm_is_in_frustum(const float* d_pos, const float* d_normal, const float* d_min_dist,
                                  const float* d_max_dist, const int32_t* d_nmp, int mp_stride, int batch,
                                  const float* d_pose, const orbm_camera* cam, float mbf, float min_x, float max_x,
                                  float min_y, float max_y, float viewing_cos_limit, float log_scale_factor,
                                  int n_levels, int fisheye_rig, uint8_t* d_in_view, float* d_proj_x, float* d_proj_y,
                                  float* d_proj_xr, float* d_depth, float* d_view_cos, int8_t* d_level, float th,
                                  const float* scale_factors_host, float* d_qr, int8_t* d_qmin, int8_t* d_qmax,
                                  void* stream) {
  if (!d_pos || !d_normal || !d_min_dist || !d_max_dist || !d_nmp || !d_pose || !cam || !d_in_view || !d_proj_x ||
      !d_proj_y || !d_level || mp_stride < 1 || batch < 1 || n_levels < 1 || n_levels > ORBX_MAX_LEVELS ||
      (cam->model != 0 && cam->model != 1))
    return ORBX_E_INVALID;
  if (d_qr && (!d_qmin || !d_qmax || !scale_factors_host || !(th > 0.0f))) return ORBX_E_INVALID;
  FrustumArgs A;
  A.pos = d_pos; A.normal = d_normal; A.min_dist = d_min_dist; A.max_dist = d_max_dist; A.nmp = d_nmp;
  A.mp_stride = mp_stride; A.pose = (const OrbxFrustumPose*)d_pose;
  A.cam.model = cam->model; A.cam.fx = cam->fx; A.cam.fy = cam->fy; A.cam.cx = cam->cx; A.cam.cy = cam->cy;
  for (int i = 0; i < 4; i++) A.cam.k[i] = cam->k[i];
  A.mbf = mbf; A.min_x = min_x; A.max_x = max_x; A.min_y = min_y; A.max_y = max_y;
  A.cos_limit = viewing_cos_limit; A.log_sf = log_scale_factor; A.n_levels = n_levels; A.fisheye_rig = fisheye_rig ? 1 : 0;
  A.th = d_qr ? th : 0.0f;
  for (int i = 0; i < ORBX_MAX_LEVELS; i++) A.scale[i] = (d_qr && i < n_levels) ? scale_factors_host[i] : 1.0f;
  A.in_view = d_in_view; A.proj_x = d_proj_x; A.proj_y = d_proj_y; A.proj_xr = d_proj_xr; A.depth = d_depth;
  A.view_cos = d_view_cos; A.level = d_level; A.qr = d_qr; A.qmin = d_qmin; A.qmax = d_qmax;
  dim3 grd((mp_stride + 255) / 256, batch);
  k_in_frustum<<<grd, 256, 0, (cudaStream_t)stream>>>(A);
  return cudaGetLastError() == cudaSuccess ? ORBX_OK : ORBX_E_CUDA;
}
