// stereo.cu — the second half of Frame::ComputeStereoMatches (src/Frame.cc:896-981) and
// Frame::ComputeStereoFromRGBD (src/Frame.cc:984-1005) for sm_100a.
//
//   k_stereo_refine      one warp per left keypoint whose Hamming best is < (TH_HIGH+TH_LOW)/2: the
//                        11x11 L1-SAD slide over +-5 px on the keypoint's pyramid level (Frame.cc:908-933),
//                        parabola fit (:939-946), rescale, disparity gate, depth (:949-964)
//   k_stereo_median_cut  one CTA per pair: order statistic n/2 of the kept SADs by a two-pass radix
//                        select (replaces the std::sort at :967), threshold 1.5f*1.4f*median, erase
//   k_rgbd_depth         one thread per keypoint: depth image gather (:984-1005)
//
// The pyramids never leave the device: both eyes' levels are read where the extractor left them.
#include "../../include/orbx.h"
#include "orbx_kernels.cuh"

#define SR_WARPS 8
#define SR_W 5   // window half size  (Frame.cc:907)
#define SR_L 5   // slide half range  (Frame.cc:913)
#define SR_PW (2 * SR_W + 1)
#define SR_NS (2 * SR_L + 1)
#define SR_RW (SR_PW + 2 * SR_L)  // right strip width: 21

__global__ void __launch_bounds__(SR_WARPS * 32)
k_stereo_refine(const OrbxKeyPointDev* __restrict__ kl, const int32_t* __restrict__ nl_arr,
                const OrbxKeyPointDev* __restrict__ kr, int kp_stride, const int32_t* __restrict__ best_idx,
                const int32_t* __restrict__ best_dist, OrbxStereoPyr pyr, float mbf, float max_d,
                float* __restrict__ u_right, float* __restrict__ depth, int32_t* __restrict__ sad_out) {
  __shared__ uint8_t s_l[SR_WARPS][SR_PW * SR_PW + 3];
  __shared__ uint8_t s_r[SR_WARPS][SR_PW * SR_RW + 1];
  __shared__ int s_sad[SR_WARPS][SR_NS + 1];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, f = blockIdx.y;
  const int iL = blockIdx.x * SR_WARPS + warp;
  if (iL >= nl_arr[f]) return;
  const size_t o = (size_t)f * kp_stride + iL;
  // mvuRight / mvDepth start at -1 (:813-814); sad < 0 marks "not in vDistIdx"
  if (lane == 0) { u_right[o] = -1.0f; depth[o] = -1.0f; sad_out[o] = -1; }
  const int bd = best_dist[o], bi = best_idx[o];
  const int thOrbDist = (100 + 50) / 2;  // (TH_HIGH + TH_LOW) / 2, :816
  if (bi < 0 || bd >= thOrbDist) return;
  const OrbxKeyPointDev kpL = kl[o];
  const int lvl = kpL.octave;
  const float uR0 = kr[(size_t)f * kp_stride + bi].x;
  const float inv = pyr.inv_scale[lvl];
  const float scaleduL = roundf(kpL.x * inv), scaledvL = roundf(kpL.y * inv), scaleduR0 = roundf(uR0 * inv);
  const int W = pyr.w[lvl];
  const float iniu = scaleduR0 + (float)SR_L - (float)SR_W, endu = scaleduR0 + (float)SR_L + (float)SR_W + 1.0f;
  if (iniu < 0 || endu >= (float)W) return;  // :916-919
  const int cu = (int)scaleduL, cv = (int)scaledvL, cr = (int)scaleduR0;
  // cv::Mat::rowRange/colRange would throw outside the level; keypoints sit >= 19 px (level units)
  // inside, so these never trigger on extractor output — guard anyway instead of reading wild.
  if (cv - SR_W < 0 || cv + SR_W >= pyr.h[lvl] || cu - SR_W < 0 || cu + SR_W >= W) return;
  const long long pl = pyr.pitch_l[lvl], pr = pyr.pitch_r[lvl];
  const uint8_t* IL = pyr.base_l[lvl] + (long long)f * pyr.stride_l[lvl] + (long long)(cv - SR_W) * pl + (cu - SR_W);
  const uint8_t* IR = pyr.base_r[lvl] + (long long)f * pyr.stride_r[lvl] + (long long)(cv - SR_W) * pr + (cr - SR_L - SR_W);
  for (int i = lane; i < SR_PW * SR_PW; i += 32) {
    const int r = i / SR_PW, c = i - r * SR_PW;
    s_l[warp][i] = __ldg(IL + r * pl + c);
  }
  for (int i = lane; i < SR_PW * SR_RW; i += 32) {
    const int r = i / SR_RW, c = i - r * SR_RW;
    s_r[warp][i] = __ldg(IR + r * pr + c);
  }
  if (lane < SR_NS) s_sad[warp][lane] = 0;
  __syncwarp();
  // cv::norm(IL, IR, NORM_L1) for the 11 shifts: task = (row, shift), 121 tasks over 32 lanes
  for (int t = lane; t < SR_PW * SR_NS; t += 32) {
    const int r = t / SR_NS, s = t - r * SR_NS;
    const uint8_t* a = &s_l[warp][r * SR_PW];
    const uint8_t* b = &s_r[warp][r * SR_RW + s];
    int acc = 0;
#pragma unroll
    for (int c = 0; c < SR_PW; c++) acc += abs((int)a[c] - (int)b[c]);
    atomicAdd(&s_sad[warp][s], acc);
  }
  __syncwarp();
  if (lane != 0) return;
  // first strict minimum over incR = -L..L (:925-931); SADs <= 121*255 are exact in float
  int bestSad = 0x7fffffff, bestinc = 0;
#pragma unroll
  for (int s = 0; s < SR_NS; s++)
    if (s_sad[warp][s] < bestSad) { bestSad = s_sad[warp][s]; bestinc = s - SR_L; }
  if (bestinc == -SR_L || bestinc == SR_L) return;  // :935-936
  const float dist1 = (float)s_sad[warp][SR_L + bestinc - 1], dist2 = (float)s_sad[warp][SR_L + bestinc],
              dist3 = (float)s_sad[warp][SR_L + bestinc + 1];
  const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2));  // :943 (NaN falls through)
  if (deltaR < -1 || deltaR > 1) return;
  float bestuR = pyr.scale[lvl] * (scaleduR0 + (float)bestinc + deltaR);  // :949
  float disparity = kpL.x - bestuR;
  if (disparity >= 0.0f && disparity < max_d) {
    if (disparity <= 0) {
      disparity = 0.01f;                          // `disparity=0.01` : double literal stored to float
      bestuR = (float)((double)kpL.x - 0.01);     // `uL-0.01` is evaluated in double (:957)
    }
    depth[o] = mbf / disparity;
    u_right[o] = bestuR;
    sad_out[o] = bestSad;
  }
}

// Median cut (:967-980).  The reference sorts (SAD, iL) pairs and takes .first of element n/2: that is
// the (n/2)-th order statistic of the SADs, independent of how ties are ordered.  SAD < 2^15, so two
// counting passes (high 8 bits, low 7 bits) select it.  Everything with SAD >= 2.1f*median is erased.
__global__ void __launch_bounds__(256)
k_stereo_median_cut(const int32_t* __restrict__ nl_arr, int kp_stride, int32_t* __restrict__ sad,
                    float* __restrict__ u_right, float* __restrict__ depth, int32_t* __restrict__ kept) {
  __shared__ int hist[256];
  __shared__ int s_sel[3];  // chosen high bucket, rank inside it, total
  const int f = blockIdx.x, tid = threadIdx.x, n = nl_arr[f];
  const size_t base = (size_t)f * kp_stride;
  hist[tid] = 0;
  __syncthreads();
  for (int i = tid; i < n; i += 256) {
    const int s = sad[base + i];
    if (s >= 0) atomicAdd(&hist[s >> 7], 1);
  }
  __syncthreads();
  if (tid == 0) {
    int tot = 0;
    for (int b = 0; b < 256; b++) tot += hist[b];
    int k = tot / 2, b = 0;
    if (tot > 0) while (k >= hist[b]) { k -= hist[b]; b++; }
    s_sel[0] = b; s_sel[1] = k; s_sel[2] = tot;
  }
  __syncthreads();
  const int hb = s_sel[0], tot = s_sel[2];
  int k = s_sel[1];
  if (tot == 0) { if (tid == 0) kept[f] = 0; return; }
  __syncthreads();
  hist[tid] = 0;
  __syncthreads();
  for (int i = tid; i < n; i += 256) {
    const int s = sad[base + i];
    if (s >= 0 && (s >> 7) == hb) atomicAdd(&hist[s & 127], 1);
  }
  __syncthreads();
  if (tid == 0) {
    int b = 0;
    while (k >= hist[b]) { k -= hist[b]; b++; }
    s_sel[0] = (hb << 7) | b;
    s_sel[1] = 0;
  }
  __syncthreads();
  const float median = (float)s_sel[0];
  const float thDist = (1.5f * 1.4f) * median;  // :969, left-to-right: (1.5f*1.4f) folds to 2.1f
  int mine = 0;
  for (int i = tid; i < n; i += 256) {
    const int s = sad[base + i];
    if (s < 0) continue;
    if ((float)s < thDist) mine++;
    else { u_right[base + i] = -1.0f; depth[base + i] = -1.0f; }
  }
  atomicAdd(&s_sel[1], mine);
  __syncthreads();
  if (tid == 0) kept[f] = s_sel[1];
}

// Frame::ComputeStereoFromRGBD (:984-1005): d = imDepth.at<float>(v,u) with v,u truncated to int;
// d>0 -> mvDepth = d, mvuRight = kpU.x - mbf/d.
__global__ void __launch_bounds__(256)
k_rgbd_depth(const OrbxKeyPointDev* __restrict__ kps, const OrbxKeyPointDev* __restrict__ kps_un,
             const int32_t* __restrict__ n_arr, int kp_stride, const float* __restrict__ depth_img,
             long long pitch_f, long long frame_stride_f, int w, int h, float mbf,
             float* __restrict__ u_right, float* __restrict__ depth) {
  const int f = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_arr[f]) return;
  const size_t o = (size_t)f * kp_stride + i;
  const OrbxKeyPointDev kp = kps[o];
  const int v = (int)kp.y, u = (int)kp.x;
  float ur = -1.0f, dp = -1.0f;
  if (u >= 0 && u < w && v >= 0 && v < h) {  // the reference indexes unchecked
    const float d = __ldg(depth_img + (long long)f * frame_stride_f + (long long)v * pitch_f + u);
    if (d > 0) { dp = d; ur = kps_un[o].x - mbf / d; }
  }
  u_right[o] = ur;
  depth[o] = dp;
}

cudaError_t orbx_launch_stereo_refine(const OrbxKeyPointDev* kl, const int32_t* nl, const OrbxKeyPointDev* kr,
                                      int kp_stride, int batch, const int32_t* best_idx, const int32_t* best_dist,
                                      const OrbxStereoPyr& pyr, float mbf, float max_d, float* u_right, float* depth,
                                      int32_t* sad, int32_t* kept, cudaStream_t st) {
  dim3 grd((kp_stride + SR_WARPS - 1) / SR_WARPS, batch);
  k_stereo_refine<<<grd, SR_WARPS * 32, 0, st>>>(kl, nl, kr, kp_stride, best_idx, best_dist, pyr, mbf, max_d, u_right,
                                                 depth, sad);
  k_stereo_median_cut<<<batch, 256, 0, st>>>(nl, kp_stride, sad, u_right, depth, kept);
  return cudaGetLastError();
}

cudaError_t orbx_launch_rgbd_depth(const OrbxKeyPointDev* kps, const OrbxKeyPointDev* kps_un, const int32_t* n,
                                   int kp_stride, int batch, const float* depth_img, long long pitch_f,
                                   long long frame_stride_f, int w, int h, float mbf, float* u_right, float* depth,
                                   cudaStream_t st) {
  dim3 grd((kp_stride + 255) / 256, batch);
  k_rgbd_depth<<<grd, 256, 0, st>>>(kps, kps_un, n, kp_stride, depth_img, pitch_f, frame_stride_f, w, h, mbf, u_right,
                                    depth);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------
// Frame::ComputeStereoFishEyeMatches (src/Frame.cc:1126-1166): brute-force knnMatch(k=2) between the lapping-area
// keypoints of the two fisheye images (cv::BFMatcher(NORM_HAMMING), ties to the lower train index), Lowe ratio 0.7,
// KannalaBrandt8::TriangulateMatches (tri_core.h) -> mvLeftToRightMatch, mvRightToLeftMatch, mvStereo3Dpoints, mvDepth.
// One warp per left keypoint; the reference's loop runs in ascending left index and simply overwrites
// mvRightToLeftMatch[train], so the last (largest) left index wins: atomicMax.
// ---------------------------------------------------------------------------------------------------------
#include "match_core.cuh"
#include "tri_core.h"

struct FishEyeArgs {
  const OrbxKeyPointDev *kl, *kr;
  const uint8_t *dl, *dr;
  const int32_t *nl, *nr, *mono_l, *mono_r;
  int kp_stride;
  OrbxFrustumCam cam_l, cam_r;
  OrbxRigPose Tlr;
  float sigma2[ORBX_MAX_LEVELS];
  int32_t *l2r, *r2l;
  float *depth, *p3d;
};

__global__ void __launch_bounds__(SR_WARPS * 32)
k_fisheye_init(const __grid_constant__ FishEyeArgs A) {
  const int f = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= A.kp_stride) return;
  const size_t o = (size_t)f * A.kp_stride + i;
  A.l2r[o] = -1; A.r2l[o] = -1; A.depth[o] = -1.0f;
  A.p3d[3 * o] = 0.f; A.p3d[3 * o + 1] = 0.f; A.p3d[3 * o + 2] = 0.f;
}

__global__ void __launch_bounds__(SR_WARPS * 32)
k_fisheye_matches(const __grid_constant__ FishEyeArgs A) {
  const int lane = threadIdx.x & 31, f = blockIdx.y;
  const int nl = A.nl[f], nr = A.nr[f], ml = A.mono_l[f], mr = A.mono_r[f];
  const int qi = ml + blockIdx.x * SR_WARPS + (threadIdx.x >> 5);
  if (qi >= nl) return;
  const size_t fb = (size_t)f * A.kp_stride;
  const Desc256 dq = load_desc(A.dl + (fb + qi) * 32);
  Top2 t;
  for (int j = mr + lane; j < nr; j += 32) {
    const int d = hamming(dq, load_desc(A.dr + (fb + j) * 32));
    t.push(((uint32_t)d << 20) | (uint32_t)j, j);
  }
  warp_top2(t);
  if (lane != 0 || t.i2 < 0) return;  // (*it).size() >= 2
  const float d0 = (float)(t.k1 >> 20), d1 = (float)(t.k2 >> 20);
  if (!((double)d0 < d_mul((double)d1, 0.7))) return;  // float * double literal
  const OrbxKeyPointDev kl = A.kl[fb + qi], kr = A.kr[fb + t.i1];
  float p[3];
  const float depth = orbx_kb8_triangulate_matches(A.cam_l, A.cam_r, kl.x, kl.y, kr.x, kr.y, A.Tlr, A.sigma2[kl.octave],
                                                   A.sigma2[kr.octave], p);
  if (depth > 0.0001f) {
    A.l2r[fb + qi] = t.i1;
    atomicMax(&A.r2l[fb + t.i1], qi);
    A.p3d[3 * (fb + qi)] = p[0]; A.p3d[3 * (fb + qi) + 1] = p[1]; A.p3d[3 * (fb + qi) + 2] = p[2];
    A.depth[fb + qi] = depth;
  }
}

extern "C" int orbm_stereo_fisheye_matches(const orbx_keypoint* d_kps_l, const uint8_t* d_desc_l, const int32_t* d_n_l,
                                           const int32_t* d_mono_l, const orbx_keypoint* d_kps_r,
                                           const uint8_t* d_desc_r, const int32_t* d_n_r, const int32_t* d_mono_r,
                                           int kp_stride, int batch, const orbm_camera* cam_l, const orbm_camera* cam_r,
                                           const float* Rlr_host, const float* tlr_host,
                                           const float* level_sigma2_host, int nlevels, int32_t* d_left_to_right,
                                           int32_t* d_right_to_left, float* d_depth, float* d_p3d, void* stream) {
  if (!d_kps_l || !d_desc_l || !d_n_l || !d_mono_l || !d_kps_r || !d_desc_r || !d_n_r || !d_mono_r || kp_stride < 1 ||
      batch < 1 || !cam_l || !cam_r || cam_l->model != 1 || cam_r->model != 1 || !Rlr_host || !tlr_host ||
      !level_sigma2_host || nlevels < 1 || nlevels > ORBX_MAX_LEVELS || !d_left_to_right || !d_right_to_left ||
      !d_depth || !d_p3d)
    return ORBX_E_INVALID;
  FishEyeArgs A;
  A.kl = (const OrbxKeyPointDev*)d_kps_l; A.kr = (const OrbxKeyPointDev*)d_kps_r; A.dl = d_desc_l; A.dr = d_desc_r;
  A.nl = d_n_l; A.nr = d_n_r; A.mono_l = d_mono_l; A.mono_r = d_mono_r; A.kp_stride = kp_stride;
  const orbm_camera* cs[2] = {cam_l, cam_r};
  OrbxFrustumCam* os[2] = {&A.cam_l, &A.cam_r};
  for (int c = 0; c < 2; c++) {
    os[c]->model = 1; os[c]->fx = cs[c]->fx; os[c]->fy = cs[c]->fy; os[c]->cx = cs[c]->cx; os[c]->cy = cs[c]->cy;
    for (int i = 0; i < 4; i++) os[c]->k[i] = cs[c]->k[i];
  }
  for (int i = 0; i < 9; i++) A.Tlr.R[i] = Rlr_host[i];
  for (int i = 0; i < 3; i++) A.Tlr.t[i] = tlr_host[i];
  for (int l = 0; l < ORBX_MAX_LEVELS; l++) A.sigma2[l] = l < nlevels ? level_sigma2_host[l] : 0.f;
  A.l2r = d_left_to_right; A.r2l = d_right_to_left; A.depth = d_depth; A.p3d = d_p3d;
  cudaStream_t st = (cudaStream_t)stream;
  k_fisheye_init<<<dim3((kp_stride + SR_WARPS * 32 - 1) / (SR_WARPS * 32), batch), SR_WARPS * 32, 0, st>>>(A);
  k_fisheye_matches<<<dim3((kp_stride + SR_WARPS - 1) / SR_WARPS, batch), SR_WARPS * 32, 0, st>>>(A);
  return cudaGetLastError() == cudaSuccess ? ORBX_OK : ORBX_E_CUDA;
}
