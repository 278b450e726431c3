// StereoMatcher.h — Frame::ComputeStereoMatches (src/Frame.cc:811-981) and Frame::ComputeStereoFromRGBD
// (src/Frame.cc:984-1005) on the device-resident outputs of the two drop-in extractors.
//
// In the reference, Frame's stereo constructor runs the two extractors (Frame.cc:122-125) and then
// ComputeStereoMatches(), which reads mvKeys / mvKeysRight / mDescriptors / mDescriptorsRight and both
// mvImagePyramid.  All of that is still on the GPU after the two operator() calls, so the body becomes:
//
//   void Frame::ComputeStereoMatches() {
//     ORB_SLAM3::ComputeStereoMatches(*mpORBextractorLeft, *mpORBextractorRight, mb, mbf, mvuRight, mvDepth);
//   }
//
// Results are bit-identical to the reference loop (tests/test_gpu_compat_cpp.py, tests/test_gpu_match.py).
#ifndef ORBX_STEREO_MATCHER_H
#define ORBX_STEREO_MATCHER_H

#include <cuda_runtime.h>

#include <vector>

#include "../orbx.h"
#include "ORBextractor.h"

namespace ORB_SLAM3 {

// Returns the number of stereo matches kept (>= 0) or a negative ORBX_E_* code.
inline int ComputeStereoMatches(ORBextractor& left, ORBextractor& right, float mb, float mbf,
                                std::vector<float>& mvuRight, std::vector<float>& mvDepth) {
  orbx_handle hl = left.handle(), hr = right.handle();
  if (!hl || !hr) return ORBX_E_INVALID;
  const orbx_keypoint *kl, *kr;
  const uint8_t *dl, *dr;
  const int32_t *nl, *nr;
  int cap = 0, cap_r = 0;
  orbx_last_outputs_device(hl, &kl, &dl, &nl, nullptr, &cap);
  orbx_last_outputs_device(hr, &kr, &dr, &nr, nullptr, &cap_r);
  if (cap != cap_r) return ORBX_E_INVALID;  // both eyes are configured alike (Tracking.cc:595-601)
  int N = 0;
  if (cudaMemcpy(&N, nl, sizeof(int), cudaMemcpyDeviceToHost) != cudaSuccess) return ORBX_E_CUDA;
  mvuRight.assign(N, -1.0f);  // Frame.cc:813-814
  mvDepth.assign(N, -1.0f);
  if (N == 0) return 0;
  const std::vector<float> sf = left.GetScaleFactors();
  const float maxD = mbf / mb;  // minZ = mb (Frame.cc:841-843)
  int32_t *bi = nullptr, *bd = nullptr, *sad = nullptr, *kept = nullptr;
  float *ur = nullptr, *dp = nullptr;
  int rc = ORBX_E_NOMEM, nkept = 0;
  const size_t ib = (size_t)cap * sizeof(int32_t), fb = (size_t)cap * sizeof(float);
  if (cudaMalloc(&bi, ib) || cudaMalloc(&bd, ib) || cudaMalloc(&sad, ib) || cudaMalloc(&kept, 4) ||
      cudaMalloc(&ur, fb) || cudaMalloc(&dp, fb))
    goto done;
  rc = orbm_stereo_rowband(kl, dl, nl, kr, dr, nr, cap, 1, sf.data(), (int)sf.size(), maxD, bi, bd, nullptr);
  if (rc == ORBX_OK) rc = orbm_stereo_refine(hl, hr, kl, nl, kr, cap, 1, bi, bd, mbf, maxD, ur, dp, sad, kept, nullptr);
  if (rc == ORBX_OK &&
      (cudaMemcpy(mvuRight.data(), ur, (size_t)N * sizeof(float), cudaMemcpyDeviceToHost) ||
       cudaMemcpy(mvDepth.data(), dp, (size_t)N * sizeof(float), cudaMemcpyDeviceToHost) ||
       cudaMemcpy(&nkept, kept, sizeof(int), cudaMemcpyDeviceToHost)))
    rc = ORBX_E_CUDA;
done:
  cudaFree(bi); cudaFree(bd); cudaFree(sad); cudaFree(kept); cudaFree(ur); cudaFree(dp);
  return rc == ORBX_OK ? nkept : rc;
}

}  // namespace ORB_SLAM3

#endif
