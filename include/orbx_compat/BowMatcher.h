// BowMatcher.h — ORBmatcher::SearchByBoW (src/ORBmatcher.cc:223-425, 765-905) and
// ORBmatcher::SearchForTriangulation (:907-1146) of the reference on liborbx.so, taking the containers the reference
// bodies already hold: std::vector<cv::KeyPoint> (mvKeysUn), the N x 32 descriptor rows (mDescriptors.data), the
// DBoW2::FeatureVector (a std::map<NodeId, std::vector<unsigned>>, Thirdparty/DBoW2/DBoW2/FeatureVector.h) and one
// byte per feature for the object state the loops consult.  The reference bodies become:
//
//   int ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame& F, vector<MapPoint*>& vpMapPointMatches) {
//     BowSide kf{&pKF->mvKeysUn, pKF->mDescriptors.data, &pKF->mFeatVec, &validKF /* pMP && !pMP->isBad() */, nullptr};
//     BowSide fr{&F.mvKeys, F.mDescriptors.data, &F.mFeatVec, nullptr, nullptr};
//     std::vector<int> m12, m21;
//     int n = SearchByBoWDevice(ORBM_BOW_KF_FRAME, kf, fr, mfNNratio, mbCheckOrientation, m12, m21);
//     for (j : F features) vpMapPointMatches[j] = m21[j] >= 0 ? vpMapPointsKF[m21[j]] : NULL;   return n; }
//
// Single-camera case (Nleft == -1, pinhole), as include/orbx.h states.  Synchronous; uploads both sides per call
// (a resident pipeline keeps them on the device and calls orbm_* directly).
#ifndef ORBX_COMPAT_BOWMATCHER_H
#define ORBX_COMPAT_BOWMATCHER_H

#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include <map>
#include <utility>
#include <vector>

#include "../orbx.h"
#include "cv_compat.h"

namespace ORB_SLAM3 {

typedef std::map<unsigned int, std::vector<unsigned int> > FeatureVectorMap;  // base of DBoW2::FeatureVector

struct BowSide {
  const std::vector<cv::KeyPoint>* keys;   // mvKeysUn
  const unsigned char* desc;               // mDescriptors.data, keys->size() x 32
  const FeatureVectorMap* fv;              // mFeatVec
  const std::vector<unsigned char>* flag;  // SearchByBoW: MapPoint valid; triangulation: GetMapPoint(i) != NULL
  const std::vector<float>* uRight;        // mvuRight (triangulation), may be null
};

namespace bowdetail {
struct DevSide {
  void* kps = nullptr; void* desc = nullptr; void* n = nullptr; void* flag = nullptr; void* ur = nullptr;
  orbm_feature_vector fv = {nullptr, nullptr, nullptr, nullptr};
  ~DevSide() {
    cudaFree(kps); cudaFree(desc); cudaFree(n); cudaFree(flag); cudaFree(ur);
    cudaFree(fv.d_node); cudaFree(fv.d_off); cudaFree(fv.d_idx); cudaFree(fv.d_nn);
  }
};
struct Buf {  // device scratch that frees itself on every path
  void* p = nullptr;
  Buf() {}
  Buf(const Buf&) = delete;
  Buf& operator=(const Buf&) = delete;
  ~Buf() { cudaFree(p); }
  bool alloc(size_t bytes) { return cudaMalloc(&p, bytes ? bytes : 1) == cudaSuccess; }
};
template <class T>
inline bool up(void** d, const T* h, size_t count) {
  if (cudaMalloc(d, (count ? count : 1) * sizeof(T)) != cudaSuccess) return false;
  return count == 0 || cudaMemcpy(*d, h, count * sizeof(T), cudaMemcpyHostToDevice) == cudaSuccess;
}
// std::map iteration = ascending node id; each vector is already in ascending feature order (addFeature)
inline bool upload(const BowSide& s, int cap, int node_stride, DevSide& d) {
  const int32_t n = (int32_t)s.keys->size();
  std::vector<int32_t> node, off(1, 0), idx;
  for (FeatureVectorMap::const_iterator it = s.fv->begin(); it != s.fv->end(); ++it) {
    node.push_back((int32_t)it->first);
    idx.insert(idx.end(), it->second.begin(), it->second.end());
    off.push_back((int32_t)idx.size());
  }
  const int32_t nn = (int32_t)node.size();
  node.resize(node_stride, 0); off.resize(node_stride + 1, off.back()); idx.resize(cap, 0);
  std::vector<cv::KeyPoint> k(*s.keys); k.resize(cap);
  std::vector<unsigned char> dsc((size_t)cap * 32, 0);
  if (n) memcpy(dsc.data(), s.desc, (size_t)n * 32);
  bool ok = up(&d.kps, k.data(), (size_t)cap) && up(&d.desc, dsc.data(), dsc.size()) && up(&d.n, &n, 1) &&
            up((void**)&d.fv.d_node, node.data(), node.size()) && up((void**)&d.fv.d_off, off.data(), off.size()) &&
            up((void**)&d.fv.d_idx, idx.data(), idx.size()) && up((void**)&d.fv.d_nn, &nn, 1);
  if (ok && s.flag) { std::vector<unsigned char> f(*s.flag); f.resize(cap, 0); ok = up(&d.flag, f.data(), f.size()); }
  if (ok && s.uRight) { std::vector<float> u(*s.uRight); u.resize(cap, -1.f); ok = up(&d.ur, u.data(), u.size()); }
  return ok;
}
inline int strides(const BowSide& a, const BowSide& b, int* node_stride) {
  size_t cap = a.keys->size() > b.keys->size() ? a.keys->size() : b.keys->size();
  size_t ns = a.fv->size() > b.fv->size() ? a.fv->size() : b.fv->size();
  *node_stride = (int)(ns ? ns : 1);
  return (int)(cap ? cap : 1);
}
}  // namespace bowdetail

// mode = ORBM_BOW_KF_FRAME (side1 keyframe, side2 frame; match21 = vpMapPointMatches as keyframe feature indices)
// or ORBM_BOW_KF_KF (match12 = vpMatches12 as side-2 feature indices).  Returns nmatches, or a negative ORBX_E_*.
inline int SearchByBoWDevice(int mode, const BowSide& s1, const BowSide& s2, float nnratio, bool checkOri,
                             std::vector<int>& match12, std::vector<int>& match21) {
  int ns = 1;
  const int cap = bowdetail::strides(s1, s2, &ns);
  bowdetail::DevSide d1, d2;
  if (!bowdetail::upload(s1, cap, ns, d1) || !bowdetail::upload(s2, cap, ns, d2)) return ORBX_E_CUDA;
  bowdetail::Buf b12, b21, bnm;
  if (!b12.alloc((size_t)cap * 4) || !b21.alloc((size_t)cap * 4) || !bnm.alloc(4)) return ORBX_E_NOMEM;
  int32_t *m12 = (int32_t*)b12.p, *m21 = (int32_t*)b21.p, *nm = (int32_t*)bnm.p;
  int rc = orbm_search_by_bow(mode, (const orbx_keypoint*)d1.kps, (const uint8_t*)d1.desc, (const int32_t*)d1.n,
                              (const uint8_t*)d1.flag, &d1.fv, (const orbx_keypoint*)d2.kps, (const uint8_t*)d2.desc,
                              (const int32_t*)d2.n, (const uint8_t*)d2.flag, &d2.fv, cap, ns, 1, nnratio,
                              checkOri ? 1 : 0, m12, m21, nm, nullptr);
  int nmatches = 0;
  match12.assign(s1.keys->size(), -1);
  match21.assign(s2.keys->size(), -1);
  if (rc == ORBX_OK &&
      (cudaMemcpy(match12.data(), m12, match12.size() * 4, cudaMemcpyDeviceToHost) ||
       cudaMemcpy(match21.data(), m21, match21.size() * 4, cudaMemcpyDeviceToHost) ||
       cudaMemcpy(&nmatches, nm, 4, cudaMemcpyDeviceToHost)))
    rc = ORBX_E_CUDA;
  return rc == ORBX_OK ? nmatches : rc;
}

// F12 = K1^-T [t12]x R12 K2^-1 row-major and ep = pKF2->mpCamera->project(T2w*Cw) are the host's (Eigen) geometry,
// ORBmatcher.cc:913-936 / Pinhole.cpp:109-112; scale2 / sigma2 = pKF2->mvScaleFactors / mvLevelSigma2.
inline int SearchForTriangulationDevice(const BowSide& s1, const BowSide& s2, const float F12[9], const float ep[2],
                                        const std::vector<float>& scale2, const std::vector<float>& sigma2,
                                        bool bOnlyStereo, bool bCoarse, bool checkOri,
                                        std::vector<std::pair<size_t, size_t> >& vMatchedPairs) {
  int ns = 1;
  const int cap = bowdetail::strides(s1, s2, &ns);
  bowdetail::DevSide d1, d2;
  if (!bowdetail::upload(s1, cap, ns, d1) || !bowdetail::upload(s2, cap, ns, d2)) return ORBX_E_CUDA;
  bowdetail::Buf b12, bnm, bF, bep;
  if (!b12.alloc((size_t)cap * 4) || !bnm.alloc(4) || !bF.alloc(36) || !bep.alloc(8)) return ORBX_E_NOMEM;
  int32_t *m12 = (int32_t*)b12.p, *nm = (int32_t*)bnm.p;
  float *dF = (float*)bF.p, *dep = (float*)bep.p;
  if (cudaMemcpy(dF, F12, 36, cudaMemcpyHostToDevice) != cudaSuccess ||
      cudaMemcpy(dep, ep, 8, cudaMemcpyHostToDevice) != cudaSuccess)
    return ORBX_E_CUDA;
  int rc = orbm_search_for_triangulation(
      (const orbx_keypoint*)d1.kps, (const uint8_t*)d1.desc, (const int32_t*)d1.n, (const uint8_t*)d1.flag,
      (const float*)d1.ur, &d1.fv, (const orbx_keypoint*)d2.kps, (const uint8_t*)d2.desc, (const int32_t*)d2.n,
      (const uint8_t*)d2.flag, (const float*)d2.ur, &d2.fv, cap, ns, 1, dF, dep, scale2.data(), sigma2.data(),
      (int)scale2.size(), bOnlyStereo ? 1 : 0, bCoarse ? 1 : 0, checkOri ? 1 : 0, m12, nm, nullptr);
  std::vector<int> v12(s1.keys->size(), -1);
  int nmatches = 0;
  if (rc == ORBX_OK && (cudaMemcpy(v12.data(), m12, v12.size() * 4, cudaMemcpyDeviceToHost) ||
                        cudaMemcpy(&nmatches, nm, 4, cudaMemcpyDeviceToHost)))
    rc = ORBX_E_CUDA;
  vMatchedPairs.clear();  // :1133-1143
  for (size_t i = 0; i < v12.size(); i++)
    if (v12[i] >= 0) vMatchedPairs.push_back(std::make_pair(i, (size_t)v12[i]));
  return rc == ORBX_OK ? nmatches : rc;
}

}  // namespace ORB_SLAM3
#endif
