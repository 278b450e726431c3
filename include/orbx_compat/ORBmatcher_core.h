// ORBmatcher_core.h — the pieces of ORB-SLAM3's ORBmatcher that this library replaces, with the
// reference's signatures (include/ORBmatcher.h:35-105).
//
//  * ORBmatcher::DescriptorDistance(const cv::Mat&, const cv::Mat&) stays a host inline with an
//    identical result (popcount of 256 bits, src/ORBmatcher.cc:2058-2074): it is called one pair at
//    a time from three threads, a launch per pair would be absurd.
//  * The inner "for each candidate: distance, keep best/second" loops of the Search* family are
//    served by the batched kernels of include/orbx.h; OrbmBatch below is the thin host helper a
//    Search* body uses to hand its candidate lists to the GPU and get the reference's
//    bestDist/bestIdx/bestLevel/bestDist2/bestLevel2 back.  The projection / bookkeeping code of
//    each Search* body is unchanged host code (INTEGRATION.md shows SearchByProjection).
#ifndef ORBMATCHER_CORE_H
#define ORBMATCHER_CORE_H

#include <stdint.h>
#include <string.h>

#include <vector>

#include "../orbx.h"
#include "cv_compat.h"

namespace ORB_SLAM3 {

struct ORBmatcherCore {
  static const int TH_HIGH = 100;       // src/ORBmatcher.cc:35
  static const int TH_LOW = 50;         // :36
  static const int HISTO_LENGTH = 30;   // :37

  // Computes the Hamming distance between two ORB descriptors (32 contiguous bytes each).
  static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b) {
    const unsigned char* pa = a.ptr<unsigned char>();
    const unsigned char* pb = b.ptr<unsigned char>();
    int dist = 0;
    for (int i = 0; i < 4; i++) {
      uint64_t x, y;
      memcpy(&x, pa + 8 * i, 8);
      memcpy(&y, pb + 8 * i, 8);
      dist += __builtin_popcountll(x ^ y);
    }
    return dist;
  }
};

// Host-side staging of one batched scoring call (explicit candidate lists).  Usage inside a
// Search* body:   batch.begin_query(desc_row);  for (idx : vIndices) if (!skip) batch.add(idx);
//                 ... batch.run(refDescriptors, refOctaves, ORBM_TIE_FIRST_MIN, results);
class OrbmBatch {
 public:
  void clear() { q_.clear(); off_.clear(); idx_.clear(); }
  void begin_query(const unsigned char* desc32) {
    off_.push_back((int32_t)idx_.size());  // start of this query's candidate list
    q_.insert(q_.end(), desc32, desc32 + 32);
  }
  void add(int ref_index) { idx_.push_back(ref_index); }
  int queries() const { return (int)(q_.size() / 32); }

  // ref: nr x 32 descriptor bytes; ref_level: nr octaves (may be null).  Synchronous.
  int run(const unsigned char* ref, int nr, const int8_t* ref_level, int tie_rule, std::vector<orbm_result>& out);

 private:
  std::vector<unsigned char> q_;
  std::vector<int32_t> off_, idx_;
};

}  // namespace ORB_SLAM3

// ---- implementation (needs the CUDA runtime only for memory management) -----------------------
#ifdef ORBM_BATCH_IMPLEMENTATION
#include <cuda_runtime.h>
namespace ORB_SLAM3 {
inline int OrbmBatch::run(const unsigned char* ref, int nr, const int8_t* ref_level, int tie_rule,
                          std::vector<orbm_result>& out) {
  const int nq = queries();
  out.resize(nq);
  if (nq == 0) return ORBX_OK;
  std::vector<int32_t> off(off_);
  off.resize(nq + 1);
  off[nq] = (int32_t)idx_.size();
  unsigned char *dq = nullptr, *dr = nullptr;
  int32_t *doff = nullptr, *didx = nullptr;
  int8_t* dlv = nullptr;
  orbm_result* dout = nullptr;
  int rc = ORBX_E_CUDA;
  if (cudaMalloc(&dq, q_.size()) || cudaMalloc(&dr, (size_t)nr * 32) || cudaMalloc(&doff, off.size() * 4) ||
      cudaMalloc(&didx, idx_.size() * 4 + 4) || cudaMalloc(&dout, (size_t)nq * sizeof(orbm_result)) ||
      (ref_level && cudaMalloc(&dlv, nr)))
    goto done;
  cudaMemcpy(dq, q_.data(), q_.size(), cudaMemcpyHostToDevice);
  cudaMemcpy(dr, ref, (size_t)nr * 32, cudaMemcpyHostToDevice);
  cudaMemcpy(doff, off.data(), off.size() * 4, cudaMemcpyHostToDevice);
  if (!idx_.empty()) cudaMemcpy(didx, idx_.data(), idx_.size() * 4, cudaMemcpyHostToDevice);
  if (ref_level) cudaMemcpy(dlv, ref_level, nr, cudaMemcpyHostToDevice);
  rc = orbm_argmin2_csr(dq, nq, dr, nr, doff, didx, dlv, tie_rule, dout, nullptr);
  if (rc == ORBX_OK && cudaMemcpy(out.data(), dout, (size_t)nq * sizeof(orbm_result), cudaMemcpyDeviceToHost))
    rc = ORBX_E_CUDA;
done:
  cudaFree(dq); cudaFree(dr); cudaFree(doff); cudaFree(didx); cudaFree(dlv); cudaFree(dout);
  return rc;
}
}  // namespace ORB_SLAM3
#endif

#endif
