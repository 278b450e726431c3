// ORBmatcher.h — drop-in for the reference's include/ORBmatcher.h:35-105 on liborbx.so.
//
// `ORBmatcherT<Frame, KeyFrame, MapPoint>` has the reference's constructor, constants and Search* signatures
// (ORBmatcher.h:40-87).  It is a template over the three object types because the reference's Frame / KeyFrame /
// MapPoint drag in Eigen, Sophus, DBoW2 and OpenCV; the bodies below only touch the members the reference bodies
// touch, under the reference's own names, so `typedef ORBmatcherT<Frame, KeyFrame, MapPoint> ORBmatcher;` compiles
// against the real classes, and the tests instantiate it with stand-in structs that carry the same members
// (tests/cpp/compat_matcher.cpp), the way cv_compat.h stands in for OpenCV.
//
// Each body keeps the reference's host-side bookkeeping (which MapPoints take part, what is written where) and hands
// the candidate loops — GetFeaturesInArea + DescriptorDistance + best/second + "already matched" — to ONE device call:
//   SearchByProjection(Frame&, vector<MapPoint*>&, th, bFarPoints, thFarPoints)   ORBmatcher.cc:43-213
//        Nleft == -1 -> orbm_search_by_projection;  Nleft != -1 (fisheye rig) -> orbm_search_by_projection_rig
//   SearchByProjection(Frame& Current, const Frame& Last, th, bMono)             ORBmatcher.cc:1676-1887 (Nleft == -1)
//   SearchByBoW(KeyFrame*, Frame&, vector<MapPoint*>&)                           ORBmatcher.cc:223-425
//        Nleft == -1 -> orbm_search_by_bow;  Nleft != -1 -> orbm_search_by_bow_rig
//   SearchByBoW(KeyFrame*, KeyFrame*, vector<MapPoint*>&)                        ORBmatcher.cc:765-905
//   SearchForInitialization(Frame&, Frame&, vbPrevMatched, vnMatches12, window)  ORBmatcher.cc:648-763
//   SearchForTriangulation(KeyFrame*, KeyFrame*, vMatchedPairs, bOnlyStereo, bCoarse)   ORBmatcher.cc:907-1146 (pinhole)
//   SearchByProjection(Frame&, KeyFrame*, sAlreadyFound, th, ORBdist)            ORBmatcher.cc:1889-2010 (relocalisation)
//   SearchByProjection(KeyFrame*, Sim3&, vpPoints, vpMatched, th, ratioHamming)  ORBmatcher.cc:427-532
//   SearchByProjection(KeyFrame*, Sim3&, vpPoints, vpPointsKFs, vpMatched, vpMatchedKF, th, ratioHamming)   :534-646
//        all three -> orbm_search_by_projection_th (ORBdist / floor(TH_LOW*ratioHamming) as the threshold)
//   Fuse(KeyFrame*, vpMapPoints, th, bRight), Fuse(KeyFrame*, Sim3&, vpPoints, th, vpReplacePoint)   :1148-1338, :1340-1455
//   SearchBySim3(KeyFrame*, KeyFrame*, vpMatches12, S12, th)                     ORBmatcher.cc:1457-1674
//        -> orbm_fuse_candidates for the scoring; the Replace / AddObservation / AddMapPoint calls and the two-way
//           agreement run on the host exactly as the reference orders them
// i.e. every method of include/ORBmatcher.h:40-87.  The Sim3 argument is a template parameter (Sophus::Sim3f in the
// reference: rotationMatrix(), translation(), scale(), inverse(), operator*).
// Members read (all reference names): Frame: N, Nleft, mvKeys, mvKeysRight, mvKeysUn, mDescriptors (.data, 32-byte
// rows), mvuRight, mvpMapPoints, mvbOutlier, mvScaleFactors, mnMinX/mnMaxX/mnMinY/mnMaxY, mFeatVec, mvLeftToRightMatch,
// mvRightToLeftMatch, mb, mbf, mpCamera->project(), GetPose();  KeyFrame: N, NLeft, mvKeys, mvKeysRight, mvKeysUn,
// mDescriptors, mFeatVec, mvuRight, mvScaleFactors, mvLevelSigma2, mvInvLevelSigma2, mnMinX.., fx, fy, cx, cy, mbf,
// mpCamera(2), GetPose(), GetRightPose(), GetCameraCenter(), GetRightCameraCenter(), IsInImage(), GetMapPointMatches(),
// GetMapPoint(i), GetMapPoints(), AddMapPoint();  MapPoint: mbTrackInView(R), mTrackProjX/Y/XR/YR, mTrackViewCos(R),
// mnTrackScaleLevel(R), mTrackDepth, isBad(), Observations(), GetDescriptor(), GetWorldPos(), GetNormal(),
// GetMin/MaxDistanceInvariance(), PredictScale(), IsInKeyFrame(), GetIndexInKeyFrame(), AddObservation(), Replace().
// Synchronous: every call uploads what it needs and downloads the result (a resident pipeline calls orbm_* directly).
#ifndef ORBX_COMPAT_ORBMATCHER_H
#define ORBX_COMPAT_ORBMATCHER_H

#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <string.h>

#include <set>
#include <utility>
#include <vector>

#include "../orbx.h"
#include "BowMatcher.h"
#include "ORBmatcher_core.h"
#include "cv_compat.h"

namespace ORB_SLAM3 {

namespace orbxm {
// device scratch that frees itself on every path
struct Dev {
  void* p = nullptr;
  Dev() {}
  Dev(const Dev&) = delete;
  Dev& operator=(const Dev&) = delete;
  ~Dev() { cudaFree(p); }
  bool alloc(size_t bytes) { return cudaMalloc(&p, bytes ? bytes : 1) == cudaSuccess; }
  template <class T> bool put(const std::vector<T>& h) {
    return alloc(h.size() * sizeof(T)) && (h.empty() || cudaMemcpy(p, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice) == cudaSuccess);
  }
  template <class T> bool get(std::vector<T>& h) const {
    return h.empty() || cudaMemcpy(h.data(), p, h.size() * sizeof(T), cudaMemcpyDeviceToHost) == cudaSuccess;
  }
  template <class T> T* as() const { return (T*)p; }
};

// the window queries of a SearchByProjection call, in MapPoint order
struct Queries {
  std::vector<float> qx, qy, qr, qur, angle, qxr, qyr, qrr;
  std::vector<int8_t> qmin, qmax, qminr, qmaxr;
  std::vector<unsigned char> desc, flags;
  void resize(size_t n) {
    qx.assign(n, 0.f); qy.assign(n, 0.f); qr.assign(n, 0.f); qur.assign(n, 0.f); angle.assign(n, 0.f);
    qxr.assign(n, 0.f); qyr.assign(n, 0.f); qrr.assign(n, 0.f);
    qmin.assign(n, 0); qmax.assign(n, 0); qminr.assign(n, 0); qmaxr.assign(n, 0);
    desc.assign(n * 32, 0); flags.assign(n, 0);
  }
  void resize_keep(size_t n) {  // truncate, keeping the first n queries
    qx.resize(n); qy.resize(n); qr.resize(n); qur.resize(n); angle.resize(n); qxr.resize(n); qyr.resize(n); qrr.resize(n);
    qmin.resize(n); qmax.resize(n); qminr.resize(n); qmaxr.resize(n); desc.resize(n * 32); flags.resize(n);
  }
};

struct DevCamera {  // one camera's features + grid on the device
  Dev kps, desc, n, cell_off, cell_idx, locked, partner;
  int count = 0;
  bool upload(const cv::KeyPoint* k, const unsigned char* d, int cnt, float mnMinX, float mnMaxX, float mnMinY,
              float mnMaxY, const std::vector<unsigned char>* lock, const std::vector<int>* pair) {
    count = cnt;
    const int cap = cnt > 0 ? cnt : 1;
    std::vector<cv::KeyPoint> kk(k, k + cnt); kk.resize(cap);
    std::vector<unsigned char> dd(d, d + (size_t)cnt * 32); dd.resize((size_t)cap * 32, 0);
    std::vector<int32_t> nn(1, cnt);
    if (!kps.put(kk) || !desc.put(dd) || !n.put(nn) || !cell_off.alloc(sizeof(int32_t) * (64 * 48 + 1)) ||
        !cell_idx.alloc(sizeof(int32_t) * cap))
      return false;
    if (lock) { std::vector<unsigned char> l(*lock); l.resize(cap, 0); if (!locked.put(l)) return false; }
    if (pair) { std::vector<int32_t> pr(pair->begin(), pair->end()); pr.resize(cap, -1); if (!partner.put(pr)) return false; }
    return orbm_grid_build(kps.as<orbx_keypoint>(), n.as<int32_t>(), cap, 1, mnMinX, mnMaxX, mnMinY, mnMaxY,
                           cell_off.as<int32_t>(), cell_idx.as<int32_t>(), nullptr) == ORBX_OK;
  }
  int cap() const { return count > 0 ? count : 1; }
};
}  // namespace orbxm

// Customisation point of SearchForTriangulation: F12 = K1^-T [t12]x R12 K2^-1 (row-major) and the epipole of KF1's
// centre in KF2 (ORBmatcher.cc:913-936, Pinhole.cpp:107-113) are the host's Eigen geometry; provide an overload for
// the real KeyFrame (INTEGRATION.md shows it) — found by argument-dependent lookup.
//   void orbx_epipolar_geometry(KeyFrame* pKF1, KeyFrame* pKF2, float F12[9], float ep[2]);

template <class FrameT, class KeyFrameT, class MapPointT>
class ORBmatcherT {
 public:
  static const int TH_LOW = 50, TH_HIGH = 100, HISTO_LENGTH = 30;  // ORBmatcher.cc:35-37

  ORBmatcherT(float nnratio = 0.6, bool checkOri = true) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}

  static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b) { return ORBmatcherCore::DescriptorDistance(a, b); }

  // ---- SearchByProjection(Frame&, const vector<MapPoint*>&, th, bFarPoints, thFarPoints)  ORBmatcher.cc:43-213 ----
  int SearchByProjection(FrameT& F, const std::vector<MapPointT*>& vpMapPoints, const float th = 3,
                         const bool bFarPoints = false, const float thFarPoints = 50.0f) {
    const bool bFactor = th != 1.0;
    const bool rig = F.Nleft != -1;
    const size_t M = vpMapPoints.size();
    orbxm::Queries q;
    q.resize(M);
    for (size_t iMP = 0; iMP < M; iMP++) {
      MapPointT* pMP = vpMapPoints[iMP];
      if (!pMP->mbTrackInView && !pMP->mbTrackInViewR) continue;
      if (bFarPoints && pMP->mTrackDepth > thFarPoints) continue;
      if (pMP->isBad()) continue;
      unsigned char fl = pMP->Observations() > 0 ? 2 : 0;
      if (pMP->mbTrackInView) {
        const int nPredictedLevel = pMP->mnTrackScaleLevel;
        float r = RadiusByViewingCos(pMP->mTrackViewCos);
        if (bFactor) r *= th;
        fl |= 1;
        q.qx[iMP] = pMP->mTrackProjX; q.qy[iMP] = pMP->mTrackProjY;
        q.qr[iMP] = r * F.mvScaleFactors[nPredictedLevel];
        q.qmin[iMP] = (int8_t)(nPredictedLevel - 1); q.qmax[iMP] = (int8_t)nPredictedLevel;
        q.qur[iMP] = pMP->mTrackProjXR;
      }
      if (rig && pMP->mbTrackInViewR && pMP->mnTrackScaleLevelR != -1) {
        const int nPredictedLevel = pMP->mnTrackScaleLevelR;
        const float r = RadiusByViewingCos(pMP->mTrackViewCosR);  // no `th` factor on the right camera (:147)
        fl |= 4;
        q.qxr[iMP] = pMP->mTrackProjXR; q.qyr[iMP] = pMP->mTrackProjYR;
        q.qrr[iMP] = r * F.mvScaleFactors[nPredictedLevel];
        q.qminr[iMP] = (int8_t)(nPredictedLevel - 1); q.qmaxr[iMP] = (int8_t)nPredictedLevel;
      }
      q.flags[iMP] = fl;
      const cv::Mat d = pMP->GetDescriptor();
      memcpy(&q.desc[iMP * 32], d.data, 32);
    }
    std::vector<unsigned char> locked(F.N, 0);
    for (int j = 0; j < F.N; j++) locked[j] = F.mvpMapPoints[j] && F.mvpMapPoints[j]->Observations() > 0;
    return rig ? ProjectRig(F, vpMapPoints, q, locked) : ProjectMono(F, vpMapPoints, q, locked);
  }

  // ---- SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, th, bMono)  ORBmatcher.cc:1676-1887 ----
  int SearchByProjection(FrameT& CurrentFrame, const FrameT& LastFrame, const float th, const bool bMono) {
    if (CurrentFrame.Nleft != -1) return ORBX_E_INVALID;  // rig: orbm_search_by_projection_rig (ORBM_PROJ_LASTFRAME)
    const auto Tcw = CurrentFrame.GetPose();
    const auto twc = Tcw.inverse().translation();
    const auto Tlw = LastFrame.GetPose();
    const auto tlc = Tlw * twc;
    const bool bForward = tlc(2) > CurrentFrame.mb && !bMono;
    const bool bBackward = -tlc(2) > CurrentFrame.mb && !bMono;
    std::vector<MapPointT*> pts;
    orbxm::Queries q;
    q.resize(LastFrame.N);
    size_t m = 0;
    for (int i = 0; i < LastFrame.N; i++) {
      MapPointT* pMP = LastFrame.mvpMapPoints[i];
      if (!pMP || LastFrame.mvbOutlier[i]) continue;
      const auto x3Dw = pMP->GetWorldPos();
      const auto x3Dc = Tcw * x3Dw;
      const float invzc = 1.0 / x3Dc(2);
      if (invzc < 0) continue;
      const auto uv = CurrentFrame.mpCamera->project(x3Dc);
      if (uv(0) < CurrentFrame.mnMinX || uv(0) > CurrentFrame.mnMaxX) continue;
      if (uv(1) < CurrentFrame.mnMinY || uv(1) > CurrentFrame.mnMaxY) continue;
      const int nLastOctave = LastFrame.mvKeys[i].octave;
      const float radius = th * CurrentFrame.mvScaleFactors[nLastOctave];
      q.qx[m] = uv(0); q.qy[m] = uv(1); q.qr[m] = radius;
      if (bForward) { q.qmin[m] = (int8_t)nLastOctave; q.qmax[m] = -1; }                  // GetFeaturesInArea(u,v,r,oct)
      else if (bBackward) { q.qmin[m] = 0; q.qmax[m] = (int8_t)nLastOctave; }
      else { q.qmin[m] = (int8_t)(nLastOctave - 1); q.qmax[m] = (int8_t)(nLastOctave + 1); }
      q.qur[m] = uv(0) - CurrentFrame.mbf * invzc;
      q.angle[m] = LastFrame.mvKeysUn[i].angle;
      q.flags[m] = (unsigned char)(1 | (pMP->Observations() > 0 ? 2 : 0));
      const cv::Mat d = pMP->GetDescriptor();
      memcpy(&q.desc[m * 32], d.data, 32);
      pts.push_back(pMP);
      m++;
    }
    q.resize_keep(m);
    std::vector<unsigned char> locked(CurrentFrame.N, 0);
    for (int j = 0; j < CurrentFrame.N; j++)
      locked[j] = CurrentFrame.mvpMapPoints[j] && CurrentFrame.mvpMapPoints[j]->Observations() > 0;
    std::vector<int> match, feat_query;
    int nmatches = 0;
    const int rc = RunProjection(CurrentFrame, q, m, locked, ORBM_PROJ_LASTFRAME, 0 /* rotation check below */, match,
                                 feat_query, nmatches);
    if (rc != ORBX_OK) return rc;
    return CommitWithRotation(CurrentFrame, pts, q, m, match);
  }

  // ---- SearchByBoW(KeyFrame*, Frame&, vector<MapPoint*>&)  ORBmatcher.cc:223-425 ----
  int SearchByBoW(KeyFrameT* pKF, FrameT& F, std::vector<MapPointT*>& vpMapPointMatches) {
    const std::vector<MapPointT*> vpMapPointsKF = pKF->GetMapPointMatches();
    vpMapPointMatches = std::vector<MapPointT*>(F.N, static_cast<MapPointT*>(NULL));
    std::vector<unsigned char> valid(vpMapPointsKF.size(), 0);
    for (size_t i = 0; i < vpMapPointsKF.size(); i++) valid[i] = vpMapPointsKF[i] && !vpMapPointsKF[i]->isBad();
    BowSide kf{&pKF->mvKeysUn, pKF->mDescriptors.data, &pKF->mFeatVec, &valid, nullptr};
    std::vector<int> m12, m21;
    int n;
    if (F.Nleft == -1) {
      BowSide fr{&F.mvKeys, F.mDescriptors.data, &F.mFeatVec, nullptr, nullptr};
      n = SearchByBoWDevice(ORBM_BOW_KF_FRAME, kf, fr, mfNNratio, mbCheckOrientation, m12, m21);
    } else {
      // the frame's features in the reference's rig numbering: mvKeys then mvKeysRight (:342-345, :371-374)
      std::vector<cv::KeyPoint> keys(F.mvKeys.begin(), F.mvKeys.begin() + F.Nleft);
      keys.insert(keys.end(), F.mvKeysRight.begin(), F.mvKeysRight.end());
      BowSide fr{&keys, F.mDescriptors.data, &F.mFeatVec, nullptr, nullptr};
      n = SearchByBoWRig(kf, fr, F.Nleft, m21);
    }
    if (n < 0) return n;
    for (int j = 0; j < F.N; j++)
      if (m21[j] >= 0) vpMapPointMatches[j] = vpMapPointsKF[m21[j]];
    return n;
  }

  // ---- SearchByBoW(KeyFrame*, KeyFrame*, vector<MapPoint*>&)  ORBmatcher.cc:765-905 ----
  int SearchByBoW(KeyFrameT* pKF1, KeyFrameT* pKF2, std::vector<MapPointT*>& vpMatches12) {
    const std::vector<MapPointT*> vpMapPoints1 = pKF1->GetMapPointMatches();
    const std::vector<MapPointT*> vpMapPoints2 = pKF2->GetMapPointMatches();
    vpMatches12 = std::vector<MapPointT*>(vpMapPoints1.size(), static_cast<MapPointT*>(NULL));
    std::vector<unsigned char> v1(vpMapPoints1.size(), 0), v2(vpMapPoints2.size(), 0);
    for (size_t i = 0; i < v1.size(); i++) v1[i] = vpMapPoints1[i] && !vpMapPoints1[i]->isBad();
    for (size_t i = 0; i < v2.size(); i++) v2[i] = vpMapPoints2[i] && !vpMapPoints2[i]->isBad();
    BowSide s1{&pKF1->mvKeysUn, pKF1->mDescriptors.data, &pKF1->mFeatVec, &v1, nullptr};
    BowSide s2{&pKF2->mvKeysUn, pKF2->mDescriptors.data, &pKF2->mFeatVec, &v2, nullptr};
    std::vector<int> m12, m21;
    const int n = SearchByBoWDevice(ORBM_BOW_KF_KF, s1, s2, mfNNratio, mbCheckOrientation, m12, m21);
    if (n < 0) return n;
    for (size_t i = 0; i < m12.size(); i++)
      if (m12[i] >= 0) vpMatches12[i] = vpMapPoints2[m12[i]];
    return n;
  }

  // ---- SearchForInitialization(Frame&, Frame&, vbPrevMatched, vnMatches12, windowSize)  ORBmatcher.cc:648-763 ----
  int SearchForInitialization(FrameT& F1, FrameT& F2, std::vector<cv::Point2f>& vbPrevMatched,
                              std::vector<int>& vnMatches12, int windowSize = 10) {
    vnMatches12 = std::vector<int>(F1.mvKeysUn.size(), -1);
    const int cap = (int)(F1.mvKeysUn.size() > F2.mvKeysUn.size() ? F1.mvKeysUn.size() : F2.mvKeysUn.size());
    if (cap == 0) return 0;
    orbxm::DevCamera c2;
    std::vector<cv::KeyPoint> k2(F2.mvKeysUn); k2.resize(cap);
    std::vector<unsigned char> d2((size_t)cap * 32, 0);
    memcpy(d2.data(), F2.mDescriptors.data, F2.mvKeysUn.size() * 32);
    if (!c2.upload(k2.data(), d2.data(), cap, F2.mnMinX, F2.mnMaxX, F2.mnMinY, F2.mnMaxY, nullptr, nullptr)) return ORBX_E_CUDA;
    std::vector<int32_t> n2(1, (int32_t)F2.mvKeysUn.size()), n1(1, (int32_t)F1.mvKeysUn.size());
    orbxm::Dev dn2, dn1, dk1, dd1, dprev, dm12, dscr, dnm;
    std::vector<cv::KeyPoint> k1(F1.mvKeysUn); k1.resize(cap);
    std::vector<unsigned char> d1((size_t)cap * 32, 0);
    memcpy(d1.data(), F1.mDescriptors.data, F1.mvKeysUn.size() * 32);
    std::vector<float> prev((size_t)cap * 2, 0.f);
    for (size_t i = 0; i < vbPrevMatched.size() && i < (size_t)cap; i++) { prev[2 * i] = vbPrevMatched[i].x; prev[2 * i + 1] = vbPrevMatched[i].y; }
    if (!dn2.put(n2) || !dn1.put(n1) || !dk1.put(k1) || !dd1.put(d1) || !dprev.put(prev) || !dm12.alloc((size_t)cap * 4) ||
        !dscr.alloc((size_t)cap * 8) || !dnm.alloc(4))
      return ORBX_E_CUDA;
    // the grid has to be rebuilt with F2's real count (c2 was uploaded with `cap` rows so the strides agree)
    if (orbm_grid_build(c2.kps.as<orbx_keypoint>(), dn2.as<int32_t>(), cap, 1, F2.mnMinX, F2.mnMaxX, F2.mnMinY, F2.mnMaxY,
                        c2.cell_off.as<int32_t>(), c2.cell_idx.as<int32_t>(), nullptr) != ORBX_OK)
      return ORBX_E_CUDA;
    int rc = orbm_search_for_initialization(dk1.as<orbx_keypoint>(), dd1.as<uint8_t>(), dn1.as<int32_t>(),
                                            c2.kps.as<orbx_keypoint>(), c2.desc.as<uint8_t>(), dn2.as<int32_t>(), cap, 1,
                                            c2.cell_off.as<int32_t>(), c2.cell_idx.as<int32_t>(), F2.mnMinX, F2.mnMaxX,
                                            F2.mnMinY, F2.mnMaxY, dprev.as<float>(), windowSize, mfNNratio,
                                            mbCheckOrientation ? 1 : 0, dm12.as<int32_t>(), dscr.as<int32_t>(),
                                            dnm.as<int32_t>(), nullptr);
    if (rc != ORBX_OK) return rc;
    std::vector<int32_t> m12(cap), nm(1);
    if (!dm12.get(m12) || !dnm.get(nm) || !dprev.get(prev)) return ORBX_E_CUDA;
    for (size_t i = 0; i < vnMatches12.size(); i++) vnMatches12[i] = m12[i];
    for (size_t i = 0; i < vbPrevMatched.size() && i < (size_t)cap; i++) { vbPrevMatched[i].x = prev[2 * i]; vbPrevMatched[i].y = prev[2 * i + 1]; }
    return nm[0];
  }

  // ---- SearchForTriangulation(KeyFrame*, KeyFrame*, vMatchedPairs, bOnlyStereo, bCoarse)  ORBmatcher.cc:907-1146 ----
  int SearchForTriangulation(KeyFrameT* pKF1, KeyFrameT* pKF2, std::vector<std::pair<size_t, size_t> >& vMatchedPairs,
                             const bool bOnlyStereo, const bool bCoarse = false) {
    float F12[9], ep[2];
    orbx_epipolar_geometry(pKF1, pKF2, F12, ep);  // host geometry (see the customisation point above)
    std::vector<unsigned char> has1(pKF1->N, 0), has2(pKF2->N, 0);
    for (int i = 0; i < pKF1->N; i++) has1[i] = pKF1->GetMapPoint(i) != NULL;
    for (int i = 0; i < pKF2->N; i++) has2[i] = pKF2->GetMapPoint(i) != NULL;
    BowSide s1{&pKF1->mvKeysUn, pKF1->mDescriptors.data, &pKF1->mFeatVec, &has1, &pKF1->mvuRight};
    BowSide s2{&pKF2->mvKeysUn, pKF2->mDescriptors.data, &pKF2->mFeatVec, &has2, &pKF2->mvuRight};
    return SearchForTriangulationDevice(s1, s2, F12, ep, pKF2->mvScaleFactors, pKF2->mvLevelSigma2, bOnlyStereo, bCoarse,
                                        mbCheckOrientation, vMatchedPairs);
  }


  // ---- SearchByProjection(Frame &CurrentFrame, KeyFrame *pKF, const set<MapPoint*> &sAlreadyFound, th, ORBdist)
  //      ORBmatcher.cc:1889-2010 (relocalisation) ----
  int SearchByProjection(FrameT& CurrentFrame, KeyFrameT* pKF, const std::set<MapPointT*>& sAlreadyFound, const float th,
                         const int ORBdist) {
    const auto Tcw = CurrentFrame.GetPose();
    const auto Ow = Tcw.inverse().translation();
    const std::vector<MapPointT*> vpMPs = pKF->GetMapPointMatches();
    std::vector<MapPointT*> pts;
    orbxm::Queries q;
    q.resize(vpMPs.size());
    size_t m = 0;
    for (size_t i = 0, iend = vpMPs.size(); i < iend; i++) {
      MapPointT* pMP = vpMPs[i];
      if (!pMP || pMP->isBad() || sAlreadyFound.count(pMP)) continue;
      const auto x3Dw = pMP->GetWorldPos();
      const auto x3Dc = Tcw * x3Dw;
      const auto uv = CurrentFrame.mpCamera->project(x3Dc);
      if (uv(0) < CurrentFrame.mnMinX || uv(0) > CurrentFrame.mnMaxX) continue;
      if (uv(1) < CurrentFrame.mnMinY || uv(1) > CurrentFrame.mnMaxY) continue;
      const auto PO = x3Dw - Ow;
      float dist3D = PO.norm();
      const float maxDistance = pMP->GetMaxDistanceInvariance();
      const float minDistance = pMP->GetMinDistanceInvariance();
      if (dist3D < minDistance || dist3D > maxDistance) continue;
      const int nPredictedLevel = pMP->PredictScale(dist3D, &CurrentFrame);
      q.qx[m] = uv(0); q.qy[m] = uv(1); q.qr[m] = th * CurrentFrame.mvScaleFactors[nPredictedLevel];
      q.qmin[m] = (int8_t)(nPredictedLevel - 1); q.qmax[m] = (int8_t)(nPredictedLevel + 1);
      q.angle[m] = pKF->mvKeysUn[i].angle;
      q.flags[m] = 3;
      const cv::Mat d = pMP->GetDescriptor();
      memcpy(&q.desc[m * 32], d.data, 32);
      pts.push_back(pMP);
      m++;
    }
    q.resize_keep(m);
    std::vector<unsigned char> locked(CurrentFrame.N, 0);  // any non-null entry blocks its feature (:1951-1952)
    for (int j = 0; j < CurrentFrame.N; j++) locked[j] = CurrentFrame.mvpMapPoints[j] != NULL;
    std::vector<int> match, feat_query;
    int nmatches = 0;
    std::vector<float> no_uright;
    const int rc = RunProjection(CurrentFrame, q, m, locked, ORBM_PROJ_LASTFRAME, 0, match, feat_query, nmatches, ORBdist,
                                 /*stereo_gate=*/false);
    if (rc != ORBX_OK) return rc;
    return CommitWithRotation(CurrentFrame, pts, q, m, match);
  }

  // ---- SearchByProjection(KeyFrame* pKF, Sim3f &Scw, vpPoints, vpMatched, th, ratioHamming)  ORBmatcher.cc:427-532 ----
  template <class Sim3T>
  int SearchByProjection(KeyFrameT* pKF, Sim3T& Scw, const std::vector<MapPointT*>& vpPoints,
                         std::vector<MapPointT*>& vpMatched, int th, float ratioHamming = 1.0) {
    std::vector<KeyFrameT*> none;
    return Sim3Projection(pKF, Scw, vpPoints, nullptr, vpMatched, nullptr, th, ratioHamming);
  }

  // ---- ... with vpPointsKFs / vpMatchedKF  ORBmatcher.cc:534-646 ----
  template <class Sim3T>
  int SearchByProjection(KeyFrameT* pKF, Sim3T& Scw, const std::vector<MapPointT*>& vpPoints,
                         const std::vector<KeyFrameT*>& vpPointsKFs, std::vector<MapPointT*>& vpMatched,
                         std::vector<KeyFrameT*>& vpMatchedKF, int th, float ratioHamming = 1.0) {
    return Sim3Projection(pKF, Scw, vpPoints, &vpPointsKFs, vpMatched, &vpMatchedKF, th, ratioHamming);
  }

  // ---- Fuse(KeyFrame *pKF, const vector<MapPoint*> &vpMapPoints, th, bRight)  ORBmatcher.cc:1148-1338 ----
  int Fuse(KeyFrameT* pKF, const std::vector<MapPointT*>& vpMapPoints, const float th = 3.0, const bool bRight = false) {
    const auto Tcw = bRight ? pKF->GetRightPose() : pKF->GetPose();
    const auto Ow = bRight ? pKF->GetRightCameraCenter() : pKF->GetCameraCenter();
    auto* pCamera = bRight ? pKF->mpCamera2 : pKF->mpCamera;
    const float bf = pKF->mbf;
    const int nMPs = (int)vpMapPoints.size();
    std::vector<int> who;
    orbxm::Queries q;
    q.resize(nMPs);
    size_t m = 0;
    for (int i = 0; i < nMPs; i++) {
      MapPointT* pMP = vpMapPoints[i];
      if (!pMP || pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;
      const auto p3Dw = pMP->GetWorldPos();
      const auto p3Dc = Tcw * p3Dw;
      if (p3Dc(2) < 0.0f) continue;
      const float invz = 1 / p3Dc(2);
      const auto uv = pCamera->project(p3Dc);
      if (!pKF->IsInImage(uv(0), uv(1))) continue;
      const float ur = uv(0) - bf * invz;
      const float maxDistance = pMP->GetMaxDistanceInvariance();
      const float minDistance = pMP->GetMinDistanceInvariance();
      const auto PO = p3Dw - Ow;
      const float dist3D = PO.norm();
      if (dist3D < minDistance || dist3D > maxDistance) continue;
      const auto Pn = pMP->GetNormal();
      if (PO.dot(Pn) < 0.5 * dist3D) continue;
      const int nPredictedLevel = pMP->PredictScale(dist3D, pKF);
      q.qx[m] = uv(0); q.qy[m] = uv(1); q.qr[m] = th * pKF->mvScaleFactors[nPredictedLevel];
      q.qur[m] = ur; q.qmin[m] = (int8_t)nPredictedLevel;
      const cv::Mat d = pMP->GetDescriptor();
      memcpy(&q.desc[m * 32], d.data, 32);
      who.push_back(i);
      m++;
    }
    q.resize_keep(m);
    std::vector<int> best_idx, best_dist;
    const int rc = RunFuse(pKF, bRight, q, m, /*chi_gate=*/true, best_idx, best_dist);
    if (rc != ORBX_OK) return rc;
    int nFused = 0;
    for (size_t k = 0; k < m; k++) {  // :1309-1331, in MapPoint order
      if (best_idx[k] < 0 || best_dist[k] > TH_LOW) continue;
      MapPointT* pMP = vpMapPoints[who[k]];
      const int bestIdx = best_idx[k] + (bRight ? pKF->NLeft : 0);
      MapPointT* pMPinKF = pKF->GetMapPoint(bestIdx);
      if (pMPinKF) {
        if (!pMPinKF->isBad()) {
          if (pMPinKF->Observations() > pMP->Observations()) pMP->Replace(pMPinKF);
          else pMPinKF->Replace(pMP);
        }
      } else {
        pMP->AddObservation(pKF, bestIdx);
        pKF->AddMapPoint(pMP, bestIdx);
      }
      nFused++;
    }
    return nFused;
  }

  // ---- Fuse(KeyFrame *pKF, Sim3f &Scw, vpPoints, th, vpReplacePoint)  ORBmatcher.cc:1340-1455 ----
  template <class Sim3T>
  int Fuse(KeyFrameT* pKF, Sim3T& Scw, const std::vector<MapPointT*>& vpPoints, float th,
           std::vector<MapPointT*>& vpReplacePoint) {
    decltype(pKF->GetPose()) Tcw(Scw.rotationMatrix(), Scw.translation() / Scw.scale());
    const auto Ow = Tcw.inverse().translation();
    const std::set<MapPointT*> spAlreadyFound = pKF->GetMapPoints();
    const int nPoints = (int)vpPoints.size();
    std::vector<int> who;
    orbxm::Queries q;
    q.resize(nPoints);
    size_t m = 0;
    for (int iMP = 0; iMP < nPoints; iMP++) {
      MapPointT* pMP = vpPoints[iMP];
      if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;
      if (!Sim3Query(pKF, Tcw, Ow, pMP, th, q, m)) continue;
      q.qmin[m] = q.qmax[m];  // RunFuse takes the predicted level itself (band [l-1, l] is applied by the kernel)
      who.push_back(iMP);
      m++;
    }
    q.resize_keep(m);
    std::vector<int> best_idx, best_dist;
    const int rc = RunFuse(pKF, false, q, m, /*chi_gate=*/false, best_idx, best_dist);
    if (rc != ORBX_OK) return rc;
    int nFused = 0;
    for (size_t k = 0; k < m; k++) {  // :1435-1449
      if (best_idx[k] < 0 || best_dist[k] > TH_LOW) continue;
      MapPointT* pMP = vpPoints[who[k]];
      MapPointT* pMPinKF = pKF->GetMapPoint(best_idx[k]);
      if (pMPinKF) {
        if (!pMPinKF->isBad()) vpReplacePoint[who[k]] = pMPinKF;
      } else {
        pMP->AddObservation(pKF, best_idx[k]);
        pKF->AddMapPoint(pMP, best_idx[k]);
      }
      nFused++;
    }
    return nFused;
  }

  // ---- SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, vpMatches12, S12, th)  ORBmatcher.cc:1457-1674 ----
  template <class Sim3T>
  int SearchBySim3(KeyFrameT* pKF1, KeyFrameT* pKF2, std::vector<MapPointT*>& vpMatches12, const Sim3T& S12, const float th) {
    const float fx = pKF1->fx, fy = pKF1->fy, cx = pKF1->cx, cy = pKF1->cy;  // (the reference uses pKF1's for both)
    const auto T1w = pKF1->GetPose();
    const auto T2w = pKF2->GetPose();
    const auto S21 = S12.inverse();
    const std::vector<MapPointT*> vpMapPoints1 = pKF1->GetMapPointMatches();
    const std::vector<MapPointT*> vpMapPoints2 = pKF2->GetMapPointMatches();
    const int N1 = (int)vpMapPoints1.size(), N2 = (int)vpMapPoints2.size();
    std::vector<bool> vbAlreadyMatched1(N1, false), vbAlreadyMatched2(N2, false);
    for (int i = 0; i < N1; i++) {
      MapPointT* pMP = vpMatches12[i];
      if (pMP) {
        vbAlreadyMatched1[i] = true;
        const int idx2 = std::get<0>(pMP->GetIndexInKeyFrame(pKF2));
        if (idx2 >= 0 && idx2 < N2) vbAlreadyMatched2[idx2] = true;
      }
    }
    std::vector<int> vnMatch1(N1, -1), vnMatch2(N2, -1);
    for (int dir = 0; dir < 2; dir++) {  // :1496-1578 KF1 -> KF2, :1580-1660 KF2 -> KF1
      const std::vector<MapPointT*>& src = dir == 0 ? vpMapPoints1 : vpMapPoints2;
      const std::vector<bool>& done = dir == 0 ? vbAlreadyMatched1 : vbAlreadyMatched2;
      KeyFrameT* dst = dir == 0 ? pKF2 : pKF1;
      std::vector<int> who;
      orbxm::Queries q;
      q.resize(src.size());
      size_t m = 0;
      for (size_t i = 0; i < src.size(); i++) {
        MapPointT* pMP = src[i];
        if (!pMP || done[i] || pMP->isBad()) continue;
        const auto p3Dw = pMP->GetWorldPos();
        const auto pa = (dir == 0 ? T1w : T2w) * p3Dw;
        const auto pb = dir == 0 ? S21 * pa : S12 * pa;
        if (pb(2) < 0.0) continue;
        const float invz = 1.0 / pb(2);
        const float x = pb(0) * invz, y = pb(1) * invz;
        const float u = fx * x + cx, v = fy * y + cy;
        if (!dst->IsInImage(u, v)) continue;
        const float maxDistance = pMP->GetMaxDistanceInvariance(), minDistance = pMP->GetMinDistanceInvariance();
        const float dist3D = pb.norm();
        if (dist3D < minDistance || dist3D > maxDistance) continue;
        const int nPredictedLevel = pMP->PredictScale(dist3D, dst);
        q.qx[m] = u; q.qy[m] = v; q.qr[m] = th * dst->mvScaleFactors[nPredictedLevel]; q.qmin[m] = (int8_t)nPredictedLevel;
        const cv::Mat d = pMP->GetDescriptor();
        memcpy(&q.desc[m * 32], d.data, 32);
        who.push_back((int)i);
        m++;
      }
      q.resize_keep(m);
      std::vector<int> best_idx, best_dist;
      const int rc = RunFuse(dst, false, q, m, /*chi_gate=*/false, best_idx, best_dist);
      if (rc != ORBX_OK) return rc;
      std::vector<int>& out = dir == 0 ? vnMatch1 : vnMatch2;
      for (size_t k = 0; k < m; k++)
        if (best_idx[k] >= 0 && best_dist[k] <= TH_HIGH) out[who[k]] = best_idx[k];
    }
    int nFound = 0;  // :1662-1677 agreement
    for (int i1 = 0; i1 < N1; i1++) {
      const int idx2 = vnMatch1[i1];
      if (idx2 >= 0 && vnMatch2[idx2] == i1) { vpMatches12[i1] = vpMapPoints2[idx2]; nFound++; }
    }
    return nFound;
  }

 protected:
  float RadiusByViewingCos(const float& viewCos) {  // ORBmatcher.cc:215-221
    if (viewCos > 0.998) return 2.5;
    else return 4.0;
  }

  void ComputeThreeMaxima(std::vector<int>* histo, const int L, int& ind1, int& ind2, int& ind3) {  // :2012-2053
    int max1 = 0, max2 = 0, max3 = 0;
    for (int i = 0; i < L; i++) {
      const int s = (int)histo[i].size();
      if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
      else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
      else if (s > max3) { max3 = s; ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
  }

  // writes in query order (:1772-1792 / :1961-1979), then the rotation histogram on the host exactly as the reference
  // (:1864-1881 / :1986-2006): the erasure nulls the pointer whoever wrote it last
  int CommitWithRotation(FrameT& CurrentFrame, const std::vector<MapPointT*>& pts, const orbxm::Queries& q, size_t m,
                         const std::vector<int>& match) {
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    int nmatches = 0;
    for (size_t i = 0; i < m; i++) {
      const int bestIdx2 = match[i];
      if (bestIdx2 < 0) continue;
      CurrentFrame.mvpMapPoints[bestIdx2] = pts[i];
      nmatches++;
      if (mbCheckOrientation) {
        float rot = q.angle[i] - CurrentFrame.mvKeysUn[bestIdx2].angle;
        if (rot < 0.0) rot += 360.0f;
        int bin = (int)roundf(rot * factor);
        if (bin == HISTO_LENGTH) bin = 0;
        rotHist[bin].push_back(bestIdx2);
      }
    }
    if (mbCheckOrientation) {
      int ind1 = -1, ind2 = -1, ind3 = -1;
      ComputeThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
      for (int i = 0; i < HISTO_LENGTH; i++) {
        if (i == ind1 || i == ind2 || i == ind3) continue;
        for (size_t j = 0, jend = rotHist[i].size(); j < jend; j++) {
          CurrentFrame.mvpMapPoints[rotHist[i][j]] = static_cast<MapPointT*>(NULL);
          nmatches--;
        }
      }
    }
    return nmatches;
  }

  // one single-camera orbm_search_by_projection_th call on a feature set
  int RunProjectionRaw(const cv::KeyPoint* keys, const unsigned char* descs, int N, float mnMinX, float mnMaxX,
                       float mnMinY, float mnMaxY, const std::vector<float>* u_right, const orbxm::Queries& q, size_t M,
                       const std::vector<unsigned char>& locked, int mode, int check_ori, int th_dist,
                       std::vector<int>& match, std::vector<int>& feat_query, int& nmatches) {
    match.assign(M, -1);
    feat_query.assign(N, -1);
    nmatches = 0;
    if (M == 0 || N == 0) return ORBX_OK;
    orbxm::DevCamera cam;
    if (!cam.upload(keys, descs, N, mnMinX, mnMaxX, mnMinY, mnMaxY, &locked, nullptr)) return ORBX_E_CUDA;
    orbxm::Dev ur, qx, qy, qr, qmin, qmax, qd, nq, qur, fl, ang, scr, dm, dfq, dnm;
    std::vector<int32_t> nqv(1, (int32_t)M);
    if (u_right) {
      std::vector<float> u(u_right->begin(), u_right->end());
      u.resize(cam.cap(), -1.f);
      if (!ur.put(u) || !qur.put(q.qur)) return ORBX_E_CUDA;
    }
    if (!qx.put(q.qx) || !qy.put(q.qy) || !qr.put(q.qr) || !qmin.put(q.qmin) || !qmax.put(q.qmax) || !qd.put(q.desc) ||
        !nq.put(nqv) || !fl.put(q.flags) || !ang.put(q.angle) || !scr.alloc(M * 32) || !dm.alloc(M * 4) ||
        !dfq.alloc((size_t)cam.cap() * 4) || !dnm.alloc(4))
      return ORBX_E_CUDA;
    const int rc = orbm_search_by_projection_th(
        cam.kps.template as<orbx_keypoint>(), cam.desc.template as<uint8_t>(), cam.n.template as<int32_t>(), cam.cap(), 1,
        cam.cell_off.template as<int32_t>(), cam.cell_idx.template as<int32_t>(), mnMinX, mnMaxX, mnMinY, mnMaxY,
        ur.as<float>(), cam.locked.template as<uint8_t>(), qx.as<float>(), qy.as<float>(), qr.as<float>(), qmin.as<int8_t>(),
        qmax.as<int8_t>(), qd.as<uint8_t>(), nq.as<int32_t>(), (int)M, qur.as<float>(), fl.as<uint8_t>(), ang.as<float>(),
        mode, mfNNratio, check_ori, th_dist, scr.p, dm.as<int32_t>(), dfq.as<int32_t>(), dnm.as<int32_t>(), nullptr);
    if (rc != ORBX_OK) return rc;
    std::vector<int32_t> fq(cam.cap()), nm(1), mt(M);
    if (!dm.get(mt) || !dfq.get(fq) || !dnm.get(nm)) return ORBX_E_CUDA;
    for (size_t i = 0; i < M; i++) match[i] = mt[i];
    for (int j = 0; j < N; j++) feat_query[j] = fq[j];
    nmatches = nm[0];
    return ORBX_OK;
  }

  int RunProjection(FrameT& F, const orbxm::Queries& q, size_t M, const std::vector<unsigned char>& locked, int mode,
                    int check_ori, std::vector<int>& match, std::vector<int>& feat_query, int& nmatches,
                    int th_dist = TH_HIGH, bool stereo_gate = true) {
    return RunProjectionRaw(F.mvKeysUn.data(), F.mDescriptors.data, F.N, F.mnMinX, F.mnMaxX, F.mnMinY, F.mnMaxY,
                            stereo_gate ? &F.mvuRight : nullptr, q, M, locked, mode, check_ori, th_dist, match, feat_query,
                            nmatches);
  }

  // the gates of the two Sim3 functions for one MapPoint (:452-492, :1366-1404): fills query slot m, false = skipped
  template <class SE3T, class Vec3T>
  bool Sim3Query(KeyFrameT* pKF, const SE3T& Tcw, const Vec3T& Ow, MapPointT* pMP, float th, orbxm::Queries& q, size_t m) {
    const auto p3Dw = pMP->GetWorldPos();
    const auto p3Dc = Tcw * p3Dw;
    if (p3Dc(2) < 0.0) return false;
    const auto uv = pKF->mpCamera->project(p3Dc);
    if (!pKF->IsInImage(uv(0), uv(1))) return false;
    const float maxDistance = pMP->GetMaxDistanceInvariance();
    const float minDistance = pMP->GetMinDistanceInvariance();
    const auto PO = p3Dw - Ow;
    const float dist = PO.norm();
    if (dist < minDistance || dist > maxDistance) return false;
    const auto Pn = pMP->GetNormal();
    if (PO.dot(Pn) < 0.5 * dist) return false;
    const int nPredictedLevel = pMP->PredictScale(dist, pKF);
    q.qx[m] = uv(0); q.qy[m] = uv(1); q.qr[m] = th * pKF->mvScaleFactors[nPredictedLevel];
    q.qmin[m] = (int8_t)(nPredictedLevel - 1); q.qmax[m] = (int8_t)nPredictedLevel;
    q.flags[m] = 3;
    const cv::Mat d = pMP->GetDescriptor();
    memcpy(&q.desc[m * 32], d.data, 32);
    return true;
  }

  template <class Sim3T>
  int Sim3Projection(KeyFrameT* pKF, Sim3T& Scw, const std::vector<MapPointT*>& vpPoints,
                     const std::vector<KeyFrameT*>* vpPointsKFs, std::vector<MapPointT*>& vpMatched,
                     std::vector<KeyFrameT*>* vpMatchedKF, int th, float ratioHamming) {
    decltype(pKF->GetPose()) Tcw(Scw.rotationMatrix(), Scw.translation() / Scw.scale());
    const auto Ow = Tcw.inverse().translation();
    std::set<MapPointT*> spAlreadyFound(vpMatched.begin(), vpMatched.end());
    spAlreadyFound.erase(static_cast<MapPointT*>(NULL));
    std::vector<int> who;
    orbxm::Queries q;
    q.resize(vpPoints.size());
    size_t m = 0;
    for (int iMP = 0, iendMP = (int)vpPoints.size(); iMP < iendMP; iMP++) {
      MapPointT* pMP = vpPoints[iMP];
      if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;
      if (!Sim3Query(pKF, Tcw, Ow, pMP, (float)th, q, m)) continue;
      who.push_back(iMP);
      m++;
    }
    q.resize_keep(m);
    const int N = (int)pKF->mvKeysUn.size();
    std::vector<unsigned char> locked(N, 0);
    for (int j = 0; j < N; j++) locked[j] = vpMatched[j] != NULL;  // :508-509
    std::vector<int> match, feat_query;
    int nmatches = 0;
    const int th_dist = (int)floorf(TH_LOW * ratioHamming);  // bestDist <= TH_LOW*ratioHamming on integer distances
    const int rc = RunProjectionRaw(pKF->mvKeysUn.data(), pKF->mDescriptors.data, N, pKF->mnMinX, pKF->mnMaxX, pKF->mnMinY,
                                    pKF->mnMaxY, nullptr, q, m, locked, ORBM_PROJ_LASTFRAME, 0, th_dist, match, feat_query,
                                    nmatches);
    if (rc != ORBX_OK) return rc;
    for (size_t k = 0; k < m; k++) {
      if (match[k] < 0) continue;
      vpMatched[match[k]] = vpPoints[who[k]];
      if (vpMatchedKF) (*vpMatchedKF)[match[k]] = (*vpPointsKFs)[who[k]];
    }
    return nmatches;
  }

  // candidate scoring of Fuse / SearchBySim3 on a keyframe's features: orbm_fuse_candidates (best index / distance per
  // query; q.qmin carries the predicted level, q.qur the stereo coordinate)
  int RunFuse(KeyFrameT* pKF, bool bRight, const orbxm::Queries& q, size_t M, bool chi_gate, std::vector<int>& best_idx,
              std::vector<int>& best_dist) {
    best_idx.assign(M, -1);
    best_dist.assign(M, 256);
    const std::vector<cv::KeyPoint>& keys = (pKF->NLeft == -1) ? pKF->mvKeysUn : (!bRight ? pKF->mvKeys : pKF->mvKeysRight);
    const int N = (int)keys.size();
    if (M == 0 || N == 0) return ORBX_OK;
    const unsigned char* descs = pKF->mDescriptors.data + (bRight ? (size_t)pKF->NLeft * 32 : 0);
    orbxm::DevCamera cam;
    if (!cam.upload(keys.data(), descs, N, pKF->mnMinX, pKF->mnMaxX, pKF->mnMinY, pKF->mnMaxY, nullptr, nullptr)) return ORBX_E_CUDA;
    const int nlevels = (int)pKF->mvScaleFactors.size();
    std::vector<float> inv(nlevels, 0.f);
    if (chi_gate) inv.assign(pKF->mvInvLevelSigma2.begin(), pKF->mvInvLevelSigma2.end());
    orbxm::Dev ur, qx, qy, qr, qur, ql, qd, nq, out;
    std::vector<int32_t> nqv(1, (int32_t)M);
    if (chi_gate) {
      std::vector<float> u(pKF->mvuRight.begin(), pKF->mvuRight.end());
      u.resize(cam.cap(), -1.f);
      if (!ur.put(u) || !qur.put(q.qur)) return ORBX_E_CUDA;
    }
    if (!qx.put(q.qx) || !qy.put(q.qy) || !qr.put(q.qr) || !ql.put(q.qmin) || !qd.put(q.desc) || !nq.put(nqv) ||
        !out.alloc(M * sizeof(orbm_result)))
      return ORBX_E_CUDA;
    const int rc = orbm_fuse_candidates(cam.kps.template as<orbx_keypoint>(), cam.desc.template as<uint8_t>(),
                                        cam.n.template as<int32_t>(), cam.cap(), cam.cell_off.template as<int32_t>(),
                                        cam.cell_idx.template as<int32_t>(), pKF->mnMinX, pKF->mnMaxX, pKF->mnMinY,
                                        pKF->mnMaxY, ur.as<float>(), inv.data(), nlevels, qx.as<float>(), qy.as<float>(),
                                        qr.as<float>(), qur.as<float>(), ql.as<int8_t>(), qd.as<uint8_t>(), nq.as<int32_t>(),
                                        (int)M, 1, out.as<orbm_result>(), nullptr);
    if (rc != ORBX_OK) return rc;
    std::vector<orbm_result> res(M);
    if (!out.get(res)) return ORBX_E_CUDA;
    for (size_t i = 0; i < M; i++) { best_idx[i] = res[i].best_idx; best_dist[i] = res[i].best_dist; }
    return ORBX_OK;
  }

  int ProjectMono(FrameT& F, const std::vector<MapPointT*>& vpMapPoints, const orbxm::Queries& q,
                  const std::vector<unsigned char>& locked) {
    std::vector<int> match, feat_query;
    int nmatches = 0;
    const int rc = RunProjection(F, q, vpMapPoints.size(), locked, ORBM_PROJ_MAPPOINTS, 0, match, feat_query, nmatches);
    if (rc != ORBX_OK) return rc;
    for (int j = 0; j < F.N; j++)
      if (feat_query[j] >= 0) F.mvpMapPoints[j] = vpMapPoints[feat_query[j]];  // the last writer holds the feature
    return nmatches;
  }

  int ProjectRig(FrameT& F, const std::vector<MapPointT*>& vpMapPoints, const orbxm::Queries& q,
                 const std::vector<unsigned char>& locked) {
    const size_t M = vpMapPoints.size();
    const int nl = F.Nleft, nr = F.N - F.Nleft;
    if (M == 0) return 0;
    std::vector<unsigned char> ll(locked.begin(), locked.begin() + nl), lr(locked.begin() + nl, locked.end());
    std::vector<int> l2r(F.mvLeftToRightMatch.begin(), F.mvLeftToRightMatch.end());
    std::vector<int> r2l(F.mvRightToLeftMatch.begin(), F.mvRightToLeftMatch.end());
    orbxm::DevCamera cl, cr;
    if (!cl.upload(F.mvKeys.data(), F.mDescriptors.data, nl, F.mnMinX, F.mnMaxX, F.mnMinY, F.mnMaxY, &ll, &l2r) ||
        !cr.upload(F.mvKeysRight.data(), F.mDescriptors.data + (size_t)nl * 32, nr, F.mnMinX, F.mnMaxX, F.mnMinY, F.mnMaxY,
                   &lr, &r2l))
      return ORBX_E_CUDA;
    orbm_rig_camera L = {cl.kps.template as<orbx_keypoint>(), cl.desc.template as<uint8_t>(), cl.n.template as<int32_t>(),
                         cl.cap(), cl.cell_off.template as<int32_t>(), cl.cell_idx.template as<int32_t>(),
                         cl.locked.template as<uint8_t>(), cl.partner.template as<int32_t>()};
    orbm_rig_camera R = {cr.kps.template as<orbx_keypoint>(), cr.desc.template as<uint8_t>(), cr.n.template as<int32_t>(),
                         cr.cap(), cr.cell_off.template as<int32_t>(), cr.cell_idx.template as<int32_t>(),
                         cr.locked.template as<uint8_t>(), cr.partner.template as<int32_t>()};
    orbxm::Dev qx, qy, qr, qmin, qmax, qxr, qyr, qrr, qminr, qmaxr, qd, nq, fl, scr, ml, mr, fql, fqr, dnm;
    std::vector<int32_t> nqv(1, (int32_t)M);
    if (!qx.put(q.qx) || !qy.put(q.qy) || !qr.put(q.qr) || !qmin.put(q.qmin) || !qmax.put(q.qmax) || !qxr.put(q.qxr) ||
        !qyr.put(q.qyr) || !qrr.put(q.qrr) || !qminr.put(q.qminr) || !qmaxr.put(q.qmaxr) || !qd.put(q.desc) ||
        !nq.put(nqv) || !fl.put(q.flags) || !scr.alloc(M * 65) || !ml.alloc(M * 4) || !mr.alloc(M * 4) ||
        !fql.alloc((size_t)cl.cap() * 4) || !fqr.alloc((size_t)cr.cap() * 4) || !dnm.alloc(4))
      return ORBX_E_CUDA;
    const int rc = orbm_search_by_projection_rig(
        ORBM_PROJ_MAPPOINTS, &L, &R, 1, F.mnMinX, F.mnMaxX, F.mnMinY, F.mnMaxY, qx.as<float>(), qy.as<float>(),
        qr.as<float>(), qmin.as<int8_t>(), qmax.as<int8_t>(), qxr.as<float>(), qyr.as<float>(), qrr.as<float>(),
        qminr.as<int8_t>(), qmaxr.as<int8_t>(), qd.as<uint8_t>(), nq.as<int32_t>(), (int)M, fl.as<uint8_t>(), nullptr,
        mfNNratio, 0, scr.p, ml.as<int32_t>(), mr.as<int32_t>(), fql.as<int32_t>(), fqr.as<int32_t>(), dnm.as<int32_t>(),
        nullptr);
    if (rc != ORBX_OK) return rc;
    std::vector<int32_t> a(cl.cap()), b(cr.cap()), nm(1);
    if (!fql.get(a) || !fqr.get(b) || !dnm.get(nm)) return ORBX_E_CUDA;
    for (int j = 0; j < nl; j++)
      if (a[j] >= 0) F.mvpMapPoints[j] = vpMapPoints[a[j]];
    for (int j = 0; j < nr; j++)
      if (b[j] >= 0) F.mvpMapPoints[j + nl] = vpMapPoints[b[j]];
    return nm[0];
  }

  // SearchByBoW on a rig frame: orbm_search_by_bow_rig through the BowMatcher.h upload helpers
  int SearchByBoWRig(const BowSide& kf, const BowSide& fr, int nleft, std::vector<int>& match21) {
    int ns = 1;
    const int cap = bowdetail::strides(kf, fr, &ns);
    bowdetail::DevSide d1, d2;
    if (!bowdetail::upload(kf, cap, ns, d1) || !bowdetail::upload(fr, cap, ns, d2)) return ORBX_E_CUDA;
    orbxm::Dev m12l, m12r, m21, nm, dnl;
    std::vector<int32_t> nlv(1, nleft);
    if (!m12l.alloc((size_t)cap * 4) || !m12r.alloc((size_t)cap * 4) || !m21.alloc((size_t)cap * 4) || !nm.alloc(4) ||
        !dnl.put(nlv))
      return ORBX_E_NOMEM;
    const int rc = orbm_search_by_bow_rig((const orbx_keypoint*)d1.kps, (const uint8_t*)d1.desc, (const int32_t*)d1.n,
                                          (const uint8_t*)d1.flag, &d1.fv, (const orbx_keypoint*)d2.kps,
                                          (const uint8_t*)d2.desc, (const int32_t*)d2.n, dnl.as<int32_t>(), &d2.fv, cap, ns,
                                          1, mfNNratio, mbCheckOrientation ? 1 : 0, m12l.as<int32_t>(), m12r.as<int32_t>(),
                                          m21.as<int32_t>(), nm.as<int32_t>(), nullptr);
    if (rc != ORBX_OK) return rc;
    std::vector<int32_t> v(cap), n(1);
    if (!m21.get(v) || !nm.get(n)) return ORBX_E_CUDA;
    match21.assign(v.begin(), v.begin() + fr.keys->size());
    return n[0];
  }

  float mfNNratio;
  bool mbCheckOrientation;
};

}  // namespace ORB_SLAM3
#endif
