// orb_oracle.cpp — CPU ORACLE for the ORB front-end hot path.
//
// TEST INFRASTRUCTURE ONLY.  Nothing under orb_slam3_vio_fixes_b200/ may link,
// import or call this file; only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs use it, as the checker / CPU baseline.
//
// What it is: a plain C++17 restatement of the reference's algorithm
//   src/ORBextractor.cc  (whole file except the dead ComputeKeyPointsOld :898-1075)
//   src/ORBmatcher.cc    (:2058-2074 distance, the best/second loops of :43-213,
//                         :907-1146 last-min rule, :1676-1887 best-only)
//   src/Frame.cc         (:385-416, :657-735 grid; :811-981 stereo matching)
// with the OpenCV primitives the reference calls (resize, copyMakeBorder, FAST,
// GaussianBlur, fastAtan2, cvRound) restated as exact integer / fp32 code.
//
// PINNED (extractor, rows a1-a11): oracle/_ref compiles the reference's own
// src/ORBextractor.cc UNMODIFIED (recipe: oracle/Makefile target `ref`, over the OpenCV
// header shim oracle/ref_shim/) and tests/test_ref_extractor.py asserts array_equal between
// that build and this file on C1-C5, the threshold/fallback cases and the committed goldens,
// with the five image primitives taken (a) from this file's exports and (b) from the real
// OpenCV 4.13 binary (cv2 callbacks).  The primitives themselves are pinned against cv2 one
// by one (tests/test_oracle_vs_cv2.py).  The reference ships no tests / golden vectors of
// its own for this path.
// PINNED (matcher rows m2-m9, m12): oracle/_ref/libref_orbmatcher.so compiles the reference's own
// src/ORBmatcher.cc UNMODIFIED (recipe: oracle/Makefile target `ref_matcher`; the reference's
// include/ORBmatcher.h as it is; Eigen / Sophus / DBoW2 / Frame / KeyFrame / MapPoint stand-ins in the
// force-included oracle/ref_shim_matcher/prelude.h, which takes the include guards of the headers
// that are not in the image) and tests/test_ref_matcher.py asserts array_equal between the reference's
// SearchByProjection x5 (MapPoints mono + rig, last frame mono + rig, relocalisation, Sim3 x2),
// SearchByBoW x3, SearchForTriangulation, SearchForInitialization, Fuse and the restatements here.
// PINNED (m11, (f).1): Frame.cc cannot be compiled as a whole here, but Frame::ComputeStereoMatches
// (:811-981) and ComputeStereoFromRGBD (:984-1005) can: oracle/Makefile target `ref_frame` cuts them
// verbatim out of the reference file at build time (scratch directory outside the repository) and compiles
// that text over the same stand-ins -> oracle/_ref/libref_frame.so; tests/test_ref_frame.py asserts
// array_equal of mvuRight / mvDepth with the restatement here at C2 and C4 (maxD = fx).
//
// Floating-point contract:
//   * cos/sin: libm sincosf, exactly what the reference's `cos(float)`/`sin(float)`
//     compile to (src/ORBextractor.cc:112).
//   * GET_VALUE (src/ORBextractor.cc:117-119): the reference's CMake flags
//     (-O3 -march=native, CMakeLists.txt:15-16) make g++ contract the two
//     expressions to  fmaf(x,b, y*a)  and  fmaf(x,a, -(y*b))  (probed with g++ 13.3,
//     32 vfmadd231ss/vfmsub132ss in computeOrbDescriptor).  flag ORACLE_STRICT_FP
//     selects the uncontracted source-order evaluation instead.
//   * everything else is integer or single fp32 operations in source order; this
//     file must be compiled with -ffp-contract=off.
#include <algorithm>
#include <cassert>
#include <chrono>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <list>
#include <thread>
#include <utility>
#include <map>
#include <memory>
#include <vector>

namespace {

static const int8_t kPattern[1024] = {
#include "orb_pattern_oracle.inc"
};

constexpr int PATCH_SIZE = 31;       // ORBextractor.cc:71
constexpr int HALF_PATCH_SIZE = 15;  // :72
constexpr int EDGE_THRESHOLD = 19;   // :73

// cv::KeyPoint memory layout (28 bytes).
struct KeyPoint {
  float x, y, size, angle, response;
  int octave, class_id;
};

struct Image {
  int w = 0, h = 0;
  std::vector<uint8_t> d;
  uint8_t* row(int y) { return d.data() + (size_t)y * w; }
  const uint8_t* row(int y) const { return d.data() + (size_t)y * w; }
};

// cvRound(float): SSE cvtss2si under the default rounding mode = round half to even.
inline int cvRound(float v) { return (int)lrintf(v); }
inline int cvRoundD(double v) { return (int)lrint(v); }
inline int cvFloor(float v) { return (int)floorf(v); }
inline int cvFloorD(double v) { return (int)floor(v); }
inline int cvCeilD(double v) { return (int)ceil(v); }

// ---------------------------------------------------------------------------
// cv::resize(src, dst, sz, 0, 0, INTER_LINEAR) for CV_8UC1  (ORBextractor.cc:1183)
// OpenCV imgproc/resize.cpp: resizeGeneric_<HResizeLinear<uchar,int,short,2048>,
// VResizeLinear<uchar,int,short,FixedPtCast<int,uchar,22>>>.
// ---------------------------------------------------------------------------
struct AxisTab {
  std::vector<int> ofs;       // source index
  std::vector<int16_t> c0, c1;  // Q11 weights
};

// Horizontal taps: coefficients are zeroed/clamped at both ends (xofs/alpha build loop).
AxisTab build_x_tab(int ssize, int dsize) {
  AxisTab t;
  t.ofs.resize(dsize); t.c0.resize(dsize); t.c1.resize(dsize);
  double inv_scale = (double)dsize / ssize;
  double scale = 1. / inv_scale;
  for (int dx = 0; dx < dsize; dx++) {
    float fx = (float)((dx + 0.5) * scale - 0.5);
    int sx = cvFloor(fx);
    fx -= sx;
    if (sx < 0) { fx = 0; sx = 0; }
    if (sx >= ssize - 1) { fx = 0; sx = ssize - 1; }
    t.ofs[dx] = sx;
    t.c0[dx] = (int16_t)cvRound((1.f - fx) * 2048.f);
    t.c1[dx] = (int16_t)cvRound(fx * 2048.f);
  }
  return t;
}
// Vertical taps: coefficients kept, row indices clipped instead.
AxisTab build_y_tab(int ssize, int dsize) {
  AxisTab t;
  t.ofs.resize(dsize); t.c0.resize(dsize); t.c1.resize(dsize);
  double inv_scale = (double)dsize / ssize;
  double scale = 1. / inv_scale;
  for (int dy = 0; dy < dsize; dy++) {
    float fy = (float)((dy + 0.5) * scale - 0.5);
    int sy = cvFloor(fy);
    fy -= sy;
    t.ofs[dy] = sy;
    t.c0[dy] = (int16_t)cvRound((1.f - fy) * 2048.f);
    t.c1[dy] = (int16_t)cvRound(fy * 2048.f);
  }
  return t;
}

void resize_linear_u8(const Image& s, Image& d, int dw, int dh) {
  d.w = dw; d.h = dh; d.d.assign((size_t)dw * dh, 0);
  AxisTab tx = build_x_tab(s.w, dw), ty = build_y_tab(s.h, dh);
  std::vector<int> h0(dw), h1(dw);
  auto hrow = [&](int sy, std::vector<int>& out) {
    sy = std::min(std::max(sy, 0), s.h - 1);
    const uint8_t* S = s.row(sy);
    for (int dx = 0; dx < dw; dx++) {
      int sx = tx.ofs[dx];
      int sx1 = std::min(sx + 1, s.w - 1);
      out[dx] = S[sx] * tx.c0[dx] + S[sx1] * tx.c1[dx];
    }
  };
  for (int dy = 0; dy < dh; dy++) {
    hrow(ty.ofs[dy], h0);
    hrow(ty.ofs[dy] + 1, h1);
    int b0 = ty.c0[dy], b1 = ty.c1[dy];
    uint8_t* D = d.row(dy);
    for (int dx = 0; dx < dw; dx++) {
      int v = (((b0 * (h0[dx] >> 4)) >> 16) + ((b1 * (h1[dx] >> 4)) >> 16) + 2) >> 2;
      D[dx] = (uint8_t)std::min(std::max(v, 0), 255);
    }
  }
}

// BORDER_REFLECT_101 index map (cv::borderInterpolate).
inline int reflect101(int p, int len) {
  if (len == 1) return 0;
  while (p < 0 || p >= len) {
    if (p < 0) p = -p;
    else p = 2 * (len - 1) - p;
  }
  return p;
}

// cv::copyMakeBorder(src, dst, b,b,b,b, BORDER_REFLECT_101 | BORDER_ISOLATED) (:1185-1191)
void copy_make_border101(const Image& s, Image& d, int b) {
  d.w = s.w + 2 * b; d.h = s.h + 2 * b; d.d.resize((size_t)d.w * d.h);
  for (int y = 0; y < d.h; y++) {
    const uint8_t* S = s.row(reflect101(y - b, s.h));
    uint8_t* D = d.row(y);
    for (int x = 0; x < b; x++) D[x] = S[reflect101(x - b, s.w)];
    memcpy(D + b, S, s.w);
    for (int x = b + s.w; x < d.w; x++) D[x] = S[reflect101(x - b, s.w)];
  }
}

// ---------------------------------------------------------------------------
// cv::GaussianBlur(m, m, Size(7,7), 2, 2, BORDER_REFLECT_101) on CV_8UC1 (:1133)
// OpenCV's 8-bit fixed-point path: Q8 taps, one rounding at the end.
// ---------------------------------------------------------------------------
static const int kGauss7[7] = {18, 34, 48, 56, 48, 34, 18};

void gaussian_blur7(const Image& s, Image& d) {
  d.w = s.w; d.h = s.h; d.d.resize(s.d.size());
  const int w = s.w, h = s.h;
  std::vector<uint16_t> tmp((size_t)w * h);
  for (int y = 0; y < h; y++) {
    const uint8_t* S = s.row(y);
    uint16_t* T = &tmp[(size_t)y * w];
    auto edge = [&](int x) {
      int acc = 0;
      for (int k = -3; k <= 3; k++) acc += kGauss7[k + 3] * S[reflect101(x + k, w)];
      T[x] = (uint16_t)acc;  // <= 255*256
    };
    if (w < 8) { for (int x = 0; x < w; x++) edge(x); continue; }
    for (int x = 0; x < 3; x++) edge(x);
    for (int x = w - 3; x < w; x++) edge(x);
    for (int x = 3; x < w - 3; x++)  // interior: no border handling, vectorises
      T[x] = (uint16_t)(18 * (S[x - 3] + S[x + 3]) + 34 * (S[x - 2] + S[x + 2]) + 48 * (S[x - 1] + S[x + 1]) +
                        56 * S[x]);
  }
  for (int y = 0; y < h; y++) {
    const uint16_t* r[7];
    for (int k = 0; k < 7; k++) r[k] = &tmp[(size_t)reflect101(y + k - 3, h) * w];
    uint8_t* D = d.row(y);
    for (int x = 0; x < w; x++) {
      uint32_t acc = 18u * (r[0][x] + r[6][x]) + 34u * (r[1][x] + r[5][x]) + 48u * (r[2][x] + r[4][x]) +
                     56u * r[3][x];
      D[x] = (uint8_t)((acc + 32768u) >> 16);
    }
  }
}

// ---------------------------------------------------------------------------
// cv::fastAtan2(y, x) scalar path (core/mathfuncs_core), fp32 Horner, no FMA.
// ---------------------------------------------------------------------------
float fast_atan2(float y, float x) {
  static const float p1 = 0.9997878412794807f * (float)(180 / 3.14159265358979323846);
  static const float p3 = -0.3258083974640975f * (float)(180 / 3.14159265358979323846);
  static const float p5 = 0.1555786518463281f * (float)(180 / 3.14159265358979323846);
  static const float p7 = -0.04432655554792128f * (float)(180 / 3.14159265358979323846);
  float ax = std::fabs(x), ay = std::fabs(y);
  float a, c, c2;
  if (ax >= ay) {
    c = ay / (ax + (float)2.2204460492503131e-16);
    c2 = c * c;
    a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
  } else {
    c = ax / (ay + (float)2.2204460492503131e-16);
    c2 = c * c;
    a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
  }
  if (x < 0) a = 180.f - a;
  if (y < 0) a = 360.f - a;
  return a;
}

// ---------------------------------------------------------------------------
// cv::FAST(img, kps, threshold, true)  TYPE_9_16 with non-max suppression,
// on a ROI of `img` (features2d/fast.cpp FAST_t<16> + cornerScore<16>).
// ROI = [x0,x1) x [y0,y1) of a plain image with row stride img.w.
// Output coordinates are ROI-relative, row-major order.
// ---------------------------------------------------------------------------
static const int kRing[16][2] = {{0, 3},  {1, 3},   {2, 2},   {3, 1},  {3, 0},  {3, -1},
                                 {2, -2}, {1, -3},  {0, -3},  {-1, -3}, {-2, -2}, {-3, -1},
                                 {-3, 0}, {-3, 1},  {-2, 2},  {-1, 3}};

// "Arc strength" A of a pixel: the largest t+1 for which the pixel is still a 9/16 corner at
// threshold t, i.e. max over the 16 contiguous arcs of 9 of max(min(d), min(-d)), d = centre -
// ring.  corner(th) <=> A > th, and cornerScore<16>() == A - 1 for every corner.
inline int fast_arc_strength(const int* d /*[25]*/) {
  // sliding window of 9 over the circular 16: 2-, 4-, 8-wide partial minima/maxima, then +1
  int mn2[16], mx2[16], mn4[16], mx4[16];
  for (int k = 0; k < 16; k++) { mn2[k] = std::min(d[k], d[k + 1]); mx2[k] = std::max(d[k], d[k + 1]); }
  for (int k = 0; k < 16; k++) {
    mn4[k] = std::min(mn2[k], mn2[(k + 2) & 15]); mx4[k] = std::max(mx2[k], mx2[(k + 2) & 15]);
  }
  int best = 0;
  for (int k = 0; k < 16; k++) {
    int mn = std::min(std::min(mn4[k], mn4[(k + 4) & 15]), d[k + 8]);
    int mx = std::max(std::max(mx4[k], mx4[(k + 4) & 15]), d[k + 8]);
    best = std::max(best, std::max(mn, -mx));
  }
  return best;
}

struct Corner { int x, y, score; };

void fast9_16_nms(const Image& img, int x0, int y0, int x1, int y1, int threshold,
                  std::vector<Corner>& out) {
  out.clear();
  int W = x1 - x0, H = y1 - y0;
  if (W < 7 || H < 7) return;
  threshold = std::min(std::max(threshold, 0), 255);
  // amap: A for corners (A > threshold), 0 elsewhere; rows/cols outside [3,dim-3) stay 0 like
  // the zero-initialised score rows of FAST_t.
  std::vector<uint8_t> amap((size_t)W * H, 0);
  const int step = img.w;
  int off[16];
  for (int k = 0; k < 16; k++) off[k] = kRing[k][1] * step + kRing[k][0];
  for (int y = 3; y < H - 3; y++) {
    const uint8_t* p = img.row(y0 + y) + x0;
    uint8_t* arow = &amap[(size_t)y * W];
    for (int x = 3; x < W - 3; x++) {
      const uint8_t* c = p + x;
      const int v = c[0], lo = v - threshold, hi = v + threshold;
      // necessary condition: every arc of 9 contains one pixel of each opposite pair
      auto cls = [&](int k) { int q = c[off[k]]; return (q < lo ? 1 : 0) | (q > hi ? 2 : 0); };
      int m = cls(0) | cls(8);
      if (!m) continue;
      m &= cls(4) | cls(12);
      if (!m) continue;
      m &= cls(2) | cls(10);
      if (!m) continue;
      m &= cls(6) | cls(14);
      if (!m) continue;
      int d[25];
      for (int k = 0; k < 25; k++) d[k] = v - c[off[k & 15]];
      int A = fast_arc_strength(d);
      if (A > threshold) arow[x] = (uint8_t)A;
    }
  }
  for (int y = 3; y < H - 3; y++)
    for (int x = 3; x < W - 3; x++) {
      const uint8_t* p = &amap[(size_t)y * W + x];
      int s = p[0];
      if (!s) continue;  // not in cornerpos[]
      // score = A-1 on both sides of every comparison; a corner with score 0 never wins
      if (s > 1 && s > p[-1] && s > p[1] && s > p[-W - 1] && s > p[-W] && s > p[-W + 1] &&
          s > p[W - 1] && s > p[W] && s > p[W + 1])
        out.push_back({x, y, s - 1});
    }
}

// ---------------------------------------------------------------------------
// ExtractorNode / DistributeOctTree  (ORBextractor.cc:480-553, 555-779)
// ---------------------------------------------------------------------------
struct Pt2i { int x = 0, y = 0; };

struct Node {
  std::vector<KeyPoint> vKeys;
  Pt2i UL, UR, BL, BR;
  std::list<Node>::iterator lit;
  bool bNoMore = false;
  void Divide(Node& n1, Node& n2, Node& n3, Node& n4) const;
};

void Node::Divide(Node& n1, Node& n2, Node& n3, Node& n4) const {  // :480-536
  const int halfX = (int)std::ceil(static_cast<float>(UR.x - UL.x) / 2);
  const int halfY = (int)std::ceil(static_cast<float>(BR.y - UL.y) / 2);
  n1.UL = UL;
  n1.UR = {UL.x + halfX, UL.y};
  n1.BL = {UL.x, UL.y + halfY};
  n1.BR = {UL.x + halfX, UL.y + halfY};
  n2.UL = n1.UR; n2.UR = UR; n2.BL = n1.BR; n2.BR = {UR.x, UL.y + halfY};
  n3.UL = n1.BL; n3.UR = n1.BR; n3.BL = BL; n3.BR = {n1.BR.x, BL.y};
  n4.UL = n3.UR; n4.UR = n2.BR; n4.BL = n3.BR; n4.BR = BR;
  for (const KeyPoint& kp : vKeys) {
    if (kp.x < n1.UR.x) {
      if (kp.y < n1.BR.y) n1.vKeys.push_back(kp); else n3.vKeys.push_back(kp);
    } else if (kp.y < n1.BR.y) n2.vKeys.push_back(kp);
    else n4.vKeys.push_back(kp);
  }
  if (n1.vKeys.size() == 1) n1.bNoMore = true;
  if (n2.vKeys.size() == 1) n2.bNoMore = true;
  if (n3.vKeys.size() == 1) n3.bNoMore = true;
  if (n4.vKeys.size() == 1) n4.bNoMore = true;
}

bool compareNodes(std::pair<int, Node*>& e1, std::pair<int, Node*>& e2) {  // :538-553
  if (e1.first < e2.first) return true;
  else if (e1.first > e2.first) return false;
  else return e1.second->UL.x < e2.second->UL.x;
}

std::vector<KeyPoint> DistributeOctTree(const std::vector<KeyPoint>& vToDistributeKeys, int minX,
                                        int maxX, int minY, int maxY, int N) {
  const int nIni = (int)std::round(static_cast<float>(maxX - minX) / (maxY - minY));  // :559
  const float hX = static_cast<float>(maxX - minX) / nIni;
  std::list<Node> lNodes;
  std::vector<Node*> vpIniNodes(nIni);
  for (int i = 0; i < nIni; i++) {
    Node ni;
    ni.UL = {(int)(hX * static_cast<float>(i)), 0};
    ni.UR = {(int)(hX * static_cast<float>(i + 1)), 0};
    ni.BL = {ni.UL.x, maxY - minY};
    ni.BR = {ni.UR.x, maxY - minY};
    lNodes.push_back(ni);
    vpIniNodes[i] = &lNodes.back();
  }
  for (const KeyPoint& kp : vToDistributeKeys) vpIniNodes[(int)(kp.x / hX)]->vKeys.push_back(kp);

  auto lit = lNodes.begin();
  while (lit != lNodes.end()) {
    if (lit->vKeys.size() == 1) { lit->bNoMore = true; lit++; }
    else if (lit->vKeys.empty()) lit = lNodes.erase(lit);
    else lit++;
  }

  bool bFinish = false;
  std::vector<std::pair<int, Node*>> vSizeAndPointerToNode;
  auto add_children = [&](Node* ch[4], int* nToExpand) {
    for (int c = 0; c < 4; c++) {
      if (ch[c]->vKeys.size() > 0) {
        lNodes.push_front(*ch[c]);
        if (ch[c]->vKeys.size() > 1) {
          if (nToExpand) (*nToExpand)++;
          vSizeAndPointerToNode.push_back(std::make_pair((int)ch[c]->vKeys.size(), &lNodes.front()));
          lNodes.front().lit = lNodes.begin();
        }
      }
    }
  };

  while (!bFinish) {
    int prevSize = (int)lNodes.size();
    lit = lNodes.begin();
    int nToExpand = 0;
    vSizeAndPointerToNode.clear();
    while (lit != lNodes.end()) {
      if (lit->bNoMore) { lit++; continue; }
      Node n1, n2, n3, n4;
      lit->Divide(n1, n2, n3, n4);
      Node* ch[4] = {&n1, &n2, &n3, &n4};
      add_children(ch, &nToExpand);
      lit = lNodes.erase(lit);
    }
    if ((int)lNodes.size() >= N || (int)lNodes.size() == prevSize) {
      bFinish = true;
    } else if (((int)lNodes.size() + nToExpand * 3) > N) {
      while (!bFinish) {
        prevSize = (int)lNodes.size();
        std::vector<std::pair<int, Node*>> vPrev = vSizeAndPointerToNode;
        vSizeAndPointerToNode.clear();
        std::sort(vPrev.begin(), vPrev.end(), compareNodes);
        for (int j = (int)vPrev.size() - 1; j >= 0; j--) {
          Node n1, n2, n3, n4;
          vPrev[j].second->Divide(n1, n2, n3, n4);
          Node* ch[4] = {&n1, &n2, &n3, &n4};
          add_children(ch, nullptr);
          lNodes.erase(vPrev[j].second->lit);
          if ((int)lNodes.size() >= N) break;
        }
        if ((int)lNodes.size() >= N || (int)lNodes.size() == prevSize) bFinish = true;
      }
    }
  }

  std::vector<KeyPoint> vResultKeys;
  for (auto it = lNodes.begin(); it != lNodes.end(); it++) {  // :757-778
    const std::vector<KeyPoint>& vNodeKeys = it->vKeys;
    const KeyPoint* pKP = &vNodeKeys[0];
    float maxResponse = pKP->response;
    for (size_t k = 1; k < vNodeKeys.size(); k++)
      if (vNodeKeys[k].response > maxResponse) { pKP = &vNodeKeys[k]; maxResponse = vNodeKeys[k].response; }
    vResultKeys.push_back(*pKP);
  }
  return vResultKeys;
}

// ---------------------------------------------------------------------------
// The extractor
// ---------------------------------------------------------------------------
enum { ORACLE_STRICT_FP = 1 };

struct Extractor {
  int nfeatures, nlevels, iniThFAST, minThFAST, flags;
  double scaleFactor;  // the member is a double in the reference (ORBextractor.h:97)
  std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;
  std::vector<int> mnFeaturesPerLevel, umax;
  std::vector<Image> pyr;       // level images without border
  std::vector<Image> pyrB;      // with the 19-px REFLECT_101 border (mvImagePyramid parents)
  std::vector<std::vector<KeyPoint>> candidates;  // per level, before the octree (debug/tests)
  std::vector<std::vector<KeyPoint>> perLevel;    // per level after octree+orientation
  // wall seconds spent per stage, accumulated over extract() calls (bench only):
  // 0 pyramid, 1 per-cell FAST, 2 octree (+fix-up), 3 orientation, 4 blur, 5 descriptors (+ordering)
  double stage_sec[6] = {0, 0, 0, 0, 0, 0};
  static double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

  Extractor(int nf, float sf, int nl, int ini, int mn, int fl)
      : nfeatures(nf), nlevels(nl), iniThFAST(ini), minThFAST(mn), flags(fl), scaleFactor(sf) {
    // ORBextractor.cc:414-430
    mvScaleFactor.resize(nlevels); mvLevelSigma2.resize(nlevels);
    mvScaleFactor[0] = 1.0f; mvLevelSigma2[0] = 1.0f;
    for (int i = 1; i < nlevels; i++) {
      mvScaleFactor[i] = (float)(mvScaleFactor[i - 1] * scaleFactor);
      mvLevelSigma2[i] = mvScaleFactor[i] * mvScaleFactor[i];
    }
    mvInvScaleFactor.resize(nlevels); mvInvLevelSigma2.resize(nlevels);
    for (int i = 0; i < nlevels; i++) {
      mvInvScaleFactor[i] = 1.0f / mvScaleFactor[i];
      mvInvLevelSigma2[i] = 1.0f / mvLevelSigma2[i];
    }
    // :434-445
    mnFeaturesPerLevel.resize(nlevels);
    float factor = (float)(1.0f / scaleFactor);
    float nDesired = nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nlevels));
    int sumFeatures = 0;
    for (int level = 0; level < nlevels - 1; level++) {
      mnFeaturesPerLevel[level] = cvRound(nDesired);
      sumFeatures += mnFeaturesPerLevel[level];
      nDesired *= factor;
    }
    mnFeaturesPerLevel[nlevels - 1] = std::max(nfeatures - sumFeatures, 0);
    // :453-468
    umax.resize(HALF_PATCH_SIZE + 1);
    int v, v0, vmax = cvFloor(HALF_PATCH_SIZE * sqrtf(2.f) / 2 + 1);
    int vmin = (int)std::ceil(HALF_PATCH_SIZE * sqrtf(2.f) / 2);
    const double hp2 = HALF_PATCH_SIZE * HALF_PATCH_SIZE;
    for (v = 0; v <= vmax; ++v) umax[v] = cvRoundD(sqrt(hp2 - v * v));
    for (v = HALF_PATCH_SIZE, v0 = 0; v >= vmin; --v) {
      while (umax[v0] == umax[v0 + 1]) ++v0;
      umax[v] = v0;
      ++v0;
    }
  }

  void level_size(int level, int cols, int rows, int* w, int* h) const {  // :1174-1175
    float scale = mvInvScaleFactor[level];
    *w = cvRound((float)cols * scale);
    *h = cvRound((float)rows * scale);
  }

  void ComputePyramid(const uint8_t* img, int w, int h, size_t stride) {  // :1170-1195
    pyr.assign(nlevels, Image()); pyrB.assign(nlevels, Image());
    for (int level = 0; level < nlevels; ++level) {
      int lw, lh; level_size(level, w, h, &lw, &lh);
      if (level == 0) {
        pyr[0].w = lw; pyr[0].h = lh; pyr[0].d.resize((size_t)lw * lh);
        for (int y = 0; y < lh; y++) memcpy(pyr[0].row(y), img + (size_t)y * stride, lw);
      } else {
        resize_linear_u8(pyr[level - 1], pyr[level], lw, lh);
      }
      copy_make_border101(pyr[level], pyrB[level], EDGE_THRESHOLD);
    }
  }

  float IC_Angle(const Image& im, int px, int py) const {  // :76-103
    int m_01 = 0, m_10 = 0;
    const int step = im.w;
    const uint8_t* center = im.row(py) + px;
    for (int u = -HALF_PATCH_SIZE; u <= HALF_PATCH_SIZE; ++u) m_10 += u * center[u];
    for (int v = 1; v <= HALF_PATCH_SIZE; ++v) {
      int v_sum = 0, d = umax[v];
      for (int u = -d; u <= d; ++u) {
        int val_plus = center[u + v * step], val_minus = center[u - v * step];
        v_sum += (val_plus - val_minus);
        m_10 += u * (val_plus + val_minus);
      }
      m_01 += v * v_sum;
    }
    return fast_atan2((float)m_01, (float)m_10);
  }

  void ComputeKeyPointsOctTree() {  // :781-896
    candidates.assign(nlevels, {}); perLevel.assign(nlevels, {});
    const float W = 35;
    for (int level = 0; level < nlevels; ++level) {
      const Image& im = pyr[level];  // coordinates below are level coords (border-free)
      const int minBorderX = EDGE_THRESHOLD - 3, minBorderY = minBorderX;
      const int maxBorderX = im.w - EDGE_THRESHOLD + 3, maxBorderY = im.h - EDGE_THRESHOLD + 3;
      std::vector<KeyPoint>& vToDistributeKeys = candidates[level];
      const float width = (float)(maxBorderX - minBorderX), height = (float)(maxBorderY - minBorderY);
      const int nCols = (int)(width / W), nRows = (int)(height / W);
      const int wCell = (int)std::ceil(width / nCols), hCell = (int)std::ceil(height / nRows);
      std::vector<Corner> cell;
      const double t_fast0 = now_s();
      for (int i = 0; i < nRows; i++) {
        const float iniY = (float)(minBorderY + i * hCell);
        float maxY = iniY + hCell + 6;
        if (iniY >= maxBorderY - 3) continue;
        if (maxY > maxBorderY) maxY = (float)maxBorderY;
        for (int j = 0; j < nCols; j++) {
          const float iniX = (float)(minBorderX + j * wCell);
          float maxX = iniX + wCell + 6;
          if (iniX >= maxBorderX - 6) continue;
          if (maxX > maxBorderX) maxX = (float)maxBorderX;
          fast9_16_nms(im, (int)iniX, (int)iniY, (int)maxX, (int)maxY, iniThFAST, cell);
          if (cell.empty()) fast9_16_nms(im, (int)iniX, (int)iniY, (int)maxX, (int)maxY, minThFAST, cell);
          for (const Corner& c : cell) {
            KeyPoint kp{(float)c.x, (float)c.y, 7.f, -1.f, (float)c.score, 0, -1};
            kp.x += j * wCell; kp.y += i * hCell;
            vToDistributeKeys.push_back(kp);
          }
        }
      }
      const double t_fast1 = now_s();
      std::vector<KeyPoint>& keypoints = perLevel[level];
      keypoints = DistributeOctTree(vToDistributeKeys, minBorderX, maxBorderX, minBorderY, maxBorderY,
                                    mnFeaturesPerLevel[level]);
      const int scaledPatchSize = (int)(PATCH_SIZE * mvScaleFactor[level]);
      for (KeyPoint& kp : keypoints) {
        kp.x += minBorderX; kp.y += minBorderY; kp.octave = level; kp.size = (float)scaledPatchSize;
      }
      const double t_oct1 = now_s();
      stage_sec[1] += t_fast1 - t_fast0; stage_sec[2] += t_oct1 - t_fast1;
    }
    const double t_or0 = now_s();
    for (int level = 0; level < nlevels; ++level)
      for (KeyPoint& kp : perLevel[level]) kp.angle = IC_Angle(pyr[level], cvRound(kp.x), cvRound(kp.y));
    stage_sec[3] += now_s() - t_or0;
  }

  void computeOrbDescriptor(const KeyPoint& kpt, const Image& img, uint8_t* desc) const {  // :107-146
    const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
    float angle = (float)kpt.angle * factorPI;
    float a, b;
    sincosf(angle, &b, &a);
    const uint8_t* center = img.row(cvRound(kpt.y)) + cvRound(kpt.x);
    const int step = img.w;
    const bool strict = flags & ORACLE_STRICT_FP;
    auto get = [&](const int8_t* p) -> int {
      float x = (float)p[0], y = (float)p[1];
      int iy, ix;
      if (strict) { iy = cvRound(x * b + y * a); ix = cvRound(x * a - y * b); }
      else { iy = cvRound(fmaf(x, b, y * a)); ix = cvRound(fmaf(x, a, -(y * b))); }
      return center[iy * step + ix];
    };
    const int8_t* pattern = kPattern;
    for (int i = 0; i < 32; ++i, pattern += 32) {
      int val = 0;
      for (int k = 0; k < 8; k++) {
        int t0 = get(pattern + 4 * k), t1 = get(pattern + 4 * k + 2);
        val |= (t0 < t1) << k;
      }
      desc[i] = (uint8_t)val;
    }
  }

  // operator()  (:1086-1168).  Returns monoIndex, or -1 for an empty image.
  int extract(const uint8_t* img, int w, int h, size_t stride, int lap0, int lap1, KeyPoint* out_kps,
              uint8_t* out_desc, int capacity, int* n_out) {
    if (!img || w <= 0 || h <= 0) return -1;
    const double t_p0 = now_s();
    ComputePyramid(img, w, h, stride);
    stage_sec[0] += now_s() - t_p0;
    ComputeKeyPointsOctTree();
    int nkeypoints = 0;
    for (int level = 0; level < nlevels; ++level) nkeypoints += (int)perLevel[level].size();
    *n_out = nkeypoints;
    if (nkeypoints > capacity) return -2;
    int monoIndex = 0, stereoIndex = nkeypoints - 1;
    Image working;
    for (int level = 0; level < nlevels; ++level) {
      std::vector<KeyPoint>& keypoints = perLevel[level];
      if (keypoints.empty()) continue;
      const double t_b0 = now_s();
      gaussian_blur7(pyr[level], working);
      const double t_b1 = now_s();
      stage_sec[4] += t_b1 - t_b0;
      float scale = mvScaleFactor[level];
      for (KeyPoint kp : keypoints) {
        uint8_t d[32];
        computeOrbDescriptor(kp, working, d);
        if (level != 0) { kp.x *= scale; kp.y *= scale; }
        int dst;
        if (kp.x >= lap0 && kp.x <= lap1) dst = stereoIndex--; else dst = monoIndex++;
        out_kps[dst] = kp;
        memcpy(out_desc + (size_t)dst * 32, d, 32);
      }
      stage_sec[5] += now_s() - t_b1;
    }
    return monoIndex;
  }
};

// ---------------------------------------------------------------------------
// ORBmatcher::DescriptorDistance (ORBmatcher.cc:2058-2074) — SWAR popcount over 8 int32.
// ---------------------------------------------------------------------------
int DescriptorDistance(const uint8_t* a, const uint8_t* b) {
  int dist = 0;
  for (int i = 0; i < 8; i++) {
    uint32_t pa, pb;
    memcpy(&pa, a + 4 * i, 4); memcpy(&pb, b + 4 * i, 4);
    unsigned int v = pa ^ pb;
    v = v - ((v >> 1) & 0x55555555);
    v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
    dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
  }
  return dist;
}

}  // namespace

// ===========================================================================
// C interface for ctypes (tests / bench only)
// ===========================================================================
extern "C" {

struct oracle_result {  // mirrors orbm_result in include/orbx.h
  int32_t best_idx;
  int16_t best_dist, second_dist;
  int8_t best_level, second_level;
  int16_t pad;
};

void* oracle_create(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh, int flags) {
  return new Extractor(nfeatures, scaleFactor, nlevels, iniTh, minTh, flags);
}
void oracle_destroy(void* h) { delete (Extractor*)h; }

int oracle_extract(void* h, const uint8_t* img, int w, int hgt, size_t stride, int lap0, int lap1,
                   void* kps, uint8_t* desc, int capacity, int* n_out) {
  return ((Extractor*)h)->extract(img, w, hgt, stride, lap0, lap1, (KeyPoint*)kps, desc, capacity, n_out);
}

void oracle_tables(void* h, float* sf, float* inv, float* sig2, float* invsig2, int* nfeat, int* umax16) {
  Extractor* e = (Extractor*)h;
  for (int i = 0; i < e->nlevels; i++) {
    sf[i] = e->mvScaleFactor[i]; inv[i] = e->mvInvScaleFactor[i];
    sig2[i] = e->mvLevelSigma2[i]; invsig2[i] = e->mvInvLevelSigma2[i];
    nfeat[i] = e->mnFeaturesPerLevel[i];
  }
  for (int i = 0; i < 16; i++) umax16[i] = e->umax[i];
}

int oracle_level_size(void* h, int level, int* w, int* hgt) {
  Extractor* e = (Extractor*)h;
  if (level < 0 || level >= (int)e->pyr.size()) return -1;
  *w = e->pyr[level].w; *hgt = e->pyr[level].h;
  return 0;
}
// Copies a pyramid level of the last extract; with_border → (w+38)x(h+38).
int oracle_get_pyramid(void* h, int level, uint8_t* dst, size_t dst_stride, int with_border) {
  Extractor* e = (Extractor*)h;
  if (level < 0 || level >= (int)e->pyr.size()) return -1;
  const Image& im = with_border ? e->pyrB[level] : e->pyr[level];
  for (int y = 0; y < im.h; y++) memcpy(dst + (size_t)y * dst_stride, im.row(y), im.w);
  return 0;
}
// Per-level intermediate lists of the last extract: stage 0 = FAST candidates in
// DistributeOctTree input order (coords relative to (16,16)), stage 1 = after the octree +
// orientation (level coords).  Returns the count; fills up to cap entries.
int oracle_level_keypoints(void* h, int level, int stage, void* kps, int cap) {
  Extractor* e = (Extractor*)h;
  const std::vector<KeyPoint>& v = stage == 0 ? e->candidates[level] : e->perLevel[level];
  int n = (int)std::min<size_t>(v.size(), (size_t)cap);
  if (kps && n) memcpy(kps, v.data(), (size_t)n * sizeof(KeyPoint));
  return (int)v.size();
}

// ---- primitives, exported one by one so tests can pin each against cv2 ----
void oracle_resize_linear(const uint8_t* src, int sw, int sh, uint8_t* dst, int dw, int dh) {
  Image s, d; s.w = sw; s.h = sh; s.d.assign(src, src + (size_t)sw * sh);
  resize_linear_u8(s, d, dw, dh);
  memcpy(dst, d.d.data(), d.d.size());
}
void oracle_gaussian_blur7(const uint8_t* src, int w, int h, uint8_t* dst) {
  Image s, d; s.w = w; s.h = h; s.d.assign(src, src + (size_t)w * h);
  gaussian_blur7(s, d);
  memcpy(dst, d.d.data(), d.d.size());
}
void oracle_border101(const uint8_t* src, int w, int h, int b, uint8_t* dst) {
  Image s, d; s.w = w; s.h = h; s.d.assign(src, src + (size_t)w * h);
  copy_make_border101(s, d, b);
  memcpy(dst, d.d.data(), d.d.size());
}
float oracle_fast_atan2(float y, float x) { return fast_atan2(y, x); }
void oracle_fast_atan2_array(const float* y, const float* x, float* out, int n) {
  for (int i = 0; i < n; i++) out[i] = fast_atan2(y[i], x[i]);
}
// FAST on ROI [x0,x1)x[y0,y1); out = (x,y,score) int triples, ROI-relative. Returns count.
int oracle_fast(const uint8_t* img, int w, int h, int x0, int y0, int x1, int y1, int threshold, int* out,
                int cap) {
  Image s; s.w = w; s.h = h; s.d.assign(img, img + (size_t)w * h);
  std::vector<Corner> c;
  fast9_16_nms(s, x0, y0, x1, y1, threshold, c);
  int n = (int)std::min<size_t>(c.size(), (size_t)cap);
  for (int i = 0; i < n; i++) { out[3 * i] = c[i].x; out[3 * i + 1] = c[i].y; out[3 * i + 2] = c[i].score; }
  return (int)c.size();
}
// DistributeOctTree on explicit (x,y,response) float triples; returns number kept, writes the
// kept indices-in-input (first index matching x,y) in list order.
int oracle_distribute(const float* xyr, int n, int minX, int maxX, int minY, int maxY, int N, float* out_xyr) {
  std::vector<KeyPoint> v(n);
  for (int i = 0; i < n; i++) v[i] = KeyPoint{xyr[3 * i], xyr[3 * i + 1], 7.f, -1.f, xyr[3 * i + 2], 0, -1};
  std::vector<KeyPoint> r = DistributeOctTree(v, minX, maxX, minY, maxY, N);
  for (size_t i = 0; i < r.size(); i++) { out_xyr[3 * i] = r[i].x; out_xyr[3 * i + 1] = r[i].y; out_xyr[3 * i + 2] = r[i].response; }
  return (int)r.size();
}
// libm sincosf over an array of angles in DEGREES as computeOrbDescriptor forms them (:111-112).
void oracle_sincos_deg(const float* angle_deg, float* cos_out, float* sin_out, long n) {
  const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
  for (long i = 0; i < n; i++) { float a = angle_deg[i] * factorPI; sincosf(a, &sin_out[i], &cos_out[i]); }
}
// libm sincosf over raw float bit patterns [lo, hi] with a stride (exhaustive sweeps).
// Per-chunk, order-independent 64-bit checksums of libm sincosf over `count` consecutive float bit patterns from
// `lo` (the device test hook orbx_debug_sincos_checksum computes the same function of its own sincosf).
void oracle_sincos_checksum(uint32_t lo, uint64_t count, uint32_t chunk, uint64_t* out, int threads) {
  const uint64_t nchunks = (count + chunk - 1) / chunk;
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; t++)
    pool.emplace_back([=]() {
      for (uint64_t k = t; k < nchunks; k += threads) {
        uint64_t acc = 0;
        const uint64_t i1 = std::min<uint64_t>(count, (k + 1) * chunk);
        for (uint64_t i = k * chunk; i < i1; i++) {
          const uint32_t bits = lo + (uint32_t)i;
          float y, s, c;
          memcpy(&y, &bits, 4);
          sincosf(y, &s, &c);
          uint32_t sb, cb;
          memcpy(&sb, &s, 4); memcpy(&cb, &c, 4);
          uint64_t h = ((uint64_t)cb << 32) | sb;
          h = h * 0x9E3779B97F4A7C15ull + bits;
          acc += h ^ (h >> 29);
        }
        out[k] = acc;
      }
    });
  for (auto& th : pool) th.join();
}
void oracle_sincos_bits(uint32_t lo, uint32_t count, uint32_t stride, float* cos_out, float* sin_out) {
  for (uint32_t i = 0; i < count; i++) {
    uint32_t u = lo + i * stride; float a; memcpy(&a, &u, 4);
    sincosf(a, &sin_out[i], &cos_out[i]);
  }
}

int oracle_descriptor_distance(const uint8_t* a, const uint8_t* b) { return DescriptorDistance(a, b); }

// Candidate-list scoring, the shared core of the ORBmatcher Search* loops.
//   tie_rule 0: strict `<` — first minimum wins, second-best tracked with levels
//               (ORBmatcher.cc:84-120 SearchByProjection; :273-325 SearchByBoW)
//   tie_rule 1: `<=` — last minimum wins, no second (SearchForTriangulation :1017)
// cand_off[nq+1] / cand_idx: CSR candidate lists in the order the reference visits them.
// ref_level[nr]: octave of each reference keypoint.
void oracle_argmin2_csr(const uint8_t* q, int nq, const uint8_t* r, int nr, const int* cand_off,
                        const int* cand_idx, const int8_t* ref_level, int tie_rule, oracle_result* out) {
  (void)nr;
  for (int i = 0; i < nq; i++) {
    int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
    for (int c = cand_off[i]; c < cand_off[i + 1]; c++) {
      int idx = cand_idx[c];
      int dist = DescriptorDistance(q + (size_t)i * 32, r + (size_t)idx * 32);
      if (tie_rule == 0) {
        if (dist < bestDist) {
          bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel;
          bestLevel = ref_level ? ref_level[idx] : 0; bestIdx = idx;
        } else if (dist < bestDist2) {
          bestLevel2 = ref_level ? ref_level[idx] : 0; bestDist2 = dist;
        }
      } else {
        if (dist <= bestDist) { bestDist = dist; bestIdx = idx; bestLevel = ref_level ? ref_level[idx] : 0; }
      }
    }
    out[i].best_idx = bestIdx; out[i].best_dist = (int16_t)bestDist; out[i].second_dist = (int16_t)bestDist2;
    out[i].best_level = (int8_t)bestLevel; out[i].second_level = (int8_t)bestLevel2; out[i].pad = 0;
  }
}

// Frame::AssignFeaturesToGrid + Frame::GetFeaturesInArea (Frame.cc:385-416, 657-735) for the
// undistorted == distorted (k1==0) case: builds the 64x48 grid over [minX,maxX)x[minY,maxY) and
// answers nq area queries, emitting CSR candidate lists in the reference's visiting order.
// Returns total candidates (fills up to cap).
int oracle_grid_queries(const void* kps_v, int n, float mnMinX, float mnMaxX, float mnMinY, float mnMaxY,
                        const float* qx, const float* qy, const float* qr, const int* qminLevel,
                        const int* qmaxLevel, int nq, int* cand_off, int* cand_idx, int cap) {
  const KeyPoint* kps = (const KeyPoint*)kps_v;
  const int COLS = 64, ROWS = 48;  // Frame.h:44-45
  const float wInv = static_cast<float>(COLS) / static_cast<float>(mnMaxX - mnMinX);  // Frame.cc:171
  const float hInv = static_cast<float>(ROWS) / static_cast<float>(mnMaxY - mnMinY);
  std::vector<std::vector<int>> grid(COLS * ROWS);
  for (int i = 0; i < n; i++) {
    int px = (int)std::round((kps[i].x - mnMinX) * wInv);
    int py = (int)std::round((kps[i].y - mnMinY) * hInv);
    if (px < 0 || px >= COLS || py < 0 || py >= ROWS) continue;
    grid[px * ROWS + py].push_back(i);
  }
  int total = 0;
  for (int q = 0; q < nq; q++) {
    cand_off[q] = total;
    const float x = qx[q], y = qy[q], r = qr[q];
    const int minLevel = qminLevel[q], maxLevel = qmaxLevel[q];
    const int nMinCellX = std::max(0, (int)std::floor((x - mnMinX - r) * wInv));
    if (nMinCellX >= COLS) continue;
    const int nMaxCellX = std::min(COLS - 1, (int)std::ceil((x - mnMinX + r) * wInv));
    if (nMaxCellX < 0) continue;
    const int nMinCellY = std::max(0, (int)std::floor((y - mnMinY - r) * hInv));
    if (nMinCellY >= ROWS) continue;
    const int nMaxCellY = std::min(ROWS - 1, (int)std::ceil((y - mnMinY + r) * hInv));
    if (nMaxCellY < 0) continue;
    const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
    for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
      for (int iy = nMinCellY; iy <= nMaxCellY; iy++)
        for (int j : grid[ix * ROWS + iy]) {
          const KeyPoint& kp = kps[j];
          if (bCheckLevels) {
            if (kp.octave < minLevel) continue;
            if (maxLevel >= 0 && kp.octave > maxLevel) continue;
          }
          const float distx = kp.x - x, disty = kp.y - y;
          if (std::fabs(distx) < r && std::fabs(disty) < r) {
            if (total < cap) cand_idx[total] = j;
            total++;
          }
        }
  }
  cand_off[nq] = total;
  return total;
}

// Frame::ComputeStereoMatches (Frame.cc:811-981).  Pyramids are passed as per-level plain
// images (border-free, the reference indexes mvImagePyramid ROIs the same way).
// Outputs: best_idx_r/best_dist = Hamming stage result per left kp (-1/100 when no candidate
// passed), uRight/depth as the reference leaves them.  Returns number of stereo matches kept.
int oracle_stereo_matches(const void* kpsL_v, const uint8_t* descL, int nL, const void* kpsR_v,
                          const uint8_t* descR, int nR, const float* scaleFactors,
                          const float* invScaleFactors, int nlevels, const uint8_t* const* pyrL,
                          const uint8_t* const* pyrR, const int* pyrW, const int* pyrH, float mb, float mbf,
                          int* best_idx_r, int* best_dist, float* uRight, float* depth) {
  (void)nlevels;
  const KeyPoint* kL = (const KeyPoint*)kpsL_v;
  const KeyPoint* kR = (const KeyPoint*)kpsR_v;
  const int TH_HIGH = 100, TH_LOW = 50;
  const int thOrbDist = (TH_HIGH + TH_LOW) / 2;
  const int nRows = pyrH[0];
  std::vector<std::vector<size_t>> vRowIndices(nRows);
  for (int iR = 0; iR < nR; iR++) {
    const float kpY = kR[iR].y;
    const float r = 2.0f * scaleFactors[kR[iR].octave];
    const int maxr = (int)std::ceil(kpY + r), minr = (int)std::floor(kpY - r);
    for (int yi = minr; yi <= maxr; yi++)
      if (yi >= 0 && yi < nRows) vRowIndices[yi].push_back(iR);  // reference indexes unchecked
  }
  const float minZ = mb, minD = 0, maxD = mbf / minZ;
  std::vector<std::pair<int, int>> vDistIdx;
  for (int iL = 0; iL < nL; iL++) {
    uRight[iL] = -1.0f; depth[iL] = -1.0f; best_idx_r[iL] = -1; best_dist[iL] = TH_HIGH;
  }
  for (int iL = 0; iL < nL; iL++) {
    const KeyPoint& kpL = kL[iL];
    const int levelL = kpL.octave;
    const float vL = kpL.y, uL = kpL.x;
    const std::vector<size_t>& vCandidates = vRowIndices[(int)vL];
    if (vCandidates.empty()) continue;
    const float minU = uL - maxD, maxU = uL - minD;
    if (maxU < 0) continue;
    int bestDist = TH_HIGH; size_t bestIdxR = 0; bool any = false;
    for (size_t iC = 0; iC < vCandidates.size(); iC++) {
      const size_t iR = vCandidates[iC];
      const KeyPoint& kpR = kR[iR];
      if (kpR.octave < levelL - 1 || kpR.octave > levelL + 1) continue;
      const float uR = kpR.x;
      if (uR >= minU && uR <= maxU) {
        const int dist = DescriptorDistance(descL + (size_t)iL * 32, descR + (size_t)iR * 32);
        if (dist < bestDist) { bestDist = dist; bestIdxR = iR; any = true; }
      }
    }
    best_dist[iL] = bestDist; best_idx_r[iL] = any ? (int)bestIdxR : -1;
    if (bestDist < thOrbDist) {
      const float uR0 = kR[bestIdxR].x;
      const float scaleFactor = invScaleFactors[kpL.octave];
      const float scaleduL = std::round(kpL.x * scaleFactor);
      const float scaledvL = std::round(kpL.y * scaleFactor);
      const float scaleduR0 = std::round(uR0 * scaleFactor);
      const int w = 5, L = 5;
      const int lvl = kpL.octave;
      const int W = pyrW[lvl];
      const uint8_t* IL = pyrL[lvl];
      const uint8_t* IR = pyrR[lvl];
      int bestSad = INT_MAX, bestincR = 0;
      float vDists[2 * 5 + 1];
      const float iniu = scaleduR0 + L - w, endu = scaleduR0 + L + w + 1;
      if (iniu < 0 || endu >= W) continue;
      for (int incR = -L; incR <= +L; incR++) {
        // cv::norm(IL, IR, NORM_L1) over the 11x11 windows
        int sad = 0;
        for (int dy = -w; dy <= w; dy++)
          for (int dx = -w; dx <= w; dx++) {
            int a = IL[(size_t)((int)scaledvL + dy) * W + ((int)scaleduL + dx)];
            int b = IR[(size_t)((int)scaledvL + dy) * W + ((int)scaleduR0 + incR + dx)];
            sad += std::abs(a - b);
          }
        float dist = (float)sad;
        if (dist < bestSad) { bestSad = (int)dist; bestincR = incR; }
        vDists[L + incR] = dist;
      }
      if (bestincR == -L || bestincR == L) continue;
      const float dist1 = vDists[L + bestincR - 1], dist2 = vDists[L + bestincR], dist3 = vDists[L + bestincR + 1];
      const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2));
      if (deltaR < -1 || deltaR > 1) continue;
      float bestuR = scaleFactors[kpL.octave] * ((float)scaleduR0 + (float)bestincR + deltaR);
      float disparity = (uL - bestuR);
      if (disparity >= minD && disparity < maxD) {
        if (disparity <= 0) { disparity = 0.01; bestuR = uL - 0.01; }  // double literals, as Frame.cc:956-957
        depth[iL] = mbf / disparity;
        uRight[iL] = bestuR;
        vDistIdx.push_back(std::pair<int, int>(bestSad, iL));
      }
    }
  }
  if (vDistIdx.empty()) return 0;
  std::sort(vDistIdx.begin(), vDistIdx.end());
  const float median = (float)vDistIdx[vDistIdx.size() / 2].first;
  const float thDist = 1.5f * 1.4f * median;
  int kept = (int)vDistIdx.size();
  for (int i = (int)vDistIdx.size() - 1; i >= 0; i--) {
    if (vDistIdx[i].first < thDist) break;
    uRight[vDistIdx[i].second] = -1; depth[vDistIdx[i].second] = -1; kept--;
  }
  return kept;
}

// Candidate scoring of ORBmatcher::Fuse (ORBmatcher.cc:1148-1338): KeyFrame::GetFeaturesInArea
// (KeyFrame.cc:704-748) then the loop :1258-1308 (octave in [l-1,l], chi-square gate, first minimum).
// best_idx/best_dist per query (-1 / 256 when no candidate passes).
void oracle_fuse_candidates(const void* kps_v, const uint8_t* desc, int n, float mnMinX, float mnMaxX,
                            float mnMinY, float mnMaxY, const float* u_right, const float* invLevelSigma2,
                            const float* qx, const float* qy, const float* qr, const float* q_ur,
                            const int8_t* qlevel, const uint8_t* qdesc, int nq, int* best_idx, int* best_dist) {
  const KeyPoint* kps = (const KeyPoint*)kps_v;
  std::vector<int> none(nq, -1), off(nq + 1), idx;
  int tot = oracle_grid_queries(kps_v, n, mnMinX, mnMaxX, mnMinY, mnMaxY, qx, qy, qr, none.data(), none.data(), nq,
                                off.data(), nullptr, 0);
  idx.resize(tot + 1);
  oracle_grid_queries(kps_v, n, mnMinX, mnMaxX, mnMinY, mnMaxY, qx, qy, qr, none.data(), none.data(), nq, off.data(),
                      idx.data(), tot);
  for (int q = 0; q < nq; q++) {
    const int nPredictedLevel = qlevel[q];
    int bestDist = 256, bestIdx = -1;
    for (int c = off[q]; c < off[q + 1]; c++) {
      const int i = idx[c];
      const KeyPoint& kp = kps[i];
      const int kpLevel = kp.octave;
      if (kpLevel < nPredictedLevel - 1 || kpLevel > nPredictedLevel) continue;
      if (u_right && u_right[i] >= 0) {
        const float ex = qx[q] - kp.x, ey = qy[q] - kp.y, er = q_ur[q] - u_right[i];
        const float e2 = ex * ex + ey * ey + er * er;
        if (e2 * invLevelSigma2[kpLevel] > 7.8) continue;
      } else {
        const float ex = qx[q] - kp.x, ey = qy[q] - kp.y;
        const float e2 = ex * ex + ey * ey;
        if (e2 * invLevelSigma2[kpLevel] > 5.99) continue;
      }
      const int dist = DescriptorDistance(qdesc + (size_t)q * 32, desc + (size_t)i * 32);
      if (dist < bestDist) { bestDist = dist; bestIdx = i; }
    }
    best_idx[q] = bestIdx; best_dist[q] = bestDist;
  }
}

static inline int rot_bin(float a1, float a2);
static void ComputeThreeMaxima(const std::vector<int>* histo, const int L, int& ind1, int& ind2, int& ind3);
// ORBmatcher::SearchForInitialization (ORBmatcher.cc:648-763) on flat arrays; prev_matched ([n1][2]) is
// vbPrevMatched and is updated in place.  Returns nmatches; match12 = vnMatches12.
int oracle_search_for_initialization(const void* kps1_v, const uint8_t* desc1, int n1, const void* kps2_v,
                                     const uint8_t* desc2, int n2, float mnMinX, float mnMaxX, float mnMinY,
                                     float mnMaxY, float* prev_matched, int windowSize, float nnratio,
                                     int check_orientation, int* match12) {
  const KeyPoint* kps1 = (const KeyPoint*)kps1_v;
  const KeyPoint* kps2 = (const KeyPoint*)kps2_v;
  const int TH_LOW = 50, HISTO_LENGTH = 30;
  int nmatches = 0;
  for (int i = 0; i < n1; i++) match12[i] = -1;
  std::vector<int> rotHist[30];
  std::vector<int> vMatchedDistance(n2, INT_MAX), vnMatches21(n2, -1);
  std::vector<int> off(2), idx(n2 + 1);
  for (int i1 = 0; i1 < n1; i1++) {
    const int level1 = kps1[i1].octave;
    if (level1 > 0) continue;
    const float x = prev_matched[2 * i1], y = prev_matched[2 * i1 + 1], r = (float)windowSize;
    const int tot = oracle_grid_queries(kps2_v, n2, mnMinX, mnMaxX, mnMinY, mnMaxY, &x, &y, &r, &level1, &level1, 1,
                                        off.data(), idx.data(), n2);
    if (tot == 0) continue;
    int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
    for (int c = 0; c < tot; c++) {
      const int i2 = idx[c];
      const int dist = DescriptorDistance(desc1 + (size_t)i1 * 32, desc2 + (size_t)i2 * 32);
      if (vMatchedDistance[i2] <= dist) continue;
      if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
      else if (dist < bestDist2) { bestDist2 = dist; }
    }
    if (bestDist <= TH_LOW) {
      if (bestDist < (float)bestDist2 * nnratio) {
        if (vnMatches21[bestIdx2] >= 0) { match12[vnMatches21[bestIdx2]] = -1; nmatches--; }
        match12[i1] = bestIdx2; vnMatches21[bestIdx2] = i1; vMatchedDistance[bestIdx2] = bestDist;
        nmatches++;
        if (check_orientation) rotHist[rot_bin(kps1[i1].angle, kps2[bestIdx2].angle)].push_back(i1);
      }
    }
  }
  if (check_orientation) {
    int ind1 = -1, ind2 = -1, ind3 = -1;
    ComputeThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (int idx1 : rotHist[i])
        if (match12[idx1] >= 0) { match12[idx1] = -1; nmatches--; }
    }
  }
  for (int i1 = 0; i1 < n1; i1++)
    if (match12[i1] >= 0) { prev_matched[2 * i1] = kps2[match12[i1]].x; prev_matched[2 * i1 + 1] = kps2[match12[i1]].y; }
  return nmatches;
}

// ORBmatcher::SearchByProjection, whole function, single-camera case (Nleft == -1), on flat arrays.
//   mode 0 = (Frame&, const vector<MapPoint*>&, th, ...)      ORBmatcher.cc:43-142  (best + second, ratio test)
//   mode 1 = (Frame& CurrentFrame, const Frame& LastFrame...) ORBmatcher.cc:1676-1887 (best only, rotation histogram)
// The projection itself (pose * point, camera model) is host geometry and arrives precomputed per query:
// qx,qy = projected pixel, qr = search radius, [qmin,qmax] = octave band handed to GetFeaturesInArea,
// q_ur = projected right coordinate, q_flags bit0 = query takes part, bit1 = its MapPoint has
// Observations()>0 (so a feature it is written to becomes invisible to later queries, :88-90 / :1741-1743),
// q_angle = kpLF.angle.  feat_locked[j] != 0: F.mvpMapPoints[j] holds such a point on entry.
// Outputs: match[i] = feature written by query i (-1 none; for mode 1 also -1 when the rotation check
// erased it), feat_query[j] = query whose MapPoint F.mvpMapPoints[j] holds on return (-1 = untouched /
// erased).  Returns nmatches.
static void ComputeThreeMaxima(const std::vector<int>* histo, const int L, int& ind1, int& ind2, int& ind3) {
  int max1 = 0, max2 = 0, max3 = 0;  // ORBmatcher.cc:2012-2053
  for (int i = 0; i < L; i++) {
    const int s = (int)histo[i].size();
    if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
    else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
    else if (s > max3) { max3 = s; ind3 = i; }
  }
  if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
  else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
}

int oracle_search_by_projection(const void* kps_v, const uint8_t* desc, int n, float mnMinX, float mnMaxX,
                                float mnMinY, float mnMaxY, const float* u_right, const uint8_t* feat_locked,
                                const float* qx, const float* qy, const float* qr, const int8_t* qmin,
                                const int8_t* qmax, const uint8_t* qdesc, int nq, const float* q_ur,
                                const uint8_t* q_flags, const float* q_angle, int mode, float nnratio,
                                int check_orientation, int th_dist, int* match, int* feat_query) {
  const KeyPoint* kps = (const KeyPoint*)kps_v;
  const int COLS = 64, ROWS = 48, HISTO_LENGTH = 30;
  const int TH_HIGH = th_dist;  // 100, or ORBdist in the relocalisation variant (ORBmatcher.cc:1961)
  const float wInv = static_cast<float>(COLS) / static_cast<float>(mnMaxX - mnMinX);
  const float hInv = static_cast<float>(ROWS) / static_cast<float>(mnMaxY - mnMinY);
  std::vector<std::vector<int>> grid(COLS * ROWS);
  for (int i = 0; i < n; i++) {
    int px = (int)std::round((kps[i].x - mnMinX) * wInv), py = (int)std::round((kps[i].y - mnMinY) * hInv);
    if (px < 0 || px >= COLS || py < 0 || py >= ROWS) continue;
    grid[px * ROWS + py].push_back(i);
  }
  std::vector<uint8_t> locked(n, 0);  // F.mvpMapPoints[j] && F.mvpMapPoints[j]->Observations()>0
  for (int j = 0; j < n; j++) { locked[j] = feat_locked ? (feat_locked[j] != 0) : 0; feat_query[j] = -1; }
  std::vector<int> rotHist[30];
  const float factor = 1.0f / HISTO_LENGTH;
  int nmatches = 0;
  for (int i = 0; i < nq; i++) {
    match[i] = -1;
    const int fl = q_flags ? q_flags[i] : 3;
    if (!(fl & 1)) continue;
    const float x = qx[i], y = qy[i], r = qr[i];
    const int minLevel = qmin[i], maxLevel = qmax[i];
    // GetFeaturesInArea (Frame.cc:657-723)
    std::vector<int> vIndices;
    do {
      const int nMinCellX = std::max(0, (int)std::floor((x - mnMinX - r) * wInv));
      if (nMinCellX >= COLS) break;
      const int nMaxCellX = std::min(COLS - 1, (int)std::ceil((x - mnMinX + r) * wInv));
      if (nMaxCellX < 0) break;
      const int nMinCellY = std::max(0, (int)std::floor((y - mnMinY - r) * hInv));
      if (nMinCellY >= ROWS) break;
      const int nMaxCellY = std::min(ROWS - 1, (int)std::ceil((y - mnMinY + r) * hInv));
      if (nMaxCellY < 0) break;
      const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
      for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
        for (int iy = nMinCellY; iy <= nMaxCellY; iy++)
          for (int j : grid[ix * ROWS + iy]) {
            const KeyPoint& kp = kps[j];
            if (bCheckLevels) {
              if (kp.octave < minLevel) continue;
              if (maxLevel >= 0 && kp.octave > maxLevel) continue;
            }
            if (std::fabs(kp.x - x) < r && std::fabs(kp.y - y) < r) vIndices.push_back(j);
          }
    } while (0);
    if (vIndices.empty()) continue;
    int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
    for (int idx : vIndices) {
      if (locked[idx]) continue;
      if (u_right && q_ur && u_right[idx] > 0) {
        const float er = std::fabs(q_ur[i] - u_right[idx]);
        if (er > r) continue;
      }
      const int dist = DescriptorDistance(qdesc + (size_t)i * 32, desc + (size_t)idx * 32);
      if (mode == 0) {
        if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = kps[idx].octave; bestIdx = idx; }
        else if (dist < bestDist2) { bestLevel2 = kps[idx].octave; bestDist2 = dist; }
      } else {
        if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
      }
    }
    if (bestDist <= TH_HIGH) {
      if (mode == 0 && bestLevel == bestLevel2 && bestDist > nnratio * bestDist2) continue;
      feat_query[bestIdx] = i;
      if (fl & 2) locked[bestIdx] = 1; else locked[bestIdx] = 0;  // the feature now holds this query's point
      match[i] = bestIdx;
      nmatches++;
      if (mode == 1 && check_orientation) {
        float rot = q_angle[i] - kps[bestIdx].angle;
        if (rot < 0.0) rot += 360.0f;
        int bin = (int)std::round(rot * factor);
        if (bin == HISTO_LENGTH) bin = 0;
        rotHist[bin].push_back(bestIdx);
      }
    }
  }
  if (mode == 1 && check_orientation) {
    int ind1 = -1, ind2 = -1, ind3 = -1;
    ComputeThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
    for (int b = 0; b < HISTO_LENGTH; b++) {
      if (b == ind1 || b == ind2 || b == ind3) continue;
      for (int j : rotHist[b]) { feat_query[j] = -1; nmatches--; }  // pointer nulled, :1870-1876
    }
    // match[i] = -1 where query i's own histogram entry was erased
    for (int i = 0; i < nq; i++)
      if (match[i] >= 0) {
        float rot = q_angle[i] - kps[match[i]].angle;
        if (rot < 0.0) rot += 360.0f;
        int bin = (int)std::round(rot * factor);
        if (bin == HISTO_LENGTH) bin = 0;
        if (!(bin == ind1 || bin == ind2 || bin == ind3)) match[i] = -1;
      }
  }
  return nmatches;
}

// Frame::ComputeStereoFromRGBD (Frame.cc:984-1005).  depth_img is float32 [h][w]; kps_un are the
// undistorted keypoints (mvKeysUn).
void oracle_rgbd_depth(const void* kps_v, const void* kps_un_v, int n, const float* depth_img, int w, int h,
                       float mbf, float* uRight, float* depth) {
  const KeyPoint* k = (const KeyPoint*)kps_v;
  const KeyPoint* ku = (const KeyPoint*)kps_un_v;
  for (int i = 0; i < n; i++) {
    uRight[i] = -1; depth[i] = -1;
    const float v = k[i].y, u = k[i].x;
    const int vi = (int)v, ui = (int)u;  // Mat::at<float>(int,int) takes the truncated floats
    if (vi < 0 || vi >= h || ui < 0 || ui >= w) continue;  // reference indexes unchecked
    const float d = depth_img[(size_t)vi * w + ui];
    if (d > 0) { depth[i] = d; uRight[i] = ku[i].x - mbf / d; }
  }
}

// cv::BFMatcher(NORM_HAMMING).knnMatch(q, r, k=2) (Frame.cc:1144): per query the two smallest
// distances, ties to the lower train index.
void oracle_knn2_dense(const uint8_t* q, int nq, const uint8_t* r, int nr, int* idx2, int* dist2) {
  for (int i = 0; i < nq; i++) {
    int b0 = INT_MAX, b1 = INT_MAX, i0 = -1, i1 = -1;
    for (int j = 0; j < nr; j++) {
      int d = DescriptorDistance(q + (size_t)i * 32, r + (size_t)j * 32);
      if (d < b0) { b1 = b0; i1 = i0; b0 = d; i0 = j; }
      else if (d < b1) { b1 = d; i1 = j; }
    }
    idx2[2 * i] = i0; idx2[2 * i + 1] = i1;
    dist2[2 * i] = i0 < 0 ? -1 : b0; dist2[2 * i + 1] = i1 < 0 ? -1 : b1;
  }
}

// ---------------------------------------------------------------------------------------------
// Vocabulary-node matching: ORBmatcher::SearchByBoW ×2 and SearchForTriangulation on flat arrays.
// A DBoW2::FeatureVector (std::map<NodeId, vector<unsigned>>) is a CSR: fv_node[nn] ascending node
// ids, fv_off[nn+1], fv_idx[] feature indices of each node in insertion (= ascending) order.
// The merge walk with lower_bound (ORBmatcher.cc:244-379) visits exactly the node ids present in
// both maps, in ascending order.
// ---------------------------------------------------------------------------------------------
static inline int rot_bin(float a1, float a2) {  // ORBmatcher.cc:332-338 and twins
  const float factor = 1.0f / 30;
  float rot = a1 - a2;
  if (rot < 0.0) rot += 360.0f;
  int bin = (int)std::round(rot * factor);
  if (bin == 30) bin = 0;
  return bin;
}

// mode 0: SearchByBoW(KeyFrame*, Frame&, ...) single-camera (F.Nleft == -1), ORBmatcher.cc:223-425.
//         side 1 = keyframe (valid1[i] = pMP && !pMP->isBad()), side 2 = frame; best <= TH_LOW.
// mode 1: SearchByBoW(KeyFrame*, KeyFrame*, ...) (NLeft == -1), ORBmatcher.cc:765-905; valid2[j] =
//         pMP2 && !pMP2->isBad(); best < TH_LOW.
// match12[i] = side-2 feature matched to side-1 feature i (-1 none / erased by the rotation check),
// match21[j] = the inverse (mode 0: this is vpMapPointMatches; mode 1: match12 is vpMatches12).
int oracle_search_by_bow(int mode, const void* kps1_v, const uint8_t* desc1, int n1, const uint8_t* valid1,
                         const int* fv1_node, const int* fv1_off, const int* fv1_idx, int nn1,
                         const void* kps2_v, const uint8_t* desc2, int n2, const uint8_t* valid2,
                         const int* fv2_node, const int* fv2_off, const int* fv2_idx, int nn2, float nnratio,
                         int check_orientation, int* match12, int* match21) {
  const KeyPoint* kps1 = (const KeyPoint*)kps1_v;
  const KeyPoint* kps2 = (const KeyPoint*)kps2_v;
  const int TH_LOW = 50, HISTO_LENGTH = 30;
  for (int i = 0; i < n1; i++) match12[i] = -1;
  for (int j = 0; j < n2; j++) match21[j] = -1;
  std::vector<int> rotHist[30];
  int nmatches = 0, a = 0, b = 0;
  while (a < nn1 && b < nn2) {
    if (fv1_node[a] == fv2_node[b]) {
      for (int i1 = fv1_off[a]; i1 < fv1_off[a + 1]; i1++) {
        const int idx1 = fv1_idx[i1];
        if (valid1 && !valid1[idx1]) continue;
        int bestDist1 = 256, bestIdx2 = -1, bestDist2 = 256;
        for (int i2 = fv2_off[b]; i2 < fv2_off[b + 1]; i2++) {
          const int idx2 = fv2_idx[i2];
          if (match21[idx2] >= 0) continue;            // vpMapPointMatches[realIdxF] / vbMatched2[idx2]
          if (mode == 1 && valid2 && !valid2[idx2]) continue;
          const int dist = DescriptorDistance(desc1 + (size_t)idx1 * 32, desc2 + (size_t)idx2 * 32);
          if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdx2 = idx2; }
          else if (dist < bestDist2) { bestDist2 = dist; }
        }
        const bool pass = mode == 0 ? bestDist1 <= TH_LOW : bestDist1 < TH_LOW;
        if (pass && static_cast<float>(bestDist1) < nnratio * static_cast<float>(bestDist2)) {
          match12[idx1] = bestIdx2; match21[bestIdx2] = idx1;
          if (check_orientation) rotHist[rot_bin(kps1[idx1].angle, kps2[bestIdx2].angle)].push_back(idx1);
          nmatches++;
        }
      }
      a++; b++;
    } else if (fv1_node[a] < fv2_node[b]) {
      a = (int)(std::lower_bound(fv1_node, fv1_node + nn1, fv2_node[b]) - fv1_node);
    } else {
      b = (int)(std::lower_bound(fv2_node, fv2_node + nn2, fv1_node[a]) - fv2_node);
    }
  }
  if (check_orientation) {
    int ind1 = -1, ind2 = -1, ind3 = -1;
    ComputeThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (int idx1 : rotHist[i]) { match21[match12[idx1]] = -1; match12[idx1] = -1; nmatches--; }
    }
  }
  return nmatches;
}

// ORBmatcher::SearchForTriangulation, pinhole single-camera case (mpCamera2 == NULL, NLeft == -1),
// ORBmatcher.cc:907-1146 with Pinhole::epipolarConstrain (CameraModels/Pinhole.cpp:107-129).
// Host geometry arrives precomputed per keyframe pair: F12 = K1^-T [t12]x R12 K2^-1 (row-major 3x3),
// ep = epipole of KF1's centre in KF2.  has_mp*[i] = GetMapPoint(i) != NULL, u_right* = mvuRight (NULL:
// monocular, every value < 0).  scale2 / sigma2_2 = pKF2->mvScaleFactors / mvLevelSigma2.
// match12[i] = vMatches12 after the rotation check.  Returns nmatches.
int oracle_search_for_triangulation(const void* kps1_v, const uint8_t* desc1, int n1, const uint8_t* has_mp1,
                                    const float* u_right1, const int* fv1_node, const int* fv1_off,
                                    const int* fv1_idx, int nn1, const void* kps2_v, const uint8_t* desc2, int /*n2*/,
                                    const uint8_t* has_mp2, const float* u_right2, const int* fv2_node,
                                    const int* fv2_off, const int* fv2_idx, int nn2, const float* F12,
                                    const float* ep, const float* scale2, const float* sigma2_2, int only_stereo,
                                    int coarse, int check_orientation, int* match12) {
  const KeyPoint* kps1 = (const KeyPoint*)kps1_v;
  const KeyPoint* kps2 = (const KeyPoint*)kps2_v;
  const int TH_LOW = 50, HISTO_LENGTH = 30;
  for (int i = 0; i < n1; i++) match12[i] = -1;
  std::vector<int> rotHist[30];
  int nmatches = 0, a = 0, b = 0;
  while (a < nn1 && b < nn2) {
    if (fv1_node[a] == fv2_node[b]) {
      for (int i1 = fv1_off[a]; i1 < fv1_off[a + 1]; i1++) {
        const int idx1 = fv1_idx[i1];
        if (has_mp1 && has_mp1[idx1]) continue;
        const bool bStereo1 = u_right1 && u_right1[idx1] >= 0;
        if (only_stereo && !bStereo1) continue;
        const KeyPoint& kp1 = kps1[idx1];
        int bestDist = TH_LOW, bestIdx2 = -1;
        for (int i2 = fv2_off[b]; i2 < fv2_off[b + 1]; i2++) {
          const int idx2 = fv2_idx[i2];
          if (has_mp2 && has_mp2[idx2]) continue;       // vbMatched2 is never set in the reference
          const bool bStereo2 = u_right2 && u_right2[idx2] >= 0;
          if (only_stereo && !bStereo2) continue;
          const int dist = DescriptorDistance(desc1 + (size_t)idx1 * 32, desc2 + (size_t)idx2 * 32);
          if (dist > TH_LOW || dist > bestDist) continue;
          const KeyPoint& kp2 = kps2[idx2];
          if (!bStereo1 && !bStereo2) {
            const float distex = ep[0] - kp2.x;
            const float distey = ep[1] - kp2.y;
            if (distex * distex + distey * distey < 100 * scale2[kp2.octave]) continue;
          }
          bool ok = coarse != 0;
          if (!ok) {  // Pinhole::epipolarConstrain
            const float la = kp1.x * F12[0] + kp1.y * F12[3] + F12[6];
            const float lb = kp1.x * F12[1] + kp1.y * F12[4] + F12[7];
            const float lc = kp1.x * F12[2] + kp1.y * F12[5] + F12[8];
            const float num = la * kp2.x + lb * kp2.y + lc;
            const float den = la * la + lb * lb;
            if (den == 0) ok = false;
            else { const float dsqr = num * num / den; ok = dsqr < 3.84 * sigma2_2[kp2.octave]; }
          }
          if (ok) { bestIdx2 = idx2; bestDist = dist; }
        }
        if (bestIdx2 >= 0) {
          match12[idx1] = bestIdx2; nmatches++;
          if (check_orientation) rotHist[rot_bin(kp1.angle, kps2[bestIdx2].angle)].push_back(idx1);
        }
      }
      a++; b++;
    } else if (fv1_node[a] < fv2_node[b]) {
      a = (int)(std::lower_bound(fv1_node, fv1_node + nn1, fv2_node[b]) - fv1_node);
    } else {
      b = (int)(std::lower_bound(fv2_node, fv2_node + nn2, fv1_node[a]) - fv2_node);
    }
  }
  if (check_orientation) {
    int ind1 = -1, ind2 = -1, ind3 = -1;
    ComputeThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (int idx1 : rotHist[i]) { match12[idx1] = -1; nmatches--; }
    }
  }
  return nmatches;
}

// DBoW2::TemplatedVocabulary<FORB::TDescriptor,FORB>::transform(features, BowVector&, FeatureVector&, levelsup)
// (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1127-1259; FORB::distance FORB.cpp:82-100; BowVector.cpp:33-88;
// FeatureVector.cpp:31-45) for TF_IDF weighting + L1 norm (the ORBvoc.txt header "10 6 0 0").  The tree is flat:
// node i has children child_idx[child_off[i] .. child_off[i+1]) (a leaf has none), descriptor node_desc[i*32],
// word id node_word[i] and weight node_weight[i] (leaves).  Outputs per feature: word / node / weight; the
// BowVector as (bow_word ascending, bow_val normalised), and the FeatureVector CSR.  Returns the number of
// BowVector entries; *fv_nn = number of FeatureVector nodes.
// PINNED: oracle/_ref/libref_dbow.so compiles the reference's own Thirdparty/DBoW2 unmodified (oracle/Makefile target
// `ref_dbow`); tests/test_ref_dbow.py asserts array_equal with it, the fp64 BowVector values included.
int oracle_bow_transform(const uint8_t* desc, int n, const int* child_off, const int* child_idx,
                         const uint8_t* node_desc, const int* node_word, const double* node_weight, int L,
                         int levelsup, int* feat_word, int* feat_node, double* feat_weight, int* bow_word,
                         double* bow_val, int* fv_node, int* fv_off, int* fv_idx, int* fv_nn) {
  std::map<unsigned, double> v;
  std::map<unsigned, std::vector<unsigned>> fv;
  const int nid_level = L - levelsup;
  for (int f = 0; f < n; f++) {
    unsigned nid = 0, final_id = 0;
    int current_level = 0;
    do {
      ++current_level;
      const int c0 = child_off[final_id], c1 = child_off[final_id + 1];
      final_id = child_idx[c0];
      double best_d = DescriptorDistance(desc + (size_t)f * 32, node_desc + (size_t)final_id * 32);
      for (int c = c0 + 1; c < c1; c++) {
        const unsigned id = child_idx[c];
        const double d = DescriptorDistance(desc + (size_t)f * 32, node_desc + (size_t)id * 32);
        if (d < best_d) { best_d = d; final_id = id; }
      }
      if (current_level == nid_level) nid = final_id;
    } while (child_off[final_id] != child_off[final_id + 1]);
    const double w = node_weight[final_id];
    feat_word[f] = node_word[final_id]; feat_node[f] = (int)nid; feat_weight[f] = w;
    if (w > 0) { v[(unsigned)node_word[final_id]] += w; fv[nid].push_back((unsigned)f); }
  }
  double norm = 0.0;
  for (auto& kv : v) norm += fabs(kv.second);
  if (norm > 0.0) for (auto& kv : v) kv.second /= norm;
  int k = 0;
  for (auto& kv : v) { bow_word[k] = (int)kv.first; bow_val[k] = kv.second; k++; }
  int nn = 0, pos = 0;
  for (auto& kv : fv) {
    fv_node[nn] = (int)kv.first; fv_off[nn] = pos;
    for (unsigned i : kv.second) fv_idx[pos++] = (int)i;
    nn++;
  }
  fv_off[nn] = pos; *fv_nn = nn;
  return k;
}

// MapPoint::ComputeDistinctiveDescriptors core (MapPoint.cc:367-401): all-pairs distances, per row the
// median = sorted[0.5*(N-1)], pick the row with the least median (first wins).  Returns the index.
int oracle_distinctive_descriptor(const uint8_t* desc, int n) {
  std::vector<std::vector<int>> D(n, std::vector<int>(n, 0));
  for (int i = 0; i < n; i++)
    for (int j = i + 1; j < n; j++) { int d = DescriptorDistance(desc + (size_t)i * 32, desc + (size_t)j * 32); D[i][j] = d; D[j][i] = d; }
  int BestMedian = INT_MAX, BestIdx = 0;
  for (int i = 0; i < n; i++) {
    std::vector<int> vDists(D[i].begin(), D[i].end());
    std::sort(vDists.begin(), vDists.end());
    int median = vDists[(size_t)(0.5 * (n - 1))];
    if (median < BestMedian) { BestMedian = median; BestIdx = i; }
  }
  return BestIdx;
}

// Multithreaded CPU baseline: runs `extract` over n_frames frames (one frame per task, T
// threads, one Extractor per thread) and returns wall seconds.  Frames are w*h tightly packed.
double oracle_bench_extract(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh, int flags,
                            const uint8_t* frames, int n_frames, int w, int h, int lap0, int lap1, int threads,
                            int* total_kps) {
  std::vector<std::thread> pool;
  std::vector<long> counts(threads, 0);
  auto t0 = std::chrono::steady_clock::now();
  for (int t = 0; t < threads; t++)
    pool.emplace_back([&, t]() {
      Extractor e(nfeatures, scaleFactor, nlevels, iniTh, minTh, flags);
      int cap = nfeatures + 3 * nlevels + 64;
      std::vector<KeyPoint> k(cap); std::vector<uint8_t> d((size_t)cap * 32);
      for (int f = t; f < n_frames; f += threads) {
        int n = 0;
        e.extract(frames + (size_t)f * w * h, w, h, w, lap0, lap1, k.data(), d.data(), cap, &n);
        counts[t] += n;
      }
    });
  for (auto& th : pool) th.join();
  auto t1 = std::chrono::steady_clock::now();
  long tot = 0; for (long c : counts) tot += c;
  if (total_kps) *total_kps = (int)tot;
  return std::chrono::duration<double>(t1 - t0).count();
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// CPU baseline of the whole benchmark step (bench.py cpu_baseline / --impl reference):
// extract every frame (one frame per task, `threads` workers, one Extractor per worker), then
// match frame f against frame f-1 (mod n) the way bench.py's GPU step does: last-frame keypoints
// shifted by the known image motion, window th*scale[octave], octave band [o-1,o+1], 64x48 grid
// lookup (Frame::GetFeaturesInArea) and the first-min loop of SearchByProjection(Frame,Frame)
// (ORBmatcher.cc:1676-1768).  Returns wall seconds; *total_matches counts best<=TH_HIGH.
// ---------------------------------------------------------------------------------------------
extern "C" double oracle_bench_pipeline(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh,
                                        int flags, const uint8_t* frames, int n_frames, int w, int h, int lap0,
                                        int lap1, const float* shift_xy, float th, int threads, int do_match,
                                        long* total_kps, long* total_matches, double* stage_out) {
  const int cap = nfeatures + 3 * nlevels + 64;
  std::vector<std::vector<KeyPoint>> K(n_frames);
  std::vector<std::vector<uint8_t>> D(n_frames);
  std::vector<float> sf(nlevels);
  // one Extractor per worker and the output buffers, built outside the timed region (the reference keeps its
  // extractors for the whole run, Tracking.cc:ParseORBParamFile)
  std::vector<std::unique_ptr<Extractor>> ex;
  for (int t = 0; t < threads; t++) ex.emplace_back(new Extractor(nfeatures, scaleFactor, nlevels, iniTh, minTh, flags));
  for (int l = 0; l < nlevels; l++) sf[l] = ex[0]->mvScaleFactor[l];
  for (int f = 0; f < n_frames; f++) { K[f].resize(cap); D[f].resize((size_t)cap * 32); }
  auto t0 = std::chrono::steady_clock::now();
  {
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; t++)
      pool.emplace_back([&, t]() {
        Extractor& e = *ex[t];
        for (int f = t; f < n_frames; f += threads) {
          K[f].resize(cap);
          int n = 0;
          e.extract(frames + (size_t)f * w * h, w, h, w, lap0, lap1, K[f].data(), D[f].data(), cap, &n);
          K[f].resize(n);
        }
      });
    for (auto& th_ : pool) th_.join();
  }
  auto t_mid = std::chrono::steady_clock::now();
  std::vector<long> matches(threads, 0);
  if (do_match) {
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; t++)
      pool.emplace_back([&, t]() {
        for (int f = t; f < n_frames; f += threads) {
          const int src = (f + n_frames - 1) % n_frames;
          const int nq = (int)K[src].size(), nr = (int)K[f].size();
          std::vector<float> qx(nq), qy(nq), qr(nq);
          std::vector<int> qmin(nq), qmax(nq), off(nq + 1), idx((size_t)nq * 256 + 16);
          for (int i = 0; i < nq; i++) {
            qx[i] = K[src][i].x + shift_xy[2 * f]; qy[i] = K[src][i].y + shift_xy[2 * f + 1];
            qr[i] = th * sf[K[src][i].octave];
            qmin[i] = K[src][i].octave - 1; qmax[i] = K[src][i].octave + 1;
          }
          int tot = oracle_grid_queries(K[f].data(), nr, 0.f, (float)w, 0.f, (float)h, qx.data(), qy.data(),
                                        qr.data(), qmin.data(), qmax.data(), nq, off.data(), idx.data(),
                                        (int)idx.size());
          if (tot > (int)idx.size()) {
            idx.resize(tot);
            oracle_grid_queries(K[f].data(), nr, 0.f, (float)w, 0.f, (float)h, qx.data(), qy.data(), qr.data(),
                                qmin.data(), qmax.data(), nq, off.data(), idx.data(), (int)idx.size());
          }
          std::vector<int8_t> lv(nr);
          for (int j = 0; j < nr; j++) lv[j] = (int8_t)K[f][j].octave;
          std::vector<oracle_result> res(nq);
          oracle_argmin2_csr(D[src].data(), nq, D[f].data(), nr, off.data(), idx.data(), lv.data(), 0, res.data());
          for (int i = 0; i < nq; i++) if (res[i].best_idx >= 0 && res[i].best_dist <= 100) matches[t]++;
        }
      });
    for (auto& th_ : pool) th_.join();
  }
  auto t1 = std::chrono::steady_clock::now();
  long tk = 0, tm = 0;
  for (int f = 0; f < n_frames; f++) tk += (long)K[f].size();
  for (long m : matches) tm += m;
  if (total_kps) *total_kps = tk;
  if (total_matches) *total_matches = tm;
  if (stage_out) {  // [0..5] per-stage seconds summed over the workers, [6] extract phase wall, [7] match phase wall
    for (int i = 0; i < 6; i++) { stage_out[i] = 0; for (auto& e : ex) stage_out[i] += e->stage_sec[i]; }
    stage_out[6] = std::chrono::duration<double>(t_mid - t0).count();
    stage_out[7] = std::chrono::duration<double>(t1 - t_mid).count();
  }
  return std::chrono::duration<double>(t1 - t0).count();
}
