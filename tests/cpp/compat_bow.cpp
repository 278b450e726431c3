// SearchByBoW (both variants) and SearchForTriangulation through the C++ drop-in helpers of
// include/orbx_compat/BowMatcher.h on std:: containers, as an ORBmatcher body would call them.
//   compat_bow <in.bin> <out_prefix>
// in.bin: int32 n1,n2,nn1,nn2 | per side: kps[n*28] desc[n*32] flag[n] uright[n*4] node[nn*4] off[(nn+1)*4] idx[n*4]
//         | F12[9] ep[2] scale[8] sigma2[8] (float32)
#include <stdio.h>
#include <stdlib.h>

#include <vector>

#include "orbx_compat/BowMatcher.h"

using namespace ORB_SLAM3;

struct Side {
  std::vector<cv::KeyPoint> keys;
  std::vector<unsigned char> desc, flag;
  std::vector<float> ur;
  FeatureVectorMap fv;
};
static bool rd(FILE* f, void* p, size_t n) { return n == 0 || fread(p, 1, n, f) == n; }
static bool read_side(FILE* f, int n, int nn, Side& s) {
  s.keys.resize(n); s.desc.resize((size_t)n * 32); s.flag.resize(n); s.ur.resize(n);
  std::vector<int> node(nn), off(nn + 1), idx(n);
  if (!rd(f, s.keys.data(), (size_t)n * 28) || !rd(f, s.desc.data(), s.desc.size()) || !rd(f, s.flag.data(), n) ||
      !rd(f, s.ur.data(), (size_t)n * 4) || !rd(f, node.data(), (size_t)nn * 4) || !rd(f, off.data(), (size_t)(nn + 1) * 4) ||
      !rd(f, idx.data(), (size_t)n * 4))
    return false;
  for (int a = 0; a < nn; a++)
    for (int i = off[a]; i < off[a + 1]; i++) s.fv[(unsigned)node[a]].push_back((unsigned)idx[i]);
  return true;
}
static void dump(const char* prefix, const char* ext, const void* p, size_t n) {
  char path[512];
  snprintf(path, sizeof path, "%s.%s", prefix, ext);
  FILE* f = fopen(path, "wb");
  if (n) fwrite(p, 1, n, f);
  fclose(f);
}

int main(int argc, char** argv) {
  if (argc < 3) return 2;
  FILE* f = fopen(argv[1], "rb");
  int hdr[4];
  if (!f || !rd(f, hdr, 16)) return 3;
  Side a, b;
  float geo[9 + 2 + 8 + 8];
  if (!read_side(f, hdr[0], hdr[2], a) || !read_side(f, hdr[1], hdr[3], b) || !rd(f, geo, sizeof geo)) return 4;
  fclose(f);
  std::vector<unsigned char> inv1(a.flag), inv2(b.flag);   // triangulation wants "has a MapPoint" = !valid here
  for (auto& v : inv1) v = !v;
  for (auto& v : inv2) v = !v;
  BowSide s1{&a.keys, a.desc.data(), &a.fv, &a.flag, nullptr}, s2{&b.keys, b.desc.data(), &b.fv, &b.flag, nullptr};
  std::vector<int> m12, m21;
  const int n0 = SearchByBoWDevice(ORBM_BOW_KF_FRAME, s1, s2, 0.7f, true, m12, m21);
  dump(argv[2], "bow0_m12", m12.data(), m12.size() * 4);
  dump(argv[2], "bow0_m21", m21.data(), m21.size() * 4);
  const int n1 = SearchByBoWDevice(ORBM_BOW_KF_KF, s1, s2, 0.8f, true, m12, m21);
  dump(argv[2], "bow1_m12", m12.data(), m12.size() * 4);
  BowSide t1{&a.keys, a.desc.data(), &a.fv, &inv1, &a.ur}, t2{&b.keys, b.desc.data(), &b.fv, &inv2, &b.ur};
  std::vector<std::pair<size_t, size_t> > pairs;
  std::vector<float> scale(geo + 11, geo + 19), sigma2(geo + 19, geo + 27);
  const int n2 = SearchForTriangulationDevice(t1, t2, geo, geo + 9, scale, sigma2, false, false, true, pairs);
  std::vector<int> flat;
  for (auto& pr : pairs) { flat.push_back((int)pr.first); flat.push_back((int)pr.second); }
  dump(argv[2], "tri_pairs", flat.data(), flat.size() * 4);
  printf("%d %d %d %d\n", n0, n1, n2, (int)pairs.size());
  return (n0 < 0 || n1 < 0 || n2 < 0) ? 1 : 0;
}
