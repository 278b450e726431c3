// The stereo half of Frame's constructor (src/Frame.cc:122-146) on the C++ drop-in classes: two extractors,
// then ComputeStereoMatches on the device-resident outputs.  Dumps mvuRight / mvDepth for the test.
//   compat_stereo <left.raw> <right.raw> <w> <h> <nfeatures> <mb> <mbf> <out_prefix>
#include <stdio.h>
#include <stdlib.h>

#include <vector>

#include "orbx_compat/ORBextractor.h"
#include "orbx_compat/StereoMatcher.h"

static bool load(const char* path, std::vector<unsigned char>& v) {
  FILE* f = fopen(path, "rb");
  const bool ok = f && fread(v.data(), 1, v.size(), f) == v.size();
  if (f) fclose(f);
  return ok;
}
static void dump(const char* prefix, const char* ext, const void* p, size_t n) {
  char path[512];
  snprintf(path, sizeof path, "%s.%s", prefix, ext);
  FILE* f = fopen(path, "wb");
  if (n) fwrite(p, 1, n, f);
  fclose(f);
}

int main(int argc, char** argv) {
  if (argc < 9) return 2;
  const int w = atoi(argv[3]), h = atoi(argv[4]), nf = atoi(argv[5]);
  const float mb = (float)atof(argv[6]), mbf = (float)atof(argv[7]);
  std::vector<unsigned char> l((size_t)w * h), r((size_t)w * h);
  if (!load(argv[1], l) || !load(argv[2], r)) return 3;
  cv::Mat imL(h, w, CV_8UC1, l.data(), (size_t)w), imR(h, w, CV_8UC1, r.data(), (size_t)w);
  ORB_SLAM3::ORBextractor left(nf, 1.2f, 8, 20, 7), right(nf, 1.2f, 8, 20, 7);
  std::vector<int> lap = {0, 0};  // Frame.cc:122-123
  std::vector<cv::KeyPoint> kL, kR;
  cv::Mat dL, dR;
  if (left(imL, cv::Mat(), kL, dL, lap) < 0 || right(imR, cv::Mat(), kR, dR, lap) < 0) return 4;
  std::vector<float> mvuRight, mvDepth;
  const int kept = ORB_SLAM3::ComputeStereoMatches(left, right, mb, mbf, mvuRight, mvDepth);
  if (kept < 0) { fprintf(stderr, "stereo failed: %d %s\n", kept, orbx_last_error()); return 5; }
  dump(argv[8], "kpsL", kL.data(), kL.size() * sizeof(cv::KeyPoint));
  dump(argv[8], "ur", mvuRight.data(), mvuRight.size() * sizeof(float));
  dump(argv[8], "depth", mvDepth.data(), mvDepth.size() * sizeof(float));
  printf("%d %d %d\n", (int)kL.size(), (int)kR.size(), kept);
  return 0;
}
