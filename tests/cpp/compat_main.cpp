// Drives the C++ drop-in classes (include/orbx_compat) the way Frame::ExtractORB does
// (src/Frame.cc:418-425) and dumps the results for tests/test_gpu_compat_cpp.py to compare with
// the oracle.  Built with -DORBX_NO_OPENCV (no OpenCV headers in this image).
//   compat_main <in.raw> <w> <h> <nfeatures> <lap0> <lap1> <out_prefix>
#define ORBM_BATCH_IMPLEMENTATION
#include <stdio.h>
#include <stdlib.h>

#include <vector>

#include "orbx_compat/ORBextractor.h"
#include "orbx_compat/ORBmatcher_core.h"

static void dump(const char* path, const void* p, size_t n) {
  FILE* f = fopen(path, "wb");
  if (n) fwrite(p, 1, n, f);
  fclose(f);
}

int main(int argc, char** argv) {
  if (argc < 8) return 2;
  const int w = atoi(argv[2]), h = atoi(argv[3]), nf = atoi(argv[4]);
  std::vector<int> lap = {atoi(argv[5]), atoi(argv[6])};
  std::vector<unsigned char> raw((size_t)w * h);
  FILE* f = fopen(argv[1], "rb");
  if (!f || fread(raw.data(), 1, raw.size(), f) != raw.size()) return 3;
  fclose(f);
  cv::Mat im(h, w, CV_8UC1, raw.data(), (size_t)w);

  ORB_SLAM3::ORBextractor* ex = new ORB_SLAM3::ORBextractor(nf, 1.2f, 8, 20, 7);  // Tracking.cc:595
  std::vector<cv::KeyPoint> keys;
  cv::Mat desc;
  int mono = (*ex)(im, cv::Mat(), keys, desc, lap);
  if (mono < 0) { fprintf(stderr, "extract failed: %s\n", orbx_last_error()); return 4; }
  // second call on the same instance (mvImagePyramid is overwritten per call, like upstream)
  std::vector<cv::KeyPoint> keys2;
  cv::Mat desc2;
  int mono2 = (*ex)(im, cv::Mat(), keys2, desc2, lap);
  if (mono2 != mono || keys2.size() != keys.size()) return 5;

  char path[512];
  snprintf(path, sizeof path, "%s.kps", argv[7]);
  dump(path, keys.data(), keys.size() * sizeof(cv::KeyPoint));
  snprintf(path, sizeof path, "%s.desc", argv[7]);
  std::vector<unsigned char> d(keys.size() * 32);
  for (size_t i = 0; i < keys.size(); i++) memcpy(&d[i * 32], desc.ptr((int)i), 32);
  dump(path, d.data(), d.size());
  // pyramid level 3 with its border, read through the public field like Frame.cc:908
  const cv::Mat& p3 = ex->mvImagePyramid[3];
  std::vector<unsigned char> pb((size_t)(p3.rows + 38) * (p3.cols + 38));
  for (int y = -19; y < p3.rows + 19; y++)
    memcpy(&pb[(size_t)(y + 19) * (p3.cols + 38)], p3.data + (long)y * (long)p3.step - 19, p3.cols + 38);
  snprintf(path, sizeof path, "%s.pyr3", argv[7]);
  dump(path, pb.data(), pb.size());

  // matcher: DescriptorDistance + one batched scoring call (every keypoint against its 8 successors)
  int dd = ORB_SLAM3::ORBmatcherCore::DescriptorDistance(desc.row(0), desc.row(1));
  ORB_SLAM3::OrbmBatch batch;
  const int n = (int)keys.size();
  for (int i = 0; i < n; i++) {
    batch.begin_query(desc.ptr(i));
    for (int k = 1; k <= 8; k++) batch.add((i + k) % n);
  }
  std::vector<int8_t> lv(n);
  for (int i = 0; i < n; i++) lv[i] = (int8_t)keys[i].octave;
  std::vector<orbm_result> res;
  if (batch.run(d.data(), n, lv.data(), ORBM_TIE_FIRST_MIN, res) != ORBX_OK) return 6;
  snprintf(path, sizeof path, "%s.match", argv[7]);
  dump(path, res.data(), res.size() * sizeof(orbm_result));
  printf("%d %d %d %d %d\n", mono, (int)keys.size(), dd, ex->GetLevels(), (int)ex->GetScaleFactors().size());
  delete ex;
  return 0;
}
