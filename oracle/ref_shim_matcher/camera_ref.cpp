// camera_ref.cpp — the scalar camera functions of the path as the REFERENCE wrote them.  TEST INFRASTRUCTURE ONLY.
// `make ref_frame` cuts
//   Eigen::Vector2f Pinhole::project(const Eigen::Vector3f&)          Pinhole.cpp:49-55
//   Eigen::Vector2f KannalaBrandt8::project(const Eigen::Vector3f&)   KannalaBrandt8.cpp:67-85
//   cv::Point3f KannalaBrandt8::unproject(const cv::Point2f&)         KannalaBrandt8.cpp:116-146
//   Eigen::Vector3f KannalaBrandt8::unprojectEig(const cv::Point2f&)  KannalaBrandt8.cpp:111-114
// verbatim out of the reference files into a scratch directory outside the repository (oracle/extract_ref_functions.py)
// and this translation unit includes that text inside namespace ORB_SLAM3.  Name lookup as in the reference's
// translation units: no `using namespace std` is visible there (GeometricCamera.h / KannalaBrandt8.h do not reach
// DBoW2's TemplatedVocabulary.h), so `cos(psi)` / `sin(psi)` on a float are the double functions of <math.h> and the
// products with them are evaluated in double; `std::tan`, `atan2f`, `sqrtf`, `fminf` are what they say.  That is why
// this file switches prelude.h's trailing using-directive off (ORBX_PRELUDE_NO_USING_STD).
#include <cmath>
#include <vector>

#define ORBX_PRELUDE_NO_USING_STD
#include "prelude.h"

namespace ORB_SLAM3 {
#include REF_PINHOLE_FUNCTIONS
#include REF_KB8_FUNCTIONS
}  // namespace ORB_SLAM3

extern "C" {
// cam = fx, fy, cx, cy (pinhole: model 0) + k0..k3 (KannalaBrandt8: model 1); xyz [n][3] -> uv [n][2]
void ref_camera_project(int model, const float* cam, const float* xyz, int n, float* uv) {
  ORB_SLAM3::Pinhole ph;
  ORB_SLAM3::KannalaBrandt8 kb;
  ph.mvParameters.assign(cam, cam + 4);
  kb.mvParameters.assign(cam, cam + 8);
  for (int i = 0; i < n; i++) {
    const Eigen::Vector3f P(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
    const Eigen::Vector2f r = model == 0 ? ph.project(P) : kb.project(P);
    uv[2 * i] = r[0]; uv[2 * i + 1] = r[1];
  }
}
// KannalaBrandt8::unprojectEig: uv [n][2] -> ray [n][3]
void ref_kb8_unproject(const float* cam, const float* uv, int n, float* ray) {
  ORB_SLAM3::KannalaBrandt8 kb;
  kb.mvParameters.assign(cam, cam + 8);
  for (int i = 0; i < n; i++) {
    const Eigen::Vector3f r = kb.unprojectEig(cv::Point2f(uv[2 * i], uv[2 * i + 1]));
    ray[3 * i] = r[0]; ray[3 * i + 1] = r[1]; ray[3 * i + 2] = r[2];
  }
}
}  // extern "C"
