// mappoint_ref.cpp — the MapPoint functions the hot path calls, as the REFERENCE wrote them.  TEST INFRASTRUCTURE ONLY.
// `make ref_matcher` cuts
//   void MapPoint::ComputeDistinctiveDescriptors()        MapPoint.cc:328-402
//   float MapPoint::GetMinDistanceInvariance()            MapPoint.cc:498-502
//   float MapPoint::GetMaxDistanceInvariance()            MapPoint.cc:504-508
//   int MapPoint::PredictScale(const float&, KeyFrame*)   MapPoint.cc:510-525
//   int MapPoint::PredictScale(const float&, Frame*)      MapPoint.cc:527-542
// verbatim out of the reference file into a scratch directory outside the repository (oracle/extract_ref_functions.py)
// and this translation unit includes that text inside namespace ORB_SLAM3 over the stand-in classes of prelude.h and the
// reference's own include/ORBmatcher.h.  `log(ratio)` resolves as in the reference build: the float overload, through
// the global `using namespace std;` (TemplatedVocabulary.h:36 there, the end of prelude.h here).
#include <algorithm>
#include <climits>
#include <cmath>

#include "prelude.h"
#include "ORBmatcher.h"

namespace ORB_SLAM3 {
#include REF_MAPPOINT_FUNCTIONS
}  // namespace ORB_SLAM3
