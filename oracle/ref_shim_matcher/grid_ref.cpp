// grid_ref.cpp — the 64 x 48 feature grid of Frame / KeyFrame as the REFERENCE wrote it.  TEST INFRASTRUCTURE ONLY.
// `make ref_matcher` cuts
//   void Frame::AssignFeaturesToGrid()                 Frame.cc:385-416
//   vector<size_t> Frame::GetFeaturesInArea(...)       Frame.cc:657-723
//   bool Frame::PosInGrid(...)                         Frame.cc:725-735
//   vector<size_t> KeyFrame::GetFeaturesInArea(...)    KeyFrame.cc:704-748
// verbatim out of the reference files into a scratch directory outside the repository
// (oracle/extract_ref_functions.py) and this translation unit includes that text inside namespace ORB_SLAM3 over the
// stand-in classes of prelude.h.  The order in which GetFeaturesInArea returns the candidates decides every tie in the
// matcher loops, so the matcher pins (tests/test_ref_matcher.py) run on the reference's own grid, not on a restatement;
// tests/test_ref_frame.py also compares the grid and the area query directly with the oracle (row m10).
#include <algorithm>
#include <cmath>

#include "prelude.h"

namespace ORB_SLAM3 {
#include REF_GRID_FUNCTIONS_FRAME
#include REF_GRID_FUNCTIONS_KEYFRAME
}  // namespace ORB_SLAM3
