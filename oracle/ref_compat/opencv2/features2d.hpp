// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Stand-in for <opencv2/features2d.hpp>; see core.hpp.
// cv::FAST (src/ORBextractor.cc:800,821) forwards to the restated, cv2-pinned FAST-9/16 + NMS of
// oracle/orb_oracle.cpp.  KeyPointsFilter::retainBest is only referenced by the dead
// ComputeKeyPointsOld (:883-1037, its call at :1072 is commented out); it has to link, not run.
#pragma once
#include "core.hpp"

namespace cv {
void FAST(InputArray image, std::vector<KeyPoint>& keypoints, int threshold,
          bool nonmaxSuppression = true);

class KeyPointsFilter {
public:
  static void retainBest(std::vector<KeyPoint>& keypoints, int npoints);
};
}  // namespace cv
