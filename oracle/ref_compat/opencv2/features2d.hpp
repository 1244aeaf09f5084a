// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Stand-in for <opencv2/features2d.hpp>; see core.hpp.
// cv::FAST (src/ORBextractor.cc:800,821) forwards to the restated, cv2-pinned FAST-9/16 + NMS of
// oracle/orb_oracle.cpp.  KeyPointsFilter::retainBest is only referenced by the dead
// ComputeKeyPointsOld (:883-1037, its call at :1072 is commented out); it has to link, not run.
#pragma once
#include "core.hpp"

namespace cv {
void FAST(InputArray image, std::vector<KeyPoint>& keypoints, int threshold,
          bool nonmaxSuppression = true);

// cv::DMatch / cv::BFMatcher(NORM_HAMMING)::knnMatch as Frame::ComputeStereoFishEyeMatches uses them
// (src/Frame.cc:1322-1329); forwards to the cv2.BFMatcher-pinned 2-NN of oracle/bfmatch_oracle.cpp
struct DMatch {
  int   queryIdx = -1, trainIdx = -1, imgIdx = -1;
  float distance = 0;
};
class BFMatcher {
public:
  BFMatcher(int normType = NORM_HAMMING, bool crossCheck = false) { (void)normType; (void)crossCheck; }
  void knnMatch(const Mat& queryDescriptors, const Mat& trainDescriptors, std::vector<std::vector<DMatch>>& matches, int k) const;
};

class KeyPointsFilter {
public:
  static void retainBest(std::vector<KeyPoint>& keypoints, int npoints);
};
}  // namespace cv
