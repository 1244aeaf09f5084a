// Drop-ins for the per-frame glue that follows ORBextractor::operator() in the reference's Frame
// constructors, templated over the Frame type so the same header compiles
// against the reference's class (member names below are the reference's):
//   void Frame::UndistortKeyPoints()                          src/Frame.cc:886-913
//   void Frame::ComputeStereoFromRGBD(const cv::Mat& imDepth) src/Frame.cc:1115-1133
// Compute is libcuorb's (bit-identical to cv::undistortPoints' double arithmetic); no CPU fallback.
#pragma once
#include <stdexcept>
#include <string>
#include <vector>

#include "../include/cuorb.h"
#include "cv_compat.h"

namespace ORB_SLAM3 {

// reads F.mvKeys, F.mK (3x3 CV_32F), F.mDistCoef (4 or 5 x 1 CV_32F), F.N; writes F.mvKeysUn
template <class FrameT>
void UndistortKeyPoints(FrameT& F, int device = 0) {
  static_assert(sizeof(cv::KeyPoint) == sizeof(cuorb_keypoint), "cv::KeyPoint must be 7 x 4 bytes");
  F.mvKeysUn.resize(F.mvKeys.size());
  if (F.mvKeys.empty()) return;
  const float K4[4] = {F.mK.template at<float>(0, 0), F.mK.template at<float>(1, 1), F.mK.template at<float>(0, 2),
                       F.mK.template at<float>(1, 2)};
  float d5[5] = {0, 0, 0, 0, 0};
  const int nd = F.mDistCoef.rows * F.mDistCoef.cols;
  for (int i = 0; i < 5 && i < nd; i++) d5[i] = F.mDistCoef.rows > 1 ? F.mDistCoef.template at<float>(i, 0) : F.mDistCoef.template at<float>(0, i);
  const int rc = cuorb_undistort_keypoints(device, (int)F.mvKeys.size(), (const cuorb_keypoint*)F.mvKeys.data(), K4, d5,
                                           (cuorb_keypoint*)F.mvKeysUn.data());
  if (rc < 0) throw std::runtime_error(std::string("UndistortKeyPoints (libcuorb): ") + cuorb_last_error());
}

// reads F.mvKeys, F.mvKeysUn, F.mbf and imDepth (CV_32F); writes F.mvuRight, F.mvDepth
template <class FrameT>
void ComputeStereoFromRGBD(FrameT& F, const cv::Mat& imDepth, int device = 0) {
  const int n = (int)F.mvKeys.size();
  F.mvuRight.assign(n, -1.0f);
  F.mvDepth.assign(n, -1.0f);
  if (n == 0) return;
  if (imDepth.type() != CV_32F) throw std::invalid_argument("ComputeStereoFromRGBD: imDepth must be CV_32F");
  const int rc = cuorb_stereo_from_rgbd(device, n, (const cuorb_keypoint*)F.mvKeys.data(), (const cuorb_keypoint*)F.mvKeysUn.data(),
                                        imDepth.template ptr<float>(0), imDepth.cols, imDepth.rows, imDepth.step, CUORB_MEM_HOST,
                                        F.mbf, F.mvuRight.data(), F.mvDepth.data());
  if (rc < 0) throw std::runtime_error(std::string("ComputeStereoFromRGBD (libcuorb): ") + cuorb_last_error());
}

}  // namespace ORB_SLAM3
