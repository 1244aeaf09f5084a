// Drop-ins for the per-frame glue that follows ORBextractor::operator() in the reference's Frame
// constructors, templated over the Frame type so the same header compiles
// against the reference's class (member names below are the reference's):
//   void Frame::UndistortKeyPoints()                          src/Frame.cc:886-913
//   void Frame::ComputeStereoFromRGBD(const cv::Mat& imDepth) src/Frame.cc:1115-1133
// Compute is libcuorb's (bit-identical to cv::undistortPoints' double arithmetic); no CPU fallback.
#pragma once
#include <stdexcept>
#include <string>
#include <type_traits>
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
  if (rc < 0) CUORB_SHIM_FAIL(std::runtime_error, std::string("UndistortKeyPoints (libcuorb): ") + cuorb_last_error());
}

// reads F.mvKeys, F.mvKeysUn, F.mbf and imDepth (CV_32F); writes F.mvuRight, F.mvDepth
template <class FrameT>
void ComputeStereoFromRGBD(FrameT& F, const cv::Mat& imDepth, int device = 0) {
  const int n = (int)F.mvKeys.size();
  F.mvuRight.assign(n, -1.0f);
  F.mvDepth.assign(n, -1.0f);
  if (n == 0) return;
  if (imDepth.type() != CV_32F) CUORB_SHIM_FAIL(std::invalid_argument, "ComputeStereoFromRGBD: imDepth must be CV_32F");
  const int rc = cuorb_stereo_from_rgbd(device, n, (const cuorb_keypoint*)F.mvKeys.data(), (const cuorb_keypoint*)F.mvKeysUn.data(),
                                        imDepth.template ptr<float>(0), imDepth.cols, imDepth.rows, imDepth.step, CUORB_MEM_HOST,
                                        F.mbf, F.mvuRight.data(), F.mvDepth.data());
  if (rc < 0) CUORB_SHIM_FAIL(std::runtime_error, std::string("ComputeStereoFromRGBD (libcuorb): ") + cuorb_last_error());
}

}  // namespace ORB_SLAM3

// ---- dense all-pairs Hamming users ----
namespace ORB_SLAM3 {

// The matching half of void Frame::ComputeStereoFishEyeMatches() (src/Frame.cc:1304-1355):
// BFmatcher.knnMatch(stereoDescLeft, stereoDescRight, matches, 2) + the 0.7 ratio test (:1329).
// accepted[i] = trainIdx of the accepted match of left row i, or -1; the caller keeps the
// triangulation loop (:1331-1352, KannalaBrandt8::TriangulateMatches) on accepted rows.
inline void FisheyeStereoCandidates(const cv::Mat& stereoDescLeft, const cv::Mat& stereoDescRight,
                                    std::vector<int>& accepted, int device = 0) {
  accepted.assign(stereoDescLeft.rows, -1);
  if (stereoDescLeft.rows == 0 || stereoDescRight.rows == 0) return;
  if ((stereoDescLeft.rows > 1 && stereoDescLeft.step != 32) || (stereoDescRight.rows > 1 && stereoDescRight.step != 32))
    CUORB_SHIM_FAIL(std::invalid_argument, "FisheyeStereoCandidates: descriptor matrices must be n x 32 contiguous");
  const int32_t qo[2] = {0, stereoDescLeft.rows}, to[2] = {0, stereoDescRight.rows};
  std::vector<int32_t> idx(2 * (size_t)stereoDescLeft.rows), dist(idx.size());
  const int rc = cuorb_knn_match2(device, 1, qo, stereoDescLeft.ptr<unsigned char>(0), to, stereoDescRight.ptr<unsigned char>(0),
                                  idx.data(), dist.data(), accepted.data());
  if (rc < 0) CUORB_SHIM_FAIL(std::runtime_error, std::string("FisheyeStereoCandidates (libcuorb): ") + cuorb_last_error());
}

// void Frame::ComputeStereoFishEyeMatches() (src/Frame.cc:1303-1355), whole: the lapping-area split (:1305-1309), the
// 2-NN + 0.7 ratio test on the device (FisheyeStereoCandidates), then - on the accepted rows, in left-row order as the
// reference walks `matches` - the caller's own camera model triangulates (CameraT = KannalaBrandt8, out of scope here)
// and the stereo tables are filled exactly as at :1345-1351 (a right feature claimed by several left rows keeps the
// last one in mvRightToLeftMatch while every left row keeps its own entry, as in the reference).
template <class CameraT, class FrameT>
inline void ComputeStereoFishEyeMatches(FrameT& F, int device = 0) {
  F.mvLeftToRightMatch.assign(F.Nleft, -1);
  F.mvRightToLeftMatch.assign(F.Nright, -1);
  F.mvDepth.assign(F.Nleft, -1.0f);
  F.mvuRight.assign(F.Nleft, -1);
  typedef typename std::decay<decltype(F.mvStereo3Dpoints[0])>::type Vec3;
  F.mvStereo3Dpoints = std::vector<Vec3>(F.Nleft);
  F.mnCloseMPs = 0;
  const int nq = F.mDescriptors.rows - F.monoLeft, nt = F.mDescriptorsRight.rows - F.monoRight;
  if (nq <= 0 || nt <= 0) return;
  const cv::Mat stereoDescLeft  = F.mDescriptors.rowRange(F.monoLeft, F.mDescriptors.rows);
  const cv::Mat stereoDescRight = F.mDescriptorsRight.rowRange(F.monoRight, F.mDescriptorsRight.rows);
  std::vector<int> accepted;
  FisheyeStereoCandidates(stereoDescLeft, stereoDescRight, accepted, device);
  for (int q = 0; q < nq; q++) {
    if (accepted[q] < 0) continue;                                 // :1329
    const int iL = q + F.monoLeft, iR = accepted[q] + F.monoRight;
    const float sigma1 = F.mvLevelSigma2[F.mvKeys[iL].octave], sigma2 = F.mvLevelSigma2[F.mvKeysRight[iR].octave];
    Vec3 p3D;
    const float depth = static_cast<CameraT*>(F.mpCamera)->TriangulateMatches(F.mpCamera2, F.mvKeys[iL], F.mvKeysRight[iR], F.mRlr,
                                                                              F.mtlr, sigma1, sigma2, p3D);   // :1335-1344
    if (depth > 0.0001f) {                                         // :1345-1351
      F.mvLeftToRightMatch[iL] = iR;
      F.mvRightToLeftMatch[iR] = iL;
      F.mvStereo3Dpoints[iL]   = p3D;
      F.mvDepth[iL]            = depth;
    }
  }
}

// void MapPoint::ComputeDistinctiveDescriptors() (src/MapPoint.cc:376-450) for a batch of MapPoints
// (LocalMapping calls it once per new/fused point): vDescriptors[p] = the observed descriptor rows of
// point p in the reference's iteration order; returns BestIdx per point.
inline std::vector<int> DistinctiveDescriptorIndices(const std::vector<std::vector<cv::Mat>>& vDescriptors, int device = 0) {
  std::vector<int32_t> off(vDescriptors.size() + 1, 0);
  for (size_t p = 0; p < vDescriptors.size(); p++) off[p + 1] = off[p] + (int32_t)vDescriptors[p].size();
  std::vector<unsigned char> flat((size_t)off.back() * 32 + 32);
  for (size_t p = 0; p < vDescriptors.size(); p++)
    for (size_t k = 0; k < vDescriptors[p].size(); k++)
      std::memcpy(&flat[((size_t)off[p] + k) * 32], vDescriptors[p][k].ptr<unsigned char>(0), 32);
  std::vector<int> best(vDescriptors.size(), -1);
  if (vDescriptors.empty()) return best;
  const int rc = cuorb_distinctive_descriptors(device, (int)vDescriptors.size(), off.data(), flat.data(), best.data());
  if (rc < 0) CUORB_SHIM_FAIL(std::runtime_error, std::string("DistinctiveDescriptorIndices (libcuorb): ") + cuorb_last_error(), std::vector<int>());
  return best;
}

}  // namespace ORB_SLAM3
