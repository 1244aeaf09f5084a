// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Stub of include/Frame.h: the data members (same
// names and types, statics included) that src/ORBmatcher.cc and the functions lifted from
// src/Frame.cc read and write.  The member FUNCTIONS declared here are defined by the reference's
// own bodies, lifted verbatim at build time (oracle/ref_extract.py -> oracle/_ref/gen/).
#pragma once
#include <climits>
#include <map>
#include <memory>
#include <mutex>
#include <vector>
#include <Eigen/Core>
#include <Thirdparty/DBoW2/DBoW2/FeatureVector.h>
#include <opencv2/core.hpp>
#include <opencv2/features2d.hpp>
#include <sophus/se3.hpp>

namespace ORB_SLAM3 {

#define FRAME_GRID_ROWS 48
#define FRAME_GRID_COLS 64

class GeometricCamera;
class KeyFrame;
class MapPoint;
class ORBextractor;

class Frame {
public:
  Frame() {}

  // --- defined by lifted reference bodies (src/Frame.cc) ---
  void ComputeStereoMatches();                                      // :944-1113
  void ComputeStereoFromRGBD(const cv::Mat& imDepth);               // :1115-1133
  void AssignFeaturesToGrid();                                      // :524-553
  bool PosInGrid(const cv::KeyPoint& kp, int& posX, int& posY);     // :867-877
  std::vector<std::size_t> GetFeaturesInArea(const float& x, const float& y, const float& r,
                                             const int minLevel = -1, const int maxLevel = -1,
                                             const bool bRight = false) const;  // :788-865
  bool isInFrustum(MapPoint* pMP, float viewingCosLimit);           // :644-721
  bool isInFrustumChecks(MapPoint* pMP, float viewingCosLimit, bool bRight = false);  // :1357-1437
  void SetPose(const Sophus::SE3<float>& Tcw);                      // :568-574
  void UpdatePoseMatrices();                                        // :608-614
  Sophus::SE3f GetRelativePoseTrl();                                // :628-630
  void ComputeStereoFishEyeMatches();                               // :1303-1355

  inline Sophus::SE3<float> GetPose() const { return mTcw; }        // Frame.h:183-186
  inline Eigen::Vector3f    GetCameraCenter() { return mOw; }

  // --- data (Frame.h:206-381) ---
  Sophus::SE3<float>         mTcw;
  Eigen::Matrix<float, 3, 3> mRwc;
  Eigen::Matrix<float, 3, 1> mOw;
  Eigen::Matrix<float, 3, 3> mRcw;
  Eigen::Matrix<float, 3, 1> mtcw;
  bool                       mbHasPose = false, mbIsSet = false;
  Sophus::SE3<float>         mTlr, mTrl;

  ORBextractor *mpORBextractorLeft = nullptr, *mpORBextractorRight = nullptr;

  static float fx, fy, cx, cy, invfx, invfy;
  float mbf = 0, mb = 0, mThDepth = 0;
  int   N = 0;

  std::vector<cv::KeyPoint> mvKeys, mvKeysRight, mvKeysUn;
  std::vector<MapPoint*>    mvpMapPoints;
  std::vector<float>        mvuRight, mvDepth;
  DBoW2::FeatureVector      mFeatVec;
  cv::Mat                   mDescriptors, mDescriptorsRight;
  std::vector<bool>         mvbOutlier;

  static float             mfGridElementWidthInv;
  static float             mfGridElementHeightInv;
  std::vector<std::size_t> mGrid[FRAME_GRID_COLS][FRAME_GRID_ROWS];

  int                mnScaleLevels = 0;
  float              mfScaleFactor = 0, mfLogScaleFactor = 0;
  std::vector<float> mvScaleFactors, mvInvScaleFactors, mvLevelSigma2, mvInvLevelSigma2;

  static float mnMinX, mnMaxX, mnMinY, mnMaxY;

  GeometricCamera *mpCamera = nullptr, *mpCamera2 = nullptr;
  int              Nleft = -1, Nright = -1;
  std::vector<int> mvLeftToRightMatch, mvRightToLeftMatch;
  std::vector<std::size_t> mGridRight[FRAME_GRID_COLS][FRAME_GRID_ROWS];
  // two-camera rigs (Frame.h:219-220, 284, 368-378)
  Eigen::Matrix<float, 3, 3>   mRlr;
  Eigen::Vector3f              mtlr;
  int                          mnCloseMPs = 0;
  int                          monoLeft = 0, monoRight = 0;
  static cv::BFMatcher         BFmatcher;
  std::vector<Eigen::Vector3f> mvStereo3Dpoints;
};

}  // namespace ORB_SLAM3
