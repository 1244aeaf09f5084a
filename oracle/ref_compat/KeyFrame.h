// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Stub of include/KeyFrame.h: the members and accessors
// src/ORBmatcher.cc uses (same names and types; the reference's `const` data members are plain
// here so a harness can fill them).  GetFeaturesInArea / IsInImage are the reference's own bodies,
// lifted verbatim from src/KeyFrame.cc:780-837 at build time.
#pragma once
#include <map>
#include <mutex>
#include <set>
#include <vector>
#include <Eigen/Core>
#include <Thirdparty/DBoW2/DBoW2/FeatureVector.h>
#include <opencv2/core.hpp>
#include <sophus/se3.hpp>

namespace ORB_SLAM3 {

class Frame;
class GeometricCamera;
class MapPoint;

class KeyFrame {
public:
  KeyFrame() {}
  // --- lifted reference bodies ---
  std::vector<std::size_t> GetFeaturesInArea(const float& x, const float& y, const float& r,
                                             const bool bRight = false) const;
  bool IsInImage(const float& x, const float& y) const;

  // --- plain accessors ---
  Sophus::SE3f    GetPose() { return mTcw; }
  Sophus::SE3f    GetPoseInverse() { return mTcw.inverse(); }
  Eigen::Vector3f GetCameraCenter() { return mTcw.inverse().translation(); }
  Sophus::SE3<float> GetRightPose() { return mTrl * mTcw; }
  Sophus::SE3<float> GetRightPoseInverse() { return (mTrl * mTcw).inverse(); }
  Eigen::Vector3f    GetRightCameraCenter() { return (mTrl * mTcw).inverse().translation(); }
  std::vector<MapPoint*> GetMapPointMatches() { return mvpMapPoints; }
  std::set<MapPoint*>    GetMapPoints() {
    std::set<MapPoint*> s;
    for (MapPoint* p : mvpMapPoints) if (p) s.insert(p);
    return s;
  }
  MapPoint* GetMapPoint(const std::size_t& idx) { return mvpMapPoints[idx]; }
  void      AddMapPoint(MapPoint* pMP, const std::size_t& idx) { mvpMapPoints[idx] = pMP; }

  // --- data (KeyFrame.h:320-521) ---
  int   mnGridCols = 64, mnGridRows = 48;
  float mfGridElementWidthInv = 0, mfGridElementHeightInv = 0;
  float fx = 0, fy = 0, cx = 0, cy = 0, invfx = 0, invfy = 0, mbf = 0, mb = 0, mThDepth = 0;
  int   N = 0;
  std::vector<cv::KeyPoint> mvKeys, mvKeysUn, mvKeysRight;
  std::vector<float>        mvuRight, mvDepth;
  cv::Mat                   mDescriptors;
  DBoW2::FeatureVector      mFeatVec;
  int                mnScaleLevels = 0;
  float              mfScaleFactor = 0, mfLogScaleFactor = 0;
  std::vector<float> mvScaleFactors, mvLevelSigma2, mvInvLevelSigma2;
  int mnMinX = 0, mnMinY = 0, mnMaxX = 0, mnMaxY = 0;
  Sophus::SE3<float> mTcw, mTlr, mTrl;
  std::vector<MapPoint*> mvpMapPoints;
  std::vector<std::vector<std::vector<std::size_t>>> mGrid, mGridRight;
  GeometricCamera *mpCamera = nullptr, *mpCamera2 = nullptr;
  int NLeft = -1, NRight = -1;
};

}  // namespace ORB_SLAM3
