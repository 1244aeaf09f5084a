// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Stub of include/MapPoint.h: the members and
// accessors src/ORBmatcher.cc and Frame::isInFrustum read, with plain storage behind them.
// PredictScale is the reference's own body (lifted from src/MapPoint.cc:557-589 at build time).
#pragma once
#include <map>
#include <mutex>
#include <tuple>
#include <Eigen/Core>
#include <opencv2/core.hpp>

namespace ORB_SLAM3 {

class Frame;
class KeyFrame;

class MapPoint {
public:
  Eigen::Vector3f GetWorldPos() { return mWorldPos; }
  Eigen::Vector3f GetNormal() { return mNormalVector; }
  int             Observations() { return nObs; }
  bool            isBad() { return mbBad; }
  cv::Mat         GetDescriptor() { return mDescriptor.clone(); }
  float           GetMinDistanceInvariance() { return 0.8f * mfMinDistance; }
  float           GetMaxDistanceInvariance() { return 1.2f * mfMaxDistance; }
  float           GetMinDistance() { return mfMinDistance; }   // the two getters INTEGRATION.md asks a maintainer to add
  float           GetMaxDistance() { return mfMaxDistance; }
  void            IncreaseVisible(int n = 1) { mnVisible += n; }
  int             PredictScale(const float& currentDist, KeyFrame* pKF);
  int             PredictScale(const float& currentDist, Frame* pF);
  std::tuple<int, int> GetIndexInKeyFrame(KeyFrame* pKF) {
    auto it = mObservations.find(pKF);
    return it != mObservations.end() ? it->second : std::tuple<int, int>(-1, -1);
  }
  bool IsInKeyFrame(KeyFrame* pKF) { return mObservations.count(pKF) != 0; }
  // bookkeeping calls of Fuse / SearchByProjection: recorded, so a harness can read them back
  void AddObservation(KeyFrame* pKF, int idx) {
    std::tuple<int, int> t(-1, -1);
    if (mObservations.count(pKF)) t = mObservations[pKF];
    std::get<0>(t) = idx;
    mObservations[pKF] = t;
    nObs++;
  }
  void      Replace(MapPoint* pMP) { mpReplaced = pMP; mbBad = true; }
  MapPoint* GetReplaced() { return mpReplaced; }

  // tracking scratch written by Frame::isInFrustum, read by SearchByProjection (MapPoint.h:166-175)
  float mTrackProjX = -1, mTrackProjY = -1, mTrackDepth = 0, mTrackDepthR = 0, mTrackProjXR = -1,
        mTrackProjYR = -1;
  bool  mbTrackInView = false, mbTrackInViewR = false;
  int   mnTrackScaleLevel = 0, mnTrackScaleLevelR = 0;
  float mTrackViewCos = 0, mTrackViewCosR = 0;
  long unsigned int mnTrackReferenceForFrame = 0, mnLastFrameSeen = 0, mnId = 0;

  // plain state
  Eigen::Vector3f mWorldPos, mNormalVector;
  cv::Mat         mDescriptor;
  int             nObs = 0, mnVisible = 1;
  bool            mbBad = false;
  float           mfMinDistance = 0, mfMaxDistance = 0;
  MapPoint*       mpReplaced = nullptr;
  std::map<KeyFrame*, std::tuple<int, int>> mObservations;
  std::mutex      mMutexPos;
};

}  // namespace ORB_SLAM3
