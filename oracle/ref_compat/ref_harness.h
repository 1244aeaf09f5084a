// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Helpers shared by the ref_capi_*.cpp harnesses: how a flat
// cuorb_frame_view / SoA of queries becomes the reference's object graph (stub Frame / KeyFrame / MapPoint
// of oracle/ref_compat, filled the way the reference's constructors fill the real ones).
#pragma once
#include <cmath>
#include <cstring>
#include <memory>
#include <vector>

#include "../../include/cuorb.h"  // POD layouts only
#include "Frame.h"
#include "GeometricCamera.h"
#include "KeyFrame.h"
#include "MapPoint.h"

namespace ref_harness {
using namespace ORB_SLAM3;

// project(x, y, z) = (x, y), K = I: a MapPoint whose world position is (u, v, z) seen from the identity pose
// projects to exactly (u, v), so the reference's own prologue returns the floats the oracle is handed.
struct PassThroughCamera : GeometricCamera {
  cv::Point2f     project(const cv::Point3f& p) override { return cv::Point2f(p.x, p.y); }
  Eigen::Vector2d project(const Eigen::Vector3d& v) override { return Eigen::Vector2d(v[0], v[1]); }
  Eigen::Vector2f project(const Eigen::Vector3f& v) override { return Eigen::Vector2f(v[0], v[1]); }
  Eigen::Matrix3f toK_() override { return Eigen::Matrix3f::Identity(); }
  bool epipolarConstrain(GeometricCamera*, const cv::KeyPoint&, const cv::KeyPoint&,
                         const Eigen::Matrix3f&, const Eigen::Vector3f&, const float,
                         const float) override {
    REF_COMPAT_UNSUPPORTED("epipolarConstrain on the pass-through camera");
    return false;
  }
};

// Pinhole with K = I: project divides by z (the reference's Pinhole::project), epipolarConstrain is the
// reference's own body (lifted), toK_ = I
struct UnitPinhole : Pinhole {
  UnitPinhole() : Pinhole(std::vector<float>{1.f, 1.f, 0.f, 0.f}) {}
};

inline cv::Mat desc_mat(const uint8_t* d, int n) {
  return n > 0 ? cv::Mat(n, 32, CV_8UC1, (void*)d, 32) : cv::Mat();
}

// Fill a stub Frame from a cuorb_frame_view the way the Frame constructors do (src/Frame.cc:
// 240-257 sets the statics; AssignFeaturesToGrid is the reference's own, called as at :262).
struct FrameFromView {
  Frame                 F;
  PassThroughCamera     cam;
  std::unique_ptr<MapPoint[]> bound;  // MapPoints standing for "already bound, Observations() > 0"
  FrameFromView(const cuorb_frame_view* f) {
    F.N = f->n;
    const cv::KeyPoint* k = (const cv::KeyPoint*)f->keys_un;
    F.mvKeysUn.assign(k, k + f->n);
    F.mvKeys = F.mvKeysUn;
    F.mDescriptors = desc_mat(f->descriptors, f->n);
    F.mvuRight.assign(f->n, -1.f);
    if (f->u_right) F.mvuRight.assign(f->u_right, f->u_right + f->n);
    F.mvpMapPoints.assign(f->n, nullptr);
    F.mvbOutlier.assign(f->n, false);
    bound.reset(new MapPoint[f->n > 0 ? f->n : 1]);
    for (int i = 0; i < f->n; i++)
      if (f->bound_with_obs && f->bound_with_obs[i]) { bound[i].nObs = 1; F.mvpMapPoints[i] = &bound[i]; }
    Frame::mnMinX = f->min_x; Frame::mnMinY = f->min_y; Frame::mnMaxX = f->max_x; Frame::mnMaxY = f->max_y;
    Frame::mfGridElementWidthInv = f->grid_w_inv; Frame::mfGridElementHeightInv = f->grid_h_inv;
    F.mb = f->mb; F.mbf = f->mbf;
    F.mvScaleFactors.assign(f->scale_factors, f->scale_factors + f->nlevels);
    F.mnScaleLevels = f->nlevels;
    F.mfScaleFactor = f->nlevels > 1 ? f->scale_factors[1] : 1.2f;
    F.mfLogScaleFactor = std::log(F.mfScaleFactor);
    F.Nleft = -1; F.Nright = -1;
    F.mpCamera = &cam;
    F.AssignFeaturesToGrid();
  }
};

// A stub KeyFrame made from a Frame as KeyFrame::KeyFrame(Frame&, Map*, KeyFrameDatabase*) does
// (src/KeyFrame.cc:70-150): the Frame's grid is copied cell by cell, the image bounds become ints.
// mvpMapPoints[i] = &bound[i] where the view says "bound" (GetMapPoint(i) != NULL); they are good points.
struct KeyFrameFromView {
  FrameFromView fv;
  KeyFrame      K;
  KeyFrameFromView(const cuorb_frame_view* f) : fv(f) {
    const Frame& F = fv.F;
    K.N = F.N; K.NLeft = -1; K.NRight = -1;
    K.mnGridCols = FRAME_GRID_COLS; K.mnGridRows = FRAME_GRID_ROWS;
    K.mfGridElementWidthInv = Frame::mfGridElementWidthInv; K.mfGridElementHeightInv = Frame::mfGridElementHeightInv;
    K.fx = 1.f; K.fy = 1.f; K.cx = 0.f; K.cy = 0.f; K.invfx = 1.f; K.invfy = 1.f;   // the pass-through camera's K
    K.mbf = F.mbf; K.mb = F.mb;
    K.mvKeys = F.mvKeys; K.mvKeysUn = F.mvKeysUn; K.mvuRight = F.mvuRight;
    K.mDescriptors = F.mDescriptors;
    K.mnScaleLevels = F.mnScaleLevels; K.mfScaleFactor = F.mfScaleFactor; K.mfLogScaleFactor = F.mfLogScaleFactor;
    K.mvScaleFactors = F.mvScaleFactors;
    K.mvLevelSigma2.resize(F.mnScaleLevels); K.mvInvLevelSigma2.resize(F.mnScaleLevels);
    for (int l = 0; l < F.mnScaleLevels; l++) {       // ORBextractor.cc:428-441: sigma2 = sf * sf, inv = 1 / sigma2
      K.mvLevelSigma2[l] = F.mvScaleFactors[l] * F.mvScaleFactors[l];
      K.mvInvLevelSigma2[l] = 1.0f / K.mvLevelSigma2[l];
    }
    K.mnMinX = Frame::mnMinX; K.mnMinY = Frame::mnMinY; K.mnMaxX = Frame::mnMaxX; K.mnMaxY = Frame::mnMaxY;
    K.mvpMapPoints = F.mvpMapPoints;
    K.mGrid.resize(K.mnGridCols);
    for (int i = 0; i < K.mnGridCols; i++) {
      K.mGrid[i].resize(K.mnGridRows);
      for (int j = 0; j < K.mnGridRows; j++) K.mGrid[i][j] = F.mGrid[i][j];
    }
    K.mpCamera = &fv.cam; K.mpCamera2 = nullptr;
  }
};

// n query MapPoints placed so that the reference's own geometric prologue, run from the identity pose through
// the pass-through camera, reproduces (u, v, level) and lets exactly the `active` ones through:
//   world position (u, v, z)                      -> Tcw * p = p, project = (u, v), depth z > 0
//   mfMaxDistance = dist * sf^(level - 0.5)       -> MapPoint::PredictScale's ceil(log(max / dist) / log sf) sits
//                                                    mid-bin and returns `level`; dist <= 1.2 * max holds
//   mfMinDistance = 0, normal = PO / dist         -> distance range and the 60 degree test pass
//   inactive                                      -> a bad point (isBad(): the first test of every prologue)
struct QueryPoints {
  std::unique_ptr<MapPoint[]> mp;
  std::vector<MapPoint*>      ptr;
  int                         n;
  QueryPoints(int n_, const uint8_t* desc, const float* u, const float* v, const float* z, const int32_t* level,
              const uint8_t* active, float scale_factor)
      : mp(new MapPoint[n_ > 0 ? n_ : 1]), ptr(n_), n(n_) {
    for (int i = 0; i < n; i++) {
      MapPoint& p = mp[i];
      const float zi = z ? z[i] : 1.f;
      p.mWorldPos = Eigen::Vector3f(u[i], v[i], zi);
      const float dist = p.mWorldPos.norm();
      p.mfMinDistance = 0.f;
      p.mfMaxDistance = (float)((double)dist * std::pow((double)scale_factor, (double)level[i] - 0.5));
      p.mNormalVector = p.mWorldPos / dist;
      p.mbBad = !active[i];
      p.mDescriptor = desc_mat(desc + (size_t)i * 32, 1);
      p.mnId = (unsigned long)i;
      ptr[i] = &p;
    }
  }
  int index_of(const MapPoint* q) const { return (q && n && q >= &mp[0] && q < &mp[0] + n) ? (int)(q - &mp[0]) : -1; }
};

}  // namespace ref_harness
