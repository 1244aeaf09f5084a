// Device-resident mirror of Tracking's local map (mvpLocalMapPoints) and a drop-in for the two hot loops of
// void Tracking::SearchLocalPoints() (src/Tracking.cc:3186-3260):
//   for (pMP : mvpLocalMapPoints) { skip if mnLastFrameSeen == frame id or bad; if (mCurrentFrame.isInFrustum(pMP, 0.5))
//     { pMP->IncreaseVisible(); nToMatch++; } }                                                  (:3207-3232)
//   matcher.SearchByProjection(mCurrentFrame, mvpLocalMapPoints, th, bFarPoints, thFarPoints);    (:3255)
// over libcuorb's cuorb_localmap_* (include/cuorb.h).  Templated over the MapPoint / Frame types so the same header
// compiles against the reference's classes (member names below are the reference's).
//
// Storage and order are separate on purpose.  Tracking::UpdateLocalPoints rebuilds mvpLocalMapPoints every frame, in
// an order that decides who gets a contested feature; the points themselves change by a few per cent per frame.  Each
// MapPoint therefore keeps a storage slot for as long as it stays in the mirror, and every frame only sends (a) the
// slots in this frame's vector order (4 B per point), (b) the points that are new or that the caller reports as changed.
// What "changed" means is the caller's to say (the reference has no version counter): a point whose SetWorldPos,
// UpdateNormalAndDepth or ComputeDistinctiveDescriptors ran since the last Sync - e.g. every point touched by local
// bundle adjustment or by LocalMapping's fusion.  `changed(pMP)` returning true for every point is always correct
// (3.4 MB per 50k points per frame); INTEGRATION.md shows the three one-line hooks that make it precise.
#pragma once
#include <algorithm>
#include <cstring>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

#include "../include/cuorb.h"
#include "cv_compat.h"

namespace ORB_SLAM3 {

template <class MapPointT>
class LocalMapMirror {
public:
  explicit LocalMapMirror(int capacity = 1 << 17, int device = 0) : capacity_(capacity) {
    h_ = cuorb_localmap_create(device, capacity);
    if (!h_) throw std::runtime_error(std::string("LocalMapMirror (libcuorb): ") + cuorb_last_error());
  }
  ~LocalMapMirror() { cuorb_localmap_destroy(h_); }
  LocalMapMirror(const LocalMapMirror&) = delete;
  LocalMapMirror& operator=(const LocalMapMirror&) = delete;

  // After Tracking::UpdateLocalPoints: bring the mirror in line with vpLocalMapPoints.  Returns the number of points
  // sent, or -1 (error text on stderr / cuorb_last_error()).
  template <class ChangedFn>
  int Sync(const std::vector<MapPointT*>& vpLocalMapPoints, ChangedFn changed) {
    const int M = (int)vpLocalMapPoints.size();
    order_.resize(M);
    points_ = vpLocalMapPoints;
    // points that left the local map give their slots back (their storage simply goes stale)
    std::unordered_map<MapPointT*, int> next;
    next.reserve((size_t)M * 2);
    std::vector<int> send;
    for (int i = 0; i < M; i++) {
      MapPointT* p = vpLocalMapPoints[i];
      auto it = slot_.find(p);
      int s;
      if (it != slot_.end()) { s = it->second; if (changed(p)) send.push_back(i); }
      else { s = -1; send.push_back(i); }
      next.emplace(p, s);
      order_[i] = s;
    }
    for (auto& kv : slot_) if (!next.count(kv.first)) free_.push_back(kv.second);
    for (int i : send) {
      if (order_[i] >= 0) continue;
      int s;
      if (!free_.empty()) { s = free_.back(); free_.pop_back(); }
      else if (used_ < capacity_) s = used_++;
      else CUORB_SHIM_FAIL(std::runtime_error, "LocalMapMirror: local map larger than the mirror's capacity", -1);
      order_[i] = s; next[vpLocalMapPoints[i]] = s;
    }
    slot_.swap(next);
    // send the new / changed points, contiguous slot runs in one call each
    std::sort(send.begin(), send.end(), [&](int a, int b) { return order_[a] < order_[b]; });
    std::vector<float> P, N, mn, mx;
    std::vector<uint8_t> desc, fl;
    size_t k = 0;
    while (k < send.size()) {
      size_t e = k + 1;
      while (e < send.size() && order_[send[e]] == order_[send[e - 1]] + 1) e++;
      const int n = (int)(e - k);
      P.resize(3 * (size_t)n); N.resize(3 * (size_t)n); mn.resize(n); mx.resize(n); desc.resize(32 * (size_t)n); fl.resize(n);
      for (int j = 0; j < n; j++) {
        MapPointT* p = vpLocalMapPoints[send[k + j]];
        const auto w = p->GetWorldPos();
        const auto nr = p->GetNormal();
        for (int c = 0; c < 3; c++) { P[3 * j + c] = w(c); N[3 * j + c] = nr(c); }
        // the raw mfMinDistance / mfMaxDistance (two one-line getters to add to MapPoint): isInFrustum's
        // range test uses 0.8f * min and 1.2f * max, PredictScale the raw max, and neither product can be undone exactly
        mn[j] = p->GetMinDistance(); mx[j] = p->GetMaxDistance();
        const cv::Mat d = p->GetDescriptor();
        std::memcpy(&desc[32 * (size_t)j], d.template ptr<unsigned char>(0), 32);
        fl[j] = (uint8_t)((p->isBad() ? CUORB_MP_BAD : 0) | (p->Observations() > 0 ? CUORB_MP_HAS_OBS : 0));
      }
      const int rc = cuorb_localmap_update(h_, order_[send[k]], n, used_, P.data(), N.data(), mn.data(), mx.data(), desc.data(), fl.data());
      if (rc < 0) CUORB_SHIM_FAIL(std::runtime_error, std::string("LocalMapMirror::Sync (libcuorb): ") + cuorb_last_error(), -1);
      k = e;
    }
    return (int)send.size();
  }

  // The two loops of Tracking::SearchLocalPoints on the mirror.  cam: the current Frame's pose and intrinsics
  // (cuorb_frustum_camera; FillCamera below builds it from the reference's Frame members); fv: the Frame as
  // ORBmatcher::view(F, bound) of shim/ORBmatcher.h flattens it.  Leaves in every MapPoint
  // what the reference leaves (mbTrackInView and, for points in view, the mTrack* fields are not needed any more by
  // the search and are NOT written back), calls IncreaseVisible() for the points in view, replays
  // F.mvpMapPoints[feature] = pMP in vector order.  Returns nmatches; *nToMatch = points in view.
  template <class FrameT>
  int SearchLocalPoints(FrameT& F, const cuorb_frustum_camera& cam, const cuorb_frame_view& fv, unsigned long frame_id, float th,
                        bool bFarPoints, float thFarPoints, float nnratio, int* nToMatch) {
    return Run(F, frame_id, nToMatch, "SearchLocalPoints", [&](int M, const std::vector<int32_t>& skip, uint8_t* in_view, int cap,
                                                                 int32_t* mp, int32_t* ft, int32_t* n_in) {
      return cuorb_localmap_search(h_, &cam, &fv, M, order_.data(), (int)skip.size(), skip.data(), 0.5f, th, bFarPoints ? 1 : 0,
                                   thFarPoints, nnratio, nullptr, in_view, cap, mp, ft, n_in);
    });
  }

  // The same, on the Frame's arrays where the extractor left them on the device (cuorb_localmap_search_resident):
  // ext = the handle behind the Frame's left ORBextractor (shim/ORBextractor.h handle()), image_index = the Frame's
  // position in that handle's last batch.  fv as above - its keys_un / descriptors / u_right host pointers are replaced
  // by the device arrays (u_right only if fv.u_right is set, i.e. a stereo Frame whose ComputeStereoMatches ran on the
  // handle).  K4 / dist5 (fx fy cx cy / k1 k2 p1 p2 k3, or both null for a rectified camera): the device keypoints are
  // mvKeys and Frame::UndistortKeyPoints is applied to them on the device.  Nothing of the Frame's feature arrays is
  // uploaded.
  template <class FrameT>
  int SearchLocalPointsResident(FrameT& F, const cuorb_frustum_camera& cam, cuorb_frame_view fv, cuorb_extractor* ext,
                                int image_index, const float* K4, const float* dist5, unsigned long frame_id, float th,
                                bool bFarPoints, float thFarPoints, float nnratio, int* nToMatch) {
    const cuorb_keypoint* dk = nullptr; const uint8_t* dd = nullptr; const float* du = nullptr;
    if (cuorb_extractor_device_features(ext, image_index, &dk, &dd, &du, nullptr) < 0 || (fv.u_right && !du))
      CUORB_SHIM_FAIL(std::runtime_error, std::string("SearchLocalPointsResident (libcuorb): ") +
                      (fv.u_right && dk ? "no stereo match has run on the handle" : cuorb_last_error()), 0);
    fv.keys_un = dk; fv.descriptors = dd; fv.u_right = fv.u_right ? du : nullptr;
    return Run(F, frame_id, nToMatch, "SearchLocalPointsResident", [&](int M, const std::vector<int32_t>& skip, uint8_t* in_view,
                                                                         int cap, int32_t* mp, int32_t* ft, int32_t* n_in) {
      return cuorb_localmap_search_resident(h_, &cam, &fv, cuorb_extractor_stream(ext), K4, dist5, nullptr, M, order_.data(),
                                            (int)skip.size(), skip.data(), 0.5f, th, bFarPoints ? 1 : 0, thFarPoints, nnratio,
                                            nullptr, in_view, cap, mp, ft, n_in);
    });
  }

  // cuorb_frustum_camera from the reference's Frame members (mRcw, mtcw, mOw, fx, fy, cx, cy, mbf, image bounds,
  // mfLogScaleFactor, mnScaleLevels) - what Frame::isInFrustum reads (src/Frame.cc:644-707)
  template <class FrameT>
  static cuorb_frustum_camera FillCamera(const FrameT& F) {
    cuorb_frustum_camera c;
    for (int r = 0; r < 3; r++) {
      for (int k = 0; k < 3; k++) c.rcw[3 * r + k] = F.mRcw(r, k);
      c.tcw[r] = F.mtcw(r); c.ow[r] = F.mOw(r);
    }
    c.fx = F.fx; c.fy = F.fy; c.cx = F.cx; c.cy = F.cy; c.mbf = F.mbf;
    c.min_x = F.mnMinX; c.max_x = F.mnMaxX; c.min_y = F.mnMinY; c.max_y = F.mnMaxY;
    c.log_scale_factor = F.mfLogScaleFactor; c.n_levels = F.mnScaleLevels;
    return c;
  }

  int size() const { return (int)points_.size(); }
  cuorb_localmap* handle() const { return h_; }

private:
  // skip list (:3214-3216) -> one library call -> what the reference's loops leave in the MapPoints and the Frame
  template <class FrameT, class CallFn>
  int Run(FrameT& F, unsigned long frame_id, int* nToMatch, const char* who, CallFn call) {
    const int M = (int)points_.size();
    if (nToMatch) *nToMatch = 0;
    if (M == 0) return 0;
    std::vector<int32_t> skip;
    for (int i = 0; i < M; i++) if (points_[i]->mnLastFrameSeen == frame_id) skip.push_back(i);
    std::vector<uint8_t> in_view(M);
    const int cap = F.N > 0 ? 4 * F.N + 64 : 64;
    std::vector<int32_t> mp(cap), ft(cap);
    int32_t n_in = 0;
    const int n = call(M, skip, in_view.data(), cap, mp.data(), ft.data(), &n_in);
    if (n < 0) CUORB_SHIM_FAIL(std::runtime_error, std::string(who) + " (libcuorb): " + cuorb_last_error(), 0);
    for (int i = 0; i < M; i++) {
      points_[i]->mbTrackInView = in_view[i] != 0;
      if (in_view[i]) points_[i]->IncreaseVisible();                                              // :3222-3225
    }
    for (int k = 0; k < n; k++) F.mvpMapPoints[ft[k]] = points_[mp[k]];                           // src/ORBmatcher.cc:136
    if (nToMatch) *nToMatch = n_in;
    return n;
  }

  cuorb_localmap* h_ = nullptr;
  int capacity_ = 0, used_ = 0;
  std::unordered_map<MapPointT*, int> slot_;
  std::vector<int> free_;
  std::vector<int32_t> order_;
  std::vector<MapPointT*> points_;
};

}  // namespace ORB_SLAM3
