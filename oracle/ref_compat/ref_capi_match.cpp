// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
// Flat C entry points (ctypes) that drive the reference's OWN matching code:
//   Frame::ComputeStereoMatches          lifted verbatim from src/Frame.cc:944-1113
//   Frame::AssignFeaturesToGrid/PosInGrid/GetFeaturesInArea   src/Frame.cc:524-553, 788-877
//   ORBmatcher::*                        the whole unmodified src/ORBmatcher.cc
//   Frame::isInFrustum, MapPoint::PredictScale, Pinhole::project   lifted verbatim
// Argument shapes mirror the oo_* calls of oracle/orb_oracle_capi.cpp (and therefore the C ABI of
// include/cuorb.h) so that tests run oracle, reference and libcuorb on identical inputs.
//
// How a flat SoA becomes the reference's object graph: every query (MapPoint / last-frame feature)
// becomes a MapPoint object whose world position IS the projected pixel, (u, v, z); the current
// frame gets the identity pose and a pass-through camera (project(x, y, z) = (x, y)), so the
// reference's own prologue (`Tcw * x3Dw`, `1.0 / z`, `project`) returns exactly the floats the
// oracle is handed.  Pose algebra parity is not claimed (sophus/se3.hpp stand-in); everything the
// reference does after the projection is its own code.
#include <atomic>
#include <chrono>
#include <cstring>
#include <memory>
#include <thread>
#include <vector>

#include "../../include/cuorb.h"  // POD layouts only (cuorb_keypoint, cuorb_frame_view, CUORB_MP_*)
#include "Frame.h"
#include "GeometricCamera.h"
#include "KeyFrame.h"
#include "MapPoint.h"
#include "ORBextractor.h"
#include "ORBmatcher.h"
#include "ref_harness.h"

using namespace ORB_SLAM3;
using namespace ref_harness;
static_assert(sizeof(cv::KeyPoint) == sizeof(cuorb_keypoint), "KeyPoint ABI");


extern "C" {

int ref_hamming(const uint8_t* a, const uint8_t* b) {
  return ORBmatcher::DescriptorDistance(desc_mat(a, 1), desc_mat(b, 1));
}

// ---- Frame::ComputeStereoMatches on the pyramids the two reference extractors hold ----
int ref_stereo(void* hl, void* hr, int nL, const cv::KeyPoint* kl, const uint8_t* dl, int nR,
               const cv::KeyPoint* kr, const uint8_t* dr, float mb, float mbf, float* uRight,
               float* depth) {
  ORBextractor *L = (ORBextractor*)hl, *R = (ORBextractor*)hr;
  Frame F;
  F.N = nL;
  F.mvKeys.assign(kl, kl + nL);
  F.mvKeysRight.assign(kr, kr + nR);
  F.mDescriptors = desc_mat(dl, nL);
  F.mDescriptorsRight = desc_mat(dr, nR);
  F.mvScaleFactors = L->GetScaleFactors();
  F.mvInvScaleFactors = L->GetInverseScaleFactors();
  F.mb = mb; F.mbf = mbf;
  F.mpORBextractorLeft = L; F.mpORBextractorRight = R;
  if (nL == 0 || nR == 0) return 0;  // the reference indexes an empty vector here (UB, :1102)
  F.ComputeStereoMatches();
  for (int i = 0; i < nL; i++) { uRight[i] = F.mvuRight[i]; depth[i] = F.mvDepth[i]; }
  return 0;
}

// ---- Frame grid on its own: GetFeaturesInArea(x, y, r, minLevel, maxLevel) ----
int ref_features_in_area(const cuorb_frame_view* frame, float x, float y, float r, int min_level,
                         int max_level, int32_t* out, int cap) {
  FrameFromView fv(frame);
  std::vector<std::size_t> v = fv.F.GetFeaturesInArea(x, y, r, min_level, max_level);
  for (size_t i = 0; i < v.size() && (int)i < cap; i++) out[i] = (int32_t)v[i];
  return (int)v.size();
}
int ref_grid_cell(const cuorb_frame_view* frame, int ix, int iy, int32_t* out, int cap) {
  FrameFromView fv(frame);
  const std::vector<std::size_t>& v = fv.F.mGrid[ix][iy];
  for (size_t i = 0; i < v.size() && (int)i < cap; i++) out[i] = (int32_t)v[i];
  return (int)v.size();
}

// ---- int ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th, bFarPoints,
// thFarPoints)  (src/ORBmatcher.cc:38-227).  final_binding_out[idx] = MapPoint index bound to
// frame feature idx when the call returns (-1: none, -2: the pre-existing binding); best_idx_out[i]
// = feature MapPoint i was bound to at the moment it was processed, obtained by feeding the
// MapPoints one call at a time to the same Frame (the loop body only communicates through
// F.mvpMapPoints, so this is the same computation). ----
int ref_search_by_projection(const cuorb_frame_view* frame, int M, const uint8_t* mp_desc,
                             const float* px, const float* py, const float* pxr,
                             const int32_t* level, const float* view_cos, const float* depth,
                             const uint8_t* flags, float th, int far_points, float th_far,
                             float nn_ratio, int32_t* best_idx_out, int32_t* final_binding_out) {
  std::vector<MapPoint> mps(M);
  std::vector<MapPoint*> ptrs(M);
  for (int i = 0; i < M; i++) {
    MapPoint& p = mps[i];
    p.mbTrackInView = (flags[i] & CUORB_MP_IN_VIEW) != 0;
    p.mbBad = (flags[i] & CUORB_MP_BAD) != 0;
    p.nObs = (flags[i] & CUORB_MP_HAS_OBS) ? 1 : 0;
    p.mTrackProjX = px[i]; p.mTrackProjY = py[i]; p.mTrackProjXR = pxr ? pxr[i] : -1.f;
    p.mnTrackScaleLevel = level[i]; p.mTrackViewCos = view_cos[i];
    p.mTrackDepth = depth ? depth[i] : 0.f;
    p.mDescriptor = desc_mat(mp_desc + (size_t)i * 32, 1);
    ptrs[i] = &p;
  }
  ORBmatcher matcher(nn_ratio, true);
  int nmatches = 0;
  {
    FrameFromView fv(frame);
    nmatches = matcher.SearchByProjection(fv.F, ptrs, th, far_points != 0, th_far);
    if (final_binding_out)
      for (int k = 0; k < fv.F.N; k++) {
        MapPoint* b = fv.F.mvpMapPoints[k];
        final_binding_out[k] = !b ? -1 : (b >= &mps[0] && b < &mps[0] + M) ? (int32_t)(b - &mps[0]) : -2;
      }
  }
  if (best_idx_out) {
    FrameFromView fv(frame);
    int total = 0;
    for (int i = 0; i < M; i++) {
      std::vector<MapPoint*> one(1, ptrs[i]);
      const int n = matcher.SearchByProjection(fv.F, one, th, far_points != 0, th_far);
      total += n;
      best_idx_out[i] = -1;
      if (n)
        for (int k = 0; k < fv.F.N; k++)
          if (fv.F.mvpMapPoints[k] == ptrs[i]) { best_idx_out[i] = k; break; }
    }
    if (total != nmatches) return -1000 - total;  // the equivalence argued above does not hold
  }
  return nmatches;
}

// ---- int ORBmatcher::SearchByProjection(Frame& Cur, const Frame& Last, th, bMono)
// (src/ORBmatcher.cc:1684-1908).  z[i] is the camera-frame depth of last-frame feature i (the
// reference computes invzc = 1.0 / z itself).  final_binding_out[i2] = last-frame feature whose
// MapPoint is bound to current feature i2 on return (-1 none, -2 pre-existing). ----
int ref_search_by_projection_frame(const cuorb_frame_view* cur, int n_last, const uint8_t* last_desc,
                                   const float* u, const float* v, const float* z,
                                   const int32_t* last_octave, const float* last_angle,
                                   const uint8_t* flags, float th, int forward, int backward,
                                   int check_ori, int32_t* final_binding_out) {
  FrameFromView fv(cur);
  Frame last;
  std::vector<MapPoint> mps(n_last);
  last.N = n_last; last.Nleft = -1;
  last.mvKeys.resize(n_last); last.mvKeysUn.resize(n_last);
  last.mvpMapPoints.assign(n_last, nullptr);
  last.mvbOutlier.assign(n_last, false);
  for (int i = 0; i < n_last; i++) {
    last.mvKeys[i].octave = last_octave[i]; last.mvKeys[i].angle = last_angle[i];
    last.mvKeysUn[i] = last.mvKeys[i];
    if (!(flags[i] & CUORB_MP_IN_VIEW)) continue;  // bit0: has a MapPoint and is not an outlier
    mps[i].mWorldPos = Eigen::Vector3f(u[i], v[i], z[i]);
    mps[i].nObs = (flags[i] & CUORB_MP_HAS_OBS) ? 1 : 0;
    mps[i].mDescriptor = desc_mat(last_desc + (size_t)i * 32, 1);
    last.mvpMapPoints[i] = &mps[i];
  }
  // bForward / bBackward (:1702-1703) come from tlc = Tlw * twc; Cur has the identity pose, so
  // tlc is the last frame's translation
  const float tz = forward ? fv.F.mb + 1.f : backward ? -(fv.F.mb + 1.f) : 0.f;
  last.mTcw = Sophus::SE3f(Eigen::Matrix3f::Identity(), Eigen::Vector3f(0.f, 0.f, tz));
  ORBmatcher matcher(0.9f, check_ori != 0);
  const int n = matcher.SearchByProjection(fv.F, last, th, false);
  if (final_binding_out)
    for (int k = 0; k < fv.F.N; k++) {
      MapPoint* b = fv.F.mvpMapPoints[k];
      final_binding_out[k] =
          !b ? -1 : (n_last && b >= &mps[0] && b < &mps[0] + n_last) ? (int32_t)(b - &mps[0]) : -2;
    }
  return n;
}

// ---- the same function with its geometric prologue live: real poses (Sophus stand-in), world points, the
// reference's own Pinhole::project.  pose = 9 floats R (row-major) + 3 floats t, world -> camera.  What
// shim/ORBmatcher.h's SearchByProjection(Cur, Last) overload is checked against (tests/test_gpu_shim.py). ----
int ref_search_by_projection_frame_pose(const cuorb_frame_view* cur, const float* pose_cur, const float* pose_last,
                                        const float* K4, int n_last, const uint8_t* last_desc, const float* world_pos,
                                        const int32_t* last_octave, const float* last_angle, const uint8_t* has_mp,
                                        const uint8_t* outlier, const int32_t* n_obs, float th, int mono, int check_ori,
                                        int32_t* final_binding_out) {
  FrameFromView fv(cur);
  Pinhole cam(std::vector<float>{K4[0], K4[1], K4[2], K4[3]});
  fv.F.mpCamera = &cam;
  auto pose = [](const float* p) {
    Eigen::Matrix3f R;
    for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) R(r, c) = p[3 * r + c];
    return Sophus::SE3f(R, Eigen::Vector3f(p[9], p[10], p[11]));
  };
  fv.F.mTcw = pose(pose_cur);
  Frame last;
  std::unique_ptr<MapPoint[]> mps(new MapPoint[n_last > 0 ? n_last : 1]);
  last.N = n_last; last.Nleft = -1;
  last.mTcw = pose(pose_last);
  last.mvKeys.resize(n_last); last.mvKeysUn.resize(n_last);
  last.mvpMapPoints.assign(n_last, nullptr);
  last.mvbOutlier.assign(n_last, false);
  for (int i = 0; i < n_last; i++) {
    last.mvKeys[i].octave = last_octave[i]; last.mvKeys[i].angle = last_angle[i];
    last.mvKeysUn[i] = last.mvKeys[i];
    last.mvbOutlier[i] = outlier[i] != 0;
    if (!has_mp[i]) continue;
    mps[i].mWorldPos = Eigen::Vector3f(world_pos[3 * i], world_pos[3 * i + 1], world_pos[3 * i + 2]);
    mps[i].nObs = n_obs[i];
    mps[i].mDescriptor = desc_mat(last_desc + (size_t)i * 32, 1);
    last.mvpMapPoints[i] = &mps[i];
  }
  ORBmatcher matcher(0.9f, check_ori != 0);
  const int n = matcher.SearchByProjection(fv.F, last, th, mono != 0);
  for (int k = 0; k < fv.F.N; k++) {
    MapPoint* b = fv.F.mvpMapPoints[k];
    final_binding_out[k] = !b ? -1 : (b >= &mps[0] && b < &mps[0] + n_last) ? (int32_t)(b - &mps[0]) : -2;
  }
  return n;
}

// ---- bool Frame::isInFrustum(MapPoint*, float) — the reference's own body (src/Frame.cc:644-721), with
// MapPoint::PredictScale (src/MapPoint.cc:574-589) and Pinhole::project (Pinhole.cpp:70-76) lifted verbatim;
// the Eigen expressions inside run on the oracle/ref_compat/Eigen stand-in. ----
int ref_is_in_frustum(const cuorb_frustum_camera* c, int n, const float* P, const float* N, const float* minD,
                      const float* maxD, float limit, uint8_t* in_view, float* px, float* py, float* pxr,
                      int32_t* level, float* view_cos, float* depth) {
  Frame F;
  Pinhole cam(std::vector<float>{c->fx, c->fy, c->cx, c->cy});
  F.mpCamera = &cam;
  F.Nleft = -1;
  for (int r = 0; r < 3; r++) {
    for (int k = 0; k < 3; k++) F.mRcw(r, k) = c->rcw[3 * r + k];
    F.mtcw(r) = c->tcw[r];
    F.mOw(r) = c->ow[r];
  }
  F.mbf = c->mbf;
  Frame::mnMinX = c->min_x; Frame::mnMaxX = c->max_x; Frame::mnMinY = c->min_y; Frame::mnMaxY = c->max_y;
  F.mfLogScaleFactor = c->log_scale_factor;
  F.mnScaleLevels = c->n_levels;
  int count = 0;
  for (int i = 0; i < n; i++) {
    MapPoint mp;
    mp.mWorldPos = Eigen::Vector3f(P[3 * i], P[3 * i + 1], P[3 * i + 2]);
    mp.mNormalVector = Eigen::Vector3f(N[3 * i], N[3 * i + 1], N[3 * i + 2]);
    mp.mfMinDistance = minD[i]; mp.mfMaxDistance = maxD[i];
    mp.mTrackProjXR = -1; mp.mnTrackScaleLevel = -1; mp.mTrackViewCos = 0; mp.mTrackDepth = 0;
    const bool ok = F.isInFrustum(&mp, limit);
    in_view[i] = ok ? 1 : 0; count += ok;
    px[i] = mp.mTrackProjX; py[i] = mp.mTrackProjY; pxr[i] = mp.mTrackProjXR;
    level[i] = mp.mnTrackScaleLevel; view_cos[i] = mp.mTrackViewCos; depth[i] = mp.mTrackDepth;
  }
  return count;
}

// ---- CPU timing loop over the reference's own per-frame stereo path (two operator() calls and
// ComputeStereoMatches, src/Frame.cc:202-205, 231), one std::thread per in-flight frame. ----
double ref_bench_stereo(int nfeatures, float scale, int nlevels, int ini, int min,
                        const uint8_t* left, const uint8_t* right, int n_frames, int w, int h,
                        float mb, float mbf, int n_threads, int64_t* total_kps,
                        int64_t* total_matches) {
  std::atomic<int>     next(0);
  std::atomic<int64_t> kps(0), matches(0);
  auto t0 = std::chrono::steady_clock::now();
  std::vector<std::thread> pool;
  for (int t = 0; t < n_threads; t++)
    pool.emplace_back([&]() {
      ORBextractor EL(nfeatures, scale, nlevels, ini, min), ER(nfeatures, scale, nlevels, ini, min);
      std::vector<int> lap = {0, 0};
      for (;;) {
        int i = next.fetch_add(1);
        if (i >= n_frames) break;
        Frame F;
        cv::Mat imL(h, w, CV_8UC1, (void*)(left + (size_t)i * w * h), w);
        cv::Mat imR(h, w, CV_8UC1, (void*)(right + (size_t)i * w * h), w);
        EL(imL, cv::Mat(), F.mvKeys, F.mDescriptors, lap);
        ER(imR, cv::Mat(), F.mvKeysRight, F.mDescriptorsRight, lap);
        F.N = (int)F.mvKeys.size();
        F.mvScaleFactors = EL.GetScaleFactors(); F.mvInvScaleFactors = EL.GetInverseScaleFactors();
        F.mb = mb; F.mbf = mbf; F.mpORBextractorLeft = &EL; F.mpORBextractorRight = &ER;
        int64_t m = 0;
        if (F.N > 0 && !F.mvKeysRight.empty()) {
          F.ComputeStereoMatches();
          for (float x : F.mvuRight) m += x >= 0;
        }
        kps += (int64_t)(F.mvKeys.size() + F.mvKeysRight.size());
        matches += m;
      }
    });
  for (auto& th : pool) th.join();
  double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  if (total_kps) *total_kps = kps.load();
  if (total_matches) *total_matches = matches.load();
  return sec;
}

}  // extern "C"
