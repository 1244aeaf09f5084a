// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
// Flat C entry points (ctypes) that drive the reference's OWN code for the ORBmatcher searches outside the
// per-frame tracking path — the whole unmodified src/ORBmatcher.cc:
//   SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches)                       :237-420
//   SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12)                          :784-914
//   SearchForInitialization(Frame&, Frame&, vbPrevMatched, vnMatches12, w)  :666-782
//   SearchByProjection(Frame&, KeyFrame*, sAlreadyFound, th, ORBdist)       :1910-2036
//   SearchByProjection(KeyFrame*, Sim3f&, vpPoints, vpMatched, th, ratio)   :423-538
//   SearchByProjection(KeyFrame*, Sim3&, vpPoints, vpPointsKFs, ...)        :540-664
//   Fuse(KeyFrame*, vpMapPoints, th, bRight)                                :1149-1328
//   Fuse(KeyFrame*, Sim3f&, vpPoints, th, vpReplacePoint)                   :1330-1452
//   SearchBySim3(KeyFrame*, KeyFrame*, vpMatches12, S12, th)                :1454-1682
//   SearchForTriangulation(KeyFrame*, KeyFrame*, vMatchedPairs, ...)        :916-1147
// over KeyFrame::GetFeaturesInArea / IsInImage, MapPoint::PredictScale, Pinhole::epipolarConstrain lifted
// verbatim.  Argument shapes follow the oo_* calls of oracle/search_oracle.cpp / bow_oracle.cpp.  The object
// graph is built as ref_harness.h describes (identity poses, pass-through camera, world position = (u, v, z)),
// so the reference's geometric prologues run for real and return exactly the floats the oracle is handed.
#include <set>
#include <utility>
#include <vector>

#include "ORBmatcher.h"
#include "ref_harness.h"
#include <sophus/sim3.hpp>

using namespace ORB_SLAM3;
using namespace ref_harness;

namespace {

void fill_featvec(DBoW2::FeatureVector& fv, int n_nodes, const int32_t* node_id, const int32_t* off, const int32_t* idx) {
  for (int k = 0; k < n_nodes; k++)
    for (int a = off[k]; a < off[k + 1]; a++) fv.addFeature((DBoW2::NodeId)node_id[k], (unsigned)idx[a]);
}

// n plain MapPoints, one per feature; features with flag 0 get no MapPoint
struct FeaturePoints {
  std::unique_ptr<MapPoint[]> mp;
  int n;
  FeaturePoints(int n_) : mp(new MapPoint[n_ > 0 ? n_ : 1]), n(n_) {}
  int index_of(const MapPoint* q) const { return (q && n && q >= &mp[0] && q < &mp[0] + n) ? (int)(q - &mp[0]) : -1; }
};

}  // namespace

extern "C" {

// ---- SearchByBoW.  mode 0: (KeyFrame, Frame); mode 1: (KeyFrame, KeyFrame).  has_mp = 1: the feature has a
// MapPoint, 2: it has one and the MapPoint is bad.  Each side's FeatureVector comes as (node ids, offsets,
// feature indices in insertion order); the merge-join with lower_bound is the reference's. ----
int ref_search_by_bow(int mode, const cuorb_frame_view* v1, const uint8_t* has_mp1, int nn1, const int32_t* node1,
                      const int32_t* off1, const int32_t* idx1, const cuorb_frame_view* v2, const uint8_t* has_mp2,
                      int nn2, const int32_t* node2, const int32_t* off2, const int32_t* idx2, float nnratio,
                      int check_ori, int32_t* match12) {
  KeyFrameFromView k1(v1);
  KeyFrameFromView k2(v2);
  FeaturePoints p1(v1->n), p2(v2->n);
  k1.K.mvpMapPoints.assign(v1->n, nullptr);
  k2.K.mvpMapPoints.assign(v2->n, nullptr);
  for (int i = 0; i < v1->n; i++)
    if (has_mp1[i]) { p1.mp[i].mbBad = has_mp1[i] == 2; k1.K.mvpMapPoints[i] = &p1.mp[i]; }
  for (int i = 0; i < v2->n; i++)
    if (has_mp2 && has_mp2[i]) { p2.mp[i].mbBad = has_mp2[i] == 2; k2.K.mvpMapPoints[i] = &p2.mp[i]; }
  fill_featvec(k1.K.mFeatVec, nn1, node1, off1, idx1);
  ORBmatcher matcher(nnratio, check_ori != 0);
  for (int i = 0; i < v1->n; i++) match12[i] = -1;
  int n;
  if (mode == 0) {
    Frame& F = k2.fv.F;
    fill_featvec(F.mFeatVec, nn2, node2, off2, idx2);
    std::vector<MapPoint*> matches;
    n = matcher.SearchByBoW(&k1.K, F, matches);
    for (int i2 = 0; i2 < (int)matches.size(); i2++) {
      const int i1 = p1.index_of(matches[i2]);
      if (matches[i2] && i1 < 0) return -1001;
      if (i1 >= 0) { if (match12[i1] != -1) return -1002; match12[i1] = i2; }
    }
  } else {
    fill_featvec(k2.K.mFeatVec, nn2, node2, off2, idx2);
    std::vector<MapPoint*> matches;
    n = matcher.SearchByBoW(&k1.K, &k2.K, matches);
    for (int i1 = 0; i1 < (int)matches.size(); i1++) {
      const int i2 = p2.index_of(matches[i1]);
      if (matches[i1] && i2 < 0) return -1001;
      match12[i1] = i2;
    }
  }
  return n;
}

// ---- SearchForInitialization.  prev_matched: [n1][2], updated in place as the reference does (:772-779). ----
int ref_search_for_initialization(const cuorb_frame_view* v1, const cuorb_frame_view* v2, float* prev_matched,
                                  int window, float nnratio, int check_ori, int32_t* match12) {
  // the two Frames share the class statics (image bounds, grid cell size): F2's grid is the one searched, so it
  // is built last
  FrameFromView f1(v1);
  FrameFromView f2(v2);
  std::vector<cv::Point2f> prev(v1->n);
  for (int i = 0; i < v1->n; i++) prev[i] = cv::Point2f(prev_matched[2 * i], prev_matched[2 * i + 1]);
  std::vector<int> m12;
  ORBmatcher matcher(nnratio, check_ori != 0);
  const int n = matcher.SearchForInitialization(f1.F, f2.F, prev, m12, window);
  for (int i = 0; i < v1->n; i++) {
    match12[i] = m12[i];
    prev_matched[2 * i] = prev[i].x; prev_matched[2 * i + 1] = prev[i].y;
  }
  return n;
}

// ---- SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, sAlreadyFound, th, ORBdist): relocalisation.
// found[i] = 1 puts query i's MapPoint into sAlreadyFound; kf_angle[i] = pKF->mvKeysUn[i].angle.
// binding_out[i2] = query bound to current feature i2 on return (-1 none, -2 the pre-existing binding). ----
int ref_search_by_projection_reloc(const cuorb_frame_view* cur, int n, const uint8_t* mp_desc, const float* u,
                                   const float* v, const int32_t* level, const float* kf_angle,
                                   const uint8_t* active, const uint8_t* found, float th, int orb_dist,
                                   int check_ori, int32_t* binding_out) {
  FrameFromView fv(cur);
  QueryPoints q(n, mp_desc, u, v, nullptr, level, active, fv.F.mfScaleFactor);
  KeyFrame kf;
  kf.N = n; kf.mvKeysUn.resize(n); kf.mvpMapPoints = q.ptr;
  for (int i = 0; i < n; i++) kf.mvKeysUn[i].angle = kf_angle[i];
  std::set<MapPoint*> already;
  for (int i = 0; i < n; i++) if (found && found[i]) already.insert(q.ptr[i]);
  ORBmatcher matcher(0.9f, check_ori != 0);
  const int r = matcher.SearchByProjection(fv.F, &kf, already, th, orb_dist);
  for (int k = 0; k < fv.F.N; k++) {
    MapPoint* b = fv.F.mvpMapPoints[k];
    binding_out[k] = !b ? -1 : q.index_of(b) >= 0 ? q.index_of(b) : -2;
  }
  return r;
}

// ---- SearchByProjection(KeyFrame*, Sim3, vpPoints, vpMatched, th, ratioHamming) (with_kfs = 0) and the overload
// that also records the source KeyFrame of every match (with_kfs = 1, :540-664).  kf->bound_with_obs[idx] = 1
// means vpMatched[idx] != NULL on entry.  matched_out[idx] = query matched to KeyFrame feature idx on return
// (-1 none, -2 the pre-existing entry).  For with_kfs = 1 the call fails (-1003) unless vpMatchedKF[idx] is the
// KeyFrame that was passed beside the matched point. ----
int ref_search_by_projection_sim3(const cuorb_frame_view* kfv, int n, const uint8_t* mp_desc, const float* u,
                                  const float* v, const int32_t* level, const uint8_t* active, int th,
                                  float ratio_hamming, int with_kfs, int32_t* matched_out) {
  KeyFrameFromView k(kfv);
  QueryPoints q(n, mp_desc, u, v, nullptr, level, active, k.K.mfScaleFactor);
  std::vector<MapPoint*> matched = k.K.mvpMapPoints;  // pre-existing matches (distinct objects from the queries)
  Sophus::Sim3f Scw;
  ORBmatcher matcher(0.9f, true);
  int r;
  if (!with_kfs) {
    r = matcher.SearchByProjection(&k.K, Scw, q.ptr, matched, th, ratio_hamming);
  } else {
    std::unique_ptr<KeyFrame[]> src(new KeyFrame[n > 0 ? n : 1]);
    std::vector<KeyFrame*> srcp(n), matched_kf(kfv->n, nullptr);
    for (int i = 0; i < n; i++) srcp[i] = &src[i];
    r = matcher.SearchByProjection(&k.K, Scw, q.ptr, srcp, matched, matched_kf, th, ratio_hamming);
    for (int idx = 0; idx < kfv->n; idx++) {
      const int i = q.index_of(matched[idx]);
      if (i >= 0 && matched_kf[idx] != srcp[i]) return -1003;
    }
  }
  for (int idx = 0; idx < kfv->n; idx++) {
    MapPoint* b = matched[idx];
    matched_out[idx] = !b ? -1 : q.index_of(b) >= 0 ? q.index_of(b) : -2;
  }
  return r;
}

// ---- Fuse.  sim3 = 0: Fuse(pKF, vpMapPoints, th, bRight = false) with the reprojection gates; sim3 = 1:
// Fuse(pKF, Scw, vpPoints, th, vpReplacePoint).  z[i] = camera-frame depth of query i: the reference computes
// ur = uv(0) - bf / z itself (:1208).  best_idx_out[i] = the KeyFrame feature query i fuses into, recovered by
// giving the queries one call at a time to a KeyFrame without MapPoints (the feature choice does not read
// pKF's MapPoints; the AddObservation the reference then makes records the index).  Returns nFused of ONE call
// over all queries against the KeyFrame as given (kf->bound_with_obs = features that already have a MapPoint,
// n_obs_kf / n_obs_q = Observations() of those and of the queries, which decide who replaces whom);
// replaced_out[i] (may be NULL): 1 = query i was Replace()d by the KeyFrame's point (sim3 = 1: recorded in
// vpReplacePoint), 2 = the KeyFrame's point was Replace()d by query i, 3 = query i was added to the KeyFrame. ----
int ref_fuse(const cuorb_frame_view* kfv, int n, const uint8_t* mp_desc, const float* u, const float* v,
             const float* z, const int32_t* level, const uint8_t* active, float th, int sim3,
             const int32_t* n_obs_kf, const int32_t* n_obs_q, int32_t* best_idx_out, int32_t* replaced_out) {
  ORBmatcher matcher(0.9f, true);
  Sophus::Sim3f Scw;
  int total_single = 0;
  {
    KeyFrameFromView k(kfv);
    k.K.mvpMapPoints.assign(kfv->n, nullptr);
    QueryPoints q(n, mp_desc, u, v, z, level, active, k.K.mfScaleFactor);
    for (int i = 0; i < n; i++) {
      std::vector<MapPoint*> one(1, q.ptr[i]);
      std::vector<MapPoint*> rep(1, nullptr);
      const int r = sim3 ? matcher.Fuse(&k.K, Scw, one, th, rep) : matcher.Fuse(&k.K, one, th, false);
      best_idx_out[i] = -1;
      if (r) {
        if (!q.mp[i].IsInKeyFrame(&k.K)) return -1004;
        best_idx_out[i] = std::get<0>(q.mp[i].GetIndexInKeyFrame(&k.K));
        k.K.mvpMapPoints[best_idx_out[i]] = nullptr;  // the next query sees an empty KeyFrame again
      }
      total_single += r;
    }
  }
  KeyFrameFromView k(kfv);
  if (n_obs_kf)
    for (int i = 0; i < kfv->n; i++) if (k.K.mvpMapPoints[i]) k.K.mvpMapPoints[i]->nObs = n_obs_kf[i];
  QueryPoints q(n, mp_desc, u, v, z, level, active, k.K.mfScaleFactor);
  if (n_obs_q) for (int i = 0; i < n; i++) q.mp[i].nObs = n_obs_q[i];
  std::vector<MapPoint*> rep(n, nullptr);
  const int r = sim3 ? matcher.Fuse(&k.K, Scw, q.ptr, th, rep) : matcher.Fuse(&k.K, q.ptr, th, false);
  if (replaced_out)
    for (int i = 0; i < n; i++) {
      replaced_out[i] = 0;
      if (sim3 ? rep[i] != nullptr : q.mp[i].GetReplaced() != nullptr) replaced_out[i] = 1;
      else if (q.mp[i].IsInKeyFrame(&k.K)) replaced_out[i] = 3;
      else if (best_idx_out[i] >= 0 && k.fv.bound[best_idx_out[i]].GetReplaced() == q.ptr[i]) replaced_out[i] = 2;
    }
  (void)total_single;
  return r;
}

// ---- SearchBySim3.  Query i1 of side 1 is KF1's MapPoint i1 projected into KF2 at (u12, v12); side 2 likewise.
// pre12[i1] (may be NULL) = index of the KF2 feature whose MapPoint vpMatches12[i1] holds on entry, -1 none. ----
int ref_search_by_sim3(const cuorb_frame_view* v1, const cuorb_frame_view* v2, const uint8_t* d1, const float* u12,
                       const float* v12, const int32_t* l12, const uint8_t* a1, const uint8_t* d2, const float* u21,
                       const float* v21, const int32_t* l21, const uint8_t* a2, float th, const int32_t* pre12,
                       int32_t* match12) {
  KeyFrameFromView k1(v1);
  KeyFrameFromView k2(v2);
  QueryPoints q1(v1->n, d1, u12, v12, nullptr, l12, a1, k2.K.mfScaleFactor);
  QueryPoints q2(v2->n, d2, u21, v21, nullptr, l21, a2, k1.K.mfScaleFactor);
  k1.K.mvpMapPoints = q1.ptr;
  k2.K.mvpMapPoints = q2.ptr;
  std::vector<MapPoint*> m12(v1->n, nullptr);
  if (pre12)
    for (int i = 0; i < v1->n; i++)
      if (pre12[i] >= 0) { m12[i] = q2.ptr[pre12[i]]; q2.mp[pre12[i]].mObservations[&k2.K] = std::make_tuple(pre12[i], -1); }
  Sophus::Sim3f S12;
  ORBmatcher matcher(0.9f, true);
  const int r = matcher.SearchBySim3(&k1.K, &k2.K, m12, S12, th);
  for (int i = 0; i < v1->n; i++) match12[i] = q2.index_of(m12[i]);
  return r;
}

// ---- SearchForTriangulation.  pose1 = T1w as 9 floats R (row-major) + 3 floats t; KF2 has the identity pose, so
// T12 = T1w and the epipole is the projection of KF1's centre.  ep_out[2] and F12_out[9] (row-major) return what
// the reference's prologue and Pinhole::epipolarConstrain computed from that pose (K = I), for the caller to hand
// to the oracle.  has_mp1/2 = GetMapPoint(idx) != NULL. ----
int ref_search_for_triangulation(const cuorb_frame_view* v1, const uint8_t* has_mp1, int nn1, const int32_t* node1,
                                 const int32_t* off1, const int32_t* idx1, const cuorb_frame_view* v2,
                                 const uint8_t* has_mp2, int nn2, const int32_t* node2, const int32_t* off2,
                                 const int32_t* idx2, const float* pose1, int only_stereo, int coarse,
                                 int check_ori, float* ep_out, float* F12_out, int32_t* match12) {
  KeyFrameFromView k1(v1);
  KeyFrameFromView k2(v2);
  UnitPinhole cam1, cam2;   // K = I; KF2's project() is only used for the epipole, z = 1 below
  k1.K.mpCamera = &cam1; k2.K.mpCamera = &cam2;
  FeaturePoints p1(v1->n), p2(v2->n);
  k1.K.mvpMapPoints.assign(v1->n, nullptr);
  k2.K.mvpMapPoints.assign(v2->n, nullptr);
  for (int i = 0; i < v1->n; i++) if (has_mp1[i]) k1.K.mvpMapPoints[i] = &p1.mp[i];
  for (int i = 0; i < v2->n; i++) if (has_mp2[i]) k2.K.mvpMapPoints[i] = &p2.mp[i];
  fill_featvec(k1.K.mFeatVec, nn1, node1, off1, idx1);
  fill_featvec(k2.K.mFeatVec, nn2, node2, off2, idx2);
  Eigen::Matrix3f R;
  for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) R(r, c) = pose1[3 * r + c];
  k1.K.mTcw = Sophus::SE3f(R, Eigen::Vector3f(pose1[9], pose1[10], pose1[11]));
  // what the reference's prologue will compute (:926-945), through the same stand-in types
  {
    const Eigen::Vector3f C2 = k2.K.GetPose() * k1.K.GetCameraCenter();
    const Eigen::Vector2f ep = k2.K.mpCamera->project(C2);
    ep_out[0] = ep(0); ep_out[1] = ep(1);
    const Sophus::SE3f T12 = k1.K.GetPose() * k2.K.GetPoseInverse();
    const Eigen::Matrix3f F12 = cam1.toK_().transpose().inverse() * Sophus::SO3f::hat(T12.translation()) *
                                T12.rotationMatrix() * cam2.toK_().inverse();   // Pinhole.cpp:153-156
    for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) F12_out[3 * r + c] = F12(r, c);
  }
  std::vector<std::pair<std::size_t, std::size_t>> pairs;
  ORBmatcher matcher(0.6f, check_ori != 0);
  const int n = matcher.SearchForTriangulation(&k1.K, &k2.K, pairs, only_stereo != 0, coarse != 0);
  for (int i = 0; i < v1->n; i++) match12[i] = -1;
  for (const auto& pr : pairs) match12[pr.first] = (int32_t)pr.second;
  return n;
}

}  // extern "C"
