// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
// Flat C entry points (ctypes) that drive the reference's OWN code for the two-camera (Frame::Nleft != -1)
// branches of the tracking-time searches — the unmodified src/ORBmatcher.cc:
//   SearchByProjection(Frame&, const vector<MapPoint*>&, th, bFarPoints, thFarPoints)   :38-227
//   SearchByProjection(Frame& Cur, const Frame& Last, th, bMono)                         :1684-1908
// over Frame::AssignFeaturesToGrid / GetFeaturesInArea / GetRelativePoseTrl lifted verbatim from src/Frame.cc.
// A rig Frame is assembled from two cuorb_frame_views: the left camera's features become indices [0, Nleft)
// (mvKeys, mGrid), the right camera's [Nleft, N) (mvKeysRight, mGridRight, descriptor rows idx + Nleft).
#include <vector>

#include "KeyFrame.h"
#include "ORBmatcher.h"
#include "fake_triangulate.h"
#include "ref_harness.h"

using namespace ORB_SLAM3;
using namespace ref_harness;

namespace {

struct RigFrame {
  Frame                       F;
  PassThroughCamera           cam;
  std::unique_ptr<MapPoint[]> bound;
  std::vector<uint8_t>        desc;
  RigFrame(const cuorb_frame_view* l, const cuorb_frame_view* r, const int32_t* l2r, const int32_t* r2l) {
    const int nl = l->n, nr = r->n, n = nl + nr;
    F.N = n; F.Nleft = nl; F.Nright = nr;
    const cv::KeyPoint* kl = (const cv::KeyPoint*)l->keys_un;
    const cv::KeyPoint* kr = (const cv::KeyPoint*)r->keys_un;
    F.mvKeys.assign(kl, kl + nl);
    F.mvKeysRight.assign(kr, kr + nr);
    F.mvKeysUn = F.mvKeys;
    desc.resize((size_t)(n > 0 ? n : 1) * 32);
    if (nl) memcpy(desc.data(), l->descriptors, (size_t)nl * 32);
    if (nr) memcpy(desc.data() + (size_t)nl * 32, r->descriptors, (size_t)nr * 32);
    F.mDescriptors = desc_mat(desc.data(), n);
    F.mvuRight.assign(nl, -1.f);                                   // src/Frame.cc:1313
    F.mvpMapPoints.assign(n, nullptr);
    F.mvbOutlier.assign(n, false);
    bound.reset(new MapPoint[n > 0 ? n : 1]);
    for (int i = 0; i < nl; i++)
      if (l->bound_with_obs && l->bound_with_obs[i]) { bound[i].nObs = 1; F.mvpMapPoints[i] = &bound[i]; }
    for (int i = 0; i < nr; i++)
      if (r->bound_with_obs && r->bound_with_obs[i]) { bound[nl + i].nObs = 1; F.mvpMapPoints[nl + i] = &bound[nl + i]; }
    Frame::mnMinX = l->min_x; Frame::mnMinY = l->min_y; Frame::mnMaxX = l->max_x; Frame::mnMaxY = l->max_y;
    Frame::mfGridElementWidthInv = l->grid_w_inv; Frame::mfGridElementHeightInv = l->grid_h_inv;
    F.mb = l->mb; F.mbf = l->mbf;
    F.mvScaleFactors.assign(l->scale_factors, l->scale_factors + l->nlevels);
    F.mnScaleLevels = l->nlevels;
    F.mvLeftToRightMatch.assign(nl, -1);
    F.mvRightToLeftMatch.assign(nr, -1);
    if (l2r) F.mvLeftToRightMatch.assign(l2r, l2r + nl);
    if (r2l) F.mvRightToLeftMatch.assign(r2l, r2l + nr);
    F.mpCamera = &cam; F.mpCamera2 = &cam;
    F.AssignFeaturesToGrid();
  }
};

}  // namespace

extern "C" {

// flags: bit0 mbTrackInView, bit1 isBad(), bit2 Observations() > 0, bit3 mbTrackInViewR.
// binding_out[k], k in [0, Nleft + Nright): MapPoint bound to rig feature k on return (-1 none, -2 pre-existing).
int ref_search_by_projection_rig(const cuorb_frame_view* left, const cuorb_frame_view* right, const int32_t* l2r,
                                 const int32_t* r2l, int M, const uint8_t* mp_desc, const float* px, const float* py,
                                 const int32_t* level, const float* view_cos, const float* pxr, const float* pyr,
                                 const int32_t* level_r, const float* view_cos_r, const float* depth,
                                 const uint8_t* flags, float th, int far_points, float th_far, float nn_ratio,
                                 int32_t* binding_out, int32_t* per_mp_out) {
  std::unique_ptr<MapPoint[]> mps(new MapPoint[M > 0 ? M : 1]);
  std::vector<MapPoint*> ptrs(M);
  for (int i = 0; i < M; i++) {
    MapPoint& p = mps[i];
    p.mbTrackInView = (flags[i] & 1) != 0; p.mbTrackInViewR = (flags[i] & 8) != 0;
    p.mbBad = (flags[i] & 2) != 0;
    p.nObs = (flags[i] & 4) ? 1 : 0;
    p.mTrackProjX = px[i]; p.mTrackProjY = py[i]; p.mnTrackScaleLevel = level[i]; p.mTrackViewCos = view_cos[i];
    p.mTrackProjXR = pxr[i]; p.mTrackProjYR = pyr[i]; p.mnTrackScaleLevelR = level_r[i]; p.mTrackViewCosR = view_cos_r[i];
    p.mTrackDepth = depth ? depth[i] : 0.f;
    p.mDescriptor = desc_mat(mp_desc + (size_t)i * 32, 1);
    ptrs[i] = &p;
  }
  RigFrame rf(left, right, l2r, r2l);
  ORBmatcher matcher(nn_ratio, true);
  const int n = matcher.SearchByProjection(rf.F, ptrs, th, far_points != 0, th_far);
  for (int k = 0; k < rf.F.N; k++) {
    MapPoint* b = rf.F.mvpMapPoints[k];
    binding_out[k] = !b ? -1 : (M && b >= &mps[0] && b < &mps[0] + M) ? (int32_t)(b - &mps[0]) : -2;
  }
  // per_mp_out[4 * i ..]: the rig features (ascending, -1 padded) that point at MapPoint i right after IT was
  // processed, obtained by feeding the MapPoints one call at a time to a second Frame (the loop body only
  // communicates through F.mvpMapPoints, so this is the same computation)
  if (per_mp_out) {
    RigFrame one(left, right, l2r, r2l);
    int total = 0;
    for (int i = 0; i < M; i++) {
      std::vector<MapPoint*> v(1, ptrs[i]);
      total += matcher.SearchByProjection(one.F, v, th, far_points != 0, th_far);
      int c = 0;
      for (int q = 0; q < 4; q++) per_mp_out[4 * i + q] = -1;
      for (int k = 0; k < one.F.N; k++)
        if (one.F.mvpMapPoints[k] == ptrs[i]) { if (c >= 4) return -1005; per_mp_out[4 * i + c++] = k; }
    }
    if (total != n) return -1000 - total;
  }
  return n;
}

// Last-frame feature i has world position (u, v, z) seen from the identity pose through the pass-through camera,
// so the left projection is (u, v) and the right one project(Trl * x3Dc) = (u + trl_x, v + trl_y) for the pure
// translation Trl given.  binding_out as above, indexed by current rig feature, holding last-frame feature ids.
int ref_search_by_projection_frame_rig(const cuorb_frame_view* left, const cuorb_frame_view* right, int n_last,
                                       const uint8_t* last_desc, const float* u, const float* v, const float* z,
                                       float trl_x, float trl_y, const int32_t* last_octave,
                                       const float* last_angle, const uint8_t* flags, float th, int forward,
                                       int backward, int check_ori, int32_t* binding_out) {
  RigFrame rf(left, right, nullptr, nullptr);
  rf.F.mTrl = Sophus::SE3f(Eigen::Matrix3f::Identity(), Eigen::Vector3f(trl_x, trl_y, 0.f));
  Frame last;
  std::unique_ptr<MapPoint[]> mps(new MapPoint[n_last > 0 ? n_last : 1]);
  last.N = n_last; last.Nleft = -1;
  last.mvKeys.resize(n_last); last.mvKeysUn.resize(n_last);
  last.mvpMapPoints.assign(n_last, nullptr);
  last.mvbOutlier.assign(n_last, false);
  for (int i = 0; i < n_last; i++) {
    last.mvKeys[i].octave = last_octave[i]; last.mvKeys[i].angle = last_angle[i];
    last.mvKeysUn[i] = last.mvKeys[i];
    if (!(flags[i] & 1)) continue;
    mps[i].mWorldPos = Eigen::Vector3f(u[i], v[i], z[i]);
    mps[i].nObs = (flags[i] & 4) ? 1 : 0;
    mps[i].mDescriptor = desc_mat(last_desc + (size_t)i * 32, 1);
    last.mvpMapPoints[i] = &mps[i];
  }
  const float tz = forward ? rf.F.mb + 1.f : backward ? -(rf.F.mb + 1.f) : 0.f;
  last.mTcw = Sophus::SE3f(Eigen::Matrix3f::Identity(), Eigen::Vector3f(0.f, 0.f, tz));
  ORBmatcher matcher(0.9f, check_ori != 0);
  const int n = matcher.SearchByProjection(rf.F, last, th, false);
  for (int k = 0; k < rf.F.N; k++) {
    MapPoint* b = rf.F.mvpMapPoints[k];
    binding_out[k] = !b ? -1 : (n_last && b >= &mps[0] && b < &mps[0] + n_last) ? (int32_t)(b - &mps[0]) : -2;
  }
  return n;
}

// ---- void Frame::ComputeStereoFishEyeMatches() (src/Frame.cc:1303-1355), the reference's own body, with
// BFMatcher::knnMatch forwarded to the cv2-pinned 2-NN and KannalaBrandt8::TriangulateMatches replaced by
// fake_triangulate.h.  Outputs: mvLeftToRightMatch [nl], mvRightToLeftMatch [nr], mvDepth [nl], mvStereo3Dpoints [nl x 3]. ----
int ref_fisheye_stereo(int nl, const cv::KeyPoint* kl, const uint8_t* dl, int mono_left, int nr, const cv::KeyPoint* kr,
                       const uint8_t* dr, int mono_right, int nlevels, const float* level_sigma2, int32_t* l2r, int32_t* r2l,
                       float* depth, float* p3d) {
  Frame F;
  KannalaBrandt8 cam, cam2;
  cam.triangulate_hook = fake_triangulate;
  F.mpCamera = &cam; F.mpCamera2 = &cam2;
  F.Nleft = nl; F.Nright = nr; F.N = nl + nr;
  F.mvKeys.assign(kl, kl + nl); F.mvKeysRight.assign(kr, kr + nr);
  F.mDescriptors = desc_mat(dl, nl); F.mDescriptorsRight = desc_mat(dr, nr);
  F.monoLeft = mono_left; F.monoRight = mono_right;
  F.mvLevelSigma2.assign(level_sigma2, level_sigma2 + nlevels);
  F.mRlr = Eigen::Matrix3f::Identity();
  F.ComputeStereoFishEyeMatches();
  int n = 0;
  for (int i = 0; i < nl; i++) {
    l2r[i] = F.mvLeftToRightMatch[i]; depth[i] = F.mvDepth[i];
    for (int c = 0; c < 3; c++) p3d[3 * i + c] = F.mvStereo3Dpoints[i](c);
    n += l2r[i] >= 0;
  }
  for (int i = 0; i < nr; i++) r2l[i] = F.mvRightToLeftMatch[i];
  return n;
}

// ---- int ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame& F, vpMapPointMatches) with BOTH a rig KeyFrame and a rig
// Frame (src/ORBmatcher.cc:237-420, the F.Nleft != -1 branch :306-331, :371-396).  The KeyFrame's features are
// [0, NLeft) = kl (mvKeys) and [NLeft, N1) = kr (mvKeysRight), has_mp1 over all N1 (1 good MapPoint, 2 bad one);
// F likewise from fl / fr.  FeatureVectors index the combined ranges.  match_l[i1] / match_r[i1] = F feature (combined
// index) holding KeyFrame feature i1's MapPoint on return, split by camera; -1 none. ----
int ref_search_by_bow_rig(const cuorb_frame_view* kl, const cuorb_frame_view* kr, const uint8_t* has_mp1, int nn1,
                          const int32_t* node1, const int32_t* off1, const int32_t* idx1, const cuorb_frame_view* fl,
                          const cuorb_frame_view* fr, int nn2, const int32_t* node2, const int32_t* off2,
                          const int32_t* idx2, float nnratio, int check_ori, int32_t* match_l, int32_t* match_r) {
  RigFrame rf(fl, fr, nullptr, nullptr);
  rf.F.mvpMapPoints.assign(rf.F.N, nullptr);
  for (int k = 0; k < nn2; k++)
    for (int a = off2[k]; a < off2[k + 1]; a++) rf.F.mFeatVec.addFeature((DBoW2::NodeId)node2[k], (unsigned)idx2[a]);
  const int n1l = kl->n, n1r = kr->n, n1 = n1l + n1r;
  KeyFrame K;
  PassThroughCamera cam;
  K.N = n1; K.NLeft = n1l; K.NRight = n1r;
  const cv::KeyPoint* a = (const cv::KeyPoint*)kl->keys_un;
  const cv::KeyPoint* b = (const cv::KeyPoint*)kr->keys_un;
  K.mvKeys.assign(a, a + n1l); K.mvKeysRight.assign(b, b + n1r); K.mvKeysUn = K.mvKeys;
  std::vector<uint8_t> desc((size_t)(n1 > 0 ? n1 : 1) * 32);
  if (n1l) memcpy(desc.data(), kl->descriptors, (size_t)n1l * 32);
  if (n1r) memcpy(desc.data() + (size_t)n1l * 32, kr->descriptors, (size_t)n1r * 32);
  K.mDescriptors = desc_mat(desc.data(), n1);
  K.mpCamera = &cam; K.mpCamera2 = &cam;
  std::unique_ptr<MapPoint[]> mps(new MapPoint[n1 > 0 ? n1 : 1]);
  K.mvpMapPoints.assign(n1, nullptr);
  for (int i = 0; i < n1; i++) if (has_mp1[i]) { mps[i].mbBad = has_mp1[i] == 2; K.mvpMapPoints[i] = &mps[i]; }
  for (int k = 0; k < nn1; k++)
    for (int q = off1[k]; q < off1[k + 1]; q++) K.mFeatVec.addFeature((DBoW2::NodeId)node1[k], (unsigned)idx1[q]);
  std::vector<MapPoint*> matches;
  ORBmatcher matcher(nnratio, check_ori != 0);
  const int n = matcher.SearchByBoW(&K, rf.F, matches);
  for (int i = 0; i < n1; i++) match_l[i] = match_r[i] = -1;
  for (int i2 = 0; i2 < (int)matches.size(); i2++) {
    if (!matches[i2]) continue;
    const int i1 = (int)(matches[i2] - &mps[0]);
    if (i1 < 0 || i1 >= n1) return -1001;
    int32_t* dst = i2 < rf.F.Nleft ? match_l : match_r;
    if (dst[i1] != -1) return -1002;
    dst[i1] = i2;
  }
  return n;
}

// ---- int ORBmatcher::Fuse(KeyFrame* pKF, vpMapPoints, th, bRight) on a rig KeyFrame, either camera
// (src/ORBmatcher.cc:1149-1328: bRight selects GetRightPose / mpCamera2 / mGridRight / mvKeysRight and adds NLeft to
// the chosen index, :1156-1164, :1236, :1270-1272).  Queries as in ref_fuse (identity poses, pass-through cameras,
// world position (u, v, z)); best_idx_out[i] = KeyFrame feature (combined index) query i fuses into, recovered one
// query at a time on a KeyFrame without MapPoints.  A rig KeyFrame's mvuRight is all -1 (src/Frame.cc:1313), so only
// the monocular 5.99 gate applies. ----
int ref_fuse_rig(const cuorb_frame_view* kl, const cuorb_frame_view* kr, int right, int n, const uint8_t* mp_desc,
                 const float* u, const float* v, const float* z, const int32_t* level, const uint8_t* active, float th,
                 int32_t* best_idx_out) {
  RigFrame rf(kl, kr, nullptr, nullptr);
  const Frame& F = rf.F;
  KeyFrame K;
  K.N = F.N; K.NLeft = F.Nleft; K.NRight = F.Nright;
  K.mnGridCols = FRAME_GRID_COLS; K.mnGridRows = FRAME_GRID_ROWS;
  K.mfGridElementWidthInv = Frame::mfGridElementWidthInv; K.mfGridElementHeightInv = Frame::mfGridElementHeightInv;
  K.fx = 1.f; K.fy = 1.f; K.cx = 0.f; K.cy = 0.f; K.mbf = F.mbf; K.mb = F.mb;
  K.mvKeys = F.mvKeys; K.mvKeysRight = F.mvKeysRight; K.mvKeysUn = F.mvKeys;
  K.mvuRight.assign(std::max(F.Nleft, F.Nright), -1.f);   // the reference indexes it with the right camera's local idx (:1248)
  K.mDescriptors = F.mDescriptors;
  K.mnScaleLevels = F.mnScaleLevels; K.mfScaleFactor = kl->nlevels > 1 ? kl->scale_factors[1] : 1.2f;
  K.mfLogScaleFactor = std::log(K.mfScaleFactor);
  K.mvScaleFactors = F.mvScaleFactors;
  K.mvLevelSigma2.resize(F.mnScaleLevels); K.mvInvLevelSigma2.resize(F.mnScaleLevels);
  for (int l = 0; l < F.mnScaleLevels; l++) {
    K.mvLevelSigma2[l] = F.mvScaleFactors[l] * F.mvScaleFactors[l];
    K.mvInvLevelSigma2[l] = 1.0f / K.mvLevelSigma2[l];
  }
  K.mnMinX = Frame::mnMinX; K.mnMinY = Frame::mnMinY; K.mnMaxX = Frame::mnMaxX; K.mnMaxY = Frame::mnMaxY;
  K.mGrid.resize(K.mnGridCols); K.mGridRight.resize(K.mnGridCols);
  for (int i = 0; i < K.mnGridCols; i++) {
    K.mGrid[i].resize(K.mnGridRows); K.mGridRight[i].resize(K.mnGridRows);
    for (int j = 0; j < K.mnGridRows; j++) { K.mGrid[i][j] = F.mGrid[i][j]; K.mGridRight[i][j] = F.mGridRight[i][j]; }
  }
  PassThroughCamera cam, cam2;
  K.mpCamera = &cam; K.mpCamera2 = &cam2;
  K.mvpMapPoints.assign(K.N, nullptr);
  QueryPoints q(n, mp_desc, u, v, z, level, active, K.mfScaleFactor);
  ORBmatcher matcher(0.9f, true);
  int total = 0;
  for (int i = 0; i < n; i++) {
    std::vector<MapPoint*> one(1, q.ptr[i]);
    const int r = matcher.Fuse(&K, one, th, right != 0);
    best_idx_out[i] = -1;
    if (r) {
      if (!q.mp[i].IsInKeyFrame(&K)) return -1004;
      best_idx_out[i] = std::get<0>(q.mp[i].GetIndexInKeyFrame(&K));
      K.mvpMapPoints[best_idx_out[i]] = nullptr;
    }
    total += r;
  }
  return total;
}

}  // extern "C"
