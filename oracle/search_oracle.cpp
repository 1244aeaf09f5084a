// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  See orb_oracle.hpp for scope and parity status.
//
// Oracle for the rest of SURVEY.md §8(f) row N3: the ORBmatcher searches that are not on the
// per-frame tracking path, restated on flattened inputs (pinhole, NLeft == -1 branches).  The
// geometric gates that need Sophus/Eigen/MapPoint state (pose transform, camera projection,
// distance-invariance range, viewing angle, PredictScale) stay with the caller: every query
// arrives as (active, u, v, predicted level, descriptor).  What is restated is everything after
// that: the window search, the per-candidate gates, the best / second-best bookkeeping with the
// reference's visiting order, the order-dependent binding, and the rotation histogram.
//   SearchByProjection(Frame&, KeyFrame*, set<MapPoint*>&, th, ORBdist)        src/ORBmatcher.cc:1910-2036
//   SearchByProjection(KeyFrame*, Sim3f&, vpPoints, vpMatched, th, ratio)      src/ORBmatcher.cc:423-538
//   SearchByProjection(KeyFrame*, Sim3&, vpPoints, vpPointsKFs, ...)           src/ORBmatcher.cc:540-664
//   Fuse(KeyFrame*, vpMapPoints, th, bRight=false)                             src/ORBmatcher.cc:1149-1328
//   Fuse(KeyFrame*, Sim3f&, vpPoints, th, vpReplacePoint)                      src/ORBmatcher.cc:1330-1452
//   SearchBySim3(KeyFrame*, KeyFrame*, vpMatches12, S12, th)                   src/ORBmatcher.cc:1454-1682
//   SearchForInitialization(Frame&, Frame&, vbPrevMatched, vnMatches12, win)   src/ORBmatcher.cc:666-782
//   SearchForTriangulation(KeyFrame*, KeyFrame*, vMatchedPairs, bOnlyStereo, bCoarse)
//                                                                              src/ORBmatcher.cc:916-1147
//   Pinhole::epipolarConstrain                                                 src/CameraModels/Pinhole.cpp:143-172
// KeyFrame::GetFeaturesInArea (src/KeyFrame.cc:780-834) is Frame::GetFeaturesInArea without the
// level filter and on the same 64x48 grid (the KeyFrame copies the Frame's grid), so
// FrameView::features_in_area(x, y, r, -1, -1) stands for it.
// Parity status: pinned to the reference's own functions (the unmodified src/ORBmatcher.cc in oracle/_ref, driven
// by oracle/ref_compat/ref_capi_search.cpp; tests/test_oracle_vs_ref_search.py), bit for bit.
#include <climits>
#include <cmath>
#include <vector>

#include "../include/cuorb.h"  // POD struct layouts only
#include "orb_oracle.hpp"

namespace orb_oracle {

namespace {

const int TH_HIGH = 100, TH_LOW = 50, HISTO_LENGTH = 30;

FrameView view_of(const cuorb_frame_view* f) {
  FrameView F;
  F.N = f->n; F.keysUn = (const KeyPoint*)f->keys_un; F.desc = f->descriptors; F.uRight = f->u_right;
  F.bound_with_obs.assign(f->n, 0);
  if (f->bound_with_obs) F.bound_with_obs.assign(f->bound_with_obs, f->bound_with_obs + f->n);
  F.minX = f->min_x; F.minY = f->min_y; F.maxX = f->max_x; F.maxY = f->max_y;
  F.gridWInv = f->grid_w_inv; F.gridHInv = f->grid_h_inv; F.mb = f->mb; F.mbf = f->mbf;
  F.scaleFactors.assign(f->scale_factors, f->scale_factors + f->nlevels);
  F.assign_features_to_grid();
  return F;
}

// ORBmatcher::ComputeThreeMaxima (src/ORBmatcher.cc:2038-2071) on bin sizes
void three_maxima(const int* histo, int& ind1, int& ind2, int& ind3) {
  int max1 = 0, max2 = 0, max3 = 0;
  ind1 = ind2 = ind3 = -1;
  for (int i = 0; i < HISTO_LENGTH; i++) {
    const int s = histo[i];
    if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
    else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
    else if (s > max3) { max3 = s; ind3 = i; }
  }
  if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
  else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
}

int rot_bin(float a1, float a2) {  // :1987-1996 and its copies
  float rot = a1 - a2;
  if (rot < 0.0) rot += 360.0f;
  int bin = (int)std::round(rot * (1.0f / HISTO_LENGTH));
  if (bin == HISTO_LENGTH) bin = 0;
  return bin;
}

}  // namespace

// :1910-2036.  active = pMP && !isBad && !sAlreadyFound && minDistance <= dist3D <= maxDistance;
// Cur.bound_with_obs = CurrentFrame.mvpMapPoints[i2] != NULL on entry.
int search_by_projection_reloc(FrameView& Cur, int n, const uint8_t* mpDesc, const float* u, const float* v,
                               const int32_t* level, const float* kfAngle, const uint8_t* active, float th,
                               int ORBdist, bool checkOri, int32_t* bestIdxOut, int32_t* rotBinOut,
                               int32_t* keepBinsOut) {
  int nmatches = 0;
  int hist[HISTO_LENGTH] = {0};
  for (int i = 0; i < n; i++) {
    bestIdxOut[i] = -1; rotBinOut[i] = -1;
    if (!active[i]) continue;
    if (u[i] < Cur.minX || u[i] > Cur.maxX) continue;
    if (v[i] < Cur.minY || v[i] > Cur.maxY) continue;
    const int nPredictedLevel = level[i];
    const float radius = th * Cur.scaleFactors[nPredictedLevel];
    const std::vector<int> vIndices2 = Cur.features_in_area(u[i], v[i], radius, nPredictedLevel - 1, nPredictedLevel + 1);
    if (vIndices2.empty()) continue;
    int bestDist = 256, bestIdx2 = -1;
    for (int i2 : vIndices2) {
      if (Cur.bound_with_obs[i2]) continue;
      const int dist = descriptor_distance(mpDesc + (size_t)i * 32, Cur.desc + (size_t)i2 * 32);
      if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
    }
    if (bestDist <= ORBdist && bestIdx2 >= 0) {
      Cur.bound_with_obs[bestIdx2] = 1;
      bestIdxOut[i] = bestIdx2;
      nmatches++;
      if (checkOri) {
        const int bin = rot_bin(kfAngle[i], Cur.keysUn[bestIdx2].angle);
        hist[bin]++; rotBinOut[i] = bin;
      }
    }
  }
  keepBinsOut[0] = keepBinsOut[1] = keepBinsOut[2] = -1;
  if (checkOri) {
    int i1, i2, i3;
    three_maxima(hist, i1, i2, i3);
    keepBinsOut[0] = i1; keepBinsOut[1] = i2; keepBinsOut[2] = i3;
    for (int b = 0; b < HISTO_LENGTH; b++)
      if (b != i1 && b != i2 && b != i3) nmatches -= hist[b];
  }
  return nmatches;
}

// :423-538 / :540-664.  active = !isBad && !spAlreadyFound && depth > 0 && IsInImage && distance
// range && viewing angle; KF.bound_with_obs = vpMatched[idx] != NULL on entry.
int search_by_projection_sim3(FrameView& KF, int n, const uint8_t* mpDesc, const float* u, const float* v,
                              const int32_t* level, const uint8_t* active, int th, float ratioHamming,
                              int32_t* bestIdxOut) {
  int nmatches = 0;
  for (int i = 0; i < n; i++) {
    bestIdxOut[i] = -1;
    if (!active[i]) continue;
    const int nPredictedLevel = level[i];
    const float radius = th * KF.scaleFactors[nPredictedLevel];
    const std::vector<int> vIndices = KF.features_in_area(u[i], v[i], radius, -1, -1);
    if (vIndices.empty()) continue;
    int bestDist = 256, bestIdx = -1;
    for (int idx : vIndices) {
      if (KF.bound_with_obs[idx]) continue;
      const int kpLevel = KF.keysUn[idx].octave;
      if (kpLevel < nPredictedLevel - 1 || kpLevel > nPredictedLevel) continue;
      const int dist = descriptor_distance(mpDesc + (size_t)i * 32, KF.desc + (size_t)idx * 32);
      if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
    }
    if (bestDist <= TH_LOW * ratioHamming) {
      KF.bound_with_obs[bestIdx] = 1;
      bestIdxOut[i] = bestIdx;
      nmatches++;
    }
  }
  return nmatches;
}

// :1149-1328 (checkReprojection, bRight = false) and :1330-1452 (no reprojection gate).
// active = the geometric gates before PredictScale; ur = uv(0) - bf * invz (:1208).
int fuse(const FrameView& KF, const float* invLevelSigma2, int n, const uint8_t* mpDesc, const float* u,
         const float* v, const float* ur, const int32_t* level, const uint8_t* active, float th,
         bool checkReprojection, int32_t* bestIdxOut) {
  int nFused = 0;
  for (int i = 0; i < n; i++) {
    bestIdxOut[i] = -1;
    if (!active[i]) continue;
    const int nPredictedLevel = level[i];
    const float radius = th * KF.scaleFactors[nPredictedLevel];
    const std::vector<int> vIndices = KF.features_in_area(u[i], v[i], radius, -1, -1);
    if (vIndices.empty()) continue;
    int bestDist = 256, bestIdx = -1;
    for (int idx : vIndices) {
      const KeyPoint& kp = KF.keysUn[idx];
      const int kpLevel = kp.octave;
      if (kpLevel < nPredictedLevel - 1 || kpLevel > nPredictedLevel) continue;
      if (checkReprojection) {
        if (KF.uRight && KF.uRight[idx] >= 0) {
          const float kpx = kp.x, kpy = kp.y, kpr = KF.uRight[idx];
          const float ex = u[i] - kpx, ey = v[i] - kpy, er = ur[i] - kpr;
          const float e2 = ex * ex + ey * ey + er * er;
          if (e2 * invLevelSigma2[kpLevel] > 7.8) continue;
        } else {
          const float kpx = kp.x, kpy = kp.y;
          const float ex = u[i] - kpx, ey = v[i] - kpy;
          const float e2 = ex * ex + ey * ey;
          if (e2 * invLevelSigma2[kpLevel] > 5.99) continue;
        }
      }
      const int dist = descriptor_distance(mpDesc + (size_t)i * 32, KF.desc + (size_t)idx * 32);
      if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
    }
    if (bestDist <= TH_LOW) { bestIdxOut[i] = bestIdx; nFused++; }
  }
  return nFused;
}

// :1454-1682.  One direction (:1501-1583 / :1586-1664); the agreement check (:1667-1679) follows.
static void sim3_direction(const FrameView& Target, int n, const uint8_t* mpDesc, const float* u, const float* v,
                           const int32_t* level, const uint8_t* active, float th, std::vector<int>& vnMatch) {
  vnMatch.assign(n, -1);
  for (int i = 0; i < n; i++) {
    if (!active[i]) continue;
    const int nPredictedLevel = level[i];
    const float radius = th * Target.scaleFactors[nPredictedLevel];
    const std::vector<int> vIndices = Target.features_in_area(u[i], v[i], radius, -1, -1);
    if (vIndices.empty()) continue;
    int bestDist = INT_MAX, bestIdx = -1;
    for (int idx : vIndices) {
      const KeyPoint& kp = Target.keysUn[idx];
      if (kp.octave < nPredictedLevel - 1 || kp.octave > nPredictedLevel) continue;
      const int dist = descriptor_distance(mpDesc + (size_t)i * 32, Target.desc + (size_t)idx * 32);
      if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
    }
    if (bestDist <= TH_HIGH) vnMatch[i] = bestIdx;
  }
}

int search_by_sim3(const FrameView& KF1, const FrameView& KF2, const uint8_t* mpDesc1, const float* u12,
                   const float* v12, const int32_t* level12, const uint8_t* active1, const uint8_t* mpDesc2,
                   const float* u21, const float* v21, const int32_t* level21, const uint8_t* active2, float th,
                   int32_t* match12Out) {
  std::vector<int> vnMatch1, vnMatch2;
  sim3_direction(KF2, KF1.N, mpDesc1, u12, v12, level12, active1, th, vnMatch1);
  sim3_direction(KF1, KF2.N, mpDesc2, u21, v21, level21, active2, th, vnMatch2);
  int nFound = 0;
  for (int i1 = 0; i1 < KF1.N; i1++) {
    match12Out[i1] = -1;
    const int idx2 = vnMatch1[i1];
    if (idx2 >= 0 && vnMatch2[idx2] == i1) { match12Out[i1] = idx2; nFound++; }
  }
  return nFound;
}

// :666-782
int search_for_initialization(const FrameView& F1, const FrameView& F2, float* vbPrevMatched, int windowSize,
                              float mfNNratio, bool checkOri, int32_t* vnMatches12) {
  int nmatches = 0;
  for (int i = 0; i < F1.N; i++) vnMatches12[i] = -1;
  std::vector<int> rotHist[HISTO_LENGTH];
  std::vector<int> vMatchedDistance(F2.N, INT_MAX), vnMatches21(F2.N, -1);
  for (int i1 = 0; i1 < F1.N; i1++) {
    const KeyPoint kp1 = F1.keysUn[i1];
    const int level1 = kp1.octave;
    if (level1 > 0) continue;
    const std::vector<int> vIndices2 =
        F2.features_in_area(vbPrevMatched[2 * i1], vbPrevMatched[2 * i1 + 1], (float)windowSize, level1, level1);
    if (vIndices2.empty()) continue;
    int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
    for (int i2 : vIndices2) {
      const int dist = descriptor_distance(F1.desc + (size_t)i1 * 32, F2.desc + (size_t)i2 * 32);
      if (vMatchedDistance[i2] <= dist) continue;
      if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
      else if (dist < bestDist2) { bestDist2 = dist; }
    }
    if (bestDist <= TH_LOW) {
      if (bestDist < (float)bestDist2 * mfNNratio) {
        if (vnMatches21[bestIdx2] >= 0) { vnMatches12[vnMatches21[bestIdx2]] = -1; nmatches--; }
        vnMatches12[i1] = bestIdx2;
        vnMatches21[bestIdx2] = i1;
        vMatchedDistance[bestIdx2] = bestDist;
        nmatches++;
        if (checkOri) rotHist[rot_bin(F1.keysUn[i1].angle, F2.keysUn[bestIdx2].angle)].push_back(i1);
      }
    }
  }
  if (checkOri) {
    int sizes[HISTO_LENGTH], ind1, ind2, ind3;
    for (int i = 0; i < HISTO_LENGTH; i++) sizes[i] = (int)rotHist[i].size();
    three_maxima(sizes, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (int idx1 : rotHist[i])
        if (vnMatches12[idx1] >= 0) { vnMatches12[idx1] = -1; nmatches--; }
    }
  }
  for (int i1 = 0; i1 < F1.N; i1++)
    if (vnMatches12[i1] >= 0) {
      vbPrevMatched[2 * i1] = F2.keysUn[vnMatches12[i1]].x;
      vbPrevMatched[2 * i1 + 1] = F2.keysUn[vnMatches12[i1]].y;
    }
  return nmatches;
}

// Pinhole::epipolarConstrain (src/CameraModels/Pinhole.cpp:143-172); F12 row-major, computed by the caller
static bool epipolar_constrain(const float* F12, const KeyPoint& kp1, const KeyPoint& kp2, float unc) {
  const float a = kp1.x * F12[0] + kp1.y * F12[3] + F12[6];
  const float b = kp1.x * F12[1] + kp1.y * F12[4] + F12[7];
  const float c = kp1.x * F12[2] + kp1.y * F12[5] + F12[8];
  const float num = a * kp2.x + b * kp2.y + c;
  const float den = a * a + b * b;
  if (den == 0) return false;
  const float dsqr = num * num / den;
  return dsqr < 3.84 * unc;
}

// :916-1147, mono/stereo pinhole branch (mpCamera2 == NULL).  KFx.bound_with_obs = GetMapPoint(idx) != NULL.
int search_for_triangulation(const FrameView& KF1, const FrameView& KF2, int n_nodes, const int32_t* off1,
                             const int32_t* idx1v, const int32_t* off2, const int32_t* idx2v, const float* F12,
                             const float* ep, const float* levelSigma2_2, bool bOnlyStereo, bool bCoarse,
                             bool checkOri, int32_t* vMatches12) {
  int nmatches = 0;
  for (int i = 0; i < KF1.N; i++) vMatches12[i] = -1;
  std::vector<int> rotHist[HISTO_LENGTH];
  for (int nd = 0; nd < n_nodes; nd++) {
    for (int a = off1[nd]; a < off1[nd + 1]; a++) {
      const int idx1 = idx1v[a];
      if (KF1.bound_with_obs[idx1]) continue;
      const bool bStereo1 = KF1.uRight && KF1.uRight[idx1] >= 0;
      if (bOnlyStereo && !bStereo1) continue;
      const KeyPoint& kp1 = KF1.keysUn[idx1];
      int bestDist = TH_LOW, bestIdx2 = -1;
      for (int b = off2[nd]; b < off2[nd + 1]; b++) {
        const int idx2 = idx2v[b];
        if (KF2.bound_with_obs[idx2]) continue;  // vbMatched2 is never set in the reference
        const bool bStereo2 = KF2.uRight && KF2.uRight[idx2] >= 0;
        if (bOnlyStereo && !bStereo2) continue;
        const int dist = descriptor_distance(KF1.desc + (size_t)idx1 * 32, KF2.desc + (size_t)idx2 * 32);
        if (dist > TH_LOW || dist > bestDist) continue;
        const KeyPoint& kp2 = KF2.keysUn[idx2];
        if (!bStereo1 && !bStereo2) {
          const float distex = ep[0] - kp2.x, distey = ep[1] - kp2.y;
          if (distex * distex + distey * distey < 100 * KF2.scaleFactors[kp2.octave]) continue;
        }
        if (bCoarse || epipolar_constrain(F12, kp1, kp2, levelSigma2_2[kp2.octave])) {
          bestIdx2 = idx2;
          bestDist = dist;
        }
      }
      if (bestIdx2 >= 0) {
        vMatches12[idx1] = bestIdx2;
        nmatches++;
        if (checkOri) rotHist[rot_bin(kp1.angle, KF2.keysUn[bestIdx2].angle)].push_back(idx1);
      }
    }
  }
  if (checkOri) {
    int sizes[HISTO_LENGTH], ind1, ind2, ind3;
    for (int i = 0; i < HISTO_LENGTH; i++) sizes[i] = (int)rotHist[i].size();
    three_maxima(sizes, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (int idx1 : rotHist[i]) { vMatches12[idx1] = -1; nmatches--; }
    }
  }
  return nmatches;
}

}  // namespace orb_oracle

using namespace orb_oracle;

extern "C" {

int oo_search_by_projection_reloc(const cuorb_frame_view* cur, int n, const uint8_t* mp_desc, const float* u,
                                  const float* v, const int32_t* level, const float* kf_angle,
                                  const uint8_t* active, float th, int orb_dist, int check_ori,
                                  int32_t* best_idx_out, int32_t* rot_bin_out, int32_t* keep_bins_out) {
  FrameView F = view_of(cur);
  return search_by_projection_reloc(F, n, mp_desc, u, v, level, kf_angle, active, th, orb_dist, check_ori != 0,
                                    best_idx_out, rot_bin_out, keep_bins_out);
}

int oo_search_by_projection_sim3(const cuorb_frame_view* kf, int n, const uint8_t* mp_desc, const float* u,
                                 const float* v, const int32_t* level, const uint8_t* active, int th,
                                 float ratio_hamming, int32_t* best_idx_out) {
  FrameView F = view_of(kf);
  return search_by_projection_sim3(F, n, mp_desc, u, v, level, active, th, ratio_hamming, best_idx_out);
}

int oo_fuse(const cuorb_frame_view* kf, const float* inv_level_sigma2, int n, const uint8_t* mp_desc,
            const float* u, const float* v, const float* ur, const int32_t* level, const uint8_t* active,
            float th, int check_reprojection, int32_t* best_idx_out) {
  FrameView F = view_of(kf);
  return fuse(F, inv_level_sigma2, n, mp_desc, u, v, ur, level, active, th, check_reprojection != 0, best_idx_out);
}

int oo_search_by_sim3(const cuorb_frame_view* kf1, const cuorb_frame_view* kf2, const uint8_t* mp_desc1,
                      const float* u12, const float* v12, const int32_t* level12, const uint8_t* active1,
                      const uint8_t* mp_desc2, const float* u21, const float* v21, const int32_t* level21,
                      const uint8_t* active2, float th, int32_t* match12_out) {
  FrameView A = view_of(kf1), B = view_of(kf2);
  return search_by_sim3(A, B, mp_desc1, u12, v12, level12, active1, mp_desc2, u21, v21, level21, active2, th,
                        match12_out);
}

int oo_search_for_initialization(const cuorb_frame_view* f1, const cuorb_frame_view* f2, float* prev_matched,
                                 int window_size, float nnratio, int check_ori, int32_t* matches12_out) {
  FrameView A = view_of(f1), B = view_of(f2);
  return search_for_initialization(A, B, prev_matched, window_size, nnratio, check_ori != 0, matches12_out);
}

int oo_search_for_triangulation(const cuorb_frame_view* kf1, const cuorb_frame_view* kf2, int n_nodes,
                                const int32_t* off1, const int32_t* idx1, const int32_t* off2, const int32_t* idx2,
                                const float* F12, const float* ep, const float* level_sigma2_2, int only_stereo,
                                int coarse, int check_ori, int32_t* match12_out) {
  FrameView A = view_of(kf1), B = view_of(kf2);
  return search_for_triangulation(A, B, n_nodes, off1, idx1, off2, idx2, F12, ep, level_sigma2_2, only_stereo != 0,
                                  coarse != 0, check_ori != 0, match12_out);
}

}  // extern "C"
