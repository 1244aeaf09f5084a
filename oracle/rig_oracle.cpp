// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  See orb_oracle.hpp for scope and parity status.
//
// Oracle for the two-camera ("fisheye rig", Frame::Nleft != -1) branches of the two tracking-time searches:
//   int ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th, bFarPoints, thFarPoints)
//        src/ORBmatcher.cc:38-227  (left part :63-150 with Nleft != -1, right part :152-224)
//   int ORBmatcher::SearchByProjection(Frame& Cur, const Frame& Last, th, bMono)
//        src/ORBmatcher.cc:1684-1908 (right part :1812-1886)
// A rig Frame keeps the left camera's features at indices [0, Nleft) (mvKeys, mGrid) and the right camera's at
// [Nleft, N) (mvKeysRight, mGridRight, descriptor rows idx + Nleft); it is passed as two FrameViews.  What differs
// from the pinhole branches, all restated here:
//   - no mvuRight gate on either side; levels come from mvKeys / mvKeysRight;
//   - MapPoints: a left match also binds the right feature mvLeftToRightMatch[best] (and vice versa), each
//     counted as a match; those cross-writes are unconditional, so a feature's state is "what the LAST writer left";
//     a ratio-test failure on the left `continue`s past the right part (:130-132); the right window is
//     RadiusByViewingCos * scale WITHOUT the th factor (:156-163); mnTrackScaleLevelR == -1 skips it (:154);
//   - last frame: the right part sits behind the left part's `continue`s (:1716-1731 and the empty-window one at
//     :1750-1752), has no bounds test of its own, and shares the rotation histogram.
// Parity status: pinned to the reference's own functions (oracle/ref_compat/ref_capi_rig.cpp,
// tests/test_oracle_vs_ref_rig.py), bit for bit.
#include <cmath>
#include <vector>

#include "../include/cuorb.h"
#include "orb_oracle.hpp"

namespace orb_oracle {

static const int TH_HIGH = 100, HISTO_LENGTH = 30;
static float RadiusByViewingCos(const float& viewCos) { return viewCos > 0.998 ? 2.5f : 4.0f; }

// flags: bit0 mbTrackInView, bit1 isBad(), bit2 Observations() > 0, bit3 mbTrackInViewR
int search_by_projection_mappoints_rig(FrameView& L, FrameView& R, const int32_t* l2r, const int32_t* r2l, int M,
                                       const uint8_t* mpDesc, const float* projX, const float* projY,
                                       const int32_t* level, const float* viewCos, const float* projXR,
                                       const float* projYR, const int32_t* levelR, const float* viewCosR,
                                       const float* depth, const uint8_t* flags, float th, bool bFarPoints,
                                       float thFarPoints, float mfNNratio, int32_t* bestL, int32_t* bestR) {
  int nmatches = 0;
  const bool bFactor = th != 1.0;
  for (int iMP = 0; iMP < M; iMP++) {
    bestL[iMP] = bestR[iMP] = -1;
    const bool inL = flags[iMP] & 1, inR = flags[iMP] & 8;
    if (!inL && !inR) continue;
    if (bFarPoints && depth[iMP] > thFarPoints) continue;
    if (flags[iMP] & 2) continue;
    const uint8_t obs = (flags[iMP] & 4) ? 1 : 0;
    const uint8_t* MPdescriptor = mpDesc + (size_t)iMP * 32;
    if (inL) {
      const int nPredictedLevel = level[iMP];
      float r = RadiusByViewingCos(viewCos[iMP]);
      if (bFactor) r *= th;
      const std::vector<int> vIndices =
          L.features_in_area(projX[iMP], projY[iMP], r * L.scaleFactors[nPredictedLevel], nPredictedLevel - 1, nPredictedLevel);
      if (!vIndices.empty()) {
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (int idx : vIndices) {
          if (L.bound_with_obs[idx]) continue;
          const int dist = descriptor_distance(MPdescriptor, L.desc + (size_t)idx * 32);
          if (dist < bestDist) {
            bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = L.keysUn[idx].octave; bestIdx = idx;
          } else if (dist < bestDist2) {
            bestLevel2 = L.keysUn[idx].octave; bestDist2 = dist;
          }
        }
        if (bestDist <= TH_HIGH) {
          if (bestLevel == bestLevel2 && bestDist > mfNNratio * bestDist2) continue;   // skips the right part too
          L.bound_with_obs[bestIdx] = obs;
          bestL[iMP] = bestIdx;
          if (l2r[bestIdx] != -1) { R.bound_with_obs[l2r[bestIdx]] = obs; nmatches++; }
          nmatches++;
        }
      }
    }
    if (inR) {
      const int nPredictedLevel = levelR[iMP];
      if (nPredictedLevel != -1) {
        float r = RadiusByViewingCos(viewCosR[iMP]);
        const std::vector<int> vIndices =
            R.features_in_area(projXR[iMP], projYR[iMP], r * R.scaleFactors[nPredictedLevel], nPredictedLevel - 1, nPredictedLevel);
        if (vIndices.empty()) continue;
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (int idx : vIndices) {
          if (R.bound_with_obs[idx]) continue;
          const int dist = descriptor_distance(MPdescriptor, R.desc + (size_t)idx * 32);
          if (dist < bestDist) {
            bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = R.keysUn[idx].octave; bestIdx = idx;
          } else if (dist < bestDist2) {
            bestLevel2 = R.keysUn[idx].octave; bestDist2 = dist;
          }
        }
        if (bestDist <= TH_HIGH) {
          if (bestLevel == bestLevel2 && bestDist > mfNNratio * bestDist2) continue;
          if (r2l[bestIdx] != -1) { L.bound_with_obs[r2l[bestIdx]] = obs; nmatches++; }
          R.bound_with_obs[bestIdx] = obs;
          bestR[iMP] = bestIdx;
          nmatches++;
        }
      }
    }
  }
  return nmatches;
}

static void three_maxima_rig(const int* histo, int& ind1, int& ind2, int& ind3) {
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

// (u, v) = left projection, (ur, vr) = project(Trl * x3Dc) (:1813-1814), invz = 1 / zc.  lastAngle = kpLF.angle.
int search_by_projection_frame_rig(FrameView& L, FrameView& R, int nLast, const uint8_t* lastDesc, const float* u,
                                   const float* v, const float* invz, const float* ur, const float* vr,
                                   const int32_t* lastOctave, const float* lastAngle, const uint8_t* flags, float th,
                                   bool bForward, bool bBackward, bool checkOri, int32_t* bestL, int32_t* bestR,
                                   int32_t* rotBinL, int32_t* rotBinR, int32_t* keepBinsOut) {
  int nmatches = 0;
  int hist[HISTO_LENGTH] = {0};
  const float factor = 1.0f / HISTO_LENGTH;
  auto bin_of = [&](float a1, float a2) {
    float rot = a1 - a2;
    if (rot < 0.0) rot += 360.0f;
    int bin = (int)std::round(rot * factor);
    if (bin == HISTO_LENGTH) bin = 0;
    return bin;
  };
  auto window = [&](const FrameView& F, float x, float y, float radius, int oct) {
    if (bForward) return F.features_in_area(x, y, radius, oct, -1);
    if (bBackward) return F.features_in_area(x, y, radius, 0, oct);
    return F.features_in_area(x, y, radius, oct - 1, oct + 1);
  };
  for (int i = 0; i < nLast; i++) {
    bestL[i] = bestR[i] = -1; rotBinL[i] = rotBinR[i] = -1;
    if (!(flags[i] & 1)) continue;
    if (invz[i] < 0) continue;
    if (u[i] < L.minX || u[i] > L.maxX) continue;
    if (v[i] < L.minY || v[i] > L.maxY) continue;
    const int nLastOctave = lastOctave[i];
    const float radius = th * L.scaleFactors[nLastOctave];
    const uint8_t obs = (flags[i] & 4) ? 1 : 0;
    const uint8_t* dMP = lastDesc + (size_t)i * 32;
    {
      const std::vector<int> vIndices2 = window(L, u[i], v[i], radius, nLastOctave);
      if (vIndices2.empty()) continue;                               // :1750-1752, skips the right part too
      int bestDist = 256, bestIdx2 = -1;
      for (int i2 : vIndices2) {
        if (L.bound_with_obs[i2]) continue;
        const int dist = descriptor_distance(dMP, L.desc + (size_t)i2 * 32);
        if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
      }
      if (bestDist <= TH_HIGH) {
        L.bound_with_obs[bestIdx2] = obs;
        bestL[i] = bestIdx2;
        nmatches++;
        if (checkOri) { rotBinL[i] = bin_of(lastAngle[i], L.keysUn[bestIdx2].angle); hist[rotBinL[i]]++; }
      }
    }
    {
      const std::vector<int> vIndices2 = window(R, ur[i], vr[i], radius, nLastOctave);
      int bestDist = 256, bestIdx2 = -1;
      for (int i2 : vIndices2) {
        if (R.bound_with_obs[i2]) continue;
        const int dist = descriptor_distance(dMP, R.desc + (size_t)i2 * 32);
        if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
      }
      if (bestDist <= TH_HIGH) {
        R.bound_with_obs[bestIdx2] = obs;
        bestR[i] = bestIdx2;
        nmatches++;
        if (checkOri) { rotBinR[i] = bin_of(lastAngle[i], R.keysUn[bestIdx2].angle); hist[rotBinR[i]]++; }
      }
    }
  }
  keepBinsOut[0] = keepBinsOut[1] = keepBinsOut[2] = -1;
  if (checkOri) {
    int i1, i2, i3;
    three_maxima_rig(hist, i1, i2, i3);
    keepBinsOut[0] = i1; keepBinsOut[1] = i2; keepBinsOut[2] = i3;
    for (int b = 0; b < HISTO_LENGTH; b++)
      if (b != i1 && b != i2 && b != i3) nmatches -= hist[b];
  }
  return nmatches;
}

FrameView make_frame_view(const cuorb_frame_view* f);   // orb_oracle_capi.cpp

}  // namespace orb_oracle

using namespace orb_oracle;

extern "C" {

int oo_search_by_projection_rig(const cuorb_frame_view* left, const cuorb_frame_view* right, const int32_t* l2r,
                                const int32_t* r2l, int M, const uint8_t* mp_desc, const float* px, const float* py,
                                const int32_t* level, const float* view_cos, const float* pxr, const float* pyr,
                                const int32_t* level_r, const float* view_cos_r, const float* depth,
                                const uint8_t* flags, float th, int far_points, float th_far, float nn_ratio,
                                int32_t* best_l, int32_t* best_r) {
  FrameView L = make_frame_view(left), R = make_frame_view(right);
  return search_by_projection_mappoints_rig(L, R, l2r, r2l, M, mp_desc, px, py, level, view_cos, pxr, pyr, level_r,
                                            view_cos_r, depth, flags, th, far_points != 0, th_far, nn_ratio, best_l, best_r);
}

int oo_search_by_projection_frame_rig(const cuorb_frame_view* left, const cuorb_frame_view* right, int n_last,
                                      const uint8_t* last_desc, const float* u, const float* v, const float* inv_z,
                                      const float* ur, const float* vr, const int32_t* last_octave,
                                      const float* last_angle, const uint8_t* flags, float th, int forward,
                                      int backward, int check_ori, int32_t* best_l, int32_t* best_r,
                                      int32_t* rot_bin_l, int32_t* rot_bin_r, int32_t* keep_bins) {
  FrameView L = make_frame_view(left), R = make_frame_view(right);
  return search_by_projection_frame_rig(L, R, n_last, last_desc, u, v, inv_z, ur, vr, last_octave, last_angle, flags,
                                        th, forward != 0, backward != 0, check_ori != 0, best_l, best_r, rot_bin_l,
                                        rot_bin_r, keep_bins);
}

}  // extern "C"
