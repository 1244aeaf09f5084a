// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  See orb_oracle.hpp for scope and parity status.
// Matching half of the oracle: Frame::ComputeStereoMatches, the Frame grid, and the two
// tracking-time ORBmatcher::SearchByProjection overloads (pinhole / Nleft == -1 branches).
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstring>
#include <utility>

#include "orb_oracle.hpp"

namespace orb_oracle {

static const int TH_HIGH = 100, TH_LOW = 50, HISTO_LENGTH = 30;  // src/ORBmatcher.cc:30-32
static const int FRAME_GRID_ROWS = 48, FRAME_GRID_COLS = 64;      // include/Frame.h:36-37

// ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:2075-2089)
int descriptor_distance(const uint8_t* a, const uint8_t* b) {
  int dist = 0;
  for (int i = 0; i < 8; i++) {
    uint32_t pa, pb;
    memcpy(&pa, a + 4 * i, 4);
    memcpy(&pb, b + 4 * i, 4);
    unsigned int v = pa ^ pb;
    v = v - ((v >> 1) & 0x55555555);
    v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
    dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
  }
  return dist;
}

// cv::norm(IL, IR, NORM_L1) on two 11x11 u8 windows (src/Frame.cc:1060)
static float norm_l1_11(const uint8_t* a, size_t sa, const uint8_t* b, size_t sb) {
  int s = 0;
  for (int y = 0; y < 11; y++)
    for (int x = 0; x < 11; x++) s += std::abs((int)a[y * sa + x] - (int)b[y * sb + x]);
  return (float)(double)s;
}

// Frame::ComputeStereoMatches (src/Frame.cc:944-1113)
void compute_stereo_matches(const ORBextractor& EL, const ORBextractor& ER,
                            const std::vector<KeyPoint>& mvKeys, const uint8_t* mDescriptors,
                            const std::vector<KeyPoint>& mvKeysRight,
                            const uint8_t* mDescriptorsRight, float mb, float mbf,
                            std::vector<float>& mvuRight, std::vector<float>& mvDepth,
                            StereoDebug* dbg) {
  const int N = (int)mvKeys.size();
  mvuRight.assign(N, -1.0f);
  mvDepth.assign(N, -1.0f);
  if (dbg) { dbg->best_idx_r.assign(N, -1); dbg->best_dist.assign(N, -1); dbg->sad.assign(N, -1); }
  const std::vector<float>& mvScaleFactors    = EL.mvScaleFactor;
  const std::vector<float>& mvInvScaleFactors = EL.mvInvScaleFactor;

  const int thOrbDist = (TH_HIGH + TH_LOW) / 2;
  const int nRows     = EL.pyr.level[0].rows;

  std::vector<std::vector<size_t>> vRowIndices(nRows);
  const int Nr = (int)mvKeysRight.size();
  for (int iR = 0; iR < Nr; iR++) {
    const KeyPoint& kp   = mvKeysRight[iR];
    const float     kpY  = kp.y;
    const float     r    = 2.0f * mvScaleFactors[kp.octave];
    const int       maxr = (int)std::ceil(kpY + r);
    const int       minr = (int)std::floor(kpY - r);
    for (int yi = minr; yi <= maxr; yi++)
      if (yi >= 0 && yi < nRows) vRowIndices[yi].push_back(iR);  // reference: unchecked (UB outside)
  }

  const float minZ = mb, minD = 0, maxD = mbf / minZ;
  std::vector<std::pair<int, int>> vDistIdx;

  for (int iL = 0; iL < N; iL++) {
    const KeyPoint& kpL    = mvKeys[iL];
    const int       levelL = kpL.octave;
    const float     vL = kpL.y, uL = kpL.x;
    if ((size_t)vL >= (size_t)nRows) continue;  // reference: unchecked
    const std::vector<size_t>& vCandidates = vRowIndices[(size_t)vL];
    if (vCandidates.empty()) continue;
    const float minU = uL - maxD, maxU = uL - minD;
    if (maxU < 0) continue;

    int    bestDist = TH_HIGH;
    size_t bestIdxR = 0;
    const uint8_t* dL = mDescriptors + (size_t)iL * 32;
    for (size_t iC = 0; iC < vCandidates.size(); iC++) {
      const size_t    iR  = vCandidates[iC];
      const KeyPoint& kpR = mvKeysRight[iR];
      if (kpR.octave < levelL - 1 || kpR.octave > levelL + 1) continue;
      const float uR = kpR.x;
      if (uR >= minU && uR <= maxU) {
        const int dist = descriptor_distance(dL, mDescriptorsRight + iR * 32);
        if (dist < bestDist) { bestDist = dist; bestIdxR = iR; }
      }
    }
    if (dbg && bestDist < TH_HIGH) { dbg->best_idx_r[iL] = (int)bestIdxR; dbg->best_dist[iL] = bestDist; }

    if (bestDist < thOrbDist) {
      const float uR0         = mvKeysRight[bestIdxR].x;
      const float scaleFactor = mvInvScaleFactors[kpL.octave];
      const float scaleduL    = std::round(kpL.x * scaleFactor);
      const float scaledvL    = std::round(kpL.y * scaleFactor);
      const float scaleduR0   = std::round(uR0 * scaleFactor);

      const int  w  = 5;
      const Img& PL = EL.pyr.level[kpL.octave];
      const Img& PR = ER.pyr.level[kpL.octave];
      const uint8_t* IL = PL.ptr((int)(scaledvL - w)) + (int)(scaleduL - w);

      int       bestDist2 = INT_MAX;  // the reference shadows bestDist here (:1043)
      int       bestincR  = 0;
      const int L         = 5;
      std::vector<float> vDists(2 * L + 1);

      const float iniu = scaleduR0 + L - w;
      const float endu = scaleduR0 + L + w + 1;
      if (iniu < 0 || endu >= PR.cols) continue;

      for (int incR = -L; incR <= +L; incR++) {
        const uint8_t* IR = PR.ptr((int)(scaledvL - w)) + (int)(scaleduR0 + incR - w);
        float dist = norm_l1_11(IL, PL.step, IR, PR.step);
        if (dist < bestDist2) { bestDist2 = (int)dist; bestincR = incR; }
        vDists[L + incR] = dist;
      }
      if (bestincR == -L || bestincR == L) continue;

      const float dist1 = vDists[L + bestincR - 1];
      const float dist2 = vDists[L + bestincR];
      const float dist3 = vDists[L + bestincR + 1];
      const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2));
      if (deltaR < -1 || deltaR > 1) continue;

      float bestuR    = mvScaleFactors[kpL.octave] * ((float)scaleduR0 + (float)bestincR + deltaR);
      float disparity = (uL - bestuR);
      if (disparity >= minD && disparity < maxD) {
        if (disparity <= 0) {
          disparity = 0.01;
          bestuR    = uL - 0.01;
        }
        mvDepth[iL]  = mbf / disparity;
        mvuRight[iL] = bestuR;
        vDistIdx.push_back(std::pair<int, int>(bestDist2, iL));
        if (dbg) dbg->sad[iL] = bestDist2;
      }
    }
  }

  if (vDistIdx.empty()) return;  // reference: reads vDistIdx[0] of an empty vector (UB)
  std::sort(vDistIdx.begin(), vDistIdx.end());
  const float median = (float)vDistIdx[vDistIdx.size() / 2].first;
  const float thDist = 1.5f * 1.4f * median;
  for (int i = (int)vDistIdx.size() - 1; i >= 0; i--) {
    if (vDistIdx[i].first < thDist) break;
    mvuRight[vDistIdx[i].second] = -1;
    mvDepth[vDistIdx[i].second]  = -1;
  }
}

// Frame::AssignFeaturesToGrid / PosInGrid (src/Frame.cc:524-553, 867-877)
void FrameView::assign_features_to_grid() {
  grid.assign((size_t)FRAME_GRID_COLS * FRAME_GRID_ROWS, {});
  for (int i = 0; i < N; i++) {
    const KeyPoint& kp = keysUn[i];
    int posX = (int)std::round((kp.x - minX) * gridWInv);
    int posY = (int)std::round((kp.y - minY) * gridHInv);
    if (posX < 0 || posX >= FRAME_GRID_COLS || posY < 0 || posY >= FRAME_GRID_ROWS) continue;
    grid[(size_t)posX * FRAME_GRID_ROWS + posY].push_back(i);
  }
}

// Frame::GetFeaturesInArea (src/Frame.cc:788-865), bRight = false, Nleft == -1
std::vector<int> FrameView::features_in_area(float x, float y, float r, int minLevel,
                                             int maxLevel) const {
  std::vector<int> vIndices;
  float factorX = r, factorY = r;
  const int nMinCellX = std::max(0, (int)std::floor((x - minX - factorX) * gridWInv));
  if (nMinCellX >= FRAME_GRID_COLS) return vIndices;
  const int nMaxCellX = std::min(FRAME_GRID_COLS - 1, (int)std::ceil((x - minX + factorX) * gridWInv));
  if (nMaxCellX < 0) return vIndices;
  const int nMinCellY = std::max(0, (int)std::floor((y - minY - factorY) * gridHInv));
  if (nMinCellY >= FRAME_GRID_ROWS) return vIndices;
  const int nMaxCellY = std::min(FRAME_GRID_ROWS - 1, (int)std::ceil((y - minY + factorY) * gridHInv));
  if (nMaxCellY < 0) return vIndices;
  const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
  for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
    for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
      const std::vector<int>& vCell = grid[(size_t)ix * FRAME_GRID_ROWS + iy];
      for (size_t j = 0; j < vCell.size(); j++) {
        const KeyPoint& kpUn = keysUn[vCell[j]];
        if (bCheckLevels) {
          if (kpUn.octave < minLevel) continue;
          if (maxLevel >= 0 && kpUn.octave > maxLevel) continue;
        }
        const float distx = kpUn.x - x, disty = kpUn.y - y;
        if (std::fabs(distx) < factorX && std::fabs(disty) < factorY) vIndices.push_back(vCell[j]);
      }
    }
  return vIndices;
}

// ORBmatcher::RadiusByViewingCos (src/ORBmatcher.cc:229-235)
static float RadiusByViewingCos(const float& viewCos) { return viewCos > 0.998 ? 2.5f : 4.0f; }

// ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th, bFarPoints, thFarPoints)
// (src/ORBmatcher.cc:38-150), Nleft == -1 branch.
int search_by_projection_mappoints(FrameView& F, int M, const uint8_t* mpDesc, const float* projX,
                                   const float* projY, const float* projXR, const int32_t* level,
                                   const float* viewCos, const float* depth, const uint8_t* flags,
                                   float th, bool bFarPoints, float thFarPoints, float mfNNratio,
                                   int32_t* bestIdxOut) {
  int nmatches = 0;
  const bool bFactor = th != 1.0;
  for (int iMP = 0; iMP < M; iMP++) {
    bestIdxOut[iMP] = -1;
    if (!(flags[iMP] & 1)) continue;                           // mbTrackInView
    if (bFarPoints && depth[iMP] > thFarPoints) continue;
    if (flags[iMP] & 2) continue;                              // isBad()
    const int nPredictedLevel = level[iMP];
    float r = RadiusByViewingCos(viewCos[iMP]);
    if (bFactor) r *= th;
    const std::vector<int> vIndices = F.features_in_area(
        projX[iMP], projY[iMP], r * F.scaleFactors[nPredictedLevel], nPredictedLevel - 1,
        nPredictedLevel);
    if (vIndices.empty()) continue;
    const uint8_t* MPdescriptor = mpDesc + (size_t)iMP * 32;
    int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
    for (int idx : vIndices) {
      if (F.bound_with_obs[idx]) continue;
      if (F.uRight && F.uRight[idx] > 0) {
        const float er = std::fabs(projXR[iMP] - F.uRight[idx]);
        if (er > r * F.scaleFactors[nPredictedLevel]) continue;
      }
      const int dist = descriptor_distance(MPdescriptor, F.desc + (size_t)idx * 32);
      if (dist < bestDist) {
        bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel;
        bestLevel = F.keysUn[idx].octave; bestIdx = idx;
      } else if (dist < bestDist2) {
        bestLevel2 = F.keysUn[idx].octave; bestDist2 = dist;
      }
    }
    if (bestDist <= TH_HIGH) {
      if (bestLevel == bestLevel2 && bestDist > mfNNratio * bestDist2) continue;
      F.bound_with_obs[bestIdx] = (flags[iMP] & 4) ? 1 : 0;     // F.mvpMapPoints[bestIdx] = pMP
      bestIdxOut[iMP] = bestIdx;
      nmatches++;
    }
  }
  return nmatches;
}

// ORBmatcher::ComputeThreeMaxima (src/ORBmatcher.cc:2038-2071) on bin sizes
static void ComputeThreeMaxima(const int* histo, const int L, int& ind1, int& ind2, int& ind3) {
  int max1 = 0, max2 = 0, max3 = 0;
  for (int i = 0; i < L; i++) {
    const int s = histo[i];
    if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
    else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
    else if (s > max3) { max3 = s; ind3 = i; }
  }
  if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
  else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
}

// ORBmatcher::SearchByProjection(Frame& Cur, const Frame& Last, th, bMono)
// (src/ORBmatcher.cc:1684-1811, 1889-1908), Nleft == -1 branch; the projection of each last-frame
// MapPoint (u, v, invz) is supplied by the caller (1710-1721).
int search_by_projection_frame(FrameView& Cur, int nLast, const uint8_t* lastDesc, const float* u,
                               const float* v, const float* invz, const int32_t* lastOctave,
                               const float* lastAngle, const uint8_t* flags, float th,
                               bool bForward, bool bBackward, bool mbCheckOrientation,
                               int32_t* bestIdxOut, int32_t* rotBinOut, int32_t* keepBinsOut) {
  int nmatches = 0;
  std::vector<int> rotHist[HISTO_LENGTH];
  const float factor = 1.0f / HISTO_LENGTH;
  for (int i = 0; i < nLast; i++) {
    bestIdxOut[i] = -1;
    rotBinOut[i]  = -1;
    if (!(flags[i] & 1)) continue;  // pMP && !mvbOutlier
    const float invzc = invz[i];
    if (invzc < 0) continue;
    if (u[i] < Cur.minX || u[i] > Cur.maxX) continue;
    if (v[i] < Cur.minY || v[i] > Cur.maxY) continue;
    const int nLastOctave = lastOctave[i];
    float radius = th * Cur.scaleFactors[nLastOctave];
    std::vector<int> vIndices2;
    if (bForward) vIndices2 = Cur.features_in_area(u[i], v[i], radius, nLastOctave, -1);
    else if (bBackward) vIndices2 = Cur.features_in_area(u[i], v[i], radius, 0, nLastOctave);
    else vIndices2 = Cur.features_in_area(u[i], v[i], radius, nLastOctave - 1, nLastOctave + 1);
    if (vIndices2.empty()) continue;
    const uint8_t* dMP = lastDesc + (size_t)i * 32;
    int bestDist = 256, bestIdx2 = -1;
    for (int i2 : vIndices2) {
      if (Cur.bound_with_obs[i2]) continue;
      if (Cur.uRight && Cur.uRight[i2] > 0) {
        const float ur = u[i] - Cur.mbf * invzc;
        const float er = std::fabs(ur - Cur.uRight[i2]);
        if (er > radius) continue;
      }
      const int dist = descriptor_distance(dMP, Cur.desc + (size_t)i2 * 32);
      if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
    }
    if (bestDist <= TH_HIGH) {
      Cur.bound_with_obs[bestIdx2] = (flags[i] & 4) ? 1 : 0;
      bestIdxOut[i] = bestIdx2;
      nmatches++;
      if (mbCheckOrientation) {
        float rot = lastAngle[i] - Cur.keysUn[bestIdx2].angle;
        if (rot < 0.0) rot += 360.0f;
        int bin = (int)std::round(rot * factor);
        if (bin == HISTO_LENGTH) bin = 0;
        rotHist[bin].push_back(bestIdx2);
        rotBinOut[i] = bin;
      }
    }
  }
  keepBinsOut[0] = keepBinsOut[1] = keepBinsOut[2] = -1;
  if (mbCheckOrientation) {
    int sizes[HISTO_LENGTH];
    for (int i = 0; i < HISTO_LENGTH; i++) sizes[i] = (int)rotHist[i].size();
    int ind1 = -1, ind2 = -1, ind3 = -1;
    ComputeThreeMaxima(sizes, HISTO_LENGTH, ind1, ind2, ind3);
    keepBinsOut[0] = ind1; keepBinsOut[1] = ind2; keepBinsOut[2] = ind3;
    for (int i = 0; i < HISTO_LENGTH; i++)
      if (i != ind1 && i != ind2 && i != ind3) nmatches -= (int)rotHist[i].size();
  }
  return nmatches;
}

}  // namespace orb_oracle
