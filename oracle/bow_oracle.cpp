// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  See orb_oracle.hpp for scope and parity status.
//
// Oracle for SURVEY.md §8(f) row N3, the two vocabulary-node searches of ORBmatcher:
//   int SearchByBoW(KeyFrame* pKF, Frame& F, std::vector<MapPoint*>& vpMapPointMatches)   src/ORBmatcher.cc:237-420
//   int SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12)  src/ORBmatcher.cc:784-916
// on flattened inputs (Nleft == -1 / NLeft == -1 branches): the merge-join of the two
// DBoW2::FeatureVectors (std::map<NodeId, std::vector<unsigned>>) is given as the list of common
// nodes with both index lists; "valid1" = side-1 feature has a non-bad MapPoint (:263-271 / :819-825),
// "eligible2" = side-2 feature may be matched at all (always for a Frame; has a non-bad MapPoint
// for a KeyFrame, :836-844).  mode 0 = KeyFrame-Frame (accept bestDist1 <= TH_LOW),
// mode 1 = KeyFrame-KeyFrame (accept bestDist1 < TH_LOW).  match12[i1] = matched side-2 feature or -1.
#include <cmath>
#include <vector>

#include "orb_oracle.hpp"

namespace orb_oracle {

static inline int ham32(const uint8_t* a, const uint8_t* b) {
  int d = 0;
  for (int i = 0; i < 32; i++) d += __builtin_popcount((unsigned)(a[i] ^ b[i]));
  return d;
}

int search_by_bow(int n1, const uint8_t* desc1, const uint8_t* valid1, const float* angle1, int n2,
                  const uint8_t* desc2, const uint8_t* eligible2, const float* angle2, int n_nodes,
                  const int32_t* off1, const int32_t* idx1, const int32_t* off2, const int32_t* idx2,
                  int mode, float nnratio, bool check_orientation, int32_t* match12) {
  const int TH_LOW = 50, HISTO_LENGTH = 30;
  const float factor = 1.0f / HISTO_LENGTH;
  std::vector<int> rotHist[HISTO_LENGTH];
  std::vector<char> matched2(n2, 0);
  for (int i = 0; i < n1; i++) match12[i] = -1;
  int nmatches = 0;
  for (int nd = 0; nd < n_nodes; nd++) {
    for (int a = off1[nd]; a < off1[nd + 1]; a++) {
      const int i1 = idx1[a];
      if (!valid1[i1]) continue;
      int bestDist1 = 256, bestIdx2 = -1, bestDist2 = 256;
      for (int b = off2[nd]; b < off2[nd + 1]; b++) {
        const int i2 = idx2[b];
        if (matched2[i2] || !eligible2[i2]) continue;
        const int dist = ham32(desc1 + 32 * (size_t)i1, desc2 + 32 * (size_t)i2);
        if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdx2 = i2; }
        else if (dist < bestDist2) { bestDist2 = dist; }
      }
      const bool close = mode == 0 ? bestDist1 <= TH_LOW : bestDist1 < TH_LOW;
      if (close && static_cast<float>(bestDist1) < nnratio * static_cast<float>(bestDist2)) {
        match12[i1] = bestIdx2;
        matched2[bestIdx2] = 1;
        if (check_orientation) {
          float rot = angle1[i1] - angle2[bestIdx2];
          if (rot < 0.0) rot += 360.0f;
          int bin = (int)std::round(rot * factor);
          if (bin == HISTO_LENGTH) bin = 0;
          rotHist[bin].push_back(i1);
        }
        nmatches++;
      }
    }
  }
  if (check_orientation) {
    int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
    for (int i = 0; i < HISTO_LENGTH; i++) {  // ComputeThreeMaxima (:2038-2071)
      const int s = (int)rotHist[i].size();
      if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
      else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
      else if (s > max3) { max3 = s; ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
    for (int i = 0; i < HISTO_LENGTH; i++) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (int i1 : rotHist[i]) { match12[i1] = -1; nmatches--; }
    }
  }
  return nmatches;
}

// The same search when F is a two-camera rig Frame (F.Nleft != -1, :306-331, :371-396): F's features [0, nleft2) are
// the left camera's, [nleft2, n2) the right camera's, all in one FeatureVector.  Every KeyFrame feature keeps a best /
// second-best pair per camera; the left match needs the ratio test, the right match sits INSIDE the left one's
// `bestDist1 <= TH_LOW` block and has its ratio test disabled (`|| true`, :373).  A KeyFrame feature can therefore
// bind one feature of each camera: matchL[i1] / matchR[i1] (indices into F's combined range, -1 none).
int search_by_bow_rig(int n1, const uint8_t* desc1, const uint8_t* valid1, const float* angle1, int n2, int nleft2,
                      const uint8_t* desc2, const float* angle2, int n_nodes, const int32_t* off1, const int32_t* idx1,
                      const int32_t* off2, const int32_t* idx2, float nnratio, bool check_orientation, int32_t* matchL,
                      int32_t* matchR) {
  const int TH_LOW = 50, HISTO_LENGTH = 30;
  const float factor = 1.0f / HISTO_LENGTH;
  std::vector<int> rotHist[HISTO_LENGTH];          // entries: F feature index, as in the reference
  std::vector<int> owner(n2, -1);                  // vpMapPointMatches[realIdxF] (the KeyFrame feature whose point it holds)
  for (int i = 0; i < n1; i++) matchL[i] = matchR[i] = -1;
  int nmatches = 0;
  auto bin_of = [&](int i1, int i2) {
    float rot = angle1[i1] - angle2[i2];
    if (rot < 0.0) rot += 360.0f;
    int bin = (int)std::round(rot * factor);
    if (bin == HISTO_LENGTH) bin = 0;
    return bin;
  };
  for (int nd = 0; nd < n_nodes; nd++) {
    for (int a = off1[nd]; a < off1[nd + 1]; a++) {
      const int i1 = idx1[a];
      if (!valid1[i1]) continue;
      int bestDist1 = 256, bestIdxF = -1, bestDist2 = 256, bestDist1R = 256, bestIdxFR = -1, bestDist2R = 256;
      for (int b = off2[nd]; b < off2[nd + 1]; b++) {
        const int i2 = idx2[b];
        if (owner[i2] >= 0) continue;
        const int dist = ham32(desc1 + 32 * (size_t)i1, desc2 + 32 * (size_t)i2);
        if (i2 < nleft2 && dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdxF = i2; }
        else if (i2 < nleft2 && dist < bestDist2) { bestDist2 = dist; }
        if (i2 >= nleft2 && dist < bestDist1R) { bestDist2R = bestDist1R; bestDist1R = dist; bestIdxFR = i2; }
        else if (i2 >= nleft2 && dist < bestDist2R) { bestDist2R = dist; }
      }
      if (bestDist1 <= TH_LOW) {
        if (static_cast<float>(bestDist1) < nnratio * static_cast<float>(bestDist2)) {
          owner[bestIdxF] = i1; matchL[i1] = bestIdxF;
          if (check_orientation) rotHist[bin_of(i1, bestIdxF)].push_back(bestIdxF);
          nmatches++;
        }
        if (bestDist1R <= TH_LOW) {
          owner[bestIdxFR] = i1; matchR[i1] = bestIdxFR;
          if (check_orientation) rotHist[bin_of(i1, bestIdxFR)].push_back(bestIdxFR);
          nmatches++;
        }
      }
    }
  }
  if (check_orientation) {
    int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
    for (int i = 0; i < HISTO_LENGTH; i++) {
      const int s = (int)rotHist[i].size();
      if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
      else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
      else if (s > max3) { max3 = s; ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
    for (int i = 0; i < HISTO_LENGTH; i++) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (int i2 : rotHist[i]) {
        const int i1 = owner[i2];
        if (i2 < nleft2) matchL[i1] = -1; else matchR[i1] = -1;
        nmatches--;
      }
    }
  }
  return nmatches;
}

}  // namespace orb_oracle

extern "C" int oo_search_by_bow(int n1, const uint8_t* desc1, const uint8_t* valid1, const float* angle1, int n2,
                                const uint8_t* desc2, const uint8_t* eligible2, const float* angle2, int n_nodes,
                                const int32_t* off1, const int32_t* idx1, const int32_t* off2, const int32_t* idx2,
                                int mode, float nnratio, int check_orientation, int32_t* match12) {
  return orb_oracle::search_by_bow(n1, desc1, valid1, angle1, n2, desc2, eligible2, angle2, n_nodes, off1, idx1, off2,
                                   idx2, mode, nnratio, check_orientation != 0, match12);
}

extern "C" int oo_search_by_bow_rig(int n1, const uint8_t* desc1, const uint8_t* valid1, const float* angle1, int n2, int nleft2,
                                    const uint8_t* desc2, const float* angle2, int n_nodes, const int32_t* off1,
                                    const int32_t* idx1, const int32_t* off2, const int32_t* idx2, float nnratio,
                                    int check_orientation, int32_t* match_l, int32_t* match_r) {
  return orb_oracle::search_by_bow_rig(n1, desc1, valid1, angle1, n2, nleft2, desc2, angle2, n_nodes, off1, idx1, off2, idx2,
                                       nnratio, check_orientation != 0, match_l, match_r);
}
