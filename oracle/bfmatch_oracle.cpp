// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  See orb_oracle.hpp for scope and parity status.
//
// Oracle for SURVEY.md §8(f) row N4 (dense all-pairs Hamming):
//   Frame::ComputeStereoFishEyeMatches   src/Frame.cc:1304-1355
//       BFmatcher.knnMatch(stereoDescLeft, stereoDescRight, matches, 2)   (cv::BFMatcher(NORM_HAMMING))
//       + Lowe ratio  matches[0].distance < matches[1].distance * 0.7     (:1329)
//   MapPoint::ComputeDistinctiveDescriptors   src/MapPoint.cc:376-450
// cv::BFMatcher::knnMatch (features2d matchers.cpp -> cv::batchDistance, core/batch_distance.cpp)
// keeps, per query, the K smallest distances by scanning the train set in index order and
// inserting with strict comparisons, i.e. the K first entries of the (distance, trainIdx)
// lexicographic order.  Pinned against cv2.BFMatcher 4.13.0 incl. tie-heavy sets
// (tests/golden/bfmatch_cv2.npz).
#include <algorithm>
#include <climits>
#include <cstring>
#include <vector>

#include "orb_oracle.hpp"

namespace orb_oracle {

static inline int hamming32(const uint8_t* a, const uint8_t* b) {
  int d = 0;
  for (int i = 0; i < 32; i++) d += __builtin_popcount((unsigned)(a[i] ^ b[i]));
  return d;
}

// idx/dist: nq x 2; missing neighbours (nt < 2) are idx -1, dist -1
void knn_match2(const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* idx, int32_t* dist) {
  for (int i = 0; i < nq; i++) {
    int d0 = INT_MAX, d1 = INT_MAX, i0 = -1, i1 = -1;
    for (int j = 0; j < nt; j++) {
      const int d = hamming32(q + 32 * (size_t)i, t + 32 * (size_t)j);
      if (d < d1) {
        if (d < d0) { d1 = d0; i1 = i0; d0 = d; i0 = j; }
        else { d1 = d; i1 = j; }
      }
    }
    idx[2 * i] = i0; idx[2 * i + 1] = i1;
    dist[2 * i] = i0 >= 0 ? d0 : -1; dist[2 * i + 1] = i1 >= 0 ? d1 : -1;
  }
}

// src/Frame.cc:1329: accepted train index per query, or -1
void fisheye_ratio_test(const int32_t* idx, const int32_t* dist, int nq, int32_t* match) {
  for (int i = 0; i < nq; i++) {
    const bool two = idx[2 * i] >= 0 && idx[2 * i + 1] >= 0;
    match[i] = (two && (float)dist[2 * i] < (float)dist[2 * i + 1] * 0.7) ? idx[2 * i] : -1;
  }
}

// MapPoint::ComputeDistinctiveDescriptors for one point: n descriptors -> BestIdx (:421-440)
int distinctive_descriptor(const uint8_t* desc, int n) {
  if (n <= 0) return -1;
  std::vector<float> D((size_t)n * n);
  for (int i = 0; i < n; i++) {
    D[(size_t)i * n + i] = 0;
    for (int j = i + 1; j < n; j++) {
      const int d = hamming32(desc + 32 * (size_t)i, desc + 32 * (size_t)j);
      D[(size_t)i * n + j] = (float)d; D[(size_t)j * n + i] = (float)d;
    }
  }
  int BestMedian = INT_MAX, BestIdx = 0;
  for (int i = 0; i < n; i++) {
    std::vector<int> v(D.begin() + (size_t)i * n, D.begin() + (size_t)(i + 1) * n);
    std::sort(v.begin(), v.end());
    const int median = v[(size_t)(0.5 * (n - 1))];
    if (median < BestMedian) { BestMedian = median; BestIdx = i; }
  }
  return BestIdx;
}

}  // namespace orb_oracle

using namespace orb_oracle;

extern "C" {
void oo_knn_match2(const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* idx, int32_t* dist) { knn_match2(q, nq, t, nt, idx, dist); }
void oo_fisheye_ratio_test(const int32_t* idx, const int32_t* dist, int nq, int32_t* match) { fisheye_ratio_test(idx, dist, nq, match); }
void oo_distinctive_descriptors(const uint8_t* desc, const int32_t* off, int n_points, int32_t* best) {
  for (int p = 0; p < n_points; p++) best[p] = distinctive_descriptor(desc + 32 * (size_t)off[p], off[p + 1] - off[p]);
}
}
