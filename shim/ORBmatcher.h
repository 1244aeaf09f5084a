// Drop-in for the tracking-time subset of the reference's class ORB_SLAM3::ORBmatcher
// (include/ORBmatcher.h:35-158) and for Frame::ComputeStereoMatches (include/Frame.h:155,
// src/Frame.cc:944-1113; BASELINE.json's "SearchStereo").  The methods are templates over the
// Frame / MapPoint types so the very same header compiles against the reference's real classes
// (member names below are the reference's) and against the small structs of tests/native.
// The pointer graph is flattened to SoA on the host, matched by libcuorb, and the bindings
// F.mvpMapPoints[idx] = pMP are replayed in the reference's order.  Pinhole (Nleft == -1) only.
#pragma once
#include <cmath>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

#include "../include/cuorb.h"
#include "ORBextractor.h"
#include "cv_compat.h"

namespace ORB_SLAM3 {

class ORBmatcher {
public:
  static const int TH_LOW = 50, TH_HIGH = 100, HISTO_LENGTH = 30;  // src/ORBmatcher.cc:30-32

  ORBmatcher(float nnratio = 0.6, bool checkOri = true, int device = 0)
      : mfNNratio(nnratio), mbCheckOrientation(checkOri), device_(device) {}

  // include/ORBmatcher.h:40, src/ORBmatcher.cc:2075-2089
  static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b) {
    return cuorb_hamming256(a.ptr<unsigned char>(0), b.ptr<unsigned char>(0));
  }

  // void Frame::ComputeStereoMatches() (src/Frame.cc:944-1113).  Reads mvKeys, mvKeysRight,
  // mDescriptors, mDescriptorsRight, mb, mbf and the two extractors' pyramids (still on the
  // device from the ExtractORB calls); writes mvuRight, mvDepth.
  template <class FrameT>
  static void ComputeStereoMatches(FrameT& F, ORBextractor& left, ORBextractor& right) {
    const int N = (int)F.mvKeys.size(), Nr = (int)F.mvKeysRight.size();
    F.mvuRight.assign(N, -1.0f);
    F.mvDepth.assign(N, -1.0f);
    if (N == 0 || Nr == 0) return;
    static_assert(sizeof(cv::KeyPoint) == sizeof(cuorb_keypoint), "cv::KeyPoint must be 7 x 4 bytes");
    const int rc = cuorb_stereo_match(left.handle(), right.handle(), N, (const cuorb_keypoint*)F.mvKeys.data(),
                                      contiguous(F.mDescriptors), Nr, (const cuorb_keypoint*)F.mvKeysRight.data(),
                                      contiguous(F.mDescriptorsRight), F.mb, F.mbf, F.mvuRight.data(), F.mvDepth.data());
    if (rc < 0) throw std::runtime_error(std::string("ComputeStereoMatches (libcuorb): ") + cuorb_last_error());
  }
  // the name BASELINE.json uses for the same operation
  template <class FrameT>
  static void SearchStereo(FrameT& F, ORBextractor& left, ORBextractor& right) { ComputeStereoMatches(F, left, right); }

  // int SearchByProjection(Frame&, const std::vector<MapPoint*>&, th, bFarPoints, thFarPoints)
  // (include/ORBmatcher.h:44-50, src/ORBmatcher.cc:38-150)
  template <class FrameT, class MapPointT>
  int SearchByProjection(FrameT& F, const std::vector<MapPointT*>& vpMapPoints, const float th = 3,
                         const bool bFarPoints = false, const float thFarPoints = 50.0f) {
    if (F.Nleft != -1) throw std::runtime_error("ORBmatcher shim: fisheye (Nleft != -1) branch is not built");
    const int M = (int)vpMapPoints.size();
    if (M == 0 || F.N == 0) return 0;
    std::vector<uint8_t> desc((size_t)M * 32), flags(M);
    std::vector<float> px(M), py(M), pxr(M), vc(M), dep(M);
    std::vector<int32_t> lvl(M), best(M, -1);
    for (int i = 0; i < M; i++) {
      MapPointT* p = vpMapPoints[i];
      flags[i] = (p->mbTrackInView ? CUORB_MP_IN_VIEW : 0) | (p->isBad() ? CUORB_MP_BAD : 0) |
                 (p->Observations() > 0 ? CUORB_MP_HAS_OBS : 0);
      px[i] = p->mTrackProjX; py[i] = p->mTrackProjY; pxr[i] = p->mTrackProjXR;
      lvl[i] = p->mnTrackScaleLevel; vc[i] = p->mTrackViewCos; dep[i] = p->mTrackDepth;
      const cv::Mat d = p->GetDescriptor();
      std::memcpy(&desc[(size_t)i * 32], d.template ptr<unsigned char>(0), 32);
    }
    std::vector<uint8_t> bound;
    const cuorb_frame_view fv = view(F, bound);
    const int n = cuorb_search_by_projection(device_, &fv, M, desc.data(), px.data(), py.data(), pxr.data(), lvl.data(),
                                             vc.data(), dep.data(), flags.data(), th, bFarPoints ? 1 : 0, thFarPoints,
                                             mfNNratio, best.data());
    if (n < 0) throw std::runtime_error(std::string("SearchByProjection (libcuorb): ") + cuorb_last_error());
    for (int i = 0; i < M; i++)
      if (best[i] >= 0) F.mvpMapPoints[best[i]] = vpMapPoints[i];  // :138, in MapPoint order
    return n;
  }

  // int SearchByBoW(KeyFrame* pKF, Frame& F, std::vector<MapPoint*>& vpMapPointMatches)
  // (include/ORBmatcher.h:66, src/ORBmatcher.cc:237-420), mono branch.  Reads pKF->GetMapPointMatches(),
  // pKF->mFeatVec, pKF->mDescriptors, pKF->mvKeysUn, F.mFeatVec, F.mDescriptors, F.mvKeys, F.N.
  template <class KeyFrameT, class FrameT, class MapPointT>
  int SearchByBoW(KeyFrameT* pKF, FrameT& F, std::vector<MapPointT*>& vpMapPointMatches) {
    if (F.Nleft != -1) throw std::runtime_error("ORBmatcher shim: fisheye (Nleft != -1) branch is not built");
    const std::vector<MapPointT*> vpMapPointsKF = pKF->GetMapPointMatches();
    vpMapPointMatches.assign(F.N, static_cast<MapPointT*>(nullptr));
    const int n1 = (int)vpMapPointsKF.size(), n2 = F.N;
    if (n1 == 0 || n2 == 0) return 0;
    std::vector<uint8_t> valid1(n1), elig2(n2, 1);
    std::vector<float> a1(n1), a2(n2);
    for (int i = 0; i < n1; i++) { valid1[i] = vpMapPointsKF[i] && !vpMapPointsKF[i]->isBad(); a1[i] = pKF->mvKeysUn[i].angle; }
    for (int j = 0; j < n2; j++) a2[j] = F.mvKeys[j].angle;
    std::vector<int32_t> match(n1, -1);
    const int n = bow(pKF->mFeatVec, F.mFeatVec, n1, contiguous(pKF->mDescriptors), valid1, a1, n2, contiguous(F.mDescriptors),
                      elig2, a2, 0, match);
    for (int i = 0; i < n1; i++)
      if (match[i] >= 0) vpMapPointMatches[match[i]] = vpMapPointsKF[i];   // :304
    return n;
  }

  // int SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12)
  // (include/ORBmatcher.h:67, src/ORBmatcher.cc:784-916), NLeft == -1.
  template <class KeyFrameT, class MapPointT>
  int SearchByBoW(KeyFrameT* pKF1, KeyFrameT* pKF2, std::vector<MapPointT*>& vpMatches12) {
    const std::vector<MapPointT*> mp1 = pKF1->GetMapPointMatches(), mp2 = pKF2->GetMapPointMatches();
    vpMatches12.assign(mp1.size(), static_cast<MapPointT*>(nullptr));
    const int n1 = (int)mp1.size(), n2 = (int)mp2.size();
    if (n1 == 0 || n2 == 0) return 0;
    std::vector<uint8_t> valid1(n1), elig2(n2);
    std::vector<float> a1(n1), a2(n2);
    for (int i = 0; i < n1; i++) { valid1[i] = mp1[i] && !mp1[i]->isBad(); a1[i] = pKF1->mvKeysUn[i].angle; }
    for (int j = 0; j < n2; j++) { elig2[j] = mp2[j] && !mp2[j]->isBad(); a2[j] = pKF2->mvKeysUn[j].angle; }
    std::vector<int32_t> match(n1, -1);
    const int n = bow(pKF1->mFeatVec, pKF2->mFeatVec, n1, contiguous(pKF1->mDescriptors), valid1, a1, n2,
                      contiguous(pKF2->mDescriptors), elig2, a2, 1, match);
    for (int i = 0; i < n1; i++)
      if (match[i] >= 0) vpMatches12[i] = mp2[match[i]];                   // :858
    return n;
  }

  // float RadiusByViewingCos(const float& viewCos)  (src/ORBmatcher.cc:229-235)
  static float RadiusByViewingCos(const float& viewCos) { return viewCos > 0.998f ? 2.5f : 4.0f; }

  float mfNNratio;
  bool  mbCheckOrientation;

private:
  int device_;

  // merge-join of two DBoW2::FeatureVectors (std::map<NodeId, std::vector<unsigned>>) + the device search
  template <class FeatVecT>
  int bow(const FeatVecT& fv1, const FeatVecT& fv2, int n1, const uint8_t* d1, const std::vector<uint8_t>& valid1,
          const std::vector<float>& a1, int n2, const uint8_t* d2, const std::vector<uint8_t>& elig2,
          const std::vector<float>& a2, int mode, std::vector<int32_t>& match) {
    std::vector<int32_t> off1(1, 0), off2(1, 0), idx1, idx2;
    auto it1 = fv1.begin(); auto it2 = fv2.begin();
    while (it1 != fv1.end() && it2 != fv2.end()) {
      if (it1->first == it2->first) {
        for (auto v : it1->second) idx1.push_back((int32_t)v);
        for (auto v : it2->second) idx2.push_back((int32_t)v);
        off1.push_back((int32_t)idx1.size()); off2.push_back((int32_t)idx2.size());
        ++it1; ++it2;
      } else if (it1->first < it2->first) it1 = fv1.lower_bound(it2->first);
      else it2 = fv2.lower_bound(it1->first);
    }
    const int n = cuorb_search_by_bow(device_, n1, d1, valid1.data(), a1.data(), n2, d2, elig2.data(), a2.data(),
                                      (int)off1.size() - 1, off1.data(), idx1.data(), off2.data(), idx2.data(), mode, mfNNratio,
                                      mbCheckOrientation ? 1 : 0, match.data());
    if (n < 0) throw std::runtime_error(std::string("SearchByBoW (libcuorb): ") + cuorb_last_error());
    return n;
  }

  static const uint8_t* contiguous(const cv::Mat& m) {
    if (m.rows > 1 && m.step != 32) throw std::runtime_error("descriptor matrix must be n x 32 contiguous");
    return m.ptr<unsigned char>(0);
  }

  // the Frame members the matchers read (src/ORBmatcher.cc:73-111; src/Frame.cc:788-877)
  template <class FrameT>
  static cuorb_frame_view view(FrameT& F, std::vector<uint8_t>& bound) {
    cuorb_frame_view v;
    v.n = F.N;
    v.keys_un = (const cuorb_keypoint*)F.mvKeysUn.data();
    v.descriptors = contiguous(F.mDescriptors);
    v.u_right = F.mvuRight.empty() ? nullptr : F.mvuRight.data();
    bound.assign(F.N, 0);
    for (int i = 0; i < F.N; i++)
      bound[i] = (F.mvpMapPoints[i] && F.mvpMapPoints[i]->Observations() > 0) ? 1 : 0;  // :96-100
    v.bound_with_obs = bound.data();
    v.min_x = F.mnMinX; v.min_y = F.mnMinY; v.max_x = F.mnMaxX; v.max_y = F.mnMaxY;
    v.grid_w_inv = F.mfGridElementWidthInv; v.grid_h_inv = F.mfGridElementHeightInv;
    v.mb = F.mb; v.mbf = F.mbf;
    v.nlevels = (int)F.mvScaleFactors.size();
    v.scale_factors = F.mvScaleFactors.data();
    return v;
  }
};

}  // namespace ORB_SLAM3
