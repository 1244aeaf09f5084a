// Drop-in for the reference's class ORB_SLAM3::ORBmatcher (pinhole branches)
// (include/ORBmatcher.h:35-158) and for Frame::ComputeStereoMatches (include/Frame.h:155,
// src/Frame.cc:944-1113; BASELINE.json's "SearchStereo").  The methods are templates over the
// Frame / MapPoint types so the very same header compiles against the reference's real classes
// (member names below are the reference's) and against the small structs of tests/native.
// The pointer graph is flattened to SoA on the host, matched by libcuorb, and the bindings
// F.mvpMapPoints[idx] = pMP are replayed in the reference's order.  The two tracking-time SearchByProjection
// overloads and SearchByBoW(KeyFrame, Frame) also take two-camera rig Frames (Nleft != -1); the other searches are
// pinhole (Nleft == -1) only.
#pragma once
#include <cmath>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <utility>
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
    if (rc < 0) CUORB_SHIM_FAIL(std::runtime_error, std::string("ComputeStereoMatches (libcuorb): ") + cuorb_last_error());
  }
  // The stereo Frame constructor's three steps in ONE library call (optional, for latency): what
  //   thread threadLeft(&Frame::ExtractORB, this, 0, imLeft, 0, 0); thread threadRight(&Frame::ExtractORB, this, 1,
  //   imRight, 0, 0); ... join ...; ComputeStereoMatches();            (src/Frame.cc:202-205, 231)
  // leaves in mvKeys, mDescriptors, mvKeysRight, mDescriptorsRight, mvuRight, mvDepth - bit-identical - without
  // sending the keypoints and descriptors back to the device for the association.  Returns the left monoIndex.
  template <class FrameT>
  static int ExtractStereo(FrameT& F, ORBextractor& left, ORBextractor& right, const cv::Mat& imLeft, const cv::Mat& imRight) {
    if (imLeft.empty() || imRight.empty() || imLeft.rows != imRight.rows || imLeft.cols != imRight.cols || imLeft.step != imRight.step)
      CUORB_SHIM_FAIL(std::invalid_argument, "ExtractStereo: two CV_8UC1 images of one size and row step needed", -1);
    const int cap = cuorb_extractor_capacity(left.handle());
    std::vector<cv::KeyPoint> kl(cap), kr(cap);
    cv::Mat dl(cap, 32, CV_8U), dr(cap, 32, CV_8U);
    std::vector<float> u(cap), d(cap);
    int nl = 0, nr = 0;
    const int mono = cuorb_extract_stereo(left.handle(), right.handle(), imLeft.data, imRight.data, imLeft.cols, imLeft.rows,
                                          imLeft.step, F.mb, F.mbf, (cuorb_keypoint*)kl.data(), dl.data, &nl,
                                          (cuorb_keypoint*)kr.data(), dr.data, &nr, cap, u.data(), d.data());
    if (mono < 0) CUORB_SHIM_FAIL(std::runtime_error, std::string("ExtractStereo (libcuorb): ") + cuorb_last_error(), -1);
    kl.resize(nl); kr.resize(nr);
    F.mvKeys.swap(kl); F.mvKeysRight.swap(kr);
    auto rows = [](const cv::Mat& src, int n, cv::Mat& dst) {
      if (n == 0) { dst.release(); return; }
      dst.create(n, 32, CV_8U);
      for (int i = 0; i < n; i++) std::memcpy(dst.ptr<unsigned char>(i), src.ptr<unsigned char>(i), 32);
    };
    rows(dl, nl, F.mDescriptors); rows(dr, nr, F.mDescriptorsRight);
    F.mvuRight.assign(u.begin(), u.begin() + nl);
    F.mvDepth.assign(d.begin(), d.begin() + nl);
    left.InvalidateImagePyramid(); right.InvalidateImagePyramid();
    return mono;
  }
  // the name BASELINE.json uses for the same operation
  template <class FrameT>
  static void SearchStereo(FrameT& F, ORBextractor& left, ORBextractor& right) { ComputeStereoMatches(F, left, right); }

  // int SearchByProjection(Frame&, const std::vector<MapPoint*>&, th, bFarPoints, thFarPoints)
  // (include/ORBmatcher.h:44-50, src/ORBmatcher.cc:38-150)
  template <class FrameT, class MapPointT>
  int SearchByProjection(FrameT& F, const std::vector<MapPointT*>& vpMapPoints, const float th = 3,
                         const bool bFarPoints = false, const float thFarPoints = 50.0f) {
    if (F.Nleft != -1) return SearchByProjectionRig(F, vpMapPoints, th, bFarPoints, thFarPoints);
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
    if (n < 0) CUORB_SHIM_FAIL(std::runtime_error, std::string("SearchByProjection (libcuorb): ") + cuorb_last_error(), 0);
    for (int i = 0; i < M; i++)
      if (best[i] >= 0) F.mvpMapPoints[best[i]] = vpMapPoints[i];  // :138, in MapPoint order
    return n;
  }

  // The same call on a two-camera rig Frame (F.Nleft != -1; src/ORBmatcher.cc:63-150 with the rig conditions and
  // the right part :152-224).  The writes are replayed as the reference makes them, in MapPoint order: the matched
  // slot and the stereo partner's slot on the other camera (:138-142, :212-216).
  template <class FrameT, class MapPointT>
  int SearchByProjectionRig(FrameT& F, const std::vector<MapPointT*>& vpMapPoints, const float th, const bool bFarPoints,
                            const float thFarPoints) {
    const int M = (int)vpMapPoints.size();
    if (M == 0 || F.N == 0) return 0;
    std::vector<uint8_t> desc((size_t)M * 32), flags(M);
    std::vector<float> px(M), py(M), pxr(M), pyr(M), vc(M), vcr(M), dep(M);
    std::vector<int32_t> lvl(M), lvlr(M), bl(M, -1), br(M, -1);
    for (int i = 0; i < M; i++) {
      MapPointT* p = vpMapPoints[i];
      flags[i] = (p->mbTrackInView ? CUORB_MP_IN_VIEW : 0) | (p->mbTrackInViewR ? CUORB_MP_IN_VIEW_R : 0) |
                 (p->isBad() ? CUORB_MP_BAD : 0) | (p->Observations() > 0 ? CUORB_MP_HAS_OBS : 0);
      px[i] = p->mTrackProjX; py[i] = p->mTrackProjY; lvl[i] = p->mnTrackScaleLevel; vc[i] = p->mTrackViewCos;
      pxr[i] = p->mTrackProjXR; pyr[i] = p->mTrackProjYR; lvlr[i] = p->mnTrackScaleLevelR; vcr[i] = p->mTrackViewCosR;
      dep[i] = p->mTrackDepth;
      const cv::Mat d = p->GetDescriptor();
      std::memcpy(&desc[(size_t)i * 32], d.template ptr<unsigned char>(0), 32);
    }
    std::vector<uint8_t> boundL, boundR;
    cuorb_frame_view vl, vr;
    rig_views(F, boundL, boundR, vl, vr);
    const int n = cuorb_search_by_projection_rig(device_, &vl, &vr, F.mvLeftToRightMatch.data(), F.mvRightToLeftMatch.data(), M,
                                                 desc.data(), px.data(), py.data(), lvl.data(), vc.data(), pxr.data(), pyr.data(),
                                                 lvlr.data(), vcr.data(), dep.data(), flags.data(), th, bFarPoints ? 1 : 0,
                                                 thFarPoints, mfNNratio, bl.data(), br.data());
    if (n < 0) CUORB_SHIM_FAIL(std::runtime_error, std::string("SearchByProjection (rig, libcuorb): ") + cuorb_last_error(), 0);
    for (int i = 0; i < M; i++) {
      if (bl[i] >= 0) {
        F.mvpMapPoints[bl[i]] = vpMapPoints[i];                                                   // :136
        if (F.mvLeftToRightMatch[bl[i]] != -1) F.mvpMapPoints[F.mvLeftToRightMatch[bl[i]] + F.Nleft] = vpMapPoints[i];
      }
      if (br[i] >= 0) {
        if (F.mvRightToLeftMatch[br[i]] != -1) F.mvpMapPoints[F.mvRightToLeftMatch[br[i]]] = vpMapPoints[i];   // :212-216
        F.mvpMapPoints[br[i] + F.Nleft] = vpMapPoints[i];
      }
    }
    return n;
  }

  // int SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th, const bool bMono)
  // (include/ORBmatcher.h:54-56, src/ORBmatcher.cc:1684-1908) - Tracking::TrackWithMotionModel.
  // The geometric prologue (:1696-1721) runs here on the caller's own Sophus / Eigen types, in the reference's
  // expression order; the window search, the "already bound" skip, the uRight gate, the rotation histogram and
  // ComputeThreeMaxima run in libcuorb; the bindings are then replayed exactly as the reference makes them:
  // EVERY accepted match binds in last-frame order (:1786 - a later MapPoint may overwrite an earlier one that
  // has no observations), and afterwards the features of every non-kept histogram bin are set to NULL (:1889-1905).
  template <class FrameT>
  int SearchByProjection(FrameT& CurrentFrame, const FrameT& LastFrame, const float th, const bool bMono) {
    const bool rig = CurrentFrame.Nleft != -1;
    const int NL = LastFrame.N;
    if (NL == 0 || CurrentFrame.N == 0) return 0;
    const auto Tcw = CurrentFrame.GetPose();                         // :1696
    const auto twc = Tcw.inverse().translation();                    // :1697
    const auto Tlw = LastFrame.GetPose();                            // :1699
    const auto tlc = Tlw * twc;                                      // :1700
    const bool bForward  = tlc(2) > CurrentFrame.mb && !bMono;       // :1702
    const bool bBackward = -tlc(2) > CurrentFrame.mb && !bMono;      // :1703
    std::vector<uint8_t> desc((size_t)NL * 32, 0), flags(NL, 0);
    std::vector<float> u(NL, 0.f), v(NL, 0.f), invz(NL, 0.f), ang(NL, 0.f), ur, vr;
    std::vector<int32_t> oct(NL, 0), best(NL, -1), bin(NL, -1);
    if (rig) { ur.assign(NL, 0.f); vr.assign(NL, 0.f); }
    for (int i = 0; i < NL; i++) {
      auto* pMP = LastFrame.mvpMapPoints[i];
      if (!pMP || LastFrame.mvbOutlier[i]) continue;                 // :1706-1708
      const auto x3Dw = pMP->GetWorldPos();                          // :1710
      const auto x3Dc = Tcw * x3Dw;                                  // :1711
      const float invzc = 1.0 / x3Dc(2);                             // :1715 (double division, rounded to float)
      if (invzc < 0) continue;                                       // :1717
      const auto uv = CurrentFrame.mpCamera->project(x3Dc);          // :1721; the image-bounds test (:1723-1728) is libcuorb's
      u[i] = uv(0); v[i] = uv(1); invz[i] = invzc;
      const bool lastRight = LastFrame.Nleft != -1 && i >= LastFrame.Nleft;
      oct[i] = lastRight ? LastFrame.mvKeysRight[i - LastFrame.Nleft].octave : LastFrame.mvKeys[i].octave;   // :1730-1732
      ang[i] = LastFrame.Nleft == -1 ? LastFrame.mvKeysUn[i].angle                                            // :1793-1795
               : lastRight ? LastFrame.mvKeysRight[i - LastFrame.Nleft].angle : LastFrame.mvKeys[i].angle;
      if (rig) {                                                     // :1813-1814 (through mpCamera, as the reference does)
        const auto x3Dr = CurrentFrame.GetRelativePoseTrl() * x3Dc;
        const auto uvr = CurrentFrame.mpCamera->project(x3Dr);
        ur[i] = uvr(0); vr[i] = uvr(1);
      }
      flags[i] = (uint8_t)(CUORB_MP_IN_VIEW | (pMP->Observations() > 0 ? CUORB_MP_HAS_OBS : 0));   // :1763-1767
      const cv::Mat d = pMP->GetDescriptor();
      std::memcpy(&desc[(size_t)i * 32], d.template ptr<unsigned char>(0), 32);
    }
    int32_t keep[3] = {-1, -1, -1};
    if (rig) {
      std::vector<uint8_t> boundL, boundR;
      cuorb_frame_view vl, vrv;
      rig_views(CurrentFrame, boundL, boundR, vl, vrv);
      std::vector<int32_t> bestR(NL, -1), binR(NL, -1);
      const int n = cuorb_search_by_projection_frame_rig(device_, &vl, &vrv, NL, desc.data(), u.data(), v.data(), invz.data(),
                                                         ur.data(), vr.data(), oct.data(), ang.data(), flags.data(), th,
                                                         bForward ? 1 : 0, bBackward ? 1 : 0, mbCheckOrientation ? 1 : 0,
                                                         best.data(), bestR.data(), bin.data(), binR.data(), keep);
      if (n < 0) CUORB_SHIM_FAIL(std::runtime_error, std::string("SearchByProjection(Cur, Last) (rig, libcuorb): ") + cuorb_last_error(), 0);
      const int nl = CurrentFrame.Nleft;
      for (int i = 0; i < NL; i++) {                                 // :1786, :1863, in last-frame order
        if (best[i] >= 0) CurrentFrame.mvpMapPoints[best[i]] = LastFrame.mvpMapPoints[i];
        if (bestR[i] >= 0) CurrentFrame.mvpMapPoints[bestR[i] + nl] = LastFrame.mvpMapPoints[i];
      }
      if (mbCheckOrientation)                                        // :1889-1905
        for (int i = 0; i < NL; i++) {
          if (best[i] >= 0 && bin[i] != keep[0] && bin[i] != keep[1] && bin[i] != keep[2]) CurrentFrame.mvpMapPoints[best[i]] = nullptr;
          if (bestR[i] >= 0 && binR[i] != keep[0] && binR[i] != keep[1] && binR[i] != keep[2])
            CurrentFrame.mvpMapPoints[bestR[i] + nl] = nullptr;
        }
      return n;
    }
    std::vector<uint8_t> bound;
    const cuorb_frame_view fv = view(CurrentFrame, bound);
    const int n = cuorb_search_by_projection_frame(device_, &fv, NL, desc.data(), u.data(), v.data(), invz.data(), oct.data(),
                                                   ang.data(), flags.data(), th, bForward ? 1 : 0, bBackward ? 1 : 0,
                                                   mbCheckOrientation ? 1 : 0, best.data(), bin.data(), keep);
    if (n < 0) CUORB_SHIM_FAIL(std::runtime_error, std::string("SearchByProjection(Cur, Last) (libcuorb): ") + cuorb_last_error(), 0);
    for (int i = 0; i < NL; i++)                                     // :1786, in last-frame order
      if (best[i] >= 0) CurrentFrame.mvpMapPoints[best[i]] = LastFrame.mvpMapPoints[i];
    if (mbCheckOrientation)                                          // :1889-1905
      for (int i = 0; i < NL; i++)
        if (best[i] >= 0 && bin[i] != keep[0] && bin[i] != keep[1] && bin[i] != keep[2])
          CurrentFrame.mvpMapPoints[best[i]] = nullptr;
    return n;
  }

  // int SearchByBoW(KeyFrame* pKF, Frame& F, std::vector<MapPoint*>& vpMapPointMatches)
  // (include/ORBmatcher.h:66, src/ORBmatcher.cc:237-420), pinhole and two-camera rig Frames.  Reads pKF->GetMapPointMatches(),
  // pKF->mFeatVec, pKF->mDescriptors, pKF->mvKeysUn, F.mFeatVec, F.mDescriptors, F.mvKeys, F.N.
  template <class KeyFrameT, class FrameT, class MapPointT>
  int SearchByBoW(KeyFrameT* pKF, FrameT& F, std::vector<MapPointT*>& vpMapPointMatches) {
    const std::vector<MapPointT*> vpMapPointsKF = pKF->GetMapPointMatches();
    vpMapPointMatches.assign(F.N, static_cast<MapPointT*>(nullptr));
    const int n1 = (int)vpMapPointsKF.size(), n2 = F.N;
    if (n1 == 0 || n2 == 0) return 0;
    std::vector<uint8_t> valid1(n1), elig2(n2, 1);
    std::vector<float> a1(n1), a2(n2);
    if (F.Nleft != -1) {
      // two-camera rig Frame (:306-331, :371-396): one best / runner-up pair per camera, both may bind
      for (int i = 0; i < n1; i++) {
        valid1[i] = vpMapPointsKF[i] && !vpMapPointsKF[i]->isBad();
        a1[i] = !pKF->mpCamera2 ? pKF->mvKeysUn[i].angle                                       // :340-343
                : i >= pKF->NLeft ? pKF->mvKeysRight[i - pKF->NLeft].angle : pKF->mvKeys[i].angle;
      }
      for (int j = 0; j < n2; j++) a2[j] = j >= F.Nleft ? F.mvKeysRight[j - F.Nleft].angle : F.mvKeys[j].angle;   // :346-348, :384-386
      std::vector<int32_t> off1, idx1, off2, idx2, ml(n1, -1), mr(n1, -1);
      join(pKF->mFeatVec, F.mFeatVec, off1, idx1, off2, idx2);
      const int n = cuorb_search_by_bow_rig(device_, n1, contiguous(pKF->mDescriptors), valid1.data(), a1.data(), n2, F.Nleft,
                                            contiguous(F.mDescriptors), a2.data(), (int)off1.size() - 1, off1.data(), idx1.data(),
                                            off2.data(), idx2.data(), mfNNratio, mbCheckOrientation ? 1 : 0, ml.data(), mr.data());
      if (n < 0) CUORB_SHIM_FAIL(std::runtime_error, std::string("SearchByBoW (rig, libcuorb): ") + cuorb_last_error(), 0);
      for (int i = 0; i < n1; i++) {
        if (ml[i] >= 0) vpMapPointMatches[ml[i]] = vpMapPointsKF[i];     // :335
        if (mr[i] >= 0) vpMapPointMatches[mr[i]] = vpMapPointsKF[i];     // :374
      }
      return n;
    }
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

  // ---- the searches outside the tracking path ----
  // One entry per query MapPoint: what the reference computes with Sophus / Eigen / MapPoint state BEFORE its
  // candidate loop (pose transform, projection, distance range, viewing angle, PredictScale; the lines are cited
  // at each method and in INTEGRATION.md), flattened.  Everything after that runs in libcuorb.
  struct ProjectedPoints {
    std::vector<uint8_t> active;   // every gate before GetFeaturesInArea passed
    std::vector<float>   u, v, ur; // projected pixel; ur = uv(0) - bf * invz (Fuse with reprojection gate)
    std::vector<int32_t> level;    // MapPoint::PredictScale
    std::vector<float>   angle;    // relocalisation: pKF->mvKeysUn[i].angle
    std::vector<uint8_t> desc;     // n x 32: MapPoint::GetDescriptor()
    int size() const { return (int)u.size(); }
    void skip() { push(false, 0.f, 0.f, 0, nullptr); }
    void push(bool ok, float u_, float v_, int level_, const unsigned char* d32, float ur_ = 0.f, float angle_ = 0.f) {
      active.push_back(ok ? 1 : 0); u.push_back(u_); v.push_back(v_); ur.push_back(ur_); level.push_back(level_);
      angle.push_back(angle_);
      desc.resize(desc.size() + 32, 0);
      if (d32) std::memcpy(&desc[desc.size() - 32], d32, 32);
    }
  };

  // int SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const std::set<MapPoint*>& sAlreadyFound, const
  // float th, const int ORBdist)  (include/ORBmatcher.h:60-66, src/ORBmatcher.cc:1910-2036).  P has one entry
  // per pKF->GetMapPointMatches()[i] (:1927-1955 is the caller's prologue); binds CurrentFrame.mvpMapPoints.
  template <class FrameT, class MapPointT>
  int SearchByProjection(FrameT& CurrentFrame, const std::vector<MapPointT*>& vpMPs, const ProjectedPoints& P,
                         const float th, const int ORBdist) {
    if (CurrentFrame.Nleft != -1) CUORB_SHIM_FAIL(std::runtime_error, "ORBmatcher shim: fisheye (Nleft != -1) branch is not built", 0);
    const int n = P.size();
    if (n == 0 || CurrentFrame.N == 0) return 0;
    std::vector<uint8_t> bound(CurrentFrame.N);
    for (int i = 0; i < CurrentFrame.N; i++) bound[i] = CurrentFrame.mvpMapPoints[i] ? 1 : 0;      // :1974
    const cuorb_frame_view fv = view_of(CurrentFrame, bound);
    std::vector<int32_t> best(n), rot(n);
    int32_t keep[3];
    const int r = cuorb_search_by_projection_reloc(device_, &fv, n, P.desc.data(), P.u.data(), P.v.data(), P.level.data(),
                                                   P.angle.data(), P.active.data(), th, ORBdist, mbCheckOrientation ? 1 : 0,
                                                   best.data(), rot.data(), keep);
    if (r < 0) CUORB_SHIM_FAIL(std::runtime_error, std::string("SearchByProjection (libcuorb): ") + cuorb_last_error(), 0);
    for (int i = 0; i < n; i++) {
      if (best[i] < 0) continue;
      const bool kept = !mbCheckOrientation || rot[i] == keep[0] || rot[i] == keep[1] || rot[i] == keep[2];
      CurrentFrame.mvpMapPoints[best[i]] = kept ? vpMPs[i] : static_cast<MapPointT*>(nullptr);     // :1985, :2027
    }
    return r;
  }

  // int SearchByProjection(KeyFrame* pKF, Sophus::Sim3f& Scw, const std::vector<MapPoint*>& vpPoints,
  // std::vector<MapPoint*>& vpMatched, int th, float ratioHamming)  (include/ORBmatcher.h:70-77,
  // src/ORBmatcher.cc:423-538); the overload with vpPointsKFs / vpMatchedKF (:540-664) passes both extra vectors.
  template <class KeyFrameT, class MapPointT>
  int SearchByProjection(KeyFrameT* pKF, const ProjectedPoints& P, const std::vector<MapPointT*>& vpPoints,
                         std::vector<MapPointT*>& vpMatched, int th, float ratioHamming,
                         const std::vector<KeyFrameT*>* vpPointsKFs = nullptr, std::vector<KeyFrameT*>* vpMatchedKF = nullptr) {
    const int n = P.size();
    if (n == 0 || pKF->N == 0) return 0;
    std::vector<uint8_t> bound(pKF->N);
    for (int i = 0; i < pKF->N; i++) bound[i] = vpMatched[i] ? 1 : 0;                              // :510
    const cuorb_frame_view fv = view_of(*pKF, bound);
    std::vector<int32_t> best(n);
    const int r = cuorb_search_by_projection_sim3(device_, &fv, n, P.desc.data(), P.u.data(), P.v.data(), P.level.data(),
                                                  P.active.data(), th, ratioHamming, best.data());
    if (r < 0) CUORB_SHIM_FAIL(std::runtime_error, std::string("SearchByProjection (libcuorb): ") + cuorb_last_error(), 0);
    for (int i = 0; i < n; i++)
      if (best[i] >= 0) {
        vpMatched[best[i]] = vpPoints[i];                                                          // :531
        if (vpPointsKFs && vpMatchedKF) (*vpMatchedKF)[best[i]] = (*vpPointsKFs)[i];               // :656
      }
    return r;
  }

  // int Fuse(KeyFrame* pKF, const std::vector<MapPoint*>& vpMapPoints, const float th, const bool bRight)
  // (include/ORBmatcher.h:129-134, src/ORBmatcher.cc:1149-1328; checkReprojection = true) and
  // int Fuse(KeyFrame* pKF, Sophus::Sim3f& Scw, vpPoints, float th, vpReplacePoint)
  // (include/ORBmatcher.h:137-143, src/ORBmatcher.cc:1330-1452; checkReprojection = false).
  // bestIdx[i] = pKF feature MapPoint i fuses into, or -1; the Replace / AddObservation / AddMapPoint
  // bookkeeping (:1301-1318, :1436-1447) stays in the caller's loop over bestIdx.
  // On a two-camera rig KeyFrame (NLeft != -1) the search runs on ONE camera's features: the left one's (mvKeys) or,
  // with bRight, the right one's (mvKeysRight, descriptor rows idx + NLeft; P projected with GetRightPose /
  // mpCamera2, :1156-1164), and bestIdx is the KeyFrame's combined index as the reference leaves it (:1270-1272).
  template <class KeyFrameT>
  int Fuse(KeyFrameT* pKF, const ProjectedPoints& P, const float th, bool checkReprojection, std::vector<int>& bestIdx,
           const bool bRight = false) {
    const int n = P.size();
    bestIdx.assign(n, -1);
    if (n == 0 || pKF->N == 0) return 0;
    if (pKF->NLeft != -1) {
      const int nl = pKF->NLeft, nr = pKF->N - pKF->NLeft, nn = bRight ? nr : nl;
      if (nn == 0) return 0;
      std::vector<uint8_t> bound(nn, 0);
      cuorb_frame_view fv;
      fv.n = nn;
      fv.keys_un = (const cuorb_keypoint*)(bRight ? pKF->mvKeysRight.data() : pKF->mvKeys.data());
      fv.descriptors = contiguous(pKF->mDescriptors) + (bRight ? (size_t)nl * 32 : 0);
      fv.u_right = nullptr;                                        // a rig KeyFrame's mvuRight is all -1 (src/Frame.cc:1313)
      fv.bound_with_obs = bound.data();
      fv.min_x = pKF->mnMinX; fv.min_y = pKF->mnMinY; fv.max_x = pKF->mnMaxX; fv.max_y = pKF->mnMaxY;
      fv.grid_w_inv = pKF->mfGridElementWidthInv; fv.grid_h_inv = pKF->mfGridElementHeightInv;
      fv.mb = pKF->mb; fv.mbf = pKF->mbf;
      fv.nlevels = (int)pKF->mvScaleFactors.size(); fv.scale_factors = pKF->mvScaleFactors.data();
      const int r = cuorb_fuse(device_, &fv, pKF->mvInvLevelSigma2.data(), n, P.desc.data(), P.u.data(), P.v.data(), P.ur.data(),
                               P.level.data(), P.active.data(), th, checkReprojection ? 1 : 0, bestIdx.data());
      if (r < 0) CUORB_SHIM_FAIL(std::runtime_error, std::string("Fuse (rig, libcuorb): ") + cuorb_last_error(), 0);
      if (bRight) for (int i = 0; i < n; i++) if (bestIdx[i] >= 0) bestIdx[i] += nl;
      return r;
    }
    std::vector<uint8_t> bound(pKF->N, 0);
    const cuorb_frame_view fv = view_of(*pKF, bound);
    const int r = cuorb_fuse(device_, &fv, pKF->mvInvLevelSigma2.data(), n, P.desc.data(), P.u.data(), P.v.data(), P.ur.data(),
                             P.level.data(), P.active.data(), th, checkReprojection ? 1 : 0, bestIdx.data());
    if (r < 0) CUORB_SHIM_FAIL(std::runtime_error, std::string("Fuse (libcuorb): ") + cuorb_last_error(), 0);
    return r;
  }

  // int SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12, const Sophus::Sim3f& S12,
  // const float th)  (include/ORBmatcher.h:120-126, src/ORBmatcher.cc:1454-1682).  P12: one entry per feature of pKF1
  // (its MapPoint projected into pKF2, :1501-1553), P21 the other way (:1586-1634).
  template <class KeyFrameT, class MapPointT>
  int SearchBySim3(KeyFrameT* pKF1, KeyFrameT* pKF2, const ProjectedPoints& P12, const ProjectedPoints& P21,
                   std::vector<MapPointT*>& vpMatches12, const float th) {
    if (P12.size() != pKF1->N || P21.size() != pKF2->N) CUORB_SHIM_FAIL(std::runtime_error, "SearchBySim3 shim: one entry per feature", 0);
    if (pKF1->N == 0 || pKF2->N == 0) return 0;
    std::vector<uint8_t> b1(pKF1->N, 0), b2(pKF2->N, 0);
    const cuorb_frame_view fv1 = view_of(*pKF1, b1), fv2 = view_of(*pKF2, b2);
    std::vector<int32_t> m12(pKF1->N);
    const int r = cuorb_search_by_sim3(device_, &fv1, &fv2, P12.desc.data(), P12.u.data(), P12.v.data(), P12.level.data(),
                                       P12.active.data(), P21.desc.data(), P21.u.data(), P21.v.data(), P21.level.data(),
                                       P21.active.data(), th, m12.data());
    if (r < 0) CUORB_SHIM_FAIL(std::runtime_error, std::string("SearchBySim3 (libcuorb): ") + cuorb_last_error(), 0);
    const std::vector<MapPointT*> vpMapPoints2 = pKF2->GetMapPointMatches();
    for (int i = 0; i < pKF1->N; i++)
      if (m12[i] >= 0) vpMatches12[i] = vpMapPoints2[m12[i]];                                      // :1674
    return r;
  }

  // int SearchForInitialization(Frame& F1, Frame& F2, std::vector<cv::Point2f>& vbPrevMatched, std::vector<int>&
  // vnMatches12, int windowSize = 10)  (include/ORBmatcher.h:99-105, src/ORBmatcher.cc:666-782)
  template <class FrameT>
  int SearchForInitialization(FrameT& F1, FrameT& F2, std::vector<cv::Point2f>& vbPrevMatched, std::vector<int>& vnMatches12,
                              int windowSize = 10) {
    vnMatches12.assign(F1.mvKeysUn.size(), -1);
    if (F1.N == 0 || F2.N == 0) return 0;
    static_assert(sizeof(cv::Point2f) == 8, "cv::Point2f must be two floats");
    std::vector<uint8_t> b1(F1.N, 0), b2(F2.N, 0);
    const cuorb_frame_view fv1 = view_of(F1, b1), fv2 = view_of(F2, b2);
    const int r = cuorb_search_for_initialization(device_, &fv1, &fv2, reinterpret_cast<float*>(vbPrevMatched.data()),
                                                  windowSize, mfNNratio, mbCheckOrientation ? 1 : 0, vnMatches12.data());
    if (r < 0) CUORB_SHIM_FAIL(std::runtime_error, std::string("SearchForInitialization (libcuorb): ") + cuorb_last_error(), 0);
    return r;
  }

  // int SearchForTriangulation(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<std::pair<size_t, size_t>>& vMatchedPairs,
  // const bool bOnlyStereo, const bool bCoarse = false)  (include/ORBmatcher.h:108-114, src/ORBmatcher.cc:916-1147;
  // mpCamera2 == NULL).  F12 (row-major K1^-T [t12]x R12 K2^-1, src/CameraModels/Pinhole.cpp:153-156) and the epipole
  // ep = pKF2->mpCamera->project(T2w * Cw) (:927-933) are the caller's Eigen prologue.
  template <class KeyFrameT>
  int SearchForTriangulation(KeyFrameT* pKF1, KeyFrameT* pKF2, const float F12[9], const float ep[2],
                             std::vector<std::pair<std::size_t, std::size_t>>& vMatchedPairs, const bool bOnlyStereo,
                             const bool bCoarse = false) {
    vMatchedPairs.clear();
    if (pKF1->N == 0 || pKF2->N == 0) return 0;
    std::vector<uint8_t> b1(pKF1->N), b2(pKF2->N);
    for (int i = 0; i < pKF1->N; i++) b1[i] = pKF1->GetMapPoint(i) ? 1 : 0;                        // :984-988
    for (int i = 0; i < pKF2->N; i++) b2[i] = pKF2->GetMapPoint(i) ? 1 : 0;                        // :1012-1017
    const cuorb_frame_view fv1 = view_of(*pKF1, b1), fv2 = view_of(*pKF2, b2);
    std::vector<int32_t> off1, idx1, off2, idx2;
    join(pKF1->mFeatVec, pKF2->mFeatVec, off1, idx1, off2, idx2);
    std::vector<int32_t> m12(pKF1->N);
    const int r = cuorb_search_for_triangulation(device_, &fv1, &fv2, (int)off1.size() - 1, off1.data(), idx1.data(), off2.data(),
                                                 idx2.data(), F12, ep, pKF2->mvLevelSigma2.data(), bOnlyStereo ? 1 : 0,
                                                 bCoarse ? 1 : 0, mbCheckOrientation ? 1 : 0, m12.data());
    if (r < 0) CUORB_SHIM_FAIL(std::runtime_error, std::string("SearchForTriangulation (libcuorb): ") + cuorb_last_error(), 0);
    vMatchedPairs.reserve(r);
    for (int i = 0; i < pKF1->N; i++)                                                              // :1133-1144
      if (m12[i] >= 0) vMatchedPairs.push_back(std::make_pair((std::size_t)i, (std::size_t)m12[i]));
    return r;
  }

  // float RadiusByViewingCos(const float& viewCos)  (src/ORBmatcher.cc:229-235)
  // the float argument is promoted and compared with the DOUBLE literal 0.998: (float)0.998 > 0.998, so
  // viewCos == 0.998f gives 2.5 (src/ORBmatcher.cc:230), as libcuorb's device path does
  static float RadiusByViewingCos(const float& viewCos) { return (double)viewCos > 0.998 ? 2.5f : 4.0f; }

  float mfNNratio;
  bool  mbCheckOrientation;

private:
  int device_;

  // merge-join of two DBoW2::FeatureVectors (std::map<NodeId, std::vector<unsigned>>) + the device search
  template <class FeatVecT>
  int bow(const FeatVecT& fv1, const FeatVecT& fv2, int n1, const uint8_t* d1, const std::vector<uint8_t>& valid1,
          const std::vector<float>& a1, int n2, const uint8_t* d2, const std::vector<uint8_t>& elig2,
          const std::vector<float>& a2, int mode, std::vector<int32_t>& match) {
    std::vector<int32_t> off1, idx1, off2, idx2;
    join(fv1, fv2, off1, idx1, off2, idx2);
    const int n = cuorb_search_by_bow(device_, n1, d1, valid1.data(), a1.data(), n2, d2, elig2.data(), a2.data(),
                                      (int)off1.size() - 1, off1.data(), idx1.data(), off2.data(), idx2.data(), mode, mfNNratio,
                                      mbCheckOrientation ? 1 : 0, match.data());
    if (n < 0) CUORB_SHIM_FAIL(std::runtime_error, std::string("SearchByBoW (libcuorb): ") + cuorb_last_error(), 0);
    return n;
  }

  // the common nodes of two DBoW2::FeatureVectors in NodeId order (src/ORBmatcher.cc:252-259, 395-399)
  template <class FeatVecT>
  static void join(const FeatVecT& fv1, const FeatVecT& fv2, std::vector<int32_t>& off1, std::vector<int32_t>& idx1,
                   std::vector<int32_t>& off2, std::vector<int32_t>& idx2) {
    off1.assign(1, 0); off2.assign(1, 0); idx1.clear(); idx2.clear();
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
  }

  // a Frame or KeyFrame (same member names) with the caller's per-feature "taken" flags
  template <class T>
  static cuorb_frame_view view_of(T& F, const std::vector<uint8_t>& bound) {
    cuorb_frame_view v;
    v.n = F.N;
    v.keys_un = (const cuorb_keypoint*)F.mvKeysUn.data();
    v.descriptors = contiguous(F.mDescriptors);
    v.u_right = F.mvuRight.empty() ? nullptr : F.mvuRight.data();
    v.bound_with_obs = bound.data();
    v.min_x = F.mnMinX; v.min_y = F.mnMinY; v.max_x = F.mnMaxX; v.max_y = F.mnMaxY;
    v.grid_w_inv = F.mfGridElementWidthInv; v.grid_h_inv = F.mfGridElementHeightInv;
    v.mb = F.mb; v.mbf = F.mbf;
    v.nlevels = (int)F.mvScaleFactors.size();
    v.scale_factors = F.mvScaleFactors.data();
    return v;
  }

  static const uint8_t* contiguous(const cv::Mat& m) {
    if (m.rows > 1 && m.step != 32) CUORB_SHIM_FAIL(std::runtime_error, "descriptor matrix must be n x 32 contiguous", nullptr);
    return m.ptr<unsigned char>(0);
  }

public:
  // a two-camera rig Frame as the two views libcuorb takes: left = features [0, Nleft) (mvKeys), right = features
  // [Nleft, N) (mvKeysRight, descriptor rows idx + Nleft)
  template <class FrameT>
  static void rig_views(FrameT& F, std::vector<uint8_t>& boundL, std::vector<uint8_t>& boundR, cuorb_frame_view& vl,
                        cuorb_frame_view& vr) {
    const int nl = F.Nleft, nr = F.N - F.Nleft;
    const unsigned char* d = contiguous(F.mDescriptors);
    boundL.assign(nl > 0 ? nl : 1, 0); boundR.assign(nr > 0 ? nr : 1, 0);
    for (int i = 0; i < F.N; i++) {
      const uint8_t b = (F.mvpMapPoints[i] && F.mvpMapPoints[i]->Observations() > 0) ? 1 : 0;
      if (i < nl) boundL[i] = b; else boundR[i - nl] = b;
    }
    vl.n = nl; vl.keys_un = (const cuorb_keypoint*)F.mvKeys.data(); vl.descriptors = d; vl.u_right = nullptr;
    vl.bound_with_obs = boundL.data();
    vl.min_x = F.mnMinX; vl.min_y = F.mnMinY; vl.max_x = F.mnMaxX; vl.max_y = F.mnMaxY;
    vl.grid_w_inv = F.mfGridElementWidthInv; vl.grid_h_inv = F.mfGridElementHeightInv;
    vl.mb = F.mb; vl.mbf = F.mbf;
    vl.nlevels = (int)F.mvScaleFactors.size(); vl.scale_factors = F.mvScaleFactors.data();
    vr = vl;
    vr.n = nr; vr.keys_un = (const cuorb_keypoint*)F.mvKeysRight.data(); vr.descriptors = d + (size_t)nl * 32;
    vr.bound_with_obs = boundR.data();
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
