// Drives shim/ORBmatcher.h's two tracking-time SearchByProjection overloads on TWO-CAMERA RIG Frames (Nleft != -1,
// src/ORBmatcher.cc:38-227 and :1684-1908) over Frame / MapPoint / Sophus / camera classes with the reference's
// member names (the stand-ins of oracle/ref_compat, headers only; nothing of the reference is linked), and writes the
// final F.mvpMapPoints to a file the pytest compares with the reference's own functions.
// usage: rig_check mappoints|lastframe|bow case.bin out.bin
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <vector>

#include "Frame.h"
#include "GeometricCamera.h"
#include "KeyFrame.h"
#include "MapPoint.h"

#include "../../shim/ORBmatcher.h"

namespace ORB_SLAM3 {
float Frame::fx = 0, Frame::fy = 0, Frame::cx = 0, Frame::cy = 0, Frame::invfx = 0, Frame::invfy = 0;
float Frame::mnMinX = 0, Frame::mnMinY = 0, Frame::mnMaxX = 0, Frame::mnMaxY = 0;
float Frame::mfGridElementWidthInv = 0, Frame::mfGridElementHeightInv = 0;
Sophus::SE3f Frame::GetRelativePoseTrl() { return mTrl; }   // src/Frame.cc:628-630
}  // namespace ORB_SLAM3

using namespace ORB_SLAM3;

struct PassThrough : GeometricCamera {
  cv::Point2f     project(const cv::Point3f& p) override { return cv::Point2f(p.x, p.y); }
  Eigen::Vector2d project(const Eigen::Vector3d& v) override { return Eigen::Vector2d(v[0], v[1]); }
  Eigen::Vector2f project(const Eigen::Vector3f& v) override { return Eigen::Vector2f(v[0], v[1]); }
  bool epipolarConstrain(GeometricCamera*, const cv::KeyPoint&, const cv::KeyPoint&, const Eigen::Matrix3f&,
                         const Eigen::Vector3f&, const float, const float) override { return false; }
};

template <class T> static std::vector<T> rd(FILE* f, size_t n) {
  std::vector<T> v(n);
  if (n && fread(v.data(), sizeof(T), n, f) != n) { fprintf(stderr, "short read\n"); exit(2); }
  return v;
}

// SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches) with a rig KeyFrame and a rig Frame (src/ORBmatcher.cc:237-420)
static int bow_main(const char* in, const char* outp) {
  FILE* f = fopen(in, "rb");
  if (!f) return 2;
  // ints: n1l, n1r, n2l, n2r, nn1, m1, nn2, m2, check_ori | floats: ratio
  const std::vector<int32_t> hi = rd<int32_t>(f, 9);
  const std::vector<float> hf = rd<float>(f, 1);
  const int n1l = hi[0], n1r = hi[1], n2l = hi[2], n2r = hi[3], n1 = n1l + n1r, n2 = n2l + n2r;
  PassThrough cam;
  KeyFrame K;
  K.N = n1; K.NLeft = n1l; K.NRight = n1r; K.mpCamera = &cam; K.mpCamera2 = &cam;
  K.mvKeys = rd<cv::KeyPoint>(f, n1l); K.mvKeysRight = rd<cv::KeyPoint>(f, n1r); K.mvKeysUn = K.mvKeys;
  std::vector<uint8_t> d1 = rd<uint8_t>(f, (size_t)n1 * 32);
  K.mDescriptors = cv::Mat(n1, 32, CV_8UC1, d1.data(), 32);
  const std::vector<uint8_t> has = rd<uint8_t>(f, n1);
  Frame F;
  F.N = n2; F.Nleft = n2l; F.Nright = n2r; F.mpCamera = &cam; F.mpCamera2 = &cam;
  F.mvKeys = rd<cv::KeyPoint>(f, n2l); F.mvKeysRight = rd<cv::KeyPoint>(f, n2r);
  std::vector<uint8_t> d2 = rd<uint8_t>(f, (size_t)n2 * 32);
  F.mDescriptors = cv::Mat(n2, 32, CV_8UC1, d2.data(), 32);
  auto featvec = [&](DBoW2::FeatureVector& fv, int nn, int m) {
    const std::vector<int32_t> node = rd<int32_t>(f, nn), off = rd<int32_t>(f, nn + 1), idx = rd<int32_t>(f, m);
    for (int k = 0; k < nn; k++) for (int a = off[k]; a < off[k + 1]; a++) fv.addFeature((DBoW2::NodeId)node[k], (unsigned)idx[a]);
  };
  featvec(K.mFeatVec, hi[4], hi[5]);
  featvec(F.mFeatVec, hi[6], hi[7]);
  fclose(f);
  std::unique_ptr<MapPoint[]> mps(new MapPoint[n1 + 1]);
  K.mvpMapPoints.assign(n1, nullptr);
  for (int i = 0; i < n1; i++) if (has[i]) { mps[i].mbBad = has[i] == 2; K.mvpMapPoints[i] = &mps[i]; }
  std::vector<MapPoint*> matches;
  ORBmatcher matcher(hf[0], hi[8] != 0);
  const int n = matcher.SearchByBoW(&K, F, matches);
  std::vector<int32_t> out(1 + 2 * (size_t)n1, -1);
  out[0] = n;
  for (int i2 = 0; i2 < (int)matches.size(); i2++)
    if (matches[i2]) out[1 + (i2 < n2l ? 0 : n1) + (matches[i2] - &mps[0])] = i2;
  f = fopen(outp, "wb");
  fwrite(out.data(), 4, out.size(), f);
  fclose(f);
  printf("rig bow ok: %d matches\n", n);
  return 0;
}

int main(int argc, char** argv) {
  if (argc != 4) { fprintf(stderr, "usage: rig_check mappoints|lastframe|bow case.bin out.bin\n"); return 2; }
  if (!strcmp(argv[1], "bow")) return bow_main(argv[2], argv[3]);
  const bool last = !strcmp(argv[1], "lastframe");
  FILE* f = fopen(argv[2], "rb");
  if (!f) return 2;
  // header ints: nl, nr, nlevels, M, far/forward, backward, check_ori | floats: th, thfar, ratio, mb, mbf, w, h, trl_x, trl_y
  const std::vector<int32_t> hi = rd<int32_t>(f, 7);
  const std::vector<float> hf = rd<float>(f, 9);
  const int nl = hi[0], nr = hi[1], nlevels = hi[2], M = hi[3], N = nl + nr;
  Frame F;
  PassThrough cam;
  F.N = N; F.Nleft = nl; F.Nright = nr; F.mb = hf[3]; F.mbf = hf[4]; F.mpCamera = &cam; F.mpCamera2 = &cam;
  Frame::mnMinX = 0; Frame::mnMinY = 0; Frame::mnMaxX = hf[5]; Frame::mnMaxY = hf[6];
  Frame::mfGridElementWidthInv = 64.0f / (Frame::mnMaxX - Frame::mnMinX);
  Frame::mfGridElementHeightInv = 48.0f / (Frame::mnMaxY - Frame::mnMinY);
  F.mvKeys = rd<cv::KeyPoint>(f, nl); F.mvKeysRight = rd<cv::KeyPoint>(f, nr); F.mvKeysUn = F.mvKeys;
  std::vector<uint8_t> desc = rd<uint8_t>(f, (size_t)N * 32);
  F.mDescriptors = cv::Mat(N, 32, CV_8UC1, desc.data(), 32);
  const std::vector<uint8_t> bound = rd<uint8_t>(f, N);
  F.mvLeftToRightMatch = rd<int32_t>(f, nl); F.mvRightToLeftMatch = rd<int32_t>(f, nr);
  F.mvScaleFactors = rd<float>(f, nlevels);
  F.mvuRight.assign(nl, -1.f);
  std::unique_ptr<MapPoint[]> pre(new MapPoint[N + 1]);
  F.mvpMapPoints.assign(N, nullptr);
  for (int i = 0; i < N; i++) if (bound[i]) { pre[i].nObs = 1; F.mvpMapPoints[i] = &pre[i]; }
  F.mTrl = Sophus::SE3f(Eigen::Matrix3f::Identity(), Eigen::Vector3f(hf[7], hf[8], 0.f));

  std::unique_ptr<MapPoint[]> mps(new MapPoint[M + 1]);
  std::vector<uint8_t> qdesc = rd<uint8_t>(f, (size_t)M * 32);
  int n = 0;
  if (!last) {
    const std::vector<float> px = rd<float>(f, M), py = rd<float>(f, M), vc = rd<float>(f, M), pxr = rd<float>(f, M),
                             pyr = rd<float>(f, M), vcr = rd<float>(f, M), dep = rd<float>(f, M);
    const std::vector<int32_t> lvl = rd<int32_t>(f, M), lvlr = rd<int32_t>(f, M);
    const std::vector<uint8_t> fl = rd<uint8_t>(f, M);
    std::vector<MapPoint*> ptrs(M);
    for (int i = 0; i < M; i++) {
      MapPoint& p = mps[i];
      p.mbTrackInView = fl[i] & 1; p.mbBad = fl[i] & 2; p.nObs = (fl[i] & 4) ? 1 : 0; p.mbTrackInViewR = fl[i] & 8;
      p.mTrackProjX = px[i]; p.mTrackProjY = py[i]; p.mnTrackScaleLevel = lvl[i]; p.mTrackViewCos = vc[i];
      p.mTrackProjXR = pxr[i]; p.mTrackProjYR = pyr[i]; p.mnTrackScaleLevelR = lvlr[i]; p.mTrackViewCosR = vcr[i];
      p.mTrackDepth = dep[i];
      p.mDescriptor = cv::Mat(1, 32, CV_8UC1, &qdesc[(size_t)i * 32], 32);
      ptrs[i] = &p;
    }
    ORBmatcher matcher(hf[2], true);
    n = matcher.SearchByProjection(F, ptrs, hf[0], hi[4] != 0, hf[1]);
  } else {
    const std::vector<float> u = rd<float>(f, M), v = rd<float>(f, M), z = rd<float>(f, M), ang = rd<float>(f, M);
    const std::vector<int32_t> oct = rd<int32_t>(f, M);
    const std::vector<uint8_t> fl = rd<uint8_t>(f, M);
    Frame L;
    L.N = M; L.Nleft = -1;
    L.mvKeys.resize(M); L.mvpMapPoints.assign(M, nullptr); L.mvbOutlier.assign(M, false);
    for (int i = 0; i < M; i++) {
      L.mvKeys[i].octave = oct[i]; L.mvKeys[i].angle = ang[i];
      if (!(fl[i] & 1)) continue;
      mps[i].mWorldPos = Eigen::Vector3f(u[i], v[i], z[i]);
      mps[i].nObs = (fl[i] & 4) ? 1 : 0;
      mps[i].mDescriptor = cv::Mat(1, 32, CV_8UC1, &qdesc[(size_t)i * 32], 32);
      L.mvpMapPoints[i] = &mps[i];
    }
    L.mvKeysUn = L.mvKeys;
    const float tz = hi[4] ? F.mb + 1.f : hi[5] ? -(F.mb + 1.f) : 0.f;
    L.mTcw = Sophus::SE3f(Eigen::Matrix3f::Identity(), Eigen::Vector3f(0.f, 0.f, tz));
    ORBmatcher matcher(0.9f, hi[6] != 0);
    n = matcher.SearchByProjection(F, L, hf[0], false);
  }
  fclose(f);
  std::vector<int32_t> out(N + 1);
  out[0] = n;
  for (int k = 0; k < N; k++) {
    MapPoint* b = F.mvpMapPoints[k];
    out[k + 1] = !b ? -1 : (b >= &mps[0] && b < &mps[0] + M) ? (int32_t)(b - &mps[0]) : -2;
  }
  f = fopen(argv[3], "wb");
  fwrite(out.data(), 4, out.size(), f);
  fclose(f);
  printf("rig %s ok: %d matches\n", argv[1], n);
  return 0;
}
