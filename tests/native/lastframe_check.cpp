// Drives shim/ORBmatcher.h's SearchByProjection(Frame& Cur, const Frame& Last, th, bMono) overload — the call
// Tracking::TrackWithMotionModel makes (src/ORBmatcher.cc:1684-1908) — over Frame / MapPoint / Sophus / camera
// classes with the reference's member names (the stand-ins of oracle/ref_compat, headers only; nothing of the
// reference is linked), and writes the final CurrentFrame.mvpMapPoints to a file the pytest compares with the
// reference's own function.  Compiled with -DCUORB_SHIM_USE_OPENCV -I oracle/ref_compat, so the shim is also
// compiled against OpenCV-like InputArray / OutputArray proxies instead of its own cv_compat.h.
// usage: lastframe_check case.bin out.bin
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <vector>

#include "Frame.h"            // oracle/ref_compat stand-ins (same member names as include/Frame.h)
#include "GeometricCamera.h"
#include "MapPoint.h"

#include "../../shim/ORBmatcher.h"

namespace ORB_SLAM3 {
float Frame::fx = 0, Frame::fy = 0, Frame::cx = 0, Frame::cy = 0, Frame::invfx = 0, Frame::invfy = 0;
float Frame::mnMinX = 0, Frame::mnMinY = 0, Frame::mnMaxX = 0, Frame::mnMaxY = 0;
float Frame::mfGridElementWidthInv = 0, Frame::mfGridElementHeightInv = 0;
Sophus::SE3f Frame::GetRelativePoseTrl() { return mTrl; }   // src/Frame.cc:628-630 (only called for rig Frames)
}  // namespace ORB_SLAM3

using namespace ORB_SLAM3;

// Pinhole::project with the reference's expression (src/CameraModels/Pinhole.cpp:70-76)
struct TestPinhole : GeometricCamera {
  TestPinhole(const std::vector<float>& p) : GeometricCamera(p) {}
  cv::Point2f project(const cv::Point3f& p) override {
    return cv::Point2f(mvParameters[0] * p.x / p.z + mvParameters[2], mvParameters[1] * p.y / p.z + mvParameters[3]);
  }
  Eigen::Vector2d project(const Eigen::Vector3d& v) override {
    return Eigen::Vector2d(mvParameters[0] * v[0] / v[2] + mvParameters[2], mvParameters[1] * v[1] / v[2] + mvParameters[3]);
  }
  Eigen::Vector2f project(const Eigen::Vector3f& v3D) override {
    Eigen::Vector2f res;
    res[0] = mvParameters[0] * v3D[0] / v3D[2] + mvParameters[2];
    res[1] = mvParameters[1] * v3D[1] / v3D[2] + mvParameters[3];
    return res;
  }
  bool epipolarConstrain(GeometricCamera*, const cv::KeyPoint&, const cv::KeyPoint&, const Eigen::Matrix3f&,
                         const Eigen::Vector3f&, const float, const float) override { return false; }
};

template <class T> static std::vector<T> rd(FILE* f, size_t n) {
  std::vector<T> v(n);
  if (n && fread(v.data(), sizeof(T), n, f) != n) { fprintf(stderr, "short read\n"); exit(2); }
  return v;
}
static Sophus::SE3f pose(const float* p) {
  Eigen::Matrix3f R;
  for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) R(r, c) = p[3 * r + c];
  return Sophus::SE3f(R, Eigen::Vector3f(p[9], p[10], p[11]));
}

int main(int argc, char** argv) {
  if (argc != 3) { fprintf(stderr, "usage: lastframe_check case.bin out.bin\n"); return 2; }
  FILE* f = fopen(argv[1], "rb");
  if (!f) return 2;
  // header: N, NL, nlevels, mono, check_ori | floats: th, mb, mbf, w, h, K4[4], pose_cur[12], pose_last[12]
  const std::vector<int32_t> hi = rd<int32_t>(f, 5);
  const std::vector<float> hf = rd<float>(f, 5 + 4 + 24);
  const int N = hi[0], NL = hi[1], nlevels = hi[2];
  const bool mono = hi[3] != 0, check = hi[4] != 0;
  const float th = hf[0];
  Frame Cur, Last;
  TestPinhole cam(std::vector<float>(hf.begin() + 5, hf.begin() + 9));
  Cur.N = N; Cur.Nleft = -1; Cur.mb = hf[1]; Cur.mbf = hf[2]; Cur.mpCamera = &cam;
  Frame::mnMinX = 0; Frame::mnMinY = 0; Frame::mnMaxX = hf[3]; Frame::mnMaxY = hf[4];
  Frame::mfGridElementWidthInv = 64.0f / (Frame::mnMaxX - Frame::mnMinX);      // src/Frame.cc:240-243
  Frame::mfGridElementHeightInv = 48.0f / (Frame::mnMaxY - Frame::mnMinY);
  Cur.mTcw = pose(&hf[9]); Last.mTcw = pose(&hf[21]);
  // current frame: keys, descriptors, uRight, bound flags (1 = already bound to a MapPoint with observations)
  Cur.mvKeysUn = rd<cv::KeyPoint>(f, N); Cur.mvKeys = Cur.mvKeysUn;
  std::vector<uint8_t> cdesc = rd<uint8_t>(f, (size_t)N * 32);
  Cur.mDescriptors = cv::Mat(N, 32, CV_8UC1, cdesc.data(), 32);
  Cur.mvuRight = rd<float>(f, N);
  const std::vector<uint8_t> cbound = rd<uint8_t>(f, N);
  Cur.mvScaleFactors = rd<float>(f, nlevels);
  std::unique_ptr<MapPoint[]> pre(new MapPoint[N + 1]);
  Cur.mvpMapPoints.assign(N, nullptr);
  for (int i = 0; i < N; i++) if (cbound[i]) { pre[i].nObs = 1; Cur.mvpMapPoints[i] = &pre[i]; }
  // last frame
  Last.N = NL; Last.Nleft = -1;
  const std::vector<float> wp = rd<float>(f, (size_t)NL * 3);
  const std::vector<int32_t> oct = rd<int32_t>(f, NL);
  const std::vector<float> ang = rd<float>(f, NL);
  const std::vector<uint8_t> has_mp = rd<uint8_t>(f, NL), outlier = rd<uint8_t>(f, NL);
  const std::vector<int32_t> nobs = rd<int32_t>(f, NL);
  std::vector<uint8_t> ldesc = rd<uint8_t>(f, (size_t)NL * 32);
  fclose(f);
  std::unique_ptr<MapPoint[]> mps(new MapPoint[NL + 1]);
  Last.mvKeys.resize(NL); Last.mvpMapPoints.assign(NL, nullptr); Last.mvbOutlier.assign(NL, false);
  for (int i = 0; i < NL; i++) {
    Last.mvKeys[i].octave = oct[i]; Last.mvKeys[i].angle = ang[i];
    Last.mvbOutlier[i] = outlier[i] != 0;
    if (!has_mp[i]) continue;
    mps[i].mWorldPos = Eigen::Vector3f(wp[3 * i], wp[3 * i + 1], wp[3 * i + 2]);
    mps[i].nObs = nobs[i];
    mps[i].mDescriptor = cv::Mat(1, 32, CV_8UC1, &ldesc[(size_t)i * 32], 32);
    Last.mvpMapPoints[i] = &mps[i];
  }
  Last.mvKeysUn = Last.mvKeys;

  ORBmatcher matcher(0.9f, check);
  const int n = matcher.SearchByProjection(Cur, Last, th, mono);

  std::vector<int32_t> out(N + 1);
  out[0] = n;
  for (int k = 0; k < N; k++) {
    MapPoint* b = Cur.mvpMapPoints[k];
    out[k + 1] = !b ? -1 : (b >= &mps[0] && b < &mps[0] + NL) ? (int32_t)(b - &mps[0]) : -2;
  }
  // public static fixed in this round (src/ORBmatcher.cc:229-235 compares against the double literal 0.998)
  const float edge = 0.998f;
  if (ORBmatcher::RadiusByViewingCos(edge) != 2.5f) { fprintf(stderr, "RadiusByViewingCos(0.998f) != 2.5\n"); return 5; }
  f = fopen(argv[2], "wb");
  fwrite(out.data(), 4, out.size(), f);
  fclose(f);
  printf("lastframe ok: %d matches\n", n);
  return 0;
}
