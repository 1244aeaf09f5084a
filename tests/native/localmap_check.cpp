// Drives shim/LocalMap.h (LocalMapMirror: Sync + SearchLocalPoints) over the stand-in Frame / MapPoint classes of
// oracle/ref_compat (headers only; nothing of the reference is linked) for TWO consecutive frames: the second frame's
// mvpLocalMapPoints is another subset in another order and some points have changed in between.  Writes, per frame,
// the return value, nToMatch and the final F.mvpMapPoints (as point ids) to a file the pytest compares with the
// reference's own Frame::isInFrustum + ORBmatcher::SearchByProjection run on the same vectors.
// usage: localmap_check case.bin out.bin
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <vector>

#include "Frame.h"
#include "GeometricCamera.h"
#include "MapPoint.h"

#include "../../shim/LocalMap.h"
#include "../../shim/ORBmatcher.h"

namespace ORB_SLAM3 {
float Frame::fx = 0, Frame::fy = 0, Frame::cx = 0, Frame::cy = 0, Frame::invfx = 0, Frame::invfy = 0;
float Frame::mnMinX = 0, Frame::mnMinY = 0, Frame::mnMaxX = 0, Frame::mnMaxY = 0;
float Frame::mfGridElementWidthInv = 0, Frame::mfGridElementHeightInv = 0;
Sophus::SE3f Frame::GetRelativePoseTrl() { return mTrl; }
}  // namespace ORB_SLAM3
using namespace ORB_SLAM3;

template <class T> static std::vector<T> rd(FILE* f, size_t n) {
  std::vector<T> v(n);
  if (n && fread(v.data(), sizeof(T), n, f) != n) { fprintf(stderr, "short read\n"); exit(2); }
  return v;
}

int main(int argc, char** argv) {
  if (argc != 3) { fprintf(stderr, "usage: localmap_check case.bin out.bin\n"); return 2; }
  FILE* f = fopen(argv[1], "rb");
  if (!f) return 2;
  // ints: N, nlevels, M, n_order1, n_skip1, n_order2, n_changed | floats: th, thfar, ratio, mb, mbf, w, h
  const std::vector<int32_t> hi = rd<int32_t>(f, 7);
  const std::vector<float> hf = rd<float>(f, 7);
  const int N = hi[0], nlevels = hi[1], M = hi[2];
  const std::vector<cuorb_frustum_camera> cams = rd<cuorb_frustum_camera>(f, 2);   // frame 1, frame 2
  const std::vector<cv::KeyPoint> keys = rd<cv::KeyPoint>(f, N);
  std::vector<uint8_t> fdesc = rd<uint8_t>(f, (size_t)N * 32);
  const std::vector<float> uright = rd<float>(f, N);
  const std::vector<uint8_t> fbound = rd<uint8_t>(f, N);
  const std::vector<float> sf = rd<float>(f, nlevels);
  // the map: two versions of every point's data (version 2 is used for the changed points in frame 2)
  std::vector<float> P[2], Nr[2], mn[2], mx[2];
  std::vector<uint8_t> desc[2], bad[2], obs[2];
  for (int v = 0; v < 2; v++) {
    P[v] = rd<float>(f, 3 * (size_t)M); Nr[v] = rd<float>(f, 3 * (size_t)M); mn[v] = rd<float>(f, M); mx[v] = rd<float>(f, M);
    desc[v] = rd<uint8_t>(f, 32 * (size_t)M); bad[v] = rd<uint8_t>(f, M); obs[v] = rd<uint8_t>(f, M);
  }
  const std::vector<int32_t> order1 = rd<int32_t>(f, hi[3]), skip1 = rd<int32_t>(f, hi[4]), order2 = rd<int32_t>(f, hi[5]),
                             changed = rd<int32_t>(f, hi[6]);
  fclose(f);

  std::unique_ptr<MapPoint[]> mps(new MapPoint[M + 1]);
  auto load = [&](int i, int v) {
    MapPoint& p = mps[i];
    p.mWorldPos = Eigen::Vector3f(P[v][3 * i], P[v][3 * i + 1], P[v][3 * i + 2]);
    p.mNormalVector = Eigen::Vector3f(Nr[v][3 * i], Nr[v][3 * i + 1], Nr[v][3 * i + 2]);
    p.mfMinDistance = mn[v][i]; p.mfMaxDistance = mx[v][i];
    p.mDescriptor = cv::Mat(1, 32, CV_8UC1, &desc[v][32 * (size_t)i], 32);
    p.mbBad = bad[v][i] != 0; p.nObs = obs[v][i] ? 1 : 0;
    p.mnLastFrameSeen = 0;
  };
  for (int i = 0; i < M; i++) load(i, 0);
  std::unique_ptr<MapPoint[]> pre(new MapPoint[N + 1]);

  auto make_frame = [&](Frame& F) {
    F.N = N; F.Nleft = -1; F.mb = hf[3]; F.mbf = hf[4];
    Frame::mnMinX = 0; Frame::mnMinY = 0; Frame::mnMaxX = hf[5]; Frame::mnMaxY = hf[6];
    Frame::mfGridElementWidthInv = 64.0f / (Frame::mnMaxX - Frame::mnMinX);
    Frame::mfGridElementHeightInv = 48.0f / (Frame::mnMaxY - Frame::mnMinY);
    F.mvKeysUn = keys; F.mvKeys = keys;
    F.mDescriptors = cv::Mat(N, 32, CV_8UC1, fdesc.data(), 32);
    F.mvuRight = uright; F.mvScaleFactors = sf;
    F.mvpMapPoints.assign(N, nullptr);
    for (int i = 0; i < N; i++) if (fbound[i]) { pre[i].nObs = 1; F.mvpMapPoints[i] = &pre[i]; }
  };

  LocalMapMirror<MapPoint> mirror(M + 16);
  std::vector<int32_t> out;
  for (int frame = 0; frame < 2; frame++) {
    const std::vector<int32_t>& order = frame ? order2 : order1;
    std::vector<char> is_changed(M, 0);
    if (frame) for (int i : changed) { load(i, 1); is_changed[i] = 1; }
    std::vector<MapPoint*> local(order.size());
    for (size_t k = 0; k < order.size(); k++) local[k] = &mps[order[k]];
    const unsigned long frame_id = 100 + frame;
    if (!frame) for (int k : skip1) local[k]->mnLastFrameSeen = frame_id;      // already matched in this frame
    const int sent = mirror.Sync(local, [&](MapPoint* p) { return is_changed[p - &mps[0]] != 0; });
    Frame F;
    make_frame(F);
    std::vector<uint8_t> bound;
    const cuorb_frame_view fv = ORBmatcher::view(F, bound);
    int nToMatch = 0;
    const int n = mirror.SearchLocalPoints(F, cams[frame], fv, frame_id, hf[0], false, hf[1], hf[2], &nToMatch);
    out.push_back(n); out.push_back(nToMatch); out.push_back(sent);
    for (int k = 0; k < N; k++) {
      MapPoint* b = F.mvpMapPoints[k];
      out.push_back(!b ? -1 : (b >= &mps[0] && b < &mps[0] + M) ? (int32_t)(b - &mps[0]) : -2);
    }
    printf("frame %d: %d matches, %d in view, %d points sent\n", frame, n, nToMatch, sent);
  }
  f = fopen(argv[2], "wb");
  fwrite(out.data(), 4, out.size(), f);
  fclose(f);
  return 0;
}
