// One frame the way a ported Tracking thread would run it with the Frame's features left on the device: libcuorb
// extracts an image (the handle keeps keypoints and descriptors in device memory), then Tracking::SearchLocalPoints'
// two loops run twice through shim/LocalMap.h over the stand-in Frame / MapPoint classes of oracle/ref_compat (headers
// only; nothing of the reference is linked):
//   A  SearchLocalPoints          - Frame flattened on the host from mvKeysUn / mDescriptors (undistorted on the host side)
//   B  SearchLocalPointsResident  - keypoints / descriptors read in place, Frame::UndistortKeyPoints on the device
// and the final F.mvpMapPoints of both go to a file the pytest compares with the reference's own functions.
// usage: localmap_resident_check image.raw w h case.bin out.bin
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
  if (argc != 6) { fprintf(stderr, "usage: localmap_resident_check image.raw w h case.bin out.bin\n"); return 2; }
  const int w = atoi(argv[2]), h = atoi(argv[3]);
  std::vector<uint8_t> img((size_t)w * h);
  FILE* f = fopen(argv[1], "rb");
  if (!f || fread(img.data(), 1, img.size(), f) != img.size()) { fprintf(stderr, "cannot read image\n"); return 2; }
  fclose(f);
  f = fopen(argv[4], "rb");
  if (!f) return 2;
  // ints: N, nlevels, M, n_order, n_skip, nfeatures | floats: th, thfar, ratio, mb, mbf, minx, miny, maxx, maxy, K4[4], D5[5]
  const std::vector<int32_t> hi = rd<int32_t>(f, 6);
  const std::vector<float> hf = rd<float>(f, 18);
  const int N = hi[0], nlevels = hi[1], M = hi[2];
  const float* K4 = &hf[9]; const float* D5 = &hf[13];
  const std::vector<cuorb_frustum_camera> cam = rd<cuorb_frustum_camera>(f, 1);
  const std::vector<uint8_t> fbound = rd<uint8_t>(f, N);
  const std::vector<float> sf = rd<float>(f, nlevels);
  const std::vector<float> P = rd<float>(f, 3 * (size_t)M), Nr = rd<float>(f, 3 * (size_t)M), mn = rd<float>(f, M), mx = rd<float>(f, M);
  std::vector<uint8_t> desc = rd<uint8_t>(f, 32 * (size_t)M);
  const std::vector<uint8_t> bad = rd<uint8_t>(f, M), obs = rd<uint8_t>(f, M);
  const std::vector<int32_t> order = rd<int32_t>(f, hi[3]), skip = rd<int32_t>(f, hi[4]);
  fclose(f);

  // ORBextractor::operator() (the C entry point behind shim/ORBextractor.h)
  cuorb_config cfg;
  cfg.nfeatures = hi[5]; cfg.scale_factor = 1.2f; cfg.nlevels = nlevels; cfg.ini_th_fast = 20; cfg.min_th_fast = 7;
  cfg.max_width = w; cfg.max_height = h; cfg.max_batch = 1; cfg.device = 0;
  cuorb_extractor* ext = cuorb_extractor_create(&cfg);
  if (!ext) { fprintf(stderr, "%s\n", cuorb_last_error()); return 3; }
  const int cap = cuorb_extractor_capacity(ext);
  std::vector<cv::KeyPoint> keys(cap);
  std::vector<uint8_t> fdesc((size_t)cap * 32);
  int n = 0;
  if (cuorb_extract(ext, img.data(), w, h, (size_t)w, 0, 0, (cuorb_keypoint*)keys.data(), fdesc.data(), cap, &n) < 0 || n != N) {
    fprintf(stderr, "extraction: %d keypoints, the case expects %d (%s)\n", n, N, cuorb_last_error());
    return 3;
  }
  keys.resize(N);
  std::vector<cv::KeyPoint> keys_un(N);
  if (cuorb_undistort_keypoints(0, N, (const cuorb_keypoint*)keys.data(), K4, D5, (cuorb_keypoint*)keys_un.data()) < 0) return 3;

  std::unique_ptr<MapPoint[]> mps(new MapPoint[M + 1]);
  for (int i = 0; i < M; i++) {
    MapPoint& p = mps[i];
    p.mWorldPos = Eigen::Vector3f(P[3 * i], P[3 * i + 1], P[3 * i + 2]);
    p.mNormalVector = Eigen::Vector3f(Nr[3 * i], Nr[3 * i + 1], Nr[3 * i + 2]);
    p.mfMinDistance = mn[i]; p.mfMaxDistance = mx[i];
    p.mDescriptor = cv::Mat(1, 32, CV_8UC1, &desc[32 * (size_t)i], 32);
    p.mbBad = bad[i] != 0; p.nObs = obs[i] ? 1 : 0;
    p.mnLastFrameSeen = 0;
  }
  std::unique_ptr<MapPoint[]> pre(new MapPoint[N + 1]);
  const unsigned long frame_id = 77;
  std::vector<MapPoint*> local(order.size());
  for (size_t k = 0; k < order.size(); k++) local[k] = &mps[order[k]];
  for (int k : skip) local[k]->mnLastFrameSeen = frame_id;

  LocalMapMirror<MapPoint> mirror(M + 16);
  if (mirror.Sync(local, [](MapPoint*) { return true; }) != (int)local.size()) return 4;

  std::vector<int32_t> out;
  out.push_back(N);
  for (int pass = 0; pass < 2; pass++) {
    Frame F;
    F.N = N; F.Nleft = -1; F.mb = hf[3]; F.mbf = hf[4];
    Frame::mnMinX = hf[5]; Frame::mnMinY = hf[6]; Frame::mnMaxX = hf[7]; Frame::mnMaxY = hf[8];
    Frame::mfGridElementWidthInv = 64.0f / (Frame::mnMaxX - Frame::mnMinX);
    Frame::mfGridElementHeightInv = 48.0f / (Frame::mnMaxY - Frame::mnMinY);
    F.mvKeys = keys;
    std::vector<uint8_t> zeros((size_t)N * 32, 0);
    if (pass == 0) { F.mvKeysUn = keys_un; F.mDescriptors = cv::Mat(N, 32, CV_8UC1, fdesc.data(), 32); }
    else { F.mvKeysUn.assign(N, cv::KeyPoint()); F.mDescriptors = cv::Mat(N, 32, CV_8UC1, zeros.data(), 32); }   // not read
    F.mvuRight.assign(N, -1.f); F.mvScaleFactors = sf;
    F.mvpMapPoints.assign(N, nullptr);
    for (int i = 0; i < N; i++) if (fbound[i]) { pre[i].nObs = 1; F.mvpMapPoints[i] = &pre[i]; }
    std::vector<uint8_t> bound;
    cuorb_frame_view fv = ORBmatcher::view(F, bound);
    fv.u_right = nullptr;                                    // monocular Frame
    int nToMatch = 0;
    const int nm = pass == 0 ? mirror.SearchLocalPoints(F, cam[0], fv, frame_id, hf[0], false, hf[1], hf[2], &nToMatch)
                             : mirror.SearchLocalPointsResident(F, cam[0], fv, ext, 0, K4, D5, frame_id, hf[0], false, hf[1],
                                                                hf[2], &nToMatch);
    out.push_back(nm); out.push_back(nToMatch);
    for (int k = 0; k < N; k++) {
      MapPoint* b = F.mvpMapPoints[k];
      out.push_back(!b ? -1 : (b >= &mps[0] && b < &mps[0] + M) ? (int32_t)(b - &mps[0]) : -2);
    }
    printf("%s: %d matches, %d in view\n", pass ? "resident" : "host frame", nm, nToMatch);
  }
  f = fopen(argv[5], "wb");
  fwrite(out.data(), 4, out.size(), f);
  fwrite(keys_un.data(), sizeof(cv::KeyPoint), N, f);
  fclose(f);
  cuorb_extractor_destroy(ext);
  return 0;
}
