// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
// Flat C entry points (ctypes) over the reference's OWN ORBextractor, i.e. the unmodified
// /root/reference/src/ORBextractor.cc compiled against the ref_compat cv:: layer.  Same shapes as
// the oo_* calls of oracle/orb_oracle_capi.cpp so tests can run oracle and reference side by side.
#include <cstring>
#include <vector>

#include "ORBextractor.h"  // the reference's header (-I $(REF)/include)

namespace {
// protected members are reachable from a derived class; nothing is overridden
struct RefExtractor : ORB_SLAM3::ORBextractor {
  using ORB_SLAM3::ORBextractor::ORBextractor;
  using ORB_SLAM3::ORBextractor::ComputeKeyPointsOctTree;
  using ORB_SLAM3::ORBextractor::DistributeOctTree;
  using ORB_SLAM3::ORBextractor::mnFeaturesPerLevel;
  using ORB_SLAM3::ORBextractor::umax;
};
}  // namespace

extern "C" {

void* ref_create(int nfeatures, float scale, int nlevels, int ini, int min) {
  return new RefExtractor(nfeatures, scale, nlevels, ini, min);
}
void ref_destroy(void* h) { delete (RefExtractor*)h; }

int ref_get_tables(void* h, float* sf, float* inv, float* s2, float* invs2, int32_t* per_level) {
  RefExtractor* e = (RefExtractor*)h;
  const int n = e->GetLevels();
  const std::vector<float> a = e->GetScaleFactors(), b = e->GetInverseScaleFactors(),
                           c = e->GetScaleSigmaSquares(), d = e->GetInverseScaleSigmaSquares();
  for (int i = 0; i < n; i++) {
    if (sf) sf[i] = a[i];
    if (inv) inv[i] = b[i];
    if (s2) s2[i] = c[i];
    if (invs2) invs2[i] = d[i];
    if (per_level) per_level[i] = e->mnFeaturesPerLevel[i];
  }
  return n;
}
int ref_get_umax(void* h, int32_t* out) {
  RefExtractor* e = (RefExtractor*)h;
  for (size_t i = 0; i < e->umax.size(); i++) out[i] = e->umax[i];
  return (int)e->umax.size();
}

// ORBextractor::operator() (src/ORBextractor.cc:1052-1138), called as Frame::ExtractORB does
// (src/Frame.cc:555-562): image Mat, empty mask, keypoint vector, descriptor Mat, lapping area
int ref_extract(void* h, const uint8_t* img, int w, int hh, size_t stride, int lap0, int lap1,
                cv::KeyPoint* kps, uint8_t* desc, int cap, int* n_out) {
  RefExtractor* e = (RefExtractor*)h;
  cv::Mat image = (img && w > 0 && hh > 0) ? cv::Mat(hh, w, CV_8UC1, (void*)img, stride) : cv::Mat();
  std::vector<cv::KeyPoint> k;
  cv::Mat                   d;
  std::vector<int>          lap = {lap0, lap1};
  int mono = (*e)(image, cv::Mat(), k, d, lap);
  if (mono < 0) return mono;
  *n_out = (int)k.size();
  if ((int)k.size() > cap) return -3;
  if (!k.empty()) {
    std::memcpy((void*)kps, k.data(), k.size() * sizeof(cv::KeyPoint));
    for (int i = 0; i < d.rows; i++) std::memcpy(desc + (size_t)i * 32, d.ptr(i), 32);
  }
  return mono;
}

int ref_level_size(void* h, int level, int* w, int* hh) {
  RefExtractor* e = (RefExtractor*)h;
  *w = e->mvImagePyramid[level].cols; *hh = e->mvImagePyramid[level].rows;
  return 0;
}
// plane 0: bordered level (`temp`, 19 px each side), 1: level ROI (mvImagePyramid[level])
int64_t ref_get_plane(void* h, int level, int plane, uint8_t* out, size_t bytes) {
  RefExtractor* e = (RefExtractor*)h;
  const cv::Mat& lv = e->mvImagePyramid[level];
  const int b = plane == 0 ? 19 : 0;
  const int W = lv.cols + 2 * b, H = lv.rows + 2 * b;
  if (bytes < (size_t)W * H) return -3;
  for (int y = 0; y < H; y++)
    std::memcpy(out + (size_t)y * W, lv.data + (ptrdiff_t)(y - b) * (ptrdiff_t)lv.step - b, W);
  return (int64_t)W * H;
}
// per-level keypoints after the quadtree and orientation (allKeypoints[level] of
// src/ORBextractor.cc:1068-1071), recomputed from the pyramid the last extract left behind
int ref_get_level_keys(void* h, int level, cv::KeyPoint* out, int cap) {
  RefExtractor* e = (RefExtractor*)h;
  std::vector<std::vector<cv::KeyPoint>> all;
  e->ComputeKeyPointsOctTree(all);
  const auto& v = all[level];
  for (size_t i = 0; i < v.size() && (int)i < cap; i++) out[i] = v[i];
  return (int)v.size();
}
// DistributeOctTree on its own (src/ORBextractor.cc:556-753)
int ref_octree(void* h, const cv::KeyPoint* keys, int n, int minX, int maxX, int minY, int maxY,
               int N, cv::KeyPoint* out, int cap) {
  RefExtractor* e = (RefExtractor*)h;
  std::vector<cv::KeyPoint> in(keys, keys + n);
  std::vector<cv::KeyPoint> r = e->DistributeOctTree(in, minX, maxX, minY, maxY, N, 0);
  for (size_t i = 0; i < r.size() && (int)i < cap; i++) out[i] = r[i];
  return (int)r.size();
}

}  // extern "C"
