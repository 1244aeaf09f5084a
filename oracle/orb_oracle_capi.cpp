// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  See orb_oracle.hpp for scope and parity status.
// Flat C entry points (ctypes) over the oracle: whole-path calls, per-stage taps, the restated
// OpenCV primitives on their own (so they can be pinned against cv2), and a threaded timing
// loop used by bench.py's cpu_baseline / --impl reference legs.
#include <atomic>
#include <chrono>
#include <cstring>
#include <thread>

#include "../include/cuorb.h"  // POD struct layouts only (cuorb_keypoint, cuorb_frame_view)
#include "orb_oracle.hpp"

using namespace orb_oracle;
static_assert(sizeof(KeyPoint) == sizeof(cuorb_keypoint) && sizeof(KeyPoint) == 28, "KeyPoint ABI");

static Img view(const uint8_t* p, int w, int h, size_t stride) {
  Img im;
  im.data = const_cast<uint8_t*>(p); im.cols = w; im.rows = h; im.step = stride;
  return im;
}

extern "C" {

void* oo_create(int nfeatures, float scale, int nlevels, int ini, int min) {
  return new ORBextractor(nfeatures, scale, nlevels, ini, min);
}
void oo_destroy(void* h) { delete (ORBextractor*)h; }

int oo_get_tables(void* h, float* sf, float* inv, float* s2, float* invs2, int32_t* per_level) {
  ORBextractor* e = (ORBextractor*)h;
  for (int i = 0; i < e->nlevels; i++) {
    if (sf) sf[i] = e->mvScaleFactor[i];
    if (inv) inv[i] = e->mvInvScaleFactor[i];
    if (s2) s2[i] = e->mvLevelSigma2[i];
    if (invs2) invs2[i] = e->mvInvLevelSigma2[i];
    if (per_level) per_level[i] = e->mnFeaturesPerLevel[i];
  }
  return e->nlevels;
}
int oo_get_umax(void* h, int32_t* out) {
  ORBextractor* e = (ORBextractor*)h;
  for (size_t i = 0; i < e->umax.size(); i++) out[i] = e->umax[i];
  return (int)e->umax.size();
}

int oo_extract(void* h, const uint8_t* img, int w, int hh, size_t stride, int lap0, int lap1,
               KeyPoint* kps, uint8_t* desc, int cap, int* n_out) {
  ORBextractor* e = (ORBextractor*)h;
  std::vector<KeyPoint> k;
  std::vector<uint8_t>  d;
  const int lap[2] = {lap0, lap1};
  int mono = (*e)(view(img, w, hh, stride), k, d, lap);
  if (mono < 0) return mono;
  *n_out = (int)k.size();
  if ((int)k.size() > cap) return -3;
  if (!k.empty()) {
    memcpy(kps, k.data(), k.size() * sizeof(KeyPoint));
    memcpy(desc, d.data(), d.size());
  }
  return mono;
}

int oo_level_size(void* h, int level, int* w, int* hh) {
  ORBextractor* e = (ORBextractor*)h;
  *w = e->pyr.level[level].cols; *hh = e->pyr.level[level].rows;
  return 0;
}
// plane 0: bordered level, 1: level ROI, 2: blurred level (tight)
int64_t oo_get_plane(void* h, int level, int plane, uint8_t* out, size_t bytes) {
  ORBextractor* e = (ORBextractor*)h;
  const Img& lv = e->pyr.level[level];
  if (plane == 0) {
    const Img& wh = e->pyr.whole[level];
    size_t n = (size_t)wh.cols * wh.rows;
    if (bytes < n) return -3;
    memcpy(out, wh.data, n);
    return (int64_t)n;
  }
  size_t n = (size_t)lv.cols * lv.rows;
  if (bytes < n) return -3;
  if (plane == 1) {
    for (int y = 0; y < lv.rows; y++) memcpy(out + (size_t)y * lv.cols, lv.ptr(y), lv.cols);
  } else {
    if (e->dbg_blurred[level].empty()) return 0;
    memcpy(out, e->dbg_blurred[level].data(), n);
  }
  return (int64_t)n;
}
int oo_get_candidates(void* h, int level, int32_t* xs, int32_t* ys, int32_t* resp, int cap) {
  ORBextractor* e = (ORBextractor*)h;
  const auto& v = e->dbg_candidates[level];
  for (size_t i = 0; i < v.size() && (int)i < cap; i++) {
    xs[i] = (int)v[i].x; ys[i] = (int)v[i].y; resp[i] = (int)v[i].response;
  }
  return (int)v.size();
}
int oo_get_level_keys(void* h, int level, KeyPoint* out, int cap) {
  ORBextractor* e = (ORBextractor*)h;
  const auto& v = e->dbg_level_keys[level];
  for (size_t i = 0; i < v.size() && (int)i < cap; i++) out[i] = v[i];
  return (int)v.size();
}

// ---- primitives on their own ----
void oo_resize(const uint8_t* src, int sw, int sh, size_t sstride, uint8_t* dst, int dw, int dh) {
  Img s = view(src, sw, sh, sstride), d = view(dst, dw, dh, dw);
  resize_linear_8u(s, d);
}
void oo_border(const uint8_t* src, int w, int h, size_t stride, uint8_t* dst, int border) {
  Img s = view(src, w, h, stride), d = view(dst, w + 2 * border, h + 2 * border, w + 2 * border);
  copy_make_border_reflect101(s, d, border);
}
int oo_fast(const uint8_t* img, int w, int h, size_t stride, int thr, int nonmax, KeyPoint* out, int cap) {
  std::vector<KeyPoint> k;
  fast9_16(view(img, w, h, stride), k, thr, nonmax != 0);
  for (size_t i = 0; i < k.size() && (int)i < cap; i++) out[i] = k[i];
  return (int)k.size();
}
void oo_blur(const uint8_t* src, int w, int h, size_t stride, uint8_t* dst) {
  Img s = view(src, w, h, stride), d = view(dst, w, h, w);
  gaussian_blur_7x7_s2(s, d);
}
void oo_fast_atan2(const float* y, const float* x, float* out, int n) {
  for (int i = 0; i < n; i++) out[i] = fast_atan2(y[i], x[i]);
}
int oo_octree(void* h, const KeyPoint* keys, int n, int minX, int maxX, int minY, int maxY, int N,
              KeyPoint* out, int cap) {
  ORBextractor* e = (ORBextractor*)h;
  std::vector<KeyPoint> in(keys, keys + n);
  std::vector<KeyPoint> r = e->DistributeOctTree(in, minX, maxX, minY, maxY, N, 0);
  for (size_t i = 0; i < r.size() && (int)i < cap; i++) out[i] = r[i];
  return (int)r.size();
}
int oo_hamming(const uint8_t* a, const uint8_t* b) { return descriptor_distance(a, b); }

// ---- stereo ----
int oo_stereo(void* hl, void* hr, int nL, const KeyPoint* kl, const uint8_t* dl, int nR,
              const KeyPoint* kr, const uint8_t* dr, float mb, float mbf, float* uRight,
              float* depth, int32_t* bestIdxR, int32_t* bestDist, int32_t* sad) {
  std::vector<KeyPoint> L(kl, kl + nL), R(kr, kr + nR);
  std::vector<float> u, d;
  StereoDebug dbg;
  compute_stereo_matches(*(ORBextractor*)hl, *(ORBextractor*)hr, L, dl, R, dr, mb, mbf, u, d, &dbg);
  for (int i = 0; i < nL; i++) {
    uRight[i] = u[i]; depth[i] = d[i];
    if (bestIdxR) bestIdxR[i] = dbg.best_idx_r[i];
    if (bestDist) bestDist[i] = dbg.best_dist[i];
    if (sad) sad[i] = dbg.sad[i];
  }
  return 0;
}

// ---- projection searches ----
extern "C++" { namespace orb_oracle { FrameView make_frame_view(const cuorb_frame_view* f); } }
static FrameView make_view(const cuorb_frame_view* f) { return orb_oracle::make_frame_view(f); }
extern "C++" FrameView orb_oracle::make_frame_view(const cuorb_frame_view* f) {
  FrameView F;
  F.N = f->n; F.keysUn = (const KeyPoint*)f->keys_un; F.desc = f->descriptors; F.uRight = f->u_right;
  F.bound_with_obs.assign(f->n, 0);
  if (f->bound_with_obs) F.bound_with_obs.assign(f->bound_with_obs, f->bound_with_obs + f->n);
  F.minX = f->min_x; F.minY = f->min_y; F.maxX = f->max_x; F.maxY = f->max_y;
  F.gridWInv = f->grid_w_inv; F.gridHInv = f->grid_h_inv; F.mb = f->mb; F.mbf = f->mbf;
  F.scaleFactors.assign(f->scale_factors, f->scale_factors + f->nlevels);
  F.assign_features_to_grid();
  return F;
}
// Frame::GetFeaturesInArea on its own (src/Frame.cc:788-865), in the reference's visiting order
int oo_features_in_area(const cuorb_frame_view* frame, float x, float y, float r, int min_level,
                        int max_level, int32_t* out, int cap) {
  FrameView F = make_view(frame);
  std::vector<int> v = F.features_in_area(x, y, r, min_level, max_level);
  for (size_t i = 0; i < v.size() && (int)i < cap; i++) out[i] = v[i];
  return (int)v.size();
}
int oo_search_by_projection(const cuorb_frame_view* frame, int M, const uint8_t* mp_desc,
                            const float* px, const float* py, const float* pxr,
                            const int32_t* level, const float* view_cos, const float* depth,
                            const uint8_t* flags, float th, int far_points, float th_far,
                            float nn_ratio, int32_t* best_idx_out) {
  FrameView F = make_view(frame);
  return search_by_projection_mappoints(F, M, mp_desc, px, py, pxr, level, view_cos, depth, flags,
                                        th, far_points != 0, th_far, nn_ratio, best_idx_out);
}
int oo_search_by_projection_frame(const cuorb_frame_view* cur, int n_last, const uint8_t* last_desc,
                                  const float* u, const float* v, const float* inv_z,
                                  const int32_t* last_octave, const float* last_angle,
                                  const uint8_t* flags, float th, int forward, int backward,
                                  int check_ori, int32_t* best_idx_out, int32_t* rot_bin_out,
                                  int32_t* keep_bins_out) {
  FrameView F = make_view(cur);
  return search_by_projection_frame(F, n_last, last_desc, u, v, inv_z, last_octave, last_angle,
                                    flags, th, forward != 0, backward != 0, check_ori != 0,
                                    best_idx_out, rot_bin_out, keep_bins_out);
}

// ---- CPU timing loop: the reference's per-frame stereo path (two extractions +
// ComputeStereoMatches), one std::thread per in-flight frame.  Returns elapsed seconds;
// *total_kps / *total_matches are checksums so the work cannot be optimised away. ----
double oo_bench_stereo(int nfeatures, float scale, int nlevels, int ini, int min,
                       const uint8_t* left, const uint8_t* right, int n_frames, int w, int h,
                       float mb, float mbf, int n_threads, int64_t* total_kps,
                       int64_t* total_matches) {
  std::atomic<int>     next(0);
  std::atomic<int64_t> kps(0), matches(0);
  auto t0 = std::chrono::steady_clock::now();
  std::vector<std::thread> pool;
  for (int t = 0; t < n_threads; t++)
    pool.emplace_back([&]() {
      ORBextractor EL(nfeatures, scale, nlevels, ini, min), ER(nfeatures, scale, nlevels, ini, min);
      const int lap[2] = {0, 0};
      std::vector<KeyPoint> kl, kr;
      std::vector<uint8_t>  dl, dr;
      std::vector<float>    u, d;
      for (;;) {
        int i = next.fetch_add(1);
        if (i >= n_frames) break;
        EL(view(left + (size_t)i * w * h, w, h, w), kl, dl, lap);
        ER(view(right + (size_t)i * w * h, w, h, w), kr, dr, lap);
        if (!kl.empty() && !kr.empty())
          compute_stereo_matches(EL, ER, kl, dl.data(), kr, dr.data(), mb, mbf, u, d, nullptr);
        else u.assign(kl.size(), -1.f);
        int64_t m = 0;
        for (float x : u) m += x >= 0;
        kps += (int64_t)(kl.size() + kr.size());
        matches += m;
      }
    });
  for (auto& th : pool) th.join();
  double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  if (total_kps) *total_kps = kps.load();
  if (total_matches) *total_matches = matches.load();
  return sec;
}

}  // extern "C"
