// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  See orb_oracle.hpp for scope and parity status.
//
// Oracle for SURVEY.md §8(f) row N1: the per-frame image operations the reference runs
// immediately BEFORE ORBextractor::operator():
//   System::TrackStereo / TrackRGBD / TrackMonocular (src/System.cc:280-300, 384-386, 476):
//       needToRectify -> cv::remap(im, out, M1, M2, cv::INTER_LINEAR)   (maps CV_32F from
//                        cv::initUndistortRectifyMap, src/Settings.cc:492-510)
//       needToResize  -> cv::resize(im, out, newImSize)                 (INTER_LINEAR default)
//   Tracking::GrabImageStereo / GrabImageRGBD / GrabImageMonocular (src/Tracking.cc:1339-1357,
//   1449-1462, 1520-1533): cv::cvtColor(RGB|BGR|RGBA|BGRA -> GRAY) when the image has 3/4 channels.
// The order is the reference's: geometric op on the multi-channel image first, gray conversion
// second.  The OpenCV arithmetic restated here (imgproc remap.cpp / resize.cpp / color_rgb) is
// pinned bit-for-bit against cv2 4.13.0 by tests/test_oracle_cv2.py and tests/golden/preproc_*.npz:
//   * remap, CV_32FC1 maps, 8U, INTER_LINEAR, BORDER_CONSTANT(0): sx = cvRound(mapx*32),
//     sy = cvRound(mapy*32); integer part >> 5 (saturated to short), 5-bit fractions; weights
//     (32-fx)(32-fy)*32, fx(32-fy)*32, (32-fx)fy*32, fx*fy*32 (sum 2^15; the one table entry that
//     saturates, fx=fy=0, cannot change an 8-bit result); out = (sum w*tap + 2^14) >> 15;
//     taps outside the source are 0.
//   * resize INTER_LINEAR 8U, any channel count: the 11-bit fixed-point path already restated in
//     orb_oracle.cpp applied per channel; an exact 2x decimation in both axes is rerouted by
//     OpenCV to INTER_AREA = (a+b+c+d+2) >> 2.
//   * cvtColor 8U: gray = (B*3735 + G*19235 + R*9798 + 2^14) >> 15.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include "orb_oracle.hpp"

namespace orb_oracle {

// cv::cvtColor(src, dst, COLOR_{RGB,BGR,RGBA,BGRA}2GRAY); channels 3|4, rgb_order: 1 = R first
void cvt_gray_8u(const uint8_t* src, int w, int h, size_t sstride, int channels, int rgb_order,
                 uint8_t* dst, size_t dstride) {
  const int ir = rgb_order ? 0 : 2, ib = rgb_order ? 2 : 0;
  for (int y = 0; y < h; y++) {
    const uint8_t* s = src + (size_t)y * sstride;
    uint8_t*       d = dst + (size_t)y * dstride;
    for (int x = 0; x < w; x++, s += channels)
      d[x] = (uint8_t)((s[ib] * 3735 + s[1] * 19235 + s[ir] * 9798 + (1 << 14)) >> 15);
  }
}

static inline int sat_s16(int v) { return std::min(std::max(v, -32768), 32767); }

// cv::remap(src, dst, mapx, mapy, INTER_LINEAR) with CV_32FC1 maps (dst_h x dst_w), C channels
void remap_linear_8u(const uint8_t* src, int sw, int sh, size_t sstride, int channels,
                     const float* mapx, const float* mapy, uint8_t* dst, int dw, int dh, size_t dstride) {
  for (int y = 0; y < dh; y++) {
    uint8_t* d = dst + (size_t)y * dstride;
    for (int x = 0; x < dw; x++) {
      const int sx = cv_round(mapx[(size_t)y * dw + x] * 32.0f);
      const int sy = cv_round(mapy[(size_t)y * dw + x] * 32.0f);
      const int ix = sat_s16(sx >> 5), iy = sat_s16(sy >> 5);
      const int fx = sx & 31, fy = sy & 31;
      const int w00 = (32 - fx) * (32 - fy) * 32, w01 = fx * (32 - fy) * 32;
      const int w10 = (32 - fx) * fy * 32, w11 = fx * fy * 32;
      const bool x0 = ix >= 0 && ix < sw, x1 = ix + 1 >= 0 && ix + 1 < sw;
      const bool y0 = iy >= 0 && iy < sh, y1 = iy + 1 >= 0 && iy + 1 < sh;
      for (int c = 0; c < channels; c++) {
        auto tap = [&](bool ok, int yy, int xx) -> int {
          return ok ? src[(size_t)yy * sstride + (size_t)xx * channels + c] : 0;
        };
        const int v = tap(y0 && x0, iy, ix) * w00 + tap(y0 && x1, iy, ix + 1) * w01 +
                      tap(y1 && x0, iy + 1, ix) * w10 + tap(y1 && x1, iy + 1, ix + 1) * w11;
        d[x * channels + c] = (uint8_t)((v + (1 << 14)) >> 15);
      }
    }
  }
}

// cv::resize(src, dst, Size(dw, dh)) (INTER_LINEAR), C interleaved channels
void resize_linear_8u_cn(const uint8_t* src, int sw, int sh, size_t sstride, int channels,
                         uint8_t* dst, int dw, int dh, size_t dstride) {
  if (sw == 2 * dw && sh == 2 * dh) {  // OpenCV reroutes the exact 2x2 case to INTER_AREA
    for (int y = 0; y < dh; y++) {
      const uint8_t* s0 = src + (size_t)(2 * y) * sstride;
      const uint8_t* s1 = s0 + sstride;
      uint8_t*       d  = dst + (size_t)y * dstride;
      for (int x = 0; x < dw; x++)
        for (int c = 0; c < channels; c++)
          d[x * channels + c] = (uint8_t)((s0[2 * x * channels + c] + s0[(2 * x + 1) * channels + c] +
                                           s1[2 * x * channels + c] + s1[(2 * x + 1) * channels + c] + 2) >> 2);
    }
    return;
  }
  // one channel at a time through the single-channel restatement (same coefficients per channel)
  std::vector<uint8_t> sp((size_t)sw * sh), dp((size_t)dw * dh);
  for (int c = 0; c < channels; c++) {
    for (int y = 0; y < sh; y++)
      for (int x = 0; x < sw; x++) sp[(size_t)y * sw + x] = src[(size_t)y * sstride + (size_t)x * channels + c];
    Img s, d;
    s.data = sp.data(); s.cols = sw; s.rows = sh; s.step = sw;
    d.data = dp.data(); d.cols = dw; d.rows = dh; d.step = dw;
    resize_linear_8u(s, d);
    for (int y = 0; y < dh; y++)
      for (int x = 0; x < dw; x++) dst[(size_t)y * dstride + (size_t)x * channels + c] = dp[(size_t)y * dw + x];
  }
}

}  // namespace orb_oracle

using namespace orb_oracle;

extern "C" {

void oo_cvt_gray(const uint8_t* src, int w, int h, size_t sstride, int channels, int rgb_order,
                 uint8_t* dst, size_t dstride) {
  cvt_gray_8u(src, w, h, sstride, channels, rgb_order, dst, dstride);
}

void oo_remap(const uint8_t* src, int sw, int sh, size_t sstride, int channels, const float* mapx,
              const float* mapy, uint8_t* dst, int dw, int dh, size_t dstride) {
  remap_linear_8u(src, sw, sh, sstride, channels, mapx, mapy, dst, dw, dh, dstride);
}

void oo_resize_cn(const uint8_t* src, int sw, int sh, size_t sstride, int channels, uint8_t* dst,
                  int dw, int dh, size_t dstride) {
  resize_linear_8u_cn(src, sw, sh, sstride, channels, dst, dw, dh, dstride);
}

// the whole pre-extraction chain of one image: mode 0 = none, 1 = remap, 2 = resize; then gray.
// fmt: 0 gray, 1 RGB, 2 BGR, 3 RGBA, 4 BGRA.  dst is dh x dw gray, tightly packed.
void oo_preprocess(const uint8_t* src, int sw, int sh, size_t sstride, int fmt, int mode,
                   const float* mapx, const float* mapy, uint8_t* dst, int dw, int dh) {
  const int ch = fmt == 0 ? 1 : (fmt <= 2 ? 3 : 4);
  std::vector<uint8_t> geo;
  const uint8_t* g = src;
  size_t gstride = sstride;
  int gw = sw, gh = sh;
  if (mode != 0) {
    geo.resize((size_t)dw * dh * ch);
    if (mode == 1) remap_linear_8u(src, sw, sh, sstride, ch, mapx, mapy, geo.data(), dw, dh, (size_t)dw * ch);
    else resize_linear_8u_cn(src, sw, sh, sstride, ch, geo.data(), dw, dh, (size_t)dw * ch);
    g = geo.data(); gstride = (size_t)dw * ch; gw = dw; gh = dh;
  }
  if (ch == 1) {
    for (int y = 0; y < gh; y++) memcpy(dst + (size_t)y * gw, g + (size_t)y * gstride, gw);
  } else {
    cvt_gray_8u(g, gw, gh, gstride, ch, fmt == 1 || fmt == 3, dst, gw);
  }
}

}  // extern "C"
