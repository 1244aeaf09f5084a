// K0 — the per-frame image operations the reference runs right before ORBextractor::operator()
// (SURVEY.md §8f row N1), fused into ONE pass that writes the gray level-0 input plane:
//   System::TrackStereo/TrackRGBD/TrackMonocular  src/System.cc:280-300, 384-386, 476
//       cv::remap(im, out, M1, M2, INTER_LINEAR)  |  cv::resize(im, out, newImSize)
//   Tracking::GrabImageStereo/RGBD/Monocular      src/Tracking.cc:1339-1357, 1449-1462, 1520-1533
//       cv::cvtColor(RGB|BGR|RGBA|BGRA -> GRAY)
// in the reference's order (geometric op on the raw channels, each channel rounded to 8 bits, then
// the gray conversion), with OpenCV's integer arithmetic:
//   remap   sx = cvRound(mapx*32), sy = cvRound(mapy*32); taps at (sx>>5, sy>>5) (+1), 5-bit
//           fractions, weights (32-fx)(32-fy)*32 ... (sum 2^15), out = (sum + 2^14) >> 15,
//           BORDER_CONSTANT 0 per tap;
//   resize  INTER_LINEAR 11-bit coefficients (host-built tables, same float sequence as the pyramid),
//           out = (((b0*(H0>>4))>>16) + ((b1*(H1>>4))>>16) + 2) >> 2; exact 2x2 decimation is
//           OpenCV's INTER_AREA reroute (a+b+c+d+2)>>2;
//   gray    (B*3735 + G*19235 + R*9798 + 2^14) >> 15.
// Lanes of a warp work on adjacent pixels, four pixels per thread 64 apart.
// Bound: L1 wavefronts of the byte gathers — src bytes + 1 byte per output pixel of HBM traffic.
#include "common.cuh"

namespace cuorb {

namespace {

template <int C>
__device__ __forceinline__ int gray_of(const int (&v)[4], int ir, int ib) {
  if (C == 1) return v[0];
  return (v[ib] * 3735 + v[1] * 19235 + v[ir] * 9798 + (1 << 14)) >> 15;
}

template <int C>
__device__ __forceinline__ int px_none(const PreArgs& a, const uint8_t* img, int x, int y) {
  const uint8_t* p = img + (size_t)y * a.sstride + (size_t)x * C;
  int v[4] = {0, 0, 0, 0};
#pragma unroll
  for (int c = 0; c < (C > 3 ? 3 : C); c++) v[c] = __ldg(p + c);
  return gray_of<C>(v, a.ir, a.ib);
}

template <int C>
__device__ __forceinline__ int px_remap(const PreArgs& a, const uint8_t* img, int x, int y) {
  const size_t m = (size_t)y * a.dw + x;
  const int sx = __float2int_rn(__fmul_rn(__ldg(a.mapx + m), 32.0f));
  const int sy = __float2int_rn(__fmul_rn(__ldg(a.mapy + m), 32.0f));
  const int ix = min(max(sx >> 5, -32768), 32767), iy = min(max(sy >> 5, -32768), 32767);
  const int fx = sx & 31, fy = sy & 31;
  const int w00 = (32 - fx) * (32 - fy) * 32, w01 = fx * (32 - fy) * 32;
  const int w10 = (32 - fx) * fy * 32, w11 = fx * fy * 32;
  const bool x0 = (unsigned)ix < (unsigned)a.sw, x1 = (unsigned)(ix + 1) < (unsigned)a.sw;
  const bool y0 = (unsigned)iy < (unsigned)a.sh, y1 = (unsigned)(iy + 1) < (unsigned)a.sh;
  const uint8_t* p = img + (ptrdiff_t)iy * (ptrdiff_t)a.sstride + (ptrdiff_t)ix * C;
  int v[4] = {0, 0, 0, 0};
#pragma unroll
  for (int c = 0; c < (C > 3 ? 3 : C); c++) {
    int s = 0;
    if (y0 && x0) s += __ldg(p + c) * w00;
    if (y0 && x1) s += __ldg(p + C + c) * w01;
    if (y1 && x0) s += __ldg(p + a.sstride + c) * w10;
    if (y1 && x1) s += __ldg(p + a.sstride + C + c) * w11;
    v[c] = (s + (1 << 14)) >> 15;
  }
  return gray_of<C>(v, a.ir, a.ib);
}

template <int C>
__device__ __forceinline__ int px_resize(const PreArgs& a, const uint8_t* img, int x, int y) {
  int v[4] = {0, 0, 0, 0};
  if (a.area2) {
    const uint8_t* p0 = img + (size_t)(2 * y) * a.sstride + (size_t)(2 * x) * C;
    const uint8_t* p1 = p0 + a.sstride;
#pragma unroll
    for (int c = 0; c < (C > 3 ? 3 : C); c++)
      v[c] = (__ldg(p0 + c) + __ldg(p0 + C + c) + __ldg(p1 + c) + __ldg(p1 + C + c) + 2) >> 2;
  } else {
    const int sx = __ldg(a.xofs + x), a0 = __ldg(a.xa + 2 * x), a1 = __ldg(a.xa + 2 * x + 1);
    const int sy = __ldg(a.yofs + y), b0 = __ldg(a.yb + 2 * y), b1 = __ldg(a.yb + 2 * y + 1);
    const int r0 = min(max(sy, 0), a.sh - 1), r1 = min(max(sy + 1, 0), a.sh - 1);
    const bool two = sx + 1 < a.sw;
    const uint8_t* p0 = img + (size_t)r0 * a.sstride + (size_t)sx * C;
    const uint8_t* p1 = img + (size_t)r1 * a.sstride + (size_t)sx * C;
#pragma unroll
    for (int c = 0; c < (C > 3 ? 3 : C); c++) {
      const int h0 = two ? __ldg(p0 + c) * a0 + __ldg(p0 + C + c) * a1 : __ldg(p0 + c) * 2048;
      const int h1 = two ? __ldg(p1 + c) * a0 + __ldg(p1 + C + c) * a1 : __ldg(p1 + c) * 2048;
      v[c] = (((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2;
    }
  }
  return gray_of<C>(v, a.ir, a.ib);
}

// grid: (ceil(dw/256), dh, n_img), block 64.  A thread produces pixels x0 + lane + 64k (k = 0..3):
// the lanes of a warp read ADJACENT source pixels (one or two L1 wavefronts per tap instead of four)
// and each byte store of a warp fills one 32-byte sector.
template <int C, int MODE>
__global__ void __launch_bounds__(64) k_preproc(const PreArgs a) {
  const int x0 = blockIdx.x * 256 + threadIdx.x, y = blockIdx.y, img = blockIdx.z;
  const uint8_t* src = a.src + (size_t)img * a.simg_stride;
  uint8_t* d = a.dst + (size_t)img * a.dimg_stride + (size_t)y * a.dstride;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const int x = x0 + 64 * k;
    if (x >= a.dw) break;
    int g;
    if (MODE == 1) g = px_remap<C>(a, src, x, y);
    else if (MODE == 2) g = px_resize<C>(a, src, x, y);
    else g = px_none<C>(a, src, x, y);
    d[x] = (uint8_t)g;
  }
}

template <int C>
void launch_c(const PreArgs& a, int mode, int n_img, cudaStream_t st) {
  const dim3 grid((a.dw + 255) / 256, a.dh, n_img);
  if (mode == 1) k_preproc<C, 1><<<grid, 64, 0, st>>>(a);
  else if (mode == 2) k_preproc<C, 2><<<grid, 64, 0, st>>>(a);
  else k_preproc<C, 0><<<grid, 64, 0, st>>>(a);
}

}  // namespace

void launch_preproc(const PreArgs& a, int channels, int mode, int n_img, cudaStream_t st) {
  if (channels == 1) launch_c<1>(a, mode, n_img, st);
  else if (channels == 3) launch_c<3>(a, mode, n_img, st);
  else launch_c<4>(a, mode, n_img, st);
  count_launch();
}

}  // namespace cuorb
