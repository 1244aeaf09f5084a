// K1 — scale pyramid (replaces ORBextractor::ComputePyramid, src/ORBextractor.cc:1140-1173).
// Level 0: copyMakeBorder(REFLECT_101) of the input.  Level l>0: cv::resize(INTER_LINEAR) of
// level l-1 (cascaded) with OpenCV's 11-bit fixed-point arithmetic, and the 19-px REFLECT_101
// border written by the same kernel (border pixels recompute the reflected interior pixel).
// One thread produces one aligned 32-bit word (4 px) of the padded plane; a warp writes 128
// contiguous bytes.  Bound: HBM (reads P_{l-1}, writes (w+38)(h+38)).
#include "common.cuh"

namespace cuorb {

__device__ __forceinline__ int reflect101(int i, int n) {
  // border (19) < n is checked on the host, so one reflection suffices
  i = i < 0 ? -i : i;
  return i >= n ? 2 * n - 2 - i : i;
}

// grid: (ceil(pitch/4/128), h+38, n_img)
__global__ void __launch_bounds__(128) k_pyr_level0(const ExtractorGeom* __restrict__ dg,
                                                    const uint8_t* __restrict__ imgs, size_t stride,
                                                    size_t image_stride, uint8_t* __restrict__ pyr) {
  const LevelGeom& L = dg->lv[0];
  const int word = blockIdx.x * blockDim.x + threadIdx.x;
  if (word * 4 >= L.pitch) return;
  const int py  = blockIdx.y;
  const int img = blockIdx.z;
  const uint8_t* src = imgs + (size_t)img * image_stride;
  const int sy = reflect101(py - EDGE_THRESHOLD, L.h);
  const uint8_t* srow = src + (size_t)sy * stride;
  unsigned out = 0;
#pragma unroll
  for (int b = 0; b < 4; b++) {
    int x = word * 4 + b - ROW_LEAD;  // level x coordinate
    x = min(max(x, -EDGE_THRESHOLD), L.w + EDGE_THRESHOLD - 1);  // pad bytes: clamp (never read)
    out |= (unsigned)__ldg(srow + reflect101(x, L.w)) << (8 * b);
  }
  uint8_t* dst = pyr + (size_t)img * dg->pyr_img_stride + L.plane_off + (size_t)py * L.pitch;
  reinterpret_cast<unsigned*>(dst)[word] = out;
}

// grid: (ceil(pitch/4/128), h+38, n_img)
__global__ void __launch_bounds__(128) k_pyr_resize(const ExtractorGeom* __restrict__ dg, int level,
                                                    uint8_t* __restrict__ pyr,
                                                    const ResizeTap* __restrict__ xtab,
                                                    const ResizeTap* __restrict__ ytab) {
  const LevelGeom& L = dg->lv[level];
  const LevelGeom& S = dg->lv[level - 1];
  const int word = blockIdx.x * blockDim.x + threadIdx.x;
  if (word * 4 >= L.pitch) return;
  const int py  = blockIdx.y;
  const int img = blockIdx.z;
  uint8_t* base = pyr + (size_t)img * dg->pyr_img_stride;
  const uint8_t* src = base + S.pix_off;
  const int y = reflect101(py - EDGE_THRESHOLD, L.h);
  const ResizeTap ty = ytab[L.ytab_off + y];
  // OpenCV clips the two source rows to [0, sh-1] (resizeGeneric_Invoker)
  const int sy0 = min(max((int)ty.s, 0), S.h - 1), sy1 = min(max((int)ty.s + 1, 0), S.h - 1);
  const uint8_t* r0 = src + (size_t)sy0 * S.pitch;
  const uint8_t* r1 = src + (size_t)sy1 * S.pitch;
  const int b0 = ty.a0, b1 = ty.a1;
  unsigned out = 0;
#pragma unroll
  for (int b = 0; b < 4; b++) {
    int x = word * 4 + b - ROW_LEAD;
    x = min(max(x, -EDGE_THRESHOLD), L.w + EDGE_THRESHOLD - 1);
    x = reflect101(x, L.w);
    const ResizeTap tx = xtab[L.xtab_off + x];
    // sx+1 may equal S.w: that byte is the stored border pixel and its weight a1 is 0 there
    const int h0 = r0[tx.s] * tx.a0 + r0[tx.s + 1] * tx.a1;
    const int h1 = r1[tx.s] * tx.a0 + r1[tx.s + 1] * tx.a1;
    const int v  = (((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2;
    out |= (unsigned)(v & 0xFF) << (8 * b);
  }
  uint8_t* dst = base + L.plane_off + (size_t)py * L.pitch;
  reinterpret_cast<unsigned*>(dst)[word] = out;
}

void launch_pyramid(const ExtractorGeom& g, const ExtractorGeom* dg, const uint8_t* imgs, int, int,
                    size_t stride, size_t image_stride, int n_img, uint8_t* pyr,
                    const ResizeTap* xtab, const ResizeTap* ytab, cudaStream_t st) {
  for (int l = 0; l < g.nlevels; l++) {
    const LevelGeom& L = g.lv[l];
    dim3 grid((L.pitch / 4 + 127) / 128, L.h + 2 * EDGE_THRESHOLD, n_img);
    if (l == 0) k_pyr_level0<<<grid, 128, 0, st>>>(dg, imgs, stride, image_stride, pyr);
    else k_pyr_resize<<<grid, 128, 0, st>>>(dg, l, pyr, xtab, ytab);
    count_launch();
  }
}

}  // namespace cuorb
