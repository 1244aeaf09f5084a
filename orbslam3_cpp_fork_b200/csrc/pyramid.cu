// K1 — scale pyramid (replaces ORBextractor::ComputePyramid, src/ORBextractor.cc:1140-1173).
// Level 0: copyMakeBorder(REFLECT_101) of the input.  Level l>0: cv::resize(INTER_LINEAR) of
// level l-1 (cascaded) with OpenCV's 11-bit fixed-point arithmetic, and the 19-px REFLECT_101
// border written by the same kernel.
//
// k_pyr_resize: one thread owns one aligned 32-bit word (4 px) of the padded row and walks down a
// strip of interior rows.  Per source row it issues three aligned 32-bit loads (the 4 output
// columns need source bytes inside a 12-byte window), extracts the two taps of each column with a
// precomputed PRMT selector and forms the horizontal sum with one dp2a (a0*p0 + a1*p1).  The
// horizontal sums of the last two source rows are kept in registers, so consecutive output rows
// (source step 1,1,1,1,2 at scale 1.2) reuse them: ~1.2 source rows are touched per output row
// instead of 2.  Border rows are not recomputed: an interior row y is also stored to the padded
// rows that REFLECT_101 maps onto it; border columns are ordinary words whose four columns are
// reflected before the tap lookup.  Bound: HBM (reads P_{l-1}, writes (w+38)(h+38)).
#include "common.cuh"

namespace cuorb {

constexpr int PYR_RS = 16;  // interior rows per thread strip

__device__ __forceinline__ int reflect101(int i, int n) {
  // border (19) < n is checked on the host, so one reflection suffices
  i = i < 0 ? -i : i;
  return i >= n ? 2 * n - 2 - i : i;
}

// grid: (ceil(pitch/4/128), ceil(h/PYR_RS), n_img)
__global__ void __launch_bounds__(128) k_pyr_level0(const ExtractorGeom* __restrict__ dg,
                                                    const uint8_t* __restrict__ imgs, size_t stride,
                                                    size_t image_stride, uint8_t* __restrict__ pyr,
                                                    int aligned) {
  const LevelGeom& L = dg->lv[0];
  const int word = blockIdx.x * blockDim.x + threadIdx.x;
  if (word * 4 >= L.pitch) return;
  const int img = blockIdx.z;
  const uint8_t* src = imgs + (size_t)img * image_stride;
  uint8_t* dst = pyr + (size_t)img * dg->pyr_img_stride + L.plane_off;
  const int xw = word * 4 - ROW_LEAD;  // level x of byte 0
  const bool inner = aligned && xw >= 0 && xw + 3 < L.w;
  int xs[4];
#pragma unroll
  for (int b = 0; b < 4; b++)
    xs[b] = reflect101(min(max(xw + b, -EDGE_THRESHOLD), L.w + EDGE_THRESHOLD - 1), L.w);  // pad bytes: clamp
  const int y0 = blockIdx.y * PYR_RS, y1 = min(y0 + PYR_RS, L.h);
  for (int y = y0; y < y1; y++) {
    const uint8_t* srow = src + (size_t)y * stride;
    unsigned out;
    if (inner) {
      out = __ldg(reinterpret_cast<const unsigned*>(srow + xw));
    } else {
      out = (unsigned)__ldg(srow + xs[0]) | ((unsigned)__ldg(srow + xs[1]) << 8) |
            ((unsigned)__ldg(srow + xs[2]) << 16) | ((unsigned)__ldg(srow + xs[3]) << 24);
    }
    reinterpret_cast<unsigned*>(dst + (size_t)(y + EDGE_THRESHOLD) * L.pitch)[word] = out;
    if (y >= 1 && y <= EDGE_THRESHOLD)
      reinterpret_cast<unsigned*>(dst + (size_t)(EDGE_THRESHOLD - y) * L.pitch)[word] = out;
    if (y >= L.h - 1 - EDGE_THRESHOLD && y <= L.h - 2)
      reinterpret_cast<unsigned*>(dst + (size_t)(2 * L.h + EDGE_THRESHOLD - 2 - y) * L.pitch)[word] = out;
  }
}

// horizontal sums (>> 4) of one source row for the thread's four columns
struct ColTaps {
  unsigned coef[4];   // a0 | a1 << 16
  unsigned sel[4];    // PRMT selector picking (p[sx], p[sx+1]) out of a word pair
  unsigned hi;        // bit b: column b's window starts in the second word pair (w1,w2)
  int      base;      // 4-aligned source x of the 12-byte window
  int      sx[4];     // only used by the wide fallback
  bool     wide;      // window does not fit 12 bytes (scale factors far above 2)
};

__device__ __forceinline__ void hrow(const uint8_t* __restrict__ srow, const ColTaps& T, int (&h)[4]) {
  if (!T.wide) {
    const unsigned* p = reinterpret_cast<const unsigned*>(srow + T.base);
    const unsigned w0 = p[0], w1 = p[1], w2 = p[2];
#pragma unroll
    for (int b = 0; b < 4; b++) {
      const bool up = (T.hi >> b) & 1;
      const unsigned v = __byte_perm(up ? w1 : w0, up ? w2 : w1, T.sel[b]);
      h[b] = (int)(__dp2a_lo(T.coef[b], v, 0u) >> 4);
    }
  } else {
#pragma unroll
    for (int b = 0; b < 4; b++)
      h[b] = (int)((srow[T.sx[b]] * (T.coef[b] & 0xFFFFu) + srow[T.sx[b] + 1] * (T.coef[b] >> 16)) >> 4);
  }
}

// grid: (ceil(pitch/4/128), ceil(h/PYR_RS), n_img)
__global__ void __launch_bounds__(128) k_pyr_resize(const ExtractorGeom* __restrict__ dg, int level,
                                                    uint8_t* __restrict__ pyr,
                                                    const ResizeTap* __restrict__ xtab,
                                                    const ResizeTap* __restrict__ ytab) {
  const LevelGeom& L = dg->lv[level];
  const LevelGeom& S = dg->lv[level - 1];
  const int word = blockIdx.x * blockDim.x + threadIdx.x;
  if (word * 4 >= L.pitch) return;
  const int img = blockIdx.z;
  uint8_t* base = pyr + (size_t)img * dg->pyr_img_stride;
  const uint8_t* src = base + S.pix_off;
  uint8_t* dst = base + L.plane_off;

  ColTaps T;
  {
    int smin = 1 << 30, smax = 0;
#pragma unroll
    for (int b = 0; b < 4; b++) {
      int x = word * 4 + b - ROW_LEAD;
      x = reflect101(min(max(x, -EDGE_THRESHOLD), L.w + EDGE_THRESHOLD - 1), L.w);
      const ResizeTap tx = xtab[L.xtab_off + x];
      // sx+1 may equal S.w: that byte is the stored border pixel and its weight a1 is 0 there
      T.sx[b] = tx.s;
      T.coef[b] = (unsigned)(unsigned short)tx.a0 | ((unsigned)(unsigned short)tx.a1 << 16);
      smin = min(smin, (int)tx.s); smax = max(smax, (int)tx.s);
    }
    T.base = smin & ~3;
    T.wide = smax + 1 - T.base > 11;
    T.hi = 0;
#pragma unroll
    for (int b = 0; b < 4; b++) {
      int t = T.sx[b] - T.base;
      if (t >= 4) { T.hi |= 1u << b; t -= 4; }
      T.sel[b] = (unsigned)t | ((unsigned)(t + 1) << 4);
    }
  }

  const int y0 = blockIdx.y * PYR_RS, y1 = min(y0 + PYR_RS, L.h);
  int rowA = -1, rowB = -1;
  int hA[4] = {0, 0, 0, 0}, hB[4] = {0, 0, 0, 0};
  for (int y = y0; y < y1; y++) {
    const ResizeTap ty = ytab[L.ytab_off + y];
    // OpenCV clips the two source rows to [0, sh-1] (resizeGeneric_Invoker)
    const int sy0 = min(max((int)ty.s, 0), S.h - 1), sy1 = min(max((int)ty.s + 1, 0), S.h - 1);
    if (sy0 == rowB) {
#pragma unroll
      for (int b = 0; b < 4; b++) hA[b] = hB[b];
      rowA = rowB;
    } else if (sy0 != rowA) {
      hrow(src + (size_t)sy0 * S.pitch, T, hA);
      rowA = sy0;
    }
    if (sy1 == rowA) {
#pragma unroll
      for (int b = 0; b < 4; b++) hB[b] = hA[b];
      rowB = rowA;
    } else if (sy1 != rowB) {
      hrow(src + (size_t)sy1 * S.pitch, T, hB);
      rowB = sy1;
    }
    const int b0 = ty.a0, b1 = ty.a1;
    unsigned out = 0;
#pragma unroll
    for (int b = 0; b < 4; b++) {
      const int v = (((b0 * hA[b]) >> 16) + ((b1 * hB[b]) >> 16) + 2) >> 2;
      out |= (unsigned)(v & 0xFF) << (8 * b);
    }
    reinterpret_cast<unsigned*>(dst + (size_t)(y + EDGE_THRESHOLD) * L.pitch)[word] = out;
    if (y >= 1 && y <= EDGE_THRESHOLD)
      reinterpret_cast<unsigned*>(dst + (size_t)(EDGE_THRESHOLD - y) * L.pitch)[word] = out;
    if (y >= L.h - 1 - EDGE_THRESHOLD && y <= L.h - 2)
      reinterpret_cast<unsigned*>(dst + (size_t)(2 * L.h + EDGE_THRESHOLD - 2 - y) * L.pitch)[word] = out;
  }
}

void launch_pyramid(const ExtractorGeom& g, const ExtractorGeom* dg, const uint8_t* imgs, int, int,
                    size_t stride, size_t image_stride, int n_img, uint8_t* pyr,
                    const ResizeTap* xtab, const ResizeTap* ytab, cudaStream_t st) {
  const int aligned = (((uintptr_t)imgs | stride | image_stride) & 3) == 0;
  for (int l = 0; l < g.nlevels; l++) {
    const LevelGeom& L = g.lv[l];
    dim3 grid((L.pitch / 4 + 127) / 128, (L.h + PYR_RS - 1) / PYR_RS, n_img);
    if (l == 0) k_pyr_level0<<<grid, 128, 0, st>>>(dg, imgs, stride, image_stride, pyr, aligned);
    else k_pyr_resize<<<grid, 128, 0, st>>>(dg, l, pyr, xtab, ytab);
    count_launch();
  }
}

}  // namespace cuorb
