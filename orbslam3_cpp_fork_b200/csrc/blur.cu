// K7 — 7x7 sigma-2 Gaussian blur of every pyramid level (replaces the per-level
// clone() + cv::GaussianBlur(7x7, 2, 2, BORDER_REFLECT_101), src/ORBextractor.cc:1103-1104).
// OpenCV's CV_8U path is fixed point: kernel {18,34,48,56,48,34,18}/256, rows in 8.8, columns in
// 16.16, result (v + 32768) >> 16.  The REFLECT_101 halo is the border K1 already stored, so the
// tile loader never reflects.  Separable in shared memory: 128x32 output tile per 256-thread CTA,
// aligned 32-bit loads, 16-bit row sums, uchar4 stores.  Bound: HBM (read P, write P).
#include "common.cuh"

namespace cuorb {

constexpr int BT_W = 128, BT_H = 32, BT_IN_H = BT_H + 6, BT_IN_WORDS = (BT_W + 8) / 4;

// grid: (tiles over all levels, n_img); tile list in constant-sized table passed by pointer
__global__ void __launch_bounds__(256) k_blur(const ExtractorGeom* __restrict__ dg,
                                              const int4* __restrict__ tiles,
                                              const uint8_t* __restrict__ pyr,
                                              uint8_t* __restrict__ blur) {
  __shared__ unsigned       s_in[BT_IN_H][BT_IN_WORDS];  // bytes x0-4 .. x0+131
  __shared__ unsigned short s_mid[BT_IN_H][BT_W];

  const int4 t = tiles[blockIdx.x];
  const LevelGeom& L = dg->lv[t.x];
  const int x0 = t.y * BT_W, y0 = t.z * BT_H;
  const int img = blockIdx.y;
  const uint8_t* src = pyr + (size_t)img * dg->pyr_img_stride + L.pix_off;
  const int tid = threadIdx.x;

  for (int i = tid; i < BT_IN_H * BT_IN_WORDS; i += 256) {
    const int r = i / BT_IN_WORDS, c = i % BT_IN_WORDS;
    const int y = min(y0 - 3 + r, L.h + EDGE_THRESHOLD - 1);  // rows past the border are unused
    const int xb = min(x0 - 4 + c * 4, L.pitch - ROW_LEAD - 4);  // stay inside the padded row
    s_in[r][c] = *reinterpret_cast<const unsigned*>(src + (ptrdiff_t)y * L.pitch + xb);
  }
  __syncthreads();

  const int lx = tid & 31, ly = tid >> 5;
  // horizontal pass: 4 px per thread per row
  for (int r = ly; r < BT_IN_H; r += 8) {
    const unsigned w0 = s_in[r][lx], w1 = s_in[r][lx + 1], w2 = s_in[r][lx + 2];
    // bytes at tile offsets lx*4+1 .. lx*4+10  (x-3 .. x+6 for x = x0 + lx*4)
    int p[10];
    p[0] = (w0 >> 8) & 0xFF; p[1] = (w0 >> 16) & 0xFF; p[2] = w0 >> 24;
    p[3] = w1 & 0xFF; p[4] = (w1 >> 8) & 0xFF; p[5] = (w1 >> 16) & 0xFF; p[6] = w1 >> 24;
    p[7] = w2 & 0xFF; p[8] = (w2 >> 8) & 0xFF; p[9] = (w2 >> 16) & 0xFF;
    unsigned short o[4];
#pragma unroll
    for (int k = 0; k < 4; k++)
      o[k] = (unsigned short)(18 * (p[k] + p[k + 6]) + 34 * (p[k + 1] + p[k + 5]) +
                              48 * (p[k + 2] + p[k + 4]) + 56 * p[k + 3]);
    *reinterpret_cast<uint2*>(&s_mid[r][lx * 4]) =
        make_uint2(o[0] | ((unsigned)o[1] << 16), o[2] | ((unsigned)o[3] << 16));
  }
  __syncthreads();

  // vertical pass: 4 rows x 4 px per thread
  unsigned m[10][4];
#pragma unroll
  for (int r = 0; r < 10; r++) {
    const uint2 v = *reinterpret_cast<const uint2*>(&s_mid[ly * 4 + r][lx * 4]);
    m[r][0] = v.x & 0xFFFF; m[r][1] = v.x >> 16; m[r][2] = v.y & 0xFFFF; m[r][3] = v.y >> 16;
  }
  uint8_t* dst = blur + (size_t)img * dg->blur_img_stride + L.blur_off;
  const int x = x0 + lx * 4;
  if (x < L.bpitch) {
#pragma unroll
    for (int r = 0; r < 4; r++) {
      const int y = y0 + ly * 4 + r;
      if (y >= L.h) break;
      unsigned out = 0;
#pragma unroll
      for (int k = 0; k < 4; k++) {
        const unsigned acc = 18u * (m[r][k] + m[r + 6][k]) + 34u * (m[r + 1][k] + m[r + 5][k]) +
                             48u * (m[r + 2][k] + m[r + 4][k]) + 56u * m[r + 3][k];
        out |= ((acc + 32768u) >> 16) << (8 * k);
      }
      *reinterpret_cast<unsigned*>(dst + (size_t)y * L.bpitch + x) = out;
    }
  }
}

void build_blur_tiles(const ExtractorGeom& g, std::vector<int4>& v) {
  v.clear();
  for (int l = 0; l < g.nlevels; l++)
    for (int ty = 0; ty < (g.lv[l].h + BT_H - 1) / BT_H; ty++)
      for (int tx = 0; tx < (g.lv[l].w + BT_W - 1) / BT_W; tx++) v.push_back(make_int4(l, tx, ty, 0));
}

void launch_blur(const ExtractorGeom*, const ExtractorGeom* dg, int n_img, const int4* tiles,
                 int n_tiles, const uint8_t* pyr, uint8_t* blur, cudaStream_t st) {
  k_blur<<<dim3(n_tiles, n_img), 256, 0, st>>>(dg, tiles, pyr, blur);
  count_launch();
}

}  // namespace cuorb
