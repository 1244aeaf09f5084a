// K7 — 7x7 sigma-2 Gaussian blur of every pyramid level (replaces the per-level
// clone() + cv::GaussianBlur(7x7, 2, 2, BORDER_REFLECT_101), src/ORBextractor.cc:1103-1104).
// OpenCV's CV_8U path is fixed point: kernel {18,34,48,56,48,34,18}/256, rows in 8.8, columns in
// 16.16, result (v + 32768) >> 16.  The REFLECT_101 halo is the border K1 already stored, so no
// reflection is computed here.
//
// One thread owns one aligned output word (4 px) and walks down a strip of rows, entirely in
// registers: per source row three aligned 32-bit loads, the four 7-tap row sums as two DP4A each
// on funnel-shifted words (taps {18,34,48,56} and {48,34,18,0}), and a rolling window of the last
// seven row sums feeds the column pass (symmetric taps: 3 adds + 4 IMAD per pixel).  No shared
// memory, no barriers; the loop is unrolled by 7 so the window rotates by register renaming.
// Bound: HBM (read P, write P); ~14 instructions per pixel.
#include "common.cuh"

namespace cuorb {

constexpr int BL_THREADS = 64;   // words (4 px each) per CTA row segment
constexpr int BL_ROWS    = 21;   // output rows per thread strip

__device__ __forceinline__ void blur_hrow(const uint8_t* __restrict__ row, unsigned (&h)[4]) {
  const unsigned* p = reinterpret_cast<const unsigned*>(row);
  const unsigned A = p[-1], B = p[0], C = p[1];
  const unsigned K0 = 18u | (34u << 8) | (48u << 16) | (56u << 24), K1 = 48u | (34u << 8) | (18u << 16);
  h[0] = __dp4a(__funnelshift_r(A, B, 8), K0, __dp4a(__funnelshift_r(B, C, 8), K1, 0u));
  h[1] = __dp4a(__funnelshift_r(A, B, 16), K0, __dp4a(__funnelshift_r(B, C, 16), K1, 0u));
  h[2] = __dp4a(__funnelshift_r(A, B, 24), K0, __dp4a(__funnelshift_r(B, C, 24), K1, 0u));
  h[3] = __dp4a(B, K0, __dp4a(C, K1, 0u));
}

// grid: (tiles over all levels, n_img); a tile = BL_THREADS consecutive (row strip, word) items of one
// level, numbered strip-major, so only the last CTA of a level has idle lanes (a lane's neighbours may
// belong to the next strip: both are contiguous row segments, still coalesced)
__global__ void __launch_bounds__(BL_THREADS) k_blur(const ExtractorGeom* __restrict__ dg,
                                                     const int4* __restrict__ tiles,
                                                     const uint8_t* __restrict__ pyr,
                                                     uint8_t* __restrict__ blur) {
  const int4 t = tiles[blockIdx.x];   // {level, first item, words per row, items in the level}
  const LevelGeom& L = dg->lv[t.x];
  const int item = t.y + threadIdx.x;
  if (item >= t.w) return;
  const int strip = item / t.z, word = item - strip * t.z;
  const int img = blockIdx.y;
  const int y0 = strip * BL_ROWS, y1 = min(y0 + BL_ROWS, L.h);
  const uint8_t* src = pyr + (size_t)img * dg->pyr_img_stride + L.pix_off + word * 4;
  uint8_t* dst = blur + (size_t)img * dg->blur_img_stride + L.blur_off + word * 4;
  const int pitch = L.pitch, bpitch = L.bpitch;

  unsigned win[7][4];  // row sums of source rows y-3 .. y+3
#pragma unroll
  for (int j = 0; j < 6; j++) blur_hrow(src + (ptrdiff_t)(y0 - 3 + j) * pitch, win[j]);
  for (int yb = y0; yb < y1; yb += 7) {
#pragma unroll
    for (int u = 0; u < 7; u++) {
      const int y = yb + u;
      if (y < y1) {
        // slot (u + 6) % 7 receives source row y + 3; the window then is slots u, u+1, ..., u+6 (mod 7)
        blur_hrow(src + (ptrdiff_t)(y + 3) * pitch, win[(u + 6) % 7]);
        unsigned out = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
          const unsigned acc = 32768u + 18u * (win[u][k] + win[(u + 6) % 7][k]) +
                               34u * (win[(u + 1) % 7][k] + win[(u + 5) % 7][k]) +
                               48u * (win[(u + 2) % 7][k] + win[(u + 4) % 7][k]) + 56u * win[(u + 3) % 7][k];
          out |= (acc >> 16) << (8 * k);
        }
        *reinterpret_cast<unsigned*>(dst + (size_t)y * bpitch) = out;
      }
    }
  }
}

void build_blur_tiles(const ExtractorGeom& g, std::vector<int4>& v) {
  v.clear();
  for (int l = 0; l < g.nlevels; l++) {
    const int words = g.lv[l].bpitch / 4;
    const int items = words * ((g.lv[l].h + BL_ROWS - 1) / BL_ROWS);
    for (int i = 0; i < items; i += BL_THREADS) v.push_back(make_int4(l, i, words, items));
  }
}

void launch_blur(const ExtractorGeom*, const ExtractorGeom* dg, int n_img, const int4* tiles,
                 int n_tiles, const uint8_t* pyr, uint8_t* blur, cudaStream_t st) {
  k_blur<<<dim3(n_tiles, n_img), BL_THREADS, 0, st>>>(dg, tiles, pyr, blur);
  count_launch();
}

}  // namespace cuorb
