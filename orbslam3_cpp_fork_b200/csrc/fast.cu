// K2 — gridded FAST-9/16 with the iniThFAST/minThFAST retry (replaces the cell loop of
// ORBextractor::ComputeKeyPointsOctTree + cv::FAST(roi, thr, nonmax=true),
// src/ORBextractor.cc:755-850; OpenCV features2d/fast.cpp FAST_t<16>, fast_score.cpp).
//
// One CTA per (cell, image).  The detection areas of the reference's overlapping cell ROIs tile
// the level exactly, so a cell's CTA only needs its own pixels + a 3-px ring halo, and NMS looks
// only at scores inside the cell (OpenCV zeroes everything outside the ROI's detection area).
//   phase 1: every pixel -> "is a corner at minThFAST" (opposite-pair reject, then 16-bit
//            brighter/darker ring masks and a run-of-9 bit test); corners are compacted into a
//            shared-memory list so phase 2 runs with full warps.
//   phase 2: cornerScore (threshold independent: max over arcs of min |diff|, minus 1) -> u8 tile
//   phase 3: strict 8-neighbour NMS inside the cell; CTA-wide "any survivor >= iniThFAST";
//            survivors at the chosen threshold are appended to the level's candidate list.
// The list is unordered; the reference order (cell-row-major, then row-major) is a pure function
// of (x, y) and is re-derived where it matters (quadtree tie-break, debug download).
#include "common.cuh"

namespace cuorb {

__constant__ int c_ring_dx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
__constant__ int c_ring_dy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};

__device__ __forceinline__ bool has_run9(unsigned m16) {
  unsigned x = m16 | (m16 << 16);
  x &= x >> 1;  // runs >= 2
  x &= x >> 2;  // >= 4
  x &= x >> 4;  // >= 8
  x &= x >> 1;  // >= 9
  return (x & 0xFFFFu) != 0;
}

// grid: (total_cells, n_img), block 128, dynamic smem = tile + score + list
__global__ void __launch_bounds__(128) k_fast_cells(const ExtractorGeom* __restrict__ dg,
                                                    const uint8_t* __restrict__ pyr,
                                                    unsigned* __restrict__ cand,
                                                    int* __restrict__ cand_count,
                                                    int* __restrict__ flags, int tile_pitch,
                                                    int tile_bytes, int score_bytes) {
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ int s_ncorner, s_any_ini, s_nout, s_base;

  // which level / cell
  int level = 0;
  const int nlv = dg->nlevels;
  while (level + 1 < nlv && (int)blockIdx.x >= dg->lv[level + 1].cell_base) level++;
  const LevelGeom& L = dg->lv[level];
  const int cell = blockIdx.x - L.cell_base;
  const int ci = cell / L.nCols, cj = cell % L.nCols;
  const int x0 = EDGE_THRESHOLD + cj * L.wCell, y0 = EDGE_THRESHOLD + ci * L.hCell;
  const int dw = min(x0 + L.wCell, L.maxBorderX - 3) - x0;
  const int dh = min(y0 + L.hCell, L.maxBorderY - 3) - y0;
  if (dw <= 0 || dh <= 0) return;  // cells the reference skips (:781, :791) or that cannot hold a corner

  const int img = blockIdx.y;
  const int tid = threadIdx.x;
  const int ini_th = dg->ini_th, min_th = dg->min_th;

  unsigned char*  s_tile  = smem;                                   // (dh+6) rows of tile_pitch
  unsigned char*  s_score = smem + tile_bytes;                      // (dh+2) x (dw+2), zero frame
  unsigned short* s_list  = reinterpret_cast<unsigned short*>(smem + tile_bytes + score_bytes);
  const int sp = dw + 2;

  if (tid == 0) { s_ncorner = 0; s_any_ini = 0; s_nout = 0; }

  // load pixel tile with aligned 32-bit words: level x in [ax, ...), rows y0-3 .. y0+dh+2
  const uint8_t* src = pyr + (size_t)img * dg->pyr_img_stride + L.pix_off;
  const int gx0 = x0 - 3, ax = gx0 & ~3, ph = gx0 - ax;
  const int wpr = (ph + dw + 6 + 3) >> 2;  // words per row
  for (int i = tid; i < wpr * (dh + 6); i += 128) {
    const int r = i / wpr, c = i - r * wpr;
    reinterpret_cast<unsigned*>(s_tile + r * tile_pitch)[c] =
        *reinterpret_cast<const unsigned*>(src + (ptrdiff_t)(y0 - 3 + r) * L.pitch + ax + c * 4);
  }
  for (int i = tid; i < (sp * (dh + 2) + 3) / 4; i += 128) reinterpret_cast<unsigned*>(s_score)[i] = 0;
  __syncthreads();

  // phase 1: corner test at minThFAST
  const int npx = dw * dh;
  for (int i = tid; i < npx; i += 128) {
    const int py = i / dw, px = i - py * dw;
    const unsigned char* p = s_tile + (py + 3) * tile_pitch + ph + px + 3;
    const int v = p[0], hi = v + min_th, lo = v - min_th;
    const int r0 = p[3 * tile_pitch], r8 = p[-3 * tile_pitch], r4 = p[3], r12 = p[-3];
    const bool qb = (r0 > hi || r8 > hi) && (r4 > hi || r12 > hi);
    const bool qd = (r0 < lo || r8 < lo) && (r4 < lo || r12 < lo);
    if (!qb && !qd) continue;
    unsigned mb = 0, md = 0;
#pragma unroll
    for (int k = 0; k < 16; k++) {
      const int r = p[c_ring_dy[k] * tile_pitch + c_ring_dx[k]];
      mb |= (unsigned)(r > hi) << k;
      md |= (unsigned)(r < lo) << k;
    }
    if (has_run9(mb) || has_run9(md)) s_list[atomicAdd(&s_ncorner, 1)] = (unsigned short)i;
  }
  __syncthreads();

  // phase 2: scores of the corners
  const int nc = s_ncorner;
  for (int n = tid; n < nc; n += 128) {
    const int i = s_list[n];
    const int py = i / dw, px = i - py * dw;
    const unsigned char* p = s_tile + (py + 3) * tile_pitch + ph + px + 3;
    const int v = p[0];
    int d[16];
#pragma unroll
    for (int k = 0; k < 16; k++) d[k] = v - (int)p[c_ring_dy[k] * tile_pitch + c_ring_dx[k]];
    int mn[16], mx[16];
#pragma unroll
    for (int k = 0; k < 16; k++) { mn[k] = min(d[k], d[(k + 1) & 15]); mx[k] = max(d[k], d[(k + 1) & 15]); }
    int mn4[16], mx4[16];
#pragma unroll
    for (int k = 0; k < 16; k++) { mn4[k] = min(mn[k], mn[(k + 2) & 15]); mx4[k] = max(mx[k], mx[(k + 2) & 15]); }
    int a = -1000, b = 1000;
#pragma unroll
    for (int k = 0; k < 16; k++) {
      const int m9 = min(min(mn4[k], mn4[(k + 4) & 15]), d[(k + 8) & 15]);
      const int x9 = max(max(mx4[k], mx4[(k + 4) & 15]), d[(k + 8) & 15]);
      a = max(a, m9);
      b = min(b, x9);
    }
    s_score[(py + 1) * sp + px + 1] = (unsigned char)(max(a, -b) - 1);
  }
  __syncthreads();

  // phase 3: NMS inside the cell, threshold choice, emit
  unsigned mine[4];  // up to 4 local maxima per thread kept in registers, rest re-evaluated
  int n_mine = 0;
  bool any_ini = false;
  for (int n = tid; n < nc; n += 128) {
    const int i = s_list[n];
    const int py = i / dw, px = i - py * dw;
    const unsigned char* q = s_score + (py + 1) * sp + px + 1;
    const int s = q[0];
    const bool lm = s > q[-1] && s > q[1] && s > q[-sp - 1] && s > q[-sp] && s > q[-sp + 1] &&
                    s > q[sp - 1] && s > q[sp] && s > q[sp + 1];
    if (lm) {
      any_ini |= s >= ini_th;
      // region coordinates of vToDistributeKeys: roi pt + (j*wCell, i*hCell)   (:842-846)
      const unsigned c = pack_cand(x0 + px - FAST_BORDER, y0 + py - FAST_BORDER, s);
      if (n_mine < 4) mine[n_mine] = c;
      n_mine++;
    }
  }
  if (any_ini) s_any_ini = 1;
  __syncthreads();
  const int thr = s_any_ini ? ini_th : min_th;

  // count survivors at thr, reserve, write
  int my_out = 0;
  if (n_mine <= 4) {
    for (int k = 0; k < n_mine; k++) my_out += cand_s(mine[k]) >= thr;
  } else {
    my_out = -1;  // rare: re-walk below
  }
  if (my_out < 0) {
    my_out = 0;
    for (int n = tid; n < nc; n += 128) {
      const int i = s_list[n];
      const int py = i / dw, px = i - py * dw;
      const unsigned char* q = s_score + (py + 1) * sp + px + 1;
      const int s = q[0];
      const bool lm = s > q[-1] && s > q[1] && s > q[-sp - 1] && s > q[-sp] && s > q[-sp + 1] &&
                      s > q[sp - 1] && s > q[sp] && s > q[sp + 1];
      my_out += lm && s >= thr;
    }
  }
  int my_off = 0;
  if (my_out) my_off = atomicAdd(&s_nout, my_out);
  __syncthreads();
  const int nout = s_nout;
  if (nout == 0) return;
  const int slot = img * nlv + level;
  if (tid == 0) s_base = atomicAdd(&cand_count[slot], nout);
  __syncthreads();
  const int base = s_base;
  if (base + nout > L.cand_cap) {  // cannot happen with the worst-case capacity; flagged anyway
    if (tid == 0) { atomicOr(flags, FLAG_CAND_OVERFLOW); }
    return;
  }
  unsigned* out = cand + (size_t)img * dg->cand_img_stride + L.cand_off + base + my_off;
  if (n_mine <= 4) {
    for (int k = 0; k < n_mine; k++)
      if (cand_s(mine[k]) >= thr) *out++ = mine[k];
  } else {
    for (int n = tid; n < nc; n += 128) {
      const int i = s_list[n];
      const int py = i / dw, px = i - py * dw;
      const unsigned char* q = s_score + (py + 1) * sp + px + 1;
      const int s = q[0];
      const bool lm = s > q[-1] && s > q[1] && s > q[-sp - 1] && s > q[-sp] && s > q[-sp + 1] &&
                      s > q[sp - 1] && s > q[sp] && s > q[sp + 1];
      if (lm && s >= thr) *out++ = pack_cand(x0 + px - FAST_BORDER, y0 + py - FAST_BORDER, s);
    }
  }
}

void launch_fast(const ExtractorGeom& g, const ExtractorGeom* dg, int n_img, const uint8_t* pyr,
                 unsigned* cand, int* cand_count, int* flags, cudaStream_t st) {
  const int tile_pitch  = ((3 + g.max_cell_w + 6 + 3) / 4) * 4;
  const int tile_bytes  = ((tile_pitch * (g.max_cell_h + 6) + 15) / 16) * 16;
  const int score_bytes = (((g.max_cell_w + 2) * (g.max_cell_h + 2) + 15) / 16) * 16;
  const int list_bytes  = 2 * g.max_cell_w * g.max_cell_h;
  const int smem        = tile_bytes + score_bytes + list_bytes;
  static bool attr_set = false;
  if (!attr_set && smem > 48 * 1024) {
    cudaFuncSetAttribute(k_fast_cells, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    attr_set = true;
  }
  k_fast_cells<<<dim3(g.total_cells, n_img), 128, smem, st>>>(dg, pyr, cand, cand_count, flags,
                                                              tile_pitch, tile_bytes, score_bytes);
  count_launch();
}

}  // namespace cuorb
