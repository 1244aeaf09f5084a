// K2 — gridded FAST-9/16 with the iniThFAST/minThFAST retry (replaces the cell loop of
// ORBextractor::ComputeKeyPointsOctTree + cv::FAST(roi, thr, nonmax=true),
// src/ORBextractor.cc:755-850; OpenCV features2d/fast.cpp FAST_t<16>, fast_score.cpp).
//
// One CTA per (cell, image).  The detection areas of the reference's overlapping cell ROIs tile
// the level exactly, so a cell's CTA only needs its own pixels + a 3-px ring halo, and NMS looks
// only at scores inside the cell (OpenCV zeroes everything outside the ROI's detection area).
//
//   phase 1: the cell's pixels (+halo) are staged in shared memory widened to u16, so one 32-bit
//            word holds two horizontally adjacent pixels.
//   phase 2: BRANCH-FREE corner score for every pixel, two pixels per thread in 16x2 SIMD:
//            d[k] = v - ring[k] (+256 bias, so a plain 32-bit subtract never borrows across
//            lanes), then the threshold-independent score
//                max_k min(d[k..k+8])  and  min_k max(d[k..k+8])        (k circular, 16 arcs)
//            as a min/max network of 3-input VIMNMX3.{U16x2} instructions (arcs k and k+1 share
//            8 of their 9 elements, so 8 shared 8-minima + 8 two-element maxima give all 16 arcs:
//            36 SIMD ops per polarity for two pixels).  "corner at t" <=> score >= t, so no
//            per-pixel test, no divergence and no compaction is needed.  Scores below minThFAST
//            are stored as 0.
//   phase 3: strict 8-neighbour NMS inside the cell on the u8 score tile (words that are all
//            zero are skipped); CTA-wide "any survivor >= iniThFAST" picks the threshold;
//            survivors at that threshold are appended to the level's candidate list.
// The list is unordered; the reference order (cell-row-major, then row-major) is a pure function
// of (x, y) and is re-derived where it matters (quadtree tie-break, debug download).
// Bound: the integer ALU pipe (min/max network), not HBM: ~45 ALU-pipe instructions per pixel.
#include "common.cuh"

namespace cuorb {

namespace {

__device__ __forceinline__ unsigned umin2(unsigned a, unsigned b) { return __vminu2(a, b); }
__device__ __forceinline__ unsigned umax2(unsigned a, unsigned b) { return __vmaxu2(a, b); }
__device__ __forceinline__ unsigned umin3(unsigned a, unsigned b, unsigned c) { return __vimin3_u16x2(a, b, c); }
__device__ __forceinline__ unsigned umax3(unsigned a, unsigned b, unsigned c) { return __vimax3_u16x2(a, b, c); }
// two u16 lanes starting at an odd column: high half of `lo` and low half of `hi`
__device__ __forceinline__ unsigned odd_pair(unsigned lo, unsigned hi) { return __byte_perm(lo, hi, 0x5432); }
// exact n / d for n * d < 2^32 with inv = magic(d)
__device__ __forceinline__ unsigned magic(int d) { return 0xFFFFFFFFu / (unsigned)d + 1u; }
__device__ __forceinline__ int fastdiv(int n, unsigned inv, int d) { return d == 1 ? n : (int)__umulhi((unsigned)n, inv); }

// score of the two pixels held in `c` (u16x2) given the 16 ring values r[k] (u16x2 each).
// Returns biased extremes: A = 256 + max_k min9(d), B = 256 + min_k max9(d), d = v - ring.
__device__ __forceinline__ void arc_extremes(unsigned c, const unsigned (&r)[16], unsigned& A, unsigned& B) {
  const unsigned cb = c + 0x01000100u;
  unsigned d[16];
#pragma unroll
  for (int k = 0; k < 16; k++) d[k] = cb - r[k];  // each lane in [1, 511]: no cross-lane borrow
  // pairs starting at odd k: m2[j] <-> d[2j+1], d[2j+2]
  unsigned lo2[8], hi2[8];
#pragma unroll
  for (int j = 0; j < 8; j++) {
    lo2[j] = umin2(d[2 * j + 1], d[(2 * j + 2) & 15]);
    hi2[j] = umax2(d[2 * j + 1], d[(2 * j + 2) & 15]);
  }
  unsigned ta[8], tb[8];
#pragma unroll
  for (int j = 0; j < 8; j++) {
    // arcs 2j and 2j+1 share d[2j+1 .. 2j+8]; they differ in d[2j] vs d[2j+9]
    const unsigned m6 = umin3(lo2[j], lo2[(j + 1) & 7], lo2[(j + 2) & 7]);
    const unsigned x6 = umax3(hi2[j], hi2[(j + 1) & 7], hi2[(j + 2) & 7]);
    const unsigned e_hi = umax2(d[2 * j], d[(2 * j + 9) & 15]);
    const unsigned e_lo = umin2(d[2 * j], d[(2 * j + 9) & 15]);
    ta[j] = umin3(m6, lo2[(j + 3) & 7], e_hi);  // max(min9(arc 2j), min9(arc 2j+1))
    tb[j] = umax3(x6, hi2[(j + 3) & 7], e_lo);  // min(max9(arc 2j), max9(arc 2j+1))
  }
  A = umax3(umax3(ta[0], ta[1], ta[2]), umax3(ta[3], ta[4], ta[5]), umax2(ta[6], ta[7]));
  B = umin3(umin3(tb[0], tb[1], tb[2]), umin3(tb[3], tb[4], tb[5]), umin2(tb[6], tb[7]));
}

}  // namespace

// grid: (total_cells, n_img), block 128, dynamic smem = u16 pixel tile + u8 score tile + lists
// TP: pixel-tile pitch in 32-bit words (two u16 pixels each).  ODD, and lanes walk down rows, so
// the 32 lanes of a warp hit 32 different banks.
template <int TP>
__global__ void __launch_bounds__(128) k_fast_cells(const ExtractorGeom* __restrict__ dg,
                                                    const uint8_t* __restrict__ pyr,
                                                    unsigned* __restrict__ cand,
                                                    int* __restrict__ cand_count,
                                                    int* __restrict__ flags, int tile_bytes,
                                                    int score_bytes, int plist_bytes, int clist_bytes) {
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ int s_np, s_nc, s_nlist, s_nout, s_base;

  // which level / cell
  int level = 0;
  const int nlv = dg->nlevels;
  while (level + 1 < nlv && (int)blockIdx.x >= dg->lv[level + 1].cell_base) level++;
  const LevelGeom& L = dg->lv[level];
  const int cell = blockIdx.x - L.cell_base;
  const int ci = cell / L.nCols, cj = cell - ci * L.nCols;
  const int x0 = EDGE_THRESHOLD + cj * L.wCell, y0 = EDGE_THRESHOLD + ci * L.hCell;
  const int dw = min(x0 + L.wCell, L.maxBorderX - 3) - x0;
  const int dh = min(y0 + L.hCell, L.maxBorderY - 3) - y0;
  if (dw <= 0 || dh <= 0) return;  // cells the reference skips (:781, :791) or that cannot hold a corner

  const int img = blockIdx.y;
  const int tid = threadIdx.x, lane = tid & 31;
  const int ini_th = dg->ini_th, min_th = dg->min_th;

  unsigned*       s_tile  = reinterpret_cast<unsigned*>(smem);      // (dh+6) rows of TP words
  unsigned char*  s_score = smem + tile_bytes;                      // (dh+2) rows of sp bytes, zero frame
  unsigned short* s_plist = reinterpret_cast<unsigned short*>(smem + tile_bytes + score_bytes);
  unsigned short* s_clist = reinterpret_cast<unsigned short*>(smem + tile_bytes + score_bytes + plist_bytes);
  unsigned*       s_list  = reinterpret_cast<unsigned*>(smem + tile_bytes + score_bytes + plist_bytes + clist_bytes);
  // score pitch: pixel px lives at column px + 4; pitch/4 odd so a column of rows spreads over banks
  const int sp = (((dw + 5 + 3) >> 2) | 1) << 2;

  if (tid == 0) { s_np = 0; s_nc = 0; s_nlist = 0; s_nout = 0; }

  // ---- phase 1: stage pixels as u16.  Tile column s <-> level x = ax + s (ax 4-aligned). ----
  const int ax  = (x0 - 4) & ~3;
  const int cx0 = x0 - ax;           // 4..7: tile column of the cell's first pixel
  const int o   = cx0 & 1;           // pairs must start on an even tile column
  const int npr = (dw + o + 1) >> 1; // pixel pairs per row (a leading / trailing pixel may be unused)
  const int gw  = (cx0 - o + 2 * npr + 4 + 3) >> 2;  // aligned 4-px global words per row (2 halo words right)
  const uint8_t* src = pyr + (size_t)img * dg->pyr_img_stride + L.pix_off + (ptrdiff_t)(y0 - 3) * L.pitch + ax;
  {
    const int n = gw * (dh + 6);
    const unsigned inv = magic(gw);
    for (int i = tid; i < n; i += 128) {
      const int r = fastdiv(i, inv, gw), c = i - r * gw;
      const unsigned w = __ldg(reinterpret_cast<const unsigned*>(src + (ptrdiff_t)r * L.pitch + c * 4));
      s_tile[r * TP + 2 * c]     = __byte_perm(w, 0, 0x4140);
      s_tile[r * TP + 2 * c + 1] = __byte_perm(w, 0, 0x4342);
    }
    const int nz = (sp * (dh + 2)) >> 2;
    for (int i = tid; i < nz; i += 128) reinterpret_cast<unsigned*>(s_score)[i] = 0;
  }
  __syncthreads();

  const int npairs = npr * dh;
  const unsigned invh = magic(dh);
  const int cbase = (cx0 - o) >> 1;

  // ---- phase 2a: necessary condition on the 4 compass ring pixels, pairs that pass are listed.
  // An arc of 9 always contains two adjacent compass points, so a corner at minThFAST needs
  // max over adjacent compass pairs of min(d_a, d_b) > t  (or the mirrored condition).
  {
    const unsigned TH = (unsigned)(257 + min_th);
    const unsigned addc = (0x8000u - TH) * 0x00010001u;
    for (int p0 = 0; p0 < npairs; p0 += 128) {
      const int p = p0 + tid;
      bool pass = false;
      if (p < npairs) {
        const int m = fastdiv(p, invh, dh), r = p - m * dh;
        const unsigned* T = s_tile + (r + 3) * TP + cbase + m;
        const unsigned cb = T[0] + 0x01000100u;
        const unsigned dN = cb - T[3 * TP], dS = cb - T[-3 * TP];
        const unsigned dE = cb - odd_pair(T[1], T[2]), dW = cb - odd_pair(T[-2], T[-1]);
        const unsigned qd = umax3(umin2(dN, dE), umin2(dE, dS), umax2(umin2(dS, dW), umin2(dW, dN)));
        const unsigned qb = umin3(umax2(dN, dE), umax2(dE, dS), umin2(umax2(dS, dW), umax2(dW, dN)));
        const unsigned x = umax2(qd, 0x02000200u - qb);
        pass = ((x + addc) & 0x80008000u) != 0;
      }
      const unsigned bal = __ballot_sync(0xffffffffu, pass);
      if (bal) {
        int base = 0;
        if (lane == 0) base = atomicAdd(&s_np, __popc(bal));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (pass) s_plist[base + __popc(bal & ((1u << lane) - 1u))] = (unsigned short)p;
      }
    }
  }
  __syncthreads();

  // ---- phase 2b: full score network on the listed pairs ----
  {
    const int np = s_np;
    for (int i = tid; i < np; i += 128) {
      const int p = s_plist[i];
      const int m = fastdiv(p, invh, dh), r = p - m * dh;
      const int px = 2 * m - o;  // first pixel of the pair, cell-relative (may be -1)
      const unsigned* T = s_tile + (r + 3) * TP + cbase + m;
      unsigned ring[16];
      {
        const unsigned a = T[3 * TP - 1], b = T[3 * TP], c = T[3 * TP + 1];
        ring[0] = b; ring[1] = odd_pair(b, c); ring[15] = odd_pair(a, b);
      }
      ring[2] = T[2 * TP + 1]; ring[14] = T[2 * TP - 1];
      ring[3] = odd_pair(T[TP + 1], T[TP + 2]); ring[13] = odd_pair(T[TP - 2], T[TP - 1]);
      ring[4] = odd_pair(T[1], T[2]); ring[12] = odd_pair(T[-2], T[-1]);
      ring[5] = odd_pair(T[-TP + 1], T[-TP + 2]); ring[11] = odd_pair(T[-TP - 2], T[-TP - 1]);
      ring[6] = T[-2 * TP + 1]; ring[10] = T[-2 * TP - 1];
      {
        const unsigned a = T[-3 * TP - 1], b = T[-3 * TP], c = T[-3 * TP + 1];
        ring[8] = b; ring[7] = odd_pair(b, c); ring[9] = odd_pair(a, b);
      }
      unsigned A, B;
      arc_extremes(T[0], ring, A, B);
      // score = max(a, -b) - 1 with a = A - 256, b = B - 256   (cornerScore<16>)
      const int s0 = max((int)(A & 0xFFFFu) - 257, 255 - (int)(B & 0xFFFFu));
      const int s1 = max((int)(A >> 16) - 257, 255 - (int)(B >> 16));
      unsigned char* q = s_score + (r + 1) * sp + px + 4;
      if (px >= 0 && s0 >= min_th) {
        q[0] = (unsigned char)s0;
        s_clist[atomicAdd(&s_nc, 1)] = (unsigned short)((r << 7) | px);
      }
      if (px + 1 < dw && s1 >= min_th) {
        q[1] = (unsigned char)s1;
        s_clist[atomicAdd(&s_nc, 1)] = (unsigned short)((r << 7) | (px + 1));
      }
    }
  }
  __syncthreads();

  // ---- phase 3: NMS inside the cell over the listed corners, threshold choice, emit ----
  bool any_ini = false;
  {
    const int nc = s_nc;
    for (int i = tid; i < nc; i += 128) {
      const int e = s_clist[i];
      const int r = e >> 7, px = e & 127;
      const unsigned char* q = s_score + (r + 1) * sp + px + 4;
      const int s = q[0];
      const bool lm = s > q[-1] && s > q[1] && s > q[-sp - 1] && s > q[-sp] && s > q[-sp + 1] &&
                      s > q[sp - 1] && s > q[sp] && s > q[sp + 1];
      if (lm) {
        any_ini |= s >= ini_th;
        // region coordinates of vToDistributeKeys: roi pt + (j*wCell, i*hCell)   (:842-846)
        s_list[atomicAdd(&s_nlist, 1)] = pack_cand(x0 + px - FAST_BORDER, y0 + r - FAST_BORDER, s);
      }
    }
  }
  const int thr = __syncthreads_or(any_ini) ? ini_th : min_th;
  const int nl = s_nlist;
  if (nl == 0) return;
  unsigned* s_out = s_tile;  // pixel tile is dead: reuse it for the compacted survivors
  for (int n = tid; n < nl; n += 128) {
    const unsigned c = s_list[n];
    if (cand_s(c) >= thr) s_out[atomicAdd(&s_nout, 1)] = c;
  }
  __syncthreads();
  const int nout = s_nout;
  if (nout == 0) return;
  const int slot = img * nlv + level;
  if (tid == 0) s_base = atomicAdd(&cand_count[slot], nout);
  __syncthreads();
  const int base = s_base;
  if (base + nout > L.cand_cap) {  // cannot happen with the worst-case capacity; flagged anyway
    if (tid == 0) atomicOr(flags, FLAG_CAND_OVERFLOW);
    return;
  }
  unsigned* out = cand + (size_t)img * dg->cand_img_stride + L.cand_off + base;
  for (int n = tid; n < nout; n += 128) out[n] = s_out[n];
}

template <int TP>
static void launch_fast_tp(const ExtractorGeom& g, const ExtractorGeom* dg, int n_img, const uint8_t* pyr,
                           unsigned* cand, int* cand_count, int* flags, cudaStream_t st) {
  const int tile_bytes  = ((TP * 4 * (g.max_cell_h + 6) + 15) / 16) * 16;
  const int sp_max      = ((((g.max_cell_w + 5 + 3) >> 2) | 1) << 2);
  const int score_bytes = ((sp_max * (g.max_cell_h + 2) + 15) / 16) * 16;
  const int plist_bytes = ((2 * ((g.max_cell_w + 2) / 2) * g.max_cell_h + 15) / 16) * 16;
  const int clist_bytes = ((2 * g.max_cell_w * g.max_cell_h + 15) / 16) * 16;
  const int list_bytes  = 4 * ((g.max_cell_w + 1) / 2) * ((g.max_cell_h + 1) / 2);
  const int smem        = tile_bytes + score_bytes + plist_bytes + clist_bytes + list_bytes;
  static int attr = 48 * 1024;
  if (smem > attr) {
    cudaFuncSetAttribute(k_fast_cells<TP>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    attr = smem;
  }
  k_fast_cells<TP><<<dim3(g.total_cells, n_img), 128, smem, st>>>(dg, pyr, cand, cand_count, flags, tile_bytes,
                                                                  score_bytes, plist_bytes, clist_bytes);
}

void launch_fast(const ExtractorGeom& g, const ExtractorGeom* dg, int n_img, const uint8_t* pyr,
                 unsigned* cand, int* cand_count, int* flags, cudaStream_t st) {
  // tile columns 0 .. (6 + max_cell_w + 2 + 4 + alignment) as u16 pairs -> words; cell height
  // below 512 and width below 128 for the packed corner list
  if (g.max_cell_w + 16 <= 2 * 29) launch_fast_tp<29>(g, dg, n_img, pyr, cand, cand_count, flags, st);
  else launch_fast_tp<71>(g, dg, n_img, pyr, cand, cand_count, flags, st);
  count_launch();
}

}  // namespace cuorb
