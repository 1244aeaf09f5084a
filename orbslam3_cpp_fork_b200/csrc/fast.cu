// K2 — gridded FAST-9/16 with the iniThFAST/minThFAST retry (replaces the cell loop of
// ORBextractor::ComputeKeyPointsOctTree + cv::FAST(roi, thr, nonmax=true),
// src/ORBextractor.cc:755-850; OpenCV features2d/fast.cpp FAST_t<16>, fast_score.cpp).
//
// The detection areas of the reference's overlapping cell ROIs tile the level exactly, so a cell
// only needs its own pixels + a 3-px ring halo, and NMS looks only at scores inside the cell
// (OpenCV zeroes everything outside the ROI's detection area).  Per cell, in shared memory:
//   2a  necessary condition on the 4 compass ring pixels, four pixels per thread in u8x4 SIMD
//       (VABSDIFF4); pixel pairs with a passing pixel are compacted into per-warp lists with
//       warp ballots (no atomics);
//   2b  BRANCH-FREE corner score on the listed pairs, two pixels per thread in 16x2 SIMD, on the raw ring
//       values: M = min_k max(ring[k..k+8]) and m = max_k min(ring[k..k+8]) (k circular, 16 arcs) as a
//       network of 3-input VIMNMX3.U16x2 (arcs k and k+1 share 8 of 9 elements: 36 SIMD ops per polarity for
//       two pixels), then score = max(v - M, m - v) - 1: the differences are taken once at the end, not per
//       ring element; "corner at t" <=> score >= t; corners go to a second list;
//   3   strict 8-neighbour NMS over the listed corners on a u8 score tile;
//   the whole thing first at iniThFAST and, only if the cell stays empty, again at minThFAST
//   (:798-826), decided CTA-wide with __syncthreads_or.
// Two kernels share these phases:
//   k_fast_persist (cells <= 54 px wide, the normal case): persistent CTAs pull (image, cell) items
//       from a global counter; the next item's u8 tile is fetched by TMA into the other half of a
//       double buffer while the current one is processed; per-cell scalars come from a host-built
//       record; bytes are widened to u16x2 with PRMT as the network reads them.
//   k_fast_cells<TP,TPB> (wider cells): one CTA per cell, tile staged with aligned 32-bit loads as a
//       u8 and a u16 copy.
// The candidate list is unordered; the reference order (cell-row-major, then row-major) is a pure
// function of (x, y) and is re-derived where it matters (quadtree tie-break, debug download).
// Bound: the integer ALU pipe (VIMNMX/PRMT/LOP3 all issue there at 2 warp-inst/clk/SM), not HBM.
#include <algorithm>
#include <cstdlib>

#include <mutex>

#include "common.cuh"
#include "tma.cuh"

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

// Arc extremes of the two pixels of a pair, on the RAW ring values r[k] (u16x2 each, lanes in 0..255):
//   M = min over the 16 arcs of 9 of max9(r),   m = max over the 16 arcs of min9(r).
// With d[k] = v - r[k]:  max_k min9(d) = v - M  and  min_k max9(d) = v - m, so the corner score
//   max(max_k min9(d), -min_k max9(d)) - 1 = max(v - M, m - v) - 1
// needs one subtraction per polarity at the end instead of sixteen up front (the min / max network is the same
// size: arcs 2j and 2j+1 share r[2j+1 .. 2j+8] and differ in r[2j] vs r[2j+9]).
__device__ __forceinline__ void arc_extremes(const unsigned (&r)[16], unsigned& M, unsigned& m) {
  // pairs starting at odd k: lo2[j], hi2[j] <-> r[2j+1], r[2j+2]
  unsigned lo2[8], hi2[8];
#pragma unroll
  for (int j = 0; j < 8; j++) {
    lo2[j] = umin2(r[2 * j + 1], r[(2 * j + 2) & 15]);
    hi2[j] = umax2(r[2 * j + 1], r[(2 * j + 2) & 15]);
  }
  unsigned tM[8], tm[8];
#pragma unroll
  for (int j = 0; j < 8; j++) {
    const unsigned x6 = umax3(hi2[j], hi2[(j + 1) & 7], hi2[(j + 2) & 7]);
    const unsigned m6 = umin3(lo2[j], lo2[(j + 1) & 7], lo2[(j + 2) & 7]);
    const unsigned e_lo = umin2(r[2 * j], r[(2 * j + 9) & 15]);
    const unsigned e_hi = umax2(r[2 * j], r[(2 * j + 9) & 15]);
    tM[j] = umax3(x6, hi2[(j + 3) & 7], e_lo);  // min(max9(arc 2j), max9(arc 2j+1))
    tm[j] = umin3(m6, lo2[(j + 3) & 7], e_hi);  // max(min9(arc 2j), min9(arc 2j+1))
  }
  M = umin3(umin3(tM[0], tM[1], tM[2]), umin3(tM[3], tM[4], tM[5]), umin2(tM[6], tM[7]));
  m = umax3(umax3(tm[0], tm[1], tm[2]), umax3(tm[3], tm[4], tm[5]), umax2(tm[6], tm[7]));
}
// score of one u16 lane: max(v - M, m - v) - 1
__device__ __forceinline__ int lane_score(int v, int M, int m) { return max(v - M, m - v) - 1; }

}  // namespace

// warp-uniform list append: `cnt` is a warp-uniform register, `seg` the warp's own segment
__device__ __forceinline__ void warp_append(bool f, unsigned short v, unsigned short* seg, int& cnt, unsigned lt) {
  const unsigned b = __ballot_sync(0xffffffffu, f);
  if (f) seg[cnt + __popc(b & lt)] = v;
  cnt += __popc(b);
}

// flat index over four per-warp segments -> entry
__device__ __forceinline__ int seg_fetch(const unsigned short* list, int cap, const int (&c)[4], int i) {
  const int w = (i >= c[0]) + (i >= c[0] + c[1]) + (i >= c[0] + c[1] + c[2]);
  const int off = (w > 0 ? c[0] : 0) + (w > 1 ? c[1] : 0) + (w > 2 ? c[2] : 0);
  return list[w * cap + (i - off)];
}

// grid: (total_cells, n_img), block 128, dynamic smem = u16 + u8 pixel tiles, u8 score tile, lists
// TP:  pitch of the u16 pixel tile in 32-bit words (two pixels each).
// TPB: pitch of the u8 pixel tile in 32-bit words (four pixels each).
// cells[]: per FAST cell {level, x0 | y0 << 16, dw | dh << 16} (host-built: no per-CTA geometry math)
template <int TP, int TPB>
__global__ void __launch_bounds__(128) k_fast_cells(const ExtractorGeom* __restrict__ dg,
                                                    const int4* __restrict__ cells,
                                                    const uint8_t* __restrict__ pyr,
                                                    unsigned* __restrict__ cand,
                                                    int* __restrict__ cand_count,
                                                    int* __restrict__ flags, int tile_bytes, int tile8_bytes,
                                                    int score_bytes, int pcap, int ccap) {
  pdl_enter();
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ int s_pcnt[4], s_ccnt[4], s_nlist, s_nout, s_base;

  const int4 cd = __ldg(&cells[blockIdx.x]);
  const int level = cd.x;
  const int x0 = cd.y & 0xFFFF, y0 = cd.y >> 16, dw = cd.z & 0xFFFF, dh = cd.z >> 16;
  if (dw <= 0 || dh <= 0) return;  // cells the reference skips (:781, :791) or that cannot hold a corner
  const LevelGeom& L = dg->lv[level];

  const int img = blockIdx.y;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const unsigned lt = (1u << lane) - 1u;
  const int ini_th = dg->ini_th, min_th = dg->min_th;

  unsigned*       s_tile  = reinterpret_cast<unsigned*>(smem);                  // (dh+6) rows of TP words (u16 px)
  unsigned*       s_tile8 = reinterpret_cast<unsigned*>(smem + tile_bytes);     // (dh+6) rows of TPB words (u8 px)
  unsigned char*  s_score = smem + tile_bytes + tile8_bytes;                    // (dh+2) rows of sp bytes, zero frame
  unsigned short* s_plist = reinterpret_cast<unsigned short*>(s_score + score_bytes);  // 4 segments of pcap
  unsigned short* s_clist = s_plist + 4 * pcap;                                         // 4 segments of ccap
  unsigned*       s_list  = reinterpret_cast<unsigned*>(s_clist + 4 * ccap);
  // score pitch: pixel px lives at column px + 4; pitch/4 odd so a column of rows spreads over banks
  const int sp = (((dw + 5 + 3) >> 2) | 1) << 2;

  if (tid == 0) { s_nlist = 0; s_nout = 0; }

  // ---- phase 1: stage pixels.  Tile column s <-> level x = ax + s (ax 4-aligned); the cell's
  // first pixel sits at column cx0 = 4..7, so the ring (3 px) and one more column fit on the left.
  const int ax  = (x0 - 4) & ~3;
  const int cx0 = x0 - ax;
  const int gw  = (cx0 + dw + 3 + 4 + 3) >> 2;  // aligned 4-px words per row: ring + one spare pixel right
  {
    const uint8_t* src = pyr + (size_t)img * dg->pyr_img_stride + L.pix_off + (ptrdiff_t)(y0 - 3) * L.pitch + ax;
    const int pitch = L.pitch;
    for (int c = tid & 15; c < gw; c += 16) {
      const uint8_t* p = src + (ptrdiff_t)(tid >> 4) * pitch + c * 4;
      for (int r = tid >> 4; r < dh + 6; r += 8, p += 8 * (ptrdiff_t)pitch) {
        const unsigned w = __ldg(reinterpret_cast<const unsigned*>(p));
        s_tile8[r * TPB + c] = w;
        *reinterpret_cast<uint2*>(s_tile + r * TP + 2 * c) = make_uint2(__byte_perm(w, 0, 0x4140), __byte_perm(w, 0, 0x4342));
      }
    }
    const int nz = (sp * (dh + 2)) >> 2;
    for (int i = tid; i < nz; i += 128) reinterpret_cast<unsigned*>(s_score)[i] = 0;
  }
  __syncthreads();

  // The reference runs FAST at iniThFAST and, only if the cell yields no keypoint, again at
  // minThFAST (:798-826).  Same here: pass 0 scores only what can reach iniThFAST (its NMS sees
  // exactly the corners FAST(ini) would see); pass 1 repeats everything at minThFAST for the
  // (rare on textured images) cells that came up empty.
  int thr = ini_th;
  for (int pass = 0; pass < 2; pass++) {
    const int cur_th = pass == 0 ? ini_th : min_th;
    thr = cur_th;
    // ---- phase 2a: necessary condition on the 4 compass ring pixels, four pixels per thread on the
    // u8 tile.  An arc of 9 always contains two adjacent compass points, so a corner at minThFAST
    // at threshold t needs (|N-c| > t or |S-c| > t) and (|E-c| > t or |W-c| > t).  Pixel pairs with a passing pixel
    // are appended to the warp's own list segment as (row << 7 | pair column).
    {
      const int nqm = ((cx0 + dw + 3) >> 2) - 1;  // quads = tile words 1 .. nqm (columns 4 .. 4*nqm+3)
      const int nquads = nqm * dh;
      const unsigned inv = magic(nqm);
      const unsigned K = (unsigned)(127 - min(cur_th, 127)) * 0x01010101u;
      unsigned short* seg = s_plist + wid * pcap;
      int cnt = 0;
      for (int i0 = 0; i0 < nquads; i0 += 128) {
        const int i = i0 + tid;
        bool f0 = false, f1 = false;
        int r = 0, q = 0;
        if (i < nquads) {
          r = fastdiv(i, inv, nqm); q = i - r * nqm + 1;
          const unsigned* T = s_tile8 + (r + 3) * TPB + q;
          const unsigned C = T[0];
          const unsigned E = __funnelshift_r(C, T[1], 24), W = __funnelshift_r(T[-1], C, 8);
          const unsigned aN = __vabsdiffu4(T[3 * TPB], C), aS = __vabsdiffu4(T[-3 * TPB], C);
          const unsigned aE = __vabsdiffu4(E, C), aW = __vabsdiffu4(W, C);
          // per byte "a > t": bit 7 of a | ((a & 0x7f) + (127 - t))
          const unsigned mN = aN | ((aN & 0x7f7f7f7fu) + K), mS = aS | ((aS & 0x7f7f7f7fu) + K);
          const unsigned mE = aE | ((aE & 0x7f7f7f7fu) + K), mW = aW | ((aW & 0x7f7f7f7fu) + K);
          const unsigned pass = (mN | mS) & (mE | mW) & 0x80808080u;
          f0 = (pass & 0x0000FFFFu) != 0; f1 = (pass & 0xFFFF0000u) != 0;
        }
        const unsigned short e = (unsigned short)((r << 7) | (2 * q - 2));   // pair column = (s - 4) / 2
        warp_append(f0, e, seg, cnt, lt);
        warp_append(f1, (unsigned short)(e + 1), seg, cnt, lt);
      }
      if (lane == 0) s_pcnt[wid] = cnt;
    }
    __syncthreads();

    // ---- phase 2b: full score network on the listed pairs; corners go to the warp's corner list ----
    {
      const int pc4[4] = {s_pcnt[0], s_pcnt[1], s_pcnt[2], s_pcnt[3]};
      const int np = pc4[0] + pc4[1] + pc4[2] + pc4[3];
      unsigned short* seg = s_clist + wid * ccap;
      int cnt = 0;
      for (int i0 = 0; i0 < np; i0 += 128) {
        const int i = i0 + tid;
        bool c0 = false, c1 = false;
        int r = 0, px = 0;
        if (i < np) {
          const int e = seg_fetch(s_plist, pcap, pc4, i);
          r = e >> 7;
          const int pc = e & 127;
          px = 4 + 2 * pc - cx0;  // first pixel of the pair, cell-relative (may be outside the cell)
          const unsigned* T = s_tile + (r + 3) * TP + 2 + pc;
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
          unsigned M, m;
          arc_extremes(ring, M, m);
          // score = max(v - M, m - v) - 1   (cornerScore<16>)
          const unsigned ctr = T[0];
          const int s0 = lane_score((int)(ctr & 0xFFFFu), (int)(M & 0xFFFFu), (int)(m & 0xFFFFu));
          const int s1 = lane_score((int)(ctr >> 16), (int)(M >> 16), (int)(m >> 16));
          unsigned char* q = s_score + (r + 1) * sp + px + 4;
          c0 = px >= 0 && px < dw && s0 >= cur_th;
          c1 = px + 1 >= 0 && px + 1 < dw && s1 >= cur_th;
          if (c0) q[0] = (unsigned char)s0;
          if (c1) q[1] = (unsigned char)s1;
        }
        warp_append(c0, (unsigned short)((r << 7) | px), seg, cnt, lt);
        warp_append(c1, (unsigned short)((r << 7) | (px + 1)), seg, cnt, lt);
      }
      if (lane == 0) s_ccnt[wid] = cnt;
    }
    __syncthreads();

    // ---- phase 3: NMS inside the cell over the listed corners, threshold choice, emit ----
    {
      const int cc4[4] = {s_ccnt[0], s_ccnt[1], s_ccnt[2], s_ccnt[3]};
      const int nc = cc4[0] + cc4[1] + cc4[2] + cc4[3];
      for (int i = tid; i < nc; i += 128) {
        const int e = seg_fetch(s_clist, ccap, cc4, i);
        const int r = e >> 7, px = e & 127;
        const unsigned char* q = s_score + (r + 1) * sp + px + 4;
        const int s = q[0];
        const bool lm = s > q[-1] && s > q[1] && s > q[-sp - 1] && s > q[-sp] && s > q[-sp + 1] &&
                        s > q[sp - 1] && s > q[sp] && s > q[sp + 1];
        if (lm) {
          // region coordinates of vToDistributeKeys: roi pt + (j*wCell, i*hCell)   (:842-846)
          s_list[atomicAdd(&s_nlist, 1)] = pack_cand(x0 + px - FAST_BORDER, y0 + r - FAST_BORDER, s);
        }
      }
    }
    if (__syncthreads_or(s_nlist > 0) || pass == 1 || min_th == ini_th) break;
    // empty at iniThFAST: clear the score tile and retry at minThFAST
    {
      const int nz = (sp * (dh + 2)) >> 2;
      for (int i = tid; i < nz; i += 128) reinterpret_cast<unsigned*>(s_score)[i] = 0;
    }
    __syncthreads();
  }
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
  const int slot = img * dg->nlevels + level;
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

// ------------------------------------------------------------------------------------------------
// Persistent variant (cells up to 54 px wide): CTAs loop over (image, cell) work items; the u8 pixel
// tile of the NEXT item is fetched by TMA (cp.async.bulk.tensor.3d, mbarrier completion) into the
// other half of a double buffer while the current item is processed, so no thread spends an
// instruction on the tile load and its latency is hidden behind compute.  Per-cell scalars come
// pre-digested from a host-built record; there is no u16 copy of the tile (the score network
// widens byte pairs with PRMT as it reads them); the score tile is zeroed once and then only the
// entries a cell wrote are cleared again.
//   rec[2c]   = {level | cx0 << 8 | qf << 16 | nq << 24,  x0 | y0 << 16,  dw | dh << 16,  magic(nq)}
//   rec[2c+1] = {plane x of tile column 0 (16-aligned) | plane row of tile row 0 << 16,  sp,  cand_off,  cand_cap}
struct FastParams {
  const int4* recs; int n_cells; int total_work;
  float inv_cells;   // 1 / n_cells (work item -> image, cell without an integer division)
  unsigned* cand; int* cand_count; int* flags; int* work_counter;
  unsigned long long cand_img_stride;
  int nlevels, ini_th, min_th;
  int box_rows, tile_bytes, score_bytes, pcap;
};

constexpr int FT_WORDS = FT_PITCH / 4;

__device__ __forceinline__ unsigned widen_even(unsigned h) { return __byte_perm(h, 0, 0x4140); }
// bytes 1 of lo and 0 of hi (two zero-extended u16 loads two bytes apart) -> u16x2
__device__ __forceinline__ unsigned widen_odd(unsigned lo, unsigned hi) { return __byte_perm(lo, hi, 0x2421); }

// work item -> (image, cell): float reciprocal + one correction step each way (the quotient is an image
// index < 65536, so the float estimate is off by at most one)
__device__ __forceinline__ void split_work(int w, int n_cells, float inv, int& img, int& cell) {
  int q = __float2int_rz(__fmul_rz(__int2float_rz(w), inv));
  int r = w - q * n_cells;
  if (r < 0) { q--; r += n_cells; }
  if (r >= n_cells) { q++; r -= n_cells; }
  img = q; cell = r;
}

template <int NW>
__global__ void __launch_bounds__(32 * NW) k_fast_persist(const __grid_constant__ FastMaps maps, const FastParams P) {
  pdl_enter();
  extern __shared__ unsigned char smem_raw[];
  __shared__ uint64_t s_bar[2];
  __shared__ int s_next[2], s_img[2], s_cell[2];

  // pointer arithmetic on the array (not an integer round trip) so the accesses stay LDS/STS, not generic LD/ST
  unsigned char* smem = smem_raw + ((128u - ((unsigned)__cvta_generic_to_shared(smem_raw) & 127u)) & 127u);
  unsigned char*  s_tiles = smem;                                              // 2 x tile_bytes
  unsigned char*  s_scores = smem + 2 * P.tile_bytes;                           // 2 x score_bytes (items alternate)
  unsigned short* s_plist = reinterpret_cast<unsigned short*>(s_scores + 2 * P.score_bytes);
  constexpr int NT = 32 * NW;   // NW warps share one cell

  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const unsigned lt = (1u << lane) - 1u;
  const int pcap = P.pcap;
  int w = blockIdx.x;
  if (w >= P.total_work) return;

  if (tid == 0) {
    mbar_init(&s_bar[0], 1); mbar_init(&s_bar[1], 1);
    mbar_fence_init();
  }
  for (int i = tid; i < (P.score_bytes >> 1); i += NT) reinterpret_cast<unsigned*>(s_scores)[i] = 0;
  __syncthreads();

  auto issue = [&](int img, int cell, int buf) {  // one thread
    const int lvl = __ldg(&P.recs[2 * cell].x) & 0xFF;
    const int org = __ldg(&P.recs[2 * cell + 1].x);
    mbar_expect_tx(&s_bar[buf], (unsigned)(FT_PITCH * P.box_rows));
    tma_load_3d(s_tiles + buf * P.tile_bytes, &maps.lv[lvl], &s_bar[buf], org & 0xFFFF, org >> 16, img);
  };
  int img, cell;
  split_work(w, P.n_cells, P.inv_cells, img, cell);
  if (tid == 0) issue(img, cell, 0);

  // items beyond the first come from a global counter (dynamic balance: cell cost varies 3x)
  for (int it = 0;; it++) {
    const int buf = it & 1;
    if (tid == 0) {
      const int nw = (int)gridDim.x + atomicAdd(P.work_counter, 1);
      s_next[buf ^ 1] = nw;
      if (nw < P.total_work) {
        int ni, nc;
        split_work(nw, P.n_cells, P.inv_cells, ni, nc);
        s_img[buf ^ 1] = ni; s_cell[buf ^ 1] = nc;
        fence_proxy_async();  // the other buffer was read/written through the generic proxy one item ago
        issue(ni, nc, buf ^ 1);
      }
    }
    const int4 c0 = __ldg(&P.recs[2 * cell]), c1 = __ldg(&P.recs[2 * cell + 1]);
    const int level = c0.x & 0xFF, cx0 = (c0.x >> 8) & 0xFF, qf = (c0.x >> 16) & 0xFF, nq = (c0.x >> 24) & 0xFF;
    const int x0 = c0.y & 0xFFFF, y0 = c0.y >> 16, dw = c0.z & 0xFFFF, dh = c0.z >> 16;
    const unsigned inv = (unsigned)c0.w;
    const int sp = c1.y;
    const unsigned char* tile = s_tiles + buf * P.tile_bytes;
    // score tile of this item; the other one is still being cleared by slower warps of the previous item
    unsigned char* s_score = s_scores + buf * P.score_bytes;
    const unsigned* s_tile8 = reinterpret_cast<const unsigned*>(tile);
    mbar_wait(&s_bar[buf], (unsigned)((it >> 1) & 1));

    for (int pass = 0; pass < 2; pass++) {
      const int cur_th = pass == 0 ? P.ini_th : P.min_th;
      // ---- 2a: compass test, four pixels per thread, passing pairs -> per-warp list ----
      int npairs;
      {
        const int nquads = nq * dh;
        const unsigned K = (unsigned)(127 - min(cur_th, 127)) * 0x01010101u;
        unsigned short* seg = s_plist + wid * pcap;
        int cnt = 0;
        for (int i0 = 0; i0 < nquads; i0 += NT) {
          // no branch around the body: out-of-range threads redo the last quad and are masked out
          const bool in = i0 + tid < nquads;
          const int i = min(i0 + tid, nquads - 1);
          const int r = (int)__umulhi((unsigned)i, inv), q = i - r * nq + qf;   // nq >= 2 (host), so the magic is exact
          const unsigned* T = s_tile8 + (r + 3) * FT_WORDS + q;
          const unsigned C = T[0];
          const unsigned E = __funnelshift_r(C, T[1], 24), W = __funnelshift_r(T[-1], C, 8);
          const unsigned aN = __vabsdiffu4(T[3 * FT_WORDS], C), aS = __vabsdiffu4(T[-3 * FT_WORDS], C);
          const unsigned aE = __vabsdiffu4(E, C), aW = __vabsdiffu4(W, C);
          const unsigned mN = aN | ((aN & 0x7f7f7f7fu) + K), mS = aS | ((aS & 0x7f7f7f7fu) + K);
          const unsigned mE = aE | ((aE & 0x7f7f7f7fu) + K), mW = aW | ((aW & 0x7f7f7f7fu) + K);
          const unsigned ok = (mN | mS) & (mE | mW) & 0x80808080u;
          const bool f0 = in && (ok & 0x0000FFFFu) != 0, f1 = in && (ok & 0xFFFF0000u) != 0;
          const unsigned short e = (unsigned short)((r << 7) | (2 * q - 2));   // pair column = (tile column - 4) / 2
          warp_append(f0, e, seg, cnt, lt);
          warp_append(f1, (unsigned short)(e + 1), seg, cnt, lt);
        }
        npairs = cnt;
      }
      __syncwarp();   // each warp consumes only the list it wrote: no CTA barrier between 2a and 2b

      // ---- 2b: score network on the warp's own pairs, bytes widened on the fly ----
      int ncorn = 0;
      {
        // corner entries (row << 8 | pair column << 2 | pixel flags) overwrite the pair list in place: a chunk of 32
        // pairs yields at most 32 entries, written at or below the chunk's own positions, after the chunk was read
        unsigned short* pl = s_plist + wid * pcap;
        for (int j0 = 0; j0 < npairs; j0 += 32) {
          const bool in = j0 + lane < npairs;
          const int e = pl[min(j0 + lane, npairs - 1)];
          const int r = e >> 7, pc = e & 127;
          const int px = 4 + 2 * pc - cx0;
          const unsigned char* B = tile + (r + 3) * FT_PITCH + 4 + 2 * pc;
#define H16(dy, dx) ((unsigned)*reinterpret_cast<const unsigned short*>(B + (dy) * FT_PITCH + (dx)))
          unsigned ring[16];
          {
            const unsigned a = H16(3, -2), b = H16(3, 0), c = H16(3, 2);
            ring[15] = widen_odd(a, b); ring[0] = widen_even(b); ring[1] = widen_odd(b, c);
          }
          ring[2] = widen_even(H16(2, 2)); ring[14] = widen_even(H16(2, -2));
          ring[3] = widen_odd(H16(1, 2), H16(1, 4)); ring[13] = widen_odd(H16(1, -4), H16(1, -2));
          unsigned ctr;
          {
            const unsigned m4 = H16(0, -4), m2 = H16(0, -2), c = H16(0, 0), p2 = H16(0, 2), p4 = H16(0, 4);
            ring[4] = widen_odd(p2, p4); ring[12] = widen_odd(m4, m2); ctr = widen_even(c);
          }
          ring[5] = widen_odd(H16(-1, 2), H16(-1, 4)); ring[11] = widen_odd(H16(-1, -4), H16(-1, -2));
          ring[6] = widen_even(H16(-2, 2)); ring[10] = widen_even(H16(-2, -2));
          {
            const unsigned a = H16(-3, -2), b = H16(-3, 0), c = H16(-3, 2);
            ring[9] = widen_odd(a, b); ring[8] = widen_even(b); ring[7] = widen_odd(b, c);
          }
#undef H16
          unsigned M, m;
          arc_extremes(ring, M, m);
          const int s0 = lane_score((int)(ctr & 0xFFFFu), (int)(M & 0xFFFFu), (int)(m & 0xFFFFu));
          const int s1 = lane_score((int)(ctr >> 16), (int)(M >> 16), (int)(m >> 16));
          unsigned char* q = s_score + (r + 1) * sp + px + 4;
          const bool k0 = in && px >= 0 && px < dw && s0 >= cur_th;
          const bool k1 = in && px + 1 >= 0 && px + 1 < dw && s1 >= cur_th;
          if (k0) q[0] = (unsigned char)s0;
          if (k1) q[1] = (unsigned char)s1;
          __syncwarp();   // every lane holds its pair entry before the list is overwritten
          warp_append(k0 || k1, (unsigned short)((r << 8) | (pc << 2) | (k1 ? 2 : 0) | (k0 ? 1 : 0)), pl, ncorn, lt);
        }
      }
      __syncthreads();   // all scores of the cell are in place

      // ---- 3: NMS over the warp's own corners; survivors go straight to the level's candidate list
      // (one global atomicAdd per warp and 32 corners): whatever a pass finds is final, because a
      // pass-0 survivor means the cell keeps its iniThFAST result and pass 1 only runs on empty cells
      const unsigned short* cl = s_plist + wid * pcap;
      int found = 0;
      for (int j0 = 0; j0 < ncorn; j0 += 32) {
        const bool in = j0 + lane < ncorn;
        const int e = cl[min(j0 + lane, ncorn - 1)];
        const int r = e >> 8, px = 4 + 2 * ((e >> 2) & 63) - cx0;
        const unsigned char* q = s_score + (r + 1) * sp + px + 4;
        // the two pixels of the pair share the 3 x 4 score neighbourhood
        const int a0 = q[-sp - 1], a1 = q[-sp], a2 = q[-sp + 1], a3 = q[-sp + 2];
        const int m0 = q[-1], s0 = q[0], s1 = q[1], m3 = q[2];
        const int b0 = q[sp - 1], b1 = q[sp], b2 = q[sp + 1], b3 = q[sp + 2];
        const bool lm0 = in && (e & 1) && s0 > m0 && s0 > s1 && s0 > a0 && s0 > a1 && s0 > a2 && s0 > b0 && s0 > b1 && s0 > b2;
        const bool lm1 = in && (e & 2) && s1 > s0 && s1 > m3 && s1 > a1 && s1 > a2 && s1 > a3 && s1 > b1 && s1 > b2 && s1 > b3;
        const unsigned bal0 = __ballot_sync(0xffffffffu, lm0), bal1 = __ballot_sync(0xffffffffu, lm1);
        if (bal0 | bal1) {
          // output addresses only here (the common iteration emits nothing)
          unsigned* out = P.cand + (size_t)img * P.cand_img_stride + (unsigned)c1.z;
          int* out_count = P.cand_count + img * P.nlevels + level;
          const int n0 = __popc(bal0), n = n0 + __popc(bal1);
          int base = 0;
          if (lane == 0) base = atomicAdd(out_count, n);
          base = __shfl_sync(0xffffffffu, base, 0);
          if (base + n > c1.w) { if (lane == 0) atomicOr(P.flags, FLAG_CAND_OVERFLOW); }
          else {
            if (lm0) out[base + __popc(bal0 & lt)] = pack_cand(x0 + px - FAST_BORDER, y0 + r - FAST_BORDER, s0);
            if (lm1) out[base + n0 + __popc(bal1 & lt)] = pack_cand(x0 + px + 1 - FAST_BORDER, y0 + r - FAST_BORDER, s1);
          }
          found += n;
        }
      }
      const bool done = __syncthreads_or(found > 0) || pass == 1 || P.min_th == P.ini_th;
      // clear the cell's score tile with 16-byte stores (all NMS reads are behind the barrier above; the tile is
      // next used two items later, or after the barrier below in a retry): ~2 stores per thread, cheaper than
      // walking the corner list again
      {
        const int n16 = (sp * (dh + 2) + 15) >> 4;
        uint4* z = reinterpret_cast<uint4*>(s_score);
        for (int j = tid; j < n16; j += NT) z[j] = make_uint4(0u, 0u, 0u, 0u);
      }
      if (done) break;
      __syncthreads();
    }

    w = s_next[buf ^ 1];
    if (w >= P.total_work) break;
    img = s_img[buf ^ 1]; cell = s_cell[buf ^ 1];
  }
}

// work records for k_fast_persist (cells the reference skips are left out); returns the cell count
int build_fast_recs(const ExtractorGeom& g, std::vector<int4>& recs) {
  recs.clear();
  for (int l = 0; l < g.nlevels; l++) {
    const LevelGeom& L = g.lv[l];
    for (int ci = 0; ci < L.nRows; ci++)
      for (int cj = 0; cj < L.nCols; cj++) {
        const int x0 = EDGE_THRESHOLD + cj * L.wCell, y0 = EDGE_THRESHOLD + ci * L.hCell;
        const int dw = std::min(x0 + L.wCell, L.maxBorderX - 3) - x0;
        const int dh = std::min(y0 + L.hCell, L.maxBorderY - 3) - y0;
        if (dw <= 0 || dh <= 0) continue;
        const int pxo = (ROW_LEAD + x0 - 4) & ~15;            // plane column of tile column 0
        const int cx0 = ROW_LEAD + x0 - pxo;                  // 4..19
        const int qf = cx0 >> 2, ql = (cx0 + dw - 1) >> 2;
        const int nq = std::max(ql - qf + 1, 2);   // >= 2 keeps the division magic exact; a surplus quad only holds masked pixels
        const int sp = (((dw + 5 + 3) >> 2) | 1) << 2;
        const unsigned inv = 0xFFFFFFFFu / (unsigned)nq + 1u;
        recs.push_back(make_int4(l | (cx0 << 8) | (qf << 16) | (nq << 24), x0 | (y0 << 16), dw | (dh << 16), (int)inv));
        recs.push_back(make_int4(pxo | ((EDGE_THRESHOLD + y0 - 3) << 16), sp, (int)L.cand_off, L.cand_cap));
      }
  }
  return (int)recs.size() / 2;
}

// per-cell geometry (src/ORBextractor.cc:777-796), cell-row-major per level
void build_fast_cells(const ExtractorGeom& g, std::vector<int4>& cells) {
  cells.clear();
  for (int l = 0; l < g.nlevels; l++) {
    const LevelGeom& L = g.lv[l];
    for (int ci = 0; ci < L.nRows; ci++)
      for (int cj = 0; cj < L.nCols; cj++) {
        const int x0 = EDGE_THRESHOLD + cj * L.wCell, y0 = EDGE_THRESHOLD + ci * L.hCell;
        const int dw = std::max(std::min(x0 + L.wCell, L.maxBorderX - 3) - x0, 0);
        const int dh = std::max(std::min(y0 + L.hCell, L.maxBorderY - 3) - y0, 0);
        cells.push_back(make_int4(l, x0 | (y0 << 16), dw | (dh << 16), 0));
      }
  }
}

template <int TP, int TPB>
static void launch_fast_tp(const ExtractorGeom& g, const ExtractorGeom* dg, const int4* cells, int n_img,
                           const uint8_t* pyr, unsigned* cand, int* cand_count, int* flags, cudaStream_t st) {
  const int rows        = g.max_cell_h + 6;
  const int tile_bytes  = ((TP * 4 * rows + 15) / 16) * 16;
  const int tile8_bytes = ((TPB * 4 * rows + 15) / 16) * 16;
  const int sp_max      = ((((g.max_cell_w + 5 + 3) >> 2) | 1) << 2);
  const int score_bytes = ((sp_max * (g.max_cell_h + 2) + 15) / 16) * 16;
  const int maxquads    = ((g.max_cell_w + 10) / 4) * g.max_cell_h;
  const int pcap        = 64 * ((maxquads + 127) / 128);          // per-warp pair list (entries)
  const int ccap        = 64 * ((2 * maxquads + 127) / 128);      // per-warp corner list (entries)
  const int list_bytes  = 4 * ((g.max_cell_w + 1) / 2) * ((g.max_cell_h + 1) / 2);
  const int smem        = tile_bytes + tile8_bytes + score_bytes + 2 * 4 * (pcap + ccap) + list_bytes;
  // the attribute is per device, not per process: set on every launch that needs it
  if (smem > 48 * 1024) cudaFuncSetAttribute(k_fast_cells<TP, TPB>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  launch_k(k_fast_cells<TP, TPB>, dim3(g.total_cells, n_img), dim3(128), (size_t)smem, st, dg, cells, pyr, cand, cand_count, flags,
           tile_bytes, tile8_bytes, score_bytes, pcap, ccap);
}

void launch_fast(const ExtractorGeom& g, const ExtractorGeom* dg, const int4* cells, int n_img, const uint8_t* pyr,
                 unsigned* cand, int* cand_count, int* flags, cudaStream_t st) {
  // tile columns 0 .. 7 + max_cell_w + 3 + 4 (+ alignment): u8 words of 4 px, u16 words of 2 px;
  // the packed lists need cell height below 512 and pair columns below 128
  if (g.max_cell_w + 17 <= 4 * 15) launch_fast_tp<38, 15>(g, dg, cells, n_img, pyr, cand, cand_count, flags, st);
  else launch_fast_tp<76, 37>(g, dg, cells, n_img, pyr, cand, cand_count, flags, st);
  count_launch();
}

// persistent TMA variant; returns false when the cell geometry does not fit its tile (caller then
// uses launch_fast)
bool fast_persist_supported(const ExtractorGeom& g) { return g.max_cell_w <= 54 && g.max_cell_h + 6 <= 256 && g.max_cell_h < 512; }

void launch_fast_persist(const ExtractorGeom& g, const FastMaps& maps, const int4* recs, int n_cells, int n_img,
                         unsigned* cand, int* cand_count, int* flags, int* work_counter, cudaStream_t st) {
  FastParams P;
  P.recs = recs; P.n_cells = n_cells; P.total_work = n_cells * n_img; P.inv_cells = 1.0f / (float)n_cells;
  P.cand = cand; P.cand_count = cand_count; P.flags = flags; P.work_counter = work_counter;
  P.cand_img_stride = g.cand_img_stride; P.nlevels = g.nlevels; P.ini_th = g.ini_th; P.min_th = g.min_th;
  P.box_rows    = g.max_cell_h + 6;
  P.tile_bytes  = ((FT_PITCH * P.box_rows + 127) / 128) * 128;
  const int sp_max = ((((g.max_cell_w + 5 + 3) >> 2) | 1) << 2);
  P.score_bytes = ((sp_max * (g.max_cell_h + 2) + 15) / 16) * 16;
  const int maxquads = ((g.max_cell_w + 3) / 4 + 1) * g.max_cell_h;
  // two warps share a cell: measured best on B200 (0.667 ms per 256 images at the time; 4 warps 0.675, 1 warp 0.810):
  // fewer warps wait at the two per-cell barriers, more cells are in flight per SM
  constexpr int NW = 2;
  P.pcap = 64 * ((maxquads + 32 * NW - 1) / (32 * NW));
  const int smem = 128 + 2 * P.tile_bytes + 2 * P.score_bytes + 2 * NW * P.pcap;   // one u16 list per warp: pairs, then corner pairs in place
  void (*kern)(const FastMaps, const FastParams) = k_fast_persist<NW>;
  if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);   // per device
  // occupancy and SM count, cached per (device, smem) under a lock: the two ExtractORB threads of the stereo
  // Frame constructor pass through here concurrently, possibly for handles on different devices
  int per_sm = 0, n_sm = 0;
  {
    static std::mutex mu;
    static int c_smem[64], c_per_sm[64], c_n_sm[64];
    int dev = 0;
    cudaGetDevice(&dev);
    std::lock_guard<std::mutex> lk(mu);
    const int slot = dev & 63;
    if (c_n_sm[slot] == 0 || c_smem[slot] != smem) {
      cudaDeviceGetAttribute(&c_n_sm[slot], cudaDevAttrMultiProcessorCount, dev);
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&c_per_sm[slot], kern, 32 * NW, smem);
      c_smem[slot] = smem;
    }
    per_sm = c_per_sm[slot]; n_sm = c_n_sm[slot];
  }
  // CTAs per SM: with the single in-place list 13-14 CTAs fit and FAST alone is fastest that way (0.594 ms per 256
  // images), but the whole step then loses 3 % (90.6k against 93k frames/s): the other handles' blur / quadtree /
  // describe / staged-pyramid CTAs need shared memory and warp slots on the same SMs.  Between 9 and 12 CTAs the step
  // is flat within 0.5 % (93.3k - 93.8k) while FAST alone goes 0.655 -> 0.596 ms, so 12 is the default.
  // CUORB_FAST_CTAS overrides (0 = as many as fit).
  static const int cap = []() { const char* e = getenv("CUORB_FAST_CTAS"); return e ? atoi(e) : 12; }();
  const int use_per_sm = cap > 0 ? std::min(cap, std::max(per_sm, 1)) : std::max(per_sm, 1);
  const int grid = std::max(1, std::min(P.total_work, n_sm * use_per_sm));
  launch_k(kern, dim3(grid), dim3(32 * NW), (size_t)smem, st, maps, P);
  count_launch();
}

}  // namespace cuorb
