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
#include <algorithm>
#include <cstdlib>

#include "common.cuh"
#include "tma.cuh"

namespace cuorb {

constexpr int PYR_RS = 16;  // interior rows per thread strip
constexpr int PYR_TX = 64;  // threads (output words) per CTA
constexpr int PYR_MIN_CTAS = 2 * 148;  // below this many CTAs a pyramid launch switches to shorter strips
constexpr int PYR_TAB_TX = 192;  // k_pyr_resize_stage: upper bound of its per-level CTA width

__device__ __forceinline__ int reflect101(int i, int n) {
  // border (19) < n is checked on the host, so one reflection suffices
  i = i < 0 ? -i : i;
  return i >= n ? 2 * n - 2 - i : i;
}

// grid: (ceil(pitch/4/128), ceil(h/PYR_RS), n_img)
template <int RS>
__global__ void __launch_bounds__(256) k_pyr_level0(const ExtractorGeom* __restrict__ dg,
                                                    const uint8_t* __restrict__ imgs, size_t stride,
                                                    size_t image_stride, uint8_t* __restrict__ pyr,
                                                    int aligned, int* __restrict__ zero_counts, int n_counts,
                                                    int* __restrict__ zero_flags) {
  pdl_enter();
  // single-frame calls: the per-level candidate counters, the overflow flags and FAST's work counter are cleared
  // here, by the first kernel of the chain, instead of by two memsets in front of it (8 us of a 130 us call)
  if (zero_counts && blockIdx.x == 0 && blockIdx.y == 0) {
    if ((int)threadIdx.x < n_counts) zero_counts[blockIdx.z * n_counts + threadIdx.x] = 0;
    if (blockIdx.z == 0 && threadIdx.x < 2) zero_flags[threadIdx.x] = 0;
  }
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
  const int y0 = blockIdx.y * RS, y1 = min(y0 + RS, L.h);
  const int dp4 = L.pitch >> 2;
  unsigned* drow = reinterpret_cast<unsigned*>(dst) + word;
  // common case (aligned interior word, full strip away from the mirrored rows): pointer-increment copy,
  // all loads of the strip issued before the stores
  if (inner && y0 > EDGE_THRESHOLD && y0 + RS < L.h - 1 - EDGE_THRESHOLD) {
    const uint8_t* sp = src + (size_t)y0 * stride + xw;
    unsigned v[RS];
#pragma unroll
    for (int r = 0; r < RS; r++, sp += stride) v[r] = __ldg(reinterpret_cast<const unsigned*>(sp));
    unsigned* dpz = drow + (size_t)(y0 + EDGE_THRESHOLD) * dp4;
#pragma unroll
    for (int r = 0; r < RS; r++, dpz += dp4) *dpz = v[r];
    return;
  }
  unsigned v[RS];
#pragma unroll
  for (int r = 0; r < RS; r++) {
    const uint8_t* srow = src + (size_t)min(y0 + r, y1 - 1) * stride;
    if (inner) {
      v[r] = __ldg(reinterpret_cast<const unsigned*>(srow + xw));
    } else {
      v[r] = (unsigned)__ldg(srow + xs[0]) | ((unsigned)__ldg(srow + xs[1]) << 8) |
             ((unsigned)__ldg(srow + xs[2]) << 16) | ((unsigned)__ldg(srow + xs[3]) << 24);
    }
  }
#pragma unroll
  for (int r = 0; r < RS; r++) {
    const int y = y0 + r;
    if (y >= y1) break;
    const unsigned out = v[r];
    drow[(size_t)(y + EDGE_THRESHOLD) * dp4] = out;
    if (y >= 1 && y <= EDGE_THRESHOLD) drow[(size_t)(EDGE_THRESHOLD - y) * dp4] = out;
    if (y >= L.h - 1 - EDGE_THRESHOLD && y <= L.h - 2) drow[(size_t)(2 * L.h + EDGE_THRESHOLD - 2 - y) * dp4] = out;
  }
}

// per-thread tap setup for its four output columns
struct ColTaps {
  unsigned coef[4];   // a0 | a1 << 16
  unsigned sel[4];    // PRMT selector picking (p[sx], p[sx+1]) out of a word pair
  unsigned hi;        // bit b: column b's window starts in the second word pair (w1,w2)
};

// horizontal sums (>> 4) of one source row (three aligned words w0..w2) for the four columns
__device__ __forceinline__ void hsum(unsigned w0, unsigned w1, unsigned w2, const ColTaps& T, unsigned (&h)[4]) {
#pragma unroll
  for (int b = 0; b < 4; b++) {
    const bool up = (T.hi >> b) & 1;
    const unsigned v = __byte_perm(up ? w1 : w0, up ? w2 : w1, T.sel[b]);
    h[b] = __dp2a_lo(T.coef[b], v, 0u) >> 4;
  }
}

// ((b0*hA)>>16 + (b1*hB)>>16 + 2) >> 2 for four columns, packed.  b0s/b1s = weight << 16, so the
// truncated product is one IMAD.HI
__device__ __forceinline__ unsigned vblend(const unsigned (&hA)[4], const unsigned (&hB)[4], unsigned b0s, unsigned b1s) {
  unsigned v[4];
#pragma unroll
  for (int b = 0; b < 4; b++) v[b] = (__umulhi(b0s, hA[b]) + __umulhi(b1s, hB[b]) + 2u) >> 2;
  return __byte_perm(__byte_perm(v[0], v[1], 0x0040), __byte_perm(v[2], v[3], 0x0040), 0x5410);
}

// grid: (ceil(pitch/4/PYR_TX), ceil(h/PYR_RS), n_img).  Source-row driven: every source row of the
// strip is summed horizontally exactly once (two register sets alternate as "upper"/"lower" row, so
// nothing is copied), and each output row whose upper source row is the current one is emitted.
// WIDE: tap windows that do not fit 12 bytes (scale factors far above 2) use byte loads.
template <bool WIDE>
__global__ void __launch_bounds__(PYR_TX) k_pyr_resize(const ExtractorGeom* __restrict__ dg, int level,
                                                       uint8_t* __restrict__ pyr,
                                                       const ResizeTap* __restrict__ xtab,
                                                       const ResizeTap* __restrict__ ytab) {
  pdl_enter();
  const LevelGeom& L = dg->lv[level];
  const LevelGeom& S = dg->lv[level - 1];
  const int word = blockIdx.x * PYR_TX + threadIdx.x;
  if (word * 4 >= L.pitch) return;
  const int img = blockIdx.z;
  uint8_t* base = pyr + (size_t)img * dg->pyr_img_stride;
  const int spitch = S.pitch, dpitch = L.pitch, sh1 = S.h - 1, lh = L.h;
  unsigned* dst = reinterpret_cast<unsigned*>(base + L.plane_off) + word;

  ColTaps T;
  int sxs[4];
  int cbase;
  {
    int smin = 1 << 30;
#pragma unroll
    for (int b = 0; b < 4; b++) {
      int x = word * 4 + b - ROW_LEAD;
      x = reflect101(min(max(x, -EDGE_THRESHOLD), L.w + EDGE_THRESHOLD - 1), L.w);
      const ResizeTap tx = xtab[L.xtab_off + x];
      // sx+1 may equal S.w: that byte is the stored border pixel and its weight a1 is 0 there
      sxs[b] = tx.s;
      T.coef[b] = (unsigned)(unsigned short)tx.a0 | ((unsigned)(unsigned short)tx.a1 << 16);
      smin = min(smin, (int)tx.s);
    }
    cbase = smin & ~3;
    T.hi = 0;
#pragma unroll
    for (int b = 0; b < 4; b++) {
      int t = sxs[b] - cbase;
      if (t >= 4) { T.hi |= 1u << b; t -= 4; }
      T.sel[b] = (unsigned)t | ((unsigned)(t + 1) << 4);
    }
  }
  const uint8_t* src = base + S.pix_off + (WIDE ? 0 : cbase);

  // raw words of one source row (WIDE: the four columns' tap pairs packed as p0 | p1 << 8)
  auto fetch = [&](int sy, unsigned (&w)[4]) {
    const uint8_t* r = src + (size_t)min(sy, sh1) * spitch;  // OpenCV clips source rows to sh-1
    if (!WIDE) {
      const unsigned* p = reinterpret_cast<const unsigned*>(r);
      w[0] = p[0]; w[1] = p[1]; w[2] = p[2]; w[3] = 0;
    } else {
#pragma unroll
      for (int b = 0; b < 4; b++) w[b] = (unsigned)r[sxs[b]] | ((unsigned)r[sxs[b] + 1] << 8);
    }
  };
  auto hsum_row = [&](const unsigned (&w)[4], unsigned (&h)[4]) {
    if (!WIDE) {
      hsum(w[0], w[1], w[2], T, h);
    } else {
#pragma unroll
      for (int b = 0; b < 4; b++) h[b] = __dp2a_lo(T.coef[b], w[b], 0u) >> 4;
    }
  };
  const int y0 = blockIdx.y * PYR_RS, y1 = min(y0 + PYR_RS, lh);
  const ResizeTap* yt = ytab + L.ytab_off;
  int y = y0;
  ResizeTap ty = yt[y];
  int s = max((int)ty.s, 0);
  // emit every output row whose upper source row is `s` (hA = row s, hB = row s+1); false when the strip is done
  auto emit_rows = [&](const unsigned (&hA)[4], const unsigned (&hB)[4]) -> bool {
    while (max((int)ty.s, 0) == s) {
      const unsigned out = vblend(hA, hB, (unsigned)ty.a0 << 16, (unsigned)ty.a1 << 16);
      dst[(size_t)(y + EDGE_THRESHOLD) * (dpitch >> 2)] = out;
      if (y >= 1 && y <= EDGE_THRESHOLD) dst[(size_t)(EDGE_THRESHOLD - y) * (dpitch >> 2)] = out;
      if (y >= lh - 1 - EDGE_THRESHOLD && y <= lh - 2) dst[(size_t)(2 * lh + EDGE_THRESHOLD - 2 - y) * (dpitch >> 2)] = out;
      if (++y >= y1) return false;
      ty = yt[y];
    }
    s++;
    return true;
  };

  unsigned h0[4], h1[4], raw[4][4];
  fetch(s, raw[0]);
  hsum_row(raw[0], h0);
  for (;;) {
    // four source rows in flight per iteration (memory-level parallelism: the small levels are
    // DRAM-latency bound); h0 holds source row s on entry
#pragma unroll
    for (int k = 0; k < 4; k++) fetch(s + 1 + k, raw[k]);
    hsum_row(raw[0], h1);
    if (!emit_rows(h0, h1)) return;
    hsum_row(raw[1], h0);
    if (!emit_rows(h1, h0)) return;
    hsum_row(raw[2], h1);
    if (!emit_rows(h0, h1)) return;
    hsum_row(raw[3], h0);
    if (!emit_rows(h1, h0)) return;
  }
}

// ---- table-driven, staged variant (all scale factors whose 4-column tap window fits 8 source bytes) ----
// Everything a thread needs about its four output columns comes from a host-built table
// (build_resize_words): the aligned source offset of its 12-byte window, the funnel shift that
// brings the first tap to byte 0, four PRMT selectors and the four coefficient pairs.  Per source
// row: 3 aligned loads, 2 funnel shifts, then per column PRMT + dp2a + shift.  The strip's vertical
// taps sit in shared memory, so the row loop has no dependent global load; the horizontal sums of
// the previous lower source row are reused whenever the source step is 1.
//   wtab[2*word]   = {window byte offset (4-aligned), funnel shift (bits), sel0 | sel1 << 16, sel2 | sel3 << 16}
//   wtab[2*word+1] = {a0 | a1 << 16} for the four columns
// blockDim.x = the row's word count rounded up to whole warps (at most PYR_TAB_TX), so few lanes idle
// on the narrow levels; `rs` = output rows per strip (<= PYR_RS).
// The row loop is written for instruction count (the kernel is issue bound, not DRAM bound): one
// 16-byte shared-memory record per output row {source row offsets, advance, vertical weights}, 32-bit
// offsets from two base pointers, border rows on a rare uniform path, and the two horizontal-sum
// register sets swap roles by code duplication (state A: H0 = upper row; state B: H1 = upper row)
// instead of register copies.
// The CTA first stages every source row its strip needs (rows s_first .. s_last+1, clipped to the last row;
// bytes [xlo, xlo + tw) of each) in shared memory with one cp.async.bulk per row, all in flight at once and
// signalled on one mbarrier: ~20 rows x ~700 bytes outstanding per CTA, and the row loop only touches shared
// memory.  (A register-only form of the same loop, ~12 bytes in flight per thread, was 5 % slower on the
// stage and was removed in round 2.)  Levels whose staged window does not fit (build_pyr_plan) take the
// generic k_pyr_resize<WIDE> above.
struct PyrCols { int xlo[PYR_MAX_COLBLOCKS], tw[PYR_MAX_COLBLOCKS]; };

__global__ void __launch_bounds__(PYR_TAB_TX, 1536 / PYR_TAB_TX) k_pyr_resize_stage(const ExtractorGeom* __restrict__ dg, int level,
                                                           uint8_t* __restrict__ pyr,
                                                           const uint4* __restrict__ wtab,
                                                           const ResizeTap* __restrict__ ytab, int rs,
                                                           const PyrCols pc) {
  pdl_enter();
  extern __shared__ __align__(128) unsigned char s_tile[];
  __shared__ uint4 s_rec[PYR_RS];   // per output row {source row offsets, advance, vertical weights}, relative to the staged tile
  __shared__ uint64_t s_bar;
  const LevelGeom& L = dg->lv[level];
  const LevelGeom& S = dg->lv[level - 1];
  const int lh = L.h, sh1 = S.h - 1;
  const int y0 = blockIdx.y * rs, n = min(rs, lh - y0);
  int tw = pc.tw[0], xlo = pc.xlo[0];
#pragma unroll
  for (int b = 1; b < PYR_MAX_COLBLOCKS; b++)
    if ((int)blockIdx.x == b) { tw = pc.tw[b]; xlo = pc.xlo[b]; }   // (constant-bank reads; a dynamic index would spill the struct)
  const ResizeTap* yt = ytab + L.ytab_off + y0;
  const int r_first = max((int)yt[0].s, 0), n_rows = max((int)yt[n - 1].s, 0) + 2 - r_first;
  uint8_t* base = pyr + (size_t)blockIdx.z * dg->pyr_img_stride;
  if (threadIdx.x == 0) {
    mbar_init(&s_bar, 1);
    mbar_fence_init();
  }
  if ((int)threadIdx.x < n) {
    const ResizeTap t = yt[threadIdx.x];
    const int sy = max((int)t.s, 0);
    const int prev = threadIdx.x == 0 ? -4 : max((int)yt[(int)threadIdx.x - 1].s, 0);
    const unsigned adv = (unsigned)min(max(sy - prev, 0), 2);
    s_rec[threadIdx.x] = make_uint4((unsigned)((sy - r_first) * tw) | adv, (unsigned)((sy + 1 - r_first) * tw),
                                    (unsigned)t.a0 << 16, (unsigned)t.a1 << 16);
  }
  __syncthreads();
  if (threadIdx.x == 0) mbar_expect_tx(&s_bar, (unsigned)(n_rows * tw));
  {
    const uint8_t* srow0 = base + S.pix_off + xlo;
    for (int j = threadIdx.x; j < n_rows; j += blockDim.x)
      bulk_load_1d(s_tile + j * tw, srow0 + (size_t)min(r_first + j, sh1) * S.pitch, (unsigned)tw, &s_bar);
  }
  const int word = blockIdx.x * blockDim.x + threadIdx.x;
  if (word * 4 >= L.pitch) return;
  const uint4 e0 = __ldg(&wtab[2 * word]), cf = __ldg(&wtab[2 * word + 1]);
  const unsigned char* src = s_tile + ((int)e0.x - xlo);
  const unsigned dp = (unsigned)L.pitch;
  uint8_t* dst = base + L.plane_off + 4 * (size_t)word;               // padded row 0 of this word column
  const unsigned shift = e0.y, sel0 = e0.z & 0xFFFFu, sel1 = e0.z >> 16, sel2 = e0.w & 0xFFFFu, sel3 = e0.w >> 16;

  auto fetch = [&](unsigned off, unsigned (&w)[3]) {
    const unsigned* p = reinterpret_cast<const unsigned*>(src + off);
    w[0] = p[0]; w[1] = p[1]; w[2] = p[2];
  };
  auto hsum4 = [&](const unsigned (&w)[3], unsigned (&h)[4]) {
    const unsigned X0 = __funnelshift_r(w[0], w[1], shift), X1 = __funnelshift_r(w[1], w[2], shift);
    h[0] = __dp2a_lo(cf.x, __byte_perm(X0, X1, sel0), 0u) >> 4;
    h[1] = __dp2a_lo(cf.y, __byte_perm(X0, X1, sel1), 0u) >> 4;
    h[2] = __dp2a_lo(cf.z, __byte_perm(X0, X1, sel2), 0u) >> 4;
    h[3] = __dp2a_lo(cf.w, __byte_perm(X0, X1, sel3), 0u) >> 4;
  };
  auto blend = [&](const unsigned (&hA)[4], const unsigned (&hB)[4], unsigned b0s, unsigned b1s) -> unsigned {
    unsigned v[4];
#pragma unroll
    for (int b = 0; b < 4; b++) v[b] = __umulhi(b0s, hA[b]) + __umulhi(b1s, hB[b]) + 2u;
    const unsigned lo = (v[0] | (v[1] << 16)) >> 2, hi = (v[2] | (v[3] << 16)) >> 2;
    return __byte_perm(lo, hi, 0x6420);
  };

  unsigned H0[4], H1[4], rA[3], rB[3];
  uint4 rec = s_rec[0];
  mbar_wait(&s_bar, 0u);
  fetch(rec.x & ~3u, rA); fetch(rec.y, rB);
  hsum4(rA, H0); hsum4(rB, H1);
  int i = 0;
  unsigned doff = (unsigned)(y0 + EDGE_THRESHOLD) * dp;
  const bool edge_strip = y0 <= EDGE_THRESHOLD || y0 + n >= lh - 1 - EDGE_THRESHOLD;
  auto row = [&](unsigned (&hA)[4], unsigned (&hB)[4]) -> unsigned {
    const unsigned b0s = rec.z, b1s = rec.w;
    unsigned adv = 0;
    if (i + 1 < n) {
      rec = s_rec[i + 1];
      adv = rec.x & 3u;
      if (adv != 0) {
        fetch(rec.y, rB);
        if (adv == 2) fetch(rec.x & ~3u, rA);
      }
    }
    const unsigned out = blend(hA, hB, b0s, b1s);
    *reinterpret_cast<unsigned*>(dst + doff) = out;
    if (edge_strip) {
      const int y = y0 + i;
      if (y >= 1 && y <= EDGE_THRESHOLD) *reinterpret_cast<unsigned*>(dst + (unsigned)(EDGE_THRESHOLD - y) * dp) = out;
      if (y >= lh - 1 - EDGE_THRESHOLD && y <= lh - 2) *reinterpret_cast<unsigned*>(dst + (unsigned)(2 * lh + EDGE_THRESHOLD - 2 - y) * dp) = out;
    }
    doff += dp;
    i++;
    if (adv == 1) hsum4(rB, hA);
    else if (adv == 2) { hsum4(rA, hA); hsum4(rB, hB); }
    return adv;
  };
  for (;;) {
    const unsigned a = row(H0, H1);
    if (i >= n) return;
    if (a != 1) continue;
    for (;;) {
      const unsigned b = row(H1, H0);
      if (i >= n) return;
      if (b == 1) break;
    }
  }
}

// host: CTA shape and staged source window per level for k_pyr_resize_stage
void build_pyr_plan(const ExtractorGeom& g, const std::vector<uint4>& wt, const int* wtab_off,
                    const std::vector<ResizeTap>& ytab, PyrPlan* plan) {
  for (int l = 0; l < MAX_LEVELS; l++) plan[l] = PyrPlan();
  for (int l = 1; l < g.nlevels; l++) {
    if (wtab_off[l] < 0) continue;
    const LevelGeom& L = g.lv[l];
    const LevelGeom& S = g.lv[l - 1];
    PyrPlan p;
    const int words = L.pitch / 4;
    p.nb = (words + PYR_TAB_TX - 1) / PYR_TAB_TX;
    if (p.nb > PYR_MAX_COLBLOCKS) continue;
    p.tx = (((words + p.nb - 1) / p.nb) + 31) / 32 * 32;
    int tw_max = 0;
    bool ok = true;
    for (int b = 0; b < p.nb; b++) {
      int lo = 1 << 30, hi = 0;
      for (int w = b * p.tx; w < std::min((b + 1) * p.tx, words); w++) {
        const int x = (int)wt[(size_t)wtab_off[l] + 2 * w].x;
        lo = std::min(lo, x); hi = std::max(hi, x + 12);
      }
      if (lo > hi) { lo = 0; hi = 16; }                       // column block without words (cannot happen: nb = ceil)
      p.xlo[b] = lo & ~15;
      // never past the end of the padded source row (the bytes a capped window loses are not selected by any tap)
      p.tw[b] = std::min((hi - p.xlo[b] + 15) & ~15, S.pitch - ROW_LEAD - p.xlo[b]);
      if (p.tw[b] <= 0) ok = false;
      tw_max = std::max(tw_max, p.tw[b]);
    }
    for (int y0 = 0; y0 < L.h; y0 += PYR_RS) {
      const int n = std::min(PYR_RS, L.h - y0);
      const int first = std::max((int)ytab[L.ytab_off + y0].s, 0), last = std::max((int)ytab[L.ytab_off + y0 + n - 1].s, 0) + 1;
      p.max_rows = std::max(p.max_rows, last - first + 1);
    }
    p.smem = p.max_rows * tw_max + 16;
    if (!ok || p.smem > 64 * 1024) continue;
    plan[l] = p;
  }
}

// host: per-word tables for k_pyr_resize_stage, levels 1..n-1 back to back; off[l] = first entry (in
// uint4) of level l, or -1 when the level needs the byte-load kernel (window wider than 8 bytes)
void build_resize_words(const ExtractorGeom& g, const std::vector<ResizeTap>& xtab, std::vector<uint4>& wt, int* off) {
  wt.clear();
  off[0] = -1;
  for (int l = 1; l < g.nlevels; l++) {
    const LevelGeom& L = g.lv[l];
    const size_t start = wt.size();
    bool ok = true;
    for (int word = 0; word * 4 < L.pitch && ok; word++) {
      int sx[4], smin = 1 << 30, smax = 0;
      unsigned coef[4];
      for (int b = 0; b < 4; b++) {
        int x = word * 4 + b - ROW_LEAD;
        x = std::min(std::max(x, -EDGE_THRESHOLD), L.w + EDGE_THRESHOLD - 1);
        x = x < 0 ? -x : x;
        x = x >= L.w ? 2 * L.w - 2 - x : x;
        const ResizeTap t = xtab[L.xtab_off + x];
        sx[b] = t.s; smin = std::min(smin, (int)t.s); smax = std::max(smax, (int)t.s);
        coef[b] = (unsigned)(unsigned short)t.a0 | ((unsigned)(unsigned short)t.a1 << 16);
      }
      if (smax - smin + 1 > 7) { ok = false; break; }
      const int cbase = smin & ~3, t0 = smin - cbase;
      unsigned sel[4];
      for (int b = 0; b < 4; b++) { const int d = sx[b] - smin; sel[b] = (unsigned)d | ((unsigned)(d + 1) << 4); }
      wt.push_back(make_uint4((unsigned)cbase, (unsigned)(8 * t0), sel[0] | (sel[1] << 16), sel[2] | (sel[3] << 16)));
      wt.push_back(make_uint4(coef[0], coef[1], coef[2], coef[3]));
    }
    if (ok) off[l] = (int)start;
    else { wt.resize(start); off[l] = -1; }
  }
}

void launch_pyramid(const ExtractorGeom& g, const ExtractorGeom* dg, const uint8_t* imgs, int, int,
                    size_t stride, size_t image_stride, int n_img, uint8_t* pyr,
                    const ResizeTap* xtab, const ResizeTap* ytab, const uint4* wtab, const int* wtab_off,
                    const PyrPlan* plan, cudaStream_t st, int* zero_counts, int* zero_flags) {
  const int aligned = (((uintptr_t)imgs | stride | image_stride) & 3) == 0;
  for (int l = 0; l < g.nlevels; l++) {
    const LevelGeom& L = g.lv[l];
    dim3 grid((L.pitch / 4 + PYR_TX - 1) / PYR_TX, (L.h + PYR_RS - 1) / PYR_RS, n_img);
    if (l == 0) {
      // CTA width = the padded row's word count split evenly over whole warps (208 words -> one CTA of 224)
      const int words = L.pitch / 4, nb = (words + 255) / 256;
      const int tx = (((words + nb - 1) / nb) + 31) / 32 * 32;
      // 16-row strips for large batches; when that would leave most of the 148 SMs without a CTA (a handful of
      // images) the strips are 4 rows and the kernel's critical path - one thread's row loop - is a quarter as long
      const int nbx = (words + tx - 1) / tx;
      PdlPause first_of_chain;   // follows the frame upload (and, for large batches, two memsets)
      if ((int)grid.y * nbx * n_img < PYR_MIN_CTAS)
        launch_k(k_pyr_level0<4>, dim3(nbx, (L.h + 3) / 4, n_img), dim3(tx), 0, st, dg, imgs, stride, image_stride, pyr, aligned, zero_counts,
                 g.nlevels, zero_flags);
      else
        launch_k(k_pyr_level0<PYR_RS>, dim3(nbx, grid.y, n_img), dim3(tx), 0, st, dg, imgs, stride, image_stride, pyr, aligned, zero_counts,
                 g.nlevels, zero_flags);
    } else {
      // four adjacent output columns span 3*scale + 2 source bytes (+3 for alignment): 12 bytes up to scale 2
      const double sx = (double)g.lv[l - 1].w / L.w;
      if (wtab && wtab_off[l] >= 0 && plan && plan[l].tx > 0) {
        const PyrPlan& p = plan[l];
        PyrCols pc;
        for (int b = 0; b < PYR_MAX_COLBLOCKS; b++) { pc.xlo[b] = p.xlo[b]; pc.tw[b] = p.tw[b]; }
        // above the 48 KB default only for very wide levels; the attribute is per device, so it is set on every
        // such launch rather than cached per process
        if (p.smem > 48 * 1024) cudaFuncSetAttribute(k_pyr_resize_stage, cudaFuncAttributeMaxDynamicSharedMemorySize, p.smem);
        // strips of PYR_RS output rows amortise the staging for large batches; with a handful of images a level is
        // only 10-30 such strips on 148 SMs and the row loop is the kernel's critical path, so the strips get shorter
        // until the grid fills the GPU (single 752x480 frame: pyramid stage 68 -> 36 us, the call 0.160 -> 0.133 ms;
        // strips of 1, 2, 3 or 4 rows within 1 us of each other)
        int rs = PYR_RS;
        while (rs > 2 && ((L.h + rs - 1) / rs) * p.nb * n_img < PYR_MIN_CTAS) rs >>= 1;
        launch_k(k_pyr_resize_stage, dim3(p.nb, (L.h + rs - 1) / rs, n_img), dim3(p.tx), (size_t)p.smem, st, dg, l, pyr, wtab + wtab_off[l], ytab, rs, pc);
      }
      else if (3.0 * sx + 5.0 < 11.5) launch_k(k_pyr_resize<false>, grid, dim3(PYR_TX), 0, st, dg, l, pyr, xtab, ytab);
      else launch_k(k_pyr_resize<true>, grid, dim3(PYR_TX), 0, st, dg, l, pyr, xtab, ytab);
    }
    count_launch();
  }
}

}  // namespace cuorb
