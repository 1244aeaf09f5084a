// K7 — 7x7 sigma-2 Gaussian blur of every pyramid level (replaces the per-level
// clone() + cv::GaussianBlur(7x7, 2, 2, BORDER_REFLECT_101), src/ORBextractor.cc:1103-1104).
// OpenCV's CV_8U path is fixed point: kernel {18,34,48,56,48,34,18}/256, rows in 8.8, columns in
// 16.16, result (v + 32768) >> 16.  Nothing is rounded between the two passes, so the sum
//   out = (sum_j k_j * sum_i k_i * p[y+j][x+i] + 32768) >> 16
// may be evaluated in either order; this kernel runs the COLUMN pass first because both passes then
// map onto the packed dot-product instructions.  The REFLECT_101 halo is the border K1 already
// stored, so no reflection is computed here.
//
// One lane owns one aligned output word (4 px) and walks down a strip in groups of four rows:
//   * four source rows of its word are loaded and byte-transposed (8 PRMT), so each column's four
//     vertically adjacent pixels share a register; a group's 7-tap column sums V (<= 255*256, 16 bit)
//     are then 2-3 DP4A per pixel on three such blocks (two of them carried over from the previous
//     group: every source row is loaded and transposed once);
//   * V of adjacent columns is packed in u16 pairs; the pairs of the left and right neighbour word
//     come from the neighbouring lanes (4 SHFL per row), and each output pixel is 4 DP2A on pairs
//     with the 32768 rounding term as the accumulator seed.
// About 10 instructions per pixel, against 16 (+29 % row halo) for a row-pass-first register
// formulation.  Lanes 0 and 31 of a warp only supply halo columns, so a warp covers 30 output words.
// No shared memory, no barriers.  Bound: issue slots (it runs beside k_fast_persist on a second
// stream), then HBM (read P, write P).
#include "common.cuh"

namespace cuorb {

constexpr int BL_WARPS = 1;     // one warp per CTA: every strip quantity is CTA-uniform, so loop control and the weights live on the uniform datapath
constexpr int BL_ROWS  = 32;    // output rows per strip (8 groups of 4)
constexpr int BL_CHUNK = 30;    // output words per warp (lanes 1..30)

namespace {

// 4x4 byte transpose: r[j] = four pixels of source row j  ->  t[c] = column c's pixels of rows 0..3
__device__ __forceinline__ void transpose4(const unsigned (&r)[4], unsigned (&t)[4]) {
  const unsigned a = __byte_perm(r[0], r[1], 0x5140), b = __byte_perm(r[2], r[3], 0x5140);
  const unsigned c = __byte_perm(r[0], r[1], 0x7362), d = __byte_perm(r[2], r[3], 0x7362);
  t[0] = __byte_perm(a, b, 0x5410); t[1] = __byte_perm(a, b, 0x7632);
  t[2] = __byte_perm(c, d, 0x5410); t[3] = __byte_perm(c, d, 0x7632);
}

__host__ __device__ constexpr unsigned pk4(unsigned a, unsigned b, unsigned c, unsigned d) { return a | (b << 8) | (c << 16) | (d << 24); }

}  // namespace

// grid: (tiles over all levels, n_img); a tile = BL_WARPS consecutive items of one level, an item =
// (row strip, chunk of BL_CHUNK words), numbered strip-major
__global__ void __launch_bounds__(BL_WARPS * 32) k_blur(const ExtractorGeom* __restrict__ dg,
                                                        const int4* __restrict__ tiles,
                                                        const uint8_t* __restrict__ pyr,
                                                        uint8_t* __restrict__ blur) {
  const int4 t = tiles[blockIdx.x];   // {level, first item, chunks per row, items in the level}
  const int lane = threadIdx.x & 31;
  const int item = BL_WARPS == 1 ? t.y : t.y + (int)(threadIdx.x >> 5);
  if (item >= t.w) return;
  const LevelGeom& L = dg->lv[t.x];
  const int strip = item / t.z, chunk = item - strip * t.z;
  const int words = L.bpitch >> 2;
  const int word = chunk * BL_CHUNK + lane - 1;                 // -1 .. words: lanes 0 / 31 are halo columns
  const int img = blockIdx.y;
  const int y0 = strip * BL_ROWS, y1 = min(y0 + BL_ROWS, L.h);
  const unsigned pitch = (unsigned)L.pitch, bpitch = (unsigned)L.bpitch;
  // word `words` (first word right of the blurred plane) still lies inside the 19-px border of the source
  const uint8_t* src = pyr + (size_t)img * dg->pyr_img_stride + L.pix_off + 4 * (ptrdiff_t)min(word, words) +
                       (ptrdiff_t)(y0 - 3) * (ptrdiff_t)pitch;
  uint8_t* dst = blur + (size_t)img * dg->blur_img_stride + L.blur_off + 4 * (ptrdiff_t)word + (size_t)y0 * bpitch;
  const bool writer = lane >= 1 && lane <= BL_CHUNK && word < words;

  // column-pass weights for output row r of a group: block A holds source rows r0-3 .. r0, B r0+1 .. r0+4, C r0+5 .. r0+8
  constexpr unsigned WA0 = pk4(18, 34, 48, 56), WB0 = pk4(48, 34, 18, 0);
  constexpr unsigned WA1 = pk4(0, 18, 34, 48),  WB1 = pk4(56, 48, 34, 18);
  constexpr unsigned WA2 = pk4(0, 0, 18, 34),   WB2 = pk4(48, 56, 48, 34), WC2 = pk4(18, 0, 0, 0);
  constexpr unsigned WA3 = pk4(0, 0, 0, 18),    WB3 = pk4(34, 48, 56, 48), WC3 = pk4(34, 18, 0, 0);
  // row-pass weights on u16 pairs (dp2a_lo uses bytes 0,1; dp2a_hi bytes 2,3)
  constexpr unsigned H0 = pk4(0, 18, 34, 48);    // lo (0,18)   hi (34,48)
  constexpr unsigned H1 = pk4(56, 48, 34, 18);   // lo (56,48)  hi (34,18)
  constexpr unsigned H2 = pk4(18, 34, 48, 56);   // lo (18,34)  hi (48,56)
  constexpr unsigned H3 = pk4(48, 34, 18, 0);    // lo (48,34)  hi (18,0)

  auto load_block = [&](int first_row, unsigned (&tb)[4]) {   // rows relative to y0 - 3
    unsigned r[4];
    const uint8_t* p = src + (unsigned)first_row * pitch;
#pragma unroll
    for (int j = 0; j < 4; j++) r[j] = __ldg(reinterpret_cast<const unsigned*>(p + (unsigned)j * pitch));
    transpose4(r, tb);
  };

  // running 32-bit offsets from the two base pointers (one add per row instead of 64-bit products)
  unsigned soff = 8u * pitch;          // next source row to load, relative to y0 - 3 (blocks 0 and 1 are loaded up front)
  unsigned doff = 0;                   // output row, relative to y0
  int rows_left = y1 - y0;

  // one output row: column sums v[0..3] of this lane's word -> pairs -> neighbours' pairs -> 4 px.
  // FULL: the strip has all BL_ROWS rows, so the store needs no row test (a single predicated STG)
  auto emit_row = [&](const unsigned (&v)[4], bool in_rows) {
    const unsigned P0 = __byte_perm(v[0], v[1], 0x5410), P1 = __byte_perm(v[2], v[3], 0x5410);
    const unsigned L0 = __shfl_up_sync(0xffffffffu, P0, 1), L1 = __shfl_up_sync(0xffffffffu, P1, 1);
    const unsigned R0 = __shfl_down_sync(0xffffffffu, P0, 1), R1 = __shfl_down_sync(0xffffffffu, P1, 1);
    // pixel c: taps V[c-3 .. c+3]; own columns 0..3 = P0, P1; left word -4..-1 = L0, L1; right word 4..7 = R0, R1
    const unsigned a0 = __dp2a_lo(L0, H0, __dp2a_hi(L1, H0, __dp2a_lo(P0, H1, __dp2a_hi(P1, H1, 32768u))));
    const unsigned a1 = __dp2a_lo(L1, H2, __dp2a_hi(P0, H2, __dp2a_lo(P1, H3, __dp2a_hi(R0, H3, 32768u))));
    const unsigned a2 = __dp2a_lo(L1, H0, __dp2a_hi(P0, H0, __dp2a_lo(P1, H1, __dp2a_hi(R0, H1, 32768u))));
    const unsigned a3 = __dp2a_lo(P0, H2, __dp2a_hi(P1, H2, __dp2a_lo(R0, H3, __dp2a_hi(R1, H3, 32768u))));
    if (writer && in_rows)
      *reinterpret_cast<unsigned*>(dst + doff) = __byte_perm(__byte_perm(a0, a1, 0x0062), __byte_perm(a2, a3, 0x0062), 0x5410);
    doff += bpitch;
  };

  // one group of four output rows; A, B, C as above
  auto group = [&](const unsigned (&A)[4], const unsigned (&B)[4], const unsigned (&C)[4], bool full) {
    unsigned v[4];
#pragma unroll
    for (int c = 0; c < 4; c++) v[c] = __dp4a(A[c], WA0, __dp4a(B[c], WB0, 0u));
    emit_row(v, true);
#pragma unroll
    for (int c = 0; c < 4; c++) v[c] = __dp4a(A[c], WA1, __dp4a(B[c], WB1, 0u));
    emit_row(v, full || rows_left > 1);
#pragma unroll
    for (int c = 0; c < 4; c++) v[c] = __dp4a(A[c], WA2, __dp4a(B[c], WB2, __dp4a(C[c], WC2, 0u)));
    emit_row(v, full || rows_left > 2);
#pragma unroll
    for (int c = 0; c < 4; c++) v[c] = __dp4a(A[c], WA3, __dp4a(B[c], WB3, __dp4a(C[c], WC3, 0u)));
    emit_row(v, full || rows_left > 3);
    rows_left -= 4;
  };
  auto load_next = [&](unsigned (&tb)[4]) {
    unsigned r[4];
    const uint8_t* p = src + soff;
#pragma unroll
    for (int j = 0; j < 4; j++) r[j] = __ldg(reinterpret_cast<const unsigned*>(p + (unsigned)j * pitch));
    transpose4(r, tb);
    soff += 4u * pitch;
  };

  // the three blocks rotate through three register sets (no copies)
  unsigned T0[4], T1[4], T2[4];
  load_block(0, T0);
  load_block(4, T1);
  const bool full = rows_left == BL_ROWS;   // the common case: no per-row test
  for (;;) {
    load_next(T2); group(T0, T1, T2, full);
    if (rows_left <= 0) break;
    load_next(T0); group(T1, T2, T0, full);
    if (rows_left <= 0) break;
    load_next(T1); group(T2, T0, T1, full);
    if (rows_left <= 0) break;
  }
}

void build_blur_tiles(const ExtractorGeom& g, std::vector<int4>& v) {
  v.clear();
  for (int l = 0; l < g.nlevels; l++) {
    const int words = g.lv[l].bpitch / 4;
    const int chunks = (words + BL_CHUNK - 1) / BL_CHUNK;
    const int items = chunks * ((g.lv[l].h + BL_ROWS - 1) / BL_ROWS);
    for (int i = 0; i < items; i += BL_WARPS) v.push_back(make_int4(l, i, chunks, items));
  }
}

void launch_blur(const ExtractorGeom*, const ExtractorGeom* dg, int n_img, const int4* tiles,
                 int n_tiles, const uint8_t* pyr, uint8_t* blur, cudaStream_t st) {
  k_blur<<<dim3(n_tiles, n_img), BL_WARPS * 32, 0, st>>>(dg, tiles, pyr, blur);
  count_launch();
}

}  // namespace cuorb
