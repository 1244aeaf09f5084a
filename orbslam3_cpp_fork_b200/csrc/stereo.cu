// K10 — stereo association (replaces Frame::ComputeStereoMatches, src/Frame.cc:944-1113; the
// path BASELINE.json calls "SearchStereo").  Device-resident: consumes the keypoints, descriptors
// and pyramids the two extractor handles left in HBM.
//
// k_stereo_rows: one CTA per image.  The reference's per-row candidate table (vRowIndices,
//   :953-971) becomes a counting sort of the right keypoints into 8-row buckets of their y; an
//   entry carries the exact membership interval [floor(y-r), ceil(y+r)], r = 2*scale[octave],
//   the octave, x and the original index iR.
// k_stereo_match: one warp per left keypoint.  Only the buckets that can hold a right keypoint
//   whose interval contains int(vL) are scanned (a few dozen entries instead of all ~1000); the
//   reference's predicates (:1007-1018) are evaluated exactly, Hamming = 8 x (XOR, POPC), and the
//   ascending-iR visiting order with strict "<" (:1021) becomes a warp arg-min on the key
//   (dist << 16 | iR), which makes the bucket order irrelevant.  The winner is refined by the
//   11-window SAD slide on the un-blurred pyramid level (:1028-1067): lane (2*row + half) owns one
//   patch row and half of the 11 shifts, rows are fetched as aligned words, re-aligned with
//   funnel shifts and compared 4 pixels per VABSDIFF4; a butterfly sums the rows.  The parabola
//   fit (:1074-1098) runs in FMA-free float.
// k_stereo_outliers: one CTA per image.  median = element size/2 of the sorted SAD list (:1101-1103)
//   by a two-level histogram select; matches with SAD >= 1.5*1.4*median are reset (:1105-1112).
#include "common.cuh"

namespace cuorb {

constexpr int TH_HIGH = 100, TH_LOW = 50;  // src/ORBmatcher.cc:30-31
constexpr int BUCKET_SHIFT = 3;

__device__ __forceinline__ int row_bucket(float y, int nb) {
  return min(max((int)y, 0) >> BUCKET_SHIFT, nb - 1);
}

// grid: n_img, block 256
__global__ void __launch_bounds__(256) k_stereo_rows(StereoArgs A) {
  pdl_enter();
  __shared__ int s_cnt[STEREO_MAX_BUCKETS + 1];
  __shared__ int s_warp[8];
  const ExtractorGeom* dg = A.dgL;
  const int img = blockIdx.x, tid = threadIdx.x;
  const int nR = A.nR[img];
  const int nb = min((dg->lv[0].h + (1 << BUCKET_SHIFT) - 1) >> BUCKET_SHIFT, STEREO_MAX_BUCKETS);
  const size_t base = (size_t)img * A.kp_cap;
  for (int b = tid; b <= STEREO_MAX_BUCKETS; b += 256) s_cnt[b] = 0;
  __syncthreads();
  for (int i = tid; i < nR; i += 256) atomicAdd(&s_cnt[row_bucket(A.kpsR[base + i].y, nb)], 1);
  __syncthreads();
  // exclusive scan of s_cnt[0..nb) (two buckets per thread)
  {
    const int b0 = 2 * tid, b1 = 2 * tid + 1;
    const int c0 = b0 < nb ? s_cnt[b0] : 0, c1 = b1 < nb ? s_cnt[b1] : 0;
    int inc = c0 + c1;
    const int lane = tid & 31, wid = tid >> 5;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
    if (lane == 31) s_warp[wid] = inc;
    __syncthreads();
    int woff = 0;
    for (int w = 0; w < wid; w++) woff += s_warp[w];
    const int ex = woff + inc - (c0 + c1);
    __syncthreads();
    if (b0 < nb) s_cnt[b0] = ex;
    if (b1 < nb) s_cnt[b1] = ex + c0;
    if (tid == 0) s_cnt[nb] = nR;
  }
  __syncthreads();
  int* gstart = A.rstart + (size_t)img * (STEREO_MAX_BUCKETS + 1);
  for (int b = tid; b <= nb; b += 256) gstart[b] = s_cnt[b];
  __syncthreads();
  for (int i = tid; i < nR; i += 256) {
    const cuorb_keypoint kp = A.kpsR[base + i];
    const float r = __fmul_rn(2.0f, dg->lv[kp.octave].scale);                         // :963
    const int maxr = (int)ceilf(__fadd_rn(kp.y, r)), minr = (int)floorf(__fsub_rn(kp.y, r));  // :965-966
    const int pos = atomicAdd(&s_cnt[row_bucket(kp.y, nb)], 1);
    A.rtab[base + pos] = make_uint4((unsigned)min(max(minr, 0), 8191) | ((unsigned)min(max(maxr, 0), 8191) << 13) |
                                        ((unsigned)kp.octave << 26),
                                    __float_as_uint(kp.x), (unsigned)i, 0u);
  }
}

// grid: (ceil(kp_cap/8), n_img), block 256
__global__ void __launch_bounds__(256, 6) k_stereo_match(StereoArgs A, int band) {
  pdl_enter();
  const ExtractorGeom* dg = A.dgL;
  const int img = blockIdx.y, lane = threadIdx.x & 31;
  const int iL = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int nL = A.nL[img];
  if (iL >= nL) return;
  const size_t base = (size_t)img * A.kp_cap;
  const cuorb_keypoint kpL = A.kpsL[base + iL];
  const float uL = kpL.x, vL = kpL.y;
  const int levelL = kpL.octave;
  float out_u = -1.f, out_d = -1.f;
  int out_sad = -1, out_idx = -1, out_dist = -1;

  const int nRows = dg->lv[0].h;
  const float minZ = A.mb, minD = 0.f, maxD = __fdiv_rn(A.mbf, minZ);
  const float minU = __fsub_rn(uL, maxD), maxU = __fsub_rn(uL, minD);
  const int row = (int)vL;
  unsigned best = (unsigned)TH_HIGH << 16;
  if (row >= 0 && row < nRows && !(maxU < 0.f)) {
    const uint4* dl = reinterpret_cast<const uint4*>(A.descL + (base + iL) * 32);
    const uint4 l0 = dl[0], l1 = dl[1];
    const int nb = min((nRows + (1 << BUCKET_SHIFT) - 1) >> BUCKET_SHIFT, STEREO_MAX_BUCKETS);
    const int* gstart = A.rstart + (size_t)img * (STEREO_MAX_BUCKETS + 1);
    const int lo = gstart[max(row - band, 0) >> BUCKET_SHIFT];
    const int hi = gstart[min(((row + band) >> BUCKET_SHIFT) + 1, nb)];
    for (int j = lo + lane; j < hi; j += 32) {
      const uint4 e = A.rtab[base + j];
      const int minr = e.x & 0x1FFF, maxr = (e.x >> 13) & 0x1FFF, octR = e.x >> 26;
      if (row < minr || row > maxr) continue;                           // vRowIndices membership
      if (octR < levelL - 1 || octR > levelL + 1) continue;             // :1011
      const float uR = __uint_as_float(e.y);
      if (!(uR >= minU && uR <= maxU)) continue;                        // :1017
      const int iR = (int)e.z;
      const uint4* dr = reinterpret_cast<const uint4*>(A.descR + (base + iR) * 32);
      const uint4 r0 = dr[0], r1 = dr[1];
      const int dist = __popc(l0.x ^ r0.x) + __popc(l0.y ^ r0.y) + __popc(l0.z ^ r0.z) + __popc(l0.w ^ r0.w) +
                       __popc(l1.x ^ r1.x) + __popc(l1.y ^ r1.y) + __popc(l1.z ^ r1.z) + __popc(l1.w ^ r1.w);
      best = min(best, ((unsigned)dist << 16) | (unsigned)iR);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
  }
  const int bestDist = (int)(best >> 16);
  const int bestIdxR = (int)(best & 0xFFFF);
  if (bestDist < TH_HIGH) { out_idx = bestIdxR; out_dist = bestDist; }

  if (bestDist < (TH_HIGH + TH_LOW) / 2) {
    const LevelGeom& G = dg->lv[levelL];
    const float uR0 = A.kpsR[base + bestIdxR].x;
    const float sf = G.inv_scale;
    const float scaleduL = roundf(__fmul_rn(uL, sf)), scaledvL = roundf(__fmul_rn(vL, sf));
    const float scaleduR0 = roundf(__fmul_rn(uR0, sf));
    const int w = 5, Lw = 5;
    const float iniu = scaleduR0 + Lw - w, endu = scaleduR0 + Lw + w + 1;
    if (!(iniu < 0 || endu >= (float)G.w)) {
      // lane = 2*row + half: half 0 owns shifts 0..5, half 1 shifts 6..10 (its 6th value is unused)
      const int prow = min(lane >> 1, 10), half = lane & 1;
      unsigned part[6] = {0, 0, 0, 0, 0, 0};
      {
        const int y = (int)(scaledvL - w) + prow;
        const size_t rowoff = (size_t)img * dg->pyr_img_stride + G.pix_off + (ptrdiff_t)y * G.pitch;
        const uint8_t* pl = A.pyrL + rowoff + (int)(scaleduL - w);
        const uint8_t* pr = A.pyrR + rowoff + (int)(scaleduR0 - Lw - w) + 6 * half;
        const unsigned offl = (unsigned)((uintptr_t)pl & 3), offr = (unsigned)((uintptr_t)pr & 3);
        const unsigned* al = reinterpret_cast<const unsigned*>(pl - offl);
        const unsigned* ar = reinterpret_cast<const unsigned*>(pr - offr);
        const unsigned a0 = al[0], a1 = al[1], a2 = al[2], a3 = al[3];
        unsigned rw[6];
#pragma unroll
        for (int k = 0; k < 6; k++) rw[k] = ar[k];
        const unsigned L0 = __funnelshift_r(a0, a1, 8 * offl), L1 = __funnelshift_r(a1, a2, 8 * offl),
                       L2 = __funnelshift_r(a2, a3, 8 * offl) & 0x00FFFFFFu;
        unsigned R[5];
#pragma unroll
        for (int k = 0; k < 5; k++) R[k] = __funnelshift_r(rw[k], rw[k + 1], 8 * offr);
#pragma unroll
        for (int s = 0; s < 6; s++) {
          const int q = s >> 2, sh = 8 * (s & 3);
          const unsigned r0 = __funnelshift_r(R[q], R[q + 1], sh), r1 = __funnelshift_r(R[q + 1], R[q + 2], sh),
                         r2 = __funnelshift_r(R[q + 2], R[(q + 3) < 5 ? q + 3 : 4], sh) & 0x00FFFFFFu;
          part[s] = __vsadu4(L0, r0) + __vsadu4(L1, r1) + __vsadu4(L2, r2);
        }
        if (lane >= 22) {
#pragma unroll
          for (int s = 0; s < 6; s++) part[s] = 0;
        }
      }
#pragma unroll
      for (int s = 0; s < 6; s++) {
#pragma unroll
        for (int o = 2; o < 32; o <<= 1) part[s] += __shfl_xor_sync(0xffffffffu, part[s], o);
      }
      int sad[11];
#pragma unroll
      for (int k = 0; k < 11; k++) sad[k] = (int)__shfl_sync(0xffffffffu, part[k < 6 ? k : k - 6], k < 6 ? 0 : 1);
      int bestSad = 0x7fffffff, bestinc = 0;
#pragma unroll
      for (int k = 0; k < 11; k++)
        if (sad[k] < bestSad) { bestSad = sad[k]; bestinc = k - Lw; }
      if (!(bestinc == -Lw || bestinc == Lw)) {
        float d1 = 0.f, d2 = 0.f, d3 = 0.f;
#pragma unroll
        for (int k = 1; k < 10; k++)
          if (k - Lw == bestinc) { d1 = (float)sad[k - 1]; d2 = (float)sad[k]; d3 = (float)sad[k + 1]; }
        const float deltaR = __fdiv_rn(__fsub_rn(d1, d3),
                                       __fmul_rn(2.0f, __fsub_rn(__fadd_rn(d1, d3), __fmul_rn(2.0f, d2))));
        if (!(deltaR < -1.f || deltaR > 1.f)) {
          float bestuR = __fmul_rn(G.scale, __fadd_rn(__fadd_rn(scaleduR0, (float)bestinc), deltaR));
          float disparity = __fsub_rn(uL, bestuR);
          if (disparity >= minD && disparity < maxD) {
            if (disparity <= 0.f) {
              disparity = 0.01f;
              bestuR = (float)((double)uL - 0.01);
            }
            out_d = __fdiv_rn(A.mbf, disparity);
            out_u = bestuR;
            out_sad = bestSad;
          }
        }
      }
    }
  }
  if (lane == 0) {
    A.u_right[base + iL] = out_u;
    A.depth[base + iL] = out_d;
    A.sad[base + iL] = out_sad;
    A.best_idx_r[base + iL] = out_idx;
    A.best_dist[base + iL] = out_dist;
  }
}

// grid: n_img, block 256
__global__ void __launch_bounds__(256) k_stereo_outliers(StereoArgs A) {
  pdl_enter();
  __shared__ int h1[256], h2[128];
  __shared__ int s_cnt, s_b1, s_rem, s_med;
  const int img = blockIdx.x, tid = threadIdx.x;
  const int nL = A.nL[img];
  const size_t base = (size_t)img * A.kp_cap;
  h1[tid] = 0;
  if (tid < 128) h2[tid] = 0;
  if (tid == 0) s_cnt = 0;
  __syncthreads();
  int local = 0;
  for (int i = tid; i < nL; i += 256) {
    const int s = A.sad[base + i];
    if (s >= 0) { atomicAdd(&h1[min(s >> 7, 255)], 1); local++; }
  }
  if (local) atomicAdd(&s_cnt, local);
  __syncthreads();
  const int cnt = s_cnt;
  if (cnt == 0) return;  // reference: UB on an empty list (:1102); defined here as "nothing to cut"
  if (tid == 0) {
    int k = cnt / 2, cum = 0, b = 0;
    for (; b < 256; b++) { if (cum + h1[b] > k) break; cum += h1[b]; }
    s_b1 = b; s_rem = k - cum;
  }
  __syncthreads();
  const int b1 = s_b1;
  for (int i = tid; i < nL; i += 256) {
    const int s = A.sad[base + i];
    if (s >= 0 && min(s >> 7, 255) == b1) atomicAdd(&h2[s & 127], 1);
  }
  __syncthreads();
  if (tid == 0) {
    int k = s_rem, cum = 0, b = 0;
    for (; b < 128; b++) { if (cum + h2[b] > k) break; cum += h2[b]; }
    s_med = (b1 << 7) | b;
  }
  __syncthreads();
  const float thDist = __fmul_rn(__fmul_rn(1.5f, 1.4f), (float)s_med);
  for (int i = tid; i < nL; i += 256) {
    const int s = A.sad[base + i];
    if (s >= 0 && !((float)s < thDist)) { A.u_right[base + i] = -1.f; A.depth[base + i] = -1.f; }
  }
}

void launch_stereo(const ExtractorGeom& g, const StereoArgs& a, cudaStream_t st) {
  // a right keypoint can register in rows up to ceil(2*scale_max)+1 away from its own
  const int band = (int)ceilf(2.0f * g.lv[g.nlevels - 1].scale) + 2;
  // a one- or two-pair call is three kernels of a few microseconds each: dependent launches (common.cuh); the first
  // one follows a cross-stream event wait and keeps plain stream order
  struct Chain { bool saved; Chain(bool on) : saved(t_pdl) { t_pdl = on; } ~Chain() { t_pdl = saved; } } chain(a.n_img <= 2 && pdl_allowed());
  {
    PdlPause first;
    launch_k(k_stereo_rows, dim3(a.n_img), dim3(256), 0, st, a);
  }
  launch_k(k_stereo_match, dim3((g.kp_cap + 7) / 8, a.n_img), dim3(256), 0, st, a, band);
  launch_k(k_stereo_outliers, dim3(a.n_img), dim3(256), 0, st, a);
  count_launch(3);
}

}  // namespace cuorb
