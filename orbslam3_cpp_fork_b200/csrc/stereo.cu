// K10 — stereo association (replaces Frame::ComputeStereoMatches, src/Frame.cc:944-1113; the
// path BASELINE.json calls "SearchStereo").  Device-resident: consumes the keypoints, descriptors
// and pyramids the two extractor handles left in HBM.
//
// k_stereo_match: one warp per left keypoint.  The reference's per-row candidate table
//   (:953-971) is replaced by evaluating its membership predicate directly
//   (floor(yR - r) <= int(vL) <= ceil(yR + r), r = 2*scale[octR]) for every right keypoint: lanes
//   stride over iR, so the ascending-iR visiting order and the strict "<" of :1021 become a warp
//   arg-min on the key (dist << 16 | iR).  Hamming = 8 x (XOR, POPC) per pair.  The winner is
//   refined by the 11-window SAD slide on the un-blurred pyramid level (:1028-1067, lanes 0..10 own
//   one patch row each) and the parabola fit (:1074-1098) in FMA-free float.
// k_stereo_outliers: one CTA per image.  median = element size/2 of the sorted SAD list (:1101-1103)
//   by a two-level histogram select; matches with SAD >= 1.5*1.4*median are reset (:1105-1112).
#include "common.cuh"

namespace cuorb {

constexpr int TH_HIGH = 100, TH_LOW = 50;  // src/ORBmatcher.cc:30-31

// grid: (ceil(kp_cap/8), n_img), block 256
__global__ void __launch_bounds__(256) k_stereo_match(StereoArgs A) {
  const ExtractorGeom* dg = A.dgL;
  const int img = blockIdx.y, lane = threadIdx.x & 31;
  const int iL = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int nL = A.nL[img], nR = A.nR[img];
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
    for (int iR = lane; iR < nR; iR += 32) {
      const cuorb_keypoint* kr = A.kpsR + base + iR;
      const int octR = kr->octave;
      if (octR < levelL - 1 || octR > levelL + 1) continue;
      const float yR = kr->y;
      const float r = __fmul_rn(2.0f, dg->lv[octR].scale);
      const int maxr = (int)ceilf(__fadd_rn(yR, r)), minr = (int)floorf(__fsub_rn(yR, r));
      if (row < minr || row > maxr) continue;
      const float uR = kr->x;
      if (!(uR >= minU && uR <= maxU)) continue;
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
      int sad[11];
#pragma unroll
      for (int k = 0; k < 11; k++) sad[k] = 0;
      if (lane < 11) {
        const int y = (int)(scaledvL - w) + lane;
        const uint8_t* pl = A.pyrL + (size_t)img * dg->pyr_img_stride + G.pix_off + (ptrdiff_t)y * G.pitch +
                            (int)(scaleduL - w);
        const uint8_t* pr = A.pyrR + (size_t)img * dg->pyr_img_stride + G.pix_off + (ptrdiff_t)y * G.pitch +
                            (int)(scaleduR0 - Lw - w);
        int lv[11], rv[21];
#pragma unroll
        for (int k = 0; k < 11; k++) lv[k] = pl[k];
#pragma unroll
        for (int k = 0; k < 21; k++) rv[k] = pr[k];
#pragma unroll
        for (int k = 0; k < 11; k++) {
          int s = 0;
#pragma unroll
          for (int x = 0; x < 11; x++) s += abs(lv[x] - rv[k + x]);
          sad[k] = s;
        }
      }
#pragma unroll
      for (int k = 0; k < 11; k++) {
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) sad[k] += __shfl_xor_sync(0xffffffffu, sad[k], o);
      }
      // lanes 0..15 now hold the totals (lanes 11..15 contributed zeros)
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
  k_stereo_match<<<dim3((g.kp_cap + 7) / 8, a.n_img), 256, 0, st>>>(a);
  k_stereo_outliers<<<a.n_img, 256, 0, st>>>(a);
  count_launch(2);
}

}  // namespace cuorb
