// K13 — the ORBmatcher searches outside the per-frame tracking path, pinhole /
// NLeft == -1 branches, on flattened inputs.  The caller keeps the Sophus/Eigen/MapPoint work (pose
// transform, projection, distance-invariance range, viewing angle, PredictScale) and passes, per
// query, (active, u, v, predicted level, descriptor); everything after that runs here:
//   SearchByProjection(Frame&, KeyFrame*, sAlreadyFound, th, ORBdist)     src/ORBmatcher.cc:1910-2036
//   SearchByProjection(KeyFrame*, Sim3, vpPoints, vpMatched[, KFs], ...)  src/ORBmatcher.cc:423-538, 540-664
//   Fuse(KeyFrame*, vpMapPoints, th) / Fuse(KeyFrame*, Sim3, ...)          src/ORBmatcher.cc:1149-1328, 1330-1452
//   SearchBySim3(KeyFrame*, KeyFrame*, vpMatches12, S12, th)               src/ORBmatcher.cc:1454-1682
//   SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, window)    src/ORBmatcher.cc:666-782
//   SearchForTriangulation(KF1, KF2, vMatchedPairs, bOnlyStereo, bCoarse)  src/ORBmatcher.cc:916-1147
//   KeyFrame::GetFeaturesInArea                                            src/KeyFrame.cc:780-834
//   Pinhole::epipolarConstrain                                             src/CameraModels/Pinhole.cpp:143-172
//
// k_area_round  one warp per query: brute-force scan of the target's features with the exact
//               GetFeaturesInArea predicate; the arg-min key carries the reference's visiting order
//               (cell ix, cell iy, feature index) so "first strict minimum wins" is reproduced.
//               Searches whose loop skips features bound earlier in the same call (reloc, Sim3) are
//               resolved by the fixed-point rounds of project.cu (binder = earliest query index).
// k_init_candidates / k_init_resolve  SearchForInitialization: the gate vMatchedDistance[i2] <= dist
//               depends on every earlier acceptance, so candidates (with distances) are listed in
//               parallel and ONE warp replays the reference loop over them in i1 order.
// k_triangulate one warp per side-1 feature of a common vocabulary node; the reference loop is order
//               independent (vbMatched2 is never set, :963/:1015), last candidate wins ties.
// Float expressions use explicit _rn intrinsics in the reference's evaluation order (no FMA contraction).
#include <algorithm>

#include "project.cuh"

namespace cuorb {

namespace {

constexpr int HISTO_LENGTH = 30, TH_LOW_S = 50;

struct AreaQuery {
  int n;
  const uint8_t* desc;
  const float *u, *v, *ur;
  const int* level;
  const uint8_t* active;
  float th;            // radius = th * scale[level]
  int lo, hi;          // accepted octaves: level + lo .. level + hi
  float max_dist;      // accept when (float)bestDist <= max_dist
  int skip_bound;      // skip features bound on entry or by an earlier query of this call
  int bounds_check;    // u, v must lie inside [min, max] of the frame (:1937-1942)
  int chi2;            // Fuse reprojection gate (:1257-1283)
  float inv_sigma2[MAX_LEVELS];
};

// grid: ceil(n/8), block 256 (warp per query)
__global__ void __launch_bounds__(256) k_area_round(ProjFrame F, AreaQuery Q, const int* __restrict__ binder_prev,
                                                    int* __restrict__ binder_next, int* __restrict__ result,
                                                    int* __restrict__ changed) {
  const int i = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (i >= Q.n) return;
  int res = -1;
  bool active = Q.active[i] != 0;
  float x = 0, y = 0, rad = 0;
  int lvl = 0;
  if (active) {
    x = Q.u[i]; y = Q.v[i]; lvl = Q.level[i];
    if (Q.bounds_check && (x < F.minX || x > F.maxX || y < F.minY || y > F.maxY)) active = false;
    rad = __fmul_rn(Q.th, F.scale[lvl]);
  }
  int cx0 = 0, cx1 = -1, cy0 = 0, cy1 = -1;
  if (active) {
    cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, F.minX), rad), F.gwinv)));
    cx1 = min(FRAME_GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, F.minX), rad), F.gwinv)));
    cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, F.minY), rad), F.ghinv)));
    cy1 = min(FRAME_GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, F.minY), rad), F.ghinv)));
    if (cx0 >= FRAME_GRID_COLS || cx1 < 0 || cy0 >= FRAME_GRID_ROWS || cy1 < 0) active = false;
  }
  if (active) {
    const int minLevel = lvl + Q.lo, maxLevel = lvl + Q.hi;
    const uint4* qd = reinterpret_cast<const uint4*>(Q.desc + (size_t)i * 32);
    const uint4 q0 = qd[0], q1 = qd[1];
    const float ur = (Q.chi2 && Q.ur) ? Q.ur[i] : 0.f;
    unsigned long long b1 = ~0ull;
    // the window's cells through the frame's cell-sorted CSR (project.cuh); the visiting order is in the key
    for (int cy = cy0; cy <= cy1; cy++) {
     const int js = F.cstart[cy * FRAME_GRID_COLS + cx0], je = F.cstart[cy * FRAME_GRID_COLS + cx1 + 1];
     for (int j = js + lane; j < je; j += 32) {
      const int f = F.order[j];
      const int cx = F.cellx[f];
      const float kx = F.x[f], ky = F.y[f];
      const float ex = __fsub_rn(kx, x), ey = __fsub_rn(ky, y);
      if (!(fabsf(ex) < rad && fabsf(ey) < rad)) continue;
      if (Q.skip_bound && (F.blocked0[f] || binder_prev[f] < i)) continue;
      const int oct = F.octave[f];
      if (oct < minLevel || oct > maxLevel) continue;
      if (Q.chi2) {
        const float dx = __fsub_rn(x, kx), dy = __fsub_rn(y, ky);
        float e2 = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
        const float kr = F.uright ? F.uright[f] : -1.f;
        if (kr >= 0.f) {
          const float er = __fsub_rn(ur, kr);
          e2 = __fadd_rn(e2, __fmul_rn(er, er));
          if ((double)__fmul_rn(e2, Q.inv_sigma2[oct]) > 7.8) continue;
        } else if ((double)__fmul_rn(e2, Q.inv_sigma2[oct]) > 5.99) continue;
      }
      const int dist = hamming256(q0, q1, reinterpret_cast<const uint4*>(F.desc + (size_t)f * 32));
      const unsigned long long key = ((unsigned long long)dist << 32) | ((unsigned)cx << 22) | ((unsigned)cy << 16) | (unsigned)f;
      b1 = min(b1, key);
     }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) b1 = min(b1, __shfl_xor_sync(0xffffffffu, b1, o));
    if (b1 != ~0ull && (float)(int)(b1 >> 32) <= Q.max_dist) res = (int)(b1 & 0xFFFF);
  }
  if (lane == 0) {
    if (result[i] != res) { result[i] = res; *changed = 1; }
    if (res >= 0 && Q.skip_bound) atomicMin(&binder_next[res], i);
  }
}

// ---- SearchForInitialization ----
// grid: ceil(nq/8), block 256.  Query q = level-0 feature qidx[q] of F1; candidates = level-0 features of
// F2 inside the window around vbPrevMatched; entry = dist << 32 | visiting order (cx, cy, f).
__global__ void __launch_bounds__(256) k_init_candidates(ProjFrame F2, int nq, const int* __restrict__ qidx,
                                                         const uint8_t* __restrict__ desc1, const float* __restrict__ prev,
                                                         float r, int cap, unsigned long long* __restrict__ cand,
                                                         int* __restrict__ cand_n) {
  const int q = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (q >= nq) return;
  const int i1 = qidx[q];
  const float x = prev[2 * i1], y = prev[2 * i1 + 1];
  const int cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, F2.minX), r), F2.gwinv)));
  const int cx1 = min(FRAME_GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, F2.minX), r), F2.gwinv)));
  const int cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, F2.minY), r), F2.ghinv)));
  const int cy1 = min(FRAME_GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, F2.minY), r), F2.ghinv)));
  int cnt = 0;
  if (!(cx0 >= FRAME_GRID_COLS || cx1 < 0 || cy0 >= FRAME_GRID_ROWS || cy1 < 0)) {
    const uint4* qd = reinterpret_cast<const uint4*>(desc1 + (size_t)i1 * 32);
    const uint4 q0 = qd[0], q1 = qd[1];
    unsigned long long* seg = cand + (size_t)q * cap;
    const unsigned lt = (1u << lane) - 1u;
    for (int f0 = 0; f0 < F2.n; f0 += 32) {
      const int f = f0 + lane;
      bool ok = f < F2.n;
      int cx = 0, cy = 0;
      if (ok) {
        cx = F2.cellx[f]; cy = F2.celly[f];
        ok = !(cx < cx0 || cx > cx1 || cy < cy0 || cy > cy1) && F2.octave[f] == 0 &&
             fabsf(__fsub_rn(F2.x[f], x)) < r && fabsf(__fsub_rn(F2.y[f], y)) < r;
      }
      const unsigned bal = __ballot_sync(0xffffffffu, ok);
      if (ok) {
        const int dist = hamming256(q0, q1, reinterpret_cast<const uint4*>(F2.desc + (size_t)f * 32));
        seg[cnt + __popc(bal & lt)] = ((unsigned long long)dist << 32) | ((unsigned)cx << 22) | ((unsigned)cy << 16) | (unsigned)f;
      }
      cnt += __popc(bal);
    }
  }
  if (lane == 0) cand_n[q] = cnt;
}

// one warp: the reference loop over i1 (:684-744) on the listed candidates
__global__ void __launch_bounds__(32) k_init_resolve(int nq, const int* __restrict__ qidx, int cap,
                                                     const unsigned long long* __restrict__ cand,
                                                     const int* __restrict__ cand_n, float nnratio,
                                                     int* __restrict__ matched_dist, int* __restrict__ m21,
                                                     int* __restrict__ m12, int* __restrict__ accepted) {
  const int lane = threadIdx.x;
  for (int q = 0; q < nq; q++) {
    const int n = cand_n[q];
    const unsigned long long* seg = cand + (size_t)q * cap;
    unsigned long long b1 = ~0ull, b2 = ~0ull;
    for (int c = lane; c < n; c += 32) {
      const unsigned long long key = seg[c];
      const int dist = (int)(key >> 32), f = (int)(key & 0xFFFF);
      if (matched_dist[f] <= dist) continue;
      if (key < b1) { b2 = b1; b1 = key; }
      else if (key < b2) b2 = key;
    }
    unsigned long long B1 = b1;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) B1 = min(B1, __shfl_xor_sync(0xffffffffu, B1, o));
    unsigned long long B2 = (b1 == B1) ? b2 : b1;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) B2 = min(B2, __shfl_xor_sync(0xffffffffu, B2, o));
    if (lane == 0 && B1 != ~0ull) {
      const int bestDist = (int)(B1 >> 32), best = (int)(B1 & 0xFFFF);
      const float second = B2 != ~0ull ? (float)(int)(B2 >> 32) : 2147483648.0f;   // (float)INT_MAX
      if (bestDist <= TH_LOW_S && (float)bestDist < __fmul_rn(second, nnratio)) {
        const int i1 = qidx[q];
        const int old = m21[best];
        if (old >= 0) m12[old] = -1;
        m12[i1] = best; m21[best] = i1; matched_dist[best] = bestDist;
        accepted[i1] = best;
      }
    }
    __syncwarp();
  }
}

// ---- SearchForTriangulation ----
struct TriArgs {
  ProjFrame A, B;             // A.blocked0 / B.blocked0 = GetMapPoint(idx) != NULL
  int n_entries;
  const int *entry_idx1, *entry_node, *off2, *idx2;
  float F12[9], ep[2], sigma2_2[MAX_LEVELS];
  int only_stereo, coarse;
};

// grid: ceil(n_entries/8), block 256 (warp per side-1 entry)
__global__ void __launch_bounds__(256) k_triangulate(TriArgs T, int* __restrict__ match12) {
  const int e = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (e >= T.n_entries) return;
  const int idx1 = T.entry_idx1[e];
  if (T.A.blocked0[idx1]) return;
  const bool stereo1 = T.A.uright && T.A.uright[idx1] >= 0.f;
  if (T.only_stereo && !stereo1) return;
  const float x1 = T.A.x[idx1], y1 = T.A.y[idx1];
  const float a = __fadd_rn(__fadd_rn(__fmul_rn(x1, T.F12[0]), __fmul_rn(y1, T.F12[3])), T.F12[6]);
  const float b = __fadd_rn(__fadd_rn(__fmul_rn(x1, T.F12[1]), __fmul_rn(y1, T.F12[4])), T.F12[7]);
  const float c = __fadd_rn(__fadd_rn(__fmul_rn(x1, T.F12[2]), __fmul_rn(y1, T.F12[5])), T.F12[8]);
  const float den = __fadd_rn(__fmul_rn(a, a), __fmul_rn(b, b));
  const uint4* qd = reinterpret_cast<const uint4*>(T.A.desc + (size_t)idx1 * 32);
  const uint4 q0 = qd[0], q1 = qd[1];
  const int nd = T.entry_node[e], lo = T.off2[nd], hi = T.off2[nd + 1];
  unsigned best = 0xFFFFFFFFu;   // dist << 16 | (0xFFFF - position): smallest distance, LAST position on ties
  for (int p = lo + lane; p < hi; p += 32) {
    const int idx2 = T.idx2[p];
    if (T.B.blocked0[idx2]) continue;
    const bool stereo2 = T.B.uright && T.B.uright[idx2] >= 0.f;
    if (T.only_stereo && !stereo2) continue;
    const int dist = hamming256(q0, q1, reinterpret_cast<const uint4*>(T.B.desc + (size_t)idx2 * 32));
    if (dist > TH_LOW_S) continue;
    const float x2 = T.B.x[idx2], y2 = T.B.y[idx2];
    const int oct2 = T.B.octave[idx2];
    if (!stereo1 && !stereo2) {
      const float dx = __fsub_rn(T.ep[0], x2), dy = __fsub_rn(T.ep[1], y2);
      if (__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)) < __fmul_rn(100.0f, T.B.scale[oct2])) continue;
    }
    if (!T.coarse) {
      if (den == 0.f) continue;
      const float num = __fadd_rn(__fadd_rn(__fmul_rn(a, x2), __fmul_rn(b, y2)), c);
      const float dsqr = __fdiv_rn(__fmul_rn(num, num), den);
      if (!((double)dsqr < 3.84 * (double)T.sigma2_2[oct2])) continue;
    }
    best = min(best, ((unsigned)dist << 16) | (unsigned)(0xFFFF - (p - lo)));
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
  if (lane == 0 && best != 0xFFFFFFFFu) match12[idx1] = T.idx2[lo + (0xFFFF - (int)(best & 0xFFFF))];
}

// ---- host side ----
int check_view(const cuorb_frame_view* f, const char* who) {
  if (!f || f->n < 0 || (f->n && (!f->keys_un || !f->descriptors)) || !f->scale_factors) {
    set_error(std::string(who) + ": null argument");
    return CUORB_E_BADARG;
  }
  if (f->n > 65535 || f->nlevels > MAX_LEVELS) {
    set_error(std::string(who) + ": frame too large for the window searches");
    return CUORB_E_UNSUPPORTED;
  }
  return 0;
}

// runs k_area_round (to its fixed point when the search binds features); Q holds HOST pointers on entry
int run_area(int device, const cuorb_frame_view* fv, AreaQuery Q, int32_t* result_out) {
  const int n = fv->n, m = Q.n;
  for (int i = 0; i < m; i++) result_out[i] = -1;
  if (n <= 0 || m <= 0) return 0;
  CUORB_CUDA_TRY(cudaSetDevice(device));
  const size_t bytes = frame_bytes(n) + (size_t)m * (32 + 4 * 4 + 1 + 4) + (size_t)n * 8 + 64 * 1024;
  if (t_arena.reserve(device, bytes) != 0) { set_error("cudaMalloc failed (search workspace)"); return CUORB_E_CUDA; }
  Arena& A = t_arena;
  cudaStream_t st = cudaStreamPerThread;
  ProjFrame F;
  upload_frame(A, fv, st, F);
  Q.desc = up(A, Q.desc, (size_t)m * 32, st);
  Q.u = up(A, Q.u, m, st); Q.v = up(A, Q.v, m, st);
  Q.ur = Q.ur ? up(A, Q.ur, m, st) : nullptr;
  Q.level = up(A, Q.level, m, st);
  Q.active = up(A, Q.active, m, st);
  int* binder[2] = {A.take<int>(n), A.take<int>(n)};
  int* result = A.take<int>(m);
  // rounds in batches, as in run_projection (project.cu): one host sync per batch; rounds past the fixed point are
  // idempotent and simply run
  constexpr int ROUND_BATCH = 6;
  int* flag = A.take<int>(ROUND_BATCH);
  k_fill<<<(n + 255) / 256, 256, 0, st>>>(binder[0], n, NO_BINDER);
  k_fill<<<(m + 255) / 256, 256, 0, st>>>(result, m, -2);
  count_launch(2);
  int rounds = 0;
  for (bool converged = false; !converged;) {
    if (rounds > m + 2) { set_error("window search did not converge"); return CUORB_E_CUDA; }
    CUORB_CUDA_TRY(cudaMemsetAsync(flag, 0, ROUND_BATCH * sizeof(int), st));
    const int nb = Q.skip_bound ? ROUND_BATCH : 1;   // no query depends on another one: one pass is the answer
    for (int k = 0; k < nb; k++, rounds++) {
      int* prev = binder[rounds & 1];
      int* next = binder[(rounds + 1) & 1];
      k_fill<<<(n + 255) / 256, 256, 0, st>>>(next, n, NO_BINDER);
      k_area_round<<<(m + 7) / 8, 256, 0, st>>>(F, Q, prev, next, result, flag + k);
      count_launch(2);
    }
    if (!Q.skip_bound) break;
    int h_flag[ROUND_BATCH];
    CUORB_CUDA_TRY(cudaMemcpyAsync(h_flag, flag, sizeof(h_flag), cudaMemcpyDeviceToHost, st));
    CUORB_CUDA_TRY(cudaStreamSynchronize(st));
    for (int k = 0; k < ROUND_BATCH; k++)
      if (!h_flag[k]) { converged = true; break; }
  }
  CUORB_CUDA_TRY(cudaMemcpyAsync(result_out, result, sizeof(int) * m, cudaMemcpyDeviceToHost, st));
  CUORB_CUDA_TRY(cudaStreamSynchronize(st));
  CUORB_CUDA_TRY(cudaGetLastError());
  return 0;
}

int rot_bin(float a1, float a2) {
  float rot = a1 - a2;
  if (rot < 0.0) rot += 360.0f;
  int bin = (int)std::round(rot * (1.0f / HISTO_LENGTH));
  if (bin == HISTO_LENGTH) bin = 0;
  return bin;
}

// ORBmatcher::ComputeThreeMaxima (src/ORBmatcher.cc:2038-2071) on bin sizes
void three_maxima(const int* hist, int keep[3]) {
  int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
  for (int i = 0; i < HISTO_LENGTH; i++) {
    const int s = hist[i];
    if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
    else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
    else if (s > max3) { max3 = s; ind3 = i; }
  }
  if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
  else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
  keep[0] = ind1; keep[1] = ind2; keep[2] = ind3;
}

bool bad_bin(int bin) {
  if (bin >= 0 && bin < HISTO_LENGTH) return false;
  set_error("keypoint angle outside [0,360)");
  return true;
}

}  // namespace

}  // namespace cuorb

using namespace cuorb;

extern "C" {

int cuorb_search_by_projection_reloc(int device, const cuorb_frame_view* cur, int n_points,
                                     const uint8_t* mp_desc, const float* u, const float* v,
                                     const int32_t* level, const float* kf_angle, const uint8_t* active,
                                     float th, int orb_dist, int check_orientation, int32_t* best_idx_out,
                                     int32_t* rot_bin_out, int32_t* keep_bins_out) {
  if (int rc = check_view(cur, "cuorb_search_by_projection_reloc")) return rc;
  if (n_points < 0 || !keep_bins_out ||
      (n_points && (!mp_desc || !u || !v || !level || !active || !best_idx_out || !rot_bin_out ||
                    (check_orientation && !kf_angle)))) {
    set_error("cuorb_search_by_projection_reloc: null argument");
    return CUORB_E_BADARG;
  }
  AreaQuery Q{};
  Q.n = n_points; Q.desc = mp_desc; Q.u = u; Q.v = v; Q.level = level; Q.active = active;
  Q.th = th; Q.lo = -1; Q.hi = 1; Q.max_dist = (float)std::min(orb_dist, 255); Q.skip_bound = 1; Q.bounds_check = 1;
  if (int rc = run_area(device, cur, Q, best_idx_out)) return rc;
  int hist[HISTO_LENGTH] = {0}, nmatches = 0;
  for (int i = 0; i < n_points; i++) {
    rot_bin_out[i] = -1;
    if (best_idx_out[i] < 0) continue;
    nmatches++;
    if (check_orientation) {
      const int bin = rot_bin(kf_angle[i], cur->keys_un[best_idx_out[i]].angle);
      if (bad_bin(bin)) return CUORB_E_BADARG;
      hist[bin]++; rot_bin_out[i] = bin;
    }
  }
  keep_bins_out[0] = keep_bins_out[1] = keep_bins_out[2] = -1;
  if (check_orientation) {
    three_maxima(hist, keep_bins_out);
    for (int b = 0; b < HISTO_LENGTH; b++)
      if (b != keep_bins_out[0] && b != keep_bins_out[1] && b != keep_bins_out[2]) nmatches -= hist[b];
  }
  return nmatches;
}

int cuorb_search_by_projection_sim3(int device, const cuorb_frame_view* kf, int n_points,
                                    const uint8_t* mp_desc, const float* u, const float* v,
                                    const int32_t* level, const uint8_t* active, int th, float ratio_hamming,
                                    int32_t* best_idx_out) {
  if (int rc = check_view(kf, "cuorb_search_by_projection_sim3")) return rc;
  if (n_points < 0 || (n_points && (!mp_desc || !u || !v || !level || !active || !best_idx_out))) {
    set_error("cuorb_search_by_projection_sim3: null argument");
    return CUORB_E_BADARG;
  }
  AreaQuery Q{};
  Q.n = n_points; Q.desc = mp_desc; Q.u = u; Q.v = v; Q.level = level; Q.active = active;
  Q.th = (float)th; Q.lo = -1; Q.hi = 0; Q.max_dist = (float)TH_LOW_S * ratio_hamming; Q.skip_bound = 1;
  if (int rc = run_area(device, kf, Q, best_idx_out)) return rc;
  int nmatches = 0;
  for (int i = 0; i < n_points; i++) nmatches += best_idx_out[i] >= 0;
  return nmatches;
}

int cuorb_fuse(int device, const cuorb_frame_view* kf, const float* inv_level_sigma2, int n_points,
               const uint8_t* mp_desc, const float* u, const float* v, const float* ur, const int32_t* level,
               const uint8_t* active, float th, int check_reprojection, int32_t* best_idx_out) {
  if (int rc = check_view(kf, "cuorb_fuse")) return rc;
  if (n_points < 0 || (check_reprojection && !inv_level_sigma2) ||
      (n_points && (!mp_desc || !u || !v || !level || !active || !best_idx_out ||
                    (check_reprojection && kf->u_right && !ur)))) {
    set_error("cuorb_fuse: null argument");
    return CUORB_E_BADARG;
  }
  AreaQuery Q{};
  Q.n = n_points; Q.desc = mp_desc; Q.u = u; Q.v = v; Q.ur = ur; Q.level = level; Q.active = active;
  Q.th = th; Q.lo = -1; Q.hi = 0; Q.max_dist = (float)TH_LOW_S; Q.chi2 = check_reprojection != 0;
  if (check_reprojection)
    for (int l = 0; l < kf->nlevels; l++) Q.inv_sigma2[l] = inv_level_sigma2[l];
  if (int rc = run_area(device, kf, Q, best_idx_out)) return rc;
  int nfused = 0;
  for (int i = 0; i < n_points; i++) nfused += best_idx_out[i] >= 0;
  return nfused;
}

int cuorb_search_by_sim3(int device, const cuorb_frame_view* kf1, const cuorb_frame_view* kf2,
                         const uint8_t* mp_desc1, const float* u12, const float* v12, const int32_t* level12,
                         const uint8_t* active1, const uint8_t* mp_desc2, const float* u21, const float* v21,
                         const int32_t* level21, const uint8_t* active2, float th, int32_t* match12_out) {
  if (int rc = check_view(kf1, "cuorb_search_by_sim3")) return rc;
  if (int rc = check_view(kf2, "cuorb_search_by_sim3")) return rc;
  if ((kf1->n && (!mp_desc1 || !u12 || !v12 || !level12 || !active1 || !match12_out)) ||
      (kf2->n && (!mp_desc2 || !u21 || !v21 || !level21 || !active2))) {
    set_error("cuorb_search_by_sim3: null argument");
    return CUORB_E_BADARG;
  }
  std::vector<int32_t> m1(std::max(kf1->n, 1)), m2(std::max(kf2->n, 1));
  AreaQuery Q{};
  Q.th = th; Q.lo = -1; Q.hi = 0; Q.max_dist = 100.0f;   // TH_HIGH
  Q.n = kf1->n; Q.desc = mp_desc1; Q.u = u12; Q.v = v12; Q.level = level12; Q.active = active1;
  if (int rc = run_area(device, kf2, Q, m1.data())) return rc;
  Q.n = kf2->n; Q.desc = mp_desc2; Q.u = u21; Q.v = v21; Q.level = level21; Q.active = active2;
  if (int rc = run_area(device, kf1, Q, m2.data())) return rc;
  int found = 0;
  for (int i1 = 0; i1 < kf1->n; i1++) {   // agreement check (:1667-1679)
    match12_out[i1] = -1;
    const int idx2 = m1[i1];
    if (idx2 >= 0 && m2[idx2] == i1) { match12_out[i1] = idx2; found++; }
  }
  return found;
}

int cuorb_search_for_initialization(int device, const cuorb_frame_view* f1, const cuorb_frame_view* f2,
                                    float* prev_matched, int window_size, float nnratio,
                                    int check_orientation, int32_t* matches12_out) {
  if (int rc = check_view(f1, "cuorb_search_for_initialization")) return rc;
  if (int rc = check_view(f2, "cuorb_search_for_initialization")) return rc;
  if (f1->n && (!prev_matched || !matches12_out)) {
    set_error("cuorb_search_for_initialization: null argument");
    return CUORB_E_BADARG;
  }
  const int n1 = f1->n, n2 = f2->n;
  for (int i = 0; i < n1; i++) matches12_out[i] = -1;
  std::vector<int> qidx;
  for (int i = 0; i < n1; i++)
    if (f1->keys_un[i].octave <= 0) qidx.push_back(i);      // level1 > 0 -> continue (:687-689)
  int cap = 0;
  for (int i = 0; i < n2; i++) cap += f2->keys_un[i].octave == 0;
  const int nq = (int)qidx.size();
  if (nq == 0 || cap == 0) return 0;
  if ((size_t)nq * cap > ((size_t)1 << 27)) { set_error("cuorb_search_for_initialization: too many candidates"); return CUORB_E_UNSUPPORTED; }
  CUORB_CUDA_TRY(cudaSetDevice(device));
  const size_t bytes = frame_bytes(n2) + (size_t)n1 * (32 + 8 + 8) + (size_t)nq * (8 + (size_t)cap * 8) +
                       (size_t)n2 * 8 + 64 * 1024;
  if (t_arena.reserve(device, bytes) != 0) { set_error("cudaMalloc failed (search workspace)"); return CUORB_E_CUDA; }
  Arena& A = t_arena;
  cudaStream_t st = cudaStreamPerThread;
  ProjFrame F2;
  upload_frame(A, f2, st, F2);
  const uint8_t* d_desc1 = up(A, f1->descriptors, (size_t)n1 * 32, st);
  const float* d_prev = up(A, prev_matched, (size_t)2 * n1, st);
  const int* d_qidx = up(A, qidx.data(), nq, st);
  unsigned long long* d_cand = A.take<unsigned long long>((size_t)nq * cap);
  int* d_cand_n = A.take<int>(nq);
  int* d_md = A.take<int>(n2);
  int* d_m21 = A.take<int>(n2);
  int* d_m12 = A.take<int>(n1);
  int* d_acc = A.take<int>(n1);
  k_fill<<<(n2 + 255) / 256, 256, 0, st>>>(d_md, n2, 0x7fffffff);
  k_fill<<<(n2 + 255) / 256, 256, 0, st>>>(d_m21, n2, -1);
  k_fill<<<(n1 + 255) / 256, 256, 0, st>>>(d_m12, n1, -1);
  k_fill<<<(n1 + 255) / 256, 256, 0, st>>>(d_acc, n1, -1);
  k_init_candidates<<<(nq + 7) / 8, 256, 0, st>>>(F2, nq, d_qidx, d_desc1, d_prev, (float)window_size, cap, d_cand, d_cand_n);
  k_init_resolve<<<1, 32, 0, st>>>(nq, d_qidx, cap, d_cand, d_cand_n, nnratio, d_md, d_m21, d_m12, d_acc);
  count_launch(6);
  std::vector<int> acc(n1);
  CUORB_CUDA_TRY(cudaMemcpyAsync(matches12_out, d_m12, sizeof(int) * n1, cudaMemcpyDeviceToHost, st));
  CUORB_CUDA_TRY(cudaMemcpyAsync(acc.data(), d_acc, sizeof(int) * n1, cudaMemcpyDeviceToHost, st));
  CUORB_CUDA_TRY(cudaStreamSynchronize(st));
  CUORB_CUDA_TRY(cudaGetLastError());
  if (check_orientation) {   // every acceptance entered the histogram, stolen matches included (:733-744)
    int hist[HISTO_LENGTH] = {0}, keep[3];
    std::vector<int> bin(n1, -1);
    for (int i = 0; i < n1; i++)
      if (acc[i] >= 0) {
        bin[i] = rot_bin(f1->keys_un[i].angle, f2->keys_un[acc[i]].angle);
        if (bad_bin(bin[i])) return CUORB_E_BADARG;
        hist[bin[i]]++;
      }
    three_maxima(hist, keep);
    for (int i = 0; i < n1; i++)
      if (bin[i] >= 0 && bin[i] != keep[0] && bin[i] != keep[1] && bin[i] != keep[2]) matches12_out[i] = -1;
  }
  int nmatches = 0;
  for (int i = 0; i < n1; i++)
    if (matches12_out[i] >= 0) {   // update prev matched (:771-776)
      prev_matched[2 * i] = f2->keys_un[matches12_out[i]].x;
      prev_matched[2 * i + 1] = f2->keys_un[matches12_out[i]].y;
      nmatches++;
    }
  return nmatches;
}

int cuorb_search_for_triangulation(int device, const cuorb_frame_view* kf1, const cuorb_frame_view* kf2,
                                   int n_nodes, const int32_t* node_off1, const int32_t* node_idx1,
                                   const int32_t* node_off2, const int32_t* node_idx2, const float* F12,
                                   const float* epipole, const float* level_sigma2_2, int only_stereo,
                                   int coarse, int check_orientation, int32_t* match12_out) {
  if (int rc = check_view(kf1, "cuorb_search_for_triangulation")) return rc;
  if (int rc = check_view(kf2, "cuorb_search_for_triangulation")) return rc;
  if (n_nodes < 0 || !F12 || !epipole || !level_sigma2_2 || (kf1->n && !match12_out) ||
      (n_nodes && (!node_off1 || !node_idx1 || !node_off2 || !node_idx2))) {
    set_error("cuorb_search_for_triangulation: null argument");
    return CUORB_E_BADARG;
  }
  const int n1 = kf1->n, n2 = kf2->n;
  for (int i = 0; i < n1; i++) match12_out[i] = -1;
  if (n_nodes == 0 || n1 == 0 || n2 == 0) return 0;
  const int e1 = node_off1[n_nodes], e2 = node_off2[n_nodes];
  std::vector<int> entry_node(std::max(e1, 1));
  for (int k = 0; k < n_nodes; k++) {
    if (node_off2[k + 1] - node_off2[k] > 65535) { set_error("cuorb_search_for_triangulation: node too large"); return CUORB_E_UNSUPPORTED; }
    for (int a = node_off1[k]; a < node_off1[k + 1]; a++) {
      if (node_idx1[a] < 0 || node_idx1[a] >= n1) { set_error("cuorb_search_for_triangulation: index out of range"); return CUORB_E_BADARG; }
      entry_node[a] = k;
    }
  }
  for (int b = 0; b < e2; b++)
    if (node_idx2[b] < 0 || node_idx2[b] >= n2) { set_error("cuorb_search_for_triangulation: index out of range"); return CUORB_E_BADARG; }
  if (e1 == 0 || e2 == 0) return 0;
  CUORB_CUDA_TRY(cudaSetDevice(device));
  const size_t bytes = frame_bytes(n1) + frame_bytes(n2) + (size_t)(2 * e1 + e2 + n_nodes + 1 + n1) * 4 + 64 * 1024;
  if (t_arena.reserve(device, bytes) != 0) { set_error("cudaMalloc failed (search workspace)"); return CUORB_E_CUDA; }
  Arena& A = t_arena;
  cudaStream_t st = cudaStreamPerThread;
  TriArgs T{};
  upload_frame(A, kf1, st, T.A);
  upload_frame(A, kf2, st, T.B);
  T.n_entries = e1;
  T.entry_idx1 = up(A, node_idx1, e1, st);
  T.entry_node = up(A, entry_node.data(), e1, st);
  T.off2 = up(A, node_off2, n_nodes + 1, st);
  T.idx2 = up(A, node_idx2, e2, st);
  for (int i = 0; i < 9; i++) T.F12[i] = F12[i];
  T.ep[0] = epipole[0]; T.ep[1] = epipole[1];
  for (int l = 0; l < kf2->nlevels; l++) T.sigma2_2[l] = level_sigma2_2[l];
  T.only_stereo = only_stereo != 0; T.coarse = coarse != 0;
  int* d_m12 = A.take<int>(n1);
  k_fill<<<(n1 + 255) / 256, 256, 0, st>>>(d_m12, n1, -1);
  k_triangulate<<<(e1 + 7) / 8, 256, 0, st>>>(T, d_m12);
  count_launch(2);
  CUORB_CUDA_TRY(cudaMemcpyAsync(match12_out, d_m12, sizeof(int) * n1, cudaMemcpyDeviceToHost, st));
  CUORB_CUDA_TRY(cudaStreamSynchronize(st));
  CUORB_CUDA_TRY(cudaGetLastError());
  int nmatches = 0;
  int hist[HISTO_LENGTH] = {0}, keep[3];
  std::vector<int> bin(n1, -1);
  for (int i = 0; i < n1; i++) {
    if (match12_out[i] < 0) continue;
    nmatches++;
    if (check_orientation) {
      bin[i] = rot_bin(kf1->keys_un[i].angle, kf2->keys_un[match12_out[i]].angle);
      if (bad_bin(bin[i])) return CUORB_E_BADARG;
      hist[bin[i]]++;
    }
  }
  if (check_orientation) {
    three_maxima(hist, keep);
    for (int i = 0; i < n1; i++)
      if (bin[i] >= 0 && bin[i] != keep[0] && bin[i] != keep[1] && bin[i] != keep[2]) { match12_out[i] = -1; nmatches--; }
  }
  return nmatches;
}

}  // extern "C"
