// K11 — projection-guided Hamming matching (replaces the tracking-time overloads
// ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th, bFar, thFar), src/ORBmatcher.cc:38-150
// and SearchByProjection(Frame& Cur, const Frame& Last, th, bMono), :1684-1811, 1889-1908; pinhole
// branches) plus Frame::GetFeaturesInArea / PosInGrid (src/Frame.cc:788-877).
//
// One warp per MapPoint; it visits the features of its GetFeaturesInArea cell window through a cell-sorted CSR
// of the frame (built on the host while the frame is flattened: a counting sort of <= a few thousand features)
// and applies the exact predicate (level window, |dx|,|dy| < r); the reference's candidate visiting order (cell
// ix, then iy, then insertion = feature index) is part of the arg-min key, which reproduces its "first strict
// minimum wins" and best/second-best bookkeeping whatever order the lanes see the candidates in.
//
// The reference loop is order dependent: a feature bound to a MapPoint with observations is
// skipped by all later MapPoints (:96-100, :1763-1767).  That is resolved by fixed-point rounds:
// round k recomputes every MapPoint against "features taken by an EARLIER MapPoint in round k-1"
// (atomicMin of the binder index per feature).  MapPoint 0 is final after round 1, MapPoint i once
// every j < i it depends on is final, and a round that changes nothing is the sequential result.
#include <algorithm>
#include <cmath>
#include <mutex>

#include "project.cuh"

namespace cuorb {

struct ProjQuery {
  int n;
  const uint8_t* desc;
  const float *px, *py, *aux;  // aux: projXR (mode 0) or invz (mode 1)
  const int* level;            // predicted level (mode 0) or last octave (mode 1)
  const float *viewcos, *depth;
  const uint8_t* flags;
  int mode;                    // 0: local-map points, 1: last-frame points, 2: last-frame points into the right camera of a rig
  int* gate_out;               // optional: 1 where GetFeaturesInArea returned a non-empty window (binding independent)
  const int* gate_in;          // mode 2: the left pass's gate_out (the right part sits behind the left part's `continue`s)
  float th, thFar, nnratio;
  int far_points, forward, backward;
};

// grid: ceil(n/8), block 256 (warp per query point)
__global__ void __launch_bounds__(256) k_proj_round(ProjFrame F, ProjQuery Q,
                                                    const int* __restrict__ binder_prev,
                                                    int* __restrict__ binder_next,
                                                    int* __restrict__ result, int* __restrict__ changed,
                                                    const unsigned* __restrict__ dirty) {
  const int i = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (i >= Q.n) return;
  int res = -1;
  const unsigned fl = Q.flags[i];
  bool active = (fl & CUORB_MP_IN_VIEW) != 0;
  float x = 0, y = 0, rad = 0;
  int minLevel = -1, maxLevel = -1;
  if (active) {
    x = Q.px[i]; y = Q.py[i];
    const int lvl = Q.level[i];
    if (Q.mode == 0) {
      if (Q.far_points && Q.depth[i] > Q.thFar) active = false;
      if (fl & CUORB_MP_BAD) active = false;
      float r = ((double)Q.viewcos[i] > 0.998) ? 2.5f : 4.0f;   // RadiusByViewingCos (:229-235)
      if ((double)Q.th != 1.0) r = __fmul_rn(r, Q.th);
      rad = __fmul_rn(r, F.scale[lvl]);
      minLevel = lvl - 1; maxLevel = lvl;
    } else {
      if (Q.mode == 1) {
        if (Q.aux[i] < 0.f) active = false;                      // invzc < 0 (:1717)
        if (x < F.minX || x > F.maxX || y < F.minY || y > F.maxY) active = false;
      } else if (!Q.gate_in[i]) active = false;                  // :1812-1886: no tests of its own
      rad = __fmul_rn(Q.th, F.scale[lvl]);
      if (Q.forward) { minLevel = lvl; maxLevel = -1; }
      else if (Q.backward) { minLevel = 0; maxLevel = lvl; }
      else { minLevel = lvl - 1; maxLevel = lvl + 1; }
    }
  }
  int cx0 = 0, cx1 = -1, cy0 = 0, cy1 = -1;
  if (active) {  // GetFeaturesInArea cell window (src/Frame.cc:802-828)
    cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, F.minX), rad), F.gwinv)));
    cx1 = min(FRAME_GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, F.minX), rad), F.gwinv)));
    cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, F.minY), rad), F.ghinv)));
    cy1 = min(FRAME_GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, F.minY), rad), F.ghinv)));
    if (cx0 >= FRAME_GRID_COLS || cx1 < 0 || cy0 >= FRAME_GRID_ROWS || cy1 < 0) active = false;
  }
  // Rounds after the first: a MapPoint's outcome is a function of "binder_prev[f] < i" over the features of its
  // window, so it can only differ from the previous round's if a feature of one of the window's cells changed its
  // binder in between.  `dirty` has one bit per grid cell holding such a feature (k_dirty_cells); a clean window
  // keeps result[i] and only re-asserts its claim in this round's binder table.
  if (dirty) {
    bool hit = false;
    if (active) {
      const unsigned long long cols = (cx1 >= 63 ? ~0ull : ((2ull << cx1) - 1ull)) & ~((1ull << cx0) - 1ull);
      for (int cy = cy0 + lane; cy <= cy1; cy += 32) {
        const unsigned long long row = ((unsigned long long)dirty[2 * cy + 1] << 32) | dirty[2 * cy];
        hit |= (row & cols) != 0;
      }
    }
    if (!__any_sync(0xffffffffu, hit)) {
      if (lane == 0) {
        const int r0 = result[i];
        if (r0 >= 0 && (fl & CUORB_MP_HAS_OBS)) atomicMin(&binder_next[r0], i);
      }
      return;
    }
  }
  if (active) {
    const bool checkLevels = (minLevel > 0) || (maxLevel >= 0);
    const uint4* qd = reinterpret_cast<const uint4*>(Q.desc + (size_t)i * 32);
    const uint4 q0 = qd[0], q1 = qd[1];
    const float aux = Q.aux ? Q.aux[i] : 0.f;
    unsigned long long b1 = ~0ull, b2 = ~0ull;  // (dist << 32 | visiting position), two smallest
    bool any_cand = false;
    // GetFeaturesInArea: only the features of the window's cells are visited (cell-sorted CSR); the reference's
    // visiting order is carried by the key, so the iteration order here is free
    for (int cy = cy0; cy <= cy1; cy++) {
     const int js = F.cstart[cy * FRAME_GRID_COLS + cx0], je = F.cstart[cy * FRAME_GRID_COLS + cx1 + 1];
     for (int j = js + lane; j < je; j += 32) {
      const int f = F.order[j];
      const int cx = F.cellx[f];
      const int oct = F.octave[f];
      if (checkLevels) {
        if (oct < minLevel) continue;
        if (maxLevel >= 0 && oct > maxLevel) continue;
      }
      if (!(fabsf(__fsub_rn(F.x[f], x)) < rad && fabsf(__fsub_rn(F.y[f], y)) < rad)) continue;
      any_cand = true;
      if (F.blocked0[f] || binder_prev[f] < i) continue;
      if (F.uright) {
        const float ur = F.uright[f];
        if (ur > 0.f) {
          const float pr = Q.mode == 0 ? aux : __fsub_rn(x, __fmul_rn(F.mbf, aux));
          if (fabsf(__fsub_rn(pr, ur)) > rad) continue;
        }
      }
      const int dist = hamming256(q0, q1, reinterpret_cast<const uint4*>(F.desc + (size_t)f * 32));
      const unsigned long long key = ((unsigned long long)dist << 32) | ((unsigned)cx << 22) | ((unsigned)cy << 16) | (unsigned)f;
      if (key < b1) { b2 = b1; b1 = key; }
      else if (key < b2) b2 = key;
     }
    }
    if (Q.gate_out) {
      const bool any = __any_sync(0xffffffffu, any_cand);
      if (lane == 0) Q.gate_out[i] = any ? 1 : 0;
    }
    unsigned long long B1 = b1;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) B1 = min(B1, __shfl_xor_sync(0xffffffffu, B1, o));
    unsigned long long B2 = (b1 == B1) ? b2 : b1;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) B2 = min(B2, __shfl_xor_sync(0xffffffffu, B2, o));
    if (B1 != ~0ull) {
      const int bestDist = (int)(B1 >> 32), bestIdx = (int)(B1 & 0xFFFF);
      if (bestDist <= TH_HIGH_P) {
        bool ok = true;
        if (Q.mode == 0) {
          int bestDist2 = 256, bestLevel2 = -1;
          if (B2 != ~0ull) { bestDist2 = (int)(B2 >> 32); bestLevel2 = F.octave[(int)(B2 & 0xFFFF)]; }
          const int bestLevel = F.octave[bestIdx];
          if (bestLevel == bestLevel2 && (float)bestDist > __fmul_rn(Q.nnratio, (float)bestDist2)) ok = false;
        }
        if (ok) res = bestIdx;
      }
    }
  }
  if (lane == 0) {
    if (result[i] != res) { result[i] = res; *changed = 1; }
    if (res >= 0 && (fl & CUORB_MP_HAS_OBS)) atomicMin(&binder_next[res], i);
  }
}

// grid: ceil(n/256): one bit per grid cell (row-major, 64 columns = 2 words per row) that holds a feature whose
// binder differs between two consecutive rounds
__global__ void k_dirty_cells(ProjFrame F, const int* __restrict__ a, const int* __restrict__ b, unsigned* __restrict__ dirty) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= F.n || a[f] == b[f]) return;
  const int cx = F.cellx[f], cy = F.celly[f];
  if (cx >= 0) atomicOr(&dirty[2 * cy + (cx >> 5)], 1u << (cx & 31));
}

// grid: ceil(n/256)
__global__ void k_hamming_pairs(const uint8_t* a, const uint8_t* b, const int* ia, const int* ib, int n, int* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint4* pa = reinterpret_cast<const uint4*>(a + (size_t)(ia ? ia[i] : i) * 32);
  const uint4 a0 = pa[0], a1 = pa[1];
  out[i] = hamming256(a0, a1, reinterpret_cast<const uint4*>(b + (size_t)(ib ? ib[i] : i) * 32));
}

// ---- host side ----
thread_local Arena t_arena;

static int run_projection(int device, const cuorb_frame_view* fv, ProjQuery Q, const uint8_t* qdesc,
                          const float* px, const float* py, const float* aux, const int32_t* level,
                          const float* viewcos, const float* depth, const uint8_t* flags,
                          int32_t* result_out, int32_t* gate_out_host = nullptr, const int32_t* gate_in_host = nullptr,
                          bool ignore_uright = false) {
  const int n = fv->n, m = Q.n;
  if (n > 65535 || fv->nlevels > MAX_LEVELS) { set_error("frame too large for projection search"); return CUORB_E_UNSUPPORTED; }
  for (int i = 0; i < m; i++) result_out[i] = -1;
  if (n <= 0 || m <= 0) return 0;
  CUORB_CUDA_TRY(cudaSetDevice(device));
  const size_t bytes = frame_bytes(n) + (size_t)m * (32 + 6 * 4 + 1 + 4 + 8) + 64 * 1024;
  if (t_arena.reserve(device, bytes) != 0) { set_error("cudaMalloc failed (projection workspace)"); return CUORB_E_CUDA; }
  Arena& A = t_arena;
  cudaStream_t st = cudaStreamPerThread;
  ProjFrame F;
  upload_frame(A, fv, st, F);
  if (ignore_uright) F.uright = nullptr;                 // rig frames: the mvuRight gate is pinhole-only (:103, :1770)
  Q.gate_out = gate_out_host ? A.take<int>(m) : nullptr;
  Q.gate_in = gate_in_host ? up(A, gate_in_host, m, st) : nullptr;
  if (Q.gate_out) CUORB_CUDA_TRY(cudaMemsetAsync(Q.gate_out, 0, sizeof(int) * m, st));
  Q.desc = up(A, qdesc, (size_t)m * 32, st);
  Q.px = up(A, px, m, st); Q.py = up(A, py, m, st);
  Q.aux = aux ? up(A, aux, m, st) : nullptr;
  Q.level = up(A, level, m, st);
  Q.viewcos = viewcos ? up(A, viewcos, m, st) : nullptr;
  Q.depth = depth ? up(A, depth, m, st) : nullptr;
  Q.flags = up(A, flags, m, st);
  int* binder[2] = {A.take<int>(n), A.take<int>(n)};
  int* result = A.take<int>(m);
  // Binding rounds to the sequential fixed point, enqueued ROUND_BATCH at a time with no host round trip in
  // between; round k of a batch sets flag[k] when a result changed and the host reads the batch's flags once.
  // Rounds past the fixed point are idempotent (equal results -> equal binder tables -> no dirty cell -> every
  // window keeps its result) and cheap (17 us at cfg4), so they simply run: a device-side "skip if the previous
  // round was quiet" test at the top of k_proj_round cost 0.07-0.15 ms per call in every form tried (per thread,
  // per block), more than the rounds it saves.  Measured at cfg4 (50k MapPoints, 5 rounds to converge): one sync
  // per round 0.556 ms per call, batches of 6 0.510 ms (scripts/ab_projection.py, profiles/r02_s9_abproj.txt).
  constexpr int ROUND_BATCH = 6;
  int* flag = A.take<int>(ROUND_BATCH);
  unsigned* dirty = A.take<unsigned>(2 * FRAME_GRID_ROWS);
  k_fill<<<(n + 255) / 256, 256, 0, st>>>(binder[0], n, NO_BINDER);
  k_fill<<<(m + 255) / 256, 256, 0, st>>>(result, m, -2);
  count_launch(2);
  int rounds = 0;
  for (bool converged = false; !converged;) {
    if (rounds > m + 2) { set_error("projection search did not converge"); return CUORB_E_CUDA; }
    CUORB_CUDA_TRY(cudaMemsetAsync(flag, 0, ROUND_BATCH * sizeof(int), st));
    for (int k = 0; k < ROUND_BATCH; k++, rounds++) {
      int* prev = binder[rounds & 1];
      int* next = binder[(rounds + 1) & 1];
      if (rounds > 0) {   // cells whose features changed binder between the last two rounds (next still holds round-2's table)
        CUORB_CUDA_TRY(cudaMemsetAsync(dirty, 0, 2 * FRAME_GRID_ROWS * sizeof(unsigned), st));
        k_dirty_cells<<<(n + 255) / 256, 256, 0, st>>>(F, prev, next, dirty);
        count_launch();
      }
      k_fill<<<(n + 255) / 256, 256, 0, st>>>(next, n, NO_BINDER);
      k_proj_round<<<(m + 7) / 8, 256, 0, st>>>(F, Q, prev, next, result, flag + k, rounds > 0 ? dirty : nullptr);
      count_launch(2);
    }
    int h_flag[ROUND_BATCH];
    CUORB_CUDA_TRY(cudaMemcpyAsync(h_flag, flag, sizeof(h_flag), cudaMemcpyDeviceToHost, st));
    CUORB_CUDA_TRY(cudaStreamSynchronize(st));
    for (int k = 0; k < ROUND_BATCH; k++)
      if (!h_flag[k]) { converged = true; rounds -= ROUND_BATCH - 1 - k; break; }   // rounds to the first quiet one, inclusive
  }
  CUORB_CUDA_TRY(cudaMemcpyAsync(result_out, result, sizeof(int) * m, cudaMemcpyDeviceToHost, st));
  if (gate_out_host) CUORB_CUDA_TRY(cudaMemcpyAsync(gate_out_host, Q.gate_out, sizeof(int) * m, cudaMemcpyDeviceToHost, st));
  CUORB_CUDA_TRY(cudaStreamSynchronize(st));
  return rounds;
}

}  // namespace cuorb

using namespace cuorb;

extern "C" {

int cuorb_search_by_projection(int device, const cuorb_frame_view* frame, int n_mappoints,
                               const uint8_t* mp_desc, const float* proj_x, const float* proj_y,
                               const float* proj_xr, const int32_t* level, const float* view_cos,
                               const float* depth, const uint8_t* flags, float th, int far_points,
                               float th_far_points, float nn_ratio, int32_t* best_idx_out) {
  if (!frame || !frame->keys_un || !frame->descriptors || !frame->scale_factors || n_mappoints < 0 ||
      (n_mappoints && (!mp_desc || !proj_x || !proj_y || !proj_xr || !level || !view_cos || !depth || !flags || !best_idx_out))) {
    set_error("cuorb_search_by_projection: null argument");
    return CUORB_E_BADARG;
  }
  ProjQuery Q{};
  Q.n = n_mappoints; Q.mode = 0; Q.th = th; Q.thFar = th_far_points; Q.nnratio = nn_ratio; Q.far_points = far_points;
  int rc = run_projection(device, frame, Q, mp_desc, proj_x, proj_y, proj_xr, level, view_cos, depth, flags, best_idx_out);
  if (rc < 0) return rc;
  int nmatches = 0;
  for (int i = 0; i < n_mappoints; i++) nmatches += best_idx_out[i] >= 0;
  return nmatches;
}

int cuorb_search_by_projection_frame(int device, const cuorb_frame_view* cur, int n_last,
                                     const uint8_t* last_desc, const float* u, const float* v,
                                     const float* inv_z, const int32_t* last_octave,
                                     const float* last_angle, const uint8_t* flags, float th,
                                     int forward, int backward, int check_orientation,
                                     int32_t* best_idx_out, int32_t* rot_bin_out, int32_t* keep_bins_out) {
  if (!cur || !cur->keys_un || !cur->descriptors || !cur->scale_factors || n_last < 0 ||
      (n_last && (!last_desc || !u || !v || !inv_z || !last_octave || !last_angle || !flags || !best_idx_out || !rot_bin_out)) || !keep_bins_out) {
    set_error("cuorb_search_by_projection_frame: null argument");
    return CUORB_E_BADARG;
  }
  ProjQuery Q{};
  Q.n = n_last; Q.mode = 1; Q.th = th; Q.forward = forward; Q.backward = backward;
  int rc = run_projection(device, cur, Q, last_desc, u, v, inv_z, last_octave, nullptr, nullptr, flags, best_idx_out);
  if (rc < 0) return rc;
  // rotation histogram + ComputeThreeMaxima (src/ORBmatcher.cc:1694, 1800-1809, 1889-1905, 2038-2071)
  const int HISTO_LENGTH = 30;
  const float factor = 1.0f / HISTO_LENGTH;
  int hist[HISTO_LENGTH] = {0};
  int nmatches = 0;
  for (int i = 0; i < n_last; i++) {
    rot_bin_out[i] = -1;
    if (best_idx_out[i] < 0) continue;
    nmatches++;
    if (check_orientation) {
      float rot = last_angle[i] - cur->keys_un[best_idx_out[i]].angle;
      if (rot < 0.0) rot += 360.0f;
      int bin = (int)std::round(rot * factor);
      if (bin == HISTO_LENGTH) bin = 0;
      if (bin < 0 || bin >= HISTO_LENGTH) { set_error("keypoint angle outside [0,360)"); return CUORB_E_BADARG; }
      hist[bin]++;
      rot_bin_out[i] = bin;
    }
  }
  keep_bins_out[0] = keep_bins_out[1] = keep_bins_out[2] = -1;
  if (check_orientation) {
    int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
    for (int i = 0; i < HISTO_LENGTH; i++) {
      const int s = hist[i];
      if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
      else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
      else if (s > max3) { max3 = s; ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
    keep_bins_out[0] = ind1; keep_bins_out[1] = ind2; keep_bins_out[2] = ind3;
    for (int i = 0; i < HISTO_LENGTH; i++)
      if (i != ind1 && i != ind2 && i != ind3) nmatches -= hist[i];
  }
  return nmatches;
}

int cuorb_search_by_projection_frame_rig(int device, const cuorb_frame_view* left, const cuorb_frame_view* right,
                                         int n_last, const uint8_t* last_desc, const float* u, const float* v,
                                         const float* inv_z, const float* ur, const float* vr,
                                         const int32_t* last_octave, const float* last_angle, const uint8_t* flags,
                                         float th, int forward, int backward, int check_orientation,
                                         int32_t* best_l_out, int32_t* best_r_out, int32_t* rot_bin_l_out,
                                         int32_t* rot_bin_r_out, int32_t* keep_bins_out) {
  if (!left || !right || !left->keys_un || !right->keys_un || !left->descriptors || !right->descriptors ||
      !left->scale_factors || !right->scale_factors || n_last < 0 || !keep_bins_out ||
      (n_last && (!last_desc || !u || !v || !inv_z || !ur || !vr || !last_octave || !last_angle || !flags || !best_l_out ||
                  !best_r_out || !rot_bin_l_out || !rot_bin_r_out))) {
    set_error("cuorb_search_by_projection_frame_rig: null argument");
    return CUORB_E_BADARG;
  }
  // The two cameras' features live in disjoint index ranges and this overload makes no cross-writes, so the two
  // parts are two independent searches; the right one only runs for points whose left window was not empty
  // (src/ORBmatcher.cc:1750-1752 `continue`s past it) and that passed the left part's depth / bounds tests.
  std::vector<int32_t> gate((size_t)std::max(n_last, 1), 0);
  ProjQuery Q{};
  Q.n = n_last; Q.mode = 1; Q.th = th; Q.forward = forward; Q.backward = backward;
  int rc = run_projection(device, left, Q, last_desc, u, v, inv_z, last_octave, nullptr, nullptr, flags, best_l_out,
                          gate.data(), nullptr, true);
  if (rc < 0) return rc;
  Q.mode = 2;
  rc = run_projection(device, right, Q, last_desc, ur, vr, nullptr, last_octave, nullptr, nullptr, flags, best_r_out,
                      nullptr, gate.data(), true);
  if (rc < 0) return rc;
  const int HISTO_LENGTH = 30;
  const float factor = 1.0f / HISTO_LENGTH;
  int hist[HISTO_LENGTH] = {0};
  int nmatches = 0;
  for (int i = 0; i < n_last; i++) {
    rot_bin_l_out[i] = rot_bin_r_out[i] = -1;
    for (int side = 0; side < 2; side++) {
      const int best = side ? best_r_out[i] : best_l_out[i];
      if (best < 0) continue;
      nmatches++;
      if (!check_orientation) continue;
      float rot = last_angle[i] - (side ? right : left)->keys_un[best].angle;    // :1793-1806, :1866-1884
      if (rot < 0.0) rot += 360.0f;
      int bin = (int)std::round(rot * factor);
      if (bin == HISTO_LENGTH) bin = 0;
      if (bin < 0 || bin >= HISTO_LENGTH) { set_error("keypoint angle outside [0,360)"); return CUORB_E_BADARG; }
      hist[bin]++;
      (side ? rot_bin_r_out : rot_bin_l_out)[i] = bin;
    }
  }
  keep_bins_out[0] = keep_bins_out[1] = keep_bins_out[2] = -1;
  if (check_orientation) {
    int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
    for (int i = 0; i < HISTO_LENGTH; i++) {
      const int s = hist[i];
      if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
      else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
      else if (s > max3) { max3 = s; ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
    keep_bins_out[0] = ind1; keep_bins_out[1] = ind2; keep_bins_out[2] = ind3;
    for (int i = 0; i < HISTO_LENGTH; i++)
      if (i != ind1 && i != ind2 && i != ind3) nmatches -= hist[i];
  }
  return nmatches;
}

int cuorb_hamming_batch(int device, const uint8_t* a, int n_a, const uint8_t* b, int n_b,
                        const int32_t* ia, const int32_t* ib, int n_pairs, int32_t* out) {
  if (!a || !b || !out || n_pairs < 0 || n_a <= 0 || n_b <= 0) { set_error("cuorb_hamming_batch: bad argument"); return CUORB_E_BADARG; }
  if (!ia && !ib && (n_pairs > n_a || n_pairs > n_b)) { set_error("cuorb_hamming_batch: n_pairs exceeds the arrays"); return CUORB_E_BADARG; }
  if (n_pairs == 0) return 0;
  CUORB_CUDA_TRY(cudaSetDevice(device));
  const size_t bytes = (size_t)(n_a + n_b) * 32 + (size_t)n_pairs * 12 + 8192;
  if (t_arena.reserve(device, bytes) != 0) { set_error("cudaMalloc failed (hamming workspace)"); return CUORB_E_CUDA; }
  cudaStream_t st = cudaStreamPerThread;
  const uint8_t* da = up(t_arena, a, (size_t)n_a * 32, st);
  const uint8_t* db = up(t_arena, b, (size_t)n_b * 32, st);
  const int* dia = ia ? up(t_arena, ia, n_pairs, st) : nullptr;
  const int* dib = ib ? up(t_arena, ib, n_pairs, st) : nullptr;
  int* dout = t_arena.take<int>(n_pairs);
  k_hamming_pairs<<<(n_pairs + 255) / 256, 256, 0, st>>>(da, db, dia, dib, n_pairs, dout);
  count_launch();
  CUORB_CUDA_TRY(cudaMemcpyAsync(out, dout, sizeof(int) * n_pairs, cudaMemcpyDeviceToHost, st));
  CUORB_CUDA_TRY(cudaStreamSynchronize(st));
  return n_pairs;
}

}  // extern "C"
