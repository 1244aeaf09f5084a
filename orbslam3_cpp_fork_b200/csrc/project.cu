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

#include "frustum.cuh"
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
  const int* slot;             // optional: descriptor row of query i (resident local map: storage slot of sequence position i)
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
  pdl_enter();
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
    const uint4* qd = reinterpret_cast<const uint4*>(Q.desc + (size_t)(Q.slot ? Q.slot[i] : i) * 32);
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
// binder differs between two consecutive rounds (a = last round's table, b = the one before); b is then reset for
// the coming round to fill (one launch instead of a compare and a fill)
__global__ void k_dirty_cells(ProjFrame F, const int* __restrict__ a, int* __restrict__ b, unsigned* __restrict__ dirty) {
  pdl_enter();
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= F.n) return;
  const bool changed = a[f] != b[f];
  b[f] = NO_BINDER;
  if (!changed) return;
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
thread_local HostStage t_stage;

// Binding rounds to the sequential fixed point on device-resident queries; result: device, m ints.  Returns the
// number of rounds (>= 1) or a negative error code; the stream is synchronised on return.
static int projection_rounds(Arena& A, cudaStream_t st, const ProjFrame& F, const ProjQuery& Q, int* result) {
  const int n = F.n, m = Q.n;
  int* binder[2] = {A.take<int>(n), A.take<int>(n)};
  // Binding rounds to the sequential fixed point, enqueued ROUND_BATCH at a time with no host round trip in
  // between; round k of a batch sets flag[k] when a result changed and the host reads the batch's flags once.
  // Rounds past the fixed point are idempotent (equal results -> equal binder tables -> no dirty cell -> every
  // window keeps its result) and cheap (17 us at cfg4), so they simply run: a device-side "skip if the previous
  // round was quiet" test at the top of k_proj_round cost 0.07-0.15 ms per call in every form tried (per thread,
  // per block), more than the rounds it saves.  Measured at cfg4 (50k MapPoints, 5 rounds to converge): one sync
  // per round 0.556 ms per call, batches of 6 0.510 ms (scripts/ab_projection.py, profiles/r02_s9_abproj.txt).
  constexpr int ROUND_BATCH = 6;   // 2 / 3 / 4 per sync: cfg4 0.558 / 0.561 / 0.581 ms against 0.525, local map 0.339 / 0.324 / 0.365 against 0.302
  constexpr int DIRTY_WORDS = 2 * FRAME_GRID_ROWS;
  int* flag = A.take<int>(ROUND_BATCH);
  unsigned* dirty = A.take<unsigned>(ROUND_BATCH * DIRTY_WORDS);   // one bitmap per round of a batch: cleared with one memset
  {
    PdlPause plain;
    k_fill<<<(n + 255) / 256, 256, 0, st>>>(binder[0], n, NO_BINDER);
    k_fill<<<(m + 255) / 256, 256, 0, st>>>(result, m, -2);
    count_launch(2);
  }
  // The rounds are chains of kernels of a few microseconds each (a table pass over the frame's features, then the
  // warp-per-point round), so they are launched as programmatic dependents (common.cuh): every kernel starts with
  // pdl_enter() and nothing but kernels sits between them inside a batch.
  const bool pdl = pdl_allowed();
  struct Chain { bool saved; Chain(bool on) : saved(t_pdl) { t_pdl = on; } ~Chain() { t_pdl = saved; } } chain(pdl);
  int rounds = 0;
  for (bool converged = false; !converged;) {
    if (rounds > m + 2) { set_error("projection search did not converge"); return CUORB_E_CUDA; }
    CUORB_CUDA_TRY(cudaMemsetAsync(flag, 0, ROUND_BATCH * sizeof(int), st));
    CUORB_CUDA_TRY(cudaMemsetAsync(dirty, 0, ROUND_BATCH * DIRTY_WORDS * sizeof(unsigned), st));
    for (int k = 0; k < ROUND_BATCH; k++, rounds++) {
      int* prev = binder[rounds & 1];
      int* next = binder[(rounds + 1) & 1];
      unsigned* dk = dirty + k * DIRTY_WORDS;
      t_pdl = pdl && k != 0;   // the batch's first kernel follows two memsets: plain stream order
      if (rounds > 0)   // cells whose features changed binder between the last two rounds (next still holds round-2's table)
        launch_k(k_dirty_cells, dim3((n + 255) / 256), dim3(256), 0, st, F, (const int*)prev, next, dk);
      else
        launch_k(k_fill, dim3((n + 255) / 256), dim3(256), 0, st, next, n, NO_BINDER);
      t_pdl = pdl;
      launch_k(k_proj_round, dim3((m + 7) / 8), dim3(256), 0, st, F, Q, (const int*)prev, next, result, flag + k,
               rounds > 0 ? (const unsigned*)dk : (const unsigned*)nullptr);
      count_launch(2);
    }
    int h_flag[ROUND_BATCH];
    CUORB_CUDA_TRY(cudaMemcpyAsync(h_flag, flag, sizeof(h_flag), cudaMemcpyDeviceToHost, st));
    CUORB_CUDA_TRY(cudaStreamSynchronize(st));
    for (int k = 0; k < ROUND_BATCH; k++)
      if (!h_flag[k]) { converged = true; rounds -= ROUND_BATCH - 1 - k; break; }   // rounds to the first quiet one, inclusive
  }
  return rounds;
}

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
  int* result = A.take<int>(m);
  const int rounds = projection_rounds(A, st, F, Q, result);
  if (rounds < 0) return rounds;
  CUORB_CUDA_TRY(cudaMemcpyAsync(result_out, result, sizeof(int) * m, cudaMemcpyDeviceToHost, st));
  if (gate_out_host) CUORB_CUDA_TRY(cudaMemcpyAsync(gate_out_host, Q.gate_out, sizeof(int) * m, cudaMemcpyDeviceToHost, st));
  CUORB_CUDA_TRY(cudaStreamSynchronize(st));
  return rounds;
}

// ---- resident local map (Tracking::SearchLocalPoints, src/Tracking.cc: the isInFrustum loop over mvpLocalMapPoints
// followed by SearchByProjection) ----
// grid: ceil(n/128).  Per MapPoint: skipped (mnLastFrameSeen == this frame) or bad points are not in view; the
// rest go through Frame::isInFrustum (frustum.cuh) and leave their tracking fields in the map's resident SoA.
__global__ void k_localmap_frustum(const cuorb_frustum_camera c, int n, const int* __restrict__ order, int n_slots,
                                   const float* __restrict__ P3, const float* __restrict__ N3,
                                   const float* __restrict__ minD, const float* __restrict__ maxD, const uint8_t* __restrict__ pflags,
                                   const uint8_t* __restrict__ skip, float limit, float* __restrict__ projX,
                                   float* __restrict__ projY, float* __restrict__ projXR, int* __restrict__ level,
                                   float* __restrict__ viewCos, float* __restrict__ depth, uint8_t* __restrict__ qflags,
                                   int* __restrict__ count) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  bool ok = false;
  if (i < n) {
    const int s = order ? order[i] : i;                           // storage slot of sequence position i
    const unsigned pf = (s >= 0 && s < n_slots) ? pflags[s] : CUORB_MP_BAD;
    FrustumOut o;
    o.px = -1.f; o.py = -1.f; o.pxr = -1.f; o.vc = 0.f; o.dz = 0.f; o.lv = 0;
    if (!skip[i] && !(pf & CUORB_MP_BAD))
      ok = frustum_point(c, P3[3 * s], P3[3 * s + 1], P3[3 * s + 2], N3[3 * s], N3[3 * s + 1], N3[3 * s + 2], minD[s], maxD[s], limit, o);
    projX[i] = o.px; projY[i] = o.py; projXR[i] = o.pxr; level[i] = ok ? o.lv : 0; viewCos[i] = o.vc; depth[i] = o.dz;
    qflags[i] = (uint8_t)((ok ? CUORB_MP_IN_VIEW : 0u) | (pf & (CUORB_MP_BAD | CUORB_MP_HAS_OBS)));
  }
  const unsigned b = __ballot_sync(0xffffffffu, ok);
  if ((threadIdx.x & 31) == 0 && b) atomicAdd(count, __popc(b));
}

// grid: ceil(n/256)
__global__ void k_mark_skip(int n_skip, const int* __restrict__ idx, int n, uint8_t* __restrict__ skip) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n_skip && idx[k] >= 0 && idx[k] < n) skip[idx[k]] = 1;
}

// grid: ceil(m/256): (MapPoint, feature) pairs of the matches, in no particular order
__global__ void k_compact_matches(int m, const int* __restrict__ result, int cap, int* __restrict__ out_mp,
                                  int* __restrict__ out_feat, int* __restrict__ count) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const bool hit = i < m && result[i] >= 0;
  const unsigned b = __ballot_sync(0xffffffffu, hit);
  if (!b) return;
  const int lane = threadIdx.x & 31;
  int base = 0;
  if (lane == 0) base = atomicAdd(count, __popc(b));
  base = __shfl_sync(0xffffffffu, base, 0);
  if (hit) {
    const int pos = base + __popc(b & ((1u << lane) - 1u));
    if (pos < cap) { out_mp[pos] = i; out_feat[pos] = result[i]; }
  }
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
  if (!left || !right || (left->n > 0 && (!left->keys_un || !left->descriptors)) || (right->n > 0 && (!right->keys_un || !right->descriptors)) ||
      left->n < 0 || right->n < 0 || !left->scale_factors || !right->scale_factors || n_last < 0 || !keep_bins_out ||
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

// ---- resident local map ----
struct cuorb_localmap {
  int device = 0, capacity = 0, count = 0;
  float *P = nullptr, *N = nullptr, *minD = nullptr, *maxD = nullptr;
  uint8_t *desc = nullptr, *pflags = nullptr, *skip = nullptr, *qflags = nullptr;
  float *px = nullptr, *py = nullptr, *pxr = nullptr, *vc = nullptr, *dz = nullptr;
  int *lv = nullptr, *result = nullptr, *match_mp = nullptr, *match_feat = nullptr, *counters = nullptr, *order = nullptr;
  int *h_counters = nullptr;    // pinned: [0] in view, [1] matches
  cudaStream_t stream = nullptr;
};

cuorb_localmap* cuorb_localmap_create(int device, int capacity) {
  if (capacity <= 0) { set_error("cuorb_localmap_create: capacity must be positive"); return nullptr; }
  if (cudaSetDevice(device) != cudaSuccess) { set_error("cuorb_localmap_create: no such CUDA device"); return nullptr; }
  cuorb_localmap* m = new cuorb_localmap();
  m->device = device; m->capacity = capacity;
  const size_t c = (size_t)capacity;
  bool ok = true;
  auto al = [&](auto** p, size_t bytes) { if (ok && cudaMalloc((void**)p, bytes) != cudaSuccess) ok = false; };
  al(&m->P, 12 * c); al(&m->N, 12 * c); al(&m->minD, 4 * c); al(&m->maxD, 4 * c); al(&m->desc, 32 * c); al(&m->pflags, c);
  al(&m->skip, c); al(&m->qflags, c); al(&m->px, 4 * c); al(&m->py, 4 * c); al(&m->pxr, 4 * c); al(&m->vc, 4 * c);
  al(&m->dz, 4 * c); al(&m->lv, 4 * c); al(&m->result, 4 * c); al(&m->match_mp, 4 * c); al(&m->match_feat, 4 * c); al(&m->order, 4 * c);
  al(&m->counters, 2 * sizeof(int));
  if (ok && cudaHostAlloc((void**)&m->h_counters, 2 * sizeof(int), cudaHostAllocDefault) != cudaSuccess) ok = false;
  if (ok && cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking) != cudaSuccess) ok = false;
  if (ok && cudaMemset(m->pflags, CUORB_MP_BAD, c) != cudaSuccess) ok = false;   // slots never written are bad points
  if (!ok) { set_error("cuorb_localmap_create: device allocation failed"); cuorb_localmap_destroy(m); return nullptr; }
  return m;
}

void cuorb_localmap_destroy(cuorb_localmap* m) {
  if (!m) return;
  cudaSetDevice(m->device);
  if (m->stream) { cudaStreamSynchronize(m->stream); cudaStreamDestroy(m->stream); }
  void* ptrs[] = {m->P, m->N, m->minD, m->maxD, m->desc, m->pflags, m->skip, m->qflags, m->px, m->py, m->pxr, m->vc, m->dz,
                  m->lv, m->result, m->match_mp, m->match_feat, m->counters, m->order};
  for (void* p : ptrs) if (p) cudaFree(p);
  if (m->h_counters) cudaFreeHost(m->h_counters);
  delete m;
}

int cuorb_localmap_update(cuorb_localmap* m, int first, int n, int new_count, const float* world_pos, const float* normal,
                          const float* min_distance, const float* max_distance, const uint8_t* descriptors,
                          const uint8_t* flags) {
  if (!m || first < 0 || n < 0 || first + n > m->capacity || new_count < 0 || new_count > m->capacity ||
      (n > 0 && (!world_pos || !normal || !min_distance || !max_distance || !descriptors || !flags))) {
    set_error("cuorb_localmap_update: bad argument or range beyond the map's capacity");
    return CUORB_E_BADARG;
  }
  CUORB_CUDA_TRY(cudaSetDevice(m->device));
  cudaStream_t st = m->stream;
  if (n > 0) {
    const size_t f = (size_t)first, c = (size_t)n;
    CUORB_CUDA_TRY(cudaMemcpyAsync(m->P + 3 * f, world_pos, 12 * c, cudaMemcpyHostToDevice, st));
    CUORB_CUDA_TRY(cudaMemcpyAsync(m->N + 3 * f, normal, 12 * c, cudaMemcpyHostToDevice, st));
    CUORB_CUDA_TRY(cudaMemcpyAsync(m->minD + f, min_distance, 4 * c, cudaMemcpyHostToDevice, st));
    CUORB_CUDA_TRY(cudaMemcpyAsync(m->maxD + f, max_distance, 4 * c, cudaMemcpyHostToDevice, st));
    CUORB_CUDA_TRY(cudaMemcpyAsync(m->desc + 32 * f, descriptors, 32 * c, cudaMemcpyHostToDevice, st));
    CUORB_CUDA_TRY(cudaMemcpyAsync(m->pflags + f, flags, c, cudaMemcpyHostToDevice, st));
    CUORB_CUDA_TRY(cudaStreamSynchronize(st));   // the caller's buffers may be pageable and reused
  }
  m->count = new_count;
  return CUORB_OK;
}

int cuorb_localmap_count(const cuorb_localmap* m) { return m ? m->count : CUORB_E_BADARG; }

int cuorb_localmap_search(cuorb_localmap* m, const cuorb_frustum_camera* cam, const cuorb_frame_view* frame, int n_order,
                          const int32_t* order, int n_skip, const int32_t* skip_pos, float viewing_cos_limit, float th,
                          int far_points, float th_far_points, float nn_ratio, int32_t* best_idx_out, uint8_t* in_view_out,
                          int match_capacity, int32_t* match_pos_out, int32_t* match_feature_out, int32_t* n_in_view_out) {
  if (!m || !cam || !frame || frame->n < 0 || (frame->n > 0 && (!frame->keys_un || !frame->descriptors)) || !frame->scale_factors || n_skip < 0 ||
      (n_skip && !skip_pos) || match_capacity < 0 || (match_capacity && (!match_pos_out || !match_feature_out)) ||
      n_order < 0 || (order && n_order > m->capacity)) {
    set_error("cuorb_localmap_search: bad argument");
    return CUORB_E_BADARG;
  }
  const int n = frame->n, M = order ? n_order : m->count;      // sequence length: the order the points are processed in
  if (n > 65535 || frame->nlevels > MAX_LEVELS) { set_error("frame too large for projection search"); return CUORB_E_UNSUPPORTED; }
  if (n_in_view_out) *n_in_view_out = 0;
  if (best_idx_out) for (int i = 0; i < M; i++) best_idx_out[i] = -1;
  if (in_view_out) for (int i = 0; i < M; i++) in_view_out[i] = 0;
  if (M == 0) return 0;
  CUORB_CUDA_TRY(cudaSetDevice(m->device));
  cudaStream_t st = m->stream;
  const size_t bytes = frame_bytes(std::max(n, 1)) + (size_t)std::max(n, 1) * 8 + (size_t)n_skip * 4 + 64 * 1024;
  if (t_arena.reserve(m->device, bytes) != 0) { set_error("cudaMalloc failed (local map search workspace)"); return CUORB_E_CUDA; }
  Arena& A = t_arena;
  CUORB_CUDA_TRY(cudaMemsetAsync(m->skip, 0, (size_t)M, st));
  CUORB_CUDA_TRY(cudaMemsetAsync(m->counters, 0, 2 * sizeof(int), st));
  if (order) CUORB_CUDA_TRY(cudaMemcpyAsync(m->order, order, sizeof(int) * (size_t)M, cudaMemcpyHostToDevice, st));
  if (n_skip) {
    const int* d_idx = up(A, skip_pos, n_skip, st);
    k_mark_skip<<<(n_skip + 255) / 256, 256, 0, st>>>(n_skip, d_idx, M, m->skip);
    count_launch();
  }
  k_localmap_frustum<<<(M + 127) / 128, 128, 0, st>>>(*cam, M, order ? m->order : nullptr, m->count, m->P, m->N, m->minD, m->maxD,
                                                      m->pflags, m->skip, viewing_cos_limit, m->px, m->py, m->pxr, m->lv,
                                                      m->vc, m->dz, m->qflags, m->counters);
  count_launch();
  int nmatches = 0;
  if (n > 0) {
    ProjFrame F;
    upload_frame(A, frame, st, F);
    ProjQuery Q{};
    Q.n = M; Q.mode = 0; Q.th = th; Q.thFar = th_far_points; Q.nnratio = nn_ratio; Q.far_points = far_points;
    Q.desc = m->desc; Q.px = m->px; Q.py = m->py; Q.aux = m->pxr; Q.level = m->lv; Q.viewcos = m->vc; Q.depth = m->dz;
    Q.flags = m->qflags; Q.slot = order ? m->order : nullptr;
    const int rounds = projection_rounds(A, st, F, Q, m->result);
    if (rounds < 0) return rounds;
    k_compact_matches<<<(M + 255) / 256, 256, 0, st>>>(M, m->result, match_capacity, m->match_mp, m->match_feat, m->counters + 1);
    count_launch();
  }
  if (in_view_out) CUORB_CUDA_TRY(cudaMemcpyAsync(in_view_out, m->qflags, (size_t)M, cudaMemcpyDeviceToHost, st));
  CUORB_CUDA_TRY(cudaMemcpyAsync(m->h_counters, m->counters, 2 * sizeof(int), cudaMemcpyDeviceToHost, st));
  if (best_idx_out && n > 0) CUORB_CUDA_TRY(cudaMemcpyAsync(best_idx_out, m->result, sizeof(int) * (size_t)M, cudaMemcpyDeviceToHost, st));
  // the pairs come back with the counters in one round trip: the first match_capacity slots, whatever they hold
  const int fetch = n > 0 ? std::min(match_capacity, M) : 0;
  if (fetch > 0) {
    CUORB_CUDA_TRY(cudaMemcpyAsync(match_pos_out, m->match_mp, sizeof(int) * (size_t)fetch, cudaMemcpyDeviceToHost, st));
    CUORB_CUDA_TRY(cudaMemcpyAsync(match_feature_out, m->match_feat, sizeof(int) * (size_t)fetch, cudaMemcpyDeviceToHost, st));
  }
  CUORB_CUDA_TRY(cudaStreamSynchronize(st));
  if (n_in_view_out) *n_in_view_out = m->h_counters[0];
  if (in_view_out) for (int i = 0; i < M; i++) in_view_out[i] &= CUORB_MP_IN_VIEW;
  nmatches = m->h_counters[1];
  if (match_capacity) {
    if (nmatches > match_capacity) { set_error("cuorb_localmap_search: more matches than match_capacity"); return CUORB_E_CAPACITY; }
    // MapPoint order = the order in which the caller replays F.mvpMapPoints[feature] = pMP
    std::vector<std::pair<int32_t, int32_t>> pairs((size_t)nmatches);
    for (int k = 0; k < nmatches; k++) pairs[k] = {match_pos_out[k], match_feature_out[k]};
    std::sort(pairs.begin(), pairs.end());
    for (int k = 0; k < nmatches; k++) { match_pos_out[k] = pairs[k].first; match_feature_out[k] = pairs[k].second; }
  }
  return nmatches;
}

}  // extern "C"
