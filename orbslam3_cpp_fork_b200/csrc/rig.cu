// Two-camera rig (Frame::Nleft != -1) branch of
//   int ORBmatcher::SearchByProjection(Frame&, const std::vector<MapPoint*>&, th, bFarPoints, thFarPoints)
//   src/ORBmatcher.cc:38-227 (left part :63-150, right part :152-224).
// (The rig branch of SearchByProjection(Cur, Last) makes no cross-writes and is two runs of k_proj_round, project.cu.)
//
// A rig Frame holds the left camera's features at [0, Nleft) and the right camera's at [Nleft, N); each MapPoint
// searches the left grid, then the right grid, and a match on one side also writes the MapPoint into the stereo
// partner's slot on the other side (mvLeftToRightMatch / mvRightToLeftMatch), unconditionally.  So a feature's
// state when MapPoint i looks at it is "does the LAST writer before i have observations" — with unconditional
// cross-writes a MapPoint without observations can un-block a feature again, and a min-binder table (project.cu) is
// no longer enough.  The writes of a round are therefore kept as an event log: one 64-bit key
// (feature << 33 | query sequence number << 1 | has-observations) per write, radix-sorted, with a per-feature
// start index; a candidate's state is the flag of the last event below the asking query's sequence number
// (sequence number = 2 * MapPoint + side, so a MapPoint's right part sees its own left part's writes).
// Rounds iterate to the sequential fixed point exactly as in project.cu: round k evaluates every MapPoint
// against the log of round k-1's results; MapPoint 0 is final after round 1, MapPoint i once everything before
// it is, and a round that changes no result is the reference's sequential outcome.
#include <cub/device/device_radix_sort.cuh>

#include <algorithm>
#include <vector>

#include "project.cuh"

namespace cuorb {

struct RigQuery {
  int n;
  const uint8_t* desc;
  const float *px, *py, *pxr, *pyr;
  const int *level, *levelR;
  const float *viewcos, *viewcosR, *depth;
  const uint8_t* flags;
  float th, thFar, nnratio;
  int far_points;
};

struct RigLog {
  const unsigned long long* ev;  // sorted; unused slots hold ~0
  const int* fstart;             // per rig feature: index of its first event
  int nev;                       // 0 in the first round
};

__device__ __forceinline__ bool rig_blocked(const RigLog& G, bool blocked0, int gf, unsigned q) {
  bool st = blocked0;
  for (int j = G.nev ? G.fstart[gf] : 0; j < G.nev; j++) {
    const unsigned long long e = G.ev[j];
    if ((int)(e >> 33) != gf || (unsigned)((e >> 1) & 0xffffffffull) >= q) break;
    st = (e & 1ull) != 0;
  }
  return st;
}

// One camera's part for one MapPoint, by one warp: best / second-best over the GetFeaturesInArea window with the
// reference's visiting order in the key (as k_proj_round).  Returns the matched feature or -1; *ratio_fail reports
// the `bestLevel == bestLevel2 && bestDist > mfNNratio * bestDist2` rejection (:130-132 `continue`s past the right part).
__device__ int rig_side(const ProjFrame& F, const RigLog& G, int gbase, unsigned q, float x, float y, float rad,
                        int lvl, const uint4& q0, const uint4& q1, float nnratio, int lane, bool* ratio_fail) {
  *ratio_fail = false;
  const int minLevel = lvl - 1, maxLevel = lvl;
  const int cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, F.minX), rad), F.gwinv)));
  const int cx1 = min(FRAME_GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, F.minX), rad), F.gwinv)));
  const int cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, F.minY), rad), F.ghinv)));
  const int cy1 = min(FRAME_GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, F.minY), rad), F.ghinv)));
  if (cx0 >= FRAME_GRID_COLS || cx1 < 0 || cy0 >= FRAME_GRID_ROWS || cy1 < 0) return -1;
  const bool checkLevels = (minLevel > 0) || (maxLevel >= 0);
  unsigned long long b1 = ~0ull, b2 = ~0ull;
  for (int cy = cy0; cy <= cy1; cy++) {
    const int js = F.cstart[cy * FRAME_GRID_COLS + cx0], je = F.cstart[cy * FRAME_GRID_COLS + cx1 + 1];
    for (int j = js + lane; j < je; j += 32) {
      const int f = F.order[j];
      const int oct = F.octave[f];
      if (checkLevels) {
        if (oct < minLevel) continue;
        if (maxLevel >= 0 && oct > maxLevel) continue;
      }
      if (!(fabsf(__fsub_rn(F.x[f], x)) < rad && fabsf(__fsub_rn(F.y[f], y)) < rad)) continue;
      if (rig_blocked(G, F.blocked0[f] != 0, gbase + f, q)) continue;
      const int dist = hamming256(q0, q1, reinterpret_cast<const uint4*>(F.desc + (size_t)f * 32));
      const unsigned long long key =
          ((unsigned long long)dist << 32) | ((unsigned)F.cellx[f] << 22) | ((unsigned)cy << 16) | (unsigned)f;
      if (key < b1) { b2 = b1; b1 = key; }
      else if (key < b2) b2 = key;
    }
  }
  unsigned long long B1 = b1;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) B1 = min(B1, __shfl_xor_sync(0xffffffffu, B1, o));
  unsigned long long B2 = (b1 == B1) ? b2 : b1;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) B2 = min(B2, __shfl_xor_sync(0xffffffffu, B2, o));
  if (B1 == ~0ull) return -1;
  const int bestDist = (int)(B1 >> 32), bestIdx = (int)(B1 & 0xFFFF);
  if (bestDist > TH_HIGH_P) return -1;
  int bestDist2 = 256, bestLevel2 = -1;
  if (B2 != ~0ull) { bestDist2 = (int)(B2 >> 32); bestLevel2 = F.octave[(int)(B2 & 0xFFFF)]; }
  if (F.octave[bestIdx] == bestLevel2 && (float)bestDist > __fmul_rn(nnratio, (float)bestDist2)) { *ratio_fail = true; return -1; }
  return bestIdx;
}

// grid: ceil(n/8), block 256 (warp per MapPoint, left part then right part)
__global__ void __launch_bounds__(256) k_rig_round(ProjFrame FL, ProjFrame FR, RigQuery Q, RigLog G,
                                                   int* __restrict__ resL, int* __restrict__ resR,
                                                   int* __restrict__ changed) {
  const int i = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (i >= Q.n) return;
  const unsigned fl = Q.flags[i];
  const bool inL = (fl & CUORB_MP_IN_VIEW) != 0, inR = (fl & CUORB_MP_IN_VIEW_R) != 0;
  int rl = -1, rr = -1;
  bool live = inL || inR;
  if (live && Q.far_points && Q.depth[i] > Q.thFar) live = false;
  if (fl & CUORB_MP_BAD) live = false;
  if (live) {
    const uint4* qd = reinterpret_cast<const uint4*>(Q.desc + (size_t)i * 32);
    const uint4 q0 = qd[0], q1 = qd[1];
    bool fail = false;
    if (inL) {
      const int lvl = Q.level[i];
      float r = ((double)Q.viewcos[i] > 0.998) ? 2.5f : 4.0f;
      if ((double)Q.th != 1.0) r = __fmul_rn(r, Q.th);
      rl = rig_side(FL, G, 0, 2u * (unsigned)i, Q.px[i], Q.py[i], __fmul_rn(r, FL.scale[lvl]), lvl, q0, q1, Q.nnratio, lane, &fail);
    }
    if (!fail && inR) {
      const int lvl = Q.levelR[i];
      if (lvl != -1) {                                             // :154
        const float r = ((double)Q.viewcosR[i] > 0.998) ? 2.5f : 4.0f;   // no th factor on this side (:156-163)
        bool f2;
        rr = rig_side(FR, G, FL.n, 2u * (unsigned)i + 1u, Q.pxr[i], Q.pyr[i], __fmul_rn(r, FR.scale[lvl]), lvl, q0, q1, Q.nnratio, lane, &f2);
      }
    }
  }
  if (lane == 0) {
    if (resL[i] != rl || resR[i] != rr) { resL[i] = rl; resR[i] = rr; *changed = 1; }
  }
}

// grid: ceil(n/256).  Four event slots per MapPoint: left direct, left's partner, right's partner, right direct.
__global__ void k_rig_events(int n, int nl, const int* __restrict__ resL, const int* __restrict__ resR,
                             const uint8_t* __restrict__ flags, const int* __restrict__ l2r, const int* __restrict__ r2l,
                             unsigned long long* __restrict__ ev) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const unsigned long long obs = (flags[i] & CUORB_MP_HAS_OBS) ? 1ull : 0ull;
  const unsigned long long qL = (unsigned long long)(2u * (unsigned)i) << 1, qR = (unsigned long long)(2u * (unsigned)i + 1u) << 1;
  unsigned long long e0 = ~0ull, e1 = ~0ull, e2 = ~0ull, e3 = ~0ull;
  const int a = resL[i], b = resR[i];
  if (a >= 0) {
    e0 = ((unsigned long long)a << 33) | qL | obs;
    const int p = l2r[a];
    if (p != -1) e1 = ((unsigned long long)(nl + p) << 33) | qL | obs;
  }
  if (b >= 0) {
    const int p = r2l[b];
    if (p != -1) e2 = ((unsigned long long)p << 33) | qR | obs;
    e3 = ((unsigned long long)(nl + b) << 33) | qR | obs;
  }
  ev[4 * (size_t)i] = e0; ev[4 * (size_t)i + 1] = e1; ev[4 * (size_t)i + 2] = e2; ev[4 * (size_t)i + 3] = e3;
}

// grid: ceil(nf/256): first event of every rig feature in the sorted log
__global__ void k_rig_index(int nf, const unsigned long long* __restrict__ ev, int nev, int* __restrict__ fstart) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= nf) return;
  const unsigned long long key = (unsigned long long)f << 33;
  int lo = 0, hi = nev;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (ev[mid] < key) lo = mid + 1; else hi = mid;
  }
  fstart[f] = lo;
}

}  // namespace cuorb

using namespace cuorb;

extern "C" int cuorb_search_by_projection_rig(int device, const cuorb_frame_view* left, const cuorb_frame_view* right,
                                              const int32_t* left_to_right, const int32_t* right_to_left, int n_mappoints,
                                              const uint8_t* mp_desc, const float* proj_x, const float* proj_y,
                                              const int32_t* level, const float* view_cos, const float* proj_xr,
                                              const float* proj_yr, const int32_t* level_r, const float* view_cos_r,
                                              const float* depth, const uint8_t* flags, float th, int far_points,
                                              float th_far_points, float nn_ratio, int32_t* best_l_out,
                                              int32_t* best_r_out) {
  const int m = n_mappoints;
  if (!left || !right || (left->n > 0 && (!left->keys_un || !left->descriptors)) || (right->n > 0 && (!right->keys_un || !right->descriptors)) ||
      left->n < 0 || right->n < 0 || !left->scale_factors || !right->scale_factors || m < 0 || (left->n && !left_to_right) || (right->n && !right_to_left) ||
      (m && (!mp_desc || !proj_x || !proj_y || !level || !view_cos || !proj_xr || !proj_yr || !level_r || !view_cos_r ||
             !depth || !flags || !best_l_out || !best_r_out))) {
    set_error("cuorb_search_by_projection_rig: null argument");
    return CUORB_E_BADARG;
  }
  const int nl = left->n, nr = right->n, nf = nl + nr;
  if (nl > 65535 || nr > 65535 || left->nlevels > MAX_LEVELS || right->nlevels > MAX_LEVELS || m > (1 << 29)) {
    set_error("cuorb_search_by_projection_rig: frame or MapPoint set too large");
    return CUORB_E_UNSUPPORTED;
  }
  for (int i = 0; i < nl; i++)
    if (left_to_right[i] < -1 || left_to_right[i] >= nr) { set_error("cuorb_search_by_projection_rig: left_to_right out of range"); return CUORB_E_BADARG; }
  for (int i = 0; i < nr; i++)
    if (right_to_left[i] < -1 || right_to_left[i] >= nl) { set_error("cuorb_search_by_projection_rig: right_to_left out of range"); return CUORB_E_BADARG; }
  for (int i = 0; i < m; i++) best_l_out[i] = best_r_out[i] = -1;
  if (m == 0 || nf == 0) return 0;
  CUORB_CUDA_TRY(cudaSetDevice(device));
  const int nev = 4 * m;
  size_t sort_bytes = 0;
  CUORB_CUDA_TRY(cub::DeviceRadixSort::SortKeys(nullptr, sort_bytes, (const unsigned long long*)nullptr,
                                                (unsigned long long*)nullptr, nev, 0, 51, cudaStreamPerThread));
  const size_t bytes = frame_bytes(nl) + frame_bytes(nr) + (size_t)m * (32 + 9 * 4 + 1 + 8 + 2 * 4 * 8) + (size_t)nf * 12 +
                       sort_bytes + 128 * 1024;
  if (t_arena.reserve(device, bytes) != 0) { set_error("cudaMalloc failed (rig projection workspace)"); return CUORB_E_CUDA; }
  Arena& A = t_arena;
  cudaStream_t st = cudaStreamPerThread;
  ProjFrame FL, FR;
  upload_frame(A, left, st, FL);
  upload_frame(A, right, st, FR);
  FL.uright = FR.uright = nullptr;
  RigQuery Q{};
  Q.n = m; Q.th = th; Q.thFar = th_far_points; Q.nnratio = nn_ratio; Q.far_points = far_points;
  Q.desc = up(A, mp_desc, (size_t)m * 32, st);
  Q.px = up(A, proj_x, m, st); Q.py = up(A, proj_y, m, st); Q.pxr = up(A, proj_xr, m, st); Q.pyr = up(A, proj_yr, m, st);
  Q.level = up(A, level, m, st); Q.levelR = up(A, level_r, m, st);
  Q.viewcos = up(A, view_cos, m, st); Q.viewcosR = up(A, view_cos_r, m, st); Q.depth = up(A, depth, m, st);
  Q.flags = up(A, flags, m, st);
  const int* d_l2r = up(A, left_to_right, std::max(nl, 1), st);
  const int* d_r2l = up(A, right_to_left, std::max(nr, 1), st);
  int* resL = A.take<int>(m);
  int* resR = A.take<int>(m);
  unsigned long long* ev_raw = A.take<unsigned long long>(nev);
  unsigned long long* ev_sorted = A.take<unsigned long long>(nev);
  int* fstart = A.take<int>(nf);
  void* sort_tmp = A.take<unsigned char>(sort_bytes);
  constexpr int ROUND_BATCH = 4;
  int* flag = A.take<int>(ROUND_BATCH);
  k_fill<<<(m + 255) / 256, 256, 0, st>>>(resL, m, -2);
  k_fill<<<(m + 255) / 256, 256, 0, st>>>(resR, m, -2);
  count_launch(2);
  RigLog G{ev_sorted, fstart, 0};
  int rounds = 0;
  for (bool converged = false; !converged;) {
    if (rounds > 2 * m + 4) { set_error("rig projection search did not converge"); return CUORB_E_CUDA; }
    CUORB_CUDA_TRY(cudaMemsetAsync(flag, 0, ROUND_BATCH * sizeof(int), st));
    for (int k = 0; k < ROUND_BATCH; k++, rounds++) {
      k_rig_round<<<(m + 7) / 8, 256, 0, st>>>(FL, FR, Q, G, resL, resR, flag + k);
      k_rig_events<<<(m + 255) / 256, 256, 0, st>>>(m, nl, resL, resR, Q.flags, d_l2r, d_r2l, ev_raw);
      CUORB_CUDA_TRY(cub::DeviceRadixSort::SortKeys(sort_tmp, sort_bytes, (const unsigned long long*)ev_raw, ev_sorted, nev, 0, 51, st));
      k_rig_index<<<(nf + 255) / 256, 256, 0, st>>>(nf, ev_sorted, nev, fstart);
      count_launch(3 + 8);
      G.nev = nev;
    }
    int h_flag[ROUND_BATCH];
    CUORB_CUDA_TRY(cudaMemcpyAsync(h_flag, flag, sizeof(h_flag), cudaMemcpyDeviceToHost, st));
    CUORB_CUDA_TRY(cudaStreamSynchronize(st));
    for (int k = 0; k < ROUND_BATCH; k++)
      if (!h_flag[k]) { converged = true; break; }
  }
  CUORB_CUDA_TRY(cudaMemcpyAsync(best_l_out, resL, sizeof(int) * m, cudaMemcpyDeviceToHost, st));
  CUORB_CUDA_TRY(cudaMemcpyAsync(best_r_out, resR, sizeof(int) * m, cudaMemcpyDeviceToHost, st));
  CUORB_CUDA_TRY(cudaStreamSynchronize(st));
  // nmatches as the reference counts it: one per written slot (:143-147, :212-221)
  int nmatches = 0;
  for (int i = 0; i < m; i++) {
    if (best_l_out[i] >= 0) nmatches += 1 + (left_to_right[best_l_out[i]] != -1);
    if (best_r_out[i] >= 0) nmatches += 1 + (right_to_left[best_r_out[i]] != -1);
  }
  return nmatches;
}
