// K11 — projection-guided Hamming matching (replaces the tracking-time overloads
// ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th, bFar, thFar), src/ORBmatcher.cc:38-150
// and SearchByProjection(Frame& Cur, const Frame& Last, th, bMono), :1684-1811, 1889-1908; pinhole
// branches) plus Frame::GetFeaturesInArea / PosInGrid (src/Frame.cc:788-877).
//
// Eight lanes per MapPoint; they visit the features of its GetFeaturesInArea cell window through a cell-sorted CSR
// of the frame (a counting sort of <= a few thousand features, made on the host while the frame is flattened -
// upload_frame - or on the device from the extractor's output - k_frame_build) and apply the exact predicate (level
// window, |dx|,|dy| < r); the reference's candidate visiting order (cell ix, then iy, then insertion = feature index)
// is part of the arg-min key, which reproduces its "first strict minimum wins" and best/second-best bookkeeping
// whatever order the lanes see the candidates in.
//
// The reference loop is order dependent: a feature bound to a MapPoint with observations is
// skipped by all later MapPoints (:96-100, :1763-1767).  That is resolved by fixed-point rounds:
// round k recomputes every MapPoint against "features taken by an EARLIER MapPoint in round k-1"
// (atomicMin of the binder index per feature).  MapPoint 0 is final after round 1, MapPoint i once
// every j < i it depends on is final, and a round that changes nothing is the sequential result.
#include <algorithm>
#include <cstdlib>
#include <cmath>
#include <mutex>

#include "frustum.cuh"
#include "project.cuh"
#include "undistort.cuh"

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

// One in-play query point, as k_proj_begin leaves it for every later round: everything the rounds need that does not
// change between rounds (window, radius, level range) in two 16-byte loads.
struct __align__(16) ProjItem {
  int      i;        // query index
  unsigned win;      // cx0 | cx1 << 6 | cy0 << 12 | cy1 << 18 | HAS_OBS << 24 (GetFeaturesInArea cell window)
  float    x, y;
  float    rad;
  float    pr;       // predicted right coordinate compared with mvuRight (mode 0: projXR, else x - mbf * invz)
  int      min_level, max_level;
};

// Control block of one batch of rounds (device ints): dirty[k] = one bit per grid cell holding a feature that a point
// joined or left in the round before round k (slot ROUND_BATCH carries over to the next batch's slot 0).
constexpr int ROUND_BATCH = 6;   // most rounds per host sync (projection_rounds picks 4 or 6)
constexpr int DIRTY_WORDS = 2 * FRAME_GRID_ROWS;
struct RoundCtl {
  unsigned dirty[ROUND_BATCH + 1][DIRTY_WORDS];
  int flag[ROUND_BATCH];     // a result changed in round k
  int count[ROUND_BATCH];    // length of round k's work list
};

// The per-point prologue of the three projection searches, once per call.  grid: ceil(max(m, n)/256).
//   mode 0  SearchByProjection(F, vpMapPoints, th, bFarPoints, thFarPoints)   src/ORBmatcher.cc:38-227
//   mode 1  SearchByProjection(CurrentFrame, LastFrame, th, bMono)            src/ORBmatcher.cc:1684-1908
//   mode 2  its right-camera half on a rig
// Points that can bind go to `items` (any order); every result starts at -1; both binder tables start empty.
__global__ void k_proj_begin(ProjFrame F, ProjQuery Q, int* __restrict__ binder0, int* __restrict__ binder1,
                             int* __restrict__ result, ProjItem* __restrict__ items, int* __restrict__ n_items) {
  pdl_enter();
  const int i = blockIdx.x * blockDim.x + threadIdx.x, lane = threadIdx.x & 31;
  if (i < F.n) { binder0[i] = NO_BINDER; binder1[i] = NO_BINDER; }
  bool active = false;
  ProjItem it{};
  if (i < Q.n) {
    result[i] = -1;
    const unsigned fl = Q.flags[i];
    active = (fl & CUORB_MP_IN_VIEW) != 0;
    if (active) {
      const float x = Q.px[i], y = Q.py[i];
      const int lvl = Q.level[i];
      const float aux = Q.aux ? Q.aux[i] : 0.f;
      float rad;
      int minLevel, maxLevel;
      if (Q.mode == 0) {
        if (Q.far_points && Q.depth[i] > Q.thFar) active = false;
        if (fl & CUORB_MP_BAD) active = false;
        float r = ((double)Q.viewcos[i] > 0.998) ? 2.5f : 4.0f;   // RadiusByViewingCos (:229-235)
        if ((double)Q.th != 1.0) r = __fmul_rn(r, Q.th);
        rad = __fmul_rn(r, F.scale[lvl]);
        minLevel = lvl - 1; maxLevel = lvl;
        it.pr = aux;
      } else {
        if (Q.mode == 1) {
          if (aux < 0.f) active = false;                           // invzc < 0 (:1717)
          if (x < F.minX || x > F.maxX || y < F.minY || y > F.maxY) active = false;
        } else if (!Q.gate_in[i]) active = false;                  // :1812-1886: no tests of its own
        rad = __fmul_rn(Q.th, F.scale[lvl]);
        if (Q.forward) { minLevel = lvl; maxLevel = -1; }
        else if (Q.backward) { minLevel = 0; maxLevel = lvl; }
        else { minLevel = lvl - 1; maxLevel = lvl + 1; }
        it.pr = __fsub_rn(x, __fmul_rn(F.mbf, aux));
      }
      if (active) {  // GetFeaturesInArea cell window (src/Frame.cc:802-828)
        const int cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, F.minX), rad), F.gwinv)));
        const int cx1 = min(FRAME_GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, F.minX), rad), F.gwinv)));
        const int cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, F.minY), rad), F.ghinv)));
        const int cy1 = min(FRAME_GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, F.minY), rad), F.ghinv)));
        if (cx0 >= FRAME_GRID_COLS || cx1 < 0 || cy0 >= FRAME_GRID_ROWS || cy1 < 0) active = false;
        // an empty range (x or y far outside the grid) visits no cell: nothing to bind, gate_out stays 0
        if (cx0 > cx1 || cy0 > cy1) active = false;
        it.i = i; it.x = x; it.y = y; it.rad = rad; it.min_level = minLevel; it.max_level = maxLevel;
        it.win = (unsigned)cx0 | ((unsigned)cx1 << 6) | ((unsigned)cy0 << 12) | ((unsigned)cy1 << 18) |
                 ((fl & CUORB_MP_HAS_OBS) ? 1u << 24 : 0u);
      }
    }
  }
  const unsigned b = __ballot_sync(0xffffffffu, active);
  if (!b) return;
  int base = 0;
  if (lane == __ffs(b) - 1) base = atomicAdd(n_items, __popc(b));
  base = __shfl_sync(0xffffffffu, base, __ffs(b) - 1);
  if (active) items[base + __popc(b & ((1u << lane) - 1u))] = it;
}

// Rounds after the first.  A point's outcome is a function of "binder_prev[f] < i" over the features of its window, so
// it can only differ from the previous round's if a feature of one of the window's cells changed its binder in
// between: `dirty` has one bit per such cell.  A point with a clean window keeps its result and only re-asserts its
// claim in this round's binder table; the others form this round's work list.  grid: ceil(m/256), thread per item.
__global__ void k_proj_select(const ProjItem* __restrict__ items, const int* __restrict__ n_items,
                              const unsigned* __restrict__ dirty, const int* __restrict__ result,
                              int* __restrict__ binder_next, int* __restrict__ work, int* __restrict__ count) {
  pdl_enter();
  const int t = blockIdx.x * blockDim.x + threadIdx.x, lane = threadIdx.x & 31;
  bool hit = false;
  if (t < *n_items) {
    const uint2 hd = *reinterpret_cast<const uint2*>(&items[t]);   // i, win
    const int cx0 = hd.y & 63, cx1 = (hd.y >> 6) & 63, cy0 = (hd.y >> 12) & 63, cy1 = (hd.y >> 18) & 63;
    const unsigned long long cols = (cx1 >= 63 ? ~0ull : ((2ull << cx1) - 1ull)) & ~((1ull << cx0) - 1ull);
    for (int cy = cy0; cy <= cy1; cy++) {
      const unsigned long long row = ((unsigned long long)dirty[2 * cy + 1] << 32) | dirty[2 * cy];
      hit |= (row & cols) != 0;
    }
    if (!hit && (hd.y >> 24)) {
      const int r0 = result[hd.x];
      if (r0 >= 0) atomicMin(&binder_next[r0], (int)hd.x);
    }
  }
  const unsigned b = __ballot_sync(0xffffffffu, hit);
  if (!b) return;
  int base = 0;
  if (lane == __ffs(b) - 1) base = atomicAdd(count, __popc(b));
  base = __shfl_sync(0xffffffffu, base, __ffs(b) - 1);
  if (hit) work[base + __popc(b & ((1u << lane) - 1u))] = t;
}

// One binding round over a work list (work == null: every item).  Persistent warps, eight lanes per item (four items
// in flight per warp).  The candidates of the item's cell window - a few features spread over 2-4 grid rows - are
// numbered across the rows (prefix sum over the lanes' row lengths) and dealt to the lanes, so a typical window is ONE
// pass of item -> cstart -> record -> binder + descriptor with all loads of a stage in flight together: the round is
// bound by the length of that chain, not by arithmetic.
// A point whose result changes marks the grid cells of the old and of the new feature in dirty_next: a feature's
// binder can only change between two rounds if a point joined or left it, so those cells cover every window that the
// next round has to look at again (k_proj_select).  `spare` is the binder table of the round after this one (nobody
// reads it now): emptied here, so the next round's claims find it clean.
__global__ void __launch_bounds__(256) k_proj_round(ProjFrame F, const uint8_t* __restrict__ qdesc, const int* __restrict__ qslot,
                                                    int* __restrict__ gate_out, int mode0, float nnratio,
                                                    const ProjItem* __restrict__ items, const int* __restrict__ work,
                                                    const int* __restrict__ count, const int* __restrict__ binder_prev,
                                                    int* __restrict__ binder_next, int* __restrict__ spare,
                                                    int* __restrict__ result, int* __restrict__ changed,
                                                    unsigned* __restrict__ dirty_next) {
  pdl_enter();
  for (int f = blockIdx.x * blockDim.x + threadIdx.x; f < F.n; f += gridDim.x * blockDim.x) spare[f] = NO_BINDER;
  const int lane = threadIdx.x & 31, sub = lane & 7;
  const unsigned mask = 0xFFu << (lane & 24);                 // this item's eight lanes
  const int n_work = *count;
  for (int w = (blockIdx.x * 8 + (threadIdx.x >> 5)) * 4 + (lane >> 3); w < n_work; w += gridDim.x * 32) {
    const uint4* ip = reinterpret_cast<const uint4*>(&items[work ? work[w] : w]);
    const uint4 h0 = ip[0], h1 = ip[1];
    const int i = (int)h0.x;
    const int cx0 = h0.y & 63, cx1 = (h0.y >> 6) & 63, cy0 = (h0.y >> 12) & 63, cy1 = (h0.y >> 18) & 63;
    const bool has_obs = (h0.y >> 24) != 0;
    const float x = __uint_as_float(h0.z), y = __uint_as_float(h0.w), rad = __uint_as_float(h1.x), pr = __uint_as_float(h1.y);
    const int minLevel = (int)h1.z, maxLevel = (int)h1.w;
    const bool checkLevels = (minLevel > 0) || (maxLevel >= 0);
    const int r_old = sub == 0 ? result[i] : 0;               // needed at the very end: in flight with everything else
    const uint4* qd = reinterpret_cast<const uint4*>(qdesc + (size_t)(qslot ? qslot[i] : i) * 32);
    const uint4 q0 = qd[0], q1 = qd[1];
    // key = dist << 32 | visiting position (cx, cy, f) << 4 | octave: smallest = the reference's best, second = its
    // second best; the iteration order here is free
    unsigned long long b1 = ~0ull, b2 = ~0ull;
    bool any_cand = false;
    for (int cyb = cy0; cyb <= cy1; cyb += 8) {
      // GetFeaturesInArea: lane r owns grid row cyb + r; the row's cells cx0..cx1 are one contiguous CSR range
      int js = 0, cnt = 0;
      if (cyb + sub <= cy1) {
        js = F.cstart[(cyb + sub) * FRAME_GRID_COLS + cx0];
        cnt = F.cstart[(cyb + sub) * FRAME_GRID_COLS + cx1 + 1] - js;
      }
      int incl = cnt;
#pragma unroll
      for (int d = 1; d < 8; d <<= 1) { const int v = __shfl_up_sync(mask, incl, d, 8); if (sub >= d) incl += v; }
      const int total = __shfl_sync(mask, incl, 7, 8);
      for (int t0 = 0; t0 < total; t0 += 8) {
        const int t = t0 + sub;
        int j = -1, cy = 0;
#pragma unroll
        for (int r = 0; r < 8; r++) {                          // which row holds candidate t
          const int e = __shfl_sync(mask, incl, r, 8), c = __shfl_sync(mask, cnt, r, 8), j0 = __shfl_sync(mask, js, r, 8);
          if (t < e && t >= e - c) { j = j0 + (t - (e - c)); cy = cyb + r; }
        }
        if (j < 0) continue;
        const float4 r = F.rec[j];
        const unsigned bits = __float_as_uint(r.w);
        const int f = bits & 0xFFFF, oct = (bits >> 16) & 31, cx = (bits >> 21) & 63;
        if (checkLevels) {
          if (oct < minLevel) continue;
          if (maxLevel >= 0 && oct > maxLevel) continue;
        }
        if (!(fabsf(__fsub_rn(r.x, x)) < rad && fabsf(__fsub_rn(r.y, y)) < rad)) continue;
        any_cand = true;
        if ((bits >> 27) & 1u) continue;                                     // blocked0
        if (F.uright && r.z > 0.f && fabsf(__fsub_rn(pr, r.z)) > rad) continue;
        const int bp = binder_prev[f];                                       // binder and descriptor: one round trip
        const uint4* fd = reinterpret_cast<const uint4*>(F.desc + (size_t)f * 32);
        const uint4 d0 = fd[0], d1 = fd[1];
        if (bp < i) continue;
        const int dist = __popc(q0.x ^ d0.x) + __popc(q0.y ^ d0.y) + __popc(q0.z ^ d0.z) + __popc(q0.w ^ d0.w) +
                         __popc(q1.x ^ d1.x) + __popc(q1.y ^ d1.y) + __popc(q1.z ^ d1.z) + __popc(q1.w ^ d1.w);
        const unsigned long long key = ((unsigned long long)dist << 32) | ((unsigned)cx << 26) | ((unsigned)cy << 20) |
                                       ((unsigned)f << 4) | (unsigned)oct;
        if (key < b1) { b2 = b1; b1 = key; }
        else if (key < b2) b2 = key;
      }
    }
    if (gate_out) {
      const bool any = (__ballot_sync(mask, any_cand) & mask) != 0;
      if (sub == 0) gate_out[i] = any ? 1 : 0;
    }
    unsigned long long B1 = b1;
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) B1 = min(B1, __shfl_xor_sync(mask, B1, o));
    unsigned long long B2 = (b1 == B1) ? b2 : b1;
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) B2 = min(B2, __shfl_xor_sync(mask, B2, o));
    if (sub == 0) {
      int res = -1;
      if (B1 != ~0ull) {
        const int bestDist = (int)(B1 >> 32), bestIdx = (int)((B1 >> 4) & 0xFFFF), bestLevel = (int)(B1 & 15);
        if (bestDist <= TH_HIGH_P) {
          bool ok = true;
          if (mode0) {
            int bestDist2 = 256, bestLevel2 = -1;
            if (B2 != ~0ull) { bestDist2 = (int)(B2 >> 32); bestLevel2 = (int)(B2 & 15); }
            if (bestLevel == bestLevel2 && (float)bestDist > __fmul_rn(nnratio, (float)bestDist2)) ok = false;
          }
          if (ok) res = bestIdx;
        }
      }
      if (r_old != res) {
        result[i] = res; *changed = 1;
        if (has_obs) {          // only points with observations bind a feature (src/ORBmatcher.cc:79-81)
          if (r_old >= 0) { const int cx = F.cellx[r_old], cy = F.celly[r_old]; atomicOr(&dirty_next[2 * cy + (cx >> 5)], 1u << (cx & 31)); }
          if (res >= 0) { const int cx = F.cellx[res], cy = F.celly[res]; atomicOr(&dirty_next[2 * cy + (cx >> 5)], 1u << (cx & 31)); }
        }
      }
      if (res >= 0 && has_obs) atomicMin(&binder_next[res], i);
    }
  }
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
// arena bytes projection_rounds takes for m query points (on top of frame_bytes(n), which covers the binder tables)
inline size_t rounds_bytes(int m) { return (size_t)std::max(m, 1) * (sizeof(ProjItem) + 4) + sizeof(RoundCtl) + 4 * 256; }
static_assert(MAX_LEVELS <= 16, "the octave rides in the low four bits of the candidate key (k_proj_round)");

struct PinnedFlags {  // per host thread: the batch's flags come back by DMA into pinned memory (a pageable target is staged)
  int* p = nullptr;
  int* get() { if (!p && cudaHostAlloc((void**)&p, 64, cudaHostAllocDefault) != cudaSuccess) p = nullptr; return p; }
};
static thread_local PinnedFlags t_flags;

// after_batch (optional): enqueued behind every batch of rounds, before the flags are read back - what the caller wants
// on the host together with the verdict (its copies ride the same synchronisation; a batch that did not converge
// simply runs it again)
template <class AfterBatch>
static int projection_rounds(Arena& A, cudaStream_t st, const ProjFrame& F, const ProjQuery& Q, int* result, AfterBatch after_batch) {
  const int n = F.n, m = Q.n;
  // binder tables: round r reads [r % 3] (the claims of round r - 1), fills [(r + 1) % 3] and empties [(r + 2) % 3]
  int* binder[3] = {A.take<int>(n), A.take<int>(n), A.take<int>(n)};
  // Binding rounds to the sequential fixed point, enqueued ROUND_BATCH at a time with no host round trip in
  // between; round k of a batch sets flag[k] when a result changed and the host reads the batch's flags once.
  // Rounds past the fixed point are idempotent (equal results -> equal binder tables -> no dirty cell -> every
  // window keeps its result) and cheap, so they simply run: a device-side "skip if the previous round was quiet"
  // test cost 0.07-0.15 ms per call in every form tried, more than the rounds it saves (profiles/r02_s9_abproj.txt).
  ProjItem* items = A.take<ProjItem>(std::max(m, 1));
  int* work = A.take<int>(std::max(m, 1));
  RoundCtl* ctl = A.take<RoundCtl>(1);
  int* n_items = A.take<int>(1);
  // A chain of kernels of a few microseconds each, launched as programmatic dependents (common.cuh): every kernel
  // starts with pdl_enter() and nothing but kernels sits between them inside a batch.
  const bool pdl = pdl_allowed();
  struct Chain { bool saved; Chain(bool on) : saved(t_pdl) { t_pdl = on; } ~Chain() { t_pdl = saved; } } chain(pdl);
  int device = 0, sms = 148;
  cudaGetDevice(&device);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  int per_sm = 4;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_proj_round, 256, 0) != cudaSuccess || per_sm < 1) per_sm = 4;
  const int round_grid = std::max(1, std::min((m + 31) / 32, sms * per_sm));   // 32 items per CTA pass, every CTA resident
  // Rounds per host sync.  How many rounds the fixed point takes grows with the number of points that chase one
  // feature: 4 (3 + the quiet one) at 2 - 10 points per feature of the frame, 6 at 50 (the 50k-point benchmark map).
  // A batch that stops short pays a second sync and a second batch; one that runs long pays ~4 us per idle round.
  // Measured (profiles/r02_s45_round_batch.txt): 2k / 4k / 10k points 0.099 / 0.107 / 0.119 ms at 4 per batch against
  // 0.110 / 0.117 / 0.130 at 6; 50k points 0.199 at 4 against 0.161 at 6.  CUORB_ROUND_BATCH overrides.
  static const int batch_env = getenv("CUORB_ROUND_BATCH") ? atoi(getenv("CUORB_ROUND_BATCH")) : 0;
  const int batch = batch_env >= 1 && batch_env <= ROUND_BATCH ? batch_env : ((long long)m <= 16LL * n ? 4 : ROUND_BATCH);
  int rounds = 0;
  for (bool converged = false; !converged;) {
    if (rounds > m + 2) { set_error("projection search did not converge"); return CUORB_E_CUDA; }
    if (rounds == 0) {
      CUORB_CUDA_TRY(cudaMemsetAsync(n_items, 0, sizeof(int), st));
      CUORB_CUDA_TRY(cudaMemsetAsync(ctl, 0, sizeof(RoundCtl), st));
    } else {   // the last round's dirty cells are the next batch's first; everything else starts from zero
      CUORB_CUDA_TRY(cudaMemcpyAsync(ctl->dirty[0], ctl->dirty[batch], sizeof(ctl->dirty[0]), cudaMemcpyDeviceToDevice, st));
      CUORB_CUDA_TRY(cudaMemsetAsync(ctl->dirty[1], 0, sizeof(RoundCtl) - sizeof(ctl->dirty[0]), st));
    }
    for (int k = 0; k < batch; k++, rounds++) {
      int* prev = binder[rounds % 3];
      int* next = binder[(rounds + 1) % 3];
      int* spare = binder[(rounds + 2) % 3];
      t_pdl = pdl && k != 0;   // the batch's first kernel follows memsets: plain stream order
      if (rounds == 0)
        launch_k(k_proj_begin, dim3((std::max(m, n) + 255) / 256), dim3(256), 0, st, F, Q, binder[0], binder[1], result, items, n_items);
      else
        launch_k(k_proj_select, dim3((m + 255) / 256), dim3(256), 0, st, (const ProjItem*)items, (const int*)n_items,
                 (const unsigned*)ctl->dirty[k], (const int*)result, next, work, ctl->count + k);
      t_pdl = pdl;
      launch_k(k_proj_round, dim3(round_grid), dim3(256), 0, st, F, Q.desc, Q.slot, Q.gate_out, Q.mode == 0 ? 1 : 0, Q.nnratio,
               (const ProjItem*)items, rounds == 0 ? (const int*)nullptr : (const int*)work,
               rounds == 0 ? (const int*)n_items : (const int*)(ctl->count + k), (const int*)prev, next, spare, result,
               ctl->flag + k, ctl->dirty[k + 1]);
      count_launch(2);
    }
    t_pdl = false;
    {
      const int rc = after_batch();
      if (rc < 0) return rc;
    }
    int stack_flag[ROUND_BATCH];
    int* h_flag = t_flags.get();
    if (!h_flag) h_flag = stack_flag;
    CUORB_CUDA_TRY(cudaMemcpyAsync(h_flag, ctl->flag, sizeof(stack_flag), cudaMemcpyDeviceToHost, st));
    CUORB_CUDA_TRY(cudaStreamSynchronize(st));
    for (int k = 0; k < batch; k++)
      if (!h_flag[k]) { converged = true; rounds -= batch - 1 - k; break; }   // rounds to the first quiet one, inclusive
  }
  return rounds;
}

static int projection_rounds(Arena& A, cudaStream_t st, const ProjFrame& F, const ProjQuery& Q, int* result) {
  return projection_rounds(A, st, F, Q, result, [] { return 0; });
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
  const size_t bytes = frame_bytes(n) + rounds_bytes(m) + (size_t)m * (32 + 6 * 4 + 1 + 4 + 8) + 64 * 1024;
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

// ---- the searched Frame built where the extractor left it ------------------------------------------------------
// Frame::UndistortKeyPoints (src/Frame.cc:886-913) + Frame::AssignFeaturesToGrid (src/Frame.cc:524-553, PosInGrid
// :867-877) on device-resident keypoints: the SoA, the per-feature cell and the cell-sorted CSR that upload_frame
// builds on the host.  One CTA (a frame has a few thousand features): count per cell in shared memory, exclusive scan,
// scatter, then each cell's range is put in ascending feature order -- the order mGrid's vectors hold, which decides
// ties between equally distant candidates.
constexpr int FB_THREADS = 1024, FB_CELLS = FRAME_GRID_ROWS * FRAME_GRID_COLS, FB_PER_THREAD = FB_CELLS / FB_THREADS;
static_assert(FB_PER_THREAD * FB_THREADS == FB_CELLS, "the scan gives each thread a whole number of cells");

__global__ void __launch_bounds__(FB_THREADS)
k_frame_build(const cuorb_keypoint* __restrict__ keys, int n, int undistort, const UndistortArgs ua, float minX, float minY,
              float gwinv, float ghinv, float* __restrict__ fx, float* __restrict__ fy, int* __restrict__ foct,
              short* __restrict__ cellx, short* __restrict__ celly, int* __restrict__ cstart,
              unsigned short* __restrict__ order, cuorb_keypoint* __restrict__ keys_un, const uint8_t* __restrict__ blocked0,
              const float* __restrict__ uright, float4* __restrict__ rec) {
  __shared__ int s_cell[FB_CELLS];   // per-cell count, then the running fill cursor
  __shared__ int s_warp[FB_THREADS / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int c = tid; c < FB_CELLS; c += FB_THREADS) s_cell[c] = 0;
  __syncthreads();
  for (int i = tid; i < n; i += FB_THREADS) {
    cuorb_keypoint kp = keys[i];
    if (undistort) undistort_point(ua, kp.x, kp.y);
    fx[i] = kp.x; fy[i] = kp.y; foct[i] = kp.octave;
    if (keys_un) keys_un[i] = kp;
    const int posX = (int)roundf(__fmul_rn(__fsub_rn(kp.x, minX), gwinv));
    const int posY = (int)roundf(__fmul_rn(__fsub_rn(kp.y, minY), ghinv));
    const bool in = !(posX < 0 || posX >= FRAME_GRID_COLS || posY < 0 || posY >= FRAME_GRID_ROWS);
    cellx[i] = in ? (short)posX : (short)-1; celly[i] = in ? (short)posY : (short)-1;
    if (in) atomicAdd(&s_cell[posY * FRAME_GRID_COLS + posX], 1);
  }
  __syncthreads();
  int cnt[FB_PER_THREAD], sum = 0;
#pragma unroll
  for (int k = 0; k < FB_PER_THREAD; k++) { cnt[k] = s_cell[tid * FB_PER_THREAD + k]; sum += cnt[k]; }
  int incl = sum;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) { const int v = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += v; }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    int w = s_warp[lane];
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const int v = __shfl_up_sync(0xffffffffu, w, d); if (lane >= d) w += v; }
    s_warp[lane] = w;
  }
  __syncthreads();
  int start = incl - sum + (warp ? s_warp[warp - 1] : 0);
#pragma unroll
  for (int k = 0; k < FB_PER_THREAD; k++) {
    cstart[tid * FB_PER_THREAD + k] = start; s_cell[tid * FB_PER_THREAD + k] = start;
    start += cnt[k];
  }
  if (tid == FB_THREADS - 1) cstart[FB_CELLS] = start;
  __syncthreads();
  for (int i = tid; i < n; i += FB_THREADS) {
    const int cx = cellx[i];
    if (cx >= 0) order[atomicAdd(&s_cell[celly[i] * FRAME_GRID_COLS + cx], 1)] = (unsigned short)i;
  }
  __syncthreads();
  for (int c = tid; c < FB_CELLS; c += FB_THREADS) {       // s_cell[c] is now the end of cell c
    const int b = cstart[c], e = s_cell[c];
    for (int a = b + 1; a < e; a++) {
      const unsigned short v = order[a];
      int j = a - 1;
      while (j >= b && order[j] > v) { order[j + 1] = order[j]; j--; }
      order[j + 1] = v;
    }
  }
  __syncthreads();
  const int total = cstart[FB_CELLS];
  for (int j = tid; j < total; j += FB_THREADS) {           // the 16-byte record per CSR position (pack_feature_bits)
    const int f = order[j];
    rec[j] = make_float4(fx[f], fy[f], uright ? uright[f] : -1.0f,
                         __uint_as_float(pack_feature_bits(f, foct[f], cellx[f], blocked0[f])));
  }
}

// ProjFrame over a cuorb_frame_view whose keys_un / descriptors / u_right are device pointers: descriptors and uRight
// are read in place, everything else is made by k_frame_build.  Returns the device array of undistorted keypoints
// when want_keys_un, else null.
static const cuorb_keypoint* build_frame_on_device(Arena& A, const cuorb_frame_view* fv, const float* K4, const float* dist5,
                                                   bool want_keys_un, cudaStream_t st, ProjFrame& F) {
  const int n = fv->n, nn = std::max(n, 1);
  int* cstart = A.take<int>(FB_CELLS + 1);
  unsigned short* order = A.take<unsigned short>(nn);
  float *x = A.take<float>(nn), *y = A.take<float>(nn);
  int* oct = A.take<int>(nn);
  short *cx = A.take<short>(nn), *cy = A.take<short>(nn);
  uint8_t* blk = A.take<uint8_t>(nn);
  cuorb_keypoint* keys_un = want_keys_un ? A.take<cuorb_keypoint>(nn) : nullptr;
  float4* rec = A.take<float4>(nn);
  if (fv->bound_with_obs) cudaMemcpyAsync(blk, fv->bound_with_obs, (size_t)n, cudaMemcpyHostToDevice, st);
  else cudaMemsetAsync(blk, 0, (size_t)nn, st);
  // src/Frame.cc:888-892: mDistCoef.at<float>(0) == 0.0 -> mvKeysUn = mvKeys
  const bool undistort = K4 && dist5 && dist5[0] != 0.0f;
  const float identK[4] = {1.f, 1.f, 0.f, 0.f}, zeroD[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
  const UndistortArgs ua = undistort ? undistort_args(K4, dist5) : undistort_args(identK, zeroD);
  k_frame_build<<<1, FB_THREADS, 0, st>>>(fv->keys_un, n, undistort ? 1 : 0, ua, fv->min_x, fv->min_y, fv->grid_w_inv,
                                          fv->grid_h_inv, x, y, oct, cx, cy, cstart, order, keys_un, blk, fv->u_right, rec);
  count_launch();
  F.n = n;
  F.cstart = cstart; F.order = order; F.rec = rec; F.x = x; F.y = y; F.uright = fv->u_right; F.octave = oct; F.cellx = cx; F.celly = cy;
  F.desc = fv->descriptors; F.blocked0 = blk;
  F.minX = fv->min_x; F.minY = fv->min_y; F.maxX = fv->max_x; F.maxY = fv->max_y;
  F.gwinv = fv->grid_w_inv; F.ghinv = fv->grid_h_inv; F.mbf = fv->mbf;
  for (int l = 0; l < MAX_LEVELS; l++) F.scale[l] = l < fv->nlevels ? fv->scale_factors[l] : 0.f;
  return keys_un;
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
  // counters ([0] in view, [1] matches), match_mp and match_feat are views of ONE allocation (out_pack), so the pair
  // list and the counters come back in one copy into h_out (pinned: out_pack's image, then qflags, then result)
  int *out_pack = nullptr;
  unsigned char* h_out = nullptr;
  // skip[] and the counters are all zero: every search clears them behind itself (off the next call's critical path);
  // false after a call that returned early
  bool scratch_clean = false;
  cudaStream_t stream = nullptr;
  cudaEvent_t producer_done = nullptr;   // orders a device-resident frame's producer stream before this map's stream
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
  al(&m->dz, 4 * c); al(&m->lv, 4 * c); al(&m->result, 4 * c); al(&m->order, 4 * c);
  al(&m->out_pack, (2 + 2 * c) * sizeof(int));
  if (ok) { m->counters = m->out_pack; m->match_mp = m->out_pack + 2; }
  if (ok && cudaHostAlloc((void**)&m->h_out, (2 + 2 * c) * sizeof(int) + c + 4 * c + 64, cudaHostAllocDefault) != cudaSuccess) ok = false;
  if (ok && cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking) != cudaSuccess) ok = false;
  if (ok && cudaEventCreateWithFlags(&m->producer_done, cudaEventDisableTiming) != cudaSuccess) ok = false;
  if (ok && cudaMemset(m->pflags, CUORB_MP_BAD, c) != cudaSuccess) ok = false;   // slots never written are bad points
  if (!ok) { set_error("cuorb_localmap_create: device allocation failed"); cuorb_localmap_destroy(m); return nullptr; }
  return m;
}

void cuorb_localmap_destroy(cuorb_localmap* m) {
  if (!m) return;
  cudaSetDevice(m->device);
  if (m->stream) { cudaStreamSynchronize(m->stream); cudaStreamDestroy(m->stream); }
  if (m->producer_done) cudaEventDestroy(m->producer_done);
  void* ptrs[] = {m->P, m->N, m->minD, m->maxD, m->desc, m->pflags, m->skip, m->qflags, m->px, m->py, m->pxr, m->vc, m->dz,
                  m->lv, m->result, m->out_pack, m->order};
  for (void* p : ptrs) if (p) cudaFree(p);
  if (m->h_out) cudaFreeHost(m->h_out);
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

// where the searched Frame's arrays live
struct FrameSource {
  bool on_device = false;               // keys_un / descriptors / u_right of the view are device pointers
  cudaStream_t producer = nullptr;      // stream that wrote them (ordered before the search); null = already complete
  const float *K4 = nullptr, *dist5 = nullptr;   // non-null: the keys are mvKeys and are undistorted on the device first
  cuorb_keypoint* keys_un_out = nullptr;          // host, optional: mvKeysUn as the search saw it
};

static int localmap_search_impl(cuorb_localmap* m, const cuorb_frustum_camera* cam, const cuorb_frame_view* frame,
                                const FrameSource& src, int n_order, const int32_t* order, int n_skip, const int32_t* skip_pos,
                                float viewing_cos_limit, float th, int far_points, float th_far_points, float nn_ratio,
                                int32_t* best_idx_out, uint8_t* in_view_out, int match_capacity, int32_t* match_pos_out,
                                int32_t* match_feature_out, int32_t* n_in_view_out) {
  if (!m || !cam || !frame || frame->n < 0 || (frame->n > 0 && (!frame->keys_un || !frame->descriptors)) || !frame->scale_factors || n_skip < 0 ||
      (n_skip && !skip_pos) || match_capacity < 0 || (match_capacity && (!match_pos_out || !match_feature_out)) ||
      n_order < 0 || (order && n_order > m->capacity) || ((src.K4 == nullptr) != (src.dist5 == nullptr))) {
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
  const size_t bytes = frame_bytes(std::max(n, 1)) + rounds_bytes(M) + (size_t)std::max(n, 1) * (8 + sizeof(cuorb_keypoint)) + (size_t)n_skip * 4 + 64 * 1024;
  if (t_arena.reserve(m->device, bytes) != 0) { set_error("cudaMalloc failed (local map search workspace)"); return CUORB_E_CUDA; }
  Arena& A = t_arena;
  if (!m->scratch_clean) {
    CUORB_CUDA_TRY(cudaMemsetAsync(m->skip, 0, (size_t)m->capacity, st));
    CUORB_CUDA_TRY(cudaMemsetAsync(m->counters, 0, 2 * sizeof(int), st));
  }
  m->scratch_clean = false;
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
  // Everything the caller gets back lands in h_out (pinned) behind ONE synchronisation: [counters | pairs] as one
  // copy of out_pack's head, then mbTrackInView, then the per-point results.  The tracking flags do not change during
  // the rounds, so their copy is enqueued right away; the rest follows each batch of rounds (projection_rounds).
  const int fetch = n > 0 ? std::min(match_capacity, M) : 0;
  int* h_pack = (int*)m->h_out;
  uint8_t* h_view = m->h_out + (2 + 2 * (size_t)m->capacity) * sizeof(int);
  int* h_best = (int*)(h_view + (((size_t)m->capacity + 63) & ~(size_t)63));
  if (in_view_out) CUORB_CUDA_TRY(cudaMemcpyAsync(h_view, m->qflags, (size_t)M, cudaMemcpyDeviceToHost, st));
  if (n > 0) {
    ProjFrame F;
    const cuorb_keypoint* d_keys_un = nullptr;
    if (src.on_device) {
      if (src.producer) {
        CUORB_CUDA_TRY(cudaEventRecord(m->producer_done, src.producer));
        CUORB_CUDA_TRY(cudaStreamWaitEvent(m->stream, m->producer_done, 0));
      }
      d_keys_un = build_frame_on_device(A, frame, src.K4, src.dist5, src.keys_un_out != nullptr, st, F);
    } else {
      upload_frame(A, frame, st, F);
    }
    if (src.keys_un_out && d_keys_un)
      CUORB_CUDA_TRY(cudaMemcpyAsync(src.keys_un_out, d_keys_un, sizeof(cuorb_keypoint) * (size_t)n, cudaMemcpyDeviceToHost, st));
    ProjQuery Q{};
    Q.n = M; Q.mode = 0; Q.th = th; Q.thFar = th_far_points; Q.nnratio = nn_ratio; Q.far_points = far_points;
    Q.desc = m->desc; Q.px = m->px; Q.py = m->py; Q.aux = m->pxr; Q.level = m->lv; Q.viewcos = m->vc; Q.depth = m->dz;
    Q.flags = m->qflags; Q.slot = order ? m->order : nullptr;
    m->match_feat = m->match_mp + fetch;          // the pair list is contiguous behind the counters: one copy
    bool first_batch = true;
    const int rounds = projection_rounds(A, st, F, Q, m->result, [&]() -> int {
      if (!first_batch) CUORB_CUDA_TRY(cudaMemsetAsync(m->counters + 1, 0, sizeof(int), st));
      first_batch = false;
      k_compact_matches<<<(M + 255) / 256, 256, 0, st>>>(M, m->result, fetch, m->match_mp, m->match_feat, m->counters + 1);
      count_launch();
      CUORB_CUDA_TRY(cudaMemcpyAsync(h_pack, m->out_pack, sizeof(int) * (2 + 2 * (size_t)fetch), cudaMemcpyDeviceToHost, st));
      if (best_idx_out) CUORB_CUDA_TRY(cudaMemcpyAsync(h_best, m->result, sizeof(int) * (size_t)M, cudaMemcpyDeviceToHost, st));
      return 0;
    });
    if (rounds < 0) return rounds;
  } else {
    CUORB_CUDA_TRY(cudaMemcpyAsync(h_pack, m->out_pack, sizeof(int) * 2, cudaMemcpyDeviceToHost, st));
    CUORB_CUDA_TRY(cudaStreamSynchronize(st));
  }
  // leave the scratch clean for the next frame (asynchronous: runs while the caller works on this frame's result)
  if ((n_skip == 0 || cudaMemsetAsync(m->skip, 0, (size_t)M, st) == cudaSuccess) &&
      cudaMemsetAsync(m->counters, 0, 2 * sizeof(int), st) == cudaSuccess)
    m->scratch_clean = true;
  if (n_in_view_out) *n_in_view_out = h_pack[0];
  if (in_view_out) for (int i = 0; i < M; i++) in_view_out[i] = h_view[i] & CUORB_MP_IN_VIEW;
  if (best_idx_out && n > 0) memcpy(best_idx_out, h_best, sizeof(int) * (size_t)M);
  const int nmatches = n > 0 ? h_pack[1] : 0;
  if (match_capacity) {
    if (nmatches > match_capacity) { set_error("cuorb_localmap_search: more matches than match_capacity"); return CUORB_E_CAPACITY; }
    // MapPoint order = the order in which the caller replays F.mvpMapPoints[feature] = pMP
    std::vector<std::pair<int32_t, int32_t>> pairs((size_t)nmatches);
    for (int k = 0; k < nmatches; k++) pairs[k] = {h_pack[2 + k], h_pack[2 + fetch + k]};
    std::sort(pairs.begin(), pairs.end());
    for (int k = 0; k < nmatches; k++) { match_pos_out[k] = pairs[k].first; match_feature_out[k] = pairs[k].second; }
  }
  return nmatches;
}

int cuorb_localmap_search(cuorb_localmap* m, const cuorb_frustum_camera* cam, const cuorb_frame_view* frame, int n_order,
                          const int32_t* order, int n_skip, const int32_t* skip_pos, float viewing_cos_limit, float th,
                          int far_points, float th_far_points, float nn_ratio, int32_t* best_idx_out, uint8_t* in_view_out,
                          int match_capacity, int32_t* match_pos_out, int32_t* match_feature_out, int32_t* n_in_view_out) {
  return localmap_search_impl(m, cam, frame, FrameSource{}, n_order, order, n_skip, skip_pos, viewing_cos_limit, th, far_points,
                              th_far_points, nn_ratio, best_idx_out, in_view_out, match_capacity, match_pos_out,
                              match_feature_out, n_in_view_out);
}

int cuorb_localmap_search_resident(cuorb_localmap* m, const cuorb_frustum_camera* cam, const cuorb_frame_view* frame,
                                   void* producer_stream, const float* K4, const float* dist5, cuorb_keypoint* keys_un_out,
                                   int n_order, const int32_t* order, int n_skip, const int32_t* skip_pos,
                                   float viewing_cos_limit, float th, int far_points, float th_far_points, float nn_ratio,
                                   int32_t* best_idx_out, uint8_t* in_view_out, int match_capacity, int32_t* match_pos_out,
                                   int32_t* match_feature_out, int32_t* n_in_view_out) {
  FrameSource src;
  src.on_device = true; src.producer = (cudaStream_t)producer_stream; src.K4 = K4; src.dist5 = dist5; src.keys_un_out = keys_un_out;
  if (frame && frame->n > 0 && frame->keys_un && frame->descriptors) {
    cudaPointerAttributes pa{}, pd{};
    const bool dev_k = cudaPointerGetAttributes(&pa, frame->keys_un) == cudaSuccess && pa.type == cudaMemoryTypeDevice;
    const bool dev_d = cudaPointerGetAttributes(&pd, frame->descriptors) == cudaSuccess && pd.type == cudaMemoryTypeDevice;
    if (!dev_k || !dev_d || (m && (pa.device != m->device || pd.device != m->device)) || ((uintptr_t)frame->descriptors & 15) ||
        ((uintptr_t)frame->keys_un & 3)) {
      cudaGetLastError();
      set_error("cuorb_localmap_search_resident: keys_un / descriptors must be device pointers on the map's device");
      return CUORB_E_BADARG;
    }
  }
  return localmap_search_impl(m, cam, frame, src, n_order, order, n_skip, skip_pos, viewing_cos_limit, th, far_points,
                              th_far_points, nn_ratio, best_idx_out, in_view_out, match_capacity, match_pos_out,
                              match_feature_out, n_in_view_out);
}

}  // extern "C"
