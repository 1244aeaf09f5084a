// K3 — quadtree keypoint distribution (replaces ORBextractor::DistributeOctTree,
// ExtractorNode::DivideNode and compareNodes, src/ORBextractor.cc:481-753).
//
// The reference is a sequential algorithm over a std::list (children push_front'ed, parent
// erased, a last phase that walks a std::sort'ed vector from the back and stops the moment the
// quota is reached), and the OUTPUT ORDER of the keypoints is the final list order.  This kernel
// produces exactly that list, but as a sequence of data-parallel ROUNDS, one CTA per
// (image, level) problem:
//
//   * A full pass of the reference's main loop (:613-675) splits EVERY node with more than one
//     key, so its result is a pure function of the list before the pass: the new list is
//     [children of the last split parent (n4..n1), ..., children of the first split parent]
//     followed by the unsplit (single-key) nodes in their old order.  Child counts come from one
//     pass over the keys (shared-memory atomics), list positions from block-wide prefix sums.
//   * A round of the final phase (:677-725) sorts the expandable nodes by (nKeys, UL.x) — with
//     a port of libstdc++'s introsort (introsort.cuh) run by one thread, because the comparator
//     has ties and std::sort is not stable — then splits from the back until the quota is
//     reached.  Child counts of ALL sorted nodes are computed in parallel first; the stopping
//     index is then a suffix-sum search, and the new list is
//     [children of the last processed node, ..., children of the first processed node]
//     followed by the old list minus the processed parents.
//   * Keys never move: every key carries the list position of its node (u16), remapped once per
//     round.  The final "max response, first one wins" pick (:735-750) is a 64-bit atomicMax per
//     node on (response, reference order of the key), where the reference order
//     (cell-row-major, then row-major inside a FAST cell) is a pure function of (x, y).
// Keys and node ids are staged in shared memory when they fit, else read from the L2-resident
// candidate list / scratch arena.  No roofline claim: latency/occupancy-bound bookkeeping.
#include <cstdlib>
#include "common.cuh"
#include "introsort.cuh"

namespace cuorb {

namespace {

constexpr int OT_THREADS = 128;       // large batches: many CTAs per SM, 4 warps each
constexpr int OT_THREADS_WIDE = 512;  // a handful of images: one CTA per (image, level) is all there is, so it gets 16 warps
constexpr int OT_KEYCAP  = 4096;  // keys staged in shared memory (6 bytes each)

struct SortEnt { int size; short ulx; short node; };
struct SortLess {
  __host__ __device__ bool operator()(const SortEnt& a, const SortEnt& b) const {
    if (a.size < b.size) return true;   // compareNodes (:542-554)
    if (a.size > b.size) return false;
    return a.ulx < b.ulx;
  }
};

struct Box { short x0, y0, x1, y1; };  // UL = (x0,y0), BR = (x1,y1)

typedef unsigned long long u64;
constexpr u64 F20 = (1ull << 20) - 1;

__host__ __device__ inline size_t octree_node_bytes(int cap) {
  // 2 Box + 2 cnt + cc[4] + remap[4] (u16) + 2 scan (u64) + 2 SortEnt + jrank (u16)
  return (size_t)cap * (2 * 8 + 2 * 4 + 16 + 8 + 2 * 8 + 2 * 8 + 2) + 128;
}

// exclusive prefix sum over a[0..n) (packed 3 x 20-bit fields), returns the total
template <int NT>
__device__ u64 block_scan(u64* a, int n, u64* s_warp) {
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int K = (n + NT - 1) / NT;
  const int b = min(tid * K, n), e = min(b + K, n);
  u64 s = 0;
  for (int i = b; i < e; i++) s += a[i];
  u64 inc = s;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const u64 t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  __syncthreads();  // s_warp may still be read from a previous call
  if (lane == 31) s_warp[wid] = inc;
  __syncthreads();
  u64 woff = 0, total = 0;
#pragma unroll
  for (int w = 0; w < NT / 32; w++) {
    const u64 v = s_warp[w];
    if (w < wid) woff += v;
    total += v;
  }
  u64 run = woff + inc - s;
  for (int i = b; i < e; i++) { const u64 v = a[i]; a[i] = run; run += v; }
  __syncthreads();
  return total;
}

__device__ __forceinline__ int quadrant(unsigned k, const Box& b) {
  const int midX = b.x0 + ((b.x1 - b.x0 + 1) >> 1);  // UL.x + ceil((UR.x-UL.x)/2)   (:483-484)
  const int midY = b.y0 + ((b.y1 - b.y0 + 1) >> 1);
  return (cand_x(k) >= midX) + 2 * (cand_y(k) >= midY);
}

__device__ __forceinline__ Box child_box(const Box& b, int q) {
  const int midX = b.x0 + ((b.x1 - b.x0 + 1) >> 1);
  const int midY = b.y0 + ((b.y1 - b.y0 + 1) >> 1);
  Box c;
  c.x0 = (short)((q & 1) ? midX : b.x0);
  c.x1 = (short)((q & 1) ? b.x1 : midX);
  c.y0 = (short)((q & 2) ? midY : b.y0);
  c.y1 = (short)((q & 2) ? b.y1 : midY);
  return c;
}

// ---- CTA-parallel std::sort -------------------------------------------------------------------
// Produces the very permutation libstdc++'s introsort (introsort.cuh) produces, ties included,
// without running it serially:
//   * the parts of __introsort_loop are disjoint ranges, so all parts of one recursion depth are
//     partitioned concurrently, one warp per part;
//   * __unguarded_partition is a Hoare scan whose k-th swap pairs the k-th element (from the
//     left) that is not less than the pivot with the k-th element (from the right) that is not
//     greater, for as long as the former lies left of the latter.  Both sequences come from
//     ballot prefix counts over the ORIGINAL contents (the pointers never re-read a swapped
//     slot before they cross), the swaps are disjoint, and the returned cut is
//     min(next left stopper, last swapped right slot);
//   * the parts left for __final_insertion_sort are at most 16 long and no element is ever
//     moved across a part boundary (everything left of a boundary is <= everything right of it
//     and the insertion only passes strictly greater elements), so it equals a stable rank
//     inside each part.
// A part that exhausts the depth limit (heapsort branch) makes the function return false and the
// caller runs the serial port instead.
typedef unsigned short u16;

__device__ __forceinline__ u64 sort_key(const SortEnt* a, int i) {
  const u64 v = reinterpret_cast<const u64*>(a)[i];   // size | ulx << 32 | node << 48
  return ((v & 0xFFFFFFFFull) << 16) | ((v >> 32) & 0xFFFFull);
}

__device__ int warp_partition(SortEnt* a, int first, int last, u16* Lpos, u16* Rpos, int lane) {
  typedef cuorb_sort::Introsort<SortEnt, SortLess> IS;
  if (lane == 0) IS::move_median_to_first(a + first, a + first + 1, a + first + (last - first) / 2, a + last - 1, SortLess());
  __syncwarp();
  const u64 pk = sort_key(a, first);
  const unsigned lt = (1u << lane) - 1u;
  int nL = 0, nR = 0;
  for (int base = first + 1; base < last; base += 32) {
    const int q = base + lane;
    const bool isL = q < last && sort_key(a, q) >= pk;
    const unsigned b = __ballot_sync(0xffffffffu, isL);
    if (isL) Lpos[first + nL + __popc(b & lt)] = (u16)q;
    nL += __popc(b);
  }
  for (int base = last - 1; base > first; base -= 32) {
    const int q = base - lane;
    const bool isR = q > first && sort_key(a, q) <= pk;
    const unsigned b = __ballot_sync(0xffffffffu, isR);
    if (isR) Rpos[first + nR + __popc(b & lt)] = (u16)q;
    nR += __popc(b);
  }
  __syncwarp();
  const int nmin = min(nL, nR);
  int K = 0;
  for (int base = 0; base < nmin; base += 32) {
    const int k = base + lane;
    const bool ok = k < nmin && Lpos[first + k] < Rpos[first + k];
    K += __popc(__ballot_sync(0xffffffffu, ok));
  }
  u64* a64 = reinterpret_cast<u64*>(a);
  for (int k = lane; k < K; k += 32) {
    const int i = Lpos[first + k], j = Rpos[first + k];
    const u64 t = a64[i]; a64[i] = a64[j]; a64[j] = t;
  }
  int cut = 0x7fffffff;
  if (K < nL) cut = Lpos[first + K];
  if (K > 0) cut = min(cut, (int)Rpos[first + K - 1]);
  __syncwarp();
  return cut;
}

// sorts a[0..n) (n <= cap); work arrays: seg (3*cap u16), pos (4*cap u16), tmp (cap entries)
template <int NT>
__device__ bool block_std_sort(SortEnt* a, int n, u16* seg, u16* pos, SortEnt* tmp, int cap) {
  __shared__ int s_nseg, s_more, s_fb;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  u16 *segF = seg, *segL = seg + cap, *segD = seg + 2 * cap;
  u16 *Lpos = pos, *Rpos = pos + cap, *lo = pos + 2 * cap, *hi = pos + 3 * cap;
  if (tid == 0) {
    segF[0] = 0; segL[0] = (u16)n; segD[0] = (u16)(2 * cuorb_sort::Introsort<SortEnt, SortLess>::lg(n));
    s_nseg = 1; s_fb = 0;
  }
  __syncthreads();
  if (n > 16) {
    for (;;) {
      const int nseg = s_nseg;
      __syncthreads();
      if (tid == 0) s_more = 0;
      __syncthreads();
      for (int si = wid; si < nseg; si += NT / 32) {
        const int f = segF[si], l = segL[si], d = segD[si];
        if (l - f <= 16) continue;
        if (d == 0) { if (lane == 0) s_fb = 1; continue; }
        const int cut = warp_partition(a, f, l, Lpos, Rpos, lane);
        if (lane == 0) {
          segL[si] = (u16)cut; segD[si] = (u16)(d - 1);
          const int j = atomicAdd(&s_nseg, 1);
          segF[j] = (u16)cut; segL[j] = (u16)l; segD[j] = (u16)(d - 1);
          if (cut - f > 16 || l - cut > 16) s_more = 1;
        }
      }
      __syncthreads();
      if (!s_more || s_fb) break;
    }
    if (s_fb) return false;
  }
  const int nseg = s_nseg;
  for (int si = tid; si < nseg; si += NT) {
    const int f = segF[si], l = segL[si];
    for (int q = f; q < l; q++) { lo[q] = (u16)f; hi[q] = (u16)l; }
  }
  __syncthreads();
  for (int q = tid; q < n; q += NT) {
    const int f = lo[q], l = hi[q];
    const u64 mk = sort_key(a, q);
    int r = f;
    for (int j = f; j < l; j++) {
      const u64 kj = sort_key(a, j);
      r += (kj < mk) || (kj == mk && j < q);
    }
    tmp[r] = a[q];
  }
  __syncthreads();
  for (int q = tid; q < n; q += NT) a[q] = tmp[q];
  __syncthreads();
  return true;
}

}  // namespace

// grid: (nlevels, n_img), block OT_THREADS
template <int NT>
__global__ void __launch_bounds__(NT) k_octree(const ExtractorGeom* __restrict__ dg,
                                                       const unsigned* __restrict__ cand,
                                                       unsigned* __restrict__ scratch,
                                                       const int* __restrict__ cand_count,
                                                       LevelKey* __restrict__ keys_out,
                                                       int* __restrict__ key_count,
                                                       int* __restrict__ flags, int pool_cap) {
  pdl_enter();
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ u64 s_warp[NT / 32];
  __shared__ int s_cnt_ok;

  const int level = blockIdx.x, img = blockIdx.y, tid = threadIdx.x;
  const LevelGeom& L = dg->lv[level];
  const int slot = img * dg->nlevels + level;
  const int M = min(cand_count[slot], L.cand_cap);
  if (M == 0) { if (tid == 0) key_count[slot] = 0; return; }
  const int cap = pool_cap;

  // ---- shared-memory carve-up ----
  unsigned char* p = smem;
  u64* scanA = (u64*)p;            p += 8 * cap;
  u64* scanB = (u64*)p;            p += 8 * cap;
  SortEnt* V  = (SortEnt*)p;       p += 8 * cap;
  SortEnt* Vs = (SortEnt*)p;       p += 8 * cap;
  Box* boxA = (Box*)p;             p += 8 * cap;
  Box* boxB = (Box*)p;             p += 8 * cap;
  int* cc   = (int*)p;             p += 16 * cap;
  int* cntA = (int*)p;             p += 4 * cap;
  int* cntB = (int*)p;             p += 4 * cap;
  unsigned short* remap = (unsigned short*)p;  p += 8 * cap;
  unsigned short* jrank = (unsigned short*)p;  p += 2 * cap;
  p = (unsigned char*)(((uintptr_t)p + 15) & ~(uintptr_t)15);
  unsigned* s_keys = (unsigned*)p;                         // OT_KEYCAP entries
  unsigned short* s_nid = (unsigned short*)(p + 4 * OT_KEYCAP);

  // keys / node ids: shared memory when they fit, else global
  const unsigned* gkeys = cand + (size_t)img * dg->cand_img_stride + L.cand_off;
  const unsigned* keys;
  unsigned short* nid;
  if (M <= OT_KEYCAP) {
    for (int i = tid; i < M; i += NT) s_keys[i] = gkeys[i];
    keys = s_keys; nid = s_nid;
  } else {
    keys = gkeys;
    nid  = reinterpret_cast<unsigned short*>(scratch + (size_t)img * 2 * dg->cand_img_stride + L.cand_off);
  }

  Box* box = boxA;  Box* nbox = boxB;
  int* cnt = cntA;  int* ncnt = cntB;
  const int N = L.quota;
  int size = 0, vcount = 0;

  // ---- roots (:566-604): key -> root int(x / hX); empty roots are erased ----
  {
    const int   nIni = L.nIni;
    const float hX   = L.hX;
    const int   H    = L.maxBorderY - FAST_BORDER;
    for (int r = tid; r < nIni; r += NT) cc[r] = 0;
    __syncthreads();
    for (int i = tid; i < M; i += NT) {
      const int r = (int)((float)cand_x(keys[i]) / hX);
      nid[i] = (unsigned short)r;
      atomicAdd(&cc[r], 1);
    }
    __syncthreads();
    if (tid == 0) {
      int n = 0;
      for (int r = 0; r < nIni; r++) {
        if (cc[r] == 0) continue;
        Box b;
        b.x0 = (short)(int)(hX * (float)r); b.x1 = (short)(int)(hX * (float)(r + 1));
        b.y0 = 0; b.y1 = (short)H;
        box[n] = b; cnt[n] = cc[r]; remap[r] = (unsigned short)n;
        n++;
      }
      s_cnt_ok = n;
    }
    __syncthreads();
    size = s_cnt_ok;
    for (int i = tid; i < M; i += NT) nid[i] = remap[nid[i]];
    __syncthreads();
  }

  bool fail = false;
  bool final_phase = false;
  for (;;) {
    const int prevSize = size;
    // ---- (final phase only) sort the expandable nodes (:686-689) ----
    if (final_phase) {
      for (int i = tid; i < vcount; i += NT) Vs[i] = V[i];
      __syncthreads();
      if (vcount > 1 && !block_std_sort<NT>(Vs, vcount, remap, reinterpret_cast<u16*>(cc), reinterpret_cast<SortEnt*>(scanA), cap)) {
        // depth limit reached somewhere (heapsort branch): run the serial port on a fresh copy
        for (int i = tid; i < vcount; i += NT) Vs[i] = V[i];
        __syncthreads();
        if (tid == 0) cuorb_sort::Introsort<SortEnt, SortLess>::sort(Vs, Vs + vcount, SortLess());
        __syncthreads();
      }
    }
    // ---- child counts of every node with more than one key ----
    for (int i = tid; i < size * 4; i += NT) cc[i] = 0;
    __syncthreads();
    for (int i = tid; i < M; i += NT) {
      const int pos = nid[i];
      if (cnt[pos] > 1) {
        const int q = quadrant(keys[i], box[pos]);
        nid[i] = (unsigned short)(pos | (q << 14));
        atomicAdd(&cc[pos * 4 + q], 1);
      }
    }
    __syncthreads();

    int newSize, newV;
    if (!final_phase) {
      // ---- one full pass of the main loop (:613-675) ----
      for (int pos = tid; pos < size; pos += NT) {
        u64 v = 1ull << 40;  // unsplit
        if (cnt[pos] > 1) {
          const int* c = cc + pos * 4;
          const int nch = (c[0] > 0) + (c[1] > 0) + (c[2] > 0) + (c[3] > 0);
          const int nex = (c[0] > 1) + (c[1] > 1) + (c[2] > 1) + (c[3] > 1);
          v = (u64)nch | ((u64)nex << 20);
        }
        scanA[pos] = v;
      }
      __syncthreads();
      const u64 tot = block_scan<NT>(scanA, size, s_warp);
      const int totCh = (int)(tot & F20), totEx = (int)((tot >> 20) & F20), nUn = (int)(tot >> 40);
      newSize = totCh + nUn; newV = totEx;
      if (newSize > cap) { fail = true; break; }
      for (int pos = tid; pos < size; pos += NT) {
        const u64 ex = scanA[pos];
        if (cnt[pos] > 1) {
          const int* c = cc + pos * 4;
          const Box b = box[pos];
          const int nch = (c[0] > 0) + (c[1] > 0) + (c[2] > 0) + (c[3] > 0);
          // later parents come first; inside a parent n4 is front-most (push_front order)
          int np = totCh - (int)(ex & F20) - nch;
          int vi = (int)((ex >> 20) & F20);
          unsigned short npos[4];
#pragma unroll
          for (int q = 3; q >= 0; q--) {
            npos[q] = 0;
            if (c[q] > 0) {
              nbox[np] = child_box(b, q); ncnt[np] = c[q];
              npos[q] = (unsigned short)np;
              np++;
            }
            remap[pos * 4 + q] = npos[q];
          }
#pragma unroll
          for (int q = 0; q < 4; q++)
            if (c[q] > 1) { V[vi] = SortEnt{c[q], (q & 1) ? (short)(b.x0 + ((b.x1 - b.x0 + 1) >> 1)) : b.x0, (short)npos[q]}; vi++; }
        } else {
          const int np = totCh + (int)(ex >> 40);
          nbox[np] = box[pos]; ncnt[np] = cnt[pos];
          remap[pos * 4] = (unsigned short)np;
        }
      }
    } else {
      // ---- one round of the final phase (:677-725) ----
      for (int j = tid; j < vcount; j += NT) {
        const int* c = cc + (int)Vs[j].node * 4;
        const int nch = (c[0] > 0) + (c[1] > 0) + (c[2] > 0) + (c[3] > 0);
        const int nex = (c[0] > 1) + (c[1] > 1) + (c[2] > 1) + (c[3] > 1);
        scanA[j] = (u64)(nch - 1) | ((u64)nch << 20) | ((u64)nex << 40);
      }
      for (int pos = tid; pos < size; pos += NT) jrank[pos] = 0xFFFF;
      if (tid == 0) s_cnt_ok = 0;
      __syncthreads();
      const u64 tot = block_scan<NT>(scanA, vcount, s_warp);
      const int totGain = (int)(tot & F20);
      // splitting goes from the back and stops as soon as size >= N (:721): j* = largest j with
      // size + sum_{j' >= j} gain >= N; the sums are non-increasing in j, so count them
      int ok = 0;
      for (int j = tid; j < vcount; j += NT) ok += (size + totGain - (int)(scanA[j] & F20)) >= N;
      if (ok) atomicAdd(&s_cnt_ok, ok);
      __syncthreads();
      const int jstar = s_cnt_ok > 0 ? s_cnt_ok - 1 : 0;
      const u64 exStar = vcount > 0 ? scanA[jstar] : 0;
      for (int j = tid + jstar; j < vcount; j += NT) jrank[Vs[j].node] = (unsigned short)j;
      __syncthreads();
      for (int pos = tid; pos < size; pos += NT) scanB[pos] = jrank[pos] == 0xFFFF ? 1 : 0;
      __syncthreads();
      const int nKeep = (int)block_scan<NT>(scanB, size, s_warp);
      const int newCh = (int)((tot >> 20) & F20) - (int)((exStar >> 20) & F20);
      const int totEx = (int)(tot >> 40);
      newSize = nKeep + newCh;
      newV = totEx - (int)(exStar >> 40);
      if (newSize > cap) { fail = true; break; }
      for (int j = tid + jstar; j < vcount; j += NT) {
        const int pos = Vs[j].node;
        const int* c = cc + pos * 4;
        const Box b = box[pos];
        const u64 ex = scanA[j];
        const int nex = (c[0] > 1) + (c[1] > 1) + (c[2] > 1) + (c[3] > 1);
        int np = (int)((ex >> 20) & F20) - (int)((exStar >> 20) & F20);  // children of j* are front-most
        int vi = totEx - (int)(ex >> 40) - nex;                          // push order: j from the back
        unsigned short npos[4];
#pragma unroll
        for (int q = 3; q >= 0; q--) {
          npos[q] = 0;
          if (c[q] > 0) {
            nbox[np] = child_box(b, q); ncnt[np] = c[q];
            npos[q] = (unsigned short)np;
            np++;
          }
          remap[pos * 4 + q] = npos[q];
        }
#pragma unroll
        for (int q = 0; q < 4; q++)
          if (c[q] > 1) { V[vi] = SortEnt{c[q], (q & 1) ? (short)(b.x0 + ((b.x1 - b.x0 + 1) >> 1)) : b.x0, (short)npos[q]}; vi++; }
      }
      for (int pos = tid; pos < size; pos += NT) {
        if (jrank[pos] != 0xFFFF) continue;
        const unsigned short np = (unsigned short)(newCh + (int)scanB[pos]);
        nbox[np] = box[pos]; ncnt[np] = cnt[pos];
        remap[pos * 4] = np; remap[pos * 4 + 1] = np; remap[pos * 4 + 2] = np; remap[pos * 4 + 3] = np;
      }
    }
    __syncthreads();
    // ---- move every key to its node's new list position ----
    for (int i = tid; i < M; i += NT) {
      const int v = nid[i];
      nid[i] = remap[(v & 0x3FFF) * 4 + (v >> 14)];
    }
    { Box* t = box; box = nbox; nbox = t; }
    { int* t = cnt; cnt = ncnt; ncnt = t; }
    size = newSize; vcount = newV;
    __syncthreads();
    // ---- loop control (:675-677, :727-728) ----
    if (size >= N || size == prevSize) break;
    if (!final_phase && size + vcount * 3 > N) final_phase = true;
  }

  if (fail) { if (tid == 0) { atomicOr(flags, FLAG_NODE_OVERFLOW); key_count[slot] = 0; } return; }
  if (size > L.key_cap) { if (tid == 0) { atomicOr(flags, FLAG_KEY_OVERFLOW); key_count[slot] = 0; } return; }

  // ---- per node keep the max-response key, first in reference order wins (:735-750) ----
  u64* best = scanA;
  for (int pos = tid; pos < size; pos += NT) best[pos] = 0;
  __syncthreads();
  {
    const int hCell = L.hCell, wCell = L.wCell, nCols = L.nCols;
    constexpr u64 ORD = (1ull << 38) - 1;
    for (int i = tid; i < M; i += NT) {
      const unsigned k = keys[i];
      const int x = cand_x(k), y = cand_y(k);
      const u64 cell = (u64)(((y - 3) / hCell) * nCols + (x - 3) / wCell);
      const u64 ord = (cell << 24) | ((u64)y << 12) | (u64)x;
      atomicMax(&best[nid[i] & 0x3FFF], ((u64)cand_s(k) << 38) | (ORD - ord));
    }
  }
  __syncthreads();
  LevelKey* out = keys_out + (size_t)img * dg->key_img_stride + L.key_off;
  for (int pos = tid; pos < size; pos += NT) {
    const u64 b = best[pos];
    const u64 ord = ((1ull << 38) - 1) - (b & ((1ull << 38) - 1));
    LevelKey kq;
    kq.x = (unsigned short)((int)(ord & 0xFFF) + FAST_BORDER);
    kq.y = (unsigned short)((int)((ord >> 12) & 0xFFF) + FAST_BORDER);
    kq.score = (unsigned short)(b >> 38);
    kq.pad = 0;
    out[pos] = kq;
  }
  if (tid == 0) key_count[slot] = size;
}

void launch_octree(const ExtractorGeom& g, const ExtractorGeom* dg, int n_img, const unsigned* cand,
                   unsigned* cand_scratch, const int* cand_count, LevelKey* keys, int* key_count,
                   int* flags, cudaStream_t st) {
  int cap = 0;
  for (int l = 0; l < g.nlevels; l++) cap = max(cap, g.lv[l].node_cap);
  const size_t smem = octree_node_bytes(cap) + (size_t)OT_KEYCAP * 6;
  // The per-candidate passes of a level are spread over the CTA's threads; with a few images the grid cannot fill
  // the GPU anyway and the level-0 CTA is the critical path of the call, so it runs 512 threads wide
  // (single-frame call on a B200: quadtree stage 46 us at 128 threads, 35 us at 256 or 512, 41 us at 1024;
  // profiles/r02_s19_cfg1_ot*.json); results do not depend on the width.
  const bool wide = g.nlevels * n_img <= 64;
  auto go = [&](auto kern, int nt) {
    if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);   // per device
    launch_k(kern, dim3(g.nlevels, n_img), dim3(nt), smem, st, dg, cand, cand_scratch, cand_count, keys, key_count, flags, cap);
  };
  if (wide) go(k_octree<OT_THREADS_WIDE>, OT_THREADS_WIDE);
  else go(k_octree<OT_THREADS>, OT_THREADS);   // at batch 256 the stage is 0.125 ms at 128, 192 or 256 threads alike
  count_launch();
}

}  // namespace cuorb
