// K3 — quadtree keypoint distribution (replaces ORBextractor::DistributeOctTree,
// ExtractorNode::DivideNode and compareNodes, src/ORBextractor.cc:481-753).
//
// The reference algorithm is sequential over a std::list: children are push_front'ed, the parent
// is erased, the last phase walks a std::sort'ed vector from the back and stops the moment the
// quota is reached.  The OUTPUT ORDER of the keypoints is the final list order, so the list
// semantics (and libstdc++'s introsort tie placement, see introsort.cuh) are reproduced exactly.
// Parallelism comes from the batch: one warp per (image, level) problem, thousands of problems
// per launch; inside a problem the 32 lanes cooperate on the 4-way key partition of each split
// (ballot + popc prefix sums) while the list bookkeeping lives in shared memory.
// A node owns a contiguous span of the level's key array; a split re-partitions that span into the
// other of two ping-pong scratch buffers (the FAST candidate list itself is left untouched).  Key order inside a node is irrelevant: the only place the
// reference depends on it is the final "max response, first one wins" pick, and "first" is a pure
// function of (x, y) (cell-row-major, then row-major), re-derived there.
#include "common.cuh"
#include "introsort.cuh"

namespace cuorb {

namespace {

struct SortEnt { int size; short ulx; short node; };
struct SortLess {
  __host__ __device__ bool operator()(const SortEnt& a, const SortEnt& b) const {
    if (a.size < b.size) return true;   // compareNodes (:542-554)
    if (a.size > b.size) return false;
    return a.ulx < b.ulx;
  }
};

struct Pool {  // shared-memory views
  short *x0, *y0, *x1, *y1, *next, *prev, *freel;
  int *beg, *cnt;
  unsigned char* buf;
  SortEnt *v, *vprev;
};

constexpr short NIL = -1;

__host__ __device__ inline size_t pool_bytes(int cap) {
  // 7 short arrays, 2 int arrays, 2 SortEnt arrays, 1 byte array (+ alignment slack)
  return (size_t)cap * (7 * 2 + 2 * 4 + 2 * 8 + 1) + 64;
}

struct State {  // warp-uniform registers
  int head, size, free_top, vcount;
  bool fail;  // node pool exhausted (cannot happen with node_cap = max(N, 4*nIni) + 8)
};

__device__ __forceinline__ void push_front(Pool& P, State& S, int id, int lane) {
  if (lane == 0) {
    P.prev[id] = NIL;
    P.next[id] = (short)S.head;
    if (S.head != NIL) P.prev[S.head] = (short)id;
  }
  S.head = id;
  S.size++;
}

__device__ __forceinline__ void erase(Pool& P, State& S, int id, int lane) {
  const int p = P.prev[id], n = P.next[id];
  __syncwarp();
  if (lane == 0) {
    if (p != NIL) P.next[p] = (short)n;
    if (n != NIL) P.prev[n] = (short)p;
    P.freel[S.free_top] = (short)id;
  }
  if (p == NIL) S.head = n;
  S.free_top++;
  S.size--;
  __syncwarp();
}

// DivideNode + the "add childs" block (:481-540, :633-664).  Returns number of children with >1 key.
__device__ int split_node(Pool& P, State& S, int id, unsigned* buf0, unsigned* buf1, int lane) {
  if (S.free_top < 4) { S.fail = true; return 0; }
  const int x0 = P.x0[id], y0 = P.y0[id], x1 = P.x1[id], y1 = P.y1[id];
  const int beg = P.beg[id], cnt = P.cnt[id];
  const int b = P.buf[id];
  const unsigned* src = b ? buf1 : buf0;
  unsigned*       dst = b ? buf0 : buf1;
  const int midX = x0 + ((x1 - x0 + 1) >> 1);  // UL.x + ceil((UR.x-UL.x)/2)
  const int midY = y0 + ((y1 - y0 + 1) >> 1);

  int c[4] = {0, 0, 0, 0};
  unsigned kreg = 0;
  int qreg = -1;
  if (cnt <= 32) {
    if (lane < cnt) {
      kreg = src[beg + lane];
      qreg = (cand_x(kreg) >= midX) + 2 * (cand_y(kreg) >= midY);
    }
#pragma unroll
    for (int q = 0; q < 4; q++) c[q] = __popc(__ballot_sync(0xffffffffu, qreg == q));
  } else {
    for (int i = lane; i < ((cnt + 31) & ~31); i += 32) {
      int q = -1;
      if (i < cnt) { const unsigned k = src[beg + i]; q = (cand_x(k) >= midX) + 2 * (cand_y(k) >= midY); }
#pragma unroll
      for (int qq = 0; qq < 4; qq++) c[qq] += __popc(__ballot_sync(0xffffffffu, q == qq));
    }
  }
  int o[4];
  o[0] = beg; o[1] = o[0] + c[0]; o[2] = o[1] + c[1]; o[3] = o[2] + c[2];
  const unsigned lt_mask = (1u << lane) - 1u;
  if (cnt <= 32) {
#pragma unroll
    for (int q = 0; q < 4; q++) {
      const unsigned m = __ballot_sync(0xffffffffu, qreg == q);
      if (qreg == q) dst[o[q] + __popc(m & lt_mask)] = kreg;
    }
  } else {
    int run[4] = {o[0], o[1], o[2], o[3]};
    for (int i = lane; i < ((cnt + 31) & ~31); i += 32) {
      int q = -1;
      unsigned k = 0;
      if (i < cnt) { k = src[beg + i]; q = (cand_x(k) >= midX) + 2 * (cand_y(k) >= midY); }
#pragma unroll
      for (int qq = 0; qq < 4; qq++) {
        const unsigned m = __ballot_sync(0xffffffffu, q == qq);
        if (q == qq) dst[run[qq] + __popc(m & lt_mask)] = k;
        run[qq] += __popc(m);
      }
    }
  }
  __syncwarp();

  int n_expand = 0;
#pragma unroll
  for (int q = 0; q < 4; q++) {
    if (c[q] == 0) continue;
    const int cid = P.freel[S.free_top - 1];
    S.free_top--;
    if (lane == 0) {
      P.x0[cid] = (short)((q & 1) ? midX : x0);
      P.x1[cid] = (short)((q & 1) ? x1 : midX);
      P.y0[cid] = (short)((q & 2) ? midY : y0);
      P.y1[cid] = (short)((q & 2) ? y1 : midY);
      P.beg[cid] = o[q];
      P.cnt[cid] = c[q];
      P.buf[cid] = (unsigned char)(b ^ 1);
    }
    push_front(P, S, cid, lane);
    if (c[q] > 1) {
      if (lane == 0) P.v[S.vcount] = SortEnt{c[q], (short)((q & 1) ? midX : x0), (short)cid};
      S.vcount++;
      n_expand++;
    }
    __syncwarp();
  }
  return n_expand;
}

}  // namespace

// grid: (nlevels, n_img), block: 32 threads (one warp per problem)
__global__ void __launch_bounds__(32) k_octree(const ExtractorGeom* __restrict__ dg,
                                               const unsigned* __restrict__ cand,
                                               unsigned* __restrict__ scratch,
                                               const int* __restrict__ cand_count,
                                               LevelKey* __restrict__ keys, int* __restrict__ key_count,
                                               int* __restrict__ flags, int pool_cap) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int level = blockIdx.x, img = blockIdx.y, lane = threadIdx.x;
  const LevelGeom& L = dg->lv[level];
  const int slot = img * dg->nlevels + level;
  const int M = min(cand_count[slot], L.cand_cap);
  if (M == 0) { if (lane == 0) key_count[slot] = 0; return; }

  Pool P;
  {
    unsigned char* p = smem;
    P.v = (SortEnt*)p;      p += sizeof(SortEnt) * pool_cap;
    P.vprev = (SortEnt*)p;  p += sizeof(SortEnt) * pool_cap;
    P.beg = (int*)p;        p += 4 * pool_cap;
    P.cnt = (int*)p;        p += 4 * pool_cap;
    P.x0 = (short*)p;       p += 2 * pool_cap;
    P.y0 = (short*)p;       p += 2 * pool_cap;
    P.x1 = (short*)p;       p += 2 * pool_cap;
    P.y1 = (short*)p;       p += 2 * pool_cap;
    P.next = (short*)p;     p += 2 * pool_cap;
    P.prev = (short*)p;     p += 2 * pool_cap;
    P.freel = (short*)p;    p += 2 * pool_cap;
    P.buf = p;
  }
  const int cap = L.node_cap;
  // the FAST candidate list (bufA) stays intact (parity/debug tap); nodes ping-pong between the
  // two halves of the scratch arena
  const unsigned* bufA = cand + (size_t)img * dg->cand_img_stride + L.cand_off;
  unsigned* bufB = scratch + (size_t)img * 2 * dg->cand_img_stride + L.cand_off;
  unsigned* bufC = bufB + dg->cand_img_stride;
  const int N = L.quota;

  // free list: ids cap-1 .. 0 (pop from the top gives 0, 1, 2, ...)
  for (int i = lane; i < cap; i += 32) P.freel[i] = (short)(cap - 1 - i);
  __syncwarp();
  State S{NIL, 0, cap, 0, false};

  // roots (:566-591): key -> root int(x / hX); partition A -> B root by root
  const int   nIni = L.nIni;
  const float hX   = L.hX;
  const int   H    = L.maxBorderY - FAST_BORDER;
  int tail = NIL, off = 0;
  for (int r = 0; r < nIni; r++) {
    int run = off;
    const unsigned lt_mask = (1u << lane) - 1u;
    for (int i = lane; i < ((M + 31) & ~31); i += 32) {
      bool mine = false;
      unsigned k = 0;
      if (i < M) { k = bufA[i]; mine = (int)((float)cand_x(k) / hX) == r; }
      const unsigned m = __ballot_sync(0xffffffffu, mine);
      if (mine) bufB[run + __popc(m & lt_mask)] = k;
      run += __popc(m);
    }
    const int c = run - off;
    if (c > 0) {  // empty roots are erased (:595-604); push_back keeps root order
      const int id = P.freel[S.free_top - 1];
      S.free_top--;
      if (lane == 0) {
        P.x0[id] = (short)(int)(hX * (float)r);
        P.x1[id] = (short)(int)(hX * (float)(r + 1));
        P.y0[id] = 0; P.y1[id] = (short)H;
        P.beg[id] = off; P.cnt[id] = c; P.buf[id] = 1;
        P.next[id] = NIL; P.prev[id] = (short)tail;
        if (tail != NIL) P.next[tail] = (short)id;
      }
      if (tail == NIL) S.head = id;
      tail = id;
      S.size++;
    }
    off = run;
    __syncwarp();
  }

  bool finish = false;
  while (!finish) {
    const int prevSize = S.size;
    int nToExpand = 0;
    S.vcount = 0;
    int cur = S.head;
    while (cur != NIL) {
      const int nxt = P.next[cur];
      if (P.cnt[cur] != 1) {  // bNoMore <=> exactly one key
        nToExpand += split_node(P, S, cur, bufC, bufB, lane);
        if (S.fail) break;
        erase(P, S, cur, lane);
      }
      cur = nxt;
    }
    if (S.fail) break;
    if (S.size >= N || S.size == prevSize) {
      finish = true;
    } else if (S.size + nToExpand * 3 > N) {
      while (!finish) {
        const int prev2 = S.size;
        const int pc = S.vcount;
        for (int i = lane; i < pc; i += 32) P.vprev[i] = P.v[i];
        S.vcount = 0;
        __syncwarp();
        if (lane == 0) cuorb_sort::Introsort<SortEnt, SortLess>::sort(P.vprev, P.vprev + pc, SortLess());
        __syncwarp();
        for (int j = pc - 1; j >= 0; j--) {
          const int id = P.vprev[j].node;
          split_node(P, S, id, bufC, bufB, lane);
          if (S.fail) { finish = true; break; }
          erase(P, S, id, lane);
          if (S.size >= N) break;
        }
        if (S.size >= N || S.size == prev2) finish = true;
      }
    }
  }

  // list order -> output order; per node keep the max-response key, first in reference order wins
  short* order = (short*)P.vprev;
  if (lane == 0) {
    int i = 0;
    for (int c = S.head; c != NIL; c = P.next[c]) order[i++] = (short)c;
  }
  __syncwarp();
  const int nout = S.size;
  if (S.fail) { if (lane == 0) { atomicOr(flags, FLAG_NODE_OVERFLOW); key_count[slot] = 0; } return; }
  if (nout > L.key_cap) { if (lane == 0) { atomicOr(flags, FLAG_KEY_OVERFLOW); key_count[slot] = 0; } return; }
  LevelKey* out = keys + (size_t)img * dg->key_img_stride + L.key_off;
  for (int i = lane; i < nout; i += 32) {
    const int id = order[i];
    const unsigned* src = (P.buf[id] ? bufB : bufC) + P.beg[id];
    const int cnt = P.cnt[id];
    unsigned best = src[0];
    unsigned long long best_ord = ~0ull;
    bool have_ord = false;
    for (int k = 1; k < cnt; k++) {
      const unsigned c = src[k];
      const int sc = cand_s(c), sb = cand_s(best);
      if (sc < sb) continue;
      auto ord = [&](unsigned q) {
        const int x = cand_x(q), y = cand_y(q);
        const unsigned long long cell = (unsigned long long)((y - 3) / L.hCell) * L.nCols + (x - 3) / L.wCell;
        return (cell << 24) | ((unsigned long long)y << 12) | (unsigned long long)x;
      };
      if (sc > sb) { best = c; have_ord = false; continue; }
      if (!have_ord) { best_ord = ord(best); have_ord = true; }
      const unsigned long long oc = ord(c);
      if (oc < best_ord) { best = c; best_ord = oc; }
    }
    LevelKey kq;
    kq.x = (unsigned short)(cand_x(best) + FAST_BORDER);
    kq.y = (unsigned short)(cand_y(best) + FAST_BORDER);
    kq.score = (unsigned short)cand_s(best);
    kq.pad = 0;
    out[i] = kq;
  }
  if (lane == 0) key_count[slot] = nout;
}

void launch_octree(const ExtractorGeom& g, const ExtractorGeom* dg, int n_img, const unsigned* cand,
                   unsigned* cand_scratch, const int* cand_count, LevelKey* keys, int* key_count,
                   int* flags, cudaStream_t st) {
  int cap = 0;
  for (int l = 0; l < g.nlevels; l++) cap = max(cap, g.lv[l].node_cap);
  const size_t smem = pool_bytes(cap);
  static size_t attr = 48 * 1024;
  if (smem > attr) {
    cudaFuncSetAttribute(k_octree, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    attr = smem;
  }
  k_octree<<<dim3(g.nlevels, n_img), 32, smem, st>>>(dg, cand, cand_scratch, cand_count, keys,
                                                     key_count, flags, cap);
  count_launch();
}

}  // namespace cuorb
