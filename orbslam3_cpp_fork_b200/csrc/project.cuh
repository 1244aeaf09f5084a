// Shared pieces of the projection / window searches (project.cu, search.cu): the device view of a
// Frame or KeyFrame, the Hamming primitive and the per-thread device scratch arena.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include "common.cuh"

namespace cuorb {

struct ProjFrame {  // device SoA of the frame members the searches read
  int n;
  const float *x, *y, *uright;
  const int *octave;
  const short *cellx, *celly;  // PosInGrid result, -1 if outside the grid
  // features inside the grid, ordered by cell (row-major: celly * 64 + cellx); cstart[c] .. cstart[c + 1] = cell c
  // (Frame::AssignFeaturesToGrid's mGrid as CSR, src/Frame.cc:524-553); the features of cells cx0..cx1 of one grid
  // row are one contiguous range
  const int* cstart;
  const unsigned short* order;
  // the same features as one 16-byte record per CSR position (pack_feature): what the projection rounds read, so a
  // window row is one coalesced load instead of an index and six gathers
  const float4* rec;
  const uint8_t* desc;
  const uint8_t* blocked0;     // mvpMapPoints[i] && Observations() > 0 on entry
  float minX, minY, maxX, maxY, gwinv, ghinv, mbf;
  float scale[MAX_LEVELS];
};

constexpr int FRAME_GRID_ROWS = 48, FRAME_GRID_COLS = 64, TH_HIGH_P = 100;
constexpr int NO_BINDER = 0x7fffffff;

// CSR record of feature f: x, y, uRight (-1 when the frame has none), and f | octave << 16 | cellx << 21 | blocked0 << 27
__host__ __device__ __forceinline__ unsigned pack_feature_bits(int f, int octave, int cellx, int blocked) {
  return (unsigned)f | ((unsigned)octave << 16) | ((unsigned)cellx << 21) | ((unsigned)(blocked != 0) << 27);
}
static_assert(MAX_LEVELS <= 32, "octave field of the CSR record is 5 bits");

__device__ __forceinline__ int hamming256(const uint4& a0, const uint4& a1, const uint4* b) {
  const uint4 b0 = b[0], b1 = b[1];
  return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
         __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

static __global__ void k_fill(int* p, int n, int v) {
  pdl_enter();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

struct Arena {  // grow-only device scratch per thread
  void* p = nullptr; size_t cap = 0, used = 0; int device = -1;
  int reserve(int dev, size_t bytes) {
    if (dev != device || bytes > cap) {
      if (p) { cudaSetDevice(device); cudaFree(p); p = nullptr; cap = 0; }
      cudaSetDevice(dev);
      if (cudaMalloc(&p, bytes) != cudaSuccess) { p = nullptr; return -1; }
      cap = bytes; device = dev;
    }
    used = 0;
    return 0;
  }
  template <typename T> T* take(size_t count) {
    used = (used + 255) & ~(size_t)255;
    T* r = reinterpret_cast<T*>(static_cast<char*>(p) + used);
    used += count * sizeof(T);
    return r;
  }
};
extern thread_local Arena t_arena;

template <typename T>
inline T* up(Arena& A, const T* host, size_t count, cudaStream_t st) {
  T* d = A.take<T>(count);
  if (host && count) cudaMemcpyAsync(d, host, count * sizeof(T), cudaMemcpyHostToDevice, st);
  return d;
}


inline size_t frame_bytes(int n) { return (size_t)n * (4 * 4 + 2 * 2 + 32 + 1 + 12 + 2 + 16) + (FRAME_GRID_ROWS * FRAME_GRID_COLS + 1) * 4 + 20 * 256; }

struct HostStage {  // grow-only pinned staging per host thread: a Frame's SoA goes up as ONE copy
  void* p = nullptr; size_t cap = 0;
  cudaEvent_t sent = nullptr; int ev_device = -1; bool pending = false;   // the last copy out of the block
  // the block, free for writing: waits for the copy enqueued by the previous user (normally long complete - every
  // search synchronises its stream before it returns)
  void* get(size_t bytes) {
    if (pending) { cudaEventSynchronize(sent); pending = false; }
    if (bytes > cap) {
      if (p) { cudaFreeHost(p); p = nullptr; cap = 0; }
      if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) { p = nullptr; return nullptr; }
      cap = bytes;
    }
    return p;
  }
  // call right after enqueueing the copy out of the block on `st`; false = no event (the caller must synchronise)
  bool mark_sent(cudaStream_t st) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (sent && dev != ev_device) { cudaEventDestroy(sent); sent = nullptr; }
    if (!sent) { if (cudaEventCreateWithFlags(&sent, cudaEventDisableTiming) != cudaSuccess) { sent = nullptr; return false; } ev_device = dev; }
    if (cudaEventRecord(sent, st) != cudaSuccess) return false;
    pending = true;
    return true;
  }
};
extern thread_local HostStage t_stage;

// host-side flattening of the Frame: SoA + PosInGrid (src/Frame.cc:867-877, std::round) + the cell-sorted CSR, laid
// out in one pinned staging block and sent with one copy (ten small pageable copies before: ~35 us of a 0.3-0.5 ms
// call); the arena must already hold frame_bytes(n) more bytes.  Does not wait for the copy: the staging block is
// guarded by an event that the next upload on this thread waits for.
inline void upload_frame(Arena& A, const cuorb_frame_view* fv, cudaStream_t st, ProjFrame& F) {
  const int n = fv->n, nn = std::max(n, 1);
  const int ncell = FRAME_GRID_ROWS * FRAME_GRID_COLS + 1;
  size_t off = 0;
  auto slot = [&](size_t bytes) { const size_t o = off; off = (off + bytes + 255) & ~(size_t)255; return o; };
  const size_t o_cstart = slot(sizeof(int) * ncell), o_order = slot(2 * (size_t)nn), o_x = slot(4 * (size_t)nn), o_y = slot(4 * (size_t)nn),
               o_ur = slot(fv->u_right ? 4 * (size_t)nn : 0), o_oct = slot(4 * (size_t)nn), o_cx = slot(2 * (size_t)nn),
               o_cy = slot(2 * (size_t)nn), o_desc = slot(32 * (size_t)nn), o_blk = slot((size_t)nn), o_rec = slot(16 * (size_t)nn);
  const size_t total = off;
  unsigned char* h = (unsigned char*)t_stage.get(total);
  std::vector<unsigned char> fallback;
  if (!h) { fallback.resize(total); h = fallback.data(); }   // pinned allocation failed: a pageable block still works
  int* cstart = (int*)(h + o_cstart); unsigned short* order = (unsigned short*)(h + o_order);
  float *fx = (float*)(h + o_x), *fy = (float*)(h + o_y);
  int* foct = (int*)(h + o_oct);
  short *cxv = (short*)(h + o_cx), *cyv = (short*)(h + o_cy);
  uint8_t* blk = h + o_blk;
  for (int c = 0; c < ncell; c++) cstart[c] = 0;
  for (int i = 0; i < n; i++) {
    const cuorb_keypoint& k = fv->keys_un[i];
    fx[i] = k.x; fy[i] = k.y; foct[i] = k.octave;
    const int posX = (int)std::round((k.x - fv->min_x) * fv->grid_w_inv);
    const int posY = (int)std::round((k.y - fv->min_y) * fv->grid_h_inv);
    const bool in = !(posX < 0 || posX >= FRAME_GRID_COLS || posY < 0 || posY >= FRAME_GRID_ROWS);
    cxv[i] = in ? (short)posX : (short)-1; cyv[i] = in ? (short)posY : (short)-1;
    blk[i] = fv->bound_with_obs ? fv->bound_with_obs[i] : 0;
    if (in) cstart[posY * FRAME_GRID_COLS + posX + 1]++;          // counting sort of the in-grid features by cell
  }
  for (int c = 1; c < ncell; c++) cstart[c] += cstart[c - 1];
  {
    std::vector<int> fill(cstart, cstart + ncell - 1);
    for (int i = 0; i < n; i++) if (cxv[i] >= 0) order[fill[cyv[i] * FRAME_GRID_COLS + cxv[i]]++] = (unsigned short)i;
  }
  {
    struct Rec { float x, y, ur; unsigned bits; };
    Rec* rec = (Rec*)(h + o_rec);
    for (int j = 0; j < cstart[ncell - 1]; j++) {
      const int f = order[j];
      rec[j] = Rec{fx[f], fy[f], fv->u_right ? fv->u_right[f] : -1.0f, pack_feature_bits(f, foct[f], cxv[f], blk[f])};
    }
  }
  if (n) memcpy(h + o_desc, fv->descriptors, 32 * (size_t)n);
  if (fv->u_right && n) memcpy(h + o_ur, fv->u_right, 4 * (size_t)n);
  unsigned char* d = A.take<unsigned char>(total);
  cudaMemcpyAsync(d, h, total, cudaMemcpyHostToDevice, st);
  if (h != t_stage.p || !t_stage.mark_sent(st)) cudaStreamSynchronize(st);   // pageable fallback block / no event: drain now
  F.n = n;
  F.cstart = (const int*)(d + o_cstart); F.order = (const unsigned short*)(d + o_order);
  F.rec = (const float4*)(d + o_rec);
  F.x = (const float*)(d + o_x); F.y = (const float*)(d + o_y);
  F.uright = fv->u_right ? (const float*)(d + o_ur) : nullptr;
  F.octave = (const int*)(d + o_oct);
  F.cellx = (const short*)(d + o_cx); F.celly = (const short*)(d + o_cy);
  F.desc = d + o_desc;
  F.blocked0 = d + o_blk;
  F.minX = fv->min_x; F.minY = fv->min_y; F.maxX = fv->max_x; F.maxY = fv->max_y;
  F.gwinv = fv->grid_w_inv; F.ghinv = fv->grid_h_inv; F.mbf = fv->mbf;
  for (int l = 0; l < MAX_LEVELS; l++) F.scale[l] = l < fv->nlevels ? fv->scale_factors[l] : 0.f;
}

}  // namespace cuorb
