// Shared pieces of the projection / window searches (project.cu, search.cu): the device view of a
// Frame or KeyFrame, the Hamming primitive and the per-thread device scratch arena.
#pragma once
#include <algorithm>
#include <cmath>
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
  const uint8_t* desc;
  const uint8_t* blocked0;     // mvpMapPoints[i] && Observations() > 0 on entry
  float minX, minY, maxX, maxY, gwinv, ghinv, mbf;
  float scale[MAX_LEVELS];
};

constexpr int FRAME_GRID_ROWS = 48, FRAME_GRID_COLS = 64, TH_HIGH_P = 100;
constexpr int NO_BINDER = 0x7fffffff;

__device__ __forceinline__ int hamming256(const uint4& a0, const uint4& a1, const uint4* b) {
  const uint4 b0 = b[0], b1 = b[1];
  return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
         __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

static __global__ void k_fill(int* p, int n, int v) {
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


inline size_t frame_bytes(int n) { return (size_t)n * (4 * 4 + 2 * 2 + 32 + 1 + 8 + 2) + (FRAME_GRID_ROWS * FRAME_GRID_COLS + 1) * 4 + 20 * 256; }

// host-side flattening of the Frame: SoA + PosInGrid (src/Frame.cc:867-877, std::round); the arena must
// already hold frame_bytes(n) more bytes
inline void upload_frame(Arena& A, const cuorb_frame_view* fv, cudaStream_t st, ProjFrame& F) {
  const int n = fv->n;
  std::vector<float> fx(n), fy(n);
  std::vector<int> foct(n);
  std::vector<short> cxv(n), cyv(n);
  std::vector<uint8_t> blk(n, 0);
  for (int i = 0; i < n; i++) {
    const cuorb_keypoint& k = fv->keys_un[i];
    fx[i] = k.x; fy[i] = k.y; foct[i] = k.octave;
    const int posX = (int)std::round((k.x - fv->min_x) * fv->grid_w_inv);
    const int posY = (int)std::round((k.y - fv->min_y) * fv->grid_h_inv);
    const bool in = !(posX < 0 || posX >= FRAME_GRID_COLS || posY < 0 || posY >= FRAME_GRID_ROWS);
    cxv[i] = in ? (short)posX : (short)-1; cyv[i] = in ? (short)posY : (short)-1;
    if (fv->bound_with_obs) blk[i] = fv->bound_with_obs[i];
  }
  // counting sort of the in-grid features by cell
  std::vector<int> cstart(FRAME_GRID_ROWS * FRAME_GRID_COLS + 1, 0);
  for (int i = 0; i < n; i++) if (cxv[i] >= 0) cstart[cyv[i] * FRAME_GRID_COLS + cxv[i] + 1]++;
  for (size_t c = 1; c < cstart.size(); c++) cstart[c] += cstart[c - 1];
  std::vector<unsigned short> order((size_t)std::max(n, 1));
  {
    std::vector<int> fill(cstart.begin(), cstart.end() - 1);
    for (int i = 0; i < n; i++) if (cxv[i] >= 0) order[fill[cyv[i] * FRAME_GRID_COLS + cxv[i]]++] = (unsigned short)i;
  }
  F.n = n;
  F.cstart = up(A, cstart.data(), cstart.size(), st);
  F.order = up(A, order.data(), order.size(), st);
  F.x = up(A, fx.data(), n, st); F.y = up(A, fy.data(), n, st);
  F.uright = fv->u_right ? up(A, fv->u_right, n, st) : nullptr;
  F.octave = up(A, foct.data(), n, st);
  F.cellx = up(A, cxv.data(), n, st); F.celly = up(A, cyv.data(), n, st);
  F.desc = up(A, fv->descriptors, (size_t)n * 32, st);
  F.blocked0 = up(A, blk.data(), n, st);
  F.minX = fv->min_x; F.minY = fv->min_y; F.maxX = fv->max_x; F.maxY = fv->max_y;
  F.gwinv = fv->grid_w_inv; F.ghinv = fv->grid_h_inv; F.mbf = fv->mbf;
  for (int l = 0; l < MAX_LEVELS; l++) F.scale[l] = l < fv->nlevels ? fv->scale_factors[l] : 0.f;
  cudaStreamSynchronize(st);  // the staging vectors above die with this scope
}

}  // namespace cuorb
