// Device-side generator of the synthetic stereo workload (SURVEY.md 8d "cfg5": 100k distinct 752x480 pairs
// produced on the device, a host-certified subset).  Every pixel is a pure INTEGER function of
// (frame id, x, y) — splitmix64 hashes, 8.8 fixed-point value noise, 320 painter's-order shapes, +-2 noise,
// right eye = left resampled through a 1/256-px disparity field — restated statement by statement from
// orbslam3_cpp_fork_b200/synth_cb.py, which is the definition; tests/test_gpu_synth.py compares the two
// byte for byte.  Not part of the reference's path: this is the benchmark's input side.
#include "common.cuh"

namespace cuorb {
namespace {

constexpr unsigned long long K1 = 0x9E3779B97F4A7C15ull, K2 = 0xD1B54A32D192ED03ull;
constexpr int N_SHAPES = 320;
constexpr int S_LATTICE = 1, S_SHAPE = 16, S_NOISE_L = 7, S_NOISE_R = 8, S_DISP = 9;
constexpr long long SEED0 = 0xC0FFEE;
constexpr int TILE_W = 64, TILE_H = 16;   // one CTA of 256 threads, 4 pixels each

__host__ __device__ __forceinline__ unsigned long long splitmix64(unsigned long long x) {
  unsigned long long z = x + 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
__host__ __device__ __forceinline__ unsigned long long rnd(unsigned long long seed, int stream, unsigned long long idx) {
  return splitmix64(seed * K1 + (unsigned long long)stream * K2 + idx);
}
__device__ __forceinline__ long long lat(unsigned long long seed, int stream, long long i, long long j) {
  return (long long)(rnd(seed, stream, ((unsigned long long)j << 16) + (unsigned long long)i) >> 56) - 128;
}
// synth_cb._value_noise
__device__ __forceinline__ long long value_noise(unsigned long long seed, int stream, int cell_log2, long long x, long long y) {
  const long long cell = 1ll << cell_log2;
  const long long ix = x >> cell_log2, iy = y >> cell_log2;
  const long long fx = (x & (cell - 1)) << (8 - cell_log2), fy = (y & (cell - 1)) << (8 - cell_log2);
  const long long a = lat(seed, stream, ix, iy), b = lat(seed, stream, ix + 1, iy);
  const long long c = lat(seed, stream, ix, iy + 1), d = lat(seed, stream, ix + 1, iy + 1);
  const long long top = a * (256 - fx) + b * fx, bot = c * (256 - fx) + d * fx;
  return (top * (256 - fy) + bot * fy) >> 16;
}

struct Shape { short x0, y0, x1, y1; short kind, g, c, pad; };   // 16 bytes

// synth_cb.shape_params: one thread per (image, shape)
__global__ void k_synth_shapes(unsigned long long first_id, int n_img, int w, int h, Shape* __restrict__ shapes) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_img * N_SHAPES) return;
  const int img = t / N_SHAPES, k = t - img * N_SHAPES;
  const unsigned long long seed = (unsigned long long)(SEED0 + (long long)(first_id + img));
  auto f = [&](int field, long long lo, unsigned long long n) { return lo + (long long)(rnd(seed, S_SHAPE + field, (unsigned long long)k) % n); };
  const long long kind = f(0, 0, 4);
  long long x0 = f(1, -10, (unsigned long long)(w - 4 + 10)), y0 = f(2, -10, (unsigned long long)(h - 4 + 10));
  const long long sw = f(3, 6, 64), sh = f(4, 6, 64), g = f(5, 0, 256), c = f(6, 4, 8);
  long long x1 = x0 + sw < w ? x0 + sw : w, y1 = y0 + sh < h ? y0 + sh : h;
  x0 = x0 > 0 ? x0 : 0; y0 = y0 > 0 ? y0 : 0;
  Shape s;
  s.x0 = (short)x0; s.y0 = (short)y0; s.x1 = (short)x1; s.y1 = (short)y1;
  s.kind = (short)kind; s.g = (short)g; s.c = (short)c; s.pad = 0;
  shapes[t] = s;
}

// synth_cb.clean_left + the left eye's noise.  grid (tiles_x, tiles_y, n_img), block 256.
__global__ void __launch_bounds__(256) k_synth_left(unsigned long long first_id, int w, int h, const Shape* __restrict__ shapes,
                                                    uint8_t* __restrict__ left, uint8_t* __restrict__ clean) {
  __shared__ Shape s_list[N_SHAPES];
  __shared__ int s_n;
  const int img = blockIdx.z;
  const int tx0 = blockIdx.x * TILE_W, ty0 = blockIdx.y * TILE_H;
  const unsigned long long seed = (unsigned long long)(SEED0 + (long long)(first_id + img));
  // shapes that touch this tile, in painter's order (warp 0, ordered ballot compaction)
  if (threadIdx.x < 32) {
    int n = 0;
    const Shape* S = shapes + (size_t)img * N_SHAPES;
    for (int k0 = 0; k0 < N_SHAPES; k0 += 32) {
      const Shape s = S[k0 + threadIdx.x];
      const bool hit = s.x1 > s.x0 && s.y1 > s.y0 && s.x0 < tx0 + TILE_W && s.x1 > tx0 && s.y0 < ty0 + TILE_H && s.y1 > ty0;
      const unsigned b = __ballot_sync(0xffffffffu, hit);
      if (hit) s_list[n + __popc(b & ((1u << threadIdx.x) - 1u))] = s;
      n += __popc(b);
    }
    if (threadIdx.x == 0) s_n = n;
  }
  __syncthreads();
  const int n = s_n;
  for (int p = threadIdx.x; p < TILE_W * TILE_H; p += 256) {
    const int x = tx0 + (p & (TILE_W - 1)), y = ty0 + (p / TILE_W);
    if (x >= w || y >= h) continue;
    long long v = 128;
    const int amps[4] = {36, 24, 14, 8};
#pragma unroll
    for (int o = 0; o < 4; o++) v += (amps[o] * value_noise(seed, S_LATTICE + o, 6 - o, x, y)) >> 7;
    v = v < 0 ? 0 : (v > 255 ? 255 : v);
    for (int k = n - 1; k >= 0; k--) {   // the last shape covering the pixel wins
      const Shape s = s_list[k];
      if (x >= s.x0 && x < s.x1 && y >= s.y0 && y < s.y1) {
        if (s.kind < 3) v = s.g;
        else v = (((y / s.c) + (x / s.c)) & 1) ? s.g : 255 - s.g;
        break;
      }
    }
    const size_t o = ((size_t)img * h + y) * w + x;
    clean[o] = (uint8_t)v;
    long long nz = v + (long long)(rnd(seed, S_NOISE_L, (unsigned long long)y * w + x) % 5ull) - 2;
    left[o] = (uint8_t)(nz < 0 ? 0 : (nz > 255 ? 255 : nz));
  }
}

// synth_cb.disparity256 + the resampling of stereo_pair.  One thread per pixel.
__global__ void __launch_bounds__(256) k_synth_right(unsigned long long first_id, int w, int h, const uint8_t* __restrict__ clean,
                                                     uint8_t* __restrict__ right) {
  const int img = blockIdx.z;
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= w) return;
  const unsigned long long seed = (unsigned long long)(SEED0 + (long long)(first_id + img));
  const long long base = 20 * 256 + (long long)(rnd(seed, S_DISP, 0) % (unsigned long long)(30 * 256));
  const long long ax = (long long)(rnd(seed, S_DISP, 1) % 2049ull) - 1024, ay = (long long)(rnd(seed, S_DISP, 2) % 2049ull) - 1024;
  long long d = base + ((ax * x) >> 8) + ((ay * y) >> 8) + 16 * value_noise(seed, S_DISP + 1, 8, x, y);
  d = d < 2 * 256 ? 2 * 256 : (d > 80 * 256 ? 80 * 256 : d);
  const long long xi = x + (d >> 8), fr = d & 255;
  const uint8_t* row = clean + ((size_t)img * h + y) * w;
  const long long a = row[xi < w - 1 ? xi : w - 1], b = row[xi + 1 < w - 1 ? xi + 1 : w - 1];
  const long long res = (a * (256 - fr) + b * fr + 128) >> 8;
  long long nz = res + (long long)(rnd(seed, S_NOISE_R, (unsigned long long)y * w + x) % 5ull) - 2;
  right[((size_t)img * h + y) * w + x] = (uint8_t)(nz < 0 ? 0 : (nz > 255 ? 255 : nz));
}

}  // namespace
}  // namespace cuorb

using namespace cuorb;

extern "C" int cuorb_synth_stereo(int device, uint64_t first_frame_id, int n_pairs, int width, int height,
                                  uint8_t* d_left, uint8_t* d_right, uint8_t* d_scratch, void* stream) {
  if (n_pairs <= 0 || width < 16 || height < 16 || width > 32767 || height > 32767 || !d_left || !d_right || !d_scratch) {
    set_error("cuorb_synth_stereo: bad argument");
    return CUORB_E_BADARG;
  }
  CUORB_CUDA_TRY(cudaSetDevice(device));
  cudaStream_t st = (cudaStream_t)stream;
  Shape* shapes = nullptr;
  CUORB_CUDA_TRY(cudaMallocAsync((void**)&shapes, sizeof(Shape) * N_SHAPES * (size_t)n_pairs, st));
  const int tot = n_pairs * N_SHAPES;
  k_synth_shapes<<<(tot + 255) / 256, 256, 0, st>>>(first_frame_id, n_pairs, width, height, shapes);
  k_synth_left<<<dim3((width + TILE_W - 1) / TILE_W, (height + TILE_H - 1) / TILE_H, n_pairs), 256, 0, st>>>(
      first_frame_id, width, height, shapes, d_left, d_scratch);
  k_synth_right<<<dim3((width + 255) / 256, height, n_pairs), 256, 0, st>>>(first_frame_id, width, height, d_scratch, d_right);
  count_launch(3);
  CUORB_CUDA_TRY(cudaGetLastError());
  CUORB_CUDA_TRY(cudaFreeAsync(shapes, st));
  return n_pairs;
}
