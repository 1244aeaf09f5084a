// Stand-alone check of the TMA path used by k_describe: rank-3 u8 tensor map, 48x37 box at an
// arbitrary byte coordinate, map passed (A) as a single __grid_constant__, (B) inside a
// __grid_constant__ array indexed at run time, (C) from global memory.
#include <cstdio>
#include <cstring>
#include <vector>
#include "../../include/cuorb.h"
#include "../../orbslam3_cpp_fork_b200/csrc/tma.cuh"
using namespace cuorb;
struct Maps { CUtensorMap m[4]; };
__device__ void fetch(const CUtensorMap* map, unsigned char* out, int x, int y, int z) {
  extern __shared__ __align__(128) unsigned char sm[];
  __shared__ uint64_t bar;
  unsigned char* dst = (unsigned char*)(((uintptr_t)sm + 127) & ~(uintptr_t)127);
  if (threadIdx.x == 0) mbar_init(&bar, 1);
  mbar_fence_init();
  __syncthreads();
  if (threadIdx.x == 0) { mbar_expect_tx(&bar, 48 * 37); tma_load_3d(dst, map, &bar, x, y, z); }
  mbar_wait(&bar, 0);
  for (int i = threadIdx.x; i < 48 * 37; i += blockDim.x) out[i] = dst[i];
}
__global__ void kA(const __grid_constant__ CUtensorMap map, unsigned char* out, int x, int y, int z) { fetch(&map, out, x, y, z); }
__global__ void kB(const __grid_constant__ Maps maps, int idx, unsigned char* out, int x, int y, int z) { fetch(&maps.m[idx], out, x, y, z); }
__global__ void kC(const CUtensorMap* map, unsigned char* out, int x, int y, int z) { fetch(map, out, x, y, z); }
namespace cuorb {
int tma_encode_u8_3d(CUtensorMap* out, void* base, uint64_t width_bytes, uint64_t rows, uint64_t images,
                     uint64_t row_pitch, uint64_t image_stride, uint32_t box_w, uint32_t box_h) {
  typedef CUresult (*encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  void* p = nullptr; cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || !p) return -1;
  const cuuint64_t dims[3] = {width_bytes, rows, images}; const cuuint64_t strides[2] = {row_pitch, image_stride};
  const cuuint32_t box[3] = {box_w, box_h, 1}; const cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = ((encode_fn)p)(out, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) printf("encode rc=%d\n", (int)r);
  return r == CUDA_SUCCESS ? 0 : -1;
}}
int main() {
  const int W = 752, H = 480, N = 3; const size_t img = (size_t)W * H;
  std::vector<unsigned char> h(img * N);
  for (size_t i = 0; i < h.size(); i++) h[i] = (unsigned char)((i * 2654435761u) >> 13);
  unsigned char *d, *o; cudaMalloc(&d, h.size()); cudaMalloc(&o, 48 * 37); cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
  Maps M; memset(&M, 0, sizeof(M));
  for (int i = 0; i < 4; i++) if (tma_encode_u8_3d(&M.m[i], d, W, H, N, W, img, 48, 37)) { printf("encode failed\n"); return 1; }
  CUtensorMap* dm; cudaMalloc(&dm, sizeof(CUtensorMap)); cudaMemcpy(dm, &M.m[0], sizeof(CUtensorMap), cudaMemcpyHostToDevice);
  const int x = 101, y = 57, z = 2;
  for (int mode = 0; mode < 3; mode++) {
    cudaMemset(o, 0, 48 * 37);
    if (mode == 0) kA<<<1, 128, 48 * 37 + 256>>>(M.m[0], o, x, y, z);
    if (mode == 1) kB<<<1, 128, 48 * 37 + 256>>>(M, 2, o, x, y, z);
    if (mode == 2) kC<<<1, 128, 48 * 37 + 256>>>(dm, o, x, y, z);
    cudaError_t e = cudaDeviceSynchronize();
    std::vector<unsigned char> r(48 * 37); cudaMemcpy(r.data(), o, r.size(), cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int j = 0; j < 37; j++) for (int i = 0; i < 48; i++) bad += r[j * 48 + i] != h[z * img + (size_t)(y + j) * W + x + i];
    printf("mode %c: %s, mismatches %d\n", 'A' + mode, cudaGetErrorString(e), bad);
    if (e != cudaSuccess) return 2;
  }
  return 0;
}
