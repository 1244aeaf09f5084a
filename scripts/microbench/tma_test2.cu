// Same check with the CUDA programming guide's libcu++ wrappers, to separate "bad tensor map" from "bad PTX".
#include <cstdio>
#include <cstring>
#include <vector>
#include <cuda.h>
#include <cuda/barrier>
using barrier = cuda::barrier<cuda::thread_scope_block>;
namespace cde = cuda::device::experimental;
__global__ void k3(const __grid_constant__ CUtensorMap map, unsigned char* out, int x, int y, int z) {
  __shared__ alignas(128) unsigned char dst[48 * 37 + 16];
#pragma nv_diag_suppress static_var_with_dynamic_init
  __shared__ barrier bar;
  if (threadIdx.x == 0) { init(&bar, blockDim.x); cde::fence_proxy_async_shared_cta(); }
  __syncthreads();
  barrier::arrival_token token;
  if (threadIdx.x == 0) {
    cde::cp_async_bulk_tensor_3d_global_to_shared(dst, &map, x, y, z, bar);
    token = cuda::device::barrier_arrive_tx(bar, 1, 48 * 37);
  } else token = bar.arrive();
  bar.wait(std::move(token));
  for (int i = threadIdx.x; i < 48 * 37; i += blockDim.x) out[i] = dst[i];
}
int main() {
  const int W = 752, H = 480, N = 3; const size_t img = (size_t)W * H;
  std::vector<unsigned char> h(img * N);
  for (size_t i = 0; i < h.size(); i++) h[i] = (unsigned char)((i * 2654435761u) >> 13);
  unsigned char *d, *o; cudaMalloc(&d, h.size()); cudaMalloc(&o, 48 * 37); cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
  typedef CUresult (*encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  void* p = nullptr; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
  for (int boxw : {48, 64, 32}) {
    CUtensorMap m;
    const cuuint64_t dims[3] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N}; const cuuint64_t strides[2] = {(cuuint64_t)W, img};
    const cuuint32_t box[3] = {(cuuint32_t)boxw, 37, 1}; const cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = ((encode_fn)p)(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("box %d encode rc=%d\n", boxw, (int)r);
    if (boxw != 48) continue;
    for (int x : {96, 101}) {
      k3<<<1, 128>>>(m, o, x, 57, 2);
      cudaError_t e = cudaDeviceSynchronize();
      std::vector<unsigned char> rr(48 * 37); cudaMemcpy(rr.data(), o, rr.size(), cudaMemcpyDeviceToHost);
      int bad = 0;
      for (int j = 0; j < 37; j++) for (int i = 0; i < 48; i++) bad += rr[j * 48 + i] != h[2 * img + (size_t)(57 + j) * W + x + i];
      printf("x=%d: %s, mismatches %d\n", x, cudaGetErrorString(e), bad);
      if (e != cudaSuccess) return 2;
    }
  }
  return 0;
}
