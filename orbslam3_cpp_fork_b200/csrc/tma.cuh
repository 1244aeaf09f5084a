// TMA (cp.async.bulk.tensor) + mbarrier helpers for sm_100a, hand-written PTX.
// Tensor maps describe the batch-major u8 planes of one pyramid level as rank-3 tensors
// (x bytes, y rows, image); a box is fetched into shared memory by one elected lane and signalled
// on an mbarrier, bypassing L1 and the register file.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace cuorb {

// boxes (bytes x rows) fetched per keypoint by k_describe
// (the innermost box coordinate must be a multiple of 16 bytes: boxes start at x & ~15 and are 15 B wider)
constexpr int IC_BOX_W = 48, IC_BOX_H = 31;    // un-blurred patch, x from (cx-15) & ~15, y from cy-15
constexpr int BR_BOX_W = 64, BR_BOX_H = 37;    // blurred patch,   x from (cx-18) & ~15, y from cy-18

// per extractor handle: one map per (plane kind, level), in device global memory
struct TmaMaps {
  CUtensorMap pyr_ic[CUORB_MAX_LEVELS];
  CUtensorMap blur_br[CUORB_MAX_LEVELS];
};

// k_fast_persist: one map per level over the padded pyramid plane, box FT_PITCH bytes x (max cell height + 6) rows
constexpr int FT_PITCH = 80;   // 20 words: rows r and r+8 share banks, so a warp's ~3 rows never collide
struct FastMaps {
  CUtensorMap lv[CUORB_MAX_LEVELS];
};

#ifdef __CUDACC__
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
  unsigned ok = 0;
  const unsigned a = smem_u32(bar);
  while (!ok) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(a), "r"(parity)
        : "memory");
  }
}
// rank-3 tiled load: box at (x, y, z) of `map` -> dst (128-byte aligned shared memory)
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* map, uint64_t* bar, int x, int y, int z) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(dst)), "l"((unsigned long long)map), "r"(smem_u32(bar)), "r"(x), "r"(y), "r"(z)
      : "memory");
}
// 1-D bulk copy global -> shared (both 16-byte aligned, bytes a multiple of 16), completion on an mbarrier
__device__ __forceinline__ void bulk_load_1d(void* dst, const void* src, unsigned bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst)), "l"((unsigned long long)src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
#endif

// host: encode a rank-3 u8 map {x bytes, rows, images}; returns 0 on success
int tma_encode_u8_3d(CUtensorMap* out, void* base, uint64_t width_bytes, uint64_t rows, uint64_t images,
                     uint64_t row_pitch, uint64_t image_stride, uint32_t box_w, uint32_t box_h);

}  // namespace cuorb
