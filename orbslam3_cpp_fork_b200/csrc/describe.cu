// K4 + K8 + K9 — orientation, rBRIEF descriptor and output assembly (replaces IC_Angle /
// computeOrientation (src/ORBextractor.cc:66-91, 470-479), computeOrbDescriptor /
// computeDescriptors (:93-139, 1039-1050) and the scatter loop of operator() (:1076-1137)).
//
// k_finalize: one CTA per image.  Level offsets, then the reference's two-ended scatter
//   (keypoints inside vLappingArea go to the back, the others to the front) turned into a block
//   scan over the "stereo" flags.
// k_describe: one warp per keypoint.  Lanes 0..30 each sum one row of the radius-15 disc
//   (int32 moments, shuffle reduction); cv::fastAtan2 is evaluated with explicit _rn intrinsics so
//   no FMA is contracted (bit-exact with OpenCV's scalar float code); lane l then builds descriptor
//   byte l from 16 rotated taps of the blurred level.  cos/sin are taken in double and rounded to
//   float (the reference uses glibc cosf/sinf; see DESIGN.md "float reproducibility").
#include <cfloat>

#include "../../include/cuorb_pattern.h"
#include "common.cuh"

namespace cuorb {

__device__ int4 d_pattern[64];  // 256 pairs x (x0,y0,x1,y1) int8; lane l reads int4 2l, 2l+1
__constant__ int c_umax[16];

// IEEE, FMA-free restatement of OpenCV's scalar fastAtan2 (core/mathfuncs_core.simd.hpp)
__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
  const float scale = (float)(180.0 / 3.14159265358979323846);
  const float p1 = __fmul_rn(0.9997878412794807f, scale), p3 = __fmul_rn(-0.3258083974640975f, scale),
              p5 = __fmul_rn(0.1555786518463281f, scale), p7 = __fmul_rn(-0.04432655554792128f, scale);
  const float ax = fabsf(x), ay = fabsf(y);
  float a, c, c2;
  if (ax >= ay) {
    c  = __fdiv_rn(ay, __fadd_rn(ax, (float)DBL_EPSILON));
    c2 = __fmul_rn(c, c);
    a  = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
  } else {
    c  = __fdiv_rn(ax, __fadd_rn(ay, (float)DBL_EPSILON));
    c2 = __fmul_rn(c, c);
    a  = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
  }
  if (x < 0) a = __fsub_rn(180.f, a);
  if (y < 0) a = __fsub_rn(360.f, a);
  return a;
}

// grid: n_img, block 256
__global__ void __launch_bounds__(256) k_finalize(const ExtractorGeom* __restrict__ dg,
                                                  const LevelKey* __restrict__ keys,
                                                  const int* __restrict__ key_count, int lap0, int lap1,
                                                  int* __restrict__ slot_map, int* __restrict__ out_n,
                                                  int* __restrict__ out_mono) {
  __shared__ int s_off[MAX_LEVELS + 1];
  __shared__ int s_warp[8];
  __shared__ int s_run;
  const int img = blockIdx.x, tid = threadIdx.x, nlv = dg->nlevels;
  if (tid == 0) {
    int o = 0;
    for (int l = 0; l < nlv; l++) { s_off[l] = o; o += key_count[img * nlv + l]; }
    s_off[nlv] = o;
    s_run = 0;
  }
  __syncthreads();
  const int n = s_off[nlv];
  const LevelKey* kimg = keys + (size_t)img * dg->key_img_stride;
  int* smap = slot_map + (size_t)img * dg->kp_cap;
  for (int base = 0; base < n; base += 256) {
    const int g = base + tid;
    int flag = 0;
    if (g < n) {
      int l = 0;
      while (g >= s_off[l + 1]) l++;
      const LevelKey k = kimg[dg->lv[l].key_off + (g - s_off[l])];
      float x = (float)k.x;
      if (l != 0) x = __fmul_rn(x, dg->lv[l].scale);       // keypoint->pt *= scale (:1120-1122)
      flag = (x >= (float)lap0 && x <= (float)lap1) ? 1 : 0;  // (:1124)
    }
    // inclusive block scan of flag
    const unsigned m = __ballot_sync(0xffffffffu, flag);
    const int lane = tid & 31, wid = tid >> 5;
    if (lane == 0) s_warp[wid] = __popc(m);
    __syncthreads();
    int before = s_run;
    for (int w = 0; w < wid; w++) before += s_warp[w];
    before += __popc(m & ((1u << lane) - 1u));  // stereo keypoints strictly before g
    if (g < n) smap[g] = flag ? (n - 1 - before) : (g - before);
    __syncthreads();
    if (tid == 0) { int t = 0; for (int w = 0; w < 8; w++) t += s_warp[w]; s_run += t; }
    __syncthreads();
  }
  if (tid == 0) { out_n[img] = n; out_mono[img] = n - s_run; }
}

// grid: (ceil(kp_cap/8), n_img), block 256 = 8 warps, warp per keypoint
__global__ void __launch_bounds__(256) k_describe(const ExtractorGeom* __restrict__ dg,
                                                  const uint8_t* __restrict__ pyr,
                                                  const uint8_t* __restrict__ blur,
                                                  const LevelKey* __restrict__ keys,
                                                  const int* __restrict__ key_count,
                                                  const int* __restrict__ slot_map,
                                                  cuorb_keypoint* __restrict__ out_kps,
                                                  uint8_t* __restrict__ out_desc) {
  const int img = blockIdx.y, nlv = dg->nlevels;
  const int lane = threadIdx.x & 31;
  const int g = blockIdx.x * 8 + (threadIdx.x >> 5);
  int l = 0, off = 0;
  {
    int c = key_count[img * nlv];
    while (g >= off + c) {
      off += c;
      if (++l == nlv) return;  // g beyond this image's keypoints
      c = key_count[img * nlv + l];
    }
  }
  const LevelGeom& L = dg->lv[l];
  const LevelKey k = keys[(size_t)img * dg->key_img_stride + L.key_off + (g - off)];
  const int cx = k.x, cy = k.y;

  // --- IC_Angle: integer moments over the disc, one row per lane ---
  const uint8_t* center = pyr + (size_t)img * dg->pyr_img_stride + L.pix_off + (size_t)cy * L.pitch + cx;
  int m01 = 0, m10 = 0;
  if (lane < 31) {
    const int v = lane - HALF_PATCH_SIZE;
    const int d = c_umax[v < 0 ? -v : v];
    const uint8_t* row = center + (ptrdiff_t)v * L.pitch;
    int s = 0, t = 0;
    for (int u = -d; u <= d; ++u) { const int val = row[u]; s += val; t += u * val; }
    m01 = v * s;
    m10 = t;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    m01 += __shfl_xor_sync(0xffffffffu, m01, o);
    m10 += __shfl_xor_sync(0xffffffffu, m10, o);
  }
  const float angle = fast_atan2_deg((float)m01, (float)m10);

  // --- steered BRIEF ---
  const float factorPI = (float)(3.14159265358979323846 / (double)180.f);
  const float ang = __fmul_rn(angle, factorPI);
  float a, b;
  {
    float ta = 0.f, tb = 0.f;
    if (lane == 0) { ta = (float)cos((double)ang); tb = (float)sin((double)ang); }
    a = __shfl_sync(0xffffffffu, ta, 0);
    b = __shfl_sync(0xffffffffu, tb, 0);
  }
  const uint8_t* bc = blur + (size_t)img * dg->blur_img_stride + L.blur_off + (size_t)cy * L.bpitch + cx;
  const int bp = L.bpitch;
  const int4 q0 = d_pattern[lane * 2], q1 = d_pattern[lane * 2 + 1];
  const int pw[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};  // one word = (x0,y0,x1,y1) int8
  unsigned val = 0;
#pragma unroll
  for (int j = 0; j < 8; j++) {
    const float x0 = (float)(signed char)(pw[j] & 0xFF), y0 = (float)(signed char)((pw[j] >> 8) & 0xFF);
    const float x1 = (float)(signed char)((pw[j] >> 16) & 0xFF), y1 = (float)(signed char)((pw[j] >> 24) & 0xFF);
    const int iy0 = __float2int_rn(__fadd_rn(__fmul_rn(x0, b), __fmul_rn(y0, a)));
    const int ix0 = __float2int_rn(__fsub_rn(__fmul_rn(x0, a), __fmul_rn(y0, b)));
    const int iy1 = __float2int_rn(__fadd_rn(__fmul_rn(x1, b), __fmul_rn(y1, a)));
    const int ix1 = __float2int_rn(__fsub_rn(__fmul_rn(x1, a), __fmul_rn(y1, b)));
    const int t0 = bc[iy0 * bp + ix0], t1 = bc[iy1 * bp + ix1];
    val |= (unsigned)(t0 < t1) << j;
  }
  // pack 4 bytes per word, lanes 0,4,8,.. store
  unsigned w = val | (__shfl_down_sync(0xffffffffu, val, 1) << 8) |
               (__shfl_down_sync(0xffffffffu, val, 2) << 16) | (__shfl_down_sync(0xffffffffu, val, 3) << 24);
  const int slot = slot_map[(size_t)img * dg->kp_cap + g];
  const size_t o = (size_t)img * dg->kp_cap + slot;
  if ((lane & 3) == 0) reinterpret_cast<unsigned*>(out_desc + o * 32)[lane >> 2] = w;
  if (lane == 0) {
    cuorb_keypoint kp;
    kp.x = (float)cx; kp.y = (float)cy;
    if (l != 0) { kp.x = __fmul_rn(kp.x, L.scale); kp.y = __fmul_rn(kp.y, L.scale); }
    kp.size = L.kp_size; kp.angle = angle; kp.response = (float)k.score;
    kp.octave = l; kp.class_id = -1;
    out_kps[o] = kp;
  }
}

int init_describe_constants() {
  // pattern: 256 pairs x 4 int8 = 1024 bytes = 64 int4
  if (cudaMemcpyToSymbol(d_pattern, CUORB_BIT_PATTERN_31, 1024) != cudaSuccess) return -1;
  // umax (src/ORBextractor.cc:446-464), same float/double sequence as the reference ctor
  int umax[16];
  int v, v0, vmax = (int)floorf(HALF_PATCH_SIZE * sqrtf(2.f) / 2 + 1);
  int vmin = (int)ceilf(HALF_PATCH_SIZE * sqrtf(2.f) / 2);
  const double hp2 = HALF_PATCH_SIZE * HALF_PATCH_SIZE;
  for (v = 0; v <= vmax; ++v) umax[v] = (int)lrint(sqrt(hp2 - v * v));
  for (v = HALF_PATCH_SIZE, v0 = 0; v >= vmin; --v) {
    while (umax[v0] == umax[v0 + 1]) ++v0;
    umax[v] = v0;
    ++v0;
  }
  if (cudaMemcpyToSymbol(c_umax, umax, sizeof(umax)) != cudaSuccess) return -1;
  return 0;
}

void launch_describe(const ExtractorGeom& g, const ExtractorGeom* dg, int n_img, const uint8_t* pyr,
                     const uint8_t* blur, const LevelKey* keys, const int* key_count, int lap0,
                     int lap1, int* slot_map, cuorb_keypoint* out_kps, uint8_t* out_desc, int* out_n,
                     int* out_mono, cudaStream_t st) {
  k_finalize<<<n_img, 256, 0, st>>>(dg, keys, key_count, lap0, lap1, slot_map, out_n, out_mono);
  k_describe<<<dim3((g.kp_cap + 7) / 8, n_img), 256, 0, st>>>(dg, pyr, blur, keys, key_count, slot_map,
                                                             out_kps, out_desc);
  count_launch(2);
}

}  // namespace cuorb
