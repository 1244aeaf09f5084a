// K4 + K8 + K9 — orientation, rBRIEF descriptor and output assembly (replaces IC_Angle /
// computeOrientation (src/ORBextractor.cc:66-91, 470-479), computeOrbDescriptor /
// computeDescriptors (:93-139, 1039-1050) and the scatter loop of operator() (:1076-1137)).
//
// k_finalize: one CTA per image.  Level offsets, then the reference's two-ended scatter
//   (keypoints inside vLappingArea go to the back, the others to the front) turned into a block
//   scan over the "stereo" flags.
// k_describe: one CTA per 16 keypoints, one warp per keypoint, three phases.
//   A  lane 0 of each warp fetches the keypoint's two patches with TMA (cp.async.bulk.tensor,
//      rank-3 maps over the batch-major planes): the 48x31-byte un-blurred box for IC_Angle and
//      the 64x37-byte blurred box that holds every rotated BRIEF tap (box x is 16-byte aligned).  The boxes land in shared
//      memory without touching L1 or registers and are signalled on the warp's mbarrier.
//      IC_Angle: lane = patch row, two 128-bit shared loads, the radius-15 disc as a per-lane
//      byte mask, the two integer moments as DP4A dot products, shuffle reduction.
//   B  one lane per keypoint: cv::fastAtan2 with explicit _rn intrinsics (no FMA contraction,
//      bit-exact with OpenCV's scalar float code) and cos/sin taken in double and rounded to float
//      (the reference uses glibc cosf/sinf; see DESIGN.md "float policy").
//   C  steered BRIEF: lane l builds descriptor byte l from 16 rotated taps read from the shared
//      blurred box; the pattern is read as pre-converted float4 pairs.
#include <cfloat>

#include "../../include/cuorb_pattern.h"
#include "common.cuh"
#include "tma.cuh"

namespace cuorb {

__device__ float4 d_patternf[256];  // pair 8l+j (x0,y0,x1,y1) as float at index 32j + l: a warp reads 512 contiguous bytes
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
  pdl_enter();
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

// exact |u| <= d disc row mask -> byte mask of word k (bit j of the nibble -> byte j = 0xFF)
__device__ __forceinline__ unsigned nibble_to_bytemask(unsigned nib) {
  return ((nib * 0x00204081u) & 0x01010101u) * 0xFFu;
}

constexpr int DESC_WARPS = 4;                                // keypoints per CTA
constexpr int IC_BYTES = IC_BOX_W * IC_BOX_H;                // 1488
constexpr int BR_BYTES = BR_BOX_W * BR_BOX_H;                // 2368
constexpr int IC_SLOT = (IC_BYTES + 127) / 128 * 128;        // 1536
constexpr int BR_SLOT = (BR_BYTES + 127) / 128 * 128;        // 2432

// grid: (ceil(kp_cap/DESC_WARPS), n_img), block = DESC_WARPS warps, warp per keypoint.  Four warps per CTA
// (0.253 ms at 16 warps, 0.226 at 8, 0.222 at 4 per 256 images); the angle is computed on lane 0 of the keypoint's
// own warp, so the only CTA barrier is the one after the level offsets
__global__ void __launch_bounds__(DESC_WARPS * 32) k_describe(const ExtractorGeom* __restrict__ dg,
                                                              const __grid_constant__ TmaMaps tmaps,
                                                              const LevelKey* __restrict__ keys,
                                                              const int* __restrict__ key_count,
                                                              const int* __restrict__ slot_map,
                                                              cuorb_keypoint* __restrict__ out_kps,
                                                              uint8_t* __restrict__ out_desc) {
  pdl_enter();
  extern __shared__ __align__(128) unsigned char smem[];     // DESC_WARPS x (IC_SLOT + BR_SLOT)
  __shared__ uint64_t s_bar[DESC_WARPS];
  __shared__ float s_angle[DESC_WARPS], s_cos[DESC_WARPS], s_sin[DESC_WARPS];
  __shared__ int   s_off[MAX_LEVELS + 1];
  const int img = blockIdx.y, nlv = dg->nlevels;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid == 0) {
    int o = 0;
    for (int l = 0; l < nlv; l++) { s_off[l] = o; o += key_count[img * nlv + l]; }
    s_off[nlv] = o;
  }
  if (lane == 0) mbar_init(&s_bar[wid], 1);
  mbar_fence_init();
  __syncthreads();
  const int n = s_off[nlv];
  const int g0 = blockIdx.x * DESC_WARPS;
  if (g0 >= n) return;
  const int g = g0 + wid;
  const bool live = g < n;

  // ---- phase A: fetch the two patches with TMA, then the integer moments of the disc ----
  int l = 0, cx = 0, cy = 0, score = 0;
  unsigned char* ic = smem + ((128u - ((unsigned)__cvta_generic_to_shared(smem) & 127u)) & 127u) + wid * (IC_SLOT + BR_SLOT);
  unsigned char* br = ic + IC_SLOT;
  if (live) {
    while (g >= s_off[l + 1]) l++;
    const LevelGeom& L = dg->lv[l];
    const LevelKey k = keys[(size_t)img * dg->key_img_stride + L.key_off + (g - s_off[l])];
    cx = k.x; cy = k.y; score = k.score;
    const int icx = ROW_LEAD + cx - HALF_PATCH_SIZE;   // padded-plane x of the patch's first column
    const int ioff = icx & 15;
    if (lane == 0) {
      mbar_expect_tx(&s_bar[wid], IC_BYTES + BR_BYTES);
      // un-blurred plane is addressed in padded-plane coordinates (ROW_LEAD / EDGE_THRESHOLD offsets)
      tma_load_3d(ic, &tmaps.pyr_ic[l], &s_bar[wid], icx - ioff, EDGE_THRESHOLD + cy - HALF_PATCH_SIZE, img);
      tma_load_3d(br, &tmaps.blur_br[l], &s_bar[wid], (cx - 18) & ~15, cy - 18, img);
    }
    mbar_wait(&s_bar[wid], 0);
    int m01 = 0, m10 = 0;
    if (lane < 31) {
      const int v = lane - HALF_PATCH_SIZE;
      const int d = c_umax[v < 0 ? -v : v];
      // byte p of the box row is u = p - ioff - 15; valid |u| <= d
      const int lo = ioff + HALF_PATCH_SIZE - d, hi = ioff + HALF_PATCH_SIZE + d;
      const unsigned long long mask = ((2ull << hi) - 1ull) & ~((1ull << lo) - 1ull);
      const uint4* rowp = reinterpret_cast<const uint4*>(ic + lane * IC_BOX_W);
      const uint4 w0 = rowp[0], w1 = rowp[1], w2 = rowp[2];
      const unsigned w[12] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w, w2.x, w2.y, w2.z, w2.w};
      unsigned S = 0, P = 0;
#pragma unroll
      for (int kk = 0; kk < 12; kk++) {
        const unsigned wm = w[kk] & nibble_to_bytemask((unsigned)(mask >> (4 * kk)) & 0xFu);
        const unsigned pos = (4u * kk) | ((4u * kk + 1u) << 8) | ((4u * kk + 2u) << 16) | ((4u * kk + 3u) << 24);
        S = __dp4a(wm, 0x01010101u, S);
        P = __dp4a(wm, pos, P);
      }
      m01 = v * (int)S;
      m10 = (int)P - (ioff + HALF_PATCH_SIZE) * (int)S;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      m01 += __shfl_xor_sync(0xffffffffu, m01, o);
      m10 += __shfl_xor_sync(0xffffffffu, m10, o);
    }
    // ---- phase B: angle and rotation on lane 0 of the keypoint's own warp (no CTA barrier) ----
    if (lane == 0) {
      const float angle = fast_atan2_deg((float)m01, (float)m10);
      const float factorPI = (float)(3.14159265358979323846 / (double)180.f);
      const float ang = __fmul_rn(angle, factorPI);
      double sn, cs;
      sincos((double)ang, &sn, &cs);
      s_angle[wid] = angle; s_cos[wid] = (float)cs; s_sin[wid] = (float)sn;
    }
    __syncwarp();
  }
  if (!live) return;

  // ---- phase C: steered BRIEF from the shared blurred box ----
  const LevelGeom& L = dg->lv[l];
  const float a = s_cos[wid], b = s_sin[wid];
  const unsigned char* bc = br + 18 * BR_BOX_W + 18 + ((cx - 18) & 15);
  unsigned val = 0;
#pragma unroll
  for (int j = 0; j < 8; j++) {
    const float4 pt = d_patternf[j * 32 + lane];
    const int iy0 = __float2int_rn(__fadd_rn(__fmul_rn(pt.x, b), __fmul_rn(pt.y, a)));
    const int ix0 = __float2int_rn(__fsub_rn(__fmul_rn(pt.x, a), __fmul_rn(pt.y, b)));
    const int iy1 = __float2int_rn(__fadd_rn(__fmul_rn(pt.z, b), __fmul_rn(pt.w, a)));
    const int ix1 = __float2int_rn(__fsub_rn(__fmul_rn(pt.z, a), __fmul_rn(pt.w, b)));
    const int t0 = bc[iy0 * BR_BOX_W + ix0], t1 = bc[iy1 * BR_BOX_W + ix1];
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
    kp.size = L.kp_size; kp.angle = s_angle[wid]; kp.response = (float)score;
    kp.octave = l; kp.class_id = -1;
    out_kps[o] = kp;
  }
}

int init_describe_constants() {
  // pattern: 256 pairs x (x0,y0,x1,y1) converted to float once
  float pf[1024];
  for (int l = 0; l < 32; l++)
    for (int j = 0; j < 8; j++)
      for (int c = 0; c < 4; c++) pf[(j * 32 + l) * 4 + c] = (float)CUORB_BIT_PATTERN_31[(l * 8 + j) * 4 + c];
  if (cudaMemcpyToSymbol(d_patternf, pf, sizeof(pf)) != cudaSuccess) return -1;
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

void launch_describe(const ExtractorGeom& g, const ExtractorGeom* dg, const TmaMaps& tmaps, int n_img, const uint8_t* pyr,
                     const uint8_t* blur, const LevelKey* keys, const int* key_count, int lap0,
                     int lap1, int* slot_map, cuorb_keypoint* out_kps, uint8_t* out_desc, int* out_n,
                     int* out_mono, cudaStream_t st) {
  (void)pyr; (void)blur;  // read through the TMA tensor maps
  const int smem = DESC_WARPS * (IC_SLOT + BR_SLOT) + 128;
  if (smem > 48 * 1024) cudaFuncSetAttribute(k_describe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);   // per device
  {
    PdlPause after_event_wait;   // the stream has just waited for the side stream's blur: a full dependency
    launch_k(k_finalize, dim3(n_img), dim3(256), 0, st, dg, keys, key_count, lap0, lap1, slot_map, out_n, out_mono);
  }
  launch_k(k_describe, dim3((g.kp_cap + DESC_WARPS - 1) / DESC_WARPS, n_img), dim3(DESC_WARPS * 32), (size_t)smem, st, dg, tmaps, keys,
           key_count, slot_map, out_kps, out_desc);
  count_launch(2);
}

}  // namespace cuorb
