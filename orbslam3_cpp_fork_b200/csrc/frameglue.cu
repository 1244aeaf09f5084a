// Per-frame glue right after ORBextractor::operator():
//   Frame::UndistortKeyPoints    src/Frame.cc:886-913  (cv::undistortPoints(mat, mat, K, D, cv::Mat(), K))
//   Frame::ComputeStereoFromRGBD src/Frame.cc:1115-1133
// cv::undistortPoints' arithmetic (calib3d cvUndistortPointsInternal: doubles, 5 fixed iterations,
// projection by P = K, result rounded to float) is evaluated with explicit round-to-nearest
// double intrinsics in the reference's operation order, so no FMA contraction can change a bit.
// One thread per keypoint; these are tiny kernels (a frame has ~1000 keypoints) whose point is to
// keep the Frame's arrays on the device between extraction and the matchers.
#include <mutex>
#include <string>

#include "common.cuh"
#include "frustum.cuh"
#include "undistort.cuh"

namespace cuorb {
void set_error(const std::string& msg);

namespace {

__global__ void k_undistort(const cuorb_keypoint* __restrict__ in, cuorb_keypoint* __restrict__ out, int n, const UndistortArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  cuorb_keypoint kp = in[i];
  undistort_point(a, kp.x, kp.y);
  out[i] = kp;
}

__global__ void k_stereo_from_rgbd(const cuorb_keypoint* __restrict__ kps, const cuorb_keypoint* __restrict__ kps_un, int n,
                                   const float* __restrict__ depth, size_t stride_floats, int w, int h, float mbf,
                                   float* __restrict__ u_right, float* __restrict__ depth_out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int u = (int)kps[i].x, v = (int)kps[i].y;  // imDepth.at<float>(v, u): float -> int truncation
  float ur = -1.0f, dz = -1.0f;
  if (u >= 0 && u < w && v >= 0 && v < h) {
    const float d = depth[(size_t)v * stride_floats + u];
    if (d > 0) { dz = d; ur = __fsub_rn(kps_un[i].x, __fdiv_rn(mbf, d)); }
  }
  u_right[i] = ur; depth_out[i] = dz;
}

// Frame::isInFrustum: one thread per MapPoint, SoA in / SoA out (the arithmetic is frustum_point, frustum.cuh)
__global__ void k_is_in_frustum(const cuorb_frustum_camera c, int n, const float* __restrict__ P3, const float* __restrict__ N3,
                                const float* __restrict__ minD, const float* __restrict__ maxD, float limit,
                                uint8_t* __restrict__ in_view, float* __restrict__ projX, float* __restrict__ projY,
                                float* __restrict__ projXR, int* __restrict__ level, float* __restrict__ viewCosOut,
                                float* __restrict__ depth, int* __restrict__ count) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  bool ok = false;
  if (i < n) {
    FrustumOut o;
    ok = frustum_point(c, P3[3 * i], P3[3 * i + 1], P3[3 * i + 2], N3[3 * i], N3[3 * i + 1], N3[3 * i + 2], minD[i], maxD[i], limit, o);
    in_view[i] = ok ? 1 : 0; projX[i] = o.px; projY[i] = o.py; projXR[i] = o.pxr; level[i] = o.lv; viewCosOut[i] = o.vc; depth[i] = o.dz;
  }
  const unsigned b = __ballot_sync(0xffffffffu, ok);
  if ((threadIdx.x & 31) == 0 && b) atomicAdd(count, __popc(b));
}

struct Scratch {  // grow-only device scratch per host thread
  void* p = nullptr; size_t cap = 0; int device = -1;
  void* get(int dev, size_t bytes) {
    if (dev != device || bytes > cap) {
      if (p) { cudaSetDevice(device); cudaFree(p); p = nullptr; cap = 0; }
      cudaSetDevice(dev);
      if (cudaMalloc(&p, bytes) != cudaSuccess) { p = nullptr; return nullptr; }
      cap = bytes; device = dev;
    }
    return p;
  }
};
thread_local Scratch t_scratch;

}  // namespace
}  // namespace cuorb

using namespace cuorb;

#define GLUE_TRY(call)                                                             \
  do {                                                                             \
    cudaError_t err__ = (call);                                                    \
    if (err__ != cudaSuccess) { set_error(std::string(#call) + ": " + cudaGetErrorString(err__)); return CUORB_E_CUDA; } \
  } while (0)

extern "C" {

int cuorb_undistort_keypoints(int device, int n, const cuorb_keypoint* kps, const float* K4, const float* dist5,
                              cuorb_keypoint* kps_un) {
  if (n < 0 || (n > 0 && (!kps || !kps_un)) || !K4 || !dist5) { set_error("cuorb_undistort_keypoints: bad argument"); return CUORB_E_BADARG; }
  if (n == 0) return 0;
  if (dist5[0] == 0.0f) {  // src/Frame.cc:887-890: mvKeysUn = mvKeys
    if (kps_un != kps) memcpy(kps_un, kps, sizeof(cuorb_keypoint) * (size_t)n);
    return n;
  }
  GLUE_TRY(cudaSetDevice(device));
  const size_t bytes = sizeof(cuorb_keypoint) * (size_t)n;
  char* d = (char*)t_scratch.get(device, 2 * bytes + 512);
  if (!d) { set_error("cudaMalloc failed (undistort workspace)"); return CUORB_E_CUDA; }
  cuorb_keypoint* din = (cuorb_keypoint*)d;
  cuorb_keypoint* dout = (cuorb_keypoint*)(d + ((bytes + 255) & ~(size_t)255));
  const UndistortArgs a = undistort_args(K4, dist5);
  cudaStream_t st = cudaStreamPerThread;
  GLUE_TRY(cudaMemcpyAsync(din, kps, bytes, cudaMemcpyHostToDevice, st));
  k_undistort<<<(n + 127) / 128, 128, 0, st>>>(din, dout, n, a);
  count_launch();
  GLUE_TRY(cudaMemcpyAsync(kps_un, dout, bytes, cudaMemcpyDeviceToHost, st));
  GLUE_TRY(cudaStreamSynchronize(st));
  return n;
}

int cuorb_stereo_from_rgbd(int device, int n, const cuorb_keypoint* kps, const cuorb_keypoint* kps_un, const float* depth,
                           int width, int height, size_t stride_bytes, cuorb_mem depth_where, float mbf, float* u_right,
                           float* depth_out) {
  if (n < 0 || (n > 0 && (!kps || !kps_un || !u_right || !depth_out)) || !depth || width <= 0 || height <= 0 ||
      stride_bytes < (size_t)width * 4 || (stride_bytes & 3)) {
    set_error("cuorb_stereo_from_rgbd: bad argument"); return CUORB_E_BADARG;
  }
  if (n == 0) return 0;
  GLUE_TRY(cudaSetDevice(device));
  const size_t kb = (sizeof(cuorb_keypoint) * (size_t)n + 255) & ~(size_t)255, fb = (4 * (size_t)n + 255) & ~(size_t)255;
  const size_t img = depth_where == CUORB_MEM_HOST ? (size_t)width * 4 * height : 0;
  char* d = (char*)t_scratch.get(device, 2 * kb + 2 * fb + img + 512);
  if (!d) { set_error("cudaMalloc failed (rgbd workspace)"); return CUORB_E_CUDA; }
  cuorb_keypoint* dk = (cuorb_keypoint*)d;
  cuorb_keypoint* dku = (cuorb_keypoint*)(d + kb);
  float* du = (float*)(d + 2 * kb);
  float* dz = (float*)(d + 2 * kb + fb);
  cudaStream_t st = cudaStreamPerThread;
  const float* ddepth = depth;
  size_t dstride = stride_bytes / 4;
  if (depth_where == CUORB_MEM_HOST) {
    float* dd = (float*)(d + 2 * kb + 2 * fb);
    GLUE_TRY(cudaMemcpy2DAsync(dd, (size_t)width * 4, depth, stride_bytes, (size_t)width * 4, height, cudaMemcpyHostToDevice, st));
    ddepth = dd; dstride = width;
  }
  GLUE_TRY(cudaMemcpyAsync(dk, kps, sizeof(cuorb_keypoint) * (size_t)n, cudaMemcpyHostToDevice, st));
  GLUE_TRY(cudaMemcpyAsync(dku, kps_un, sizeof(cuorb_keypoint) * (size_t)n, cudaMemcpyHostToDevice, st));
  k_stereo_from_rgbd<<<(n + 127) / 128, 128, 0, st>>>(dk, dku, n, ddepth, dstride, width, height, mbf, du, dz);
  count_launch();
  GLUE_TRY(cudaMemcpyAsync(u_right, du, 4 * (size_t)n, cudaMemcpyDeviceToHost, st));
  GLUE_TRY(cudaMemcpyAsync(depth_out, dz, 4 * (size_t)n, cudaMemcpyDeviceToHost, st));
  GLUE_TRY(cudaStreamSynchronize(st));
  return n;
}

int cuorb_is_in_frustum(int device, const cuorb_frustum_camera* cam, int n, const float* world_pos, const float* normal,
                        const float* min_distance, const float* max_distance, float viewing_cos_limit, uint8_t* in_view,
                        float* proj_x, float* proj_y, float* proj_xr, int32_t* scale_level, float* view_cos, float* depth) {
  if (!cam || n < 0 || (n > 0 && (!world_pos || !normal || !min_distance || !max_distance || !in_view || !proj_x || !proj_y ||
                                  !proj_xr || !scale_level || !view_cos || !depth))) {
    set_error("cuorb_is_in_frustum: bad argument"); return CUORB_E_BADARG;
  }
  if (n == 0) return 0;
  GLUE_TRY(cudaSetDevice(device));
  const size_t a4 = (4 * (size_t)n + 255) & ~(size_t)255, a12 = (12 * (size_t)n + 255) & ~(size_t)255;
  const size_t a1 = ((size_t)n + 255) & ~(size_t)255;
  char* d = (char*)t_scratch.get(device, 2 * a12 + 2 * a4 + a1 + 6 * a4 + 512);
  if (!d) { set_error("cudaMalloc failed (frustum workspace)"); return CUORB_E_CUDA; }
  float* dP = (float*)d; float* dN = (float*)(d + a12);
  float* dmin = (float*)(d + 2 * a12); float* dmax = (float*)(d + 2 * a12 + a4);
  char* o = d + 2 * a12 + 2 * a4;
  uint8_t* dview = (uint8_t*)o; o += a1;
  float* dpx = (float*)o; o += a4; float* dpy = (float*)o; o += a4; float* dpxr = (float*)o; o += a4;
  int* dlv = (int*)o; o += a4; float* dvc = (float*)o; o += a4; float* ddz = (float*)o; o += a4;
  int* dcount = (int*)o;
  cudaStream_t st = cudaStreamPerThread;
  GLUE_TRY(cudaMemcpyAsync(dP, world_pos, 12 * (size_t)n, cudaMemcpyHostToDevice, st));
  GLUE_TRY(cudaMemcpyAsync(dN, normal, 12 * (size_t)n, cudaMemcpyHostToDevice, st));
  GLUE_TRY(cudaMemcpyAsync(dmin, min_distance, 4 * (size_t)n, cudaMemcpyHostToDevice, st));
  GLUE_TRY(cudaMemcpyAsync(dmax, max_distance, 4 * (size_t)n, cudaMemcpyHostToDevice, st));
  GLUE_TRY(cudaMemsetAsync(dcount, 0, sizeof(int), st));
  k_is_in_frustum<<<(n + 127) / 128, 128, 0, st>>>(*cam, n, dP, dN, dmin, dmax, viewing_cos_limit, dview, dpx, dpy, dpxr, dlv,
                                                    dvc, ddz, dcount);
  count_launch();
  GLUE_TRY(cudaGetLastError());
  int count = 0;
  GLUE_TRY(cudaMemcpyAsync(in_view, dview, (size_t)n, cudaMemcpyDeviceToHost, st));
  GLUE_TRY(cudaMemcpyAsync(proj_x, dpx, 4 * (size_t)n, cudaMemcpyDeviceToHost, st));
  GLUE_TRY(cudaMemcpyAsync(proj_y, dpy, 4 * (size_t)n, cudaMemcpyDeviceToHost, st));
  GLUE_TRY(cudaMemcpyAsync(proj_xr, dpxr, 4 * (size_t)n, cudaMemcpyDeviceToHost, st));
  GLUE_TRY(cudaMemcpyAsync(scale_level, dlv, 4 * (size_t)n, cudaMemcpyDeviceToHost, st));
  GLUE_TRY(cudaMemcpyAsync(view_cos, dvc, 4 * (size_t)n, cudaMemcpyDeviceToHost, st));
  GLUE_TRY(cudaMemcpyAsync(depth, ddz, 4 * (size_t)n, cudaMemcpyDeviceToHost, st));
  GLUE_TRY(cudaMemcpyAsync(&count, dcount, sizeof(int), cudaMemcpyDeviceToHost, st));
  GLUE_TRY(cudaStreamSynchronize(st));
  return count;
}

// ---- allocation-free variant (no counterpart in the reference) ----
// cuorb_is_in_frustum above stages its SoA in a grow-only cudaMalloc'ed workspace.  Callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it) use this entry point instead: the
// same kernel, the same results, with the staging area in a file-scope device array of fixed size.
//   * limit: CUORB_FRUSTUM_NOALLOC_MAX MapPoints per call (larger n is rejected with CUORB_E_CAPACITY);
//   * cost: every GPU that loads libcuorb holds the array (about 3.6 MiB) until the process exits;
//   * the array is shared by ALL calls on a device and keeps the previous call's contents: the one word that has
//     to start at zero (the in-view counter) is cleared by every call, everything else is overwritten before it is
//     read; two calls that overlap on the same GPU (two host threads, two streams) race on it - serialise them.
// No cudaMalloc / cudaFree / cudaMallocAsync and no allocating library routine is called here.
}  // extern "C"

namespace {
constexpr size_t FRUSTUM_NA_A4  = ((size_t)4 * CUORB_FRUSTUM_NOALLOC_MAX + 255) & ~(size_t)255;
constexpr size_t FRUSTUM_NA_A12 = ((size_t)12 * CUORB_FRUSTUM_NOALLOC_MAX + 255) & ~(size_t)255;
constexpr size_t FRUSTUM_NA_A1  = ((size_t)CUORB_FRUSTUM_NOALLOC_MAX + 255) & ~(size_t)255;
constexpr size_t FRUSTUM_NA_BYTES = 2 * FRUSTUM_NA_A12 + 2 * FRUSTUM_NA_A4 + FRUSTUM_NA_A1 + 6 * FRUSTUM_NA_A4 + 256;
}  // namespace
// In HOST code this name is only a stand-in for the symbol, not a device address: the address a kernel may be
// given comes from cudaGetSymbolAddress (frustum_scratch_ptr below), once per device with that device current.
static __device__ __align__(256) unsigned char g_frustum_scratch[FRUSTUM_NA_BYTES];

namespace {
// device address of g_frustum_scratch on `device` (which must be the current device); nullptr + error text on failure
unsigned char* frustum_scratch_ptr(int device) {
  static std::mutex mu;
  static unsigned char* ptrs[64];
  if (device < 0 || device >= 64) { set_error("cuorb_is_in_frustum_noalloc: device ordinal beyond 63"); return nullptr; }
  std::lock_guard<std::mutex> lk(mu);
  if (!ptrs[device]) {
    void* p = nullptr;
    const cudaError_t err = cudaGetSymbolAddress(&p, g_frustum_scratch);
    if (err != cudaSuccess || !p) {
      set_error(std::string("cudaGetSymbolAddress(g_frustum_scratch): ") + cudaGetErrorString(err));
      return nullptr;
    }
    ptrs[device] = (unsigned char*)p;
  }
  return ptrs[device];
}
}  // namespace

extern "C" {

int cuorb_is_in_frustum_noalloc(int device, const cuorb_frustum_camera* cam, int n, const float* world_pos,
                                const float* normal, const float* min_distance, const float* max_distance,
                                float viewing_cos_limit, uint8_t* in_view, float* proj_x, float* proj_y, float* proj_xr,
                                int32_t* scale_level, float* view_cos, float* depth) {
  if (!cam || n < 0 || (n > 0 && (!world_pos || !normal || !min_distance || !max_distance || !in_view || !proj_x || !proj_y ||
                                  !proj_xr || !scale_level || !view_cos || !depth))) {
    set_error("cuorb_is_in_frustum_noalloc: bad argument"); return CUORB_E_BADARG;
  }
  if (n > CUORB_FRUSTUM_NOALLOC_MAX) {
    set_error("cuorb_is_in_frustum_noalloc: n = " + std::to_string(n) + " MapPoints exceeds the fixed scratch array (" +
              "CUORB_FRUSTUM_NOALLOC_MAX = " + std::to_string(CUORB_FRUSTUM_NOALLOC_MAX) +
              "); split the call or use cuorb_is_in_frustum");
    return CUORB_E_CAPACITY;
  }
  if (n == 0) return 0;
  GLUE_TRY(cudaSetDevice(device));
  unsigned char* d = frustum_scratch_ptr(device);
  if (!d) return CUORB_E_CUDA;
  // same layout as cuorb_is_in_frustum's workspace, sized for this n
  const size_t a4 = (4 * (size_t)n + 255) & ~(size_t)255, a12 = (12 * (size_t)n + 255) & ~(size_t)255;
  const size_t a1 = ((size_t)n + 255) & ~(size_t)255;
  float* dP = (float*)d; float* dN = (float*)(d + a12);
  float* dmin = (float*)(d + 2 * a12); float* dmax = (float*)(d + 2 * a12 + a4);
  unsigned char* o = d + 2 * a12 + 2 * a4;
  uint8_t* dview = (uint8_t*)o; o += a1;
  float* dpx = (float*)o; o += a4; float* dpy = (float*)o; o += a4; float* dpxr = (float*)o; o += a4;
  int* dlv = (int*)o; o += a4; float* dvc = (float*)o; o += a4; float* ddz = (float*)o; o += a4;
  int* dcount = (int*)o;
  cudaStream_t st = cudaStreamPerThread;
  GLUE_TRY(cudaMemcpyAsync(dP, world_pos, 12 * (size_t)n, cudaMemcpyHostToDevice, st));
  GLUE_TRY(cudaMemcpyAsync(dN, normal, 12 * (size_t)n, cudaMemcpyHostToDevice, st));
  GLUE_TRY(cudaMemcpyAsync(dmin, min_distance, 4 * (size_t)n, cudaMemcpyHostToDevice, st));
  GLUE_TRY(cudaMemcpyAsync(dmax, max_distance, 4 * (size_t)n, cudaMemcpyHostToDevice, st));
  GLUE_TRY(cudaMemsetAsync(dcount, 0, sizeof(int), st));   // the array keeps the previous call's count
  k_is_in_frustum<<<(n + 127) / 128, 128, 0, st>>>(*cam, n, dP, dN, dmin, dmax, viewing_cos_limit, dview, dpx, dpy, dpxr, dlv,
                                                    dvc, ddz, dcount);
  count_launch();
  GLUE_TRY(cudaGetLastError());
  int count = 0;
  GLUE_TRY(cudaMemcpyAsync(in_view, dview, (size_t)n, cudaMemcpyDeviceToHost, st));
  GLUE_TRY(cudaMemcpyAsync(proj_x, dpx, 4 * (size_t)n, cudaMemcpyDeviceToHost, st));
  GLUE_TRY(cudaMemcpyAsync(proj_y, dpy, 4 * (size_t)n, cudaMemcpyDeviceToHost, st));
  GLUE_TRY(cudaMemcpyAsync(proj_xr, dpxr, 4 * (size_t)n, cudaMemcpyDeviceToHost, st));
  GLUE_TRY(cudaMemcpyAsync(scale_level, dlv, 4 * (size_t)n, cudaMemcpyDeviceToHost, st));
  GLUE_TRY(cudaMemcpyAsync(view_cos, dvc, 4 * (size_t)n, cudaMemcpyDeviceToHost, st));
  GLUE_TRY(cudaMemcpyAsync(depth, ddz, 4 * (size_t)n, cudaMemcpyDeviceToHost, st));
  GLUE_TRY(cudaMemcpyAsync(&count, dcount, sizeof(int), cudaMemcpyDeviceToHost, st));
  GLUE_TRY(cudaStreamSynchronize(st));
  return count;
}

}  // extern "C"
