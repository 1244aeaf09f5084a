// Vocabulary-node searches of ORBmatcher:
//   int SearchByBoW(KeyFrame* pKF, Frame& F, std::vector<MapPoint*>&)      src/ORBmatcher.cc:237-420   (mode 0)
//   int SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>&) src/ORBmatcher.cc:784-916 (mode 1)
// Features of the two sides are only compared inside a common vocabulary node, and a side-2 feature
// belongs to exactly one node, so the reference's greedy "first come, first served" assignment
// (vpMapPointMatches[realIdxF] / vbMatched2) only couples features of the SAME node: nodes run in
// parallel, one warp each; inside a node the side-1 features are taken in list order and the lanes
// scan the side-2 list for the best (distance, position) and the second-best distance (the
// reference's strict-< updates are order-independent once ties go to the earliest position).
// The rotation-consistency cut (HISTO_LENGTH = 30 bins, ComputeThreeMaxima :2038-2071) runs on the
// host over the few hundred matches.
#include <cmath>

#include "common.cuh"

namespace cuorb {
void set_error(const std::string& msg);

namespace {

__device__ __forceinline__ int ham256w(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1) {
  return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
         __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

// grid: ceil(n_nodes / 4), block 128 (one warp per node)
__global__ void __launch_bounds__(128) k_bow(const uint4* __restrict__ d1, const uint8_t* __restrict__ valid1,
                                             const uint4* __restrict__ d2, const uint8_t* __restrict__ elig2,
                                             uint8_t* matched2, int n_nodes, const int* __restrict__ off1,
                                             const int* __restrict__ idx1, const int* __restrict__ off2,
                                             const int* __restrict__ idx2, int th_low, float nnratio,
                                             int* __restrict__ match12) {
  const int node = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (node >= n_nodes) return;
  const int a_beg = off1[node], a_end = off1[node + 1], b_beg = off2[node], b_end = off2[node + 1];
  for (int a = a_beg; a < a_end; a++) {
    const int i1 = idx1[a];
    if (!valid1[i1]) continue;                                   // no (good) MapPoint (:263-271, :819-825)
    const uint4 A0 = d1[2 * (size_t)i1], A1 = d1[2 * (size_t)i1 + 1];
    unsigned key1 = (256u << 16) | 0xFFFFu;                      // dist << 16 | position in the node's side-2 list
    unsigned sec = 256u;
    for (int b = b_beg + lane; b < b_end; b += 32) {
      const int i2 = idx2[b];
      if (matched2[i2] || !elig2[i2]) continue;                  // (:290-292, :841-847)
      const unsigned dist = (unsigned)ham256w(A0, A1, d2[2 * (size_t)i2], d2[2 * (size_t)i2 + 1]);
      const unsigned key = (dist << 16) | (unsigned)(b - b_beg);
      if (key < key1) { sec = key1 >> 16; key1 = key; }
      else sec = min(sec, dist);
    }
    unsigned kmin = key1;
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) kmin = min(kmin, __shfl_xor_sync(0xffffffffu, kmin, s));
    unsigned second = key1 == kmin ? sec : (key1 >> 16);         // the winner lane offers its runner-up
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) second = min(second, __shfl_xor_sync(0xffffffffu, second, s));
    const int best1 = (int)(kmin >> 16);
    if (best1 <= th_low && (kmin & 0xFFFFu) != 0xFFFFu &&
        static_cast<float>(best1) < __fmul_rn(nnratio, static_cast<float>((int)second))) {
      const int i2 = idx2[b_beg + (int)(kmin & 0xFFFFu)];
      if (lane == 0) { match12[i1] = i2; matched2[i2] = 1; }
    }
    __syncwarp();                                                // the claim is visible to the next scan
  }
}

// The (KeyFrame, Frame) search when F is a two-camera rig Frame (F.Nleft != -1; src/ORBmatcher.cc:306-331, :371-396):
// F's features [0, nleft2) belong to the left camera, the rest to the right one, all in the same node lists.  The lanes
// keep a best / runner-up pair per camera; the left match needs the ratio test, the right match sits inside the left
// one's `bestDist1 <= TH_LOW` block with its ratio test disabled (`|| true`).  grid: ceil(n_nodes / 4), block 128.
__global__ void __launch_bounds__(128) k_bow_rig(const uint4* __restrict__ d1, const uint8_t* __restrict__ valid1,
                                                 const uint4* __restrict__ d2, int nleft2, uint8_t* matched2, int n_nodes,
                                                 const int* __restrict__ off1, const int* __restrict__ idx1,
                                                 const int* __restrict__ off2, const int* __restrict__ idx2, int th_low,
                                                 float nnratio, int* __restrict__ matchL, int* __restrict__ matchR) {
  const int node = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (node >= n_nodes) return;
  const int a_beg = off1[node], a_end = off1[node + 1], b_beg = off2[node], b_end = off2[node + 1];
  for (int a = a_beg; a < a_end; a++) {
    const int i1 = idx1[a];
    if (!valid1[i1]) continue;
    const uint4 A0 = d1[2 * (size_t)i1], A1 = d1[2 * (size_t)i1 + 1];
    unsigned keyL = (256u << 16) | 0xFFFFu, secL = 256u, keyR = (256u << 16) | 0xFFFFu;
    for (int b = b_beg + lane; b < b_end; b += 32) {
      const int i2 = idx2[b];
      if (matched2[i2]) continue;
      const unsigned dist = (unsigned)ham256w(A0, A1, d2[2 * (size_t)i2], d2[2 * (size_t)i2 + 1]);
      const unsigned key = (dist << 16) | (unsigned)(b - b_beg);
      if (i2 < nleft2) {
        if (key < keyL) { secL = keyL >> 16; keyL = key; }
        else secL = min(secL, dist);
      } else if (key < keyR) keyR = key;
    }
    unsigned kminL = keyL, kminR = keyR;
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
      kminL = min(kminL, __shfl_xor_sync(0xffffffffu, kminL, s));
      kminR = min(kminR, __shfl_xor_sync(0xffffffffu, kminR, s));
    }
    unsigned second = keyL == kminL ? secL : (keyL >> 16);
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) second = min(second, __shfl_xor_sync(0xffffffffu, second, s));
    const int best1 = (int)(kminL >> 16), best1R = (int)(kminR >> 16);
    if (best1 <= th_low) {                                        // (:333) - also gates the right camera's match
      if ((kminL & 0xFFFFu) != 0xFFFFu && static_cast<float>(best1) < __fmul_rn(nnratio, static_cast<float>((int)second))) {
        const int i2 = idx2[b_beg + (int)(kminL & 0xFFFFu)];
        if (lane == 0) { matchL[i1] = i2; matched2[i2] = 1; }
      }
      if (best1R <= th_low && (kminR & 0xFFFFu) != 0xFFFFu) {     // (:371-373)
        const int i2 = idx2[b_beg + (int)(kminR & 0xFFFFu)];
        if (lane == 0) { matchR[i1] = i2; matched2[i2] = 1; }
      }
    }
    __syncwarp();
  }
}

struct Scratch {
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
thread_local Scratch t_bow;

}  // namespace
}  // namespace cuorb

using namespace cuorb;

#define BOW_TRY(call)                                                                                  \
  do {                                                                                                 \
    cudaError_t err__ = (call);                                                                        \
    if (err__ != cudaSuccess) { set_error(std::string(#call) + ": " + cudaGetErrorString(err__)); return CUORB_E_CUDA; } \
  } while (0)

extern "C" int cuorb_search_by_bow(int device, int n1, const uint8_t* desc1, const uint8_t* valid1, const float* angle1,
                                   int n2, const uint8_t* desc2, const uint8_t* eligible2, const float* angle2,
                                   int n_nodes, const int32_t* node_off1, const int32_t* node_idx1,
                                   const int32_t* node_off2, const int32_t* node_idx2, int mode, float nnratio,
                                   int check_orientation, int32_t* match12_out) {
  if (n1 < 0 || n2 < 0 || n_nodes < 0 || !match12_out || (n1 > 0 && (!desc1 || !valid1)) || (n2 > 0 && (!desc2 || !eligible2)) ||
      (n_nodes > 0 && (!node_off1 || !node_off2)) || (check_orientation && (!angle1 || !angle2)) || mode < 0 || mode > 1) {
    set_error("cuorb_search_by_bow: bad argument"); return CUORB_E_BADARG;
  }
  for (int i = 0; i < n1; i++) match12_out[i] = -1;
  if (n1 == 0 || n2 == 0 || n_nodes == 0) return 0;
  const int m1 = node_off1[n_nodes], m2 = node_off2[n_nodes];
  if (m1 < 0 || m2 < 0 || (m1 > 0 && !node_idx1) || (m2 > 0 && !node_idx2)) { set_error("cuorb_search_by_bow: bad node lists"); return CUORB_E_BADARG; }
  for (int nd = 0; nd < n_nodes; nd++)
    if (node_off2[nd + 1] - node_off2[nd] >= 0xFFFF || node_off1[nd + 1] < node_off1[nd] || node_off2[nd + 1] < node_off2[nd]) {
      set_error("cuorb_search_by_bow: node list not ascending or longer than 65534"); return CUORB_E_BADARG;
    }
  for (int k = 0; k < m1; k++) if ((unsigned)node_idx1[k] >= (unsigned)n1) { set_error("cuorb_search_by_bow: side-1 index out of range"); return CUORB_E_BADARG; }
  for (int k = 0; k < m2; k++) if ((unsigned)node_idx2[k] >= (unsigned)n2) { set_error("cuorb_search_by_bow: side-2 index out of range"); return CUORB_E_BADARG; }
  if (m1 == 0 || m2 == 0) return 0;
  BOW_TRY(cudaSetDevice(device));
  const size_t bytes = 33 * ((size_t)n1 + n2) + n2 + 4 * ((size_t)2 * (n_nodes + 1) + m1 + m2 + n1) + 4096;
  if (t_bow.reserve(device, bytes) != 0) { set_error("cudaMalloc failed (BoW search workspace)"); return CUORB_E_CUDA; }
  cudaStream_t st = cudaStreamPerThread;
  uint4* dd1 = t_bow.take<uint4>(2 * (size_t)n1);
  uint4* dd2 = t_bow.take<uint4>(2 * (size_t)n2);
  uint8_t* dv1 = t_bow.take<uint8_t>(n1);
  uint8_t* de2 = t_bow.take<uint8_t>(n2);
  uint8_t* dm2 = t_bow.take<uint8_t>(n2);
  int* do1 = t_bow.take<int>(n_nodes + 1);
  int* do2 = t_bow.take<int>(n_nodes + 1);
  int* di1 = t_bow.take<int>(m1);
  int* di2 = t_bow.take<int>(m2);
  int* dm = t_bow.take<int>(n1);
  BOW_TRY(cudaMemcpyAsync(dd1, desc1, 32 * (size_t)n1, cudaMemcpyHostToDevice, st));
  BOW_TRY(cudaMemcpyAsync(dd2, desc2, 32 * (size_t)n2, cudaMemcpyHostToDevice, st));
  BOW_TRY(cudaMemcpyAsync(dv1, valid1, n1, cudaMemcpyHostToDevice, st));
  BOW_TRY(cudaMemcpyAsync(de2, eligible2, n2, cudaMemcpyHostToDevice, st));
  BOW_TRY(cudaMemsetAsync(dm2, 0, n2, st));
  BOW_TRY(cudaMemsetAsync(dm, 0xFF, 4 * (size_t)n1, st));
  BOW_TRY(cudaMemcpyAsync(do1, node_off1, 4 * (size_t)(n_nodes + 1), cudaMemcpyHostToDevice, st));
  BOW_TRY(cudaMemcpyAsync(do2, node_off2, 4 * (size_t)(n_nodes + 1), cudaMemcpyHostToDevice, st));
  BOW_TRY(cudaMemcpyAsync(di1, node_idx1, 4 * (size_t)m1, cudaMemcpyHostToDevice, st));
  BOW_TRY(cudaMemcpyAsync(di2, node_idx2, 4 * (size_t)m2, cudaMemcpyHostToDevice, st));
  const int TH_LOW = 50;
  k_bow<<<(n_nodes + 3) / 4, 128, 0, st>>>(dd1, dv1, dd2, de2, dm2, n_nodes, do1, di1, do2, di2,
                                            mode == 0 ? TH_LOW : TH_LOW - 1, nnratio, dm);
  count_launch();
  BOW_TRY(cudaGetLastError());
  BOW_TRY(cudaMemcpyAsync(match12_out, dm, 4 * (size_t)n1, cudaMemcpyDeviceToHost, st));
  BOW_TRY(cudaStreamSynchronize(st));
  // rotation consistency (:311-323, :399-417 / :865-876, :896-912)
  const int HISTO_LENGTH = 30;
  const float factor = 1.0f / HISTO_LENGTH;
  int hist[HISTO_LENGTH] = {0};
  std::vector<int> bin(check_orientation ? n1 : 0, -1);
  int nmatches = 0;
  for (int i = 0; i < n1; i++) {
    if (match12_out[i] < 0) continue;
    nmatches++;
    if (check_orientation) {
      float rot = angle1[i] - angle2[match12_out[i]];
      if (rot < 0.0) rot += 360.0f;
      int b = (int)std::round(rot * factor);
      if (b == HISTO_LENGTH) b = 0;
      if (b < 0 || b >= HISTO_LENGTH) { set_error("keypoint angle outside [0,360)"); return CUORB_E_BADARG; }
      bin[i] = b; hist[b]++;
    }
  }
  if (check_orientation) {
    int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
    for (int i = 0; i < HISTO_LENGTH; i++) {
      const int s = hist[i];
      if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
      else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
      else if (s > max3) { max3 = s; ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
    for (int i = 0; i < n1; i++)
      if (bin[i] >= 0 && bin[i] != ind1 && bin[i] != ind2 && bin[i] != ind3) { match12_out[i] = -1; nmatches--; }
  }
  return nmatches;
}

extern "C" int cuorb_search_by_bow_rig(int device, int n1, const uint8_t* desc1, const uint8_t* valid1, const float* angle1,
                                       int n2, int nleft2, const uint8_t* desc2, const float* angle2, int n_nodes,
                                       const int32_t* node_off1, const int32_t* node_idx1, const int32_t* node_off2,
                                       const int32_t* node_idx2, float nnratio, int check_orientation,
                                       int32_t* match_left_out, int32_t* match_right_out) {
  if (n1 < 0 || n2 < 0 || nleft2 < 0 || nleft2 > n2 || n_nodes < 0 || !match_left_out || !match_right_out ||
      (n1 > 0 && (!desc1 || !valid1)) || (n2 > 0 && !desc2) || (n_nodes > 0 && (!node_off1 || !node_off2)) ||
      (check_orientation && (!angle1 || !angle2))) {
    set_error("cuorb_search_by_bow_rig: bad argument"); return CUORB_E_BADARG;
  }
  for (int i = 0; i < n1; i++) match_left_out[i] = match_right_out[i] = -1;
  if (n1 == 0 || n2 == 0 || n_nodes == 0) return 0;
  const int m1 = node_off1[n_nodes], m2 = node_off2[n_nodes];
  if (m1 < 0 || m2 < 0 || (m1 > 0 && !node_idx1) || (m2 > 0 && !node_idx2)) { set_error("cuorb_search_by_bow_rig: bad node lists"); return CUORB_E_BADARG; }
  for (int nd = 0; nd < n_nodes; nd++)
    if (node_off2[nd + 1] - node_off2[nd] >= 0xFFFF || node_off1[nd + 1] < node_off1[nd] || node_off2[nd + 1] < node_off2[nd]) {
      set_error("cuorb_search_by_bow_rig: node list not ascending or longer than 65534"); return CUORB_E_BADARG;
    }
  for (int k = 0; k < m1; k++) if ((unsigned)node_idx1[k] >= (unsigned)n1) { set_error("cuorb_search_by_bow_rig: side-1 index out of range"); return CUORB_E_BADARG; }
  for (int k = 0; k < m2; k++) if ((unsigned)node_idx2[k] >= (unsigned)n2) { set_error("cuorb_search_by_bow_rig: side-2 index out of range"); return CUORB_E_BADARG; }
  if (m1 == 0 || m2 == 0) return 0;
  BOW_TRY(cudaSetDevice(device));
  const size_t bytes = 33 * ((size_t)n1 + n2) + n2 + 4 * ((size_t)2 * (n_nodes + 1) + m1 + m2 + 2 * (size_t)n1) + 8192;
  if (t_bow.reserve(device, bytes) != 0) { set_error("cudaMalloc failed (BoW search workspace)"); return CUORB_E_CUDA; }
  cudaStream_t st = cudaStreamPerThread;
  uint4* dd1 = t_bow.take<uint4>(2 * (size_t)n1);
  uint4* dd2 = t_bow.take<uint4>(2 * (size_t)n2);
  uint8_t* dv1 = t_bow.take<uint8_t>(n1);
  uint8_t* dm2 = t_bow.take<uint8_t>(n2);
  int* do1 = t_bow.take<int>(n_nodes + 1);
  int* do2 = t_bow.take<int>(n_nodes + 1);
  int* di1 = t_bow.take<int>(m1);
  int* di2 = t_bow.take<int>(m2);
  int* dml = t_bow.take<int>(n1);
  int* dmr = t_bow.take<int>(n1);
  BOW_TRY(cudaMemcpyAsync(dd1, desc1, 32 * (size_t)n1, cudaMemcpyHostToDevice, st));
  BOW_TRY(cudaMemcpyAsync(dd2, desc2, 32 * (size_t)n2, cudaMemcpyHostToDevice, st));
  BOW_TRY(cudaMemcpyAsync(dv1, valid1, n1, cudaMemcpyHostToDevice, st));
  BOW_TRY(cudaMemsetAsync(dm2, 0, n2, st));
  BOW_TRY(cudaMemsetAsync(dml, 0xFF, 4 * (size_t)n1, st));
  BOW_TRY(cudaMemsetAsync(dmr, 0xFF, 4 * (size_t)n1, st));
  BOW_TRY(cudaMemcpyAsync(do1, node_off1, 4 * (size_t)(n_nodes + 1), cudaMemcpyHostToDevice, st));
  BOW_TRY(cudaMemcpyAsync(do2, node_off2, 4 * (size_t)(n_nodes + 1), cudaMemcpyHostToDevice, st));
  BOW_TRY(cudaMemcpyAsync(di1, node_idx1, 4 * (size_t)m1, cudaMemcpyHostToDevice, st));
  BOW_TRY(cudaMemcpyAsync(di2, node_idx2, 4 * (size_t)m2, cudaMemcpyHostToDevice, st));
  k_bow_rig<<<(n_nodes + 3) / 4, 128, 0, st>>>(dd1, dv1, dd2, nleft2, dm2, n_nodes, do1, di1, do2, di2, 50, nnratio, dml, dmr);
  count_launch();
  BOW_TRY(cudaGetLastError());
  BOW_TRY(cudaMemcpyAsync(match_left_out, dml, 4 * (size_t)n1, cudaMemcpyDeviceToHost, st));
  BOW_TRY(cudaMemcpyAsync(match_right_out, dmr, 4 * (size_t)n1, cudaMemcpyDeviceToHost, st));
  BOW_TRY(cudaStreamSynchronize(st));
  // rotation consistency over the matches of both cameras (:344-361, :381-393, :399-417)
  const int HISTO_LENGTH = 30;
  const float factor = 1.0f / HISTO_LENGTH;
  int hist[HISTO_LENGTH] = {0};
  std::vector<int> binL(check_orientation ? n1 : 0, -1), binR(check_orientation ? n1 : 0, -1);
  int nmatches = 0;
  for (int i = 0; i < n1; i++)
    for (int side = 0; side < 2; side++) {
      const int m = side ? match_right_out[i] : match_left_out[i];
      if (m < 0) continue;
      nmatches++;
      if (!check_orientation) continue;
      float rot = angle1[i] - angle2[m];
      if (rot < 0.0) rot += 360.0f;
      int b = (int)std::round(rot * factor);
      if (b == HISTO_LENGTH) b = 0;
      if (b < 0 || b >= HISTO_LENGTH) { set_error("keypoint angle outside [0,360)"); return CUORB_E_BADARG; }
      (side ? binR : binL)[i] = b; hist[b]++;
    }
  if (check_orientation) {
    int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
    for (int i = 0; i < HISTO_LENGTH; i++) {
      const int s = hist[i];
      if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
      else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
      else if (s > max3) { max3 = s; ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
    for (int i = 0; i < n1; i++) {
      if (binL[i] >= 0 && binL[i] != ind1 && binL[i] != ind2 && binL[i] != ind3) { match_left_out[i] = -1; nmatches--; }
      if (binR[i] >= 0 && binR[i] != ind1 && binR[i] != ind2 && binR[i] != ind3) { match_right_out[i] = -1; nmatches--; }
    }
  }
  return nmatches;
}
