// Dense all-pairs Hamming, the one place on the path where the integer pipe
// does nothing but XOR + POPC:
//   cuorb_knn_match2             cv::BFMatcher(NORM_HAMMING).knnMatch(query, train, matches, 2) as used by
//                                Frame::ComputeStereoFishEyeMatches (src/Frame.cc:1304-1355): per query the
//                                two first entries of the (distance, trainIdx) order (cv::batchDistance's
//                                strict-comparison insertion), optionally followed by the 0.7 ratio test (:1329)
//   cuorb_distinctive_descriptors MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:376-450): per
//                                MapPoint the observation whose median distance to all observations
//                                (itself included, index floor(0.5*(N-1)) of the sorted row) is smallest,
//                                first one wins.
// k_knn2: one thread per query (its 256 bits in eight registers), train descriptors staged through
// shared memory in tiles of 128 and read as broadcast uint4; sets are batched CSR-style so many
// frames fill the GPU.  k_distinctive: one warp per MapPoint, descriptors in shared memory, each
// lane owns rows i = lane, lane+32, ...; the row median comes from a 257-bin counting histogram.
#include "common.cuh"

namespace cuorb {
void set_error(const std::string& msg);

namespace {

constexpr int KNN_THREADS = 128, KNN_TILE = 128;

// POPC issues on the quarter-rate XU pipe (measured: 8 POPC per pair saturate it), so the eight XOR
// words are first compressed with carry-save adders (Harley-Seal: LOP3 sum 0x96 / majority 0xE8 on
// the full-rate ALU pipe) into ones/twos/fours/eights planes: 4 POPC instead of 8.
__device__ __forceinline__ void csa(unsigned a, unsigned b, unsigned c, unsigned& sum, unsigned& carry) {
  sum = a ^ b ^ c;
  carry = (a & b) | (a & c) | (b & c);
}
__device__ __forceinline__ int ham256(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1) {
  const unsigned x0 = a0.x ^ b0.x, x1 = a0.y ^ b0.y, x2 = a0.z ^ b0.z, x3 = a0.w ^ b0.w;
  const unsigned x4 = a1.x ^ b1.x, x5 = a1.y ^ b1.y, x6 = a1.z ^ b1.z, x7 = a1.w ^ b1.w;
  unsigned o1, t1, o2, t2, o3, t3, tS, f1;
  csa(x0, x1, x2, o1, t1);
  csa(o1, x3, x4, o2, t2);
  csa(o2, x5, x6, o3, t3);
  const unsigned ones = o3 ^ x7, t4 = o3 & x7;
  csa(t1, t2, t3, tS, f1);
  const unsigned twos = tS ^ t4, f2 = tS & t4;
  const unsigned fours = f1 ^ f2, eights = f1 & f2;
  return __popc(ones) + 2 * __popc(twos) + 4 * __popc(fours) + 8 * __popc(eights);
}

// grid: (max query blocks per set, n_sets)
__global__ void __launch_bounds__(KNN_THREADS) k_knn2(const uint4* __restrict__ q, const int* __restrict__ qoff,
                                                      const uint4* __restrict__ t, const int* __restrict__ toff,
                                                      int* __restrict__ idx, int* __restrict__ dist, int* __restrict__ match) {
  __shared__ uint4 s_t[KNN_TILE * 2];
  const int set = blockIdx.y;
  const int q0 = qoff[set], nq = qoff[set + 1] - q0, t0 = toff[set], nt = toff[set + 1] - t0;
  if ((int)blockIdx.x * KNN_THREADS >= nq) return;
  const int i = blockIdx.x * KNN_THREADS + threadIdx.x;
  const bool live = i < nq;
  uint4 a0 = make_uint4(0, 0, 0, 0), a1 = a0;
  if (live) { a0 = q[2 * (size_t)(q0 + i)]; a1 = q[2 * (size_t)(q0 + i) + 1]; }
  // key = dist << 20 | train index: the two smallest keys are the two first (distance, index) entries
  unsigned k0 = 0xFFFFFFFFu, k1 = 0xFFFFFFFFu;
  for (int base = 0; base < nt; base += KNN_TILE) {
    const int m = min(KNN_TILE, nt - base);
    __syncthreads();
    for (int w = threadIdx.x; w < 2 * m; w += KNN_THREADS) s_t[w] = t[2 * (size_t)(t0 + base) + w];
    __syncthreads();
    if (live) {
#pragma unroll 4
      for (int j = 0; j < m; j++) {
        const unsigned key = ((unsigned)ham256(a0, a1, s_t[2 * j], s_t[2 * j + 1]) << 20) | (unsigned)(base + j);
        const unsigned lo = min(key, k0);
        k1 = min(k1, max(key, k0));
        k0 = lo;
      }
    }
  }
  if (!live) return;
  const int i0 = k0 == 0xFFFFFFFFu ? -1 : (int)(k0 & 0xFFFFFu), i1 = k1 == 0xFFFFFFFFu ? -1 : (int)(k1 & 0xFFFFFu);
  const int d0 = i0 < 0 ? -1 : (int)(k0 >> 20), d1 = i1 < 0 ? -1 : (int)(k1 >> 20);
  const size_t o = 2 * (size_t)(q0 + i);
  idx[o] = i0; idx[o + 1] = i1; dist[o] = d0; dist[o + 1] = d1;
  // src/Frame.cc:1329: (*it)[0].distance < (*it)[1].distance * 0.7  (float * double literal -> double)
  if (match) match[q0 + i] = (i1 >= 0 && (double)(float)d0 < (double)(float)d1 * 0.7) ? i0 : -1;
}

constexpr int DD_MAXN = 1024;   // observations per MapPoint held in shared memory (32 KB)

// grid: n_points, block 32
__global__ void __launch_bounds__(32) k_distinctive(const uint4* __restrict__ desc, const int* __restrict__ off,
                                                    int* __restrict__ best) {
  extern __shared__ __align__(16) unsigned char smem[];
  uint4* s_d = reinterpret_cast<uint4*>(smem);                                   // n descriptors
  const int p = blockIdx.x, lane = threadIdx.x;
  const int o = off[p], n = off[p + 1] - o;
  if (n <= 0) { if (lane == 0) best[p] = -1; return; }
  unsigned short* hist = reinterpret_cast<unsigned short*>(smem + 32 * (size_t)n) + lane * 258;  // per-lane 257 bins
  for (int w = lane; w < 2 * n; w += 32) s_d[w] = desc[2 * (size_t)o + w];
  __syncwarp();
  const int rank = (int)(0.5 * (n - 1));     // vDists[0.5 * (N - 1)]  (:432)
  unsigned bestkey = 0xFFFFFFFFu;            // median << 16 | row: smallest median, first row wins (:434)
  for (int i = lane; i < n; i += 32) {
    for (int b = 0; b < 257; b++) hist[b] = 0;
    const uint4 a0 = s_d[2 * i], a1 = s_d[2 * i + 1];
    for (int j = 0; j < n; j++) hist[ham256(a0, a1, s_d[2 * j], s_d[2 * j + 1])]++;   // j == i contributes distance 0 (:417)
    int acc = 0, med = 0;
    for (int b = 0; b < 257; b++) { acc += hist[b]; if (acc > rank) { med = b; break; } }
    bestkey = min(bestkey, ((unsigned)med << 16) | (unsigned)i);
  }
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) bestkey = min(bestkey, __shfl_xor_sync(0xffffffffu, bestkey, s));
  if (lane == 0) best[p] = (int)(bestkey & 0xFFFFu);
}

struct Scratch {  // grow-only device scratch per host thread
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
thread_local Scratch t_bf;

}  // namespace
}  // namespace cuorb

using namespace cuorb;

#define BF_TRY(call)                                                                                   \
  do {                                                                                                 \
    cudaError_t err__ = (call);                                                                        \
    if (err__ != cudaSuccess) { set_error(std::string(#call) + ": " + cudaGetErrorString(err__)); return CUORB_E_CUDA; } \
  } while (0)

extern "C" {

int cuorb_knn_match2(int device, int n_sets, const int32_t* query_off, const uint8_t* query_desc,
                     const int32_t* train_off, const uint8_t* train_desc, int32_t* idx_out, int32_t* dist_out,
                     int32_t* ratio_match_out) {
  if (n_sets <= 0 || !query_off || !train_off || !query_desc || !train_desc || !idx_out || !dist_out) {
    set_error("cuorb_knn_match2: bad argument"); return CUORB_E_BADARG;
  }
  const int nq = query_off[n_sets], nt = train_off[n_sets];
  int maxq = 0;
  for (int s = 0; s < n_sets; s++) {
    const int a = query_off[s + 1] - query_off[s], b = train_off[s + 1] - train_off[s];
    if (a < 0 || b < 0 || b >= (1 << 20)) { set_error("cuorb_knn_match2: offsets not ascending or a train set of 2^20 or more"); return CUORB_E_BADARG; }
    maxq = std::max(maxq, a);
  }
  if (nq == 0) return 0;
  BF_TRY(cudaSetDevice(device));
  const size_t bytes = 32 * ((size_t)nq + nt) + 4 * (size_t)(2 * (n_sets + 1) + 5 * (size_t)nq) + 4096;
  if (t_bf.reserve(device, bytes) != 0) { set_error("cudaMalloc failed (knn workspace)"); return CUORB_E_CUDA; }
  cudaStream_t st = cudaStreamPerThread;
  uint4* dq = t_bf.take<uint4>(2 * (size_t)nq);
  uint4* dt = t_bf.take<uint4>(2 * (size_t)std::max(nt, 1));
  int* dqo = t_bf.take<int>(n_sets + 1);
  int* dto = t_bf.take<int>(n_sets + 1);
  int* didx = t_bf.take<int>(2 * (size_t)nq);
  int* ddist = t_bf.take<int>(2 * (size_t)nq);
  int* dmatch = ratio_match_out ? t_bf.take<int>(nq) : nullptr;
  BF_TRY(cudaMemcpyAsync(dq, query_desc, 32 * (size_t)nq, cudaMemcpyHostToDevice, st));
  if (nt > 0) BF_TRY(cudaMemcpyAsync(dt, train_desc, 32 * (size_t)nt, cudaMemcpyHostToDevice, st));
  BF_TRY(cudaMemcpyAsync(dqo, query_off, 4 * (size_t)(n_sets + 1), cudaMemcpyHostToDevice, st));
  BF_TRY(cudaMemcpyAsync(dto, train_off, 4 * (size_t)(n_sets + 1), cudaMemcpyHostToDevice, st));
  k_knn2<<<dim3((maxq + KNN_THREADS - 1) / KNN_THREADS, n_sets), KNN_THREADS, 0, st>>>(dq, dqo, dt, dto, didx, ddist, dmatch);
  count_launch();
  BF_TRY(cudaGetLastError());
  BF_TRY(cudaMemcpyAsync(idx_out, didx, 8 * (size_t)nq, cudaMemcpyDeviceToHost, st));
  BF_TRY(cudaMemcpyAsync(dist_out, ddist, 8 * (size_t)nq, cudaMemcpyDeviceToHost, st));
  if (dmatch) BF_TRY(cudaMemcpyAsync(ratio_match_out, dmatch, 4 * (size_t)nq, cudaMemcpyDeviceToHost, st));
  BF_TRY(cudaStreamSynchronize(st));
  return nq;
}

int cuorb_distinctive_descriptors(int device, int n_points, const int32_t* off, const uint8_t* desc, int32_t* best_out) {
  if (n_points <= 0 || !off || !desc || !best_out) { set_error("cuorb_distinctive_descriptors: bad argument"); return CUORB_E_BADARG; }
  int maxn = 0;
  for (int p = 0; p < n_points; p++) {
    const int n = off[p + 1] - off[p];
    if (n < 0) { set_error("cuorb_distinctive_descriptors: offsets not ascending"); return CUORB_E_BADARG; }
    maxn = std::max(maxn, n);
  }
  if (maxn > DD_MAXN) { set_error("cuorb_distinctive_descriptors: more than 1024 observations of one MapPoint"); return CUORB_E_UNSUPPORTED; }
  const int total = off[n_points];
  BF_TRY(cudaSetDevice(device));
  if (t_bf.reserve(device, 32 * (size_t)std::max(total, 1) + 4 * (size_t)(2 * n_points + 1) + 2048) != 0) {
    set_error("cudaMalloc failed (distinctive-descriptor workspace)"); return CUORB_E_CUDA;
  }
  cudaStream_t st = cudaStreamPerThread;
  uint4* dd = t_bf.take<uint4>(2 * (size_t)std::max(total, 1));
  int* doff = t_bf.take<int>(n_points + 1);
  int* dbest = t_bf.take<int>(n_points);
  if (total > 0) BF_TRY(cudaMemcpyAsync(dd, desc, 32 * (size_t)total, cudaMemcpyHostToDevice, st));
  BF_TRY(cudaMemcpyAsync(doff, off, 4 * (size_t)(n_points + 1), cudaMemcpyHostToDevice, st));
  const int smem = 32 * std::max(maxn, 1) + 32 * 258 * 2;
  if (smem > 48 * 1024) cudaFuncSetAttribute(k_distinctive, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);   // per device
  k_distinctive<<<n_points, 32, smem, st>>>(dd, doff, dbest);
  count_launch();
  BF_TRY(cudaGetLastError());
  BF_TRY(cudaMemcpyAsync(best_out, dbest, 4 * (size_t)n_points, cudaMemcpyDeviceToHost, st));
  BF_TRY(cudaStreamSynchronize(st));
  return n_points;
}

}  // extern "C"
