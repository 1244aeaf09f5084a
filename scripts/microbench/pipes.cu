// Microbenchmark: which issue pipe do VIMNMX3.U16x2 / HMNMX2 / PRMT / IMAD share on sm_100a?
// Each kernel runs N iterations of 8 independent chains; reports warp-instructions / clk / SM.
#include <cstdio>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#define ITER 4096
template <int MODE>
__global__ void k(unsigned* out, unsigned seed) {
  unsigned a[8], b[8];
#pragma unroll
  for (int i = 0; i < 8; i++) { a[i] = seed * (threadIdx.x + i + 1); b[i] = a[i] ^ 0x12345678u; }
  for (int it = 0; it < ITER; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      if (MODE == 0) a[i] = __vimin3_u16x2(a[i], b[i], a[(i + 1) & 7]);                  // VIMNMX3
      if (MODE == 1) { __half2 h = __hmin2(*(__half2*)&a[i], *(__half2*)&b[i]); a[i] = *(unsigned*)&h ^ b[(i + 1) & 7]; }  // HMNMX2 + LOP3
      if (MODE == 2) { __half2 h = __hmin2(*(__half2*)&a[i], *(__half2*)&b[i]); a[i] = *(unsigned*)&h; b[i] = __vimax3_u16x2(b[i], a[(i + 3) & 7], b[(i + 1) & 7]); }  // both
      if (MODE == 3) a[i] = __vminu2(a[i], b[i]) + 1;                                      // VIMNMX 2-input + add
      if (MODE == 4) a[i] = a[i] * b[i] + a[(i + 1) & 7];                                  // IMAD
      if (MODE == 5) { a[i] = a[i] * b[i] + a[(i + 1) & 7]; b[i] = __vimax3_u16x2(b[i], a[(i + 3) & 7], b[(i + 1) & 7]); }  // IMAD + VIMNMX3
      if (MODE == 6) { __half2 h = __hmin2(*(__half2*)&a[i], *(__half2*)&b[i]); a[i] = *(unsigned*)&h; b[i] = b[i] * a[(i + 3) & 7] + b[(i + 1) & 7]; }  // HMNMX2 + IMAD
      if (MODE == 7) a[i] = __byte_perm(a[i], b[i], 0x5432) ;                              // PRMT
      if (MODE == 8) { __half2 h = __hmin2(*(__half2*)&a[i], *(__half2*)&b[i]); __half2 g = __hmax2(h, *(__half2*)&a[(i+1)&7]); a[i] = *(unsigned*)&g; }  // HMNMX2 x2
    }
  }
  unsigned s = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) s += a[i] + b[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE>
void run(const char* name, int ops_per_iter) {
  unsigned* d; cudaMalloc(&d, 148 * 8 * 512 * 4);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MODE><<<148 * 4, 512>>>(d, 3); cudaDeviceSynchronize();
  cudaEventRecord(e0); k<MODE><<<148 * 4, 512>>>(d, 3); cudaEventRecord(e1); cudaDeviceSynchronize();
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  double winst = 148.0 * 4 * 16 * ITER * 8.0 * ops_per_iter;
  printf("%-28s %8.3f ms  %6.2f warp-inst/clk/SM (at %d MHz nominal)\n", name, ms, winst / (ms * 1e-3) / (clk * 1e3) / 148, clk / 1000);
  cudaFree(d);
}
int main() {
  run<0>("VIMNMX3.U16x2", 1); run<1>("HMNMX2+LOP3", 2); run<2>("HMNMX2+VIMNMX3", 2); run<3>("VIMNMX+IADD", 2);
  run<4>("IMAD", 1); run<5>("IMAD+VIMNMX3", 2); run<6>("HMNMX2+IMAD", 2); run<7>("PRMT", 1); run<8>("HMNMX2 x2", 2);
  return 0;
}
