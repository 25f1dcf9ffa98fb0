// Cycles of the diagonal-block factorisation and the panel substitution of the LM kernel (qsb_ch4_diag, qsb_ch4_panel), one CTA:
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -Iqsearch_b200/csrc -o /tmp/d tools/diag_bench.cu && /tmp/d
#include <cstdio>
#include "qsb_lm2.cuh"
#include "qsb_lm3.cuh"
__global__ void k(long long* out, int reps, int mode) {
  __shared__ double Ls[2100];
  __shared__ double Dk[2 * QSB_CH4_DSZ];
  __shared__ int ok;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  long long best = 1ll << 60, best2 = 1ll << 60;
  for (int r = 0; r < reps; ++r) {
    for (int e = threadIdx.x; e < 2016; e += blockDim.x) Ls[e] = 0.01 * ((e * 7) % 13);
    __syncthreads();
    for (int i = threadIdx.x; i < 63; i += blockDim.x) Ls[QSB_TRI32(i, i)] = 50.0 + i;
    __syncthreads();
    long long t0 = clock64();
    if (warp == 0) qsb_ch4_diag(Ls, 0, 16, Dk, &ok, lane);
    long long t1 = clock64();
    __syncthreads();
    long long t2 = clock64();
    if (t1 - t0 < best) best = t1 - t0;
    if (t2 - t0 < best2) best2 = t2 - t0;
    // panel
    long long t3 = clock64();
    qsb_ch4_panel(Ls, 16, 0, 16, 47, Dk);
    __syncthreads();
    long long t4 = clock64();
    if (threadIdx.x == 0 && r == reps - 1) { out[0] = best; out[1] = best2; out[2] = t4 - t3; }
  }
  if (threadIdx.x == 0) { out[0] = best; out[1] = best2; }
}
int main() {
  long long* d; cudaMalloc(&d, 64);
  k<<<1, 256>>>(d, 20, 0);
  long long h[4]; cudaMemcpy(h, d, 32, cudaMemcpyDeviceToHost);
  printf("diag (warp 0 alone): %lld cycles; with barrier: %lld; panel+barrier: %lld; err=%s\n", h[0], h[1], h[2], cudaGetErrorString(cudaGetLastError()));
  return 0;
}
