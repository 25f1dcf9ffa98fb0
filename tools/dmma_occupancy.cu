// FP64 pipe throughput against the number of resident warps, for the analysis of the hot kernel:
// qsb_k_warp_jac<4> runs 12 one-warp CTAs per SM (3 warps per scheduler), each issuing mma.sync.m8n8k4.f64 (DMMA.8x8x4)
// into 8 accumulators plus DFMA work on the same pipe.  Does that many warps saturate the pipe?
//   nvcc -O2 -gencode arch=compute_100a,code=sm_100a -o /tmp/dmma_occ tools/dmma_occupancy.cu && /tmp/dmma_occ
// Modes: dmma (CH accumulator chains, back to back), dfma (CH chains), mix (per step: 8 DMMA into CH accumulators + 12 DFMA in
// 12 chains, the hot kernel's ratio).  Output: TFLOP/s per (mode, warps per SM, chains).
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

template <int CH>
__global__ void __launch_bounds__(32) k_dmma(double* out, int iters, double a, double b) {
  double c0[CH], c1[CH];
#pragma unroll
  for (int i = 0; i < CH; ++i) { c0[i] = threadIdx.x; c1[i] = i; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 16 / CH; ++r)
#pragma unroll
      for (int i = 0; i < CH; ++i) dmma(c0[i], c1[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < CH; ++i) s += c0[i] + c1[i];
  out[(size_t)blockIdx.x * 32 + threadIdx.x] = s;
}

template <int CH>
__global__ void __launch_bounds__(32) k_dfma(double* out, int iters, double a, double b) {
  double acc[CH];
#pragma unroll
  for (int i = 0; i < CH; ++i) acc[i] = threadIdx.x + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 48 / CH; ++r)
#pragma unroll
      for (int i = 0; i < CH; ++i) acc[i] = fma(acc[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < CH; ++i) s += acc[i];
  out[(size_t)blockIdx.x * 32 + threadIdx.x] = s;
}

// one step = what one half-size derivative product of the hot kernel issues per k-step pair: 24 DFMA (the factors), then
// 16 DMMA into 8 accumulators (two per accumulator, 8 apart) that consume them
template <int CH>
__global__ void __launch_bounds__(32) k_mix(double* out, int iters, double a, double b) {
  double c0[CH], c1[CH], f[12];
#pragma unroll
  for (int i = 0; i < CH; ++i) { c0[i] = threadIdx.x; c1[i] = i; }
#pragma unroll
  for (int i = 0; i < 12; ++i) f[i] = threadIdx.x + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 12; ++i) f[i] = fma(f[i], a, b);
#pragma unroll
    for (int i = 0; i < 12; ++i) f[i] = fma(f[i], a, b);
#pragma unroll
    for (int r = 0; r < 16 / CH; ++r)
#pragma unroll
      for (int i = 0; i < CH; ++i) dmma(c0[i], c1[i], f[(i + r) % 12], f[(i + 3 * r + 5) % 12]);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < CH; ++i) s += c0[i] + c1[i];
#pragma unroll
  for (int i = 0; i < 12; ++i) s += f[i];
  out[(size_t)blockIdx.x * 32 + threadIdx.x] = s;
}

template <typename K>
static double run(K kernel, int grid, int iters, double flop_per_warp_iter, double* buf) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  double best = 0;
  for (int rep = 0; rep < 4; ++rep) {
    cudaEventRecord(e0);
    kernel<<<grid, 32>>>(buf, iters, 1.0000001, 1e-9);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double tf = (double)grid * iters * flop_per_warp_iter / (ms * 1e-3) / 1e12;
    if (rep >= 1 && tf > best) best = tf;
  }
  return best;
}

int main() {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  double* buf = nullptr;
  cudaMalloc(&buf, sizeof(double) * 32 * (size_t)sms * 64);
  const int iters = 20000;
  printf("%-6s %-9s %-7s %s\n", "mode", "warps/SM", "chains", "TFLOP/s");
  const int ws[] = {4, 8, 12, 16, 24, 32, 64};
  for (int w : ws) {
    const int grid = sms * w;
    printf("dmma   %-9d %-7d %.2f\n", w, 8, run(k_dmma<8>, grid, iters, 16 * 512.0, buf));
    printf("dmma   %-9d %-7d %.2f\n", w, 16, run(k_dmma<16>, grid, iters, 16 * 512.0, buf));
    printf("dfma   %-9d %-7d %.2f\n", w, 8, run(k_dfma<8>, grid, iters, 48 * 64.0, buf));
    printf("dfma   %-9d %-7d %.2f\n", w, 16, run(k_dfma<16>, grid, iters, 48 * 64.0, buf));
    printf("mix    %-9d %-7d %.2f\n", w, 8, run(k_mix<8>, grid, iters, 16 * 512.0 + 24 * 64.0, buf));
    printf("mix    %-9d %-7d %.2f\n", w, 16, run(k_mix<16>, grid, iters, 16 * 512.0 + 24 * 64.0, buf));
    fflush(stdout);
  }
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return 1; }
  return 0;
}
