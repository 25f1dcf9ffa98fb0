// FP64 peak micro-benchmarks (own translation unit).
#include "qsb_host.h"
#include "qsb_warp.cuh"
// FP64 peak micro-benchmarks: the roofline denominator for the complex-fp64 kernels is not in MEASURED_PEAKS.json
// (it holds HBM and bf16 only), so bench.py measures it on the box with these two kernels:
//   use_dmma = 0: independent DFMA chains on the FP64 pipe;  use_dmma = 1: mma.sync.m8n8k4.f64 (DMMA) chains.
template <int CHAINS>
QSB_GLOBAL qsb_k_peak_dfma(double* out, int iters, double a, double b) {
  double acc[CHAINS];
#pragma unroll
  for (int i = 0; i < CHAINS; ++i) acc[i] = (double)(threadIdx.x + i);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) acc[i] = fma(acc[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < CHAINS; ++i) s += acc[i];
  out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

#ifndef QSB_EMU
template <int CHAINS>
__global__ void qsb_k_peak_dmma(double* out, int iters, double a, double b) {
  double c0[CHAINS], c1[CHAINS];
#pragma unroll
  for (int i = 0; i < CHAINS; ++i) { c0[i] = threadIdx.x; c1[i] = i; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) qsb_dmma884(c0[i], c1[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < CHAINS; ++i) s += c0[i] + c1[i];
  out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}
#endif

extern "C" int qsb_measure_fp64_peak(qsb_ctx* ctx, int slot, int use_dmma, double* out_tflops) {
  if (!ctx || !out_tflops || slot < 0 || slot >= (int)ctx->devs.size()) return qsb_fail(QSB_ERR_INVALID, "bad argument");
  const QsbDevice& d = ctx->devs[slot];
  QSB_CUDA(cudaSetDevice(d.id));
  const int threads = 256, grid = d.sm_count * 8, iters = 4096;
  constexpr int CH = 8;
  double* buf = nullptr;
  QSB_CUDA(cudaMalloc((void**)&buf, sizeof(double) * (size_t)grid * threads));
  double best = 0;
  for (int rep = 0; rep < 6; ++rep) {
    QSB_CUDA(cudaEventRecord(d.ev0, d.stream));
#ifndef QSB_EMU
    if (use_dmma) qsb_k_peak_dmma<CH><<<grid, threads, 0, d.stream>>>(buf, iters, 1.0000001, 1e-9);
    else
#endif
      QSB_LAUNCH((qsb_k_peak_dfma<CH>), (rep < 0 ? 1 : grid), threads, 0, d.stream, buf, iters, 1.0000001, 1e-9);
    QSB_CUDA(cudaEventRecord(d.ev1, d.stream));
    QSB_CUDA(cudaEventSynchronize(d.ev1));
    float ms = 0;
    QSB_CUDA(cudaEventElapsedTime(&ms, d.ev0, d.ev1));
    QSB_CUDA(cudaGetLastError());
    // DFMA: 2 flop per lane-op.  DMMA m8n8k4: 8*8*4 FMA = 512 flop per warp instruction.
    double flops = use_dmma ? (double)grid * (threads / 32) * iters * CH * 512.0
                            : (double)grid * threads * iters * CH * 2.0;
    if (rep >= 1) best = std::max(best, flops / (ms * 1e-3) / 1e12);
  }
  cudaFree(buf);
  ctx->launches += 6;
  *out_tflops = best;
  return QSB_OK;
}
