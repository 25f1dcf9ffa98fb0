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

// ---------------------------------------------------------------------------- write-only HBM bandwidth
// The LM-form kernel only WRITES to HBM (258 KB of r and J_r per instance against 0.5 KB read), while the copy figure in
// MEASURED_PEAKS.json is half reads, half writes.  These two streams measure what a store-only stream reaches on the box:
// mode 0: 16-byte stores, fully coalesced, grid-stride (a kernel like the product's own epilogue); mode 1: cudaMemsetAsync.
QSB_GLOBAL qsb_k_store_stream(double2* out, size_t n16, double v) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) out[i] = make_double2(v, v + (double)i);
}

// mode 2: the store pattern of the LM-form epilogue at n = 4 (qsb_jac_param, RESID): a warp owns a 16 x 16 tile of doubles
// (2 KB, rows of 128 B); lane (g, t) holds columns {2t, 2t+1} of two 8-wide column tiles, so one store instruction writes
// eight 64-byte half rows and the other halves follow in the next instruction.  mode 3: the same bytes staged through shared
// memory and written as 512 contiguous bytes per instruction.
QSB_GLOBAL qsb_k_store_tiles(double* out, size_t n_tiles, double v, int staged) {
#ifndef QSB_EMU
  __shared__ __align__(16) double stage[8][256];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
  const size_t wstride = (size_t)gridDim.x * (blockDim.x >> 5);
  for (size_t tile = (size_t)blockIdx.x * (blockDim.x >> 5) + warp; tile < n_tiles; tile += wstride) {
    double* o = out + tile * 256;
    if (!staged) {
#pragma unroll
      for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int nt = 0; nt < 2; ++nt)
          *reinterpret_cast<double2*>(o + (8 * mt + g) * 16 + 8 * nt + 2 * t) = make_double2(v + mt, v + nt);
    } else {
#pragma unroll
      for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int nt = 0; nt < 2; ++nt)
          *reinterpret_cast<double2*>(&stage[warp][(8 * mt + g) * 16 + 8 * nt + 2 * t]) = make_double2(v + mt, v + nt);
      __syncwarp();
#pragma unroll
      for (int i = 0; i < 4; ++i) *reinterpret_cast<double2*>(o + 64 * i + 2 * lane) = *reinterpret_cast<const double2*>(&stage[warp][64 * i + 2 * lane]);
      __syncwarp();
    }
  }
#endif
}

extern "C" int qsb_measure_hbm_write(qsb_ctx* ctx, int slot, int mode, double* out_gbs) {
  if (!ctx || !out_gbs || slot < 0 || slot >= (int)ctx->devs.size()) return qsb_fail(QSB_ERR_INVALID, "bad argument");
  const QsbDevice& d = ctx->devs[slot];
  QSB_CUDA(cudaSetDevice(d.id));
#ifdef QSB_EMU
  const size_t bytes = (size_t)1 << 20;
#else
  const size_t bytes = (size_t)4 << 30;  // 4 GiB >> 126 MB of L2
#endif
  double2* buf = nullptr;
  QSB_CUDA(cudaMalloc((void**)&buf, bytes));
  double best = 0;
  for (int rep = 0; rep < 5; ++rep) {
    QSB_CUDA(cudaEventRecord(d.ev0, d.stream));
    if (mode == 1) QSB_CUDA(cudaMemsetAsync(buf, rep, bytes, d.stream));
    else if (mode >= 2) QSB_LAUNCH(qsb_k_store_tiles, d.sm_count * 8, 256, 0, d.stream, (double*)buf, bytes / 2048, (double)rep, mode == 3 ? 1 : 0);
    else QSB_LAUNCH(qsb_k_store_stream, d.sm_count * 16, 256, 0, d.stream, buf, bytes / 16, (double)rep);
    QSB_CUDA(cudaEventRecord(d.ev1, d.stream));
    QSB_CUDA(cudaEventSynchronize(d.ev1));
    float ms = 0;
    QSB_CUDA(cudaEventElapsedTime(&ms, d.ev0, d.ev1));
    QSB_CUDA(cudaGetLastError());
    if (rep >= 1) best = std::max(best, (double)bytes / (ms * 1e-3) / 1e9);
  }
  cudaFree(buf);
  *out_gbs = best;
  return QSB_OK;
}
