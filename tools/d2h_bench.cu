// Plain pinned device-to-host copy bandwidth of ONE process / GPU, for the multi-rank e2e analysis:
//   nvcc -O2 -o /tmp/d2h_bench tools/d2h_bench.cu && /tmp/d2h_bench <device> [MiB per copy] [copies]
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
int main(int argc, char** argv) {
  const int dev = argc > 1 ? atoi(argv[1]) : 0;
  const size_t mib = argc > 2 ? atoi(argv[2]) : 512;
  const int copies = argc > 3 ? atoi(argv[3]) : 8;
  const size_t bytes = mib << 20;
  void *d = nullptr, *h = nullptr;
  cudaSetDevice(dev);
  if (cudaMalloc(&d, bytes) != cudaSuccess || cudaHostAlloc(&h, bytes, cudaHostAllocPortable) != cudaSuccess) { printf("alloc failed\n"); return 1; }
  cudaMemset(d, 1, bytes);
  cudaStream_t st; cudaStreamCreate(&st);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, st); cudaStreamSynchronize(st);  // first touch
  cudaEventRecord(e0, st);
  for (int i = 0; i < copies; ++i) cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, st);
  cudaEventRecord(e1, st); cudaEventSynchronize(e1);
  float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
  printf("device %d: %d x %zu MiB D2H pinned: %.1f GB/s\n", dev, copies, mib, copies * (double)bytes / (ms * 1e-3) / 1e9);
  return 0;
}
