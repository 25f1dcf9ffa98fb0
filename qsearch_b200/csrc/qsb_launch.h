// Launch helpers shared by every translation unit of the library (kernel attribute set-up, occupancy query).
// The launch geometry of the evaluation kernels themselves lives in qsb_launch_eval.h, which only qsb_api.cu includes: a
// non-template function that names a kernel template instantiates it, so keeping those functions out of this header keeps
// the evaluation kernels from being compiled once per translation unit.
#pragma once
#include <algorithm>

#include "qsb_host.h"

template <class K>
static inline int qsb_prepare_kernel(K kernel, size_t smem) {
  if (smem > 48 * 1024) QSB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  return QSB_OK;
}

template <class K>
static inline int qsb_resident_ctas(K kernel, int threads, size_t smem) {
#ifdef QSB_EMU
  (void)kernel; (void)threads; (void)smem;
  return 2;
#else
  int n = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, threads, smem) != cudaSuccess || n < 1) n = 1;
  return n;
#endif
}
