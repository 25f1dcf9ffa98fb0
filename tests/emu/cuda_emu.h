// TEST INFRASTRUCTURE ONLY -- never part of the product.
//
// A tiny host emulation of the CUDA execution model so that the *same* kernel source in
// qsearch_b200/csrc/ can be compiled with g++ (-DQSB_EMU) and exercised by the CPU test-suite in the
// build container, which has no GPU.  Every CUDA thread of a block is a real std::thread; __syncthreads
// and warp primitives are std::barrier exchanges.  Blocks of a launch run one after another.
// The product library (libqsb200.so) is built by nvcc without QSB_EMU and never sees this file.
#pragma once
#include <atomic>
#include <barrier>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <chrono>
#include <functional>
#include <memory>
#include <thread>
#include <vector>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __restrict__
#define __launch_bounds__(...)
#define __align__(n) alignas(n)

struct double2 { double x, y; };
static inline double2 make_double2(double x, double y) { return double2{x, y}; }
struct dim3 { unsigned x = 1, y = 1, z = 1; dim3(unsigned a = 1, unsigned b = 1, unsigned c = 1) : x(a), y(b), z(c) {} };

namespace qsb_emu {
struct BlockState {
  std::unique_ptr<std::barrier<>> block_bar;
  std::vector<std::unique_ptr<std::barrier<>>> warp_bar;
  std::vector<std::vector<uint64_t>> warp_buf;  // [warp][32]
  unsigned char* smem = nullptr;
};
inline BlockState*& cur() { static BlockState* p = nullptr; return p; }
}  // namespace qsb_emu

extern thread_local dim3 threadIdx, blockIdx, blockDim, gridDim;
#ifdef QSB_EMU_DEFINE
thread_local dim3 threadIdx, blockIdx, blockDim, gridDim;
#endif

static inline void __syncthreads() { qsb_emu::cur()->block_bar->arrive_and_wait(); }
static inline void __syncwarp(unsigned = 0xffffffffu) { qsb_emu::cur()->warp_bar[threadIdx.x / 32]->arrive_and_wait(); }

template <class T>
static inline T qsb_emu_xchg(T v, int src_lane) {
  static_assert(sizeof(T) <= 8, "shuffle of <= 8 bytes");
  auto* st = qsb_emu::cur();
  int w = threadIdx.x / 32, l = threadIdx.x % 32;
  uint64_t raw = 0;
  std::memcpy(&raw, &v, sizeof(T));
  st->warp_buf[w][l] = raw;
  st->warp_bar[w]->arrive_and_wait();
  uint64_t got = st->warp_buf[w][src_lane & 31];
  st->warp_bar[w]->arrive_and_wait();
  T out;
  std::memcpy(&out, &got, sizeof(T));
  return out;
}
template <class T> static inline T __shfl_xor_sync(unsigned, T v, int m) { return qsb_emu_xchg(v, (threadIdx.x % 32) ^ m); }
template <class T> static inline T __shfl_sync(unsigned, T v, int src) { return qsb_emu_xchg(v, src); }
template <class T> static inline T __shfl_down_sync(unsigned, T v, int d) {
  int l = threadIdx.x % 32;
  return qsb_emu_xchg(v, l + d < 32 ? l + d : l);
}
template <class T> static inline T __ldg(const T* p) { return *p; }
static inline double atomicAdd(double* p, double v) {
  static std::atomic_flag lock = ATOMIC_FLAG_INIT;
  while (lock.test_and_set(std::memory_order_acquire)) {}
  double old = *p; *p = old + v;
  lock.clear(std::memory_order_release);
  return old;
}
static inline int atomicAdd(int* p, int v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
static inline double rsqrt(double x) { return 1.0 / std::sqrt(x); }
static inline double __longlong_as_double(long long v) { double d; std::memcpy(&d, &v, 8); return d; }

// ------------------------------------------------------------------ runtime API subset
typedef int cudaError_t;
typedef struct qsb_emu_stream* cudaStream_t;
struct qsb_emu_event { std::chrono::steady_clock::time_point t; };
typedef qsb_emu_event* cudaEvent_t;
enum { cudaSuccess = 0, cudaErrorInvalidValue = 1 };
enum cudaMemcpyKind { cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice, cudaMemcpyDefault };
enum { cudaFuncAttributeMaxDynamicSharedMemorySize = 8 };
enum { cudaDevAttrMultiProcessorCount = 16 };
static inline cudaError_t cudaGetDeviceCount(int* n) { *n = 1; return 0; }
static inline cudaError_t cudaSetDevice(int) { return 0; }
static inline cudaError_t cudaDeviceGetAttribute(int* v, int, int) { *v = 4; return 0; }
static inline cudaError_t cudaMalloc(void** p, size_t n) { *p = std::malloc(n ? n : 1); return *p ? 0 : 2; }
static inline cudaError_t cudaFree(void* p) { std::free(p); return 0; }
static inline cudaError_t cudaMallocHost(void** p, size_t n) { return cudaMalloc(p, n); }
static inline cudaError_t cudaFreeHost(void* p) { return cudaFree(p); }
static inline cudaError_t cudaMemcpy(void* d, const void* s, size_t n, cudaMemcpyKind) { std::memcpy(d, s, n); return 0; }
static inline cudaError_t cudaMemcpyAsync(void* d, const void* s, size_t n, cudaMemcpyKind, cudaStream_t = nullptr) { std::memcpy(d, s, n); return 0; }
static inline cudaError_t cudaMemsetAsync(void* d, int v, size_t n, cudaStream_t = nullptr) { std::memset(d, v, n); return 0; }
static inline cudaError_t cudaStreamCreate(cudaStream_t* s) { *s = nullptr; return 0; }
static inline cudaError_t cudaStreamCreateWithFlags(cudaStream_t* s, unsigned) { *s = nullptr; return 0; }
static inline cudaError_t cudaStreamDestroy(cudaStream_t) { return 0; }
static inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return 0; }
static inline cudaError_t cudaDeviceSynchronize() { return 0; }
static inline cudaError_t cudaEventCreate(cudaEvent_t* e) { *e = new qsb_emu_event(); return 0; }
static inline cudaError_t cudaEventDestroy(cudaEvent_t e) { delete e; return 0; }
static inline cudaError_t cudaEventRecord(cudaEvent_t e, cudaStream_t = nullptr) { e->t = std::chrono::steady_clock::now(); return 0; }
static inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return 0; }
static inline cudaError_t cudaEventElapsedTime(float* ms, cudaEvent_t a, cudaEvent_t b) {
  *ms = std::chrono::duration<float, std::milli>(b->t - a->t).count(); return 0;
}
static inline cudaError_t cudaGetLastError() { return 0; }
static inline const char* cudaGetErrorString(cudaError_t) { return "emulated"; }
template <class F> static inline cudaError_t cudaFuncSetAttribute(F, int, int) { return 0; }
static inline cudaError_t cudaMemGetInfo(size_t* f, size_t* t) { *f = *t = (size_t)8 << 30; return 0; }
#define cudaStreamNonBlocking 1

// ------------------------------------------------------------------ kernel launch
namespace qsb_emu {
template <class F>
inline void launch(unsigned grid, unsigned block, size_t smem_bytes, F body) {
  BlockState st;
  st.block_bar = std::make_unique<std::barrier<>>(block);
  unsigned nw = (block + 31) / 32;
  for (unsigned w = 0; w < nw; ++w) {
    unsigned cnt = (w + 1) * 32 <= block ? 32 : block - w * 32;
    st.warp_bar.push_back(std::make_unique<std::barrier<>>(cnt));
    st.warp_buf.emplace_back(32, 0);
  }
  std::vector<unsigned char> sm(smem_bytes + 64);
  st.smem = (unsigned char*)(((uintptr_t)sm.data() + 63) & ~(uintptr_t)63);
  cur() = &st;
  std::vector<std::thread> th;
  for (unsigned t = 0; t < block; ++t)
    th.emplace_back([&, t] {
      threadIdx = dim3(t); blockDim = dim3(block); gridDim = dim3(grid);
      for (unsigned b = 0; b < grid; ++b) {
        blockIdx = dim3(b);
        body();
        st.block_bar->arrive_and_wait();
      }
    });
  for (auto& x : th) x.join();
  cur() = nullptr;
}
}  // namespace qsb_emu
#define QSB_DYN_SMEM(name) unsigned char* name = qsb_emu::cur()->smem
#define QSB_LAUNCH(kernel, grid, block, smem, stream, ...) \
  qsb_emu::launch((grid), (block), (smem), [&] { kernel(__VA_ARGS__); })
