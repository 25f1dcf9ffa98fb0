// TEST INFRASTRUCTURE ONLY -- never part of the product.
//
// A tiny host emulation of the CUDA execution model so that the *same* kernel source in
// qsearch_b200/csrc/ can be compiled with g++ (-DQSB_EMU) and exercised by the CPU test-suite in the
// build container, which has no GPU.  Every CUDA thread of a block is a user-level fiber (own stack, hand-written
// x86-64 context switch) scheduled round-robin inside one OS thread; __syncthreads and the warp primitives are
// barrier counters that yield to the next fiber.  Blocks of a launch are spread over a few OS worker threads.
// The product library (libqsb200.so) is built by nvcc without QSB_EMU and never sees this file.
#pragma once
#include <algorithm>
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
#define __align__(n) __attribute__((aligned(n)))

struct double2 { double x, y; };
static inline double2 make_double2(double x, double y) { return double2{x, y}; }
struct float2 { float x, y; };
static inline float2 make_float2(float x, float y) { return float2{x, y}; }
struct alignas(16) float4 { float x, y, z, w; };

namespace qsb_emu {
struct Fiber {
  void* sp = nullptr;
  unsigned tid = 0;
  bool done = false;
  std::vector<unsigned char> stack;
};
struct BlockState {
  unsigned nthreads = 0, block = 0, grid = 0;
  std::vector<Fiber> fibers;
  unsigned current = 0, alive = 0;
  void* sched_sp = nullptr;
  unsigned block_arrived = 0;
  unsigned long block_gen = 0;
  std::vector<unsigned> warp_arrived, warp_alive;
  std::vector<unsigned long> warp_gen;
  std::vector<std::vector<uint64_t>> warp_buf, warp_buf2;  // [warp][32]
  unsigned char* smem = nullptr;
  const std::function<void()>* body = nullptr;
  std::barrier<>* cluster_bar = nullptr;  // shared by the CTAs of one emulated thread-block cluster
};
inline BlockState*& cur() { static thread_local BlockState* p = nullptr; return p; }
extern "C" void qsb_emu_switch(void** save_sp, void* load_sp);
inline void yield_to_scheduler() {
  BlockState* st = cur();
  qsb_emu_switch(&st->fibers[st->current].sp, st->sched_sp);
}
}  // namespace qsb_emu

struct qsb_emu_idx { unsigned x, y, z; };
#define threadIdx (qsb_emu_idx{qsb_emu::cur()->fibers[qsb_emu::cur()->current].tid, 0u, 0u})
#define blockIdx (qsb_emu_idx{qsb_emu::cur()->block, 0u, 0u})
#define blockDim (qsb_emu_idx{qsb_emu::cur()->nthreads, 1u, 1u})
#define gridDim (qsb_emu_idx{qsb_emu::cur()->grid, 1u, 1u})

static inline void __syncthreads() {
  auto* st = qsb_emu::cur();
  const unsigned long gen = st->block_gen;
  if (++st->block_arrived >= st->alive) { st->block_arrived = 0; ++st->block_gen; return; }
  while (st->block_gen == gen) qsb_emu::yield_to_scheduler();
}
static inline void __syncwarp(unsigned = 0xffffffffu) {
  auto* st = qsb_emu::cur();
  const unsigned w = st->fibers[st->current].tid / 32;
  const unsigned long gen = st->warp_gen[w];
  if (++st->warp_arrived[w] >= st->warp_alive[w]) { st->warp_arrived[w] = 0; ++st->warp_gen[w]; return; }
  while (st->warp_gen[w] == gen) qsb_emu::yield_to_scheduler();
}

template <class T>
static inline T qsb_emu_xchg(T v, int src_lane) {
  static_assert(sizeof(T) <= 8, "shuffle of <= 8 bytes");
  auto* st = qsb_emu::cur();
  int w = threadIdx.x / 32, l = threadIdx.x % 32;
  uint64_t raw = 0;
  std::memcpy(&raw, &v, sizeof(T));
  st->warp_buf[w][l] = raw;
  __syncwarp();
  uint64_t got = st->warp_buf[w][src_lane & 31];
  __syncwarp();
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
static inline int atomicMax(int* p, int v) {
  int old = __atomic_load_n(p, __ATOMIC_SEQ_CST);
  while (old < v && !__atomic_compare_exchange_n(p, &old, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) {}
  return old;
}
static inline unsigned long long atomicAdd(unsigned long long* p, unsigned long long v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
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
static inline cudaError_t cudaGetDeviceCount(int* n) {  // QSB_EMU_DEVICES: pretend to have several GPUs (shard tests)
  const char* e = std::getenv("QSB_EMU_DEVICES");
  *n = e ? std::max(1, std::atoi(e)) : 1;
  return 0;
}
static inline cudaError_t cudaSetDevice(int) { return 0; }
static inline cudaError_t cudaDeviceGetAttribute(int* v, int, int) { *v = 4; return 0; }
static inline cudaError_t cudaMalloc(void** p, size_t n) { *p = std::malloc(n ? n : 1); return *p ? 0 : 2; }
static inline cudaError_t cudaFree(void* p) { std::free(p); return 0; }
static inline cudaError_t cudaMallocHost(void** p, size_t n) { return cudaMalloc(p, n); }
static inline cudaError_t cudaFreeHost(void* p) { return cudaFree(p); }
static inline cudaError_t cudaMemcpy(void* d, const void* s, size_t n, cudaMemcpyKind) { std::memcpy(d, s, n); return 0; }
static inline cudaError_t cudaMemcpyAsync(void* d, const void* s, size_t n, cudaMemcpyKind, cudaStream_t = nullptr) { std::memcpy(d, s, n); return 0; }
static inline cudaError_t cudaMemset(void* d, int v, size_t n) { std::memset(d, v, n); return 0; }
static inline cudaError_t cudaMemsetAsync(void* d, int v, size_t n, cudaStream_t = nullptr) { std::memset(d, v, n); return 0; }
static inline cudaError_t cudaStreamCreate(cudaStream_t* s) { *s = nullptr; return 0; }
static inline cudaError_t cudaStreamCreateWithFlags(cudaStream_t* s, unsigned) { *s = nullptr; return 0; }
static inline cudaError_t cudaStreamDestroy(cudaStream_t) { return 0; }
static inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return 0; }
static inline cudaError_t cudaDeviceSynchronize() { return 0; }
static inline cudaError_t cudaEventCreate(cudaEvent_t* e) { *e = new qsb_emu_event(); return 0; }
static inline cudaError_t cudaEventDestroy(cudaEvent_t e) { delete e; return 0; }
static inline cudaError_t cudaEventCreateWithFlags(cudaEvent_t* e, unsigned) { *e = new qsb_emu_event(); return 0; }
static inline cudaError_t cudaStreamWaitEvent(cudaStream_t, cudaEvent_t, unsigned) { return 0; }
static inline cudaError_t cudaHostAlloc(void** p, size_t n, unsigned) { return cudaMalloc(p, n); }
#define cudaEventDisableTiming 2
#define cudaHostAllocPortable 1
static inline cudaError_t cudaEventRecord(cudaEvent_t e, cudaStream_t = nullptr) { e->t = std::chrono::steady_clock::now(); return 0; }
static inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return 0; }
static inline cudaError_t cudaEventElapsedTime(float* ms, cudaEvent_t a, cudaEvent_t b) {
  *ms = std::chrono::duration<float, std::milli>(b->t - a->t).count(); return 0;
}
static inline cudaError_t cudaGetLastError() { return 0; }
template <class T> static inline cudaError_t cudaGetSymbolAddress(void** p, T& sym) { *p = (void*)&sym; return 0; }  // emulated: host address
static inline const char* cudaGetErrorString(cudaError_t) { return "emulated"; }
template <class F> static inline cudaError_t cudaFuncSetAttribute(F, int, int) { return 0; }
static inline cudaError_t cudaMemGetInfo(size_t* f, size_t* t) { *f = *t = (size_t)8 << 30; return 0; }
#define cudaStreamNonBlocking 1

// ------------------------------------------------------------------ kernel launch
namespace qsb_emu {
#ifdef QSB_EMU_DEFINE
asm(R"(
.text
.globl qsb_emu_switch
.type qsb_emu_switch,@function
qsb_emu_switch:
  pushq %rbp
  pushq %rbx
  pushq %r12
  pushq %r13
  pushq %r14
  pushq %r15
  movq %rsp, (%rdi)
  movq %rsi, %rsp
  popq %r15
  popq %r14
  popq %r13
  popq %r12
  popq %rbx
  popq %rbp
  ret
)");
extern "C" void qsb_emu_fiber_main() {
  BlockState* st = cur();
  (*st->body)();
  st = cur();
  Fiber& f = st->fibers[st->current];
  f.done = true;
  --st->alive;
  const unsigned w = f.tid / 32;
  --st->warp_alive[w];
  // a finished fiber may complete a barrier the others are already waiting on
  if (st->alive && st->block_arrived >= st->alive) { st->block_arrived = 0; ++st->block_gen; }
  if (st->warp_alive[w] && st->warp_arrived[w] >= st->warp_alive[w]) { st->warp_arrived[w] = 0; ++st->warp_gen[w]; }
  yield_to_scheduler();
  std::abort();  // a finished fiber is never resumed
}
#else
extern "C" void qsb_emu_fiber_main();
#endif

inline void run_block(BlockState& st, unsigned block_index) {
  constexpr size_t STACK = 96 * 1024;
  st.block = block_index;
  st.alive = st.nthreads;
  st.block_arrived = 0;
  const unsigned nw = (st.nthreads + 31) / 32;
  st.warp_arrived.assign(nw, 0);
  st.warp_gen.assign(nw, 0);
  st.warp_alive.assign(nw, 0);
  for (unsigned t = 0; t < st.nthreads; ++t) {
    Fiber& f = st.fibers[t];
    if (f.stack.empty()) f.stack.resize(STACK);
    f.tid = t;
    f.done = false;
    ++st.warp_alive[t / 32];
    uintptr_t top = ((uintptr_t)f.stack.data() + STACK) & ~(uintptr_t)15;
    void** sp = (void**)top;
    *--sp = nullptr;                              // fake return address of the entry function
    *--sp = (void*)&qsb_emu_fiber_main;           // popped by `ret` of the first switch
    for (int i = 0; i < 6; ++i) *--sp = nullptr;  // rbp rbx r12..r15
    f.sp = sp;
  }
  cur() = &st;
  while (st.alive) {
    for (unsigned t = 0; t < st.nthreads && st.alive; ++t) {
      if (st.fibers[t].done) continue;
      st.current = t;
      qsb_emu_switch(&st.sched_sp, st.fibers[t].sp);
    }
  }
}

inline void launch(unsigned grid, unsigned block, size_t smem_bytes, const std::function<void()>& body) {
  unsigned workers = std::min<unsigned>(grid, std::max(1u, std::thread::hardware_concurrency()));
  std::atomic<unsigned> next{0};
  auto work = [&] {
    BlockState st;
    st.nthreads = block;
    st.grid = grid;
    st.fibers.resize(block);
    st.warp_buf.assign((block + 31) / 32, std::vector<uint64_t>(32, 0));
    st.warp_buf2 = st.warp_buf;
    std::vector<unsigned char> sm(smem_bytes + 64);
    st.smem = (unsigned char*)(((uintptr_t)sm.data() + 63) & ~(uintptr_t)63);
    st.body = &body;
    for (unsigned b = next.fetch_add(1); b < grid; b = next.fetch_add(1)) run_block(st, b);
    cur() = nullptr;
  };
  if (workers <= 1) { work(); return; }
  std::vector<std::thread> th;
  for (unsigned w = 0; w < workers; ++w) th.emplace_back(work);
  for (auto& x : th) x.join();
}
// Thread-block clusters: every block gets its own OS thread (tests launch few blocks), blocks of one cluster share
// a std::barrier.  qsb_emu_cluster_sync = block barrier, one fiber waits on the cluster barrier, block barrier.
inline void launch_cluster(unsigned grid, unsigned block, size_t smem_bytes, unsigned csize, const std::function<void()>& body) {
  std::vector<std::unique_ptr<std::barrier<>>> bars;
  for (unsigned c = 0; c < grid / csize; ++c) bars.emplace_back(new std::barrier<>(csize));
  std::vector<std::thread> th;
  for (unsigned b = 0; b < grid; ++b)
    th.emplace_back([&, b] {
      BlockState st;
      st.nthreads = block;
      st.grid = grid;
      st.fibers.resize(block);
      st.warp_buf.assign((block + 31) / 32, std::vector<uint64_t>(32, 0));
      st.warp_buf2 = st.warp_buf;
      std::vector<unsigned char> sm(smem_bytes + 64);
      st.smem = (unsigned char*)(((uintptr_t)sm.data() + 63) & ~(uintptr_t)63);
      st.body = &body;
      st.cluster_bar = bars[b / csize].get();
      run_block(st, b);
      cur() = nullptr;
    });
  for (auto& x : th) x.join();
}
}  // namespace qsb_emu
static inline void qsb_emu_cluster_sync() {
  __syncthreads();
  auto* st = qsb_emu::cur();
  if (st->cluster_bar && st->fibers[st->current].tid == 0) st->cluster_bar->arrive_and_wait();
  __syncthreads();
}
#define QSB_LAUNCH_CLUSTER(kernel, grid, block, smem, csize, ...) \
  qsb_emu::launch_cluster((grid), (block), (smem), (csize), std::function<void()>([&] { kernel(__VA_ARGS__); }))
#define QSB_DYN_SMEM(name) unsigned char* name = qsb_emu::cur()->smem
#define QSB_LAUNCH(kernel, grid, block, smem, stream, ...) \
  qsb_emu::launch((grid), (block), (smem), std::function<void()>([&] { kernel(__VA_ARGS__); }))
