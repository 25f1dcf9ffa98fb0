// Host-side objects behind the opaque handles of include/qsb200.h.
#pragma once
#include <cstdint>
#include <mutex>
#include <string>
#include <utility>
#include <vector>

#include "qsb_device.cuh"

struct QsbBuf {
  void* p = nullptr;
  size_t bytes = 0;
};

// Per-device cache of freed device blocks: the host-buffer calls and the solvers build a short-lived batch per call, and a
// cudaMalloc / cudaFree pair per buffer per call (cudaFree synchronises the device) showed up as 0.1-0.8x first calls and
// run-to-run spread.  Blocks are reused for requests of the same rounded size.
struct QsbPool {
  std::mutex mu;
  std::vector<std::pair<size_t, void*>> free_blocks;
  size_t cached_bytes = 0;
};

struct QsbDevice {
  int id = 0;
  int sm_count = 148;
  cudaStream_t stream = nullptr;       // compute
  cudaStream_t copy_stream = nullptr;  // device->host copies of the host-buffer calls (overlap with the next chunk)
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  // solver scratch (slabs, Gram matrices, snapshots: up to a few hundred MB) kept across qsb_solve_* calls instead of a
  // cudaMalloc / cudaFree pair per frontier expansion; solve_mu serialises the solves of one device
  QsbBuf solve_scratch;
  std::mutex* solve_mu = nullptr;
  QsbPool* pool = nullptr;
};

struct qsb_ctx {
  std::vector<QsbDevice> devs;
  uint64_t launches = 0;
  bool force_generic = false;  // QSB_KERNEL=generic in the environment at qsb_init
  std::mutex mu;
};

struct qsb_program {
  int n_qubits = 0, n_ops = 0, n_params = 0;   // n_qubits counts qudits of size radix
  int radix = 2;
  size_t dim = 0;                               // N = radix^n_qubits
  std::vector<qsb_op> ops;
  std::vector<double> consts;
  bool fast_ok = false;             // only U3/XZXZ/.../CONST1/CNOT ops: eligible for the register-tiled kernels
  std::vector<qsb_op*> d_ops;       // per device slot
  std::vector<double*> d_consts;
  std::vector<double*> d_pscale;
  std::vector<int32_t*> d_slotmap;
};

struct qsb_batch {
  qsb_ctx* ctx = nullptr;
  int slot = 0;
  int n_qubits = 0, n_instances = 0, n_programs = 0;
  int radix = 2;
  size_t dim = 0;                    // N = radix^n_qubits
  int max_ops = 0, max_params = 0;
  int max_pops = 0;                 // most ops with parameters in one program (snapshots of the body-frame LM's forward pass)
  bool fast_ok = false;
  bool target_unitary = false;      // set by qsb_batch_set_target: selects the body-frame LM forms (qsb_lm3.cuh)
  bool force_generic = false;       // tests / profiling: run the shared-memory interpreter even where a warp kernel exists
  std::vector<long long> param_off;  // host copy, n_instances + 1
  std::vector<int32_t> prog_idx;
  QsbDevProgram* d_progs = nullptr;
  int32_t* d_prog_idx = nullptr;
  long long* d_param_off = nullptr;
  long long* d_sq_off = nullptr;     // prefix sums of P_i^2 (QSB_RUN_LMSYS output offsets), built on first use
  double* d_params = nullptr;
  double* d_target = nullptr;
  size_t sz_progs = 0, sz_prog_idx = 0, sz_param_off = 0, sz_params = 0, sz_target = 0;  // bytes requested from the pool
  QsbBuf out_a, out_b, out_c;  // grown on demand
  int last_first = 0, last_count = 0, last_what = 0;
  void* solver_scratch = nullptr;
  size_t solver_scratch_bytes = 0;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;  // timed runs: own pair, so two threads timing batches of one device do not race
};

// number of ops of a program that have parameters (= snapshots of a forward pass); host side
static inline int qsb_count_param_ops(const qsb_op* ops, int n_ops) {
  int n = 0;
  for (int j = 0; j < n_ops; ++j) {
    switch (ops[j].opcode) {
      case QSB_OP_RX: case QSB_OP_RY: case QSB_OP_RZ: case QSB_OP_U1: case QSB_OP_U2: case QSB_OP_U3: case QSB_OP_XZXZ: case QSB_OP_ZXZXZ: ++n; break;
      default: break;
    }
  }
  return n;
}

void qsb_set_error(const std::string& msg);
const std::string& qsb_get_error();
int qsb_buf_reserve(QsbBuf& b, size_t bytes);  // cudaMalloc-backed, for long-lived buffers
int qsb_pool_alloc(const QsbDevice& d, void** out, size_t bytes);   // the device of d must be current
void qsb_pool_free(const QsbDevice& d, void* p, size_t bytes);
int qsb_pool_reserve(const QsbDevice& d, QsbBuf& b, size_t bytes);  // grow-only buffer on the pool
bool qsb_target_is_unitary(const double* T, int dim);  // dim = N
int qsb_fail(int code, const std::string& msg);
int qsb_cuda_fail(cudaError_t e, const char* what);

#define QSB_CUDA(call)                                            \
  do {                                                            \
    cudaError_t _e = (call);                                      \
    if (_e != cudaSuccess) return qsb_cuda_fail(_e, #call);       \
  } while (0)
