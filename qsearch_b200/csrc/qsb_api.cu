// C ABI of libqsb200.so (include/qsb200.h): context, gate programs, device-resident batches and the
// host-buffer evaluation calls.  Replaces the PyO3 boundary of the reference (qsrs/src/lib.rs).
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <thread>

#include "qsb_host.h"
#include "qsb_generic.cuh"
#include "qsb_warp.cuh"
#include "qsb_launch_eval.h"

// ---------------------------------------------------------------------------- errors
static thread_local std::string g_err;
void qsb_set_error(const std::string& msg) { g_err = msg; }
const std::string& qsb_get_error() { return g_err; }
int qsb_fail(int code, const std::string& msg) { g_err = msg; return code; }
int qsb_cuda_fail(cudaError_t e, const char* what) {
  g_err = std::string("CUDA error: ") + cudaGetErrorString(e) + " in " + what;
  return QSB_ERR_CUDA;
}
extern "C" const char* qsb_last_error(void) { return g_err.c_str(); }
extern "C" const char* qsb_version(void) {
#ifdef QSB_EMU
  return "qsb200 0.1 (host emulation: tests only)";
#else
  return "qsb200 0.1 (sm_100a)";
#endif
}

// ---------------------------------------------------------------------------- context
extern "C" int qsb_device_count(int* out_count) {
  if (!out_count) return qsb_fail(QSB_ERR_INVALID, "out_count is NULL");
  QSB_CUDA(cudaGetDeviceCount(out_count));
  return QSB_OK;
}

extern "C" int qsb_init(int n_devices, const int* device_ids, qsb_ctx** out_ctx) {
  if (!out_ctx) return qsb_fail(QSB_ERR_INVALID, "out_ctx is NULL");
  int avail = 0;
  QSB_CUDA(cudaGetDeviceCount(&avail));
  if (avail <= 0) return qsb_fail(QSB_ERR_CUDA, "no CUDA device visible (this library has no CPU fallback)");
  std::vector<int> ids;
  if (n_devices <= 0) { for (int i = 0; i < avail; ++i) ids.push_back(i); }
  else {
    for (int i = 0; i < n_devices; ++i) {
      int id = device_ids ? device_ids[i] : i;
      if (id < 0 || id >= avail) return qsb_fail(QSB_ERR_INVALID, "device id out of range");
      ids.push_back(id);
    }
  }
  qsb_ctx* ctx = new (std::nothrow) qsb_ctx();
  if (!ctx) return qsb_fail(QSB_ERR_NOMEM, "out of host memory");
  if (const char* k = getenv("QSB_KERNEL")) ctx->force_generic = (strcmp(k, "generic") == 0);
  for (int id : ids) {
    QsbDevice d;
    d.id = id;
    cudaError_t e;
    if ((e = cudaSetDevice(id)) != cudaSuccess || (e = cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, id)) != cudaSuccess ||
        (e = cudaStreamCreateWithFlags(&d.stream, cudaStreamNonBlocking)) != cudaSuccess ||
        (e = cudaStreamCreateWithFlags(&d.copy_stream, cudaStreamNonBlocking)) != cudaSuccess ||
        (e = cudaEventCreate(&d.ev0)) != cudaSuccess || (e = cudaEventCreate(&d.ev1)) != cudaSuccess) {
      const int rc = qsb_cuda_fail(e, "qsb_init (device setup)");
      const std::string keep = qsb_get_error();
      if (d.stream) cudaStreamDestroy(d.stream);
      if (d.copy_stream) cudaStreamDestroy(d.copy_stream);
      if (d.ev0) cudaEventDestroy(d.ev0);
      qsb_shutdown(ctx);  // the devices set up before this one
      qsb_set_error(keep);
      return rc;
    }
    d.solve_mu = new std::mutex();
    d.pool = new QsbPool();
    ctx->devs.push_back(d);
  }
  *out_ctx = ctx;
  return QSB_OK;
}

extern "C" int qsb_shutdown(qsb_ctx* ctx) {
  if (!ctx) return QSB_OK;
  for (auto& d : ctx->devs) {
    cudaSetDevice(d.id);
    cudaStreamSynchronize(d.stream);
    cudaEventDestroy(d.ev0);
    cudaEventDestroy(d.ev1);
    cudaStreamDestroy(d.stream);
    cudaStreamDestroy(d.copy_stream);
    if (d.solve_scratch.p) cudaFree(d.solve_scratch.p);
    if (d.pool) { for (auto& fb : d.pool->free_blocks) cudaFree(fb.second); delete d.pool; }
    delete d.solve_mu;
  }
  delete ctx;
  return QSB_OK;
}

extern "C" int qsb_host_alloc(qsb_ctx* ctx, size_t bytes, void** out) {
  if (!ctx || !out) return qsb_fail(QSB_ERR_INVALID, "NULL argument");
  QSB_CUDA(cudaSetDevice(ctx->devs[0].id));
  cudaError_t e = cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocPortable);
  if (e != cudaSuccess) { *out = nullptr; return qsb_fail(QSB_ERR_NOMEM, std::string("pinned host allocation failed: ") + cudaGetErrorString(e)); }
  return QSB_OK;
}
extern "C" int qsb_host_free(qsb_ctx* ctx, void* p) {
  (void)ctx;
  if (p) cudaFreeHost(p);
  return QSB_OK;
}

extern "C" int qsb_ctx_num_devices(const qsb_ctx* ctx) { return ctx ? (int)ctx->devs.size() : 0; }
extern "C" int qsb_ctx_launch_count(const qsb_ctx* ctx, uint64_t* out) {
  if (!ctx || !out) return qsb_fail(QSB_ERR_INVALID, "NULL argument");
  *out = ctx->launches;
  return QSB_OK;
}

// ---------------------------------------------------------------------------- programs
static int op_params(int opcode) {
  switch (opcode) {
    case QSB_OP_RX: case QSB_OP_RY: case QSB_OP_RZ: case QSB_OP_U1: case QSB_OP_RXX: case QSB_OP_RYY: case QSB_OP_RZZ: return 1;
    case QSB_OP_U2: case QSB_OP_XZXZ: return 2;
    case QSB_OP_U3: case QSB_OP_ZXZXZ: return 3;
    case QSB_OP_QUTRIT1: return 8;
    default: return 0;
  }
}

static int program_create(qsb_ctx* ctx, int radix, int n_qubits, int n_ops, const qsb_op* ops, int n_params,
                          const double* const_pool, size_t const_len, qsb_program** out_prog) {
  if (!ctx || !out_prog || (n_ops > 0 && !ops)) return qsb_fail(QSB_ERR_INVALID, "NULL argument");
  if (radix != 2 && radix != 3) return qsb_fail(QSB_ERR_UNSUPPORTED, "radix must be 2 (qubits) or 3 (qutrits)");
  if (radix == 2 && (n_qubits < 1 || n_qubits > QSB_MAX_QUBITS))
    return qsb_fail(QSB_ERR_UNSUPPORTED, "n_qubits must be in 1.." + std::to_string(QSB_MAX_QUBITS));
  if (radix == 3 && (n_qubits < 1 || n_qubits > QSB_MAX_QUTRITS))
    return qsb_fail(QSB_ERR_UNSUPPORTED, "n_qudits must be in 1.." + std::to_string(QSB_MAX_QUTRITS) + " for radix 3");
  if (n_ops < 0 || n_params < 0) return qsb_fail(QSB_ERR_INVALID, "negative size");
  std::vector<char> seen(n_params, 0);
  bool fast = radix == 2;
  for (int j = 0; j < n_ops; ++j) {
    const qsb_op& o = ops[j];
    int np = op_params(o.opcode);
    if (radix == 3 && o.opcode != QSB_OP_QUTRIT1 && o.opcode != QSB_OP_CONSTK)
      return qsb_fail(QSB_ERR_INVALID, "op " + std::to_string(j) + ": a radix-3 program takes QSB_OP_QUTRIT1 and QSB_OP_CONSTK ops only");
    switch (o.opcode) {
      case QSB_OP_QUTRIT1:
        if (radix != 3) return qsb_fail(QSB_ERR_INVALID, "op " + std::to_string(j) + ": QSB_OP_QUTRIT1 needs a radix-3 program");
        if (o.q0 < 0 || o.q0 >= n_qubits) return qsb_fail(QSB_ERR_INVALID, "op " + std::to_string(j) + ": qutrit out of range");
        if (o.param_off < 0 || o.param_off + np > n_params)
          return qsb_fail(QSB_ERR_INVALID, "op " + std::to_string(j) + ": parameter range out of bounds");
        for (int k = 0; k < np; ++k) {
          if (seen[o.param_off + k]) return qsb_fail(QSB_ERR_INVALID, "parameter used by two ops");
          seen[o.param_off + k] = 1;
        }
        break;
      case QSB_OP_RX: case QSB_OP_RY: case QSB_OP_RZ: case QSB_OP_U1: case QSB_OP_U2: case QSB_OP_U3:
      case QSB_OP_XZXZ: case QSB_OP_ZXZXZ:
        if (o.q0 < 0 || o.q0 >= n_qubits) return qsb_fail(QSB_ERR_INVALID, "op " + std::to_string(j) + ": qubit out of range");
        if (o.param_off < 0 || o.param_off + np > n_params)
          return qsb_fail(QSB_ERR_INVALID, "op " + std::to_string(j) + ": parameter range out of bounds");
        for (int k = 0; k < np; ++k) {
          if (seen[o.param_off + k]) return qsb_fail(QSB_ERR_INVALID, "parameter used by two ops");
          seen[o.param_off + k] = 1;
        }
        break;
      case QSB_OP_CONST1:
        if (o.q0 < 0 || o.q0 >= n_qubits) return qsb_fail(QSB_ERR_INVALID, "op " + std::to_string(j) + ": qubit out of range");
        if (o.const_off < 0 || (size_t)o.const_off + 8 > const_len) return qsb_fail(QSB_ERR_INVALID, "constant out of bounds");
        break;
      case QSB_OP_CNOT:
        if (o.q0 < 0 || o.q0 >= n_qubits || o.q1 < 0 || o.q1 >= n_qubits || o.q0 == o.q1)
          return qsb_fail(QSB_ERR_INVALID, "op " + std::to_string(j) + ": bad CNOT qubits");
        break;
      case QSB_OP_RXX: case QSB_OP_RYY: case QSB_OP_RZZ:
        if (o.q0 < 0 || o.q0 >= n_qubits || o.q1 < 0 || o.q1 >= n_qubits || o.q0 == o.q1)
          return qsb_fail(QSB_ERR_INVALID, "op " + std::to_string(j) + ": bad two-qubit rotation qubits");
        if (o.param_off < 0 || o.param_off + 1 > n_params) return qsb_fail(QSB_ERR_INVALID, "op " + std::to_string(j) + ": parameter range out of bounds");
        if (seen[o.param_off]) return qsb_fail(QSB_ERR_INVALID, "parameter used by two ops");
        seen[o.param_off] = 1;
        fast = false;  // interpreter kernels only (the warp kernels are specialised to one-qubit leaves + CNOT)
        break;
      case QSB_OP_CONSTK: {
        int k = o.q1;
        if (k < 1 || o.q0 < 0 || o.q0 + k > n_qubits) return qsb_fail(QSB_ERR_INVALID, "op " + std::to_string(j) + ": bad CONSTK range");
        const size_t kk = qsb_geom_dim(radix, k);
        if (o.const_off < 0 || (size_t)o.const_off + 2 * kk * kk > const_len)
          return qsb_fail(QSB_ERR_INVALID, "constant out of bounds");
        fast = false;
      } break;
      default:
        return qsb_fail(QSB_ERR_INVALID, "op " + std::to_string(j) + ": unknown opcode " + std::to_string(o.opcode));
    }
  }
  for (int k = 0; k < n_params; ++k)
    if (!seen[k]) return qsb_fail(QSB_ERR_INVALID, "parameter " + std::to_string(k) + " is not used by any op");
  qsb_program* p = new (std::nothrow) qsb_program();
  if (!p) return qsb_fail(QSB_ERR_NOMEM, "out of host memory");
  p->n_qubits = n_qubits; p->n_ops = n_ops; p->n_params = n_params; p->fast_ok = fast;
  p->radix = radix; p->dim = qsb_geom_dim(radix, n_qubits);
  p->ops.assign(ops, ops + n_ops);
  if (const_len) p->consts.assign(const_pool, const_pool + const_len);
  std::vector<double> pscale(std::max(1, n_params), 1.0);  // angle scale per parameter (half angles: gates.py:324-346 etc.)
  for (int j = 0; j < n_ops; ++j) {
    const qsb_op& o = ops[j];
    switch (o.opcode) {
      case QSB_OP_RX: case QSB_OP_RY: case QSB_OP_RZ: case QSB_OP_RXX: case QSB_OP_RYY: case QSB_OP_RZZ: pscale[o.param_off] = 0.5; break;
      case QSB_OP_U3: pscale[o.param_off] = 0.5; break;
      case QSB_OP_XZXZ: pscale[o.param_off] = pscale[o.param_off + 1] = 0.5; break;
      case QSB_OP_ZXZXZ: pscale[o.param_off] = pscale[o.param_off + 1] = pscale[o.param_off + 2] = 0.5; break;
      default: break;
    }
  }
  // gate-table slots: op j owns slots [reserved, reserved + 1 + np): its 2x2 and the np parameter derivatives
  std::vector<int32_t> slotmap((size_t)std::max(1, n_ops + n_params), -1);
  {
    int used = 0;
    for (int j = 0; j < n_ops; ++j) {
      qsb_op& o = p->ops[j];
      const int np = op_params(o.opcode);
      o.reserved = j + used;
      if (o.opcode != QSB_OP_CNOT && o.opcode != QSB_OP_CONSTK && o.opcode != QSB_OP_RXX && o.opcode != QSB_OP_RYY && o.opcode != QSB_OP_RZZ &&
          o.opcode != QSB_OP_QUTRIT1)
        for (int w = 0; w <= np; ++w) slotmap[o.reserved + w] = (j << 2) | w;
      used += np;
    }
  }
  ops = p->ops.data();
  // on any CUDA failure: release what the devices visited so far hold, and the host object (no leak, no half-built handle)
  auto fail = [&](cudaError_t e, const char* what) {
    const int rc = qsb_cuda_fail(e, what);
    const std::string keep = qsb_get_error();
    qsb_program_destroy(ctx, p);
    qsb_set_error(keep);
    return rc;
  };
#define QSB_P(call) do { cudaError_t _e = (call); if (_e != cudaSuccess) return fail(_e, #call); } while (0)
  for (auto& d : ctx->devs) {
    QSB_P(cudaSetDevice(d.id));
    p->d_ops.push_back(nullptr); p->d_consts.push_back(nullptr); p->d_pscale.push_back(nullptr); p->d_slotmap.push_back(nullptr);
    QSB_P(cudaMalloc((void**)&p->d_ops.back(), sizeof(qsb_op) * std::max(1, n_ops)));
    QSB_P(cudaMalloc((void**)&p->d_consts.back(), sizeof(double) * std::max<size_t>(1, const_len)));
    QSB_P(cudaMalloc((void**)&p->d_pscale.back(), sizeof(double) * pscale.size()));
    QSB_P(cudaMalloc((void**)&p->d_slotmap.back(), sizeof(int32_t) * slotmap.size()));
    if (n_ops) QSB_P(cudaMemcpy(p->d_ops.back(), ops, sizeof(qsb_op) * n_ops, cudaMemcpyHostToDevice));
    if (const_len) QSB_P(cudaMemcpy(p->d_consts.back(), const_pool, sizeof(double) * const_len, cudaMemcpyHostToDevice));
    QSB_P(cudaMemcpy(p->d_pscale.back(), pscale.data(), sizeof(double) * pscale.size(), cudaMemcpyHostToDevice));
    QSB_P(cudaMemcpy(p->d_slotmap.back(), slotmap.data(), sizeof(int32_t) * slotmap.size(), cudaMemcpyHostToDevice));
  }
#undef QSB_P
  *out_prog = p;
  return QSB_OK;
}

extern "C" int qsb_program_create(qsb_ctx* ctx, int n_qubits, int n_ops, const qsb_op* ops, int n_params,
                                  const double* const_pool, size_t const_len, qsb_program** out_prog) {
  return program_create(ctx, 2, n_qubits, n_ops, ops, n_params, const_pool, const_len, out_prog);
}
extern "C" int qsb_program_create_qudits(qsb_ctx* ctx, int radix, int n_qudits, int n_ops, const qsb_op* ops, int n_params,
                                         const double* const_pool, size_t const_len, qsb_program** out_prog) {
  return program_create(ctx, radix, n_qudits, n_ops, ops, n_params, const_pool, const_len, out_prog);
}

extern "C" int qsb_program_destroy(qsb_ctx* ctx, qsb_program* prog) {
  if (!prog) return QSB_OK;
  if (ctx) {
    for (size_t s = 0; s < ctx->devs.size() && s < prog->d_ops.size(); ++s) {
      cudaSetDevice(ctx->devs[s].id);
      cudaFree(prog->d_ops[s]);
      cudaFree(prog->d_consts[s]);
      cudaFree(prog->d_pscale[s]);
      cudaFree(prog->d_slotmap[s]);
    }
  }
  delete prog;
  return QSB_OK;
}
extern "C" int qsb_program_num_params(const qsb_program* p) { return p ? p->n_params : -1; }
extern "C" int qsb_program_num_qubits(const qsb_program* p) { return p ? p->n_qubits : -1; }
extern "C" int qsb_program_radix(const qsb_program* p) { return p ? p->radix : -1; }

// max |T T^dagger - I| <= 1e-9 (N <= 32: a few thousand flops): the body-frame LM forms (qsb_lm3.cuh) need a unitary target
bool qsb_target_is_unitary(const double* T, int N) {
  for (int i = 0; i < N; ++i)
    for (int j = 0; j <= i; ++j) {
      double re = (i == j) ? -1.0 : 0.0, im = 0.0;
      for (int k = 0; k < N; ++k) {
        const double ar = T[2 * (i * N + k)], ai = T[2 * (i * N + k) + 1], br = T[2 * (j * N + k)], bi = T[2 * (j * N + k) + 1];
        re += ar * br + ai * bi; im += ai * br - ar * bi;
      }
      if (!(fabs(re) <= 1e-9) || !(fabs(im) <= 1e-9)) return false;
    }
  return true;
}

// ---------------------------------------------------------------------------- device-resident batches
static size_t pool_round(size_t bytes) {
  if (bytes < 256) return 256;
  if (bytes <= ((size_t)1 << 20)) return (bytes + 255) & ~(size_t)255;
  return (bytes + (((size_t)1 << 21) - 1)) & ~(((size_t)1 << 21) - 1);  // 2 MiB granules above 1 MiB
}
static const size_t QSB_POOL_LIMIT = (size_t)8 << 30;  // cached (idle) bytes per device

int qsb_pool_alloc(const QsbDevice& d, void** out, size_t bytes) {
  const size_t want = pool_round(bytes);
  {
    std::lock_guard<std::mutex> g(d.pool->mu);
    auto& fb = d.pool->free_blocks;
    for (size_t i = 0; i < fb.size(); ++i)
      if (fb[i].first == want) {
        *out = fb[i].second;
        d.pool->cached_bytes -= want;
        fb[i] = fb.back(); fb.pop_back();
        return QSB_OK;
      }
  }
  cudaError_t e = cudaMalloc(out, want);
  if (e != cudaSuccess) {  // give the cache back and retry once
    cudaGetLastError();
    {
      std::lock_guard<std::mutex> g(d.pool->mu);
      for (auto& b : d.pool->free_blocks) cudaFree(b.second);
      d.pool->free_blocks.clear(); d.pool->cached_bytes = 0;
    }
    e = cudaMalloc(out, want);
  }
  if (e != cudaSuccess) {
    *out = nullptr;
    return qsb_fail(QSB_ERR_NOMEM, std::string("device allocation of ") + std::to_string(want) + " bytes failed: " + cudaGetErrorString(e));
  }
  return QSB_OK;
}

void qsb_pool_free(const QsbDevice& d, void* p, size_t bytes) {
  if (!p) return;
  const size_t have = pool_round(bytes);
  std::lock_guard<std::mutex> g(d.pool->mu);
  if (d.pool->cached_bytes + have > QSB_POOL_LIMIT) { cudaFree(p); return; }
  d.pool->free_blocks.emplace_back(have, p);
  d.pool->cached_bytes += have;
}

int qsb_pool_reserve(const QsbDevice& d, QsbBuf& b, size_t bytes) {
  if (bytes <= b.bytes) return QSB_OK;
  if (b.p) { qsb_pool_free(d, b.p, b.bytes); b.p = nullptr; b.bytes = 0; }
  int rc = qsb_pool_alloc(d, &b.p, bytes);
  if (rc) return rc;
  b.bytes = bytes;
  return QSB_OK;
}

int qsb_buf_reserve(QsbBuf& b, size_t bytes) {
  if (bytes <= b.bytes) return QSB_OK;
  // grow by at least half: a search asks for a little more with every layer it adds (the solver scratch is proportional to P),
  // and a cudaFree + cudaMalloc of several hundred MB per frontier expansion is not free
  bytes = std::max(bytes, b.bytes + b.bytes / 2);
  bytes = (bytes + ((size_t)2 << 20) - 1) & ~(((size_t)2 << 20) - 1);
  if (b.p) { cudaFree(b.p); b.p = nullptr; b.bytes = 0; }
  cudaError_t e = cudaMalloc(&b.p, bytes);
  if (e != cudaSuccess) {
    b.p = nullptr;
    return qsb_fail(QSB_ERR_NOMEM, std::string("device allocation of ") + std::to_string(bytes) + " bytes failed: " + cudaGetErrorString(e));
  }
  b.bytes = bytes;
  return QSB_OK;
}

extern "C" int qsb_batch_create(qsb_ctx* ctx, int slot, int n_programs, const qsb_program* const* progs,
                                int n_instances, const int32_t* prog_of_instance, qsb_batch** out_batch) {
  if (!ctx || !progs || !out_batch || (n_instances > 0 && !prog_of_instance)) return qsb_fail(QSB_ERR_INVALID, "NULL argument");
  if (slot < 0 || slot >= (int)ctx->devs.size()) return qsb_fail(QSB_ERR_INVALID, "device slot out of range");
  if (n_programs <= 0 || n_instances < 0) return qsb_fail(QSB_ERR_INVALID, "empty program list");
  for (int p = 0; p < n_programs; ++p) {
    if (!progs[p]) return qsb_fail(QSB_ERR_INVALID, "NULL program");
    if (progs[p]->n_qubits != progs[0]->n_qubits || progs[p]->radix != progs[0]->radix)
      return qsb_fail(QSB_ERR_INVALID, "all programs of a batch must act on the same number of qubits");
    if ((int)progs[p]->d_ops.size() <= slot) return qsb_fail(QSB_ERR_INVALID, "program was created on another context");
  }
  qsb_batch* b = new (std::nothrow) qsb_batch();
  if (!b) return qsb_fail(QSB_ERR_NOMEM, "out of host memory");
  b->ctx = ctx; b->slot = slot; b->n_qubits = progs[0]->n_qubits; b->n_instances = n_instances; b->n_programs = n_programs;
  b->radix = progs[0]->radix; b->dim = progs[0]->dim;
  b->fast_ok = true;
  b->force_generic = ctx->force_generic;
  std::vector<QsbDevProgram> dp(n_programs);
  for (int p = 0; p < n_programs; ++p) {
    dp[p].ops = progs[p]->d_ops[slot]; dp[p].consts = progs[p]->d_consts[slot]; dp[p].pscale = progs[p]->d_pscale[slot]; dp[p].slotmap = progs[p]->d_slotmap[slot];
    dp[p].n_ops = progs[p]->n_ops; dp[p].n_params = progs[p]->n_params;
    b->max_ops = std::max(b->max_ops, progs[p]->n_ops);
    b->max_params = std::max(b->max_params, progs[p]->n_params);
    b->max_pops = std::max(b->max_pops, qsb_count_param_ops(progs[p]->ops.data(), progs[p]->n_ops));
    b->fast_ok = b->fast_ok && progs[p]->fast_ok;
  }
  b->param_off.resize((size_t)n_instances + 1);
  b->prog_idx.assign(prog_of_instance, prog_of_instance + n_instances);
  long long off = 0;
  for (int i = 0; i < n_instances; ++i) {
    int pi = prog_of_instance[i];
    if (pi < 0 || pi >= n_programs) { delete b; return qsb_fail(QSB_ERR_INVALID, "prog_of_instance out of range"); }
    b->param_off[i] = off;
    off += progs[pi]->n_params;
  }
  b->param_off[n_instances] = off;
  const QsbDevice& d = ctx->devs[slot];
  int rc = QSB_OK;
  auto fail = [&](int code) { qsb_batch_destroy(b); return code; };
  if (cudaSetDevice(d.id) != cudaSuccess) return fail(qsb_fail(QSB_ERR_CUDA, "cudaSetDevice failed"));
#define QSB_B(call) do { cudaError_t _e = (call); if (_e != cudaSuccess) { rc = qsb_cuda_fail(_e, #call); return fail(rc); } } while (0)
  b->sz_progs = sizeof(QsbDevProgram) * n_programs;
  b->sz_prog_idx = sizeof(int32_t) * std::max(1, n_instances);
  b->sz_param_off = sizeof(long long) * ((size_t)n_instances + 1);
  b->sz_params = sizeof(double) * std::max<long long>(1, off);
  b->sz_target = sizeof(double) * 2 * b->dim * b->dim;
  if ((rc = qsb_pool_alloc(d, (void**)&b->d_progs, b->sz_progs)) || (rc = qsb_pool_alloc(d, (void**)&b->d_prog_idx, b->sz_prog_idx)) ||
      (rc = qsb_pool_alloc(d, (void**)&b->d_param_off, b->sz_param_off)) || (rc = qsb_pool_alloc(d, (void**)&b->d_params, b->sz_params)) ||
      (rc = qsb_pool_alloc(d, (void**)&b->d_target, b->sz_target)))
    return fail(rc);
  QSB_B(cudaMemcpy(b->d_progs, dp.data(), sizeof(QsbDevProgram) * n_programs, cudaMemcpyHostToDevice));
  if (n_instances) QSB_B(cudaMemcpy(b->d_prog_idx, prog_of_instance, sizeof(int32_t) * n_instances, cudaMemcpyHostToDevice));
  QSB_B(cudaMemcpy(b->d_param_off, b->param_off.data(), sizeof(long long) * ((size_t)n_instances + 1), cudaMemcpyHostToDevice));
#undef QSB_B
  *out_batch = b;
  return QSB_OK;
}

extern "C" int qsb_batch_destroy(qsb_batch* b) {
  if (!b) return QSB_OK;
  cudaSetDevice(b->ctx->devs[b->slot].id);
  cudaStreamSynchronize(b->ctx->devs[b->slot].stream);
  const QsbDevice& d = b->ctx->devs[b->slot];  // the stream is idle: the blocks can be handed to the next batch
  qsb_pool_free(d, b->d_progs, b->sz_progs); qsb_pool_free(d, b->d_prog_idx, b->sz_prog_idx); qsb_pool_free(d, b->d_param_off, b->sz_param_off);
  qsb_pool_free(d, b->d_params, b->sz_params); qsb_pool_free(d, b->d_target, b->sz_target);
  qsb_pool_free(d, b->out_a.p, b->out_a.bytes); qsb_pool_free(d, b->out_b.p, b->out_b.bytes); qsb_pool_free(d, b->out_c.p, b->out_c.bytes);
  cudaFree(b->d_sq_off); cudaFree(b->solver_scratch);
  if (b->ev0) { cudaEventDestroy(b->ev0); cudaEventDestroy(b->ev1); }
  delete b;
  return QSB_OK;
}

extern "C" int qsb_batch_set_kernel(qsb_batch* b, int kind) {
  if (!b || kind < 0 || kind > 1) return qsb_fail(QSB_ERR_INVALID, "bad argument");
  b->force_generic = (kind == QSB_KERNEL_GENERIC);
  return QSB_OK;
}

extern "C" int qsb_batch_set_params(qsb_batch* b, const double* params, size_t n_doubles) {
  if (!b || (!params && n_doubles)) return qsb_fail(QSB_ERR_INVALID, "NULL argument");
  if ((long long)n_doubles != b->param_off[b->n_instances])
    return qsb_fail(QSB_ERR_INVALID, "parameter count mismatch: expected " + std::to_string(b->param_off[b->n_instances]));
  const QsbDevice& d = b->ctx->devs[b->slot];
  QSB_CUDA(cudaSetDevice(d.id));
  if (n_doubles) QSB_CUDA(cudaMemcpyAsync(b->d_params, params, sizeof(double) * n_doubles, cudaMemcpyHostToDevice, d.stream));
  QSB_CUDA(cudaStreamSynchronize(d.stream));
  return QSB_OK;
}

extern "C" int qsb_batch_set_target(qsb_batch* b, const double* target) {
  if (!b || !target) return qsb_fail(QSB_ERR_INVALID, "NULL argument");
  const QsbDevice& d = b->ctx->devs[b->slot];
  QSB_CUDA(cudaSetDevice(d.id));
  QSB_CUDA(cudaMemcpyAsync(b->d_target, target, sizeof(double) * 2 * b->dim * b->dim, cudaMemcpyHostToDevice, d.stream));
  QSB_CUDA(cudaStreamSynchronize(d.stream));
  b->target_unitary = qsb_target_is_unitary(target, (int)b->dim);
  return QSB_OK;
}

// sizes (in doubles) of the three output buffers of a run over [first, first+count)
static void out_sizes(const qsb_batch* b, int what, int first, int count, size_t* a, size_t* bb, size_t* c) {
  const size_t NN2 = 2 * b->dim * b->dim;
  const size_t np = (size_t)(b->param_off[first + count] - b->param_off[first]);
  *a = *bb = *c = 0;
  switch (what) {
    case QSB_RUN_MATRIX: *a = count * NN2; break;
    case QSB_RUN_DIST: *a = count; break;
    case QSB_RUN_MATJAC: case QSB_RUN_RESID: *a = count * NN2; *bb = np * NN2; break;
    case QSB_RUN_HS: *a = count; *bb = np; break;
    case QSB_RUN_LMSYS: {
      *a = count; *bb = np;
      size_t sq = 0;
      for (int i = first; i < first + count; ++i) { size_t p = (size_t)(b->param_off[i + 1] - b->param_off[i]); sq += p * p; }
      *c = sq;
    } break;
    default: break;
  }
}

extern "C" int qsb_batch_output_bytes(qsb_batch* b, int what, int count, uint64_t* out_bytes) {
  if (!b || !out_bytes || count < 0 || count > b->n_instances) return qsb_fail(QSB_ERR_INVALID, "bad argument");
  size_t a, bb, c;
  out_sizes(b, what, 0, count, &a, &bb, &c);
  *out_bytes = (uint64_t)(a + bb + c) * sizeof(double);
  return QSB_OK;
}

int qsb_batch_enqueue(qsb_batch* b, int what, int first, int count);  // defined below (kernel selection)

extern "C" int qsb_batch_run(qsb_batch* b, int what, int first, int count, float* elapsed_ms) {
  if (!b) return qsb_fail(QSB_ERR_INVALID, "NULL batch");
  if (first < 0 || count < 0 || first + count > b->n_instances) return qsb_fail(QSB_ERR_INVALID, "instance range out of bounds");
  const QsbDevice& d = b->ctx->devs[b->slot];
  QSB_CUDA(cudaSetDevice(d.id));
  size_t a, bb, c;
  out_sizes(b, what, first, count, &a, &bb, &c);
  int rc;
  if ((rc = qsb_pool_reserve(d, b->out_a, a * sizeof(double)))) return rc;
  if ((rc = qsb_pool_reserve(d, b->out_b, bb * sizeof(double)))) return rc;
  if ((rc = qsb_pool_reserve(d, b->out_c, c * sizeof(double)))) return rc;
  b->last_first = first; b->last_count = count; b->last_what = what;
  if (count == 0) { if (elapsed_ms) *elapsed_ms = 0.f; return QSB_OK; }
  if (elapsed_ms) {
    if (!b->ev0) { QSB_CUDA(cudaEventCreate(&b->ev0)); QSB_CUDA(cudaEventCreate(&b->ev1)); }
    QSB_CUDA(cudaEventRecord(b->ev0, d.stream));
  }
  if ((rc = qsb_batch_enqueue(b, what, first, count))) return rc;
  if (elapsed_ms) {
    QSB_CUDA(cudaEventRecord(b->ev1, d.stream));
    QSB_CUDA(cudaEventSynchronize(b->ev1));
    QSB_CUDA(cudaEventElapsedTime(elapsed_ms, b->ev0, b->ev1));
    QSB_CUDA(cudaGetLastError());
  }
  return QSB_OK;
}

extern "C" int qsb_batch_sync(qsb_batch* b) {
  if (!b) return qsb_fail(QSB_ERR_INVALID, "NULL batch");
  const QsbDevice& d = b->ctx->devs[b->slot];
  QSB_CUDA(cudaSetDevice(d.id));
  QSB_CUDA(cudaStreamSynchronize(d.stream));
  QSB_CUDA(cudaGetLastError());
  return QSB_OK;
}

extern "C" int qsb_batch_fetch(qsb_batch* b, int what, double* a, double* bb, double* c) {
  if (!b) return qsb_fail(QSB_ERR_INVALID, "NULL batch");
  if (what != b->last_what) return qsb_fail(QSB_ERR_INVALID, "fetch does not match the last run of this batch");
  const QsbDevice& d = b->ctx->devs[b->slot];
  QSB_CUDA(cudaSetDevice(d.id));
  size_t na, nb, nc;
  out_sizes(b, what, b->last_first, b->last_count, &na, &nb, &nc);
  if (a && na) QSB_CUDA(cudaMemcpyAsync(a, b->out_a.p, na * sizeof(double), cudaMemcpyDeviceToHost, d.stream));
  if (bb && nb) QSB_CUDA(cudaMemcpyAsync(bb, b->out_b.p, nb * sizeof(double), cudaMemcpyDeviceToHost, d.stream));
  if (c && nc) QSB_CUDA(cudaMemcpyAsync(c, b->out_c.p, nc * sizeof(double), cudaMemcpyDeviceToHost, d.stream));
  QSB_CUDA(cudaStreamSynchronize(d.stream));
  return QSB_OK;
}

int qsb_launch_lmsys(qsb_batch* b, const QsbBatchArgs& args, cudaStream_t st, int sm_count);  // qsb_lmsys.cu

int qsb_batch_enqueue(qsb_batch* b, int what, int first, int count) {
  const QsbDevice& d = b->ctx->devs[b->slot];
  QsbBatchArgs args;
  args.progs = b->d_progs; args.prog_idx = b->d_prog_idx; args.param_off = b->d_param_off; args.params = b->d_params;
  args.target = b->d_target; args.first = first; args.count = count; args.max_ops = b->max_ops;
  args.max_params = std::max(1, b->max_params);
  args.out_a = (double*)b->out_a.p; args.out_b = (double*)b->out_b.p; args.out_c = (double*)b->out_c.p;
  int rc = (what == QSB_RUN_LMSYS) ? qsb_launch_lmsys(b, args, d.stream, d.sm_count) : qsb_launch(b, what, args, d.stream, d.sm_count);
  if (rc == QSB_OK) {
    std::lock_guard<std::mutex> g(b->ctx->mu);
    b->ctx->launches += 1;
  }
  return rc;
}

// ---------------------------------------------------------------------------- host-buffer evaluation calls
// Instances are split into one contiguous range per device of the context; each range is processed by its own
// host thread in chunks whose device outputs stay under QSB_CHUNK_BYTES, so arbitrarily large batches stream
// through a bounded device footprint.  No collective: the units are independent.
static const size_t QSB_CHUNK_BYTES = (size_t)1 << 30;

struct HostEvalJob {
  qsb_ctx* ctx; int n_programs; const qsb_program* const* progs; int n_instances; const int32_t* prog_of_instance;
  const double* params; const double* target; int what;
  double *a, *b, *c;
};

static int run_shard(const HostEvalJob& job, int slot, int i0, int i1, const std::vector<long long>& poff) {
  if (i1 <= i0) return QSB_OK;
  // two batches = two sets of device output buffers: the kernel of chunk c+1 runs while chunk c is copied to the host
  qsb_batch* batch[2] = {nullptr, nullptr};
  cudaEvent_t ev_kernel[2] = {nullptr, nullptr}, ev_copy[2] = {nullptr, nullptr};
  const QsbDevice& d = job.ctx->devs[slot];
  auto done = [&](int code) {
    std::string keep = g_err;
    cudaStreamSynchronize(d.copy_stream);
    for (int k = 0; k < 2; ++k) {
      if (batch[k]) qsb_batch_destroy(batch[k]);
      if (ev_kernel[k]) cudaEventDestroy(ev_kernel[k]);
      if (ev_copy[k]) cudaEventDestroy(ev_copy[k]);
    }
    g_err = keep;
    return code;
  };
  int rc;
  const int n = i1 - i0;
  // chunk list (output bytes per chunk bounded; always at least one instance per chunk)
  std::vector<int> starts;
  {
    qsb_batch probe;  // only param_off / the dimension are needed by out_sizes
    probe.n_qubits = job.progs[0]->n_qubits; probe.radix = job.progs[0]->radix; probe.dim = job.progs[0]->dim;
    probe.param_off.assign(poff.begin() + i0, poff.begin() + i1 + 1);
    size_t sa, sb, sc;
    out_sizes(&probe, job.what, 0, n, &sa, &sb, &sc);
    const size_t total = (sa + sb + sc) * sizeof(double);
    size_t budget = std::min(QSB_CHUNK_BYTES, std::max<size_t>((size_t)32 << 20, total / 8));
    if (const char* e = getenv("QSB_CHUNK_BYTES")) { long long v = atoll(e); if (v > 0) budget = (size_t)v; }  // tests
    int first = 0;
    while (first < n) {
      int count = 0;
      size_t bytes = 0;
      while (first + count < n) {
        out_sizes(&probe, job.what, first + count, 1, &sa, &sb, &sc);
        const size_t add = (sa + sb + sc) * sizeof(double);
        if (count > 0 && bytes + add > budget) break;
        bytes += add; ++count;
      }
      starts.push_back(first);
      first += count;
    }
    starts.push_back(n);
  }
  const int n_chunks = (int)starts.size() - 1;
  const int n_sets = n_chunks > 1 ? 2 : 1;
  for (int k = 0; k < n_sets; ++k) {
    if ((rc = qsb_batch_create(job.ctx, slot, job.n_programs, job.progs, n, job.prog_of_instance + i0, &batch[k]))) return done(rc);
    if ((rc = qsb_batch_set_params(batch[k], job.params + poff[i0], (size_t)(poff[i1] - poff[i0])))) return done(rc);
    if (job.target && (rc = qsb_batch_set_target(batch[k], job.target))) return done(rc);
    if (cudaEventCreateWithFlags(&ev_kernel[k], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&ev_copy[k], cudaEventDisableTiming) != cudaSuccess)
      return done(qsb_fail(QSB_ERR_CUDA, "cudaEventCreate failed"));
  }
  const size_t NN2 = 2 * batch[0]->dim * batch[0]->dim;
#define QSB_R(call) do { cudaError_t _e = (call); if (_e != cudaSuccess) return done(qsb_cuda_fail(_e, #call)); } while (0)
  auto launch = [&](int c) -> int {
    const int k = c % n_sets;
    if (c >= n_sets) { cudaError_t e = cudaStreamWaitEvent(d.stream, ev_copy[k], 0); if (e != cudaSuccess) return qsb_cuda_fail(e, "cudaStreamWaitEvent"); }
    int r = qsb_batch_run(batch[k], job.what, starts[c], starts[c + 1] - starts[c], nullptr);
    if (r) return r;
    cudaError_t e = cudaEventRecord(ev_kernel[k], d.stream);
    return e == cudaSuccess ? QSB_OK : qsb_cuda_fail(e, "cudaEventRecord");
  };
  if ((rc = launch(0))) return done(rc);
  for (int c = 0; c < n_chunks; ++c) {
    if (c + 1 < n_chunks && (rc = launch(c + 1))) return done(rc);
    const int k = c % n_sets, first = starts[c], count = starts[c + 1] - starts[c];
    const long long g0 = i0 + first;  // global instance index of the chunk start
    const long long p0 = poff[g0];
    double *pa = nullptr, *pb = nullptr;
    switch (job.what) {
      case QSB_RUN_MATRIX: pa = job.a + g0 * NN2; break;
      case QSB_RUN_DIST: pa = job.a + g0; break;
      case QSB_RUN_MATJAC: case QSB_RUN_RESID: pa = job.a + g0 * NN2; pb = job.b ? job.b + p0 * NN2 : nullptr; break;
      case QSB_RUN_HS: pa = job.a + g0; pb = job.b + p0; break;
      default: break;
    }
    size_t na, nb, nc;
    out_sizes(batch[k], job.what, first, count, &na, &nb, &nc);
    QSB_R(cudaStreamWaitEvent(d.copy_stream, ev_kernel[k], 0));
    if (pa && na) QSB_R(cudaMemcpyAsync(pa, batch[k]->out_a.p, na * sizeof(double), cudaMemcpyDeviceToHost, d.copy_stream));
    if (pb && nb) QSB_R(cudaMemcpyAsync(pb, batch[k]->out_b.p, nb * sizeof(double), cudaMemcpyDeviceToHost, d.copy_stream));
    QSB_R(cudaEventRecord(ev_copy[k], d.copy_stream));
  }
  QSB_R(cudaStreamSynchronize(d.copy_stream));
  QSB_R(cudaStreamSynchronize(d.stream));
  QSB_R(cudaGetLastError());
#undef QSB_R
  return done(QSB_OK);
}

static int host_eval(const HostEvalJob& job) {
  if (!job.ctx || !job.progs || (job.n_instances > 0 && (!job.prog_of_instance || !job.a)))
    return qsb_fail(QSB_ERR_INVALID, "NULL argument");
  if (job.n_instances < 0 || job.n_programs <= 0) return qsb_fail(QSB_ERR_INVALID, "bad batch size");
  if (job.n_instances == 0) return QSB_OK;
  std::vector<long long> poff((size_t)job.n_instances + 1);
  long long off = 0;
  for (int i = 0; i < job.n_instances; ++i) {
    int pi = job.prog_of_instance[i];
    if (pi < 0 || pi >= job.n_programs || !job.progs[pi]) return qsb_fail(QSB_ERR_INVALID, "prog_of_instance out of range");
    poff[i] = off;
    off += job.progs[pi]->n_params;
  }
  poff[job.n_instances] = off;
  if (off > 0 && !job.params) return qsb_fail(QSB_ERR_INVALID, "params is NULL");
  const int nd = (int)job.ctx->devs.size();
  const int used = std::min(nd, job.n_instances);
  if (used == 1) return run_shard(job, 0, 0, job.n_instances, poff);
  std::vector<int> rcs(used, QSB_OK);
  std::vector<std::string> errs(used);
  std::vector<std::thread> th;
  for (int s = 0; s < used; ++s) {
    int i0 = (int)((long long)job.n_instances * s / used), i1 = (int)((long long)job.n_instances * (s + 1) / used);
    th.emplace_back([&, s, i0, i1] { rcs[s] = run_shard(job, s, i0, i1, poff); if (rcs[s]) errs[s] = g_err; });
  }
  for (auto& t : th) t.join();
  for (int s = 0; s < used; ++s) if (rcs[s]) return qsb_fail(rcs[s], errs[s]);
  return QSB_OK;
}

extern "C" int qsb_eval_matrix(qsb_ctx* ctx, int n_programs, const qsb_program* const* progs, int n_instances,
                               const int32_t* prog_of_instance, const double* params, double* out_U) {
  HostEvalJob j{ctx, n_programs, progs, n_instances, prog_of_instance, params, nullptr, QSB_RUN_MATRIX, out_U, nullptr, nullptr};
  return host_eval(j);
}
extern "C" int qsb_eval_mat_jac(qsb_ctx* ctx, int n_programs, const qsb_program* const* progs, int n_instances,
                                const int32_t* prog_of_instance, const double* params, double* out_U, double* out_J) {
  HostEvalJob j{ctx, n_programs, progs, n_instances, prog_of_instance, params, nullptr, QSB_RUN_MATJAC, out_U, out_J, nullptr};
  return host_eval(j);
}
extern "C" int qsb_eval_hs_grad(qsb_ctx* ctx, int n_programs, const qsb_program* const* progs, int n_instances,
                                const int32_t* prog_of_instance, const double* params, const double* target,
                                double* out_f, double* out_grad) {
  if (!target || !out_grad) return qsb_fail(QSB_ERR_INVALID, "NULL argument");
  HostEvalJob j{ctx, n_programs, progs, n_instances, prog_of_instance, params, target, QSB_RUN_HS, out_f, out_grad, nullptr};
  return host_eval(j);
}
extern "C" int qsb_eval_distance(qsb_ctx* ctx, int n_programs, const qsb_program* const* progs, int n_instances,
                                 const int32_t* prog_of_instance, const double* params, const double* target, double* out_f) {
  if (!target) return qsb_fail(QSB_ERR_INVALID, "NULL argument");
  HostEvalJob j{ctx, n_programs, progs, n_instances, prog_of_instance, params, target, QSB_RUN_DIST, out_f, nullptr, nullptr};
  return host_eval(j);
}
extern "C" int qsb_eval_residuals_jac(qsb_ctx* ctx, int n_programs, const qsb_program* const* progs, int n_instances,
                                      const int32_t* prog_of_instance, const double* params, const double* target,
                                      double* out_r, double* out_Jr) {
  if (!target) return qsb_fail(QSB_ERR_INVALID, "NULL argument");
  HostEvalJob j{ctx, n_programs, progs, n_instances, prog_of_instance, params, target, QSB_RUN_RESID, out_r, out_Jr, nullptr};
  return host_eval(j);
}

