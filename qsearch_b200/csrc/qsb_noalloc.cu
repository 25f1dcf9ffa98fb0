// qsb_eval_lm_system_static: the LM-system evaluation (r^T r, J_r^T r, J_r^T J_r per instance) for callers that must not
// allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
// No counterpart in the reference.
//
// It is a variant of qsb_k_lmsys3 (qsb_lm3.cuh) whose whole device working set - staged inputs, staged outputs, the
// per-CTA slab / packed Gram matrix / snapshots - lives in ONE file-scope static __device__ array of
// QSB_STATIC_SCRATCH_BYTES (64 MiB; every GPU that loads the library holds it until the process exits).  The entry point
// calls no cudaMalloc, cudaFree or allocating library routine: only cudaMemcpyAsync, kernel launches and a stream
// synchronise on the context's existing stream.
//
// In host code the array's NAME is a host-side stand-in, not a device address.  The device address is obtained once per
// device, with that device current, by cudaGetSymbolAddress (static_scratch_of below) and passed to the kernel.
#include <algorithm>
#include <cstring>
#include <mutex>

#include "qsb_host.h"
#include "qsb_lm3.cuh"
#include "qsb_launch.h"

#define QSB_STATIC_IO_BYTES ((size_t)16 << 20)  // staged inputs and outputs of one call
static __device__ __align__(256) unsigned char qsb_static_scratch[QSB_STATIC_SCRATCH_BYTES];

// device address of the array per CUDA device ordinal, resolved on first use with that device current
#define QSB_STATIC_MAX_DEVICES 64
static std::mutex g_static_mu[QSB_STATIC_MAX_DEVICES];  // one call at a time per device: every call shares the array
static unsigned char* g_static_ptr[QSB_STATIC_MAX_DEVICES];
static std::mutex g_static_setup_mu;

static int static_scratch_of(int device_id, unsigned char** out) {
  if (device_id < 0 || device_id >= QSB_STATIC_MAX_DEVICES) return qsb_fail(QSB_ERR_UNSUPPORTED, "device ordinal beyond QSB_STATIC_MAX_DEVICES");
  std::lock_guard<std::mutex> g(g_static_setup_mu);
  if (!g_static_ptr[device_id]) {
    void* p = nullptr;
    QSB_CUDA(cudaSetDevice(device_id));
    QSB_CUDA(cudaGetSymbolAddress(&p, qsb_static_scratch));
    if (!p) return qsb_fail(QSB_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the static scratch array");
    g_static_ptr[device_id] = (unsigned char*)p;
  }
  *out = g_static_ptr[device_id];
  return QSB_OK;
}

// One program, uniform P: params [count][P], outputs out_rr [count], out_g [count][P], out_A [count][P][P], all inside the
// static array (the kernel never dereferences a caller pointer).  Same phases as qsb_k_lmsys3.
template <int NQ>
QSB_GLOBAL __launch_bounds__(QsbQGeom<NQ>::WARPS * 32, (NQ <= 4 ? 2 : 1))
qsb_k_lmsys3_static(QsbDevProgram prog, int count, const double* params, const double* target, double* out_rr, double* out_g,
                    double* out_A, double* scratch, size_t scratch_per_cta) {
  using G = QsbQGeom<NQ>;
  constexpr int N = G::N, K2 = G::K2, NW = G::WARPS;
  QSB_DYN_SMEM(smem_raw);
  const int P = prog.n_params, rows = P + 1;
  double* trig = (double*)smem_raw;                       // 2 * P
  double* part = trig + 2 * P;                            // QSB_ROWS_PART
  cplx* gtab = (cplx*)(part + QSB_ROWS_PART);
  cplx* vsm = gtab + (size_t)4 * (prog.n_ops + P);        // N * N
  int* seg = (int*)(vsm + N * N);                         // 16
  qsb_op* sops = (qsb_op*)(seg + 16);
  const int lane = threadIdx.x & 31, warp = qsb_warp_index();
  double* Rg = scratch + (size_t)blockIdx.x * scratch_per_cta;
  double* A = Rg + (size_t)rows * K2;
  double* snap = A + (((size_t)rows * (rows + 1) / 2 + 1) & ~(size_t)1);
  for (int e = threadIdx.x; e < prog.n_ops * 6; e += blockDim.x) ((int32_t*)sops)[e] = ((const int32_t*)prog.ops)[e];
  prog.ops = sops;
  __syncthreads();
  for (int inst = blockIdx.x; inst < count; inst += gridDim.x) {
    __syncthreads();
    const double F = qsb_cta_eval_rows<NQ>(prog, params + (size_t)inst * P, target, trig, gtab, part, snap, vsm, nullptr);
    qsb_cta_qrows_snap<NQ>(prog, gtab, snap, vsm, Rg);
    // the array keeps the previous call's contents: whatever is accumulated into starts from an explicit zero
    if (qsb_syrk_ksplit3(rows, NW) == 2) { for (int e = threadIdx.x; e < rows * (rows + 1) / 2; e += blockDim.x) A[e] = 0.0; __syncthreads(); }
    qsb_syrk_packed(Rg, rows, K2, A, warp, NW, lane);
    __syncthreads();
    double* og = out_g + (size_t)inst * P;
    double* oA = out_A + (size_t)inst * P * P;
    for (int k = threadIdx.x; k < P; k += blockDim.x) og[k] = QSB_LDCG(&A[QSB_TRI(P, k)]);
    for (int e = threadIdx.x; e < P * P; e += blockDim.x) {
      const int k = e / P, l = e - k * P;
      oA[e] = QSB_LDCG(&A[k >= l ? QSB_TRI(k, l) : QSB_TRI(l, k)]);
    }
    if (threadIdx.x == 0) out_rr[inst] = 2.0 * F;
  }
}

template <int NQ>
static int run_static(const QsbDevice& d, const qsb_program* prog, int slot, int n, const double* params, const double* target,
                      double* out_rr, double* out_g, double* out_A, unsigned char* base) {
  const size_t NN = (size_t)1 << (2 * NQ), P = (size_t)prog->n_params, rows = P + 1;
  const size_t smem = qsb_lmsys3_smem_bytes(NQ, prog->n_ops, (int)P);
  if (smem > 200 * 1024) return qsb_fail(QSB_ERR_UNSUPPORTED, "program too long for qsb_eval_lm_system_static");
  // staging region: params, target, then the three outputs; every piece starts on a 256-byte boundary
  auto up = [](size_t b) { return (b + 255) & ~(size_t)255; };
  const size_t o_par = 0, o_tgt = o_par + up(n * P * 8), o_rr = o_tgt + up(2 * NN * 8), o_g = o_rr + up((size_t)n * 8),
               o_A = o_g + up(n * P * 8), io_end = o_A + up(n * P * P * 8);
  if (io_end > QSB_STATIC_IO_BYTES)
    return qsb_fail(QSB_ERR_UNSUPPORTED, "qsb_eval_lm_system_static: " + std::to_string(n) + " instances x " + std::to_string(P) +
                    " parameters need " + std::to_string(io_end) + " bytes of staging, the static limit is " +
                    std::to_string(QSB_STATIC_IO_BYTES) + " (use qsb_batch_run(QSB_RUN_LMSYS) for larger batches)");
  size_t per_cta = rows * NN + ((rows * (rows + 1) / 2 + 1) & ~(size_t)1) + (size_t)qsb_count_param_ops(prog->ops.data(), prog->n_ops) * 2 * NN;
  per_cta = (per_cta + 31) & ~(size_t)31;  // 256-byte slabs
  const size_t avail = (QSB_STATIC_SCRATCH_BYTES - QSB_STATIC_IO_BYTES) / 8;
  if (per_cta > avail)
    return qsb_fail(QSB_ERR_UNSUPPORTED, "qsb_eval_lm_system_static: one instance needs " + std::to_string(per_cta * 8) +
                    " bytes of scratch, the static limit is " + std::to_string(avail * 8));
  int rc = qsb_prepare_kernel(qsb_k_lmsys3_static<NQ>, smem);
  if (rc) return rc;
  const int threads = QsbQGeom<NQ>::WARPS * 32;
  const int resident = d.sm_count * qsb_resident_ctas(qsb_k_lmsys3_static<NQ>, threads, smem);
  const int grid = (int)std::max<size_t>(1, std::min<size_t>(std::min<size_t>((size_t)n, (size_t)resident), avail / per_cta));
  QsbDevProgram dp;
  dp.ops = prog->d_ops[slot]; dp.consts = prog->d_consts[slot]; dp.pscale = prog->d_pscale[slot]; dp.slotmap = prog->d_slotmap[slot];
  dp.n_ops = prog->n_ops; dp.n_params = prog->n_params;
  double* s_par = (double*)(base + o_par);
  double* s_tgt = (double*)(base + o_tgt);
  double* s_rr = (double*)(base + o_rr);
  double* s_g = (double*)(base + o_g);
  double* s_A = (double*)(base + o_A);
  double* s_scratch = (double*)(base + QSB_STATIC_IO_BYTES);
  // caller buffers may be host or device memory: cudaMemcpyDefault resolves the direction from the pointers
  if (n * P) QSB_CUDA(cudaMemcpyAsync(s_par, params, n * P * 8, cudaMemcpyDefault, d.stream));
  QSB_CUDA(cudaMemcpyAsync(s_tgt, target, 2 * NN * 8, cudaMemcpyDefault, d.stream));
  QSB_LAUNCH((qsb_k_lmsys3_static<NQ>), grid, threads, smem, d.stream, dp, n, (const double*)s_par, (const double*)s_tgt, s_rr, s_g, s_A,
             s_scratch, per_cta);
  QSB_CUDA(cudaGetLastError());
  QSB_CUDA(cudaMemcpyAsync(out_rr, s_rr, (size_t)n * 8, cudaMemcpyDefault, d.stream));
  if (n * P) {
    QSB_CUDA(cudaMemcpyAsync(out_g, s_g, n * P * 8, cudaMemcpyDefault, d.stream));
    QSB_CUDA(cudaMemcpyAsync(out_A, s_A, n * P * P * 8, cudaMemcpyDefault, d.stream));
  }
  QSB_CUDA(cudaStreamSynchronize(d.stream));
  QSB_CUDA(cudaGetLastError());
  return QSB_OK;
}

extern "C" int qsb_eval_lm_system_static(qsb_ctx* ctx, int device_slot, const qsb_program* prog, int n_instances,
                                         const double* params, const double* target, double* out_rr, double* out_g, double* out_A) {
  if (!ctx || !prog || !target || !out_rr) return qsb_fail(QSB_ERR_INVALID, "NULL argument");
  if (device_slot < 0 || device_slot >= (int)ctx->devs.size()) return qsb_fail(QSB_ERR_INVALID, "device slot out of range");
  if (n_instances < 0) return qsb_fail(QSB_ERR_INVALID, "negative instance count");
  if ((int)prog->d_ops.size() <= device_slot) return qsb_fail(QSB_ERR_INVALID, "program was created on another context");
  if (prog->n_params > 0 && n_instances > 0 && (!params || !out_g || !out_A)) return qsb_fail(QSB_ERR_INVALID, "NULL argument");
  if (!prog->fast_ok || prog->n_qubits < 3)
    return qsb_fail(QSB_ERR_UNSUPPORTED, "qsb_eval_lm_system_static needs a one-qubit-leaf + CNOT program on 3..5 qubits");
  if (n_instances == 0) return QSB_OK;
  const QsbDevice& d = ctx->devs[device_slot];
  unsigned char* base = nullptr;
  int rc = static_scratch_of(d.id, &base);
  if (rc) return rc;
  std::lock_guard<std::mutex> one_call(g_static_mu[d.id]);
  QSB_CUDA(cudaSetDevice(d.id));
  {  // the body-frame form needs a unitary target: check a host copy (16 KB at most, on the stack; the pointer may be device memory)
    double t_host[2 * 32 * 32];
    const size_t NN = (size_t)1 << (2 * prog->n_qubits);
    QSB_CUDA(cudaMemcpyAsync(t_host, target, 2 * NN * 8, cudaMemcpyDefault, d.stream));
    QSB_CUDA(cudaStreamSynchronize(d.stream));
    if (!qsb_target_is_unitary(t_host, (int)prog->dim))
      return qsb_fail(QSB_ERR_UNSUPPORTED, "qsb_eval_lm_system_static needs a unitary target (use qsb_batch_run(QSB_RUN_LMSYS) otherwise)");
  }
  switch (prog->n_qubits) {
    case 3: return run_static<3>(d, prog, device_slot, n_instances, params, target, out_rr, out_g, out_A, base);
    case 4: return run_static<4>(d, prog, device_slot, n_instances, params, target, out_rr, out_g, out_A, base);
    default: return run_static<5>(d, prog, device_slot, n_instances, params, target, out_rr, out_g, out_A, base);
  }
}
