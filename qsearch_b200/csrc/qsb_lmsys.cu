// QSB_RUN_LMSYS launches (own translation unit: the library is compiled as several units in parallel, build.py).
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "qsb_host.h"
#include "qsb_lm2.cuh"
#include "qsb_lm3.cuh"
#include "qsb_launch.h"

// QSB_RUN_LMSYS: tensor-core kernels only (one-qubit leaves + CNOT, 3..5 qubits)
template <int NQ>
static int qsb_launch_lmsys_nq(qsb_batch* b, const QsbBatchArgs& args, cudaStream_t st, int sm_count) {
  const size_t NN = (size_t)1 << (2 * NQ), maxP = (size_t)args.max_params;
  bool body = b->target_unitary;  // body-frame form (qsb_lm3.cuh); QSB_LM_KERNEL=lm2 forces the general one
  if (const char* e = getenv("QSB_LM_KERNEL")) { if (strcmp(e, "lm2") == 0) body = false; }
  const size_t smem = body ? qsb_lmsys3_smem_bytes(NQ, args.max_ops, args.max_params) : qsb_lmsys_smem_bytes(args.max_ops, args.max_params);
  if (smem > 200 * 1024) return qsb_fail(QSB_ERR_UNSUPPORTED, "program too long for QSB_RUN_LMSYS");
  int rc = body ? qsb_prepare_kernel(qsb_k_lmsys3<NQ>, smem) : qsb_prepare_kernel(qsb_k_lmsys<NQ>, smem);
  if (rc) return rc;
  const int threads = body ? QsbQGeom<NQ>::WARPS * 32 : QSB_LM_WARPS * 32;
  const int resident = body ? qsb_resident_ctas(qsb_k_lmsys3<NQ>, threads, smem) : qsb_resident_ctas(qsb_k_lmsys<NQ>, threads, smem);
  const int grid = std::max(1, std::min(args.count, sm_count * resident));
  size_t per_cta = (maxP + 1) * (body ? 1 : 2) * NN + (maxP + 1) * (maxP + 1) + (body ? 2 + (size_t)b->max_pops * 2 * NN : 0);
  per_cta = (per_cta + 1) & ~(size_t)1;
  const size_t need = (size_t)grid * per_cta * sizeof(double);
  if (need > b->solver_scratch_bytes) {
    if (b->solver_scratch) cudaFree(b->solver_scratch);
    b->solver_scratch = nullptr; b->solver_scratch_bytes = 0;
    if (cudaMalloc(&b->solver_scratch, need) != cudaSuccess) return qsb_fail(QSB_ERR_NOMEM, "device allocation of the LM-system scratch failed");
    b->solver_scratch_bytes = need;
  }
  if (!b->d_sq_off) {
    std::vector<long long> sq((size_t)b->n_instances + 1);
    long long off = 0;
    for (int i = 0; i < b->n_instances; ++i) { sq[i] = off; const long long p = b->param_off[i + 1] - b->param_off[i]; off += p * p; }
    sq[b->n_instances] = off;
    QSB_CUDA(cudaMalloc((void**)&b->d_sq_off, sizeof(long long) * sq.size()));
    QSB_CUDA(cudaMemcpy(b->d_sq_off, sq.data(), sizeof(long long) * sq.size(), cudaMemcpyHostToDevice));
  }
  if (body) QSB_LAUNCH((qsb_k_lmsys3<NQ>), grid, threads, smem, st, args, (double*)b->solver_scratch, per_cta, (const long long*)b->d_sq_off);
  else QSB_LAUNCH((qsb_k_lmsys<NQ>), grid, threads, smem, st, args, (double*)b->solver_scratch, per_cta, (const long long*)b->d_sq_off);
  QSB_CUDA(cudaGetLastError());
  return QSB_OK;
}

int qsb_launch_lmsys(qsb_batch* b, const QsbBatchArgs& args, cudaStream_t st, int sm_count) {
  if (!b->fast_ok || b->n_qubits < 3)
    return qsb_fail(QSB_ERR_UNSUPPORTED, "QSB_RUN_LMSYS needs a one-qubit-leaf + CNOT program on 3..5 qubits");
  switch (b->n_qubits) {
    case 3: return qsb_launch_lmsys_nq<3>(b, args, st, sm_count);
    case 4: return qsb_launch_lmsys_nq<4>(b, args, st, sm_count);
    default: return qsb_launch_lmsys_nq<5>(b, args, st, sm_count);
  }
}
