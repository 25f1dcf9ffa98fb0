// Host side of the batched solvers (C ABI: qsb_solve_lm / qsb_solve_lbfgs).  One source, three translation units
// (build.py compiles them in parallel): QSB_SOLVE_TU = 0: Levenberg-Marquardt, 1..4 qubits + qsb_solve_lm;
// 1: L-BFGS + qsb_solve_lbfgs; 2: Levenberg-Marquardt, 5 qubits; 3: the same with the TF32 Gram product (qsb_k_lm3<5, true>);
// 4: the 4-warp, two-CTAs-per-SM shape of the body-frame LM kernel for 3 and 4 qubits (qsb_k_lm3<n, tf32, 4>).
#ifndef QSB_SOLVE_TU
#error "compile with -DQSB_SOLVE_TU=0|1|2|3|4"
#endif
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>

#include "qsb_solve_host.h"
#include "qsb_launch.h"
#if QSB_SOLVE_TU == 1
#include "qsb_lbfgs.cuh"
#include "qsb_lbfgs2.cuh"
#define QSB_TU_METHOD QSB_SOLVE_LBFGS
#else
#include "qsb_lm2.cuh"
#include "qsb_lm3.cuh"
#define QSB_TU_METHOD QSB_SOLVE_LM
#endif

#if QSB_SOLVE_TU == 4
// ---------------------------------------------------------------------------- body-frame LM, 4 warps per instance
template <int NQ, bool TF32>
static int launch_lm3_w4(const SolveJob& job, qsb_batch* b, QsbSolveArgs& args, const QsbDevice& d) {
  const size_t NN = (size_t)1 << (2 * NQ), maxP = (size_t)b->max_params;
  const int threads = 4 * 32;
  const size_t smem = qsb_lm3_smem_bytes(NQ, b->max_ops, (int)maxP);
  size_t per_cta = (maxP + 1) * NN + (maxP + 1) * (maxP + 2) / 2 + 2 + (size_t)b->max_pops * 2 * NN;  // slab, packed Gram, snapshots
  per_cta = (per_cta + 1) & ~(size_t)1;
  int rc = qsb_prepare_kernel(qsb_k_lm3<NQ, TF32, 4>, smem);
  if (rc) return rc;
  const int ctas = qsb_resident_ctas(qsb_k_lm3<NQ, TF32, 4>, threads, smem);
  const int grid = std::max(1, std::min(args.count, d.sm_count * ctas));
  QsbBuf& scratch = job.ctx->devs[b->slot].solve_scratch;
  if ((rc = qsb_buf_reserve(scratch, (size_t)grid * per_cta * sizeof(double)))) return rc;
  args.scratch = (double*)scratch.p;
  args.scratch_per_cta = per_cta;
  QSB_LAUNCH((qsb_k_lm3<NQ, TF32, 4>), grid, threads, smem, d.stream, args);
  QSB_CUDA(cudaGetLastError());
  return QSB_OK;
}

int qsb_launch_solver_lm_w4(const SolveJob& job, qsb_batch* b, QsbSolveArgs& args, const QsbDevice& d, int nq, bool tf32) {
  if (nq == 3) return tf32 ? launch_lm3_w4<3, true>(job, b, args, d) : launch_lm3_w4<3, false>(job, b, args, d);
  if (nq == 4) return tf32 ? launch_lm3_w4<4, true>(job, b, args, d) : launch_lm3_w4<4, false>(job, b, args, d);
  return qsb_fail(QSB_ERR_UNSUPPORTED, "the 4-warp LM kernel exists for 3 and 4 qubits");
}
#elif QSB_SOLVE_TU != 1
// ---------------------------------------------------------------------------- Levenberg-Marquardt launches
// TF32: the body-frame kernel with the TF32 Gram product (qsb_solver_opts.gram_tf32); a separate instantiation, so that
// neither variant pays registers for the other's code
template <int NQ, bool TF32 = false>
static int launch_solver_nq(const SolveJob& job, qsb_batch* b, QsbSolveArgs& args, const QsbDevice& d) {
  const size_t NN = (size_t)1 << (2 * NQ);
  const size_t maxP = (size_t)b->max_params;
  const bool tensor_lm = NQ >= 3 && b->fast_ok && !b->force_generic && maxP + 1 <= QSB_CH_MAXROWS;
  // body-frame LM (qsb_lm3.cuh): unitary target, packed factor in shared memory; QSB_LM_KERNEL=lm2 forces the general kernel
  bool body_lm = tensor_lm && job.target_unitary && qsb_lm3_smem_bytes(NQ, b->max_ops, (int)maxP) <= 227 * 1024;
  if (const char* e = getenv("QSB_LM_KERNEL")) { if (strcmp(e, "lm2") == 0) body_lm = false; }
  if constexpr (NQ == 3 || NQ == 4) {
    // more than one instance per SM: two 4-warp CTAs per SM instead of one 8-warp CTA (QSB_LM_WARPS=4 / 8 forces either)
    // Only with the straggler rule: without it the wall time of a batch is its slowest instance (thousands of iterations against
    // a mean of 200, profiles/r2/w_solve_w4_ab.log), i.e. per-iteration latency, where the 8-warp shape is 15-20 % better.
    bool dense = body_lm && args.straggler_c > 0.0 && args.count > d.sm_count && qsb_lm3_smem_bytes(NQ, b->max_ops, (int)maxP) <= 110 * 1024;
    if (const char* e = getenv("QSB_LM_WARPS")) { const int v = atoi(e); if (v == 4) dense = body_lm; else if (v == 8) dense = false; }
    if (dense) return qsb_launch_solver_lm_w4(job, b, args, d, NQ, TF32);
  }
  int threads = (int)std::max<size_t>(32, NN / 2);
  size_t smem, per_cta;
  if (body_lm) {
    if constexpr (NQ >= 3) threads = QsbQGeom<NQ>::WARPS * 32;
    smem = qsb_lm3_smem_bytes(NQ, b->max_ops, (int)maxP);
    per_cta = (maxP + 1) * NN + (maxP + 1) * (maxP + 2) / 2 + 2 + (size_t)b->max_pops * 2 * NN;  // slab, packed Gram, snapshots
  } else if (tensor_lm) {
    threads = QSB_LM_WARPS * 32;
    args.chol_smem = qsb_lm2_smem_bytes(b->max_ops, (int)maxP, true) <= 224 * 1024;
    if (const char* e = getenv("QSB_LM_CHOL")) { if (strcmp(e, "global") == 0) args.chol_smem = 0; }  // profiling
    smem = qsb_lm2_smem_bytes(b->max_ops, (int)maxP, args.chol_smem != 0);
    // small batches: several SMs per instance (thread-block cluster); the Jacobian and the SYRK split across them
    int csize = 1;
#ifndef QSB_EMU
    {  // as many SMs per instance as the batch leaves free, but at least ~4 parameters per warp of the Jacobian phase
      const int nseg = QSB_LM_WARPS / QsbWarpGeom<(NQ >= 3 ? NQ : 3)>::WARPS;
      const int by_work = std::max(1, (int)maxP / (nseg * 4));
      csize = std::max(1, std::min(std::min(QSB_LM_MAX_CLUSTER, by_work), d.sm_count / std::max(1, args.count)));
      // A cluster lives inside one GPC, so fewer clusters are co-resident than SMs / size suggests.  All instances must
      // run at once: an instance that waits for a free cluster slot adds its whole solve to the wall time.
      if constexpr (NQ >= 3) {
        if (csize > 1 && qsb_prepare_kernel(qsb_k_lm2<NQ>, smem) == QSB_OK) {
          for (; csize > 1; --csize) {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(csize * args.count); cfg.blockDim = dim3(threads); cfg.dynamicSmemBytes = smem;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeClusterDimension;
            attr[0].val.clusterDim.x = csize; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
            cfg.attrs = attr; cfg.numAttrs = 1;
            int active = 0;
            if (cudaOccupancyMaxActiveClusters(&active, qsb_k_lm2<NQ>, &cfg) == cudaSuccess && active >= args.count) break;
          }
          cudaGetLastError();
        }
      }
    }
#endif
    // the batch-level stop rules need one CTA per instance (qsb_k_lm2); a capped tail beats a faster iteration by far
    if (args.straggler_c > 0.0 || args.sibling_f > 0.0) csize = 1;
    if (const char* e = getenv("QSB_LM_CLUSTER")) { int v = atoi(e); if (v >= 1 && v <= QSB_LM_MAX_CLUSTER) csize = v; }
    args.cluster = csize;
    per_cta = (maxP + 1) * 2 * NN + 2 * (maxP + 1) * (maxP + 1) + (args.chol_smem ? 0 : (size_t)csize * (maxP + 1) * (maxP + 1));
  } else {
    smem = qsb_lm_smem_bytes(NN, QsbDims<NQ>::TRIG, b->max_ops, maxP);
    per_cta = qsb_lm_scratch_doubles(NN, maxP);
  }
  if (smem > 227 * 1024) return qsb_fail(QSB_ERR_UNSUPPORTED, "circuit too large for the solver's shared-memory working set");
  int rc = QSB_OK;
  int ctas;
  if (body_lm) {
    if constexpr (NQ >= 3) {
      if ((rc = qsb_prepare_kernel(qsb_k_lm3<NQ, TF32>, smem))) return rc;
      ctas = qsb_resident_ctas(qsb_k_lm3<NQ, TF32>, threads, smem);
    } else ctas = 1;
  } else if (tensor_lm) {
    if constexpr (NQ >= 3) {
      if ((rc = qsb_prepare_kernel(qsb_k_lm2<NQ>, smem))) return rc;
      ctas = qsb_resident_ctas(qsb_k_lm2<NQ>, threads, smem);
    } else ctas = 1;
  } else {
    int by_threads = 2048 / threads, by_smem = (int)((size_t)220 * 1024 / smem);
    ctas = std::max(1, std::min(std::min(by_threads, by_smem), 16));
  }
  per_cta = (per_cta + 1) & ~(size_t)1;  // 16-byte alignment of every CTA's slab (vector stores of the Jacobian rows)
  int grid = std::max(1, std::min(args.count, d.sm_count * ctas));
  if (tensor_lm && !body_lm) grid = std::max(1, std::min(args.count, d.sm_count * ctas / args.cluster)) * args.cluster;
  QsbBuf& scratch = job.ctx->devs[b->slot].solve_scratch;  // kept across calls (solve_shard holds the device's solve_mu)
  rc = qsb_buf_reserve(scratch, (size_t)(grid / args.cluster) * per_cta * sizeof(double));  // per cluster / CTA
  if (rc) return rc;
  args.scratch = (double*)scratch.p;
  args.scratch_per_cta = per_cta;
  if (body_lm) {
    if constexpr (NQ >= 3) QSB_LAUNCH((qsb_k_lm3<NQ, TF32>), grid, threads, smem, d.stream, args);
  } else if (tensor_lm) {
    if constexpr (NQ >= 3) {
#ifdef QSB_EMU
      if (args.cluster > 1) QSB_LAUNCH_CLUSTER((qsb_k_lm2<NQ>), grid, threads, smem, args.cluster, args);
      else QSB_LAUNCH((qsb_k_lm2<NQ>), grid, threads, smem, d.stream, args);
#else
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(grid); cfg.blockDim = dim3(threads); cfg.dynamicSmemBytes = smem; cfg.stream = d.stream;
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeClusterDimension;
      attr[0].val.clusterDim.x = args.cluster; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
      cfg.attrs = attr; cfg.numAttrs = 1;
      QSB_CUDA(cudaLaunchKernelEx(&cfg, qsb_k_lm2<NQ>, args));
#endif
    }
  } else {
    if ((rc = qsb_prepare_kernel(qsb_k_lm<NQ>, smem))) return rc;
    QSB_LAUNCH((qsb_k_lm<NQ>), grid, threads, smem, d.stream, args);
  }
  QSB_CUDA(cudaGetLastError());
  return QSB_OK;
}
#else
// ---------------------------------------------------------------------------- L-BFGS launches
template <int NQ>
static int launch_solver_nq(const SolveJob& job, qsb_batch* b, QsbSolveArgs& args, const QsbDevice& d) {
  const size_t NN = (size_t)1 << (2 * NQ);
  const int mem = job.opts.lbfgs_memory > 0 ? job.opts.lbfgs_memory : 10;
  const bool warp_lbfgs = NQ >= 3 && b->fast_ok && !b->force_generic && qsb_lbfgs2_smem_bytes(NQ, b->max_ops, b->max_params, mem) <= 200 * 1024;
  int threads = (int)std::max<size_t>(32, NN / 2);
  size_t smem;
  if (warp_lbfgs) {
    threads = 32 * ((1 << NQ) / 8);  // one warp per 8-column block
    smem = qsb_lbfgs2_smem_bytes(NQ, b->max_ops, b->max_params, mem);
  } else {
    smem = qsb_lbfgs_smem_bytes(NN, QsbDims<NQ>::TRIG, b->max_ops, b->max_params, mem);
  }
  if (smem > 227 * 1024) return qsb_fail(QSB_ERR_UNSUPPORTED, "circuit too large for the solver's shared-memory working set");
  int rc = QSB_OK;
  int ctas;
  if (warp_lbfgs) {
    if constexpr (NQ >= 3) {
      if ((rc = qsb_prepare_kernel(qsb_k_lbfgs2<NQ>, smem))) return rc;
      ctas = qsb_resident_ctas(qsb_k_lbfgs2<NQ>, threads, smem);
    } else ctas = 1;
  } else {
    int by_threads = 2048 / threads, by_smem = (int)((size_t)220 * 1024 / smem);
    ctas = std::max(1, std::min(std::min(by_threads, by_smem), 16));
  }
  const int grid = std::max(1, std::min(args.count, d.sm_count * ctas));
  QsbBuf& scratch = job.ctx->devs[b->slot].solve_scratch;
  rc = qsb_buf_reserve(scratch, (size_t)grid * 2 * sizeof(double));
  if (rc) return rc;
  args.scratch = (double*)scratch.p;
  args.scratch_per_cta = 2;
  if (warp_lbfgs) {
    if constexpr (NQ >= 3) QSB_LAUNCH((qsb_k_lbfgs2<NQ>), grid, threads, smem, d.stream, args);
  } else {
    if ((rc = qsb_prepare_kernel(qsb_k_lbfgs<NQ>, smem))) return rc;
    QSB_LAUNCH((qsb_k_lbfgs<NQ>), grid, threads, smem, d.stream, args);
  }
  QSB_CUDA(cudaGetLastError());
  return QSB_OK;
}
#endif

#if QSB_SOLVE_TU < 2
// Interpreter solver kernels for a geometry code GQ (qsb_generic.cuh): the radix-3 programs, which have no warp kernels
template <int GQ>
static int launch_solver_interp(const SolveJob& job, qsb_batch* b, QsbSolveArgs& args, const QsbDevice& d) {
  const size_t NN = (size_t)QsbDims<GQ>::NN, maxP = (size_t)b->max_params;
  const int threads = QsbGeom<GQ>::T;
#if QSB_SOLVE_TU == 0
  const size_t smem = qsb_lm_smem_bytes(NN, QsbDims<GQ>::TRIG, b->max_ops, maxP);
  size_t per_cta = qsb_lm_scratch_doubles(NN, maxP);
#else
  const int mem = job.opts.lbfgs_memory > 0 ? job.opts.lbfgs_memory : 10;
  const size_t smem = qsb_lbfgs_smem_bytes(NN, QsbDims<GQ>::TRIG, b->max_ops, b->max_params, mem);
  size_t per_cta = 2;
#endif
  if (smem > 227 * 1024) return qsb_fail(QSB_ERR_UNSUPPORTED, "circuit too large for the solver's shared-memory working set");
  per_cta = (per_cta + 1) & ~(size_t)1;
  const int by_threads = 2048 / threads, by_smem = (int)((size_t)220 * 1024 / smem);
  const int ctas = std::max(1, std::min(std::min(by_threads, by_smem), 16));
  const int grid = std::max(1, std::min(args.count, d.sm_count * ctas));
  QsbBuf& scratch = job.ctx->devs[b->slot].solve_scratch;
  int rc = qsb_buf_reserve(scratch, (size_t)grid * per_cta * sizeof(double));
  if (rc) return rc;
  args.scratch = (double*)scratch.p;
  args.scratch_per_cta = per_cta;
#if QSB_SOLVE_TU == 0
  if ((rc = qsb_prepare_kernel(qsb_k_lm<GQ>, smem))) return rc;
  QSB_LAUNCH((qsb_k_lm<GQ>), grid, threads, smem, d.stream, args);
#else
  if ((rc = qsb_prepare_kernel(qsb_k_lbfgs<GQ>, smem))) return rc;
  QSB_LAUNCH((qsb_k_lbfgs<GQ>), grid, threads, smem, d.stream, args);
#endif
  QSB_CUDA(cudaGetLastError());
  return QSB_OK;
}
#endif

#if QSB_SOLVE_TU == 4
// (launch function above)
#elif QSB_SOLVE_TU == 2
int qsb_launch_solver_lm5(const SolveJob& job, qsb_batch* b, QsbSolveArgs& args, const QsbDevice& d) {
  return launch_solver_nq<5, false>(job, b, args, d);
}
#elif QSB_SOLVE_TU == 3
int qsb_launch_solver_lm5t(const SolveJob& job, qsb_batch* b, QsbSolveArgs& args, const QsbDevice& d) {
  return launch_solver_nq<5, true>(job, b, args, d);
}
#else
static int solve_shard(const SolveJob& job, int slot, int i0, int i1, const std::vector<long long>& poff) {
  if (i1 <= i0) return QSB_OK;
  std::lock_guard<std::mutex> device_lock(*job.ctx->devs[slot].solve_mu);
  qsb_batch* b = nullptr;
  int rc = qsb_batch_create(job.ctx, slot, job.n_programs, job.progs, i1 - i0, job.prog_of_instance + i0, &b);
  if (rc) return rc;
  auto done = [&](int code) { std::string keep = qsb_get_error(); qsb_batch_destroy(b); qsb_set_error(keep); return code; };
  const size_t np = (size_t)(poff[i1] - poff[i0]);
  const int n = i1 - i0;
  if ((rc = qsb_batch_set_params(b, job.x0 + poff[i0], np))) return done(rc);
  if ((rc = qsb_batch_set_target(b, job.target))) return done(rc);
  // outputs: out_a = x [np] then f [n]; out_b = iters [n] + status [n] as int32
  if ((rc = qsb_pool_reserve(job.ctx->devs[slot], b->out_a, (np + n + 1) * sizeof(double)))) return done(rc);
  const size_t n_ctl = (size_t)QSB_CTL_SOLVED + job.n_programs;
  if ((rc = qsb_pool_reserve(job.ctx->devs[slot], b->out_b, ((size_t)(2 * n + 2) + n_ctl) * sizeof(int32_t)))) return done(rc);
  const QsbDevice& d = job.ctx->devs[slot];
  QsbSolveArgs args;
  args.progs = b->d_progs; args.prog_idx = b->d_prog_idx; args.param_off = b->d_param_off; args.x0 = b->d_params;
  args.target = b->d_target; args.count = n; args.max_ops = b->max_ops; args.max_params = std::max(1, b->max_params);
  args.max_iters = job.opts.max_iters;
  args.ftol = job.opts.ftol > 0 ? job.opts.ftol : 1e-8;
  args.xtol = job.opts.xtol > 0 ? job.opts.xtol : 1e-8;
  args.gtol = job.opts.gtol > 0 ? job.opts.gtol : 1e-8;
  args.stop_f = job.opts.stop_value;
  args.lbfgs_memory = job.opts.lbfgs_memory > 0 ? job.opts.lbfgs_memory : 10;
  args.out_x = (double*)b->out_a.p; args.out_f = args.out_x + np;
  args.out_iters = (int32_t*)b->out_b.p; args.out_status = args.out_iters + n;
  args.ctl = (int*)(args.out_status + n + 2);
  args.gram_tf32 = job.opts.gram_tf32 > 0 ? 1 : 0;
  args.straggler_c = job.opts.straggler_factor;
  args.sibling_f = job.opts.sibling_stop;
  args.straggler_min = job.opts.straggler_min_iters > 0 ? (int)job.opts.straggler_min_iters : 32;
  if (cudaMemsetAsync(args.ctl, 0, n_ctl * sizeof(int32_t), job.ctx->devs[slot].stream) != cudaSuccess)
    return done(qsb_fail(QSB_ERR_CUDA, "cudaMemsetAsync of the solver control block failed"));
  b->max_params = args.max_params;
  if (args.straggler_c > 0.0) { args.mu_down = 0.7; args.mu_streak = 1; }  // capped batches: the cheaper schedule (QsbSolveArgs)
  if (const char* e = getenv("QSB_LM_MU_DOWN")) { const double v = atof(e); if (v > 0.0 && v <= 1.0) args.mu_down = v; }
  if (const char* e = getenv("QSB_LM_MU_UP")) { const double v = atof(e); if (v > 1.0) args.mu_up = v; }
  if (const char* e = getenv("QSB_LM_MU_STREAK")) args.mu_streak = atoi(e);
  if (const char* e = getenv("QSB_LM_MU_FAST")) args.mu_fast = atoi(e);
  args.prof = nullptr;
  args.chol_smem = 0;
  args.cluster = 1;
  long long* d_prof = nullptr;
  if (getenv("QSB_LM_PROFILE")) {
    if (cudaMalloc((void**)&d_prof, 12 * sizeof(long long)) == cudaSuccess) { cudaMemset(d_prof, 0, 12 * sizeof(long long)); args.prof = d_prof; }
  }
  if (b->radix == 3) {
    switch (b->n_qubits) {
      case 1: rc = launch_solver_interp<QSB_GEOM_QUTRIT + 1>(job, b, args, d); break;
      case 2: rc = launch_solver_interp<QSB_GEOM_QUTRIT + 2>(job, b, args, d); break;
      case 3: rc = launch_solver_interp<QSB_GEOM_QUTRIT + 3>(job, b, args, d); break;
      default: rc = qsb_fail(QSB_ERR_UNSUPPORTED, "unsupported qutrit count");
    }
  } else
  switch (b->n_qubits) {
    case 1: rc = launch_solver_nq<1>(job, b, args, d); break;
    case 2: rc = launch_solver_nq<2>(job, b, args, d); break;
#if QSB_SOLVE_TU == 0
    case 3: rc = args.gram_tf32 ? launch_solver_nq<3, true>(job, b, args, d) : launch_solver_nq<3, false>(job, b, args, d); break;
    case 4: rc = args.gram_tf32 ? launch_solver_nq<4, true>(job, b, args, d) : launch_solver_nq<4, false>(job, b, args, d); break;
    case 5: rc = args.gram_tf32 ? qsb_launch_solver_lm5t(job, b, args, d) : qsb_launch_solver_lm5(job, b, args, d); break;
#else
    case 3: rc = launch_solver_nq<3>(job, b, args, d); break;
    case 4: rc = launch_solver_nq<4>(job, b, args, d); break;
    case 5: rc = launch_solver_nq<5>(job, b, args, d); break;
#endif
    default: rc = qsb_fail(QSB_ERR_UNSUPPORTED, "unsupported qubit count");
  }
  if (rc) return done(rc);
  { std::lock_guard<std::mutex> g(job.ctx->mu); job.ctx->launches += 1; }
#define QSB_S(call) do { cudaError_t _e = (call); if (_e != cudaSuccess) return done(qsb_cuda_fail(_e, #call)); } while (0)
  if (np) QSB_S(cudaMemcpyAsync(job.out_x + poff[i0], args.out_x, np * sizeof(double), cudaMemcpyDeviceToHost, d.stream));
  QSB_S(cudaMemcpyAsync(job.out_f + i0, args.out_f, n * sizeof(double), cudaMemcpyDeviceToHost, d.stream));
  if (job.out_iters) QSB_S(cudaMemcpyAsync(job.out_iters + i0, args.out_iters, n * sizeof(int32_t), cudaMemcpyDeviceToHost, d.stream));
  if (job.out_status) QSB_S(cudaMemcpyAsync(job.out_status + i0, args.out_status, n * sizeof(int32_t), cudaMemcpyDeviceToHost, d.stream));
  QSB_S(cudaStreamSynchronize(d.stream));
  QSB_S(cudaGetLastError());
  if (d_prof) {
    long long hp[12];
    cudaMemcpy(hp, d_prof, sizeof(hp), cudaMemcpyDeviceToHost);
    cudaFree(d_prof);
    const double it = hp[5] > 0 ? (double)hp[5] : 1.0;
    fprintf(stderr, "[qsb lm profile] instances=%d jac_evals=%lld attempts=%lld cycles/jac_eval: jac=%.0f syrk=%.0f cholesky=%.0f (fill=%.0f diag=%.0f panel=%.0f trailing=%.0f) solve=%.0f trial=%.0f\n", n,
            hp[5], hp[10], hp[0] / it, hp[1] / it, hp[2] / it, hp[6] / it, hp[7] / it, hp[8] / it, hp[9] / it, hp[3] / it, hp[4] / it);
  }
#undef QSB_S
  return done(QSB_OK);
}

static int solve_dispatch(SolveJob job) {
  if (!job.ctx || !job.progs || !job.target || !job.out_f || (job.n_instances > 0 && !job.prog_of_instance))
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
  if (off > 0 && (!job.x0 || !job.out_x)) return qsb_fail(QSB_ERR_INVALID, "x0 / out_x is NULL");
  job.target_unitary = job.method == QSB_SOLVE_LM && qsb_target_is_unitary(job.target, (int)job.progs[0]->dim);
  const int used = std::min((int)job.ctx->devs.size(), job.n_instances);
  if (used == 1) return solve_shard(job, 0, 0, job.n_instances, poff);
  // Contiguous shards of equal COST, not equal count: an LM iteration is dominated by the Gram
  // product, (P + 1)^2 N^2 / 2, so an instance weighs (P + 1)^2 (all instances of a batch share N).  LEAP children have
  // equal P; the beams of SearchCompiler and the multistart axis mix depths.
  std::vector<double> cum((size_t)job.n_instances + 1, 0.0);
  for (int i = 0; i < job.n_instances; ++i) { const double p1 = (double)(poff[i + 1] - poff[i]) + 1.0; cum[i + 1] = cum[i] + p1 * p1; }
  std::vector<int> cut((size_t)used + 1, 0);
  cut[used] = job.n_instances;
  for (int s = 1; s < used; ++s) {
    const double want = cum[job.n_instances] * s / used;
    int i = (int)(std::lower_bound(cum.begin(), cum.end(), want) - cum.begin());
    if (i > 0 && want - cum[i - 1] < cum[i] - want) --i;  // nearer boundary
    cut[s] = std::min(std::max(i, cut[s - 1] + 1), job.n_instances - (used - s));  // every shard keeps at least one instance
  }
  std::vector<int> rcs(used, QSB_OK);
  std::vector<std::string> errs(used);
  std::vector<std::thread> th;
  for (int s = 0; s < used; ++s) {
    const int i0 = cut[s], i1 = cut[s + 1];
    th.emplace_back([&, s, i0, i1] { rcs[s] = solve_shard(job, s, i0, i1, poff); if (rcs[s]) errs[s] = qsb_get_error(); });
  }
  for (auto& t : th) t.join();
  for (int s = 0; s < used; ++s) if (rcs[s]) return qsb_fail(rcs[s], errs[s]);
  return QSB_OK;
}

#if QSB_SOLVE_TU == 0
extern "C" int qsb_solve_lm(qsb_ctx* ctx, int n_programs, const qsb_program* const* progs, int n_instances,
                            const int32_t* prog_of_instance, const double* x0, const double* target,
                            const qsb_solver_opts* opts, double* out_x, double* out_f, int32_t* out_iters, int32_t* out_status) {
  qsb_solver_opts o{};
  if (opts) o = *opts;
  SolveJob j{ctx, n_programs, progs, n_instances, prog_of_instance, x0, target, o, QSB_SOLVE_LM, out_x, out_f, out_iters, out_status};
  return solve_dispatch(j);
}
#else
extern "C" int qsb_solve_lbfgs(qsb_ctx* ctx, int n_programs, const qsb_program* const* progs, int n_instances,
                               const int32_t* prog_of_instance, const double* x0, const double* target,
                               const qsb_solver_opts* opts, double* out_x, double* out_f, int32_t* out_iters, int32_t* out_status) {
  qsb_solver_opts o{};
  if (opts) o = *opts;
  SolveJob j{ctx, n_programs, progs, n_instances, prog_of_instance, x0, target, o, QSB_SOLVE_LBFGS, out_x, out_f, out_iters, out_status};
  return solve_dispatch(j);
}
#endif
#endif  // QSB_SOLVE_TU < 2
