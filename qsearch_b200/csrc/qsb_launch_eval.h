// Kernel selection and launch geometry for one batch run (host side).  Included by qsb_api.cu ONLY (see qsb_launch.h).
#pragma once
#include <algorithm>

#include "qsb_host.h"
#include "qsb_launch.h"
#include "qsb_generic.cuh"
#include "qsb_warp.cuh"
#include "qsb_hs_warp.cuh"

struct QsbLaunchShape {
  int threads;
  size_t smem;
  int ctas_per_sm;
};

// NQ: geometry code (qsb_generic.cuh)
template <int NQ>
static inline QsbLaunchShape qsb_shape_generic(int what, int max_ops, int max_params) {
  const size_t NN = (size_t)QsbDims<NQ>::NN;
  QsbLaunchShape s;
  s.threads = QsbGeom<NQ>::T;
  const size_t trig = (size_t)max_ops * QsbDims<NQ>::TRIG * sizeof(double);
  switch (what) {
    case QSB_RUN_MATRIX: case QSB_RUN_DIST: s.smem = (NN + 32) * 16 + trig; break;
    case QSB_RUN_HS: s.smem = (2 * NN + 32 + QSB_HS_WPART + (size_t)max_params) * 16 + trig; break;
    default: s.smem = 3 * NN * 16 + trig; break;
  }
  s.smem += 16;
  int by_threads = 2048 / s.threads, by_smem = (int)((size_t)220 * 1024 / s.smem);
  s.ctas_per_sm = std::max(1, std::min(std::min(by_threads, by_smem), 32));
  return s;
}

template <int NQ>
static int qsb_launch_generic_nq(qsb_batch* b, int what, const QsbBatchArgs& args, cudaStream_t st, int sm_count) {
  QsbLaunchShape sh = qsb_shape_generic<NQ>(what, b->max_ops, b->max_params);
  if (sh.smem > 227 * 1024) return qsb_fail(QSB_ERR_UNSUPPORTED, "program too long for the shared-memory gate cache");
  int grid = std::max(1, std::min(args.count, sm_count * sh.ctas_per_sm));
  int rc = QSB_OK;
  switch (what) {
    case QSB_RUN_MATRIX:
      if ((rc = qsb_prepare_kernel(qsb_k_forward<NQ, QSB_RUN_MATRIX>, sh.smem))) return rc;
      QSB_LAUNCH((qsb_k_forward<NQ, QSB_RUN_MATRIX>), grid, sh.threads, sh.smem, st, args);
      break;
    case QSB_RUN_DIST:
      if ((rc = qsb_prepare_kernel(qsb_k_forward<NQ, QSB_RUN_DIST>, sh.smem))) return rc;
      QSB_LAUNCH((qsb_k_forward<NQ, QSB_RUN_DIST>), grid, sh.threads, sh.smem, st, args);
      break;
    case QSB_RUN_MATJAC:
      if ((rc = qsb_prepare_kernel(qsb_k_jac<NQ, QSB_RUN_MATJAC>, sh.smem))) return rc;
      QSB_LAUNCH((qsb_k_jac<NQ, QSB_RUN_MATJAC>), grid, sh.threads, sh.smem, st, args);
      break;
    case QSB_RUN_RESID:
      if ((rc = qsb_prepare_kernel(qsb_k_jac<NQ, QSB_RUN_RESID>, sh.smem))) return rc;
      QSB_LAUNCH((qsb_k_jac<NQ, QSB_RUN_RESID>), grid, sh.threads, sh.smem, st, args);
      break;
    case QSB_RUN_HS:
      if ((rc = qsb_prepare_kernel(qsb_k_hs<NQ>, sh.smem))) return rc;
      QSB_LAUNCH((qsb_k_hs<NQ>), grid, sh.threads, sh.smem, st, args);
      break;
    default:
      return qsb_fail(QSB_ERR_INVALID, "unknown run kind");
  }
  QSB_CUDA(cudaGetLastError());
  return QSB_OK;
}

static inline int qsb_launch_generic(qsb_batch* b, int what, const QsbBatchArgs& args, cudaStream_t st, int sm_count) {
  if (b->radix == 3) {
    switch (b->n_qubits) {
      case 1: return qsb_launch_generic_nq<QSB_GEOM_QUTRIT + 1>(b, what, args, st, sm_count);
      case 2: return qsb_launch_generic_nq<QSB_GEOM_QUTRIT + 2>(b, what, args, st, sm_count);
      case 3: return qsb_launch_generic_nq<QSB_GEOM_QUTRIT + 3>(b, what, args, st, sm_count);
      default: return qsb_fail(QSB_ERR_UNSUPPORTED, "unsupported qutrit count");
    }
  }
  switch (b->n_qubits) {
    case 1: return qsb_launch_generic_nq<1>(b, what, args, st, sm_count);
    case 2: return qsb_launch_generic_nq<2>(b, what, args, st, sm_count);
    case 3: return qsb_launch_generic_nq<3>(b, what, args, st, sm_count);
    case 4: return qsb_launch_generic_nq<4>(b, what, args, st, sm_count);
    case 5: return qsb_launch_generic_nq<5>(b, what, args, st, sm_count);
    default: return qsb_fail(QSB_ERR_UNSUPPORTED, "unsupported qubit count");
  }
}

// ---------------------------------------------------------------------------- warp (register / DMMA) kernels
// one CTA per instance (1 warp, n = 5: 4 quadrant warps), grid = SMs x resident CTAs (persistent CTAs stride over the instances).
#define QSB_WARP_NOFIT 1  // internal: the program's tables do not fit the warp kernel's shared memory -> use the interpreter

template <int NQ, int WHAT>
static int qsb_launch_warp_kernel(const QsbBatchArgs& args, cudaStream_t st, int sm_count) {
  constexpr int wpb = QsbWarpGeom<NQ>::WARPS, threads = 32 * wpb;  // one CTA = the warps of one instance
  const size_t smem = (qsb_warp_jac_smem_doubles(args.max_ops, args.max_params) + (size_t)(wpb - 1) * QSB_WTILE_DOUBLES) * sizeof(double) + 16;
  if (smem > 200 * 1024) return QSB_WARP_NOFIT;
  int rc = qsb_prepare_kernel(qsb_k_warp_jac<NQ, WHAT>, smem);
  if (rc) return rc;
  const long long want = args.count;
  const int per_sm = qsb_resident_ctas(qsb_k_warp_jac<NQ, WHAT>, threads, smem), resident = sm_count * per_sm;
  const int grid = (int)std::max<long long>(1, std::min<long long>(want, resident));
  QsbBatchArgs a2 = args;
  a2.group = grid == resident ? per_sm : 0;
  QSB_LAUNCH((qsb_k_warp_jac<NQ, WHAT>), grid, threads, smem, st, a2);
  QSB_CUDA(cudaGetLastError());
  return QSB_OK;
}

static inline bool qsb_warp_path(const qsb_batch* b, int what) {
  return b->fast_ok && !b->force_generic && b->radix == 2 && b->n_qubits >= 3 &&
         (what == QSB_RUN_MATJAC || what == QSB_RUN_RESID || what == QSB_RUN_HS);
}

template <int NQ>
static int qsb_launch_warp_hs(const QsbBatchArgs& args, cudaStream_t st, int sm_count) {
  constexpr int threads = 32 * QsbHsGeom<NQ>::NCB, ipc = QsbHsGeom<NQ>::IPC;
  size_t slot = qsb_hs_slot_doubles(NQ, args.max_ops, args.max_params);
  slot = (slot + 1) & ~(size_t)1;
  const size_t smem = ipc * slot * sizeof(double) + 16;
  if (smem > 200 * 1024) return QSB_WARP_NOFIT;
  int rc = qsb_prepare_kernel(qsb_k_warp_hs<NQ>, smem);
  if (rc) return rc;
  const int want = (args.count + ipc - 1) / ipc;
  const int per_sm = qsb_resident_ctas(qsb_k_warp_hs<NQ>, threads, smem), resident = sm_count * per_sm;
  const int grid = std::max(1, std::min(want, resident));
  QsbBatchArgs a2 = args;
  a2.group = 0;  // plain instance order: the SM-local order of the Jacobian kernel is 1-2 % slower here (profiles/r1/aq_ab_hs.log)
  (void)per_sm;
  QSB_LAUNCH((qsb_k_warp_hs<NQ>), grid, threads, smem, st, a2);
  QSB_CUDA(cudaGetLastError());
  return QSB_OK;
}

static inline int qsb_launch_warp(qsb_batch* b, int what, const QsbBatchArgs& args, cudaStream_t st, int sm_count) {
  if (what == QSB_RUN_HS) {
    switch (b->n_qubits) {
      case 3: return qsb_launch_warp_hs<3>(args, st, sm_count);
      case 4: return qsb_launch_warp_hs<4>(args, st, sm_count);
      default: return qsb_launch_warp_hs<5>(args, st, sm_count);
    }
  }
  const bool mj = what == QSB_RUN_MATJAC;
  switch (b->n_qubits) {
    case 3: return mj ? qsb_launch_warp_kernel<3, QSB_RUN_MATJAC>(args, st, sm_count) : qsb_launch_warp_kernel<3, QSB_RUN_RESID>(args, st, sm_count);
    case 4: return mj ? qsb_launch_warp_kernel<4, QSB_RUN_MATJAC>(args, st, sm_count) : qsb_launch_warp_kernel<4, QSB_RUN_RESID>(args, st, sm_count);
    default: return mj ? qsb_launch_warp_kernel<5, QSB_RUN_MATJAC>(args, st, sm_count) : qsb_launch_warp_kernel<5, QSB_RUN_RESID>(args, st, sm_count);
  }
}

static inline int qsb_launch(qsb_batch* b, int what, const QsbBatchArgs& args, cudaStream_t st, int sm_count) {
  if (qsb_warp_path(b, what)) {
    const int rc = qsb_launch_warp(b, what, args, st, sm_count);
    if (rc != QSB_WARP_NOFIT) return rc;
  }
  return qsb_launch_generic(b, what, args, st, sm_count);
}
